// cifar_simt.cuh -- CUDA-core fp32 kernels for the CIFAR parts (B2P_MATH_FP32).
//
// Reference-grade parity mode: every contraction is an fp32 FMA chain, so the
// result differs from the reference's oneDNN/MKL CPU path only by summation
// order.  The tensor-core path (cifar_tc.cuh) is checked against this one at
// full batch size on the device and against the CPU oracle at small sizes.
//
//   part0_fused_f32   ModelPart0_2Node.forward  cifar_model_parts.py:37-42
//   fc1_splitk_f32    fc1 partial products      cifar_model_parts.py:55
//   fc_tail_f32       reduce + bias + ReLU + fc2 + softmax   :55-57
#pragma once
#include "common.cuh"
#include "ring_hooks.cuh"

namespace b2p {

// ---------------------------------------------------------------------------------
// Part 0, one sample per CTA iteration, everything between the input image and the
// flattened [4096] vector stays in shared memory:
//   conv1(3->32,3x3,p1)+ReLU+maxpool2  ->  s_a1 [32][18][18] (zero halo)
//   conv2(32->64,3x3,p1)+ReLU+maxpool2 ->  out[b][c*64 + h*8 + w]   (NCHW flatten,
//                                          the order fc1's weights assume)
// HBM traffic per sample = 12 288 B in + 16 384 B out; weights (77.6 KB) are read
// once per CTA and kept in shared memory across the CTA's samples.
// ---------------------------------------------------------------------------------
struct Part0Smem {
  float in[3][34][34];        // zero-padded input
  float w1[27][32];           // [ic*9+tap][oc]
  float b1[32];
  float a1[32][18][18];       // zero-padded conv1 output after pool
  float w2[32 * 9][64];       // [ic*9+tap][oc]
  float b2[64];
};

__global__ void __launch_bounds__(256, 1)
part0_fused_f32(const float* __restrict__ x,      // [B,3,32,32]
                const float* __restrict__ w1t,    // [27][32]  pre-transposed
                const float* __restrict__ b1,
                const float* __restrict__ w2t,    // [288][64] pre-transposed
                const float* __restrict__ b2,
                float* __restrict__ out,          // [B,4096]
                int batch) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Part0Smem& s = *reinterpret_cast<Part0Smem*>(smem_raw);
  const int tid = threadIdx.x;

  // weights -> smem (once per CTA)
  for (int i = tid; i < 27 * 32; i += 256) (&s.w1[0][0])[i] = w1t[i];
  for (int i = tid; i < 288 * 64 / 4; i += 256)
    reinterpret_cast<float4*>(&s.w2[0][0])[i] = reinterpret_cast<const float4*>(w2t)[i];
  if (tid < 32) s.b1[tid] = b1[tid];
  if (tid < 64) s.b2[tid] = b2[tid];
  // zero the halos once: interior cells are overwritten for every sample
  for (int i = tid; i < 3 * 34 * 34; i += 256) (&s.in[0][0][0])[i] = 0.f;
  for (int i = tid; i < 32 * 18 * 18; i += 256) (&s.a1[0][0][0])[i] = 0.f;
  __syncthreads();

  for (int b = blockIdx.x; b < batch; b += gridDim.x) {
    // ---- load image (coalesced float4) -------------------------------------------
    const float4* xin = reinterpret_cast<const float4*>(x + (size_t)b * 3072);
    for (int i = tid; i < 768; i += 256) {
      float4 v = __ldg(xin + i);
      int e = i * 4;
      int c = e >> 10, r = (e >> 5) & 31, col = e & 31;
      float* d = &s.in[c][r + 1][col + 1];
      d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
    }
    __syncthreads();

    // ---- conv1 + ReLU + pool: thread = one pooled pixel, all 32 channels -----------
    {
      const int py = tid >> 4, px = tid & 15;
      float win[3][4][4];
#pragma unroll
      for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
          for (int q = 0; q < 4; ++q) win[c][r][q] = s.in[c][2 * py + r][2 * px + q];
#pragma unroll 4
      for (int oc = 0; oc < 32; ++oc) {
        float a00 = s.b1[oc], a01 = a00, a10 = a00, a11 = a00;
#pragma unroll
        for (int c = 0; c < 3; ++c)
#pragma unroll
          for (int kh = 0; kh < 3; ++kh)
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) {
              float w = s.w1[c * 9 + kh * 3 + kw][oc];
              a00 = fmaf(win[c][kh][kw], w, a00);
              a01 = fmaf(win[c][kh][kw + 1], w, a01);
              a10 = fmaf(win[c][kh + 1][kw], w, a10);
              a11 = fmaf(win[c][kh + 1][kw + 1], w, a11);
            }
        float m = fmax_nan(fmax_nan(a00, a01), fmax_nan(a10, a11));
        s.a1[oc][py + 1][px + 1] = fmax_nan(m, 0.f);      // relu(max) == max(relu)
      }
    }
    __syncthreads();

    // ---- conv2 + ReLU + pool: thread = one pooled pixel x 16 output channels ---------
    {
      const int p = tid & 63;            // pooled pixel 0..63
      const int og = tid >> 6;           // oc group 0..3  (warp-uniform)
      const int py = p >> 3, px = p & 7;
      float acc[4][16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        float bv = s.b2[og * 16 + j];
        acc[0][j] = bv; acc[1][j] = bv; acc[2][j] = bv; acc[3][j] = bv;
      }
      for (int ic = 0; ic < 32; ++ic) {
        float win[4][4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const float2* rowp = reinterpret_cast<const float2*>(&s.a1[ic][2 * py + r][2 * px]);
          float2 lo = rowp[0], hi = rowp[1];
          win[r][0] = lo.x; win[r][1] = lo.y; win[r][2] = hi.x; win[r][3] = hi.y;
        }
#pragma unroll
        for (int kh = 0; kh < 3; ++kh)
#pragma unroll
          for (int kw = 0; kw < 3; ++kw) {
            const float4* wp = reinterpret_cast<const float4*>(&s.w2[ic * 9 + kh * 3 + kw][og * 16]);
#pragma unroll
            for (int j4 = 0; j4 < 4; ++j4) {
              float4 w = wp[j4];
              float wv[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
              for (int jj = 0; jj < 4; ++jj) {
                int j = j4 * 4 + jj;
                acc[0][j] = fmaf(win[kh][kw], wv[jj], acc[0][j]);
                acc[1][j] = fmaf(win[kh][kw + 1], wv[jj], acc[1][j]);
                acc[2][j] = fmaf(win[kh + 1][kw], wv[jj], acc[2][j]);
                acc[3][j] = fmaf(win[kh + 1][kw + 1], wv[jj], acc[3][j]);
              }
            }
          }
      }
      float* o = out + (size_t)b * 4096 + p;
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        float m = fmax_nan(fmax_nan(acc[0][j], acc[1][j]), fmax_nan(acc[2][j], acc[3][j]));
        o[(og * 16 + j) * 64] = fmax_nan(m, 0.f);          // 32 lanes -> 128 B contiguous
      }
    }
    __syncthreads();     // s.in / s.a1 are rewritten by the next sample
  }
}

// ---------------------------------------------------------------------------------
// fc1 as a split-K fp32 GEMM:  ws[z][m][n] = sum_{k in slice z} x[m][k] * W1[n][k]
// 64x64 output tile, 16-wide K steps, 256 threads x (4x4) register tile.
// M = batch (guarded), N = 512, K = 4096.  Deterministic (no atomics): slices are
// summed in a fixed order by fc_tail_f32.
// ---------------------------------------------------------------------------------
constexpr int FC1_N = 512, FC1_K = 4096;

__global__ void __launch_bounds__(256)
fc1_splitk_f32(const float* __restrict__ x,     // [M,4096]
               const float* __restrict__ w1,    // [512,4096]
               float* __restrict__ ws,          // [splits][M][512]
               int M, int k_per_split) {
  __shared__ float As[16][64 + 4];
  __shared__ float Bs[16][64 + 4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  const int kbeg = blockIdx.z * k_per_split, kend = kbeg + k_per_split;
  const int lr = tid >> 2, lc = (tid & 3) * 4;          // loader: row 0..63, k offset 0,4,8,12
  const int ty = tid >> 4, tx = tid & 15;               // compute: rows ty*4.., cols tx*4..
  float acc[4][4] = {};
  const bool a_ok = (m0 + lr) < M;
  const float* ap = x + (size_t)(m0 + lr) * FC1_K + lc;
  const float* bp = w1 + (size_t)(n0 + lr) * FC1_K + lc;
  for (int k0 = kbeg; k0 < kend; k0 += 16) {
    float4 av = a_ok ? __ldg(reinterpret_cast<const float4*>(ap + k0)) : make_float4(0, 0, 0, 0);
    float4 bv = __ldg(reinterpret_cast<const float4*>(bp + k0));
    As[lc + 0][lr] = av.x; As[lc + 1][lr] = av.y; As[lc + 2][lr] = av.z; As[lc + 3][lr] = av.w;
    Bs[lc + 0][lr] = bv.x; Bs[lc + 1][lr] = bv.y; Bs[lc + 2][lr] = bv.z; Bs[lc + 3][lr] = bv.w;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      float ar[4] = {a.x, a.y, a.z, a.w}, br[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(ar[i], br[j], acc[i][j]);
    }
    __syncthreads();
  }
  float* wz = ws + (size_t)blockIdx.z * M * FC1_N;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int m = m0 + ty * 4 + i;
    if (m < M)
      *reinterpret_cast<float4*>(wz + (size_t)m * FC1_N + n0 + tx * 4) =
          make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
  }
}

// ---------------------------------------------------------------------------------
// Tail of part 1, one CTA (128 threads) per sample row:
//   h = relu(b1 + sum_z ws[z][m][:])      (512)
//   l = b2 + W2 h                          (10)
//   out[m] = softmax(l)                    cifar_model_parts.py:55-57
// ---------------------------------------------------------------------------------
// SPLITS is compile-time so the partial-sum loads are all in flight together (they were a serial chain).
// GROUP: the partials are first added in groups of GROUP, left to right (((p0+p1)+p2)+p3, ...), the association the fc1
// kernels of fc_stage_tc.cuh use when they sum the K slices of an item themselves -- so 32 single-slice partials with
// GROUP = 4 and 8 four-slice partials with GROUP = 1 give the same bits (GROUP = 2: the 3xTF32 kernel's slice pairs).
// (Tried: four consecutive rows per CTA, fc2's 20 KB of weights loaded once per four rows instead of once per row:
//  part 1 went from 30.7 to 33.8 us per 1024 rows -- the 96 registers of that variant and a quarter of the CTAs hide the
//  L2 latency of the partial-sum loads worse than 1024 small CTAs do.)
template <int SPLITS, int GROUP = 1>
__global__ void __launch_bounds__(128)
fc_tail_f32(const float* __restrict__ ws, int M,
            const float* __restrict__ b1,
            const float* __restrict__ w2,     // [10,512]
            const float* __restrict__ b2,
            float* __restrict__ out,          // [M,10]
            RingSpan span) {                  // span.data: fc1's input came from handoff-ring slots: release them
  const int m = blockIdx.x, tid = threadIdx.x;
  pdl_trigger();
  float4 h = __ldg(reinterpret_cast<const float4*>(b1) + tid);      // constant weight: may precede the wait
  pdl_wait();
  if (m >= M) {              // one extra CTA when ring slots have to be released: the fc1 kernel (their only reader)
    const int msgs = (M + span.rows_per_msg - 1) / span.rows_per_msg;      // has completed -- the wait above waited for that
    for (int i = tid; i < msgs; i += blockDim.x) {      // a ring may hold up to 1024 slots: more messages than threads
      const int slot = span_slot(span, i);
      const unsigned long long k = ld_count(span.count + slot) + 1;
      span.count[slot] = k;
      st_relaxed_sys(span.released + slot, k);    // nothing to publish but "my reads are over"
    }
    return;
  }
  float4 p[SPLITS];
#pragma unroll
  for (int z = 0; z < SPLITS; ++z)      // plain loads, NOT __ldg: ws was written by the previous kernel, and the compiler
    p[z] = ld_produced(reinterpret_cast<const float4*>(ws + ((size_t)z * M + m) * FC1_N) + tid);   // may hoist ld.global.nc above pdl_wait()
#pragma unroll
  for (int z = 0; z < SPLITS; z += GROUP) {      // fixed order
    float4 g = p[z];
#pragma unroll
    for (int i = 1; i < GROUP; ++i) { g.x += p[z + i].x; g.y += p[z + i].y; g.z += p[z + i].z; g.w += p[z + i].w; }
    h.x += g.x; h.y += g.y; h.z += g.z; h.w += g.w;
  }
  h.x = fmax_nan(h.x, 0.f); h.y = fmax_nan(h.y, 0.f); h.z = fmax_nan(h.z, 0.f); h.w = fmax_nan(h.w, 0.f);   // nn.ReLU keeps NaN
  __shared__ float part[4][10];
  float l[10];
  float4 w[10];
#pragma unroll
  for (int j = 0; j < 10; ++j) w[j] = __ldg(reinterpret_cast<const float4*>(w2 + j * FC1_N) + tid);
#pragma unroll
  for (int j = 0; j < 10; ++j) {
    float v = fmaf(h.x, w[j].x, fmaf(h.y, w[j].y, fmaf(h.z, w[j].z, h.w * w[j].w)));
    v = warp_sum(v);
    if ((tid & 31) == 0) part[tid >> 5][j] = v;
  }
  __syncthreads();
  if (tid == 0) {
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < 10; ++j) {
      l[j] = b2[j] + ((part[0][j] + part[1][j]) + (part[2][j] + part[3][j]));
      mx = fmaxf(mx, l[j]);
    }
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < 10; ++j) { l[j] = expf(l[j] - mx); sum += l[j]; }
#pragma unroll
    for (int j = 0; j < 10; ++j) out[(size_t)m * 10 + j] = l[j] / sum;
  }
}

}  // namespace b2p
