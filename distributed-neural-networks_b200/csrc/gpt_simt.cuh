// gpt_simt.cuh -- CUDA-core fp32 kernels for the GPT parts (B2P_MATH_FP32) plus the
// bandwidth-bound kernels every math mode shares (embedding, LayerNorm, arg-max).
//
// Arithmetic follows oracle/gpt_oracle.py::_block (the third-party ``model.Block``
// the reference imports at partitions/gpt_model_parts.py:4; semantics pinned in
// oracle/gpt_model.py).
#pragma once
#include <cuda_bf16.h>

#include "common.cuh"

namespace b2p {

// ---------------------------------------------------------------------------------
// Embedding: out[m,:] = wte[idx[m],:] + wpe[m % T,:]       gpt_model_parts.py:16-19
// One warp per token, 128-bit accesses.  Out-of-range ids (a Python IndexError in
// the reference) set *bad_flag and are clamped so the kernel never faults.
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
embed_f32(const int64_t* __restrict__ idx, const float* __restrict__ wte,
          const float* __restrict__ wpe, float* __restrict__ out,
          int M, int T, int C, int V, int* __restrict__ bad_flag, int pos0,        // pos0: position of column 0 (KV-cache decoding)
          const int* pos_dev, int pos_limit) {      // pos_dev != null: the position comes from device memory (clamped below pos_limit)
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= M) return;
  if (pos_dev) pos0 = min(max(*pos_dev, 0), pos_limit - T);
  long long tok = idx[warp];
  if (tok < 0 || tok >= V) {
    if (lane == 0) atomicExch(bad_flag, 1);
    tok = tok < 0 ? 0 : V - 1;
  }
  const float4* a = reinterpret_cast<const float4*>(wte + (size_t)tok * C);
  const float4* p = reinterpret_cast<const float4*>(wpe + (size_t)(pos0 + warp % T) * C);
  float4* o = reinterpret_cast<float4*>(out + (size_t)warp * C);
  for (int i = lane; i < C / 4; i += 32) {
    float4 x = __ldg(a + i), y = __ldg(p + i);
    o[i] = make_float4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w);
  }
}

// ---------------------------------------------------------------------------------
// LayerNorm over the last dim (eps 1e-5, affine), one warp per row, row held in
// registers, two-pass (mean, then centred variance) like ATen's CPU kernel.
// C must be a multiple of 128 and <= 2048.
// ---------------------------------------------------------------------------------
template <typename OutT>
__device__ __forceinline__ void store4(OutT* p, float a, float b, float c, float d);
template <>
__device__ __forceinline__ void store4<float>(float* p, float a, float b, float c, float d) {
  *reinterpret_cast<float4*>(p) = make_float4(a, b, c, d);
}
template <>
__device__ __forceinline__ void store4<__nv_bfloat16>(__nv_bfloat16* p, float a, float b, float c, float d) {
  __nv_bfloat162 lo = __floats2bfloat162_rn(a, b), hi = __floats2bfloat162_rn(c, d);
  uint2 v;
  v.x = *reinterpret_cast<uint32_t*>(&lo);
  v.y = *reinterpret_cast<uint32_t*>(&hi);
  *reinterpret_cast<uint2*>(p) = v;
}
__device__ __forceinline__ void store1(float* p, float v) { *p = v; }
__device__ __forceinline__ void store1(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

// fp32 -> bf16 copy (weights at finalize time)
__global__ void __launch_bounds__(256)
f32_to_bf16(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) out[i] = __float2bfloat16_rn(in[i]);
}

// x = hi + lo with hi, lo exactly representable in tf32 (10-bit mantissa): the operand
// planes of the 3xTF32 contraction.  |x - hi - lo| <= 2^-22 |x|.
__device__ __forceinline__ float round_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
__global__ void __launch_bounds__(256)
split_tf32(const float* __restrict__ in, float* __restrict__ hi, float* __restrict__ lo, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  pdl_trigger();
  pdl_wait();
  for (; i < n; i += stride) {
    const float x = ld_produced(in + i), h = round_tf32(x);
    hi[i] = h;
    lo[i] = round_tf32(x - h);
  }
}

template <typename OutT, int NVEC>          // NVEC = C / 128 float4 per lane (compile-time: row lives in NVEC*4 registers)
__global__ void __launch_bounds__(256)
layernorm_rows(const float* __restrict__ x, const float* __restrict__ gamma,
               const float* __restrict__ beta, OutT* __restrict__ y, int M, int C) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  pdl_trigger();
  pdl_wait();
  if (row >= M) return;
  const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * C);
  float4 v[NVEC];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < NVEC; ++i) {
    v[i] = ld_produced(xr + lane + 32 * i);      // x comes from the previous kernel (see pdl_wait)
    sum += (v[i].x + v[i].y) + (v[i].z + v[i].w);
  }
  const float mean = warp_sum(sum) / (float)C;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < NVEC; ++i) {
    float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
    sq += (a * a + b * b) + (c * c + d * d);
  }
  const float rstd = rsqrtf(warp_sum(sq) / (float)C + 1e-5f);
#pragma unroll
  for (int i = 0; i < NVEC; ++i) {
    float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + lane + 32 * i);
    float4 b = beta ? __ldg(reinterpret_cast<const float4*>(beta) + lane + 32 * i) : make_float4(0, 0, 0, 0);
    store4<OutT>(y + (size_t)row * C + (lane + 32 * i) * 4,
                 (v[i].x - mean) * rstd * g.x + b.x, (v[i].y - mean) * rstd * g.y + b.y,
                 (v[i].z - mean) * rstd * g.z + b.z, (v[i].w - mean) * rstd * g.w + b.w);
  }
}

// ---------------------------------------------------------------------------------
// Generic fp32 "linear":  out[m,n] = epi( sum_k A[m,k] * W[n,k] + bias[n] ) (+ res[m,n])
// A [M,K] row-major, W [N,K] row-major (nn.Linear layout).  K % 16 == 0.
// 64x64 tile, 256 threads, 4x4 per thread; M and N edges guarded.
// ---------------------------------------------------------------------------------
enum { EPI_NONE = 0, EPI_RELU = 1, EPI_GELU = 2 };

template <int EPI>
__global__ void __launch_bounds__(256)
linear_f32(const float* __restrict__ A, const float* __restrict__ W,
           const float* __restrict__ bias, const float* __restrict__ res,
           float* __restrict__ out, int M, int N, int K) {
  __shared__ float As[16][64 + 4];
  __shared__ float Bs[16][64 + 4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  const int lr = tid >> 2, lc = (tid & 3) * 4;
  const int ty = tid >> 4, tx = tid & 15;
  float acc[4][4] = {};
  const bool a_ok = (m0 + lr) < M, b_ok = (n0 + lr) < N;
  const float* ap = A + (size_t)(a_ok ? m0 + lr : 0) * K + lc;
  const float* bp = W + (size_t)(b_ok ? n0 + lr : 0) * K + lc;
  for (int k0 = 0; k0 < K; k0 += 16) {
    float4 av = a_ok ? __ldg(reinterpret_cast<const float4*>(ap + k0)) : make_float4(0, 0, 0, 0);
    float4 bv = b_ok ? __ldg(reinterpret_cast<const float4*>(bp + k0)) : make_float4(0, 0, 0, 0);
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
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= N) continue;
      float v = acc[i][j] + (bias ? __ldg(bias + n) : 0.f);
      if (EPI == EPI_RELU) v = fmaxf(v, 0.f);
      if (EPI == EPI_GELU) v = gelu_erf(v);
      if (res) v += res[(size_t)m * N + n];
      out[(size_t)m * N + n] = v;
    }
  }
}

// ---------------------------------------------------------------------------------
// Causal self-attention, head dim 64, fp32, flash-style (online softmax):
//   y[b,t,h*64:(h+1)*64] = softmax_j<=t( q_t . k_j / 8 ) v_j
// qkv is the c_attn output [B*T, 3C] (q | k | v, each C wide, head h at h*64).
// One CTA (256 threads) per (64-query tile, head, batch); key tiles of 64 up to the
// diagonal.  Thread (ty,tx): rows 4ty..4ty+3, columns tx+16j.
// ---------------------------------------------------------------------------------
struct AttnSmem {
  float V[64][64];         // [key][d]   (first: float4 stores need 16 B alignment)
  float Q[64][64 + 1];     // [query][d]
  float K[64][64 + 1];     // [key][d]
  float P[64][64 + 1];     // [query][key]
};

template <typename OutT>
__global__ void __launch_bounds__(256)
attention_causal_f32(const float* __restrict__ qkv, OutT* __restrict__ y, int T, int C) {
  extern __shared__ __align__(16) unsigned char attn_smem_raw[];
  AttnSmem& sm = *reinterpret_cast<AttnSmem*>(attn_smem_raw);
  float (*Qs)[65] = sm.Q;
  float (*Ks)[65] = sm.K;
  float (*Vs)[64] = sm.V;
  float (*Ps)[65] = sm.P;
  const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
  const int q0 = blockIdx.x * 64, h = blockIdx.y, b = blockIdx.z;
  const size_t row_stride = (size_t)3 * C;
  const float* base = qkv + (size_t)b * T * row_stride + h * 64;

  for (int i = tid; i < 64 * 16; i += 256) {                    // Q tile, float4 loads
    int r = i >> 4, c4 = (i & 15) * 4;
    float4 v = (q0 + r < T) ? __ldg(reinterpret_cast<const float4*>(base + (size_t)(q0 + r) * row_stride + c4))
                            : make_float4(0, 0, 0, 0);
    Qs[r][c4] = v.x; Qs[r][c4 + 1] = v.y; Qs[r][c4 + 2] = v.z; Qs[r][c4 + 3] = v.w;
  }
  float o[4][4] = {};
  float mrow[4], lrow[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) { mrow[i] = -INFINITY; lrow[i] = 0.f; }

  for (int k0 = 0; k0 <= q0 && k0 < T; k0 += 64) {
    __syncthreads();                                            // previous tile fully consumed
    for (int i = tid; i < 64 * 16; i += 256) {
      int r = i >> 4, c4 = (i & 15) * 4;
      bool ok = (k0 + r) < T;
      const float* rp = base + (size_t)(k0 + r) * row_stride + c4;
      float4 kv = ok ? __ldg(reinterpret_cast<const float4*>(rp + C)) : make_float4(0, 0, 0, 0);
      float4 vv = ok ? __ldg(reinterpret_cast<const float4*>(rp + 2 * C)) : make_float4(0, 0, 0, 0);
      Ks[r][c4] = kv.x; Ks[r][c4 + 1] = kv.y; Ks[r][c4 + 2] = kv.z; Ks[r][c4 + 3] = kv.w;
      *reinterpret_cast<float4*>(&Vs[r][c4]) = vv;
    }
    __syncthreads();
    // S = Q K^T / 8
    float sacc[4][4] = {};
#pragma unroll 8
    for (int d = 0; d < 64; ++d) {
      float qv[4], kv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) qv[i] = Qs[ty * 4 + i][d];
#pragma unroll
      for (int j = 0; j < 4; ++j) kv[j] = Ks[tx + 16 * j][d];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) sacc[i][j] = fmaf(qv[i], kv[j], sacc[i][j]);
    }
    // mask + online softmax (row statistics shared by the 16 lanes with equal ty)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int qi = q0 + ty * 4 + i;
      float mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int kj = k0 + tx + 16 * j;
        sacc[i][j] = (kj <= qi && kj < T) ? sacc[i][j] * 0.125f : -INFINITY;
        mx = fmaxf(mx, sacc[i][j]);
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
      const float mnew = fmaxf(mrow[i], mx);
      // rows beyond T (padding) keep mnew = -inf: guard the subtraction
      const float corr = (mnew == -INFINITY) ? 1.f : expf(mrow[i] - mnew);
      float ps = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float p = (mnew == -INFINITY) ? 0.f : expf(sacc[i][j] - mnew);
        Ps[ty * 4 + i][tx + 16 * j] = p;
        ps += p;
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) ps += __shfl_xor_sync(0xffffffffu, ps, off);
      lrow[i] = lrow[i] * corr + ps;
      mrow[i] = mnew;
#pragma unroll
      for (int j = 0; j < 4; ++j) o[i][j] *= corr;
    }
    __syncthreads();
    // O += P V
#pragma unroll 8
    for (int k = 0; k < 64; ++k) {
      float pv[4], vv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) pv[i] = Ps[ty * 4 + i][k];
#pragma unroll
      for (int j = 0; j < 4; ++j) vv[j] = Vs[k][tx + 16 * j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) o[i][j] = fmaf(pv[i], vv[j], o[i][j]);
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int qi = q0 + ty * 4 + i;
    if (qi >= T) continue;
    OutT* yr = y + ((size_t)b * T + qi) * C + h * 64;
#pragma unroll
    for (int j = 0; j < 4; ++j) store1(yr + tx + 16 * j, o[i][j] / lrow[i]);
  }
}

// ---------------------------------------------------------------------------------
// Row arg-max over logits [M,V] -> int64 ids.  First maximal index wins, like
// torch.argmax / np.argmax.  One CTA per row.
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
argmax_rows(const float* __restrict__ logits, int64_t* __restrict__ ids, int V, int64_t ld) {
  const int row = blockIdx.x, tid = threadIdx.x;
  const float* p = logits + (size_t)row * ld;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int i = tid; i < V; i += 256) {
    float v = __ldg(p + i);
    if (v > best || (v == best && i < bi)) { best = v; bi = i; }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    float ov = __shfl_xor_sync(0xffffffffu, best, off);
    int oi = __shfl_xor_sync(0xffffffffu, bi, off);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  __shared__ float sv[8];
  __shared__ int si[8];
  if ((tid & 31) == 0) { sv[tid >> 5] = best; si[tid >> 5] = bi; }
  __syncthreads();
  if (tid == 0) {
    for (int w = 1; w < 8; ++w)
      if (sv[w] > best || (sv[w] == best && si[w] < bi)) { best = sv[w]; bi = si[w]; }
    ids[row] = bi;
  }
}

// Final reduction of the LM-head epilogue partials: [M][slots] (max, first arg-max) -> int64 ids.
// Slots are in ascending column order, so on equal values the lower slot / lower index wins,
// which keeps torch.argmax / np.argmax's first-maximum convention.  One warp per row.
__global__ void __launch_bounds__(256)
argmax_partials(const float* __restrict__ val, const int* __restrict__ idx, int64_t* __restrict__ ids, int M,
                int slots) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  pdl_trigger();
  pdl_wait();
  if (row >= M) return;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int s = lane; s < slots; s += 32) {
    const float v = ld_produced(val + (size_t)row * slots + s);
    const int i = ld_produced(idx + (size_t)row * slots + s);
    if (v > best || (v == best && i < bi)) { best = v; bi = i; }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, off);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, off);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  if (lane == 0) ids[row] = bi;
}

}  // namespace b2p
