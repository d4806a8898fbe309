// cifar_tc.cuh -- CIFAR part 0 (ModelPart0_2Node.forward, cifar_model_parts.py:37-42)
// with conv2 as an implicit GEMM on the 5th-gen tensor cores.
//
// One persistent CTA per SM, one sample at a time, three warp roles running concurrently:
//
//   warps 0..7   conv1 producers (CUDA cores, packed fp32 FMA):
//                  conv1(3->32)+bias+ReLU+maxpool for one pooled pixel per thread, exact
//                  fp32; the 32-channel result is written as ONE 128-byte K-major row
//                  [hi fp16 x32 | lo fp16 x32] into three x-shifted, y-padded NHWC copies
//                  of the activation in shared memory (copy kw holds pixel (y, x+kw-1)).
//   warp 8       MMA issuer: conv2 = sum over 9 taps of [128 px x 32 ic] x [32 ic x 64 oc].
//                  Because copy kw is shifted by the tap's x offset and rows are 16 px
//                  wide, the A operand of tap (kh,kw) for output rows y0..y0+7 is simply
//                  the 128 consecutive smem rows starting at row (y0+kh)*16 of copy kw --
//                  a plain 128B-swizzled K-major UMMA descriptor, no im2col is ever built.
//   warps 9..12  epilogue: TMEM -> registers, + bias, ReLU, 2x2 maxpool with two warp
//                  shuffles (a warp holds exactly two image rows), NCHW-flatten order
//                  (index c*64 + h*8 + w) staged in smem, 16 KB coalesced store per sample.
//
// fp32-grade arithmetic on fp16 tensor-core inputs ("split fp16"): every fp32 operand x is
// carried as hi = fp16(x) and lo = fp16((x - hi) * 2^11); three MMAs per K step
// (hi*hi -> acc0, hi*lo and lo*hi -> acc1) and the epilogue forms acc0 + 2^-11 * acc1.
// fp16 products are exact in the fp32 accumulator, so the only loss is the dropped lo*lo
// term and the 2^-22 representation error: same tolerance as the fp32 CUDA-core kernel.
// Compared with 3xTF32 this halves the shared-memory footprint (what makes the weights
// AND a whole sample of activations fit) and doubles the MMA rate.  Range: |x| < 65504
// (inputs are normalised images; conv1 outputs of a sane model are far below that).
#pragma once
#include <cuda_fp16.h>

#include "common.cuh"
#include "tc_gemm.cuh"
#include "ring_hooks.cuh"

namespace b2p {
namespace cifar_tc {

using namespace tc;

// 16 producer warps: thread = (pooled pixel, half of the channel group).  Eight warps (2 per SM sub-partition) left
// the FP32 pipe 49 % busy -- ncu: the FFMA2s stall on shared-memory loads (short scoreboard) and fixed latencies with
// nobody else to issue; twice the warps at half the accumulators per thread hide that.
constexpr int kProducerWarps = 16, kProducerThreads = 512;
constexpr int kMmaWarp = 16;
constexpr int kEpiWarp0 = 17, kEpiThreads = 128;
constexpr int kAcqWarp = kEpiWarp0 + 4, kPubWarp = kAcqWarp + 1;   // handoff-ring handshake, off the epilogue's critical path
constexpr int kThreads = (kPubWarp + 1) * 32;        // 736

constexpr int W2_TAP_BYTES = 128 * 64;               // [64 oc hi rows ; 64 oc lo rows][64 B]
constexpr int W2_BYTES = 9 * W2_TAP_BYTES;           // 73 728
constexpr int ACT_COPY_BYTES = 18 * 16 * 128;        // 36 864: 18 rows (y halo) x 16 px x 128 B
constexpr int ACT_BYTES = 3 * ACT_COPY_BYTES;        // 110 592

// conv1 weights [c*9+kh*3+kw][oc] and bias travel as a kernel parameter (3.5 KB) and are copied to shared memory in
// the prologue.  Reading them straight from the parameter bank in the FMA loop (LDC, warp-uniform addresses) was
// measured slower: 55.9 vs 52.7 us per 1024 samples -- ncu showed the FFMA2s waiting on the constant loads.
struct Conv1Const {
  float w[27][32];
  float b[32];
};

struct Smem {
  unsigned char w2[W2_BYTES];                        // pre-swizzled image from global
  unsigned char act[ACT_BYTES];                      // 3 shifted copies, swizzled rows
  float out[64 * 64];                                // one sample, flatten order
  float in[3][34][34];                               // zero-padded input image
  float b2[64];
  float w1[27][32];                                  // [c*9+kh*3+kw][oc]
  uint64_t bar_w2, bar_act_full[2], bar_act_empty, bar_tmem_full[2][2], bar_tmem_empty[2];
  uint32_t tmem_slot;
  int acq_upto;                                      // ring: samples of this CTA whose slot may be written (acquirer -> epilogue)
  int staged;                                        // ring: samples whose 16 KB sit complete in `out` (epilogue -> publisher)
  int out_free;                                      // ring: samples whose bulk copy has finished READING `out` (publisher -> epilogue)
};
constexpr int SMEM_BYTES = sizeof(Smem) + 1024;

__device__ __forceinline__ void named_bar(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
// 32 lanes x 16 consecutive 32-bit columns
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
// shared -> global bulk copy (TMA store path); the destination may be another GPU's memory
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// packed fp32 FMA (two independent fp32 FMAs per lane per instruction, sm_100)
__device__ __forceinline__ void ffma2(unsigned long long& acc, unsigned long long a, unsigned long long b) {
#ifdef B2P_SCALAR_FMA   // measured: same producer time as FFMA2 (6.0 k vs 6.1 k cycles per pass), slower overall
  float a0, a1, b0, b1, c0, c1;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a0), "=f"(a1) : "l"(a));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(b0), "=f"(b1) : "l"(b));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(c0), "=f"(c1) : "l"(acc));
  c0 = fmaf(a0, b0, c0);
  c1 = fmaf(a1, b1, c1);
  asm("mov.b64 %0, {%1, %2};" : "=l"(acc) : "f"(c0), "f"(c1));
#else
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(a), "l"(b));
#endif
}
__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(unsigned long long v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
// x -> (hi, lo) with x ~= hi + lo * 2^-11
__device__ __forceinline__ void split_f16(float x, __half& hi, __half& lo) {
  hi = __float2half_rn(x);
  lo = __float2half_rn((x - __half2float(hi)) * 2048.0f);
}
__host__ __device__ constexpr uint32_t make_idesc_f16(int m, int n) {   // kind::f16, A/B = F16 (format 0), fp32 accumulate
  return (1u << 4) | (static_cast<uint32_t>(n >> 3) << 17) | (static_cast<uint32_t>(m >> 4) << 24);
}

// K-major operand tile with 64-byte rows in the 64B-swizzle layout: 8-row groups 512 B apart (SBO), 16-byte chunk c
// of row r stored at chunk c ^ ((r >> 1) & 3)  (address bits [4,6) ^= bits [7,9): cute Swizzle<2,4,3>).
__device__ __forceinline__ uint64_t make_desc_sw64(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(1) << 16;
  d |= static_cast<uint64_t>(512 >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(4) << 61;                        // SWIZZLE_64B
  return d;
}

// One-off weight repack (b2p_stage_finalize): conv2.weight [64,32,3,3] fp32 -> the exact byte image of the
// B-operand tiles in shared memory: per tap a [128 rows][64 B] K-major tile, rows 0..63 = fp16 hi part of oc's 32
// input channels, rows 64..127 = the lo part (x ~= hi + lo * 2^-11), 64B-swizzled.  One descriptor over all 128
// rows makes  act_hi x [W_hi ; W_lo]  a single N=128 MMA (hi*hi and hi*lo together): an MMA this small costs
// ~130 cycles for its operand fetch whatever N is, so two instructions per tap and K-step instead of three.
__global__ void pack_conv2_f16x2(const float* __restrict__ w2, unsigned char* __restrict__ img) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;       // one 16-byte chunk each
  if (i >= 9 * 128 * 4) return;
  const int tap = i / 512, r = (i / 4) % 128, c = i % 4;
  const int n = r & 63;
  const int kh = tap / 3, kw = tap % 3;
  __half v[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int ic = 8 * c + e;
    __half hi, lo;
    split_f16(w2[((n * 32 + ic) * 3 + kh) * 3 + kw], hi, lo);
    v[e] = r < 64 ? hi : lo;
  }
  *reinterpret_cast<uint4*>(img + tap * W2_TAP_BYTES + r * 64 + ((c ^ ((r >> 1) & 3)) * 16)) =
      *reinterpret_cast<const uint4*>(v);
}

__global__ void __launch_bounds__(kThreads, 1)
part0_tc(const float* __restrict__ x,                // [B,3,32,32]
         const __grid_constant__ Conv1Const c1,      // conv1.weight (transposed) and conv1.bias
         const unsigned char* __restrict__ w2img,
         const float* __restrict__ b2, float* __restrict__ out,   // [B,4096]
         int batch, long long* __restrict__ dbg,     // dbg: optional [3 roles][8 samples][8 events] clock64 trace of CTA 0
         RingSpan span) {                            // span.data: rows go into handoff-ring slots instead of out (ring_hooks.cuh)
#define B2P_TRACE(role, n, ev, cond) \
  do { if (dbg && blockIdx.x == 0 && (n) < 8 && (cond)) dbg[((role) * 8 + (n)) * 8 + (ev)] = clock64(); } while (0)
  // (offset arithmetic on the __shared__ array itself, not on a uintptr_t: the compiler keeps
  //  the shared address space and emits LDS/STS instead of generic LD/ST -- with generic
  //  accesses it also has to assume every store may alias every load, which serialises them)
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  Smem& s = *reinterpret_cast<Smem*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  // ---- one-time set-up (touches only constant weights: runs before pdl_wait) --------------
  pdl_trigger();
  for (int i = tid; i < ACT_BYTES / 16; i += kThreads) reinterpret_cast<uint4*>(s.act)[i] = make_uint4(0, 0, 0, 0);
  for (int i = tid; i < 3 * 34 * 34; i += kThreads) (&s.in[0][0][0])[i] = 0.f;
  if (tid < 64) s.b2[tid] = b2[tid];
  for (int i = tid; i < 27 * 32; i += kThreads) (&s.w1[0][0])[i] = (&c1.w[0][0])[i];
  if (tid == 0) { s.acq_upto = 0; s.staged = 0; s.out_free = 0; }
  if (warp == kMmaWarp) {
    if (lane == 0) {
      mbar_init(&s.bar_w2, 1);
      mbar_init(&s.bar_act_full[0], kProducerWarps);
      mbar_init(&s.bar_act_full[1], kProducerWarps);
      mbar_init(&s.bar_act_empty, 1);
      for (int i = 0; i < 2; ++i) {
        mbar_init(&s.bar_tmem_full[i][0], 1);
        mbar_init(&s.bar_tmem_full[i][1], 1);
        mbar_init(&s.bar_tmem_empty[i], 4);
      }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(&s.tmem_slot, 512);
    tmem_relinquish();
  }
  fence_proxy_async();               // the zeroed activation halo must be visible to the tensor core
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = s.tmem_slot;
  if (dbg && blockIdx.x == 0 && tid == 0) dbg[3 * 64] = clock64();      // end of set-up
  if (warp != kMmaWarp) pdl_wait();   // producers read the input, the epilogue writes the output: both may belong
                                      // to the previous kernel of the stream; the MMA warp meanwhile streams in W2

  if (warp < kProducerWarps) {
    // =========================== conv1 producers ===========================================
    const int pix = tid & 255, half = tid >> 8;              // pooled pixel; which 8 of the pass's 16 channels
    const int py = pix >> 4, px = pix & 15;
    float4 pre[2];
    if ((int)blockIdx.x < batch) {
      const float4* xin = reinterpret_cast<const float4*>(x + (size_t)blockIdx.x * 3072);
      pre[0] = ld_produced(xin + tid);                      // the image may come from the previous kernel
      if (tid < 256) pre[1] = ld_produced(xin + tid + 512);
    }
    uint32_t n = 0;
    for (int b = blockIdx.x; b < batch; b += gridDim.x, ++n) {
      B2P_TRACE(0, n, 0, tid == 0);
      named_bar(1, kProducerThreads);                       // previous sample's window reads are done
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        if (j == 1 && tid >= 256) break;
        const int e = (tid + 512 * j) * 4;
        const int c = e >> 10, r = (e >> 5) & 31, col = e & 31;
        float* d = &s.in[c][r + 1][col + 1];
        d[0] = pre[j].x; d[1] = pre[j].y; d[2] = pre[j].z; d[3] = pre[j].w;
      }
      named_bar(1, kProducerThreads);
      if (b + (int)gridDim.x < batch) {                     // next image: in flight during the FMAs
        const float4* xin = reinterpret_cast<const float4*>(x + (size_t)(b + gridDim.x) * 3072);
        pre[0] = ld_produced(xin + tid);
        if (tid < 256) pre[1] = ld_produced(xin + tid + 512);
      }
      B2P_TRACE(0, n, 1, tid == 0);
      // ---- conv1 + ReLU + pool, 32 channels in two passes of 16; within a pass this thread owns 8 channels
      //      (4 packed pairs) of its pixel.  Each pass publishes its part of the pixel's row (one hi and one lo
      //      16-byte chunk) into the three shifted copies right away: the tensor core starts on channels 0..15
      //      (the first K=16 MMA of every tap) while the CUDA cores still compute channels 16..31.
#pragma unroll
      for (int pass = 0; pass < 2; ++pass) {
        const int ch0 = pass * 16 + half * 8;                // first channel of this thread's group
        unsigned long long acc[4][4];
#pragma unroll
        for (int p = 0; p < 4; ++p) {
          const float2 bb = *reinterpret_cast<const float2*>(&c1.b[ch0 + 2 * p]);
          const unsigned long long bv = pack2(bb.x, bb.y);
          acc[p][0] = bv; acc[p][1] = bv; acc[p][2] = bv; acc[p][3] = bv;
        }
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          float win[4][4];
#pragma unroll
          for (int r = 0; r < 4; ++r) {
            const float2* rp = reinterpret_cast<const float2*>(&s.in[c][2 * py + r][2 * px]);
            const float2 a = rp[0], bq = rp[1];
            win[r][0] = a.x; win[r][1] = a.y; win[r][2] = bq.x; win[r][3] = bq.y;
          }
#pragma unroll
          for (int kh = 0; kh < 3; ++kh)
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) {
              const unsigned long long x00 = pack2(win[kh][kw], win[kh][kw]);
              const unsigned long long x01 = pack2(win[kh][kw + 1], win[kh][kw + 1]);
              const unsigned long long x10 = pack2(win[kh + 1][kw], win[kh + 1][kw]);
              const unsigned long long x11 = pack2(win[kh + 1][kw + 1], win[kh + 1][kw + 1]);
              const unsigned long long* wrow =
                  reinterpret_cast<const unsigned long long*>(&s.w1[c * 9 + kh * 3 + kw][ch0]);
#pragma unroll
              for (int p = 0; p < 4; ++p) {
                const unsigned long long w = wrow[p];
                ffma2(acc[p][0], x00, w);
                ffma2(acc[p][1], x01, w);
                ffma2(acc[p][2], x10, w);
                ffma2(acc[p][3], x11, w);
              }
            }
        }
        uint32_t hi_pk[4], lo_pk[4];                          // fp16x2 words: channels ch0 + (2i, 2i+1)
#pragma unroll
        for (int p = 0; p < 4; ++p) {
          float a0, a1, b0, b1v, c0, c1, d0, d1;
          unpack2(acc[p][0], a0, a1); unpack2(acc[p][1], b0, b1v);
          unpack2(acc[p][2], c0, c1); unpack2(acc[p][3], d0, d1);
          const float v0 = fmaxf(fmaxf(fmaxf(a0, b0), fmaxf(c0, d0)), 0.f);   // relu(max) == max(relu)
          const float v1 = fmaxf(fmaxf(fmaxf(a1, b1v), fmaxf(c1, d1)), 0.f);
          __half h0, l0, h1, l1;
          split_f16(v0, h0, l0);
          split_f16(v1, h1, l1);
          __half2 hh = __halves2half2(h0, h1), ll = __halves2half2(l0, l1);
          hi_pk[p] = *reinterpret_cast<uint32_t*>(&hh);
          lo_pk[p] = *reinterpret_cast<uint32_t*>(&ll);
        }
        if (pass == 0) {
          B2P_TRACE(0, n, 2, tid == 0);
          mbar_wait(&s.bar_act_empty, (n & 1) ^ 1);           // MMAs of the previous sample are done
          B2P_TRACE(0, n, 3, tid == 0);
        }
        const int ch = 2 * pass + half, cl = 4 + 2 * pass + half;      // hi chunk, lo chunk of this channel group
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) {
          const int xc = px - kw + 1;                          // column of this pixel in copy kw
          if (xc < 0 || xc > 15) continue;
          const int R = (py + 1) * 16 + xc;
          unsigned char* row = s.act + kw * ACT_COPY_BYTES + R * 128;
          *reinterpret_cast<uint4*>(row + ((ch ^ (R & 7)) * 16)) = make_uint4(hi_pk[0], hi_pk[1], hi_pk[2], hi_pk[3]);
          *reinterpret_cast<uint4*>(row + ((cl ^ (R & 7)) * 16)) = make_uint4(lo_pk[0], lo_pk[1], lo_pk[2], lo_pk[3]);
        }
        fence_proxy_async();                                   // generic-proxy writes -> async proxy (UMMA)
        __syncwarp();
        if (lane == 0) mbar_arrive(&s.bar_act_full[pass]);
      }
      B2P_TRACE(0, n, 4, tid == 0);
    }
  } else if (warp == kMmaWarp) {
    // =========================== MMA issuer ====================================================
    if (lane == 0) {
      mbar_expect_tx(&s.bar_w2, W2_BYTES);
      for (int t = 0; t < 9; ++t) bulk_g2s(s.w2 + t * W2_TAP_BYTES, w2img + t * W2_TAP_BYTES, W2_TAP_BYTES, &s.bar_w2);
    }
    __syncwarp();
    mbar_wait(&s.bar_w2, 0);
    constexpr uint32_t idesc = make_idesc_f16(128, 64), idesc128 = make_idesc_f16(128, 128);
    const uint32_t act_addr = smem_u32(s.act), w2_addr = smem_u32(s.w2);
    uint32_t n = 0;
    for (int b = blockIdx.x; b < batch; b += gridDim.x, ++n) {
      const uint32_t set = n & 1, use = n >> 1;
      B2P_TRACE(1, n, 0, lane == 0);
      mbar_wait(&s.bar_tmem_empty[set], (use & 1) ^ 1);
      B2P_TRACE(1, n, 1, lane == 0);
      // channels 0..15 (first K=16 step of every tap) as soon as the first conv1 pass has landed
      mbar_wait(&s.bar_act_full[0], n & 1);
      tc_fence_after();
      B2P_TRACE(1, n, 2, lane == 0);
      if (lane == 0) {
        // Consecutive MMAs go to DIFFERENT accumulators (tile 0 / tile 1 alternate): a 128x64x16 MMA is ~32 cycles of
        // tensor work, far less than the pipeline latency of a dependent accumulate into the same TMEM tile.
        const uint32_t a0 = tmem_base + set * 256, a1 = a0 + 128;            // tile 0 / tile 1: acc0 at +0, acc1 at +64
#pragma unroll 1
        for (int tap = 0; tap < 9; ++tap) {
          const int kh = tap / 3, kw = tap % 3;
          const uint64_t da0 = make_desc_sw128(act_addr + kw * ACT_COPY_BYTES + kh * 16 * 128);
          const uint64_t da1 = make_desc_sw128(act_addr + kw * ACT_COPY_BYTES + (8 + kh) * 16 * 128);
          const uint64_t db = make_desc_sw64(w2_addr + tap * W2_TAP_BYTES);
          // descriptor units are 16 B: +2 = the second K=16 step (channels 16..31); act rows: +4 = the lo half
          umma<0>(a0, da0, db, idesc128, tap != 0);               // hi * [W_hi ; W_lo] -> acc0 | acc1
          umma<0>(a1, da1, db, idesc128, tap != 0);
          umma<0>(a0 + 64, da0 + 4, db, idesc, 1u);               // lo * W_hi -> acc1
          umma<0>(a1 + 64, da1 + 4, db, idesc, 1u);
        }
      }
      __syncwarp();
      // channels 16..31, tile by tile, so the epilogue of tile 0 overlaps the MMAs of tile 1
      mbar_wait(&s.bar_act_full[1], n & 1);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t a0 = tmem_base + set * 256, a1 = a0 + 128;
#pragma unroll 1
        for (int tap = 0; tap < 9; ++tap) {
          const int kh = tap / 3, kw = tap % 3;
          const uint64_t da0 = make_desc_sw128(act_addr + kw * ACT_COPY_BYTES + kh * 16 * 128);
          const uint64_t da1 = make_desc_sw128(act_addr + kw * ACT_COPY_BYTES + (8 + kh) * 16 * 128);
          const uint64_t db = make_desc_sw64(w2_addr + tap * W2_TAP_BYTES) + 2;
          umma<0>(a0, da0 + 2, db, idesc128, 1u);
          umma<0>(a1, da1 + 2, db, idesc128, 1u);
          umma<0>(a0 + 64, da0 + 6, db, idesc, 1u);
          umma<0>(a1 + 64, da1 + 6, db, idesc, 1u);
        }
        umma_commit(&s.bar_tmem_full[set][0]);
        umma_commit(&s.bar_tmem_full[set][1]);
        umma_commit(&s.bar_act_empty);                       // activation copies may be overwritten
      }
      B2P_TRACE(1, n, 3, lane == 0);
      __syncwarp();
    }
  } else if (warp < kAcqWarp) {
    // =========================== epilogue =======================================================
    const int q = warp & 3;                                  // TMEM lane quarter: image rows 2q, 2q+1 of the tile
    const int et = tid - kEpiWarp0 * 32;                     // 0..127
    int acquired = -1;                                       // last message whose slot this CTA has waited for
    uint32_t n = 0;
    for (int b = blockIdx.x; b < batch; b += gridDim.x, ++n) {
      const uint32_t set = n & 1, use = n >> 1;
      B2P_TRACE(2, n, 0, et == 0);
      float* dst_row = out + (size_t)b * 4096;
      if (span.data) {           // row b travels in message b / rows_per_msg, in that message's ring slot
        const int msg = b / span.rows_per_msg;
        dst_row = reinterpret_cast<float*>(span.data + (size_t)span_slot(span, msg) * span.slot_bytes) +
                  (size_t)(b - msg * span.rows_per_msg) * 4096;
      }
      const bool xo = (lane & 1) != 0, yo = (lane & 16) != 0;
#pragma unroll 1
      for (int t = 0; t < 2; ++t) {
        mbar_wait(&s.bar_tmem_full[set][t], use & 1);
        tc_fence_after();
        if (t == 0) {
          B2P_TRACE(2, n, 1, et == 0);
          if (span.data) {                                  // the previous sample's bulk copy must be done reading `out`
            while (ld_acquire_cta(&s.out_free) < (int)n) { }
          }
        }
#pragma unroll 1
        for (int ch = 0; ch < 4; ++ch) {                     // 16 output channels at a time (register budget: 88 per thread)
          uint32_t r0[16], r1[16];
          const uint32_t ta = tmem_base + set * 256 + t * 128 + ch * 16 + (static_cast<uint32_t>(q * 32) << 16);
          tmem_ld16(ta, r0);
          tmem_ld16(ta + 64, r1);
          tmem_ld_wait();
          // 2x2 max-pool across lanes.  A lane owns one conv2 pixel (x = lane&15, row = lane>>4);
          // pool partners are lane^1 (x) and lane^16 (y).  Instead of two shuffles per channel
          // with 3/4 of the lanes ending up redundant, partners SWAP channels: after the x step
          // the even-x lane keeps channel 2p and the odd-x lane channel 2p+1, after the y step
          // likewise -- 12 shuffles per 16 channels and every lane ends with 4 distinct results.
          float u[8];
#pragma unroll
          for (int p = 0; p < 8; ++p) {
            const float a = fmaf(__uint_as_float(r1[2 * p]), 1.0f / 2048.0f, __uint_as_float(r0[2 * p]));
            const float b = fmaf(__uint_as_float(r1[2 * p + 1]), 1.0f / 2048.0f, __uint_as_float(r0[2 * p + 1]));
            const float recv = __shfl_xor_sync(0xffffffffu, xo ? a : b, 1);
            u[p] = fmaxf(xo ? b : a, recv);                  // channel 2p + xo, pooled in x
          }
          const int pooled = (t * 4 + q) * 8 + ((lane & 15) >> 1);    // h*8 + w of the pooled pixel
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const float recv = __shfl_xor_sync(0xffffffffu, yo ? u[2 * g] : u[2 * g + 1], 16);
            const float m = fmaxf(yo ? u[2 * g + 1] : u[2 * g], recv);
            const int c = ch * 16 + 4 * g + (yo ? 2 : 0) + (xo ? 1 : 0);
            s.out[c * 64 + pooled] = fmaxf(m + s.b2[c], 0.f);         // max(a)+b == max(a+b)
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&s.bar_tmem_empty[set]);
      B2P_TRACE(2, n, 2, et == 0);
      if (span.data) {
        // Ring mode: the 16 KB leave through ONE bulk copy issued by the publisher warp.  With per-thread peer stores
        // the four epilogue warps sat in the LSU queue while all 148 CTAs pushed their samples over NVLink at the same
        // moment (2.4 MB bursts): +19 us on a 53 us step.
        fence_proxy_async();                                  // this thread's writes to `out` -> visible to the copy engine
        named_bar(2, kEpiThreads);                            // the sample's 4096 outputs are staged
        if (et == 0) st_release_cta(&s.staged, (int)n + 1);
      } else {
        named_bar(2, kEpiThreads);                            // the sample's 4096 outputs are staged
        float4* dst = reinterpret_cast<float4*>(dst_row);
        const float4* src = reinterpret_cast<const float4*>(s.out);
#pragma unroll
        for (int i = 0; i < 8; ++i) dst[et + 128 * i] = src[et + 128 * i];
        named_bar(2, kEpiThreads);                            // staging buffer free for the next sample
      }
      B2P_TRACE(2, n, 3, et == 0);
    }
  } else if (span.data && lane == 0) {
    // =========================== handoff-ring handshake (ring_hooks.cuh) ========================
    // Two single-thread roles kept OFF the epilogue warps: with one sample per CTA per message, every sample
    // needs a poll of the consumer's release counter (an NVLink round trip) and an arrival ticket (fence + L2
    // atomic); done by the epilogue they added ~40 us to a 64 us step.
    uint32_t n = 0;
    if (warp == kAcqWarp) {
      // acquirer: runs ahead of the pipeline.  The consumer's release counter lives on the consumer's GPU, and
      // sys-scope reads of one remote line serialise: with every CTA polling for every message (1036 polls per
      // step) the acquirers fell behind the epilogue (+19 us per 53 us step).  So ONE CTA per message -- the one
      // that holds the message's first row -- polls over NVLink and posts what it saw in local memory; the others
      // wait on that local word.  (Acyclic: the poller of message m only ever waits for messages < m.)
      int acquired = -1;
      for (int b = blockIdx.x; b < batch; b += gridDim.x, ++n) {
        const int msg = b / span.rows_per_msg, slot = span_slot(span, msg);
        if (msg != acquired) {
          const unsigned long long k = ld_count(span.count + slot);      // messages already published into this slot
          if (k > 0) {
            if (b == msg * span.rows_per_msg) {
              spin_until_ge_remote(span.released + slot, k);
              st_release_gpu(span.acq_seen + slot, k);
            } else {
              while (ld_acquire_gpu(span.acq_seen + slot) < k) __nanosleep(64);
            }
          }
          acquired = msg;
        }
        st_release_cta(&s.acq_upto, (int)n + 1);
      }
    } else {
      // publisher: when sample n is staged and its slot acquired, one bulk copy (shared -> the consumer GPU's slot);
      // as soon as the copy engine has READ the staging buffer the epilogue may refill it; when the copy has
      // completed, fence at gpu scope and count the row in.  Whoever counts a message's last row has observed every
      // other arrival; its release store of the flag, the only system-scope fence on this path, carries all of them.
      // (Awaiting copy n-1 only after issuing copy n was measured slower: 68.5 vs 62.9 us per step.)
      for (int b = blockIdx.x; b < batch; b += gridDim.x, ++n) {
        const int msg = b / span.rows_per_msg, slot = span_slot(span, msg);
        float* dst_row = reinterpret_cast<float*>(span.data + (size_t)slot * span.slot_bytes) +
                         (size_t)(b - msg * span.rows_per_msg) * 4096;
        while (ld_acquire_cta(&s.staged) <= (int)n) __nanosleep(32);
        while (ld_acquire_cta(&s.acq_upto) <= (int)n) { }     // the slot must have been released by the consumer
        bulk_s2g(dst_row, s.out, 64 * 64 * 4);
        bulk_wait_read();
        st_release_cta(&s.out_free, (int)n + 1);
        bulk_wait_all();
        __threadfence();
        const int rows = min(span.rows_per_msg, batch - msg * span.rows_per_msg);
        if (atomicAdd(span.tickets + slot, 1) == rows - 1) {
          span.tickets[slot] = 0;                             // ready for the slot's next message
          __threadfence();
          const unsigned long long k = ld_count(span.count + slot) + 1;
          span.count[slot] = k;
          st_release_sys(span.published + slot, k);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) tmem_dealloc(tmem_base, 512);
#undef B2P_TRACE
}

}  // namespace cifar_tc
}  // namespace b2p
