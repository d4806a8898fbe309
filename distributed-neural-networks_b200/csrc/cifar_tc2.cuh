// cifar_tc2.cuh -- CIFAR part 0 (ModelPart0_2Node.forward, cifar_model_parts.py:37-42) with BOTH convolutions on
// the 5th-gen tensor cores, weights stationary, pixels on the N dimension.
//
// Measured first (scripts/micro/mma_probe.cu, profiles/r2_mma_probe.log): one tcgen05.mma kind::f16 K=16 costs
//   A from shared memory:  ~81 cycles for N <= 128, 130 for N = 256   (M = 64 or 128 alike)
//   A from tensor memory:  128 cycles for N <= 128, 138 for N = 256
// so only N = 256 runs the tensor pipe near its rate, and round 1's "pixels on M, 64 output channels on N" layout
// (72 narrow MMAs per sample) could never do better than ~40 %.  Here the roles are swapped:
//
//   conv2  D[128 x 256] += A[128 x 16] * B[256 x 16]^T  per 3x3 tap and 16 input channels, where
//          A = rows 2c, 2c+1 = fp16 hi part, lo part of output channel c (64 channels -> M = 128), in TENSOR MEMORY
//              for taps 0..7 (128 columns; tap 8 is read from shared memory: 144 columns would not leave room for
//              the accumulators), and
//          B = 128 pixels (8 image rows) of the conv1 activation, hi rows and lo rows interleaved in groups of 8,
//              so ONE instruction yields  W_hi*a_hi | W_hi*a_lo  (even lanes)  and  W_lo*a_hi | (W_lo*a_lo, unused)
//              (odd lanes): the three products of the split-fp16 scheme in 36 wide MMAs per sample, and the cross
//              term sits in the NEIGHBOURING lane of the same warp -- one shuffle away in the epilogue.
//          The B operand of tap (kh,kw) is a plain shifted window of the activation (three x-shifted copies with a
//          zero halo, as in round 1): no im2col.  The accumulator comes out [channel][pixel], i.e. already in the
//          reference's flatten order c*64+h*8+w, and the 2x2 pool is between a thread's own columns.
//   conv1  as a "pool-window" GEMM: one ROW per POOLED pixel (M = the 128 pooled pixels of half a sample), one column per
//          (channel, dy, dx) of its 2x2 window (32 x 4 = 128 columns), K = 3 channels x 4x4 input window = 48 (weights
//          zero-padded 3x3 -> 4x4).  Three split-fp16 products into ONE accumulator: lo terms first, then the first
//          hi*hi MMA carries scale-input-d = 2^-11 (D = A*B + D*2^-11).  18 MMAs (N = 128) per sample instead of 6.9 k
//          FP32-pipe cycles.  In the epilogue a thread owns a pixel and eight channels; the four window positions of a
//          channel are its own neighbouring columns, so pool, bias, ReLU, the fp16 hi/lo split and the 16-byte rows of
//          the activation copies need no shuffle (with the channels on M it took 48 per thread).
//
// Warp roles (20 warps, one persistent CTA per SM, samples blockIdx.x, +gridDim.x, ...):
//   0..1     conv1 patch producers (input image -> 4x4 windows, fp16 hi/lo); they start on the first image while the
//            other warps still set the kernel up
//   2        publisher: four lanes send the finished samples in turn, one 16 KB bulk copy each, to HBM or straight into
//            the next GPU's ring slot (+ the ring handshake: all slots of the step acquired before the first copy)
//   3        MMA issuer: all 32 lanes in uniform control flow, tensor-core instructions predicated on elect.sync
//   4..19    epilogue warps: drain conv2's accumulator (-> staging tile) and run conv1's epilogue, alternately
// Set-up: conv2's tensor-memory image arrives by ONE bulk copy into the not-yet-used activation buffer and goes to
// tensor memory from there.  Ring mode with bit 2 of the stage's PDL mask: no griddepcontrol.wait, consecutive steps
// overlap (b200pipe.h).
//
// Arithmetic and range are those of round 1's kernel: x ~= hi + lo * 2^-11 with fp16 hi/lo, products exact in the
// fp32 accumulator, the lo*lo term dropped.  |input| and |conv1 output| must stay below 65504 (fp16 range); beyond
// it the result is non-finite where the reference's would be finite.  The kernel guards the INPUT (a per-stage device
// flag, reported by b2p_stage_check as B2P_ELIMIT): every input is compared with the magnitude from which on a conv1
// output could reach the limit, |x| >= (65504 - max|b1|) / max_oc sum|w1[oc]| -- conservative, never misses an overflow,
// and costs one max per input pixel (checking the conv1 outputs themselves cost 1.5 us per 1024 samples: the conv1
// epilogue is on the kernel's critical chain).  Use math="fp32" for such data (b200pipe.h).
// NaN inputs propagate like torch's ReLU / MaxPool2d (max.NaN), not like fmaxf; an infinite input comes out as NaN
// where the reference keeps +-inf (inf * 0 in the zero-padded pool-window weights, inf - inf in the split).
#pragma once
#include "cifar_tc.cuh"

namespace b2p {
namespace cifar_tc2 {

using namespace tc;
using cifar_tc::bulk_g2s;
using cifar_tc::make_desc_sw64;
using cifar_tc::make_idesc_f16;
using cifar_tc::named_bar;

constexpr int kPatchWarp0 = 0, kPatchThreads = 64;
constexpr int kPubWarp = 2, kMmaWarp = 3;
constexpr int kPubLanes = 4;                          // lanes of the publisher warp that send samples (power of two)
constexpr int kE1Warp0 = 4, kEpiWarps = 16;           // epilogue warps 4..19
constexpr int kThreads = 20 * 32;                     // 640 threads: 96 registers each (at 736 threads the 80-register limit decided everything;
                                                      // a 21st warp is accounted as four: 80 registers again)

// tensor-memory map (512 columns)
constexpr uint32_t TM_A2 = 0;                         // conv2 weights, taps 0..7: 8 taps x 2 K-steps x 8 columns
constexpr uint32_t TM_D2 = 128;                       // conv2 accumulator of one half sample: 256 columns
constexpr uint32_t TM_D1 = 384;                       // conv1 accumulator of one half sample: 128 columns

constexpr int ACT_ROW_BYTES = 2048;                   // one image row of one copy: 2 groups x (8 hi rows + 8 lo rows) x 64 B
constexpr int ACT_COPY_BYTES = 17 * ACT_ROW_BYTES;    // rows -1..15; the zero row 16 of a copy is row -1 of the next one
constexpr int ACT_BYTES = (3 * 17 + 1) * ACT_ROW_BYTES;   // 106 496
constexpr int PLANE_BYTES = 128 * 128;                // a 128-row operand tile with 128-byte rows
constexpr int W2T_WORDS = 128;                        // tensor-memory image of conv2's weights: 32-bit words per row
constexpr int W2T_BYTES = 128 * W2T_WORDS * 4;        // 65 536

struct Smem {
  unsigned char act[ACT_BYTES];                       // conv2's B operand: 3 x-shifted copies, hi/lo interleaved by 8 pixels
  unsigned char patch[2 * PLANE_BYTES];               // conv1's B operand: [hi | lo][128 pooled pixels][128 B], 96 B used
  unsigned char a1[2 * PLANE_BYTES];                  // conv1's A operand: [hi | lo][128 rows (oc,dy,dx)][128 B]
  unsigned char w2tap8[cifar_tc::W2_TAP_BYTES];       // conv2 tap 8: 128 rows (2c | 2c+1 = hi | lo of channel c) of 64 B, 64B-swizzled
  float out[2][64 * 64];                              // finished samples in flatten order [channel][h*8+w], double-buffered
  uint32_t in[3][34][34];                             // zero-padded input image, every pixel already split: fp16 hi | fp16 lo << 16
  float b1[32], b2[64];
  uint64_t bar_w, bar_w2t, bar_patch_full, bar_patch_empty, bar_d1_full, bar_d1_empty, bar_act_ready[2], bar_act_free,
      bar_d2_full, bar_d2_empty;
  uint32_t tmem_slot;
  int staged, out_free[2];                            // staging hand-offs between the epilogue warps and the publisher (per buffer)
};
constexpr int SMEM_BYTES = sizeof(Smem) + 1024;
static_assert(SMEM_BYTES <= 232448, "shared memory budget of one CTA on sm_100");

// ---- PTX wrappers this kernel adds -------------------------------------------------------------------------------
__device__ __forceinline__ void umma_ts_f16(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// D = A*B + D * 2^-11   (scale-input-d; checked on the device by scripts/micro/mma_probe.cu)
__device__ __forceinline__ void umma_ss_f16_scale11(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc) {
  const uint32_t z = 0;
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, {%4, %4, %4, %4}, p, 11;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(z)
      : "memory");
}
// tensor-core instructions issued from warp-uniform code (tc_gemm.cuh: elect_one, umma_p, umma_commit_p), plus the
// two forms only this kernel uses: A from tensor memory, and scale-input-d
__device__ __forceinline__ void umma_ts_p(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate, uint32_t lead) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\tsetp.ne.b32 q, %5, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(lead)
      : "memory");
}
// D = A*B + D * 2^-11   (scale-input-d; checked on the device by scripts/micro/mma_probe.cu)
__device__ __forceinline__ void umma_ss_scale11_p(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t lead) {
  const uint32_t z = 0;
  asm volatile(
      "{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, 1, 0;\n\tsetp.ne.b32 q, %5, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, {%4, %4, %4, %4}, p, 11;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(z), "r"(lead)
      : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
        "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
        "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
}
// torch's ReLU and MaxPool2d propagate NaN; fmaxf would drop it
__device__ __forceinline__ float max_nan(float a, float b) {
  float r;
  asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}
// x -> fp16 hi in the low half, fp16 lo = (x - hi) * 2^11 in the high half
__device__ __forceinline__ uint32_t split_pack(float x) {
  const __half hi = __float2half_rn(x);
  const __half lo = __float2half_rn((x - __half2float(hi)) * 2048.0f);
  return (uint32_t)__half_as_ushort(hi) | ((uint32_t)__half_as_ushort(lo) << 16);
}
// two values at once: (a, b) -> fp16x2 hi parts and fp16x2 lo parts (packed conversions: 4 instructions per value)
__device__ __forceinline__ void split2(float a, float b, uint32_t& hi, uint32_t& lo) {
  const __half2 h = __floats2half2_rn(a, b);
  const float2 hf = __half22float2(h);
  const __half2 l = __floats2half2_rn((a - hf.x) * 2048.0f, (b - hf.y) * 2048.0f);
  hi = *reinterpret_cast<const uint32_t*>(&h);
  lo = *reinterpret_cast<const uint32_t*>(&l);
}

// ---- one-off weight repacks (b2p_stage_finalize) -----------------------------------------------------------------
// conv2.weight [64,32,3,3] -> the tensor-memory image of A for taps 0..7: row r = hi part (r even) or lo part (r odd)
// of channel r/2, 32-bit word tap*16 + ks*8 + j = input channels ks*16 + 2j | 2j+1 (low | high half): exactly what
// tcgen05.mma reads as A[r][k] from lane r, column j of a K=16 step (column = word index).
__global__ void pack_conv2_tmem(const float* __restrict__ w2, uint32_t* __restrict__ img) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 128 * W2T_WORDS) return;
  const int r = i / W2T_WORDS, word = i % W2T_WORDS;
  const int tap = word >> 4, ks = (word >> 3) & 1, j = word & 7;
  const int oc = r >> 1, kh = tap / 3, kw = tap % 3;
  uint32_t v = 0;
#pragma unroll
  for (int e = 0; e < 2; ++e) {
    const int ic = ks * 16 + 2 * j + e;
    const uint32_t p = split_pack(w2[((oc * 32 + ic) * 3 + kh) * 3 + kw]);
    v |= ((r & 1) ? (p >> 16) : (p & 0xffffu)) << (16 * e);
  }
  // stored chunk-major -- 16-byte chunk j of row r at uint4 index j * 128 + r -- so that the kernel, which stages the
  // image in shared memory with one bulk copy, reads a row's chunk per thread without bank conflicts
  img[((word >> 2) * 128 + r) * 4 + (word & 3)] = v;
}
// ... and tap 8 (kh = kw = 2) as the byte image of its shared-memory tile: the same 128 rows, 64 B (32 input channels)
// each, 64B-swizzled K-major (16-byte chunk c of row r at chunk c ^ ((r >> 1) & 3)).
__global__ void pack_conv2_tap8(const float* __restrict__ w2, unsigned char* __restrict__ img) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;      // one 16-byte chunk each
  if (i >= 128 * 4) return;
  const int r = i >> 2, c = i & 3, oc = r >> 1;
  unsigned short v[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const uint32_t p = split_pack(w2[((oc * 32 + 8 * c + e) * 3 + 2) * 3 + 2]);
    v[e] = (unsigned short)((r & 1) ? (p >> 16) : (p & 0xffffu));
  }
  *reinterpret_cast<uint4*>(img + r * 64 + ((c ^ ((r >> 1) & 3)) * 16)) = *reinterpret_cast<const uint4*>(v);
}
// conv1.weight [32,3,3,3] -> A operand of the pool-window GEMM: two 128B-swizzled K-major tiles (hi, lo) of 128 rows;
// row m = oc*4 + dy*2 + dx, k = c*16 + r*4 + s multiplies input pixel (2py-1+r, 2px-1+s) of channel c, so its weight is
// w1[oc][c][r-dy][s-dx] where that exists and zero elsewhere.
__global__ void pack_conv1_poolwin(const float* __restrict__ w1, unsigned char* __restrict__ img) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;      // one (row, k) each
  if (i >= 128 * 64) return;
  const int m = i >> 6, k = i & 63;
  const int oc = m >> 2, dy = (m >> 1) & 1, dx = m & 1;
  float w = 0.f;
  if (k < 48) {
    const int c = k >> 4, r = (k >> 2) & 3, s = k & 3;
    const int kh = r - dy, kw = s - dx;
    if (kh >= 0 && kh < 3 && kw >= 0 && kw < 3) w = w1[((oc * 3 + c) * 3 + kh) * 3 + kw];
  }
  const uint32_t p = split_pack(w);
  const int off = m * 128 + (((k >> 3) ^ (m & 7)) * 16) + (k & 7) * 2;
  *reinterpret_cast<unsigned short*>(img + off) = (unsigned short)(p & 0xffffu);
  *reinterpret_cast<unsigned short*>(img + PLANE_BYTES + off) = (unsigned short)(p >> 16);
}

// ---- the kernel ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 1)
part0_tc2(const float* __restrict__ x,                    // [B,3,32,32]
          const unsigned char* __restrict__ a1img,        // pack_conv1_poolwin: 2 x 16 KB
          const uint32_t* __restrict__ w2t,               // pack_conv2_tmem: [128][128]
          const unsigned char* __restrict__ w2tap8,       // pack_conv2_tap8: 8 KB
          const float* __restrict__ b1, const float* __restrict__ b2,
          float* __restrict__ out,                        // [B,4096]
          int batch, int* __restrict__ range_flag,         // set to 1 when a value may have left the split-fp16 range (b2p_stage_check)
          float in_limit,                                  // |input| from which on that can happen: min(65504, (65504 - max|b1|) / max_oc sum|w1[oc]|)
          long long* __restrict__ dbg,         // dbg: optional clock64 trace of CTA 0, [role 0..3][sample 0..7][event 0..7], entry 192 = end of set-up
          RingSpan span,                                  // span.data: rows go into handoff-ring slots instead of out
          int no_wait) {                                  // ring mode only: do not wait for the preceding kernel of the stream (b2p_stage_set_pdl bit 2)
#ifndef B2P_TRACE_CTA
#define B2P_TRACE_CTA 0          // the CTA whose clock64 timeline b2p_stage_trace records (developer builds: -DB2P_TRACE_CTA=8 ...)
#endif
#define B2P_TRACE2(role, n, ev, cond) \
  do { if (dbg && blockIdx.x == B2P_TRACE_CTA && (n) < 8 && (cond)) dbg[((role) * 8 + (n)) * 8 + (ev)] = clock64(); } while (0)
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  Smem& s = *reinterpret_cast<Smem*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;   // warp index: uniform for the compiler

  // ---- one-time set-up (touches only constant weights: runs before pdl_wait) ----------------------------------
  if (dbg && tid == 0 && blockIdx.x < 256) dbg[256 + 3 * blockIdx.x] = (long long)global_ns();      // entry
  if (dbg && tid == 0 && blockIdx.x == B2P_TRACE_CTA) dbg[710] = clock64();
  pdl_trigger();
  // (the first 64 KB of `act` first serve as the landing zone of conv2's tensor-memory image, zeroed after the fill)
  for (int i = W2T_BYTES / 16 + tid; i < ACT_BYTES / 16; i += kThreads) reinterpret_cast<uint4*>(s.act)[i] = make_uint4(0, 0, 0, 0);
  for (int i = tid; i < 3 * 34 * 34; i += kThreads) (&s.in[0][0][0])[i] = 0u;
  if (tid < 32) s.b1[tid] = b1[tid];
  if (tid >= 32 && tid < 96) s.b2[tid - 32] = b2[tid - 32];
  if (tid == 0) { s.staged = 0; s.out_free[0] = 0; s.out_free[1] = 0; }
  if (warp == kMmaWarp) {
    if (lane == 0) {
      mbar_init(&s.bar_w, 1);
      mbar_init(&s.bar_patch_full, kPatchThreads / 32);
      mbar_init(&s.bar_patch_empty, 1);
      mbar_init(&s.bar_d1_full, 1);
      mbar_init(&s.bar_d1_empty, kEpiWarps);
      mbar_init(&s.bar_act_ready[0], kEpiWarps);
      mbar_init(&s.bar_act_ready[1], kEpiWarps);
      mbar_init(&s.bar_act_free, 1);
      mbar_init(&s.bar_d2_full, 1);
      mbar_init(&s.bar_d2_empty, kEpiWarps);
      mbar_init(&s.bar_w2t, 1);
      fence_barrier_init();
      // constant weights, requested first thing: conv2's tensor-memory image into the (not yet used) activation buffer,
      // conv1's A tiles and conv2's tap 8 to their places
      mbar_expect_tx(&s.bar_w2t, W2T_BYTES);
      bulk_g2s(s.act, w2t, W2T_BYTES, &s.bar_w2t);
      mbar_expect_tx(&s.bar_w, 2 * PLANE_BYTES + cifar_tc::W2_TAP_BYTES);
      bulk_g2s(s.a1, a1img, PLANE_BYTES, &s.bar_w);
      bulk_g2s(s.a1 + PLANE_BYTES, a1img + PLANE_BYTES, PLANE_BYTES, &s.bar_w);
      bulk_g2s(s.w2tap8, w2tap8, cifar_tc::W2_TAP_BYTES, &s.bar_w);
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
  // The conv1 patch producers need nothing of what follows (tensor memory, weights): they go straight to the first
  // image, whose load, split and first patches (~3 k cycles) then overlap the rest of the set-up.
  if (warp >= kPubWarp) {
    if (warp >= kE1Warp0 && warp < kE1Warp0 + 4) {           // conv2's A, taps 0..7 -> tensor memory: thread = row = lane
      // The image arrives in shared memory by ONE bulk copy issued at the top of the kernel (pack_conv2_tmem stores it
      // chunk-major).  Reading it from global memory here -- four rounds of eight dependent 16-byte loads -- took four
      // L2 round trips, 6 k cycles of every launch.  (Sixteen warps filling 32 columns each in one round was slower
      // still, +1.5 us per launch: 32 more live registers on the epilogue warps' path change the allocation of the
      // whole kernel.)
      mbar_wait(&s.bar_w2t, 0);
      const uint4* src = reinterpret_cast<const uint4*>(s.act) + 32 * (warp & 3) + lane;
      const uint32_t row = tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16) + TM_A2;
#pragma unroll 1
      for (int i = 0; i < 4; ++i) {
        uint32_t v[32];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint4 w = src[(i * 8 + j) * 128];
          v[4 * j] = w.x; v[4 * j + 1] = w.y; v[4 * j + 2] = w.z; v[4 * j + 3] = w.w;
        }
        tmem_st32(row + 32 * i, v);
      }
      tmem_st_wait();
    }
    named_bar(3, kThreads - kPatchThreads);                  // the landing zone has been read: zero it (activation halo)
    for (int i = tid - kPatchThreads; i < W2T_BYTES / 16; i += kThreads - kPatchThreads)
      reinterpret_cast<uint4*>(s.act)[i] = make_uint4(0, 0, 0, 0);
    fence_proxy_async();               // the zeroed activation halo must be visible to the tensor core
    tc_fence_before();
    named_bar(3, kThreads - kPatchThreads);
    tc_fence_after();
    if (dbg && blockIdx.x == B2P_TRACE_CTA && tid == kPubWarp * 32) dbg[3 * 64] = clock64();      // end of set-up
  }
  // input readers and output writers: both may touch the previous kernel's data.  Not so in ring mode when the caller
  // says the input is not that kernel's output (no_wait): consecutive steps write different ring slots, at most two
  // consecutive grids are resident (grids dispatch in order), and the slots' counters were last written two steps ago.
  // The previous step's stragglers -- the CTAs that publish its last messages, 5-8 us of fences -- and the flush of its
  // peer writes then overlap this step's first samples instead of standing between the steps (52 -> 4x us per step).
  if (warp != kMmaWarp && !no_wait) pdl_wait();
  if (dbg && tid == 0 && blockIdx.x < 256) dbg[257 + 3 * blockIdx.x] = (long long)global_ns();      // released by the previous kernel
  if (dbg && tid == 0 && blockIdx.x == B2P_TRACE_CTA) {
    dbg[711] = clock64();
    const unsigned long long k = atomicAdd(reinterpret_cast<unsigned long long*>(dbg + 719), 1ull);   // launches seen by this trace buffer
    dbg[720 + (k & 15)] = (long long)global_ns();            // CTA 0 released, exit: per launch (ring of 16), for back-to-back periods
  }

  if (warp == kMmaWarp) {
    // =========================== MMA issuer ========================================================================
    {
      const uint32_t lead = elect_one();
      const uint32_t tmem_base = __shfl_sync(0xffffffffu, s.tmem_slot, 0);     // warp-uniform for the compiler
      mbar_wait(&s.bar_w, 0);
      constexpr uint32_t idesc128 = make_idesc_f16(128, 128), idesc256 = make_idesc_f16(128, 256);
      const uint64_t d_act = make_desc_sw64(smem_u32(s.act));
      const uint64_t d_w8 = make_desc_sw64(smem_u32(s.w2tap8));
      const uint64_t d_phi = make_desc_sw128(smem_u32(s.patch)), d_plo = make_desc_sw128(smem_u32(s.patch) + PLANE_BYTES);
      const uint64_t d_ahi = make_desc_sw128(smem_u32(s.a1)), d_alo = make_desc_sw128(smem_u32(s.a1) + PLANE_BYTES);
      const uint32_t t_d1 = tmem_base + TM_D1, t_d2 = tmem_base + TM_D2, t_a2 = tmem_base + TM_A2;
      int n_samples = 0;
      for (int b = blockIdx.x; b < batch; b += gridDim.x) ++n_samples;

      // conv1 of half `k1 & 1` of sample k1 >> 1: 9 MMAs, lo terms first, the first hi*hi one scales them by 2^-11
      auto issue_c1 = [&](uint32_t k1) {
        mbar_wait(&s.bar_patch_full, k1 & 1);
        B2P_TRACE2(1, (k1 >> 1) - 1, 4, lead && (k1 & 1) && k1 > 1);
        mbar_wait(&s.bar_d1_empty, (k1 & 1) ^ 1);
        B2P_TRACE2(1, (k1 >> 1) - 1, 5, lead && (k1 & 1) && k1 > 1);
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < 3; ++c) {                          // A = patches (M = pixels), B = weights (N = (oc,dy,dx))
          umma_p<0>(t_d1, d_phi + 2 * c, d_alo + 2 * c, idesc128, c != 0, lead);
          umma_p<0>(t_d1, d_plo + 2 * c, d_ahi + 2 * c, idesc128, 1u, lead);
        }
        umma_ss_scale11_p(t_d1, d_phi, d_ahi, idesc128, lead);
        umma_p<0>(t_d1, d_phi + 2, d_ahi + 2, idesc128, 1u, lead);
        umma_p<0>(t_d1, d_phi + 4, d_ahi + 4, idesc128, 1u, lead);
        B2P_TRACE2(1, (k1 >> 1) - 1, 6, lead && (k1 & 1) && k1 > 1);
        umma_commit_p(&s.bar_d1_full, lead);
        B2P_TRACE2(1, (k1 >> 1) - 1, 7, lead && (k1 & 1) && k1 > 1);
        umma_commit_p(&s.bar_patch_empty, lead);
      };
      // conv2 of image rows 8h..8h+7 of sample n: 18 MMAs of N = 256 (128 pixels x hi|lo).  Tap row kh reads activation
      // rows 8h+kh-1 .. 8h+kh+6, so of the upper half only kh = 2 touches row 8, the first row conv1's SECOND half
      // produces: those six MMAs wait for it separately, the other twelve start as soon as the first half is there.
      auto issue_q2 = [&](uint32_t n, uint32_t h) {
        const uint32_t k2 = 2 * n + h;
        if (h == 0) {
          mbar_wait(&s.bar_act_ready[0], n & 1);
          B2P_TRACE2(0, n, 1, lead);
        }
        mbar_wait(&s.bar_d2_empty, (k2 & 1) ^ 1);
        B2P_TRACE2(0, n, h ? 5 : 2, lead);
        tc_fence_after();
        const uint64_t d_half = d_act + ((h * 8 * ACT_ROW_BYTES) >> 4);
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
          const int kh = tap / 3, kw = tap % 3;
          if (tap == 6 && h == 0) {
            mbar_wait(&s.bar_act_ready[1], n & 1);
            tc_fence_after();
          }
#pragma unroll
          for (int ks = 0; ks < 2; ++ks) {
            const uint64_t db = d_half + ((kw * ACT_COPY_BYTES + kh * ACT_ROW_BYTES + ks * 32) >> 4);
            if (tap < 8) umma_ts_p(t_d2, t_a2 + tap * 16 + ks * 8, db, idesc256, (tap | ks) != 0, lead);
            else umma_p<0>(t_d2, d_w8 + 2 * ks, db, idesc256, 1u, lead);
          }
        }
        umma_commit_p(&s.bar_d2_full, lead);
        if (h == 1) umma_commit_p(&s.bar_act_free, lead);
      };

      if (n_samples > 0) { issue_c1(0); issue_c1(1); }
      for (int n = 0; n < n_samples; ++n) {
        B2P_TRACE2(0, n, 0, lead);
        issue_q2(n, 0);
        B2P_TRACE2(0, n, 3, lead);
        if (n + 1 < n_samples) issue_c1(2 * (n + 1));
        B2P_TRACE2(0, n, 4, lead);
        issue_q2(n, 1);
        B2P_TRACE2(0, n, 6, lead);
        if (n + 1 < n_samples) issue_c1(2 * (n + 1) + 1);
        B2P_TRACE2(0, n, 7, lead);
      }
    }
    __syncwarp();
  } else if (warp < kPubWarp) {
    // =========================== conv1 patch producers ================================================================
    // thread = pooled pixel of the current half; per channel its 4x4 input window becomes 16 K-values (32 B hi, 32 B lo)
    const int t = tid - kPatchWarp0 * 32;                    // 0..63: pooled pixels t and t + 64 of the half
    constexpr int NLD = 768 / kPatchThreads;                 // float4 loads per thread and image
    float4 pre[NLD];
    if ((int)blockIdx.x < batch) {
      const float4* xin = reinterpret_cast<const float4*>(x + (size_t)blockIdx.x * 3072);
#pragma unroll
      for (int j = 0; j < NLD; ++j) pre[j] = ld_produced(xin + t + kPatchThreads * j);
    }
    uint32_t n = 0;
    float in_max = 0.f;
    for (int b = blockIdx.x; b < batch; b += gridDim.x, ++n) {
      named_bar(1, kPatchThreads);                           // the previous sample's window reads are done
#pragma unroll
      for (int j = 0; j < NLD; ++j) {
        const int e = (t + kPatchThreads * j) * 4;
        // split ONCE per pixel here (every pixel belongs to four pool windows): the window loop below only moves halves
        uint32_t* d = &s.in[e >> 10][((e >> 5) & 31) + 1][(e & 31) + 1];
        uint32_t h01, l01, h23, l23;
        split2(pre[j].x, pre[j].y, h01, l01);
        split2(pre[j].z, pre[j].w, h23, l23);
        d[0] = __byte_perm(h01, l01, 0x5410); d[1] = __byte_perm(h01, l01, 0x7632);
        d[2] = __byte_perm(h23, l23, 0x5410); d[3] = __byte_perm(h23, l23, 0x7632);
#ifndef B2P_NO_RANGE_CHECK
        in_max = fmaxf(in_max, fmaxf(fmaxf(fabsf(pre[j].x), fabsf(pre[j].y)), fmaxf(fabsf(pre[j].z), fabsf(pre[j].w))));   // once per pixel
#endif
      }
      named_bar(1, kPatchThreads);
      B2P_TRACE2(3, n, 1, t == 0);
      if (b + (int)gridDim.x < batch) {                      // next image: in flight while this one is processed
        const float4* xin = reinterpret_cast<const float4*>(x + (size_t)(b + gridDim.x) * 3072);
#pragma unroll
        for (int j = 0; j < NLD; ++j) pre[j] = ld_produced(xin + t + kPatchThreads * j);
      }
#pragma unroll 1
      for (uint32_t h = 0; h < 2; ++h) {
        const uint32_t k1 = 2 * n + h;
        mbar_wait(&s.bar_patch_empty, (k1 & 1) ^ 1);         // the MMAs that read the previous half's patches are done
        B2P_TRACE2(3, n, 2 + 2 * h, t == 0);
#pragma unroll 1
        for (int i = t; i < 128; i += kPatchThreads) {       // pooled pixel i of this half
          const int py = 8 * h + (i >> 4), pc = i & 15;
          unsigned char* row_hi = s.patch + i * 128;
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            uint32_t hi[8], lo[8];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
              const uint2* rp = reinterpret_cast<const uint2*>(&s.in[c][2 * py + r][2 * pc]);
              const uint2 a = rp[0], bq = rp[1];
              hi[2 * r] = __byte_perm(a.x, a.y, 0x5410);       lo[2 * r] = __byte_perm(a.x, a.y, 0x7632);
              hi[2 * r + 1] = __byte_perm(bq.x, bq.y, 0x5410); lo[2 * r + 1] = __byte_perm(bq.x, bq.y, 0x7632);

            }
#pragma unroll
            for (int half = 0; half < 2; ++half) {           // 16-byte chunk 2c + half = window rows 2*half, 2*half+1
              const int off = ((2 * c + half) ^ (i & 7)) * 16;
              *reinterpret_cast<uint4*>(row_hi + off) = make_uint4(hi[4 * half], hi[4 * half + 1], hi[4 * half + 2], hi[4 * half + 3]);
              *reinterpret_cast<uint4*>(row_hi + PLANE_BYTES + off) = make_uint4(lo[4 * half], lo[4 * half + 1], lo[4 * half + 2], lo[4 * half + 3]);
            }
          }
        }
#ifndef B2P_NO_RANGE_CHECK
        if (in_max >= in_limit) *range_flag = 1;             // see in_limit above (NaN input: not flagged, it is data)
#endif
        fence_proxy_async();                                 // generic-proxy writes -> async proxy (UMMA)
        __syncwarp();
        if (lane == 0) mbar_arrive(&s.bar_patch_full);
        B2P_TRACE2(3, n, 3 + 2 * h, t == 0);
      }
    }
  } else if (warp >= kE1Warp0) {
    // =========================== epilogue warps (16): conv2 drain and conv1 epilogue in the tensor pipe's order =========
    // The MMA stream is  Q2(n,0) C1(n+1,0) Q2(n,1) C1(n+1,1) ...  and conv2 has ONE accumulator, so the time to drain it
    // is what the tensor pipe waits for.  All 16 warps therefore drain it together (E2, one image-row pair each), and
    // between two drains do their share of the conv1 epilogue (E1, 32 pooled pixels each): read and pool right away,
    // but write the activation only after the second drain -- it may not be overwritten before Q2(n,1) has completed,
    // which is also the moment that drain starts, and the drain goes first.
    const int q = warp & 3, part = (warp - kE1Warp0) >> 2;    // TMEM lane quarter; 0..3
    // ---- E2 geometry: lane 2c = channel 16q + c (even lanes finish), columns 64*part .. +63 = image rows 8h+2*part, +1
    const bool even = (lane & 1) == 0;
    const int oc2 = 16 * q + (lane >> 1);
    const float bias2 = s.b2[oc2];
    const uint32_t t_row2 = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + TM_D2 + 64 * part;
    // ---- E1 geometry: lane = pooled pixel 32q + lane of the half (pooled rows 8h+2q, +1), columns 32*part .. +31 =
    //      channels 8*part .. +7, the four window positions of a channel side by side: pooling stays inside the thread
    const uint32_t t_row1 = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + TM_D1 + 32 * part;
    const bool tracer = warp == kE1Warp0 && lane == 0;

    uint4 chunk_hi, chunk_lo;                                  // E1 result held between read and write
    // conv1 epilogue, first part: TMEM -> pooled, biased, rectified, split (this pixel's channels 8*part .. +7)
    auto e1_read = [&](uint32_t k1) {
      mbar_wait(&s.bar_d1_full, k1 & 1);
      tc_fence_after();
      uint32_t v[32];
      tmem_ld32(t_row1, v);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&s.bar_d1_empty);           // the accumulator may be overwritten (next half's MMAs)
      const float4 ba = *reinterpret_cast<const float4*>(&s.b1[8 * part]);
      const float4 bb = *reinterpret_cast<const float4*>(&s.b1[8 * part + 4]);
      const float bias[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
      float u[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {                          // max(a)+b == max(a+b); ReLU
        const float m = max_nan(max_nan(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1])),
                                max_nan(__uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3])));
        u[j] = max_nan(m + bias[j], 0.f);
      }
      split2(u[0], u[1], chunk_hi.x, chunk_lo.x);
      split2(u[2], u[3], chunk_hi.y, chunk_lo.y);
      split2(u[4], u[5], chunk_hi.z, chunk_lo.z);
      split2(u[6], u[7], chunk_hi.w, chunk_lo.w);
    };
    // conv1 epilogue, second part: this lane's pixel (16-byte hi and lo chunk of its 8 channels) into the three copies
    auto e1_write = [&](uint32_t h) {
      const int y = 8 * h + 2 * q + (lane >> 4), xc = lane & 15;
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const int xs = xc - kw + 1;                            // its column in copy kw (copy kw holds pixel (y, x+kw-1) at x)
        if (xs < 0 || xs > 15) continue;
        unsigned char* dst = s.act + kw * ACT_COPY_BYTES + ((y + 1) * 2 + (xs >> 3)) * 1024 + (xs & 7) * 64 +
                             ((part ^ ((xs & 7) >> 1)) * 16);
        *reinterpret_cast<uint4*>(dst) = chunk_hi;
        *reinterpret_cast<uint4*>(dst + 512) = chunk_lo;
      }
      fence_proxy_async();                                     // generic-proxy writes -> async proxy (UMMA)
      __syncwarp();
      if (lane == 0) mbar_arrive(&s.bar_act_ready[h]);
    };
    // conv2 drain, first part: TMEM -> registers, accumulator released at once.  Kept per pixel: the even lanes'
    // W_hi*a_hi + 2^-11 W_hi*a_lo, the odd lanes' 2^-11 W_lo*a_hi (32 values for 2 rows x 16 pixels).
    float dv[4][8];
    auto e2_load = [&](uint32_t n, uint32_t h, bool rows67) {
      mbar_wait(&s.bar_d2_full, (2 * n + h) & 1);
      tc_fence_after();
      uint32_t r[4][16];
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        cifar_tc::tmem_ld16(t_row2 + 16 * g, r[2 * g]);            // row y:   hi x8 | lo x8   (left / right 8 pixels)
        cifar_tc::tmem_ld16(t_row2 + 32 + 16 * g, r[2 * g + 1]);   // row y+1
      }
      if (rows67) e1_write(0);                                 // Q2(n,1) has completed: rows 6, 7 go out while the loads fly
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&s.bar_d2_empty);             // the accumulator may be overwritten (other half's MMAs)
#pragma unroll
      for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float lo_or_0 = even ? __uint_as_float(r[k][8 + j]) : 0.f;       // odd lanes: their lo columns are not used
          dv[k][j] = fmaf(lo_or_0, 1.0f / 2048.0f, __uint_as_float(r[k][j]) * (even ? 1.0f : 1.0f / 2048.0f));
        }
    };
    // second part: the odd lanes hand their term to the even lane below (one shuffle per value); the even lanes pool 2x2
    // inside the thread, add the bias, apply ReLU and store 4 consecutive outputs of the flatten order
    auto e2_finish = [&](uint32_t n, uint32_t h) {
      if (h == 0 && n > 1) {                                   // the bulk copy of sample n-2 must be done reading this buffer
        while (ld_acquire_cta(&s.out_free[n & 1]) < (int)(n >> 1)) { }
      }
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        float v0[8], v1[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          v0[j] = dv[2 * g][j] + __shfl_down_sync(0xffffffffu, dv[2 * g][j], 1);
          v1[j] = dv[2 * g + 1][j] + __shfl_down_sync(0xffffffffu, dv[2 * g + 1][j], 1);
        }
        float4 o4;
        float* op = &o4.x;
#pragma unroll
        for (int j = 0; j < 4; ++j)
          op[j] = max_nan(max_nan(max_nan(v0[2 * j], v0[2 * j + 1]), max_nan(v1[2 * j], v1[2 * j + 1])) + bias2, 0.f);
        if (even) *reinterpret_cast<float4*>(&s.out[n & 1][oc2 * 64 + (4 * h + part) * 8 + 4 * g]) = o4;
      }
    };

    int n_samples = 0;
    for (int b = blockIdx.x; b < batch; b += gridDim.x) ++n_samples;
    if (n_samples > 0) {                                       // sample 0's conv1: nothing to wait for
      e1_read(0); e1_write(0);
      e1_read(1); e1_write(1);
    }
    for (int n = 0; n < n_samples; ++n) {
      const bool more = n + 1 < n_samples;
      B2P_TRACE2(2, n, 0, tracer);
      e2_load(n, 0, false);
      e2_finish(n, 0);
      B2P_TRACE2(2, n, 1, tracer);
      if (more) {
        e1_read(2 * (n + 1));
        B2P_TRACE2(1, n, 0, tracer);
        // Q2(n,0) has completed and Q2(n,1) reads rows 7..16: rows 0..5 of the next sample can be written now, only
        // the warps holding rows 6, 7 (lane quarter 3) wait for Q2(n,1) -- inside the second drain
        if (q != 3) e1_write(0);
        B2P_TRACE2(1, n, 1, tracer);
      }
      e2_load(n, 1, more && q == 3);                           // Q2(n,1) has completed: the accumulator first
      B2P_TRACE2(2, n, 2, tracer);
      if (more) {
        mbar_wait(&s.bar_act_free, n & 1);                     // (passes at once: committed together with d2_full)
        e1_read(2 * (n + 1) + 1);
        B2P_TRACE2(1, n, 2, tracer);
        e1_write(1);
        B2P_TRACE2(1, n, 3, tracer);
      }
      e2_finish(n, 1);
      fence_proxy_async();                                     // this thread's writes to `out` -> visible to the copy engine
      named_bar(2, 16 * 32);                                   // all 4096 outputs are staged
      if (tracer) st_release_cta(&s.staged, n + 1);
      B2P_TRACE2(2, n, 3, tracer);
    }
  } else {
    // =========================== publisher (+ ring handshake, ring_hooks.cuh) =========================================
    // One finished sample leaves as ONE 16 KB bulk copy into HBM or -- ring mode -- straight into the consumer GPU's slot.
    // Ring mode: before the copy the slot must have been released by the consumer; ONE CTA per message -- the one that
    // holds the message's first row -- polls the consumer's counter over NVLink and posts what it saw in local memory,
    // the others wait on that local word (sys-scope reads of one remote line serialise).  After the copy the row is
    // counted into the message's ticket; whoever counts the last row publishes the message (one system-scope release).
    // Two lanes share the work, lane l taking the samples whose staging buffer is l: each sample's chain -- copy, its
    // completion over NVLink, a gpu-scope fence, the ticket -- takes about as long as a sample's arithmetic, so one
    // thread doing them back to back set the pace of the whole kernel in ring mode.
    // Slot handshake with the consumer GPU, done for the WHOLE step before the first copy: at kernel start the consumer
    // has normally released every slot this step will use (the ring holds two steps), so the NVLink round trip of the
    // poll and the system-scope fence behind it (2.5 us and more with peer traffic outstanding) pass while the first
    // sample is still being computed.  The message this CTA leads (it holds the message's first row; at most one, 148
    // CTAs > 128 rows) is polled first and the value posted for the other CTAs, then the CTA waits for what the other
    // leaders posted.  (Acquiring per sample inside the copy loop stalled a leader's lane for 5-8 us per message and
    // everybody else behind it; a third lane of this warp running ahead was starved by its spinning siblings.)
    if (span.data) {
      if (lane == 0) {
        for (int b = blockIdx.x; b < batch; b += gridDim.x) {
          const int msg = b / span.rows_per_msg;
          if (b != msg * span.rows_per_msg) continue;
          const int slot = span_slot(span, msg);
          const unsigned long long k = ld_count(span.count + slot);      // messages already published into this slot
          if (k > 0) {
            spin_until_ge_remote(span.released + slot, k);
            st_release_gpu(span.acq_seen + slot, k);
          }
        }
        int acquired = -1;
        for (int b = blockIdx.x; b < batch; b += gridDim.x) {
          const int msg = b / span.rows_per_msg;
          if (msg == acquired) continue;
          const int slot = span_slot(span, msg);
          const unsigned long long k = ld_count(span.count + slot);
          if (k > 0 && b != msg * span.rows_per_msg)
            while (ld_acquire_gpu(span.acq_seen + slot) < k) __nanosleep(64);
          acquired = msg;
        }
      }
      __syncwarp();
    }
    // Four lanes take the samples in turn (sample n: lane n & 3, staging buffer n & 1).  In ring mode a sample's chain
    // after the copy -- gpu-scope fence + ticket 1-3.4 us, and for whoever counts a message's last row another fence and
    // the system-scope release, 3-6.6 us with peer traffic outstanding -- is longer than a sample's arithmetic; with two
    // lanes the CTA that published one message fell behind, was last again for the next one, and ended the kernel 10 us
    // after the others (two CTAs published all eight messages of a step).
    if (lane < kPubLanes) {
      uint32_t n = 0;
      for (int b = blockIdx.x; b < batch; b += gridDim.x, ++n) {
        if ((n & (kPubLanes - 1)) != (uint32_t)lane) continue;
        float* dst_row = out + (size_t)b * 4096;
        int msg = 0, slot = 0;
        if (span.data) {
          msg = b / span.rows_per_msg;
          slot = span_slot(span, msg);
          dst_row = reinterpret_cast<float*>(span.data + (size_t)slot * span.slot_bytes) + (size_t)(b - msg * span.rows_per_msg) * 4096;
        }
        B2P_TRACE2(2, n, 4, true);                              // slot acquired
        while (ld_acquire_cta(&s.staged) <= (int)n) __nanosleep(32);
        B2P_TRACE2(2, n, 5, true);                              // sample staged
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_row), "r"(smem_u32(s.out[n & 1])),
                     "r"(64 * 64 * 4) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        cifar_tc::bulk_wait_read();                             // the staging buffer has been read
        st_release_cta(&s.out_free[n & 1], (int)(n >> 1) + 1);
        B2P_TRACE2(2, n, 6, true);                              // staging buffer read
        if (span.data) {
          const long long tq0 = dbg ? clock64() : 0;
          cifar_tc::bulk_wait_all();                            // the row has arrived at the consumer
          const long long tq1 = dbg ? clock64() : 0;
          __threadfence();
          const int rows = min(span.rows_per_msg, batch - msg * span.rows_per_msg);
          if (atomicAdd(span.tickets + slot, 1) == rows - 1) {
            const long long tq2 = dbg ? clock64() : 0;
            span.tickets[slot] = 0;                             // ready for the slot's next message
            __threadfence();
            const unsigned long long k = ld_count(span.count + slot) + 1;
            span.count[slot] = k;
            st_release_sys(span.published + slot, k);
            if (dbg && msg < 8) {                               // developer trace: who published message msg, and what it cost
              dbg[760 + 4 * msg] = blockIdx.x * 16 + n;
              dbg[761 + 4 * msg] = tq1 - tq0;                   // copy completion
              dbg[762 + 4 * msg] = tq2 - tq1;                   // gpu-scope fence + ticket
              dbg[763 + 4 * msg] = clock64() - tq2;             // fence + system-scope release
            }
          }
          B2P_TRACE2(2, n, 7, true);                            // row counted (and the message published, if it was the last)
        }
      }
      cifar_tc::bulk_wait_all();
    }
    __syncwarp();
  }

  // (no_wait: executing the skipped griddepcontrol.wait HERE, so that the grid at least completes after its predecessor,
  //  was measured to cost the whole gain -- 49.9 against 39.3 us per step on two GPUs -- and nothing on the producer GPU
  //  needs that order.  Tried on ONE GPU too -- alternating activation buffers, the next step's stage-0 kernel not waiting
  //  for the tail kernel of this one: no gain, 67.8 us per step either way; the tail's 1024 small CTAs hold every SM's
  //  registers until they have drained, and a slim persistent tail that would leave room took 18 us instead of 7.)
  tc_fence_before();
  __syncthreads();
  if (dbg && tid == kMmaWarp * 32 && blockIdx.x < 256) {     // exit (stamped by a thread that has work until the end)
    dbg[258 + 3 * blockIdx.x] = (long long)global_ns();
    if (blockIdx.x == B2P_TRACE_CTA) {
      dbg[712] = clock64();
      dbg[736 + ((dbg[719] - 1) & 15)] = (long long)global_ns();
    }
  }
  if (warp == kMmaWarp) tmem_dealloc(tmem_base, 512);
#undef B2P_TRACE2
}

}  // namespace cifar_tc2
}  // namespace b2p
