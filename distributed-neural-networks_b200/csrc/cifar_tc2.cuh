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
//          A = [W_hi ; W_lo]  (64 output channels, fp16 hi part stacked on the lo part) sits in TENSOR MEMORY for taps
//              0..7 (128 columns; tap 8 is read from shared memory: 144 columns would not leave room for the
//              accumulators), and
//          B = 128 pixels (8 image rows) of the conv1 activation, hi rows and lo rows interleaved in groups of 8,
//              so ONE instruction yields  W_hi*a_hi | W_hi*a_lo  (lanes 0..63)  and  W_lo*a_hi | (W_lo*a_lo, unused)
//              (lanes 64..127): the three products of the split-fp16 scheme in 36 wide MMAs per sample.
//          The B operand of tap (kh,kw) is a plain shifted window of the activation (three x-shifted copies with a
//          zero halo, as in round 1): no im2col.  The accumulator comes out [channel][pixel], i.e. already in the
//          reference's flatten order c*64+h*8+w, and the 2x2 pool is between a thread's own columns.
//   conv1  as a "pool-window" GEMM: one column per POOLED pixel, one row per (channel, dy, dx) of its 2x2 window
//          (32 x 4 = 128 rows: a full M), K = 3 channels x 4x4 input window = 48 (weights zero-padded 3x3 -> 4x4).
//          Three split-fp16 products into ONE accumulator: lo terms first, then the first hi*hi MMA carries
//          scale-input-d = 2^-11 (D = A*B + D*2^-11).  18 MMAs (N = 128) per sample instead of 6.9 k FP32-pipe cycles.
//          The epilogue pools across the 4 lanes of a channel with lane-swapping shuffles, transposes 8x8 across
//          lanes and writes the fp16 hi/lo activation rows for conv2.
//
// Warp roles (23 warps, one persistent CTA per SM, samples blockIdx.x, +gridDim.x, ...):
//   0        MMA issuer            1..4    conv1 patch producers (input image -> 4x4 windows, fp16 hi/lo)
//   5..12    conv1 epilogue        13..20  conv2 epilogue (lanes 64..127 hand their cross term to lanes 0..63
//   21       ring acquirer                  through shared memory) -> padded staging tile
//   22       publisher: bulk copies of the finished sample to HBM or straight into the next GPU's ring slot
//
// Arithmetic and range are those of round 1's kernel: x ~= hi + lo * 2^-11 with fp16 hi/lo, products exact in the
// fp32 accumulator, the lo*lo term dropped.  |input| and |conv1 output| must stay below 65504 (fp16 range); beyond
// it the result is non-finite where the reference's would be finite -- use math="fp32" for such data (b200pipe.h).
// NaN inputs propagate like torch's ReLU / MaxPool2d (max.NaN), not like fmaxf.
#pragma once
#include "cifar_tc.cuh"

namespace b2p {
namespace cifar_tc2 {

using namespace tc;
using cifar_tc::bulk_g2s;
using cifar_tc::make_desc_sw64;
using cifar_tc::make_idesc_f16;
using cifar_tc::named_bar;

constexpr int kMmaWarp = 0;
constexpr int kPatchWarp0 = 1, kPatchThreads = 128;
constexpr int kE1Warp0 = 5, kE1Warps = 8;
constexpr int kE2Warp0 = 13, kE2Warps = 8;
constexpr int kAcqWarp = 21, kPubWarp = 22;
constexpr int kThreads = 23 * 32;                     // 736

// tensor-memory map (512 columns)
constexpr uint32_t TM_A2 = 0;                         // conv2 weights, taps 0..7: 8 taps x 2 K-steps x 8 columns
constexpr uint32_t TM_D2 = 128;                       // conv2 accumulator of one half sample: 256 columns
constexpr uint32_t TM_D1 = 384;                       // conv1 accumulator of one half sample: 128 columns

constexpr int ACT_ROW_BYTES = 2048;                   // one image row of one copy: 2 groups x (8 hi rows + 8 lo rows) x 64 B
constexpr int ACT_COPY_BYTES = 18 * ACT_ROW_BYTES;    // rows -1..16
constexpr int ACT_BYTES = 3 * ACT_COPY_BYTES;         // 110 592
constexpr int PLANE_BYTES = 128 * 128;                // a 128-row operand tile with 128-byte rows
constexpr int OUT_PITCH = 68;                         // floats per channel row of the staging tile (64 + 4: no bank conflicts)
constexpr int W2T_WORDS = 128;                        // tensor-memory image of conv2's weights: 32-bit words per row

struct Smem {
  unsigned char act[ACT_BYTES];                       // conv2's B operand: 3 x-shifted copies, hi/lo interleaved by 8 pixels
  unsigned char patch[2 * PLANE_BYTES];               // conv1's B operand: [hi | lo][128 pooled pixels][128 B], 96 B used
  unsigned char a1[2 * PLANE_BYTES];                  // conv1's A operand: [hi | lo][128 rows (oc,dy,dx)][128 B]
  unsigned char w2tap8[cifar_tc::W2_TAP_BYTES];       // conv2 tap 8: [W_hi ; W_lo] rows of 64 B, 64B-swizzled
  float out[64 * OUT_PITCH];                          // one finished sample, [channel][h*8+w]
  unsigned char xch[4 * 2048];                        // conv2 epilogue: cross terms of lanes 64..127 on their way to lanes 0..63
  float in[3][34][34];                                // zero-padded input image
  float b1[32], b2[64];
  uint64_t bar_w, bar_patch_full, bar_patch_empty, bar_d1_full, bar_d1_empty, bar_act_ready, bar_act_free,
      bar_d2_full, bar_d2_empty;
  uint32_t tmem_slot;
  int acq_upto, staged, out_free;                     // ring / staging hand-offs (see cifar_tc.cuh)
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
__device__ __forceinline__ void bar_arrive(int id, int threads) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// ---- one-off weight repacks (b2p_stage_finalize) -----------------------------------------------------------------
// conv2.weight [64,32,3,3] -> the tensor-memory image of A = [W_hi ; W_lo] for taps 0..7: row r (0..63 hi part of
// channel r, 64..127 lo part of channel r-64), 32-bit word tap*16 + ks*8 + j = input channels ks*16 + 2j | 2j+1
// (low | high half): exactly what tcgen05.mma reads as A[r][k] from lane r, column j of a K=16 step.
__global__ void pack_conv2_tmem(const float* __restrict__ w2, uint32_t* __restrict__ img) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 128 * W2T_WORDS) return;
  const int r = i / W2T_WORDS, word = i % W2T_WORDS;
  const int tap = word >> 4, ks = (word >> 3) & 1, j = word & 7;
  const int oc = r & 63, kh = tap / 3, kw = tap % 3;
  uint32_t v = 0;
#pragma unroll
  for (int e = 0; e < 2; ++e) {
    const int ic = ks * 16 + 2 * j + e;
    const uint32_t p = split_pack(w2[((oc * 32 + ic) * 3 + kh) * 3 + kw]);
    v |= (r < 64 ? (p & 0xffffu) : (p >> 16)) << (16 * e);
  }
  img[i] = v;
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

