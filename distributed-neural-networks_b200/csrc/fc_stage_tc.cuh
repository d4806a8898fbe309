// fc_stage_tc.cuh -- fc1 of CIFAR part 1 in the tensor-core parity mode (math = tf32x3), split-K 32.
//
// Reference: ModelPart1_2Node.forward, cifar_model_parts.py:53-58
//     x = relu(fc1(x)); x = fc2(x); return softmax(x, dim=1)          [B,4096] -> [B,10]
//
// Part 1 used to be three launches per microbatch: split_tf32 (x -> hi/lo planes in HBM), the generic
// linear_tc<tf32x3> with 16 K slices (64 CTAs, each pulling 512 KB through one SM), and fc_tail.  This kernel
// replaces the first two:
//   * 128 CTAs per 128 rows -- CTA (z, j, mt): K slice z (128 of 4096), hidden-unit tile j (128 of 512), row tile
//     mt -- so the 16.8 MB of weight planes and the activation are spread over 128 SMs (192 KB each);
//   * x is loaded ONCE as raw fp32 by TMA and split in shared memory by four converter warps: hi = rna_tf32(x)
//     in place, lo = rna_tf32(x - hi) beside it (elementwise, so TMA's 128-byte swizzle is preserved);
//   * the weight planes of the first three k-blocks are requested before griddepcontrol.wait (they are constants),
//     so most of the CTA's bytes are in flight while the previous kernel of the stream drains;
//   * tcgen05.mma kind::tf32, three products per k-step (lo*hi, hi*lo, hi*hi) into one 128 x 128 accumulator,
//     written as partial sums ws[z][row][512]; fc_tail_f32<32> adds the 32 slices in ascending order.
// A row's result depends only on that row (fixed K slicing, fixed summation order): micro-batching is bit-exact.
//
// Measured dead end, kept as a note: a fully fused variant (clusters of 8 K slices reducing over distributed
// shared memory, then two levels of last-arrival tickets in L2 for the remaining sums, fc2 and the soft-max) was
// correct but took 34 us per launch against 16 us for the three kernels: 56 KB of DSMEM reads per CTA ran at
// ~4 B/clk (13.7 k cycles), and each ticket level is a chain of three to four dependent L2 round trips
// (fence, atomic, reload) -- a dependent launch is the cheaper grid-wide barrier.
#pragma once
#include "tc_gemm2.cuh"
#include "gpt_simt.cuh"      // round_tf32
#include "ring_hooks.cuh"

namespace b2p {
namespace fcs {

using namespace tc;

constexpr int BM = 128, BN = 128;           // rows x hidden units per CTA
constexpr int N_HIDDEN = 512, K_IN = 4096;
constexpr int KB = 32;                      // fp32 per 128-byte swizzle row
constexpr int KSLICE = 128, KBLOCKS = KSLICE / KB;
constexpr int NSLICES = K_IN / KSLICE;      // 32 partial sums per output element
constexpr int NTILES = N_HIDDEN / BN;       // 4
constexpr int PLANE = BM * 128;             // 16 KB: 128 rows x 32 fp32
constexpr int STAGE_BYTES = 4 * PLANE;      // x hi | x lo | W hi | W lo
constexpr int STAGES = 3;
constexpr int BAR_BYTES = 256;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + BAR_BYTES + 1024;
constexpr int THREADS = 192;                // TMA warp, MMA warp, 4 convert/epilogue warps

// optional per-CTA timeline of row tile 0 (clock64 at 14 points, thread 64), read back by b2p_debug_fc_trace
__device__ long long g_fc_trace[NSLICES * NTILES * 16];
#define FC_TRACE(ev) do { if (trace_on && threadIdx.x == 64 && blockIdx.z == 0) \
    g_fc_trace[(blockIdx.y * NSLICES + blockIdx.x) * 16 + (ev)] = clock64(); } while (0)

__global__ void __launch_bounds__(THREADS, 1)
fc1_x3_splitk(const __grid_constant__ CUtensorMap tmap_x,       // x [M,4096] fp32, box [32 k][128 rows]
            const __grid_constant__ CUtensorMap tmap_w_hi,    // fc1.weight hi plane [512,4096], box [32 k][128 rows]
            const __grid_constant__ CUtensorMap tmap_w_lo,    // fc1.weight lo plane
            float* __restrict__ ws,                           // partial sums [32][M][512]
            int M, int trace_on,
            RingSpan span) {                                  // span.data: tmap_x covers the whole ring, row tile mt is message mt
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  unsigned char* tiles = smem;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* full = bars;                  // [STAGES] TMA -> converters, MMA
  uint64_t* conv = bars + STAGES;         // [STAGES] converters -> MMA
  uint64_t* empty = bars + 2 * STAGES;    // [STAGES] MMA -> TMA
  uint64_t* tmem_full = bars + 3 * STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int z = blockIdx.x, j = blockIdx.y, mt = blockIdx.z;
  const int m0 = mt * BM, n0 = j * BN, k0 = z * KSLICE;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_x)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_w_hi)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_w_lo)) : "memory");
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&conv[i], 4); mbar_init(&empty[i], 1); }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, BN);
    tmem_relinquish();
  }
  FC_TRACE(0);
  pdl_trigger();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  FC_TRACE(1);

  if (warp == 0) {
    // ===================== TMA producer =====================
    // fc1.weight is constant: its first STAGES k-blocks are requested BEFORE griddepcontrol.wait, so two thirds of
    // this CTA's bytes are already on their way while the previous kernel in the stream drains.
    if (lane == 0) {
#pragma unroll
      for (int kb = 0; kb < (KBLOCKS < STAGES ? KBLOCKS : STAGES); ++kb) {
        unsigned char* sa = tiles + kb * STAGE_BYTES;
        mbar_expect_tx(&full[kb], 3 * PLANE);
        tma_load_2d(sa + 2 * PLANE, &tmap_w_hi, &full[kb], k0 + kb * KB, n0);
        tma_load_2d(sa + 3 * PLANE, &tmap_w_lo, &full[kb], k0 + kb * KB, n0);
      }
    }
    __syncwarp();
    pdl_wait();          // x belongs to the previous kernel until here
    int x_row0 = m0;
    if (span.data) {
      // row tile mt is message mt of this launch: wait until the producer GPU has published it
      const int slot = span_slot(span, mt);
      x_row0 = slot * BM;                // the ring's slots are 128 rows each, back to back (checked by the host)
      if (lane == 0) {
        spin_until_ge(span.published + slot, ld_count(span.count + slot) + 1);     // acquire: the peer's stores are visible ...
        fence_proxy_async();             // ... also to the TMA unit (async proxy)
      }
      __syncwarp();
    }
    for (int kb = 0; kb < KBLOCKS; ++kb) {
      const int stage = kb % STAGES, round = kb / STAGES;
      mbar_wait(&empty[stage], (round & 1) ^ 1);
      if (lane == 0) {
        unsigned char* sa = tiles + stage * STAGE_BYTES;
        if (kb >= STAGES) {
          mbar_expect_tx(&full[stage], 3 * PLANE);
          tma_load_2d(sa + 2 * PLANE, &tmap_w_hi, &full[stage], k0 + kb * KB, n0);
          tma_load_2d(sa + 3 * PLANE, &tmap_w_lo, &full[stage], k0 + kb * KB, n0);
        }
        tma_load_2d(sa, &tmap_x, &full[stage], k0 + kb * KB, x_row0);                   // raw fp32, becomes the hi plane
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = make_idesc(1, BM, BN);
    for (int kb = 0; kb < KBLOCKS; ++kb) {
      const int stage = kb % STAGES, round = kb / STAGES;
      mbar_wait(&full[stage], round & 1);
      mbar_wait(&conv[stage], round & 1);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t base = smem_u32(tiles + stage * STAGE_BYTES);
        const uint64_t xa = make_desc_sw128(base), xl = make_desc_sw128(base + PLANE);
        const uint64_t wh = make_desc_sw128(base + 2 * PLANE), wl = make_desc_sw128(base + 3 * PLANE);
#pragma unroll
        for (int k = 0; k < KB / 8; ++k) {              // tf32 k-step = 8 elements = 32 B = +2 in the descriptor
          umma<1>(tmem_base, xl + 2 * k, wh + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);      // small terms first
          umma<1>(tmem_base, xa + 2 * k, wl + 2 * k, idesc, 1u);
          umma<1>(tmem_base, xa + 2 * k, wh + 2 * k, idesc, 1u);
        }
        umma_commit(&empty[stage]);
        if (kb == KBLOCKS - 1) umma_commit(tmem_full);
      }
      __syncwarp();
    }
  } else {
    // ===================== converters, then epilogue (warps 2..5) =====================
    const int et = threadIdx.x - 64;                  // 0..127
    pdl_wait();          // ws is still being read by the previous microbatch's tail until here
    FC_TRACE(2);
    for (int kb = 0; kb < KBLOCKS; ++kb) {
      const int stage = kb % STAGES, round = kb / STAGES;
      mbar_wait(&full[stage], round & 1);
      if (kb == 0) FC_TRACE(3);
      float4* xs = reinterpret_cast<float4*>(tiles + stage * STAGE_BYTES);
      float4 v[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = xs[et + 128 * i];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float4 h, l;
        h.x = round_tf32(v[i].x); l.x = round_tf32(v[i].x - h.x);
        h.y = round_tf32(v[i].y); l.y = round_tf32(v[i].y - h.y);
        h.z = round_tf32(v[i].z); l.z = round_tf32(v[i].z - h.z);
        h.w = round_tf32(v[i].w); l.w = round_tf32(v[i].w - h.w);
        xs[et + 128 * i] = h;
        xs[et + 128 * i + PLANE / 16] = l;
      }
      fence_proxy_async();                            // generic-proxy stores -> visible to the tensor core's smem reads
      __syncwarp();
      if (lane == 0) mbar_arrive(&conv[stage]);
    }
    // accumulator -> partial sums of K slice z (thread = row; 128 contiguous bytes per 32-column chunk)
    const int q = warp & 3;
    FC_TRACE(4);
    mbar_wait(tmem_full, 0);
    FC_TRACE(5);
    tc_fence_after();
    const int row = m0 + q * 32 + lane;
    float* dst = ws + ((size_t)z * M + row) * N_HIDDEN + n0;
#pragma unroll 1
    for (int ch = 0; ch < BN / 32; ++ch) {
      uint32_t r[32];
      tmem_ld32(tmem_base + ch * 32 + (static_cast<uint32_t>(q * 32) << 16), r);
      tmem_ld_wait();
      if (row < M) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
          *reinterpret_cast<float4*>(dst + ch * 32 + 4 * i) =
              make_float4(__uint_as_float(r[4 * i]), __uint_as_float(r[4 * i + 1]), __uint_as_float(r[4 * i + 2]),
                          __uint_as_float(r[4 * i + 3]));
      }
    }
    tc_fence_before();
    FC_TRACE(6);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, BN);
}

inline size_t ws_bytes(int64_t batch) { return (size_t)NSLICES * batch * N_HIDDEN * sizeof(float); }

// x [batch,4096] fp32; w_hi / w_lo: the tf32 planes of fc1.weight [512,4096]; ws: ws_bytes(batch).  One launch.
inline int launch_fc1_x3_splitk(const float* x, const float* w_hi, const float* w_lo, float* ws, int64_t batch,
                                cudaStream_t s, RingSpan span = RingSpan{}) {
  static bool configured[64] = {};
  int dev = 0;
  B2P_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !configured[dev]) {
    B2P_CUDA(cudaFuncSetAttribute(fc1_x3_splitk, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    configured[dev] = true;
  }
  CUtensorMap tx, th, tl;
  if (span.data) {       // every slot of the ring as one [slots*128, 4096] matrix
    if (span.rows_per_msg != BM || span.slot_bytes != (unsigned long long)BM * K_IN * 4)
      return fail(B2P_EINVAL, "fc1: a ring feeds this kernel in messages of exactly %d rows in %d-byte slots", BM, BM * K_IN * 4);
    B2P_TRY(make_tmap(&tx, 1, span.data, (int64_t)span.slots * BM, K_IN, K_IN, BM));
  } else {
    B2P_TRY(make_tmap(&tx, 1, x, batch, K_IN, K_IN, BM));
  }
  B2P_TRY(make_tmap(&th, 1, w_hi, N_HIDDEN, K_IN, K_IN, BN));
  B2P_TRY(make_tmap(&tl, 1, w_lo, N_HIDDEN, K_IN, K_IN, BN));
  dim3 grid(NSLICES, NTILES, ceil_div(batch, BM));
  static const int trace_on = [] { const char* e = getenv("B2P_FC_TRACE"); return (e && e[0] == '1') ? 1 : 0; }();
  B2P_LAUNCH_PDL(fc1_x3_splitk, grid, THREADS, SMEM_BYTES, s, tx, th, tl, ws, (int)batch, trace_on, span);
  return 0;
}
inline int read_fc_trace(long long* host_out, int n) {
  if (n != NSLICES * NTILES * 16) return fail(B2P_EINVAL, "fc trace: expected %d slots", NSLICES * NTILES * 16);
  B2P_CUDA(cudaDeviceSynchronize());
  B2P_CUDA(cudaMemcpyFromSymbol(host_out, g_fc_trace, sizeof(long long) * NSLICES * NTILES * 16));
  return 0;
}

}  // namespace fcs
}  // namespace b2p
