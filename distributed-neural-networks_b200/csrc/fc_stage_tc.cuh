// fc_stage_tc.cuh -- fc1 of CIFAR part 1 in the tensor-core parity mode (math = tf32x3), split-K 32.
// Two kernels: fc1_f16x3_splitk (split fp16, second half of this file) is what runs; fc1_x3_splitk (3xTF32, round 1,
// described first) stays behind B2P_FC1_TF32=1 for A/B measurements.
//
// Reference: ModelPart1_2Node.forward, cifar_model_parts.py:53-58
//     x = relu(fc1(x)); x = fc2(x); return softmax(x, dim=1)          [B,4096] -> [B,10]
//
// Part 1 used to be three launches per microbatch: split_tf32 (x -> hi/lo planes in HBM), the generic
// linear_tc<tf32x3> with 16 K slices (64 CTAs, each pulling 512 KB through one SM), and fc_tail.  This kernel
// replaces the first two:
//   * persistent CTAs over work items (K slice(s), 128 hidden units, 128 rows): a single 128-row tile already gives
//     128 items, so the 16.8 MB of weight planes and the activation are spread over 128 SMs;
//   * x is loaded ONCE as raw fp32 by TMA.  kind::tf32 reads only the top 19 bits of a word, so the raw tile IS the
//     hi plane; four converter warps write lo = rna_tf32(x - trunc_tf32(x)) beside it (elementwise, so TMA's
//     128-byte swizzle is preserved) -- no split kernel, no hi/lo planes in HBM;
//   * the weight planes of the first item's first three k-blocks are requested before griddepcontrol.wait (they are
//     constants), so most of its bytes are in flight while the previous kernel of the stream drains;
//   * tcgen05.mma kind::tf32, three products per k-step (lo*hi, hi*lo, hi*hi) into a 128 x 128 accumulator stage;
//     partial sums leave as ws[slice or pair][row][512]; fc_tail_f32 adds them pairwise in ascending order.
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

__device__ __forceinline__ float trunc_tf32(float x) { return __uint_as_float(__float_as_uint(x) & 0xffffe000u); }

constexpr int BM = 128, BN = 128;           // rows x hidden units per CTA
constexpr int N_HIDDEN = 512, K_IN = 4096;
constexpr int KB = 32;                      // fp32 per 128-byte swizzle row
constexpr int KSLICE = 128, KBLOCKS = KSLICE / KB;
constexpr int NSLICES = K_IN / KSLICE;      // 32 partial sums per output element
constexpr int NTILES = N_HIDDEN / BN;       // 4
constexpr int PLANE = BM * 128;             // 16 KB: 128 rows x 32 fp32
constexpr int STAGE_BYTES = 4 * PLANE;      // x hi | x lo | W hi | W lo
constexpr int STAGES = 3;
constexpr int EPI_PITCH = 36;               // floats: 32 columns + pad, 16-byte aligned rows
constexpr int EPI_BYTES = 4 * 32 * EPI_PITCH * 4;      // one [32][36] transpose buffer per epilogue warp
constexpr int BAR_BYTES = 256;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + BAR_BYTES + 1024;
constexpr int THREADS = 320;                // TMA warp, MMA warp, 4 converter warps, 4 epilogue warps

// optional timeline of the first work item of each CTA (clock64 at a few points), read back by b2p_debug_fc_trace
__device__ long long g_fc_trace[NSLICES * NTILES * 16];
#define FC_TRACE(ev) do { if (trace_on && lane == 0 && blockIdx.x < NSLICES * NTILES) \
    g_fc_trace[blockIdx.x * 16 + (ev)] = clock64(); } while (0)

// Persistent: CTA c takes work items c, c + gridDim.x, ...; item = (K slice z, hidden tile j, row tile mt) with z
// fastest, so the 128 items of a row tile (= one ring message) are adjacent.  The smem stage ring and the two
// accumulator stages in tensor memory run on across items: the loads and MMAs of item i+1 overlap the epilogue of
// item i, and only the first item of a CTA sees the ~1.2 us TMA start-up latency.
// GROUP = 1: one K slice per item, 32 partial sums per output element (fills the GPU from a single 128-row tile).
// GROUP = 2: an item covers two consecutive K slices, each into its own accumulator stage; the epilogue adds the pair
//            (slice 2g first) and writes 16 partial sums -- half the partial-sum traffic once there are enough row
//            tiles to fill the GPU anyway.  fc_tail adds unpaired partials pairwise in the same order, so both
//            groupings give bit-identical results.
// (Tried and dropped: 256-wide items with the pair's accumulators filling all 512 TMEM columns -- N = 256 MMAs
//  are bit-identical and balance math against operand fetch, but without double buffering across items part 1
//  only went from 36.6 to 35.6 us per 1024 rows.)
template <int GROUP>
__global__ void __launch_bounds__(THREADS, 1)
fc1_x3_splitk(const __grid_constant__ CUtensorMap tmap_x,       // x [M,4096] fp32, box [32 k][128 rows]
              const __grid_constant__ CUtensorMap tmap_w_hi,    // fc1.weight hi plane [512,4096], box [32 k][128 rows]
              const __grid_constant__ CUtensorMap tmap_w_lo,    // fc1.weight lo plane
              float* __restrict__ ws,                           // partial sums [32 / GROUP][M][512]
              int M, int trace_on,
              RingSpan span) {                                  // span.data: tmap_x covers the whole ring, row tile mt is message mt
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  unsigned char* tiles = smem;
  float* epi_smem = reinterpret_cast<float*>(smem + STAGES * STAGE_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES + EPI_BYTES);
  uint64_t* full = bars;                  // [STAGES] TMA -> converters, MMA
  uint64_t* conv = bars + STAGES;         // [STAGES] converters -> MMA
  uint64_t* empty = bars + 2 * STAGES;    // [STAGES] MMA -> TMA
  constexpr int ACC = 2 * GROUP;             // accumulator stages in tensor memory (an item uses GROUP of them)
  constexpr int ZI = NSLICES / GROUP;        // items along K
  uint64_t* tmem_full = bars + 3 * STAGES;   // [ACC] MMA -> epilogue
  uint64_t* tmem_empty = tmem_full + ACC;    // [ACC] epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + ACC);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;      // warp-uniform for the compiler
  const int m_tiles = (M + BM - 1) / BM;
  const int num_items = ZI * NTILES * m_tiles;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_x)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_w_hi)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_w_lo)) : "memory");
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&conv[i], 4); mbar_init(&empty[i], 1); }
    for (int i = 0; i < ACC; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 4); }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, ACC * BN);
    tmem_relinquish();
  }
  FC_TRACE(0);
  pdl_trigger();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    // fc1.weight is constant: the first item's first STAGES k-blocks are requested BEFORE griddepcontrol.wait, so
    // most of that item's bytes are on their way while the previous kernel in the stream drains.
    const int first = blockIdx.x;
    if (lane == 0 && first < num_items) {
      const int zi = first % ZI, j = (first / ZI) % NTILES;
#pragma unroll
      for (int kb = 0; kb < STAGES; ++kb) {          // (KBLOCKS * GROUP >= STAGES)
        unsigned char* sa = tiles + kb * STAGE_BYTES;
        mbar_expect_tx(&full[kb], 3 * PLANE);
        tma_load_2d(sa + 2 * PLANE, &tmap_w_hi, &full[kb], zi * GROUP * KSLICE + kb * KB, j * BN);
        tma_load_2d(sa + 3 * PLANE, &tmap_w_lo, &full[kb], zi * GROUP * KSLICE + kb * KB, j * BN);
      }
    }
    __syncwarp();
    pdl_wait();          // x belongs to the previous kernel until here
    uint32_t stage = 0, phase = 0;
    int waited_mt = -1, x_row0 = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
      const int zi = item % ZI, j = (item / ZI) % NTILES, mt = item / (ZI * NTILES);
      if (mt != waited_mt) {
        x_row0 = mt * BM;
        if (span.data) {
          // row tile mt is message mt of this launch: wait until the producer GPU has published it
          const int slot = span_slot(span, mt);
          x_row0 = slot * BM;              // the ring's slots are 128 rows each, back to back (checked by the host)
          if (lane == 0) {
            spin_until_ge(span.published + slot, ld_count(span.count + slot) + 1);     // acquire: the peer's stores are visible ...
            fence_proxy_async();           // ... also to the TMA unit (async proxy)
          }
          __syncwarp();
        }
        waited_mt = mt;
      }
      for (int kb = 0; kb < KBLOCKS * GROUP; ++kb) {          // the item's K range is contiguous: GROUP slices back to back
        mbar_wait(&empty[stage], phase ^ 1);
        if (lane == 0) {
          unsigned char* sa = tiles + stage * STAGE_BYTES;
          const int kx = zi * GROUP * KSLICE + kb * KB;
          if (item != first || kb >= STAGES) {
            mbar_expect_tx(&full[stage], 3 * PLANE);
            tma_load_2d(sa + 2 * PLANE, &tmap_w_hi, &full[stage], kx, j * BN);
            tma_load_2d(sa + 3 * PLANE, &tmap_w_lo, &full[stage], kx, j * BN);
          }
          tma_load_2d(sa, &tmap_x, &full[stage], kx, x_row0);      // raw fp32, becomes the hi plane
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // all 32 lanes in uniform control flow, the tensor-core instructions predicated on elect.sync (tc_gemm.cuh)
    constexpr uint32_t idesc = make_idesc(1, BM, BN);
    const uint32_t lead = elect_one();
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t tiles_u = __shfl_sync(0xffffffffu, smem_u32(tiles), 0);
    uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x)
    for (int sl = 0; sl < GROUP; ++sl) {             // one accumulator stage per K slice of the item
      mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_u + acc * BN;
      for (int kb = 0; kb < KBLOCKS; ++kb) {
        mbar_wait(&full[stage], phase);
        mbar_wait(&conv[stage], phase);
        tc_fence_after();
        const uint32_t base = tiles_u + stage * STAGE_BYTES;
        const uint64_t xa = make_desc_sw128(base), xl = make_desc_sw128(base + PLANE);
        const uint64_t wh = make_desc_sw128(base + 2 * PLANE), wl = make_desc_sw128(base + 3 * PLANE);
#pragma unroll
        for (int k = 0; k < KB / 8; ++k) {              // tf32 k-step = 8 elements = 32 B = +2 in the descriptor
          umma_p<1>(tmem_d, xl + 2 * k, wh + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u, lead);      // small terms first
          umma_p<1>(tmem_d, xa + 2 * k, wl + 2 * k, idesc, 1u, lead);
          umma_p<1>(tmem_d, xa + 2 * k, wh + 2 * k, idesc, 1u, lead);
        }
        umma_commit_p(&empty[stage], lead);
        if (kb == KBLOCKS - 1) umma_commit_p(&tmem_full[acc], lead);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if (++acc == ACC) { acc = 0; acc_phase ^= 1; }
    }
  } else if (warp < 6) {
    // ===================== converters (warps 2..5) =====================
    const int ct = threadIdx.x - 64;                  // 0..127
    uint32_t stage = 0, phase = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
      for (int kb = 0; kb < KBLOCKS * GROUP; ++kb) {
        mbar_wait(&full[stage], phase);
        if (item == (int)blockIdx.x && kb == 0 && warp == 2) FC_TRACE(1);
        float4* xs = reinterpret_cast<float4*>(tiles + stage * STAGE_BYTES);
        float4 v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = xs[ct + 128 * i];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          // kind::tf32 reads the top 19 bits of each fp32 word, so the raw tile already IS the hi plane
          // (hi = x with the low 13 mantissa bits dropped); only lo = rna_tf32(x - hi) has to be written.
          float4 l;
          l.x = round_tf32(v[i].x - trunc_tf32(v[i].x));
          l.y = round_tf32(v[i].y - trunc_tf32(v[i].y));
          l.z = round_tf32(v[i].z - trunc_tf32(v[i].z));
          l.w = round_tf32(v[i].w - trunc_tf32(v[i].w));
          xs[ct + 128 * i + PLANE / 16] = l;
        }
        fence_proxy_async();                            // generic-proxy stores -> visible to the tensor core's smem reads
        __syncwarp();
        if (lane == 0) mbar_arrive(&conv[stage]);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if (item == (int)blockIdx.x && warp == 2) FC_TRACE(2);
    }
  } else {
    // ===================== epilogue (warps 6..9) =====================
    // accumulator -> partial sums of K slice z.  tcgen05.ld gives a thread one ROW (32 consecutive columns); a
    // per-warp [32][36] shared-memory transpose turns that into stores where 8 lanes write one row's 128 bytes.
    pdl_wait();          // ws is still being read by the previous microbatch's tail until here
    const int q = warp & 3;                           // TMEM lane quarter this warp may read
    float* stg = epi_smem + (warp - 6) * (32 * EPI_PITCH);
    uint32_t acc = 0, acc_phase = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
      const int zi = item % ZI, j = (item / ZI) % NTILES, mt = item / (ZI * NTILES);
      mbar_wait(&tmem_full[acc], acc_phase);
      if (GROUP == 2) mbar_wait(&tmem_full[acc + 1], acc_phase);      // (ACC is a multiple of GROUP: same phase)
      tc_fence_after();
      if (item == (int)blockIdx.x && warp == 6) FC_TRACE(3);
      const int row0 = mt * BM + q * 32;
      float* dst = ws + ((size_t)zi * M + row0) * N_HIDDEN + j * BN;
#pragma unroll 1
      for (int ch = 0; ch < BN / 32; ++ch) {
        uint32_t r[32];
        tmem_ld32(tmem_base + acc * BN + ch * 32 + (static_cast<uint32_t>(q * 32) << 16), r);
        if (GROUP == 2) {
          uint32_t r2[32];
          tmem_ld32(tmem_base + (acc + 1) * BN + ch * 32 + (static_cast<uint32_t>(q * 32) << 16), r2);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) + __uint_as_float(r2[i]));   // slice 2g + slice 2g+1
        } else {
          tmem_ld_wait();
        }
#pragma unroll
        for (int i = 0; i < 8; ++i)
          *reinterpret_cast<float4*>(stg + lane * EPI_PITCH + 4 * i) =
              make_float4(__uint_as_float(r[4 * i]), __uint_as_float(r[4 * i + 1]), __uint_as_float(r[4 * i + 2]),
                          __uint_as_float(r[4 * i + 3]));
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int rr = 4 * i + (lane >> 3), c4 = (lane & 7) * 4;
          const float4 v = *reinterpret_cast<const float4*>(stg + rr * EPI_PITCH + c4);
          if (row0 + rr < M) *reinterpret_cast<float4*>(dst + (size_t)rr * N_HIDDEN + ch * 32 + c4) = v;
        }
        __syncwarp();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&tmem_empty[acc]);
        if (GROUP == 2) mbar_arrive(&tmem_empty[acc + 1]);
      }
      if (item == (int)blockIdx.x && warp == 6) FC_TRACE(4);
      acc += GROUP;
      if (acc == ACC) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, ACC * BN);
}

inline size_t ws_bytes(int64_t batch) { return (size_t)NSLICES * batch * N_HIDDEN * sizeof(float); }
// rows from which pairs of K slices still give every SM an item (16 x 4 items per 128-row tile)
inline int fc1_group_for(int64_t batch) { return ceil_div(batch, BM) >= 3 ? 2 : 1; }

// x [batch,4096] fp32; w_hi / w_lo: the tf32 planes of fc1.weight [512,4096]; ws: ws_bytes(batch).  One launch;
// returns the number of partial sums per output element it wrote (32 or 16) or a negative error.
inline int launch_fc1_x3_splitk(const float* x, const float* w_hi, const float* w_lo, float* ws, int64_t batch,
                                cudaStream_t s, RingSpan span = RingSpan{}) {
  static bool configured[64] = {};
  int dev = 0;
  B2P_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !configured[dev]) {
    B2P_CUDA(cudaFuncSetAttribute(fc1_x3_splitk<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    B2P_CUDA(cudaFuncSetAttribute(fc1_x3_splitk<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
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
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int group = fc1_group_for(batch);
  const int grid = std::min((NSLICES / group) * NTILES * ceil_div(batch, BM), sms);
  static const int trace_on = [] { const char* e = getenv("B2P_FC_TRACE"); return (e && e[0] == '1') ? 1 : 0; }();
  if (group == 2)
    B2P_LAUNCH_PDL(fc1_x3_splitk<2>, grid, THREADS, SMEM_BYTES, s, tx, th, tl, ws, (int)batch, trace_on, span);
  else
    B2P_LAUNCH_PDL(fc1_x3_splitk<1>, grid, THREADS, SMEM_BYTES, s, tx, th, tl, ws, (int)batch, trace_on, span);
  return NSLICES / group;
}
inline int read_fc_trace(long long* host_out, int n) {
  if (n != NSLICES * NTILES * 16) return fail(B2P_EINVAL, "fc trace: expected %d slots", NSLICES * NTILES * 16);
  B2P_CUDA(cudaDeviceSynchronize());
  B2P_CUDA(cudaMemcpyFromSymbol(host_out, g_fc_trace, sizeof(long long) * NSLICES * NTILES * 16));
  return 0;
}


// =====================================================================================================================
// fc1 in split fp16 (the scheme both convolutions of part 0 use, cifar_tc2.cuh): x ~= hi + lo * 2^-11 with fp16 hi / lo,
// products exact in the fp32 accumulator, lo*lo dropped.  Against the 3xTF32 kernel above:
//   * kind::f16 runs at twice the tf32 rate and a 128-byte swizzle row holds 64 k instead of 32: per 128 x 128 x 64
//     block the tensor core reads 96 KB of operands instead of 192 KB -- the 3xTF32 kernel was bound by exactly that
//     shared-memory traffic (352 KB per block, 28 us per 1024 rows);
//   * the weights travel as fp16 planes: 8.4 MB per pass instead of 16.8 MB;
//   * x never sits in shared memory as fp32: eight converter warps load it from global memory (L2: the stage-0 kernel
//     has just written it) into registers, one k-block ahead, and store the two fp16 tiles in the 128B-swizzle layout;
//   * the lo terms carry a factor 2^11, so they accumulate in their own tensor-memory stage (main | cross) and the
//     epilogue forms main + cross * 2^-11;
//   * the epilogue adds the K slices of an item in REGISTERS (fp32, round to nearest; the tensor core's own fp32
//     accumulation truncates, which is why a slice stays 128 long) and writes one partial sum per item: with four slices
//     per item 8 partials per output element instead of 16 -- half the traffic of the tail.
// Summation order of an output element: slices ((s0 + s1) + s2) + s3 inside an item, items in ascending order in
// fc_tail_f32; with one slice per item (small batches: enough items to fill the GPU) the tail adds groups of four
// first -- both give the same bits, so micro-batching stays bit-exact.
// Range: |x| and |W| below 65504 (fp16); beyond it the result is not finite where the reference's is: the converters
// and the weight repack set the stage's range flag (b2p_stage_check -> B2P_ELIMIT).
constexpr int KB16 = 64;                                 // fp16 per 128-byte swizzle row
constexpr int KBLOCKS16 = KSLICE / KB16;                 // 2 k-blocks per K slice
constexpr int STAGES16 = 3;                              // x hi | x lo | W hi | W lo, 16 KB each (STAGE_BYTES)
constexpr int CONV_WARPS = 8, EPI_WARPS16 = 8;
constexpr int THREADS16 = (2 + CONV_WARPS + EPI_WARPS16) * 32;      // 576: TMA, MMA, 8 converter, 8 epilogue warps
constexpr int EPI16_PITCH = 20;                          // floats: 16 columns + pad, 16-byte aligned rows
constexpr int EPI16_BYTES = EPI_WARPS16 * 32 * EPI16_PITCH * 4;     // one [32][20] transpose buffer per epilogue warp
constexpr int SMEM16_BYTES = STAGES16 * STAGE_BYTES + EPI16_BYTES + BAR_BYTES + 1024;
static_assert(SMEM16_BYTES <= 232448, "shared memory budget of one CTA on sm_100");

// (a, b) -> fp16x2 hi parts and fp16x2 lo parts, lo = (x - hi) * 2^11
__device__ __forceinline__ void split2h(float a, float b, uint32_t& hi, uint32_t& lo) {
  const __half2 h = __floats2half2_rn(a, b);
  const float2 hf = __half22float2(h);
  const __half2 l = __floats2half2_rn((a - hf.x) * 2048.0f, (b - hf.y) * 2048.0f);
  hi = *reinterpret_cast<const uint32_t*>(&h);
  lo = *reinterpret_cast<const uint32_t*>(&l);
}
// one-off repack of fc1.weight (b2p_stage_finalize): fp32 [512,4096] -> fp16 hi plane and fp16 lo plane (x 2^11)
__global__ void split_f16_planes(const float* __restrict__ w, __half* __restrict__ hi, __half* __restrict__ lo, size_t n,
                                 int* __restrict__ range_flag) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float v = w[i];
    const __half h = __float2half_rn(v);
    hi[i] = h;
    lo[i] = __float2half_rn((v - __half2float(h)) * 2048.0f);
    if (fabsf(v) >= 65504.f && range_flag) *range_flag = 1;
  }
}

__device__ __forceinline__ void tmem_ld16f(uint32_t taddr, uint32_t (&r)[16]) {      // 32 lanes x 16 consecutive columns
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__host__ __device__ constexpr uint32_t idesc_f16_f32acc(int m, int n) {   // kind::f16, A/B = F16 (format 0), K-major, fp32 accumulate
  return (1u << 4) | (static_cast<uint32_t>(n >> 3) << 17) | (static_cast<uint32_t>(m >> 4) << 24);
}

// Persistent: CTA c takes work items c, c + gridDim.x, ...; item = (S consecutive K slices zi, hidden tile j, row
// tile mt) with zi fastest, so the items of a row tile (= one ring message) are adjacent.
__global__ void __launch_bounds__(THREADS16, 1)
fc1_f16x3_splitk(const float* __restrict__ x,                     // [M,4096] fp32 (unused when span.data)
                 const __grid_constant__ CUtensorMap tmap_w_hi,   // fc1.weight fp16 hi plane [512,4096], box [64 k][128 rows]
                 const __grid_constant__ CUtensorMap tmap_w_lo,   // fp16 lo plane (x 2^11)
                 float* __restrict__ ws,                          // partial sums [32 / S][M][512]
                 int M, int S,                                    // S = K slices per item (1 or 4)
                 int* __restrict__ range_flag,                    // set when |x| reaches the fp16 range limit (b2p_stage_check)
                 RingSpan span) {                                 // span.data: row tile mt is message mt of a handoff ring
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  unsigned char* tiles = smem;
  float* epi_smem = reinterpret_cast<float*>(smem + STAGES16 * STAGE_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES16 * STAGE_BYTES + EPI16_BYTES);
  uint64_t* full = bars;                    // [STAGES16] TMA (W planes) -> MMA
  uint64_t* conv = bars + STAGES16;         // [STAGES16] converters (x planes) -> MMA
  uint64_t* empty = bars + 2 * STAGES16;    // [STAGES16] MMA -> TMA and converters
  uint64_t* tmem_full = bars + 3 * STAGES16;   // [2] MMA -> epilogue   (stage a: main at column 256 a, cross at 256 a + 128)
  uint64_t* tmem_empty = tmem_full + 2;        // [2] epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;      // warp-uniform for the compiler
  const int m_tiles = (M + BM - 1) / BM;
  const int ZI = NSLICES / S;
  const int num_items = ZI * NTILES * m_tiles;
  const int nkb = S * KBLOCKS16;                   // k-blocks per item

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_w_hi)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_w_lo)) : "memory");
    for (int i = 0; i < STAGES16; ++i) { mbar_init(&full[i], 1); mbar_init(&conv[i], CONV_WARPS / 2); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], EPI_WARPS16); }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  pdl_trigger();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer: the weight planes (constants: no griddepcontrol.wait) =====================
    uint32_t stage = 0, phase = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
      const int zi = item % ZI, j = (item / ZI) % NTILES;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(&empty[stage], phase ^ 1);
        if (lane == 0) {
          unsigned char* sa = tiles + stage * STAGE_BYTES;
          const int kx = zi * S * KSLICE + kb * KB16;
          mbar_expect_tx(&full[stage], 2 * PLANE);
          tma_load_2d(sa + 2 * PLANE, &tmap_w_hi, &full[stage], kx, j * BN);
          tma_load_2d(sa + 3 * PLANE, &tmap_w_lo, &full[stage], kx, j * BN);
        }
        __syncwarp();
        if (++stage == STAGES16) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (uniform control flow, elect.sync predicate: tc_gemm.cuh) =====================
    constexpr uint32_t idesc = idesc_f16_f32acc(BM, BN);
    const uint32_t lead = elect_one();
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t tiles_u = __shfl_sync(0xffffffffu, smem_u32(tiles), 0);
    uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x)
    for (int sl = 0; sl < S; ++sl) {                 // one (main | cross) accumulator stage per K slice
      mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t t_main = tmem_u + acc * 256, t_cross = t_main + 128;
      for (int kb = 0; kb < KBLOCKS16; ++kb) {
        mbar_wait(&full[stage], phase);
        mbar_wait(&conv[stage], phase);
        tc_fence_after();
        const uint32_t base = tiles_u + stage * STAGE_BYTES;
        const uint64_t xh = make_desc_sw128(base), xl = make_desc_sw128(base + PLANE);
        const uint64_t wh = make_desc_sw128(base + 2 * PLANE), wl = make_desc_sw128(base + 3 * PLANE);
#pragma unroll
        for (int k = 0; k < KB16 / 16; ++k) {           // fp16 k-step = 16 elements = 32 B = +2 in the descriptor
          umma_p<0>(t_cross, xl + 2 * k, wh + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u, lead);
          umma_p<0>(t_cross, xh + 2 * k, wl + 2 * k, idesc, 1u, lead);
          umma_p<0>(t_main, xh + 2 * k, wh + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u, lead);
        }
        umma_commit_p(&empty[stage], lead);
        if (kb == KBLOCKS16 - 1) umma_commit_p(&tmem_full[acc], lead);
        if (++stage == STAGES16) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else if (warp < 2 + CONV_WARPS) {
    // ===================== converters (warps 2..9): x fp32 from global -> fp16 hi / lo tiles =====================
    // Two groups of four warps take the k-blocks in turn, so two blocks' loads (L2 latency) are always in flight
    // while the other group converts.  (One group with a one-block look-ahead still had only one block's loads in
    // flight at a time: 2.5 k cycles per k-block against 1 k of tensor-core time.)
    // thread = float4 column c4 (4 k) of rows r0, r0 + 8, ...: a warp reads two rows' 256 contiguous bytes per load
    // and the swizzle term (row & 7) is the same for all of a thread's rows.
    const int g = (warp - 2) >> 2;                    // group 0 / 1
    const int gt = threadIdx.x - 64 - 128 * g;        // 0..127 within the group
    const int c4 = gt & 15, r0 = gt >> 4;
    const uint32_t off0 = (uint32_t)r0 * 128u + ((((uint32_t)c4 >> 1) ^ ((uint32_t)r0 & 7u)) << 4) + ((uint32_t)c4 & 1u) * 8u;
    pdl_wait();          // x belongs to the previous kernel until here
    int waited_mt = -1;
    const float* tile_base = nullptr;
    int tile_rows = 0;
    float in_max = 0.f;
    int item = blockIdx.x, kb = g;                    // this group's first block: t = g
    if (kb >= nkb) { kb -= nkb; item += gridDim.x; }  // (nkb >= 2 always)
    for (uint32_t t = g; item < num_items; t += 2) {
      const int zi = item % ZI, mt = item / (ZI * NTILES);
      if (mt != waited_mt) {                          // (same control flow in all 128 threads of the group)
        tile_base = x + (size_t)mt * BM * K_IN;
        if (span.data) {
          // row tile mt is message mt of this launch: wait until the producer GPU has published it
          const int slot = span_slot(span, mt);
          tile_base = reinterpret_cast<const float*>(span.data + (size_t)slot * span.slot_bytes);
          if (gt == 0) spin_until_ge(span.published + slot, ld_count(span.count + slot) + 1);      // acquire: the peer's stores are visible
          asm volatile("bar.sync %0, 128;" ::"r"(1 + g) : "memory");
        }
        tile_rows = min(BM, M - mt * BM);
        waited_mt = mt;
      }
      const float* p = tile_base + (size_t)r0 * K_IN + zi * S * KSLICE + kb * KB16 + c4 * 4;
      float4 v[16];
#pragma unroll
      for (int i = 0; i < 16; ++i)
        v[i] = (r0 + 8 * i < tile_rows) ? ld_produced(reinterpret_cast<const float4*>(p + (size_t)8 * i * K_IN)) : make_float4(0.f, 0.f, 0.f, 0.f);
      const uint32_t stage = t % STAGES16, phase = (t / STAGES16) & 1u;
      mbar_wait(&empty[stage], phase ^ 1);            // the MMAs that read this stage's previous tiles are done
      unsigned char* hi_t = tiles + stage * STAGE_BYTES + off0;
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        uint32_t h01, l01, h23, l23;
        split2h(v[i].x, v[i].y, h01, l01);
        split2h(v[i].z, v[i].w, h23, l23);
        *reinterpret_cast<uint2*>(hi_t + i * 8 * 128) = make_uint2(h01, h23);
        *reinterpret_cast<uint2*>(hi_t + PLANE + i * 8 * 128) = make_uint2(l01, l23);
        in_max = fmaxf(in_max, fmaxf(fmaxf(fabsf(v[i].x), fabsf(v[i].y)), fmaxf(fabsf(v[i].z), fabsf(v[i].w))));
      }
      fence_proxy_async();                            // generic-proxy stores -> visible to the tensor core's smem reads
      __syncwarp();
      if (lane == 0) mbar_arrive(&conv[stage]);
      kb += 2;                                        // the group's next block
      if (kb >= nkb) { kb -= nkb; item += gridDim.x; }
    }
    if (in_max >= 65504.f && range_flag) *range_flag = 1;      // (NaN input: not flagged, it is data)
  } else {
    // ===================== epilogue (warps 10..17) =====================
    // warp (q, half): TMEM lanes 32 q .. +31 (rows), columns 64 half .. +63 of the 128-wide tile.  The item's K slices
    // are added in registers; the sum leaves through a per-warp [32][20] transpose as 64-byte row segments.
    pdl_wait();          // ws is still being read by the previous step's tail until here
    const int q = warp & 3, half = (warp - (2 + CONV_WARPS)) >> 2;
    float* stg = epi_smem + (warp - (2 + CONV_WARPS)) * (32 * EPI16_PITCH);
    const uint32_t t_warp = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + half * 64;
    uint32_t acc = 0, acc_phase = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
      const int zi = item % ZI, j = (item / ZI) % NTILES, mt = item / (ZI * NTILES);
      float sum[64];
#pragma unroll
      for (int i = 0; i < 64; ++i) sum[i] = 0.f;
      for (int sl = 0; sl < S; ++sl) {
        mbar_wait(&tmem_full[acc], acc_phase);
        tc_fence_after();
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          uint32_t rm[16], rc[16];
          tmem_ld16f(t_warp + acc * 256 + ch * 16, rm);
          tmem_ld16f(t_warp + acc * 256 + 128 + ch * 16, rc);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i)
            sum[ch * 16 + i] += fmaf(__uint_as_float(rc[i]), 1.0f / 2048.0f, __uint_as_float(rm[i]));
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
      const int row0 = mt * BM + q * 32;
      float* dst = ws + ((size_t)zi * M + row0) * N_HIDDEN + j * BN + half * 64;
#pragma unroll
      for (int ch = 0; ch < 4; ++ch) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
          *reinterpret_cast<float4*>(stg + lane * EPI16_PITCH + 4 * i) =
              make_float4(sum[ch * 16 + 4 * i], sum[ch * 16 + 4 * i + 1], sum[ch * 16 + 4 * i + 2], sum[ch * 16 + 4 * i + 3]);
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int rr = 8 * i + (lane >> 2), cc = (lane & 3) * 4;
          const float4 v = *reinterpret_cast<const float4*>(stg + rr * EPI16_PITCH + cc);
          if (row0 + rr < M) *reinterpret_cast<float4*>(dst + (size_t)rr * N_HIDDEN + ch * 16 + cc) = v;
        }
        __syncwarp();
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// rows from which four K slices per item still give every SM work (8 x 4 items per 128-row tile)
inline int fc1_slices_per_item(int64_t batch) { return ceil_div(batch, BM) >= 3 ? 4 : 1; }

// x [batch,4096] fp32; w_hi / w_lo: the fp16 planes of fc1.weight [512,4096] (split_f16_planes); ws: ws_bytes(batch).
// One launch; returns the number of K slices per partial sum it wrote (4 or 1) or a negative error.
inline int launch_fc1_f16x3_splitk(const float* x, const __half* w_hi, const __half* w_lo, float* ws, int64_t batch,
                                   int* range_flag, cudaStream_t s, RingSpan span = RingSpan{}) {
  static bool configured[64] = {};
  int dev = 0;
  B2P_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !configured[dev]) {
    B2P_CUDA(cudaFuncSetAttribute(fc1_f16x3_splitk, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM16_BYTES));
    configured[dev] = true;
  }
  if (span.data && (span.rows_per_msg != BM || span.slot_bytes != (unsigned long long)BM * K_IN * 4))
    return fail(B2P_EINVAL, "fc1: a ring feeds this kernel in messages of exactly %d rows in %d-byte slots", BM, BM * K_IN * 4);
  if (!span.data && (reinterpret_cast<uintptr_t>(x) & 15))
    return fail(B2P_EINVAL, "fc1: the activation must be 16-byte aligned");
  CUtensorMap th, tl;
  B2P_TRY(make_tmap(&th, 0, w_hi, N_HIDDEN, K_IN, K_IN, BN));
  B2P_TRY(make_tmap(&tl, 0, w_lo, N_HIDDEN, K_IN, K_IN, BN));
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int S = fc1_slices_per_item(batch);
  const int grid = std::min((NSLICES / S) * NTILES * ceil_div(batch, BM), sms);
  B2P_LAUNCH_PDL(fc1_f16x3_splitk, grid, THREADS16, SMEM16_BYTES, s, x, th, tl, ws, (int)batch, S, range_flag, span);
  return S;
}
}  // namespace fcs
}  // namespace b2p
