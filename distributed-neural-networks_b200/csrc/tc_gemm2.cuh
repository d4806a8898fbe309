// tc_gemm2.cuh -- the bf16 linear with CTA pairs (tcgen05 cta_group::2).
//
// Same contraction and epilogue as tc_gemm.cuh, but two CTAs on the two SMs of a TPC form a
// cluster and share one 256 x 256 output tile: each CTA loads ITS 128 rows of A and ITS 128 of the
// tile's 256 W rows; the leader CTA issues tcgen05.mma.cta_group::2 (M = 256), which feeds both
// SMs' tensor cores from both SMs' shared memory; each CTA ends up with its 128 x 256 half of the
// accumulator in its own tensor memory and runs its own epilogue.
//
// Why: with one CTA per tile the tensor core is fed at ~72 B/clk of shared-memory operand reads
// (measured), so a 128x256x16 MMA (12 KB of operands) takes ~160 cycles instead of the 128 it
// needs -- and every k-block costs 48 KB of L2->smem traffic per SM.  In pair mode each SM reads
// 8 KB per MMA step (its A half + its W half, W is broadcast to the partner) and loads 32 KB per
// k-block: the MMA becomes tensor-bound and the L2 traffic per FLOP drops by a third.
//
// Protocol (per smem stage s, per accumulator stage a):
//   full[s]      lives in the LEADER; both CTAs' TMA loads complete_tx on it (cta_group::2 loads
//                with the barrier address' peer bit cleared), the leader's producer arms it.
//   empty[s]     one per CTA; tcgen05.commit multicasts the "MMAs retired" arrival to both.
//   tmem_full[a] one per CTA (commit multicast); tmem_empty[a] lives in the leader and counts the
//                epilogue warps of BOTH CTAs (the partner arrives remotely).
#pragma once
#include "tc_gemm.cuh"

namespace b2p {
namespace tc {

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t local_smem_addr, uint32_t rank) {
  uint32_t a;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(a) : "r"(local_smem_addr), "r"(rank));
  return a;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;     // clears the CTA-pair bit of a shared::cluster address -> even CTA
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* map, uint64_t* leader_bar, int c0,
                                                 int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(leader_bar) & kPeerBitMask), "r"(c0),
      "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d_pair_p(uint32_t smem_dst, const CUtensorMap* map, uint32_t leader_bar_addr, int c0,
                                                   int c1, uint32_t lead) {      // from warp-uniform code (tc_gemm.cuh, elect_one)
  asm volatile(
      "{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %5, 0;\n\t"
      "@q cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n\t}"
      ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(leader_bar_addr & kPeerBitMask), "r"(c0), "r"(c1), "r"(lead)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish_pair() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t addr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_pair_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                               uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// the same two, issued from warp-uniform code and predicated on elect.sync (tc_gemm.cuh, elect_one)
__device__ __forceinline__ void umma_pair_bf16_p(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                                 uint32_t accumulate, uint32_t lead) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\tsetp.ne.b32 q, %5, 0;\n\t"
      "@q tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(lead)
      : "memory");
}
__device__ __forceinline__ void umma_commit_pair_p(uint64_t* bar, uint32_t lead) {
  asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %2, 0;\n\t"
               "@q tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n\t}"
               ::"r"(smem_u32(bar)), "h"(static_cast<uint16_t>(3)), "r"(lead)
               : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {     // arrives on `bar` in BOTH CTAs of the pair
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"(static_cast<uint16_t>(3))
               : "memory");
}

struct PairCfg {
  static constexpr int BN = 256;                    // tile columns
  static constexpr int BM_CTA = 128, BM_PAIR = 256;
  static constexpr int BK = 64, UMMA_K = 16;
  static constexpr int A_BYTES = BM_CTA * 128;      // this CTA's 128 rows of A
  static constexpr int B_BYTES = (BN / 2) * 128;    // this CTA's 128 of the tile's 256 W rows
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = 6;
  static constexpr int EPI_PITCH = 20;
  static constexpr int EPI_WARP_FLOATS = 32 * EPI_PITCH + BN / 2;
  static constexpr int EPI_BYTES = 8 * EPI_WARP_FLOATS * 4;
  static constexpr int BAR_BYTES = 256;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + BAR_BYTES + 1024;
  static constexpr int TMEM_COLS = 2 * BN;          // two accumulator stages of 256 columns
  static constexpr int THREADS = 320;
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(320, 1)
linear_tc_pair(const __grid_constant__ CUtensorMap tmap_a,      // box [64 K, 128 rows]
               const __grid_constant__ CUtensorMap tmap_w,      // box [64 K, 128 rows]
               EpiParams ep, int M, int N, int K) {
  using Cfg = PairCfg;
  constexpr int BN = Cfg::BN;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  unsigned char* tiles = smem;
  float* epi_smem = reinterpret_cast<float*>(smem + Cfg::STAGES * Cfg::STAGE_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::STAGES * Cfg::STAGE_BYTES + Cfg::EPI_BYTES);
  uint64_t* full = bars;                         // [STAGES]  (used in the leader)
  uint64_t* empty = bars + Cfg::STAGES;          // [STAGES]  (one per CTA)
  uint64_t* tmem_full = bars + 2 * Cfg::STAGES;  // [2]       (one per CTA)
  uint64_t* tmem_empty = tmem_full + 2;          // [2]       (used in the leader)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;     // warp index: uniform for the compiler
  const uint32_t rank = cluster_ctarank();       // 0 = leader
  const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  const int tiles_m = (M + Cfg::BM_PAIR - 1) / Cfg::BM_PAIR, tiles_n = (N + BN - 1) / BN;
  const int num_tiles = tiles_m * tiles_n;
  const int num_kb = (K + Cfg::BK - 1) / Cfg::BK;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_a)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_w)) : "memory");
    for (int i = 0; i < Cfg::STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 16); }   // 8 warps x 2 CTAs
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc_pair(tmem_slot, Cfg::TMEM_COLS);
    tmem_relinquish_pair();
  }
  pdl_trigger();
  tc_fence_before();
  cluster_sync_all();                            // both CTAs' barriers exist before anyone signals across the pair
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    const uint32_t lead = elect_one();                       // uniform control flow, loads predicated on elect.sync
    const uint32_t tiles_u = __shfl_sync(0xffffffffu, smem_u32(tiles), 0);
    const uint32_t full_u = __shfl_sync(0xffffffffu, smem_u32(full), 0);
    uint32_t stage = 0, phase = 0;
    for (int tile = pair; tile < num_tiles; tile += n_pairs) {
      const int m0 = (tile % tiles_m) * Cfg::BM_PAIR + rank * Cfg::BM_CTA;
      const int n0 = (tile / tiles_m) * BN + rank * (BN / 2);
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&empty[stage], phase ^ 1);
        const uint32_t sa = tiles_u + stage * Cfg::STAGE_BYTES, fb = full_u + stage * 8;
        if (rank == 0) mbar_expect_tx_p(fb, 2 * Cfg::STAGE_BYTES, lead);          // both CTAs' bytes land here
        tma_load_2d_pair_p(sa, &tmap_a, fb, kb * Cfg::BK, m0, lead);
        tma_load_2d_pair_p(sa + Cfg::A_BYTES, &tmap_w, fb, kb * Cfg::BK, n0, lead);
        if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader only) =====================
    if (rank == 0) {
      constexpr uint32_t idesc = make_idesc(0, Cfg::BM_PAIR, BN);
      const uint32_t lead = elect_one();
      const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
      const uint32_t tiles_u = __shfl_sync(0xffffffffu, smem_u32(tiles), 0);
      uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
      for (int tile = pair; tile < num_tiles; tile += n_pairs) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_u + acc * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = tiles_u + stage * Cfg::STAGE_BYTES;
          const uint64_t da = make_desc_sw128(a_addr), db = make_desc_sw128(a_addr + Cfg::A_BYTES);
#pragma unroll
          for (int k = 0; k < Cfg::BK / Cfg::UMMA_K; ++k)
            umma_pair_bf16_p(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u, lead);
          umma_commit_pair_p(&empty[stage], lead);                 // frees the slot in both CTAs
          if (kb == num_kb - 1) umma_commit_pair_p(&tmem_full[acc], lead);
          if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue (warps 2..9, both CTAs) =====================
    constexpr int P = Cfg::EPI_PITCH;
    const int q = warp & 3, ew = warp - 2, col_half = ew >> 2;
    float* stg = epi_smem + ew * Cfg::EPI_WARP_FLOATS;
    float* bias_s = stg + 32 * P;
    const bool vec_f32 = !ep.out_bf16 && (ep.ldo & 3) == 0 && (reinterpret_cast<uintptr_t>(ep.out) & 15) == 0 &&
                         (!ep.res || ((ep.ld_res & 3) == 0 && (reinterpret_cast<uintptr_t>(ep.res) & 15) == 0));
    const bool vec_bf16 = ep.out_bf16 && (ep.ldo & 7) == 0 && (reinterpret_cast<uintptr_t>(ep.out) & 15) == 0 && !ep.res;
    const uint32_t leader_empty[2] = {map_to_cta(smem_u32(&tmem_empty[0]), 0), map_to_cta(smem_u32(&tmem_empty[1]), 0)};
    uint32_t acc = 0, acc_phase = 0;
    for (int tile = pair; tile < num_tiles; tile += n_pairs) {
      const int m0 = (tile % tiles_m) * Cfg::BM_PAIR + rank * Cfg::BM_CTA;     // this CTA's rows
      const int n0 = (tile / tiles_m) * BN;                                     // all 256 columns of the tile
      if (ep.bias) {
        const int cbase = n0 + col_half * (BN / 2);
#pragma unroll
        for (int i = lane; i < BN / 2; i += 32) bias_s[i] = (cbase + i < N) ? __ldg(ep.bias + cbase + i) : 0.f;
        __syncwarp();
      }
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      epilogue_tile<BN, P>(ep, stg, bias_s, tmem_base + acc * BN, m0, n0, M, N, q, col_half, lane, 0, (tile / tiles_m) * 2,
                           vec_f32, vec_bf16);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(leader_empty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  cluster_sync_all();                            // nobody leaves while the partner may still touch its smem / barriers
  if (warp == 1) tmem_dealloc_pair(tmem_base, Cfg::TMEM_COLS);
}

// bf16 A [M,K], W [N,K]; returns 0 or a negative error.  The caller decides when the pair kernel pays off.
inline int launch_linear_tc_pair(const void* A, int64_t lda, const void* W, int64_t ldw, const EpiParams& ep, int M,
                                 int N, int K, int sms, cudaStream_t s) {
  using Cfg = PairCfg;
  static bool configured[64] = {};
  int dev = 0;
  B2P_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !configured[dev]) {
    B2P_CUDA(cudaFuncSetAttribute(linear_tc_pair, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    configured[dev] = true;
  }
  CUtensorMap ta, tw;
  B2P_TRY(make_tmap(&ta, 0, A, M, K, lda, Cfg::BM_CTA));
  B2P_TRY(make_tmap(&tw, 0, W, N, K, ldw, Cfg::BN / 2));
  const int num_tiles = ceil_div(M, Cfg::BM_PAIR) * ceil_div(N, Cfg::BN);
  const int pairs = std::max(1, std::min(num_tiles, sms / 2));
  // plain launch (cluster dims are compiled in); PDL attribute as for the other kernels
  B2P_LAUNCH_PDL(linear_tc_pair, 2 * pairs, Cfg::THREADS, Cfg::SMEM_BYTES, s, ta, tw, ep, M, N, K);
  return 0;
}

// Front door for every tcgen05 linear: bn == kPairTile selects the CTA-pair kernel (bf16 operands, no split-K);
// bn == 0 picks it when the 256 x 256 tiles still fill every pair of SMs (set B2P_NO_PAIR=1 to rule it out).
constexpr int kPairTile = 2;
inline bool pair_allowed() {
  static const bool allowed = [] { const char* e = getenv("B2P_NO_PAIR"); return !(e && e[0] == '1'); }();
  return allowed;
}
inline int linear_dispatch(int kind, const void* A, const void* A_lo, int64_t lda, const void* W, const void* W_lo,
                           int64_t ldw, const EpiParams& ep, int M, int N, int K, int splits, int bn, cudaStream_t s) {
  if (bn == kPairTile || (bn == 0 && kind == 0 && splits <= 1 && pair_allowed())) {
    if (M <= 0 || N <= 0 || K <= 0) return fail(B2P_EINVAL, "linear: M, N, K must be positive");
    int dev = 0, sms = 148;
    B2P_CUDA(cudaGetDevice(&dev));
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t pair_tiles = (int64_t)ceil_div(M, PairCfg::BM_PAIR) * ceil_div(N, PairCfg::BN);
    if (bn == kPairTile) {
      if (kind != 0 || splits > 1) return fail(B2P_EINVAL, "linear: the CTA-pair kernel takes bf16 operands and no split-K");
    }
    if (bn == kPairTile || pair_tiles >= sms / 2) {
      if (ep.out_bf16 && ((N & 1) || (ep.ldo & 1))) return fail(B2P_EINVAL, "linear: bf16 output needs even N and ldo");
      B2P_TRY(launch_linear_tc_pair(A, lda, W, ldw, ep, M, N, K, sms, s));
      return 1;
    }
  }
  return linear_tc_dispatch(kind, A, A_lo, lda, W, W_lo, ldw, ep, M, N, K, splits, bn == kPairTile ? 0 : bn, s);
}

}  // namespace tc
}  // namespace b2p
