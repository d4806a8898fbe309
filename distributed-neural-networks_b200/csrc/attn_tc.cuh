// attn_tc.cuh -- causal self-attention on the 5th-gen tensor cores (bf16, head dim 64).
//
//   y[b,t,h*64:(h+1)*64] = softmax_{j<=t}( q_t . k_j / 8 ) v_j          (oracle/gpt_model.py,
//   the CausalSelfAttention of the block the reference imports at gpt_model_parts.py:4)
//
// qkv is the c_attn output as bf16 [B*T, 3C] (q | k | v, head h at column h*64); nothing
// is transposed or re-laid-out: Q, K and V tiles are fetched by TMA straight out of that
// buffer (128B-swizzled boxes of 64 columns).
//
// One CTA = 256 queries (two 128-row tiles) of one (head, batch element); key/value tiles of 128:
//   warp 0       TMA producer: Q0, Q1 once, then K_j / V_j through a 2-stage ring (shared by both tiles)
//   warp 1       MMA issuer  : S_w = Q_w K_j^T  (M128 N128 K64, A and B K-major)        -> TMEM S_w
//                              O_w = P_w V_j    (M128 N64 K128, A K-major from smem,
//                                                B = V tile as MN-major operand)        -> TMEM O_w
//                              alternating w = 0, 1: PV_0(j), S_0(j+1), PV_1(j), S_1(j+1), ...
//   warps 2..5   softmax warpgroup of query tile 0, warps 6..9 of query tile 1: one query row per
//                thread.  Row max and probabilities are computed from TMEM in two passes (TMEM
//                reads are cheap, registers are not), P is written as a swizzled K-major A tile,
//                and the fresh O_j is folded into fp32 registers with the running-max correction,
//                so the tensor core never waits for a TMEM read-modify-write.
// While one warpgroup runs its softmax the tensor core works for the other (ping-pong), and
// each K/V tile is fetched once for 256 queries.  Heavy (late) query tiles are scheduled first.
#pragma once
#include <cuda_bf16.h>

#include <type_traits>

#include "common.cuh"
#include "tc_gemm.cuh"

namespace b2p {
namespace attn {

using namespace tc;

constexpr int BQ = 128, BKV = 128, HD = 64;
constexpr int QW = 2;                                  // query tiles per CTA (ping-pong)
constexpr int Q_BYTES = BQ * 128, KV_BYTES = BKV * 128, P_ATOM_BYTES = BQ * 128, P_BYTES = (BKV / 64) * P_ATOM_BYTES;
constexpr int KV_STAGES = 2;
constexpr int TMEM_COLS = 512;                         // S0 @0, S1 @128, O0 @256, O1 @320

struct Smem {
  unsigned char q[QW][Q_BYTES];
  unsigned char k[KV_STAGES][KV_BYTES];
  unsigned char v[KV_STAGES][KV_BYTES];
  unsigned char p[QW][P_BYTES];                        // per query tile: two 64-key swizzle atoms of [128 rows x 128 B]
  uint64_t q_full, kv_full[KV_STAGES], kv_empty[KV_STAGES], s_full[QW], p_full[QW], o_full[QW];
  uint32_t tmem_slot;
};
constexpr int SMEM_SLACK = 1024;
constexpr int SMEM_BYTES = sizeof(Smem) + SMEM_SLACK;  // ~161 KB: one CTA per SM
constexpr int THREADS = 320;                           // TMA warp, MMA warp, 2 x 4 softmax warps

// bf16 operands, fp32 accumulate; b_mn_major selects an MN-major B operand (V tile: d contiguous)
__host__ __device__ constexpr uint32_t idesc_bf16(int m, int n, bool b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (b_mn_major ? (1u << 16) : 0u) |
         (static_cast<uint32_t>(n >> 3) << 17) | (static_cast<uint32_t>(m >> 4) << 24);
}
__device__ __forceinline__ float ex2_fast(float x) {   // MUFU.EX2; exp2(-inf) = 0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Developer timeline (b2p_debug_trace_read): clock64 stamps of CTA (0,0,0), [role 0..2][j 0..7][event 0..7];
// role 0 = MMA warp, 1 = softmax warpgroup 0, 2 = warpgroup 1.  Costs one predicate per event.
__device__ long long g_attn_trace[3 * 8 * 8 + 1];
#define ATTN_TRACE(role, j, ev, cond)                                                              \
  do {                                                                                             \
    if (trace_cta && (j) < 8 && (cond)) g_attn_trace[((role) * 8 + (j)) * 8 + (ev)] = clock64();   \
  } while (0)

// One CTA = 256 queries (two 128-row tiles Q0, Q1) of one (head, batch element).  The two tiles
// have their own softmax warpgroup and their own S/O accumulators in tensor memory; the single
// MMA warp alternates between them, so while one warpgroup runs its softmax the tensor core
// works for the other ("ping-pong"), and every K/V tile is fetched once for both.
__global__ void __launch_bounds__(THREADS, 1)
attention_causal_tc(const __grid_constant__ CUtensorMap tmap_q,     // box [64 cols, 128 rows] over qkv
                    const __grid_constant__ CUtensorMap tmap_kv,    // box [64 cols, 128 rows] over qkv
                    __nv_bfloat16* __restrict__ y, int T, int C) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  Smem& s = *reinterpret_cast<Smem*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;     // warp index: uniform for the compiler
  const int pair = gridDim.x - 1 - blockIdx.x;          // heavy (late) query tiles first
  const int h = blockIdx.y, b = blockIdx.z;
  const int q0 = pair * (QW * BQ);
  const int row0 = b * T;                                // first row of this sequence in qkv
  const int kv_tiles_total = (T + BKV - 1) / BKV;
  // causal: query tile w needs key tiles 0 .. last[w]; a tile that starts beyond T does not exist
  int last[QW];
#pragma unroll
  for (int w = 0; w < QW; ++w)
    last[w] = (q0 + w * BQ < T) ? min((q0 + (w + 1) * BQ + BKV - 1) / BKV, kv_tiles_total) - 1 : -1;
  const int n_kv = max(last[0], last[1]) + 1;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_q)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_kv)) : "memory");
    mbar_init(&s.q_full, 1);
    for (int i = 0; i < KV_STAGES; ++i) { mbar_init(&s.kv_full[i], 1); mbar_init(&s.kv_empty[i], 1); }
    for (int w = 0; w < QW; ++w) { mbar_init(&s.s_full[w], 1); mbar_init(&s.p_full[w], 4); mbar_init(&s.o_full[w], 1); }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(&s.tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  pdl_trigger();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = s.tmem_slot;
  pdl_wait();
  const bool trace_cta = (blockIdx.x | blockIdx.y | blockIdx.z) == 0;
  if (trace_cta && threadIdx.x == 0) g_attn_trace[3 * 64] = clock64();

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      mbar_expect_tx(&s.q_full, QW * Q_BYTES);
      for (int w = 0; w < QW; ++w) tma_load_2d(s.q[w], &tmap_q, &s.q_full, h * HD, row0 + q0 + w * BQ);
    }
    uint32_t stage = 0, phase = 0;
    for (int j = 0; j < n_kv; ++j) {
      mbar_wait(&s.kv_empty[stage], phase ^ 1);
      if (lane == 0) {
        mbar_expect_tx(&s.kv_full[stage], 2 * KV_BYTES);
        tma_load_2d(s.k[stage], &tmap_kv, &s.kv_full[stage], C + h * HD, row0 + j * BKV);
        tma_load_2d(s.v[stage], &tmap_kv, &s.kv_full[stage], 2 * C + h * HD, row0 + j * BKV);
      }
      __syncwarp();
      if (++stage == KV_STAGES) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc_s = idesc_bf16(BQ, BKV, false);
    constexpr uint32_t idesc_o = idesc_bf16(BQ, HD, true);
    // all 32 lanes in uniform control flow, the tensor-core instructions predicated on elect.sync (tc_gemm.cuh, elect_one)
    const uint32_t lead = elect_one();
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    auto issue_s = [&](int w, int st) {                  // S_w = Q_w K^T  (4 MMAs, K = 64)
      const uint64_t dq = make_desc_sw128(smem_u32(s.q[w])), dk = make_desc_sw128(smem_u32(s.k[st]));
#pragma unroll
      for (int k = 0; k < HD / 16; ++k) umma_p<0>(tmem_u + w * 128, dq + 2 * k, dk + 2 * k, idesc_s, k != 0, lead);
      umma_commit_p(&s.s_full[w], lead);
    };
    auto issue_pv = [&](int w, int st) {                 // O_w = P_w V  (8 MMAs, 16 keys each)
      const uint64_t dv = make_desc_sw128(smem_u32(s.v[st]));
#pragma unroll
      for (int k = 0; k < BKV / 16; ++k) {
        // A = P: 64-key atom (k>>2), +32 B per 16 keys inside it; B = V (MN-major): +16 key rows = 2048 B
        const uint64_t dp = make_desc_sw128(smem_u32(s.p[w]) + (k >> 2) * P_ATOM_BYTES) + 2 * (k & 3);
        umma_p<0>(tmem_u + 256 + w * 64, dp, dv + 128 * k, idesc_o, k != 0, lead);
      }
      umma_commit_p(&s.o_full[w], lead);
    };
    mbar_wait(&s.q_full, 0);
    if (n_kv > 0) {
      mbar_wait(&s.kv_full[0], 0);
      tc_fence_after();
      for (int w = 0; w < QW; ++w)
        if (last[w] >= 0) issue_s(w, 0);
    }
    uint32_t stage = 0, phase = 0;
    for (int j = 0; j < n_kv; ++j) {
      const uint32_t nstage = (stage + 1 == KV_STAGES) ? 0 : stage + 1;
      const uint32_t nphase = (stage + 1 == KV_STAGES) ? phase ^ 1 : phase;
      bool next_ready = false;
      for (int w = 0; w < QW; ++w) {
        if (j <= last[w]) {
          ATTN_TRACE(0, j, w * 4 + 0, lane == 0);
          mbar_wait(&s.p_full[w], j & 1);                // P_w(j) written; S_w and O_w(j-1) already consumed
          tc_fence_after();
          ATTN_TRACE(0, j, w * 4 + 1, lane == 0);
          issue_pv(w, stage);
          ATTN_TRACE(0, j, w * 4 + 2, lane == 0);
        }
        if (w == QW - 1) {                               // both PVs of this stage are in flight
          umma_commit_p(&s.kv_empty[stage], lead);
        }
        if (j + 1 <= last[w]) {                          // start the next score tile of this query tile right away
          if (!next_ready) {
            mbar_wait(&s.kv_full[nstage], nphase);
            tc_fence_after();
            next_ready = true;
          }
          issue_s(w, nstage);
          ATTN_TRACE(0, j, w * 4 + 3, lane == 0);
        }
      }
      stage = nstage;
      phase = nphase;
    }
  } else {
    // ===================== softmax / accumulate: warpgroup w owns query tile w ====================
    const int w = (warp - 2) >> 2;
    const int qw = warp & 3;                             // TMEM lane quarter
    const int r = qw * 32 + lane;                        // row within the tile
    const int qbase = q0 + w * BQ;
    const int qi = qbase + r;                            // query position in the sequence
    const uint32_t lane_sel = static_cast<uint32_t>(qw * 32) << 16;
    const uint32_t tmem_s = tmem_base + w * 128, tmem_o = tmem_base + 256 + w * 64;
    constexpr float kScale = 0.125f * 1.4426950408889634f;   // 1/sqrt(64) * log2(e)
    float o[HD];
#pragma unroll
    for (int d = 0; d < HD; ++d) o[d] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;
    const int my_last = last[w];
    for (int j = 0; j <= my_last; ++j) {
      ATTN_TRACE(1 + w, j, 0, r == 0);
      mbar_wait(&s.s_full[w], j & 1);
      tc_fence_after();
      ATTN_TRACE(1 + w, j, 1, r == 0);
      const int k0 = j * BKV;
      const bool diag = k0 + BKV - 1 > qbase;            // tile may contain keys beyond some query of the tile
      // pass 1: row maximum.  S stays in tensor memory and is simply read twice (TMEM reads are
      // cheap) instead of parking 128 scores per thread in registers.
      float mx = m_run;
#pragma unroll 1
      for (int c2 = 0; c2 < BKV / 64; ++c2) {            // two 32-column TMEM loads in flight per wait
        uint32_t a[32], bq[32];
        tmem_ld32(tmem_s + lane_sel + c2 * 64, a);
        tmem_ld32(tmem_s + lane_sel + c2 * 64 + 32, bq);
        tmem_ld_wait();
        if (diag) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            if (k0 + c2 * 64 + i <= qi) mx = fmaxf(mx, __uint_as_float(a[i]) * kScale);
            if (k0 + c2 * 64 + 32 + i <= qi) mx = fmaxf(mx, __uint_as_float(bq[i]) * kScale);
          }
        } else {
          float t0 = -INFINITY, t1 = -INFINITY;
#pragma unroll
          for (int i = 0; i < 32; ++i) { t0 = fmaxf(t0, __uint_as_float(a[i])); t1 = fmaxf(t1, __uint_as_float(bq[i])); }
          mx = fmaxf(mx, fmaxf(t0, t1) * kScale);
        }
      }
      // (key 0 is never masked for a real query, so mx is finite whenever it matters)
      const float corr = ex2_fast(m_run - mx);
      float psum = 0.f;
      ATTN_TRACE(1 + w, j, 2, r == 0);
      // pass 2: probabilities -> bf16 -> swizzled K-major A tile in shared memory.  Two copies of the
      // loop: only the tile on the diagonal pays for the causal compare/select per element.
      auto pass2 = [&](auto diag_tag) {
        constexpr bool kDiag = decltype(diag_tag)::value;
#pragma unroll 1
        for (int c = 0; c < BKV / 32; ++c) {
          uint32_t a[32];
          tmem_ld32(tmem_s + lane_sel + c * 32, a);
          tmem_ld_wait();
          uint32_t pk[16];
          float ps0 = 0.f, ps1 = 0.f;                    // two partial sums: shorter dependency chains
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float p0 = ex2_fast(fmaf(__uint_as_float(a[2 * i]), kScale, -mx));
            float p1 = ex2_fast(fmaf(__uint_as_float(a[2 * i + 1]), kScale, -mx));
            if (kDiag) {
              if (k0 + c * 32 + 2 * i > qi) p0 = 0.f;
              if (k0 + c * 32 + 2 * i + 1 > qi) p1 = 0.f;
            }
            ps0 += p0;                                   // row sum in fp32 before the bf16 rounding (unbiased)
            ps1 += p1;
            const __nv_bfloat162 pb = __floats2bfloat162_rn(p0, p1);
            pk[i] = *reinterpret_cast<const uint32_t*>(&pb);
          }
          psum += ps0 + ps1;
          unsigned char* prow = s.p[w] + (c >> 1) * P_ATOM_BYTES + r * 128;
#pragma unroll
          for (int cc = 0; cc < 4; ++cc) {
            const int chunk = (c & 1) * 4 + cc;
            *reinterpret_cast<uint4*>(prow + ((chunk ^ (r & 7)) * 16)) =
                make_uint4(pk[4 * cc], pk[4 * cc + 1], pk[4 * cc + 2], pk[4 * cc + 3]);
          }
        }
      };
      if (diag) pass2(std::true_type{});
      else pass2(std::false_type{});
      l_run = l_run * corr + psum;
      fence_proxy_async();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&s.p_full[w]);
      ATTN_TRACE(1 + w, j, 3, r == 0);
      // rescale what is accumulated so far while the tensor core runs P_j V_j
#pragma unroll
      for (int d = 0; d < HD; ++d) o[d] *= corr;
      m_run = mx;
      ATTN_TRACE(1 + w, j, 4, r == 0);
      mbar_wait(&s.o_full[w], j & 1);
      tc_fence_after();
      ATTN_TRACE(1 + w, j, 5, r == 0);
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t a[32];
        tmem_ld32(tmem_o + lane_sel + c * 32, a);
        tmem_ld_wait();
#pragma unroll
        for (int d = 0; d < 32; ++d) o[c * 32 + d] += __uint_as_float(a[d]);
      }
      ATTN_TRACE(1 + w, j, 6, r == 0);
    }
    if (qi < T && my_last >= 0) {
      const float inv = 1.0f / l_run;
      __nv_bfloat16* yr = y + (static_cast<size_t>(row0 + qi)) * C + h * HD;
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        uint32_t wv[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const __nv_bfloat162 t2 = __floats2bfloat162_rn(o[8 * c + 2 * e] * inv, o[8 * c + 2 * e + 1] * inv);
          wv[e] = *reinterpret_cast<const uint32_t*>(&t2);
        }
        *reinterpret_cast<uint4*>(yr + 8 * c) = make_uint4(wv[0], wv[1], wv[2], wv[3]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

#undef ATTN_TRACE
inline int read_attn_trace(long long* host_out, int n) {
  if (n < 3 * 64 + 1) return fail(B2P_EINVAL, "trace buffer must hold at least 193 entries");
  B2P_CUDA(cudaDeviceSynchronize());
  B2P_CUDA(cudaMemcpyFromSymbol(host_out, g_attn_trace, (3 * 64 + 1) * sizeof(long long)));
  return 0;
}

// qkv: bf16 [M, 3C]; y: bf16 [M, C]
inline int launch_attention_tc(const __nv_bfloat16* qkv, __nv_bfloat16* y, int B, int T, int C, int H,
                               cudaStream_t st) {
  static bool configured[64] = {};
  int dev = 0;
  B2P_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !configured[dev]) {
    B2P_CUDA(cudaFuncSetAttribute(attention_causal_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    configured[dev] = true;
  }
  const int64_t M = static_cast<int64_t>(B) * T;
  CUtensorMap tq, tkv;
  {
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return fail(B2P_ECUDA, "cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t gdim[2] = {(cuuint64_t)(3 * C), (cuuint64_t)M};
    cuuint64_t gstride[1] = {(cuuint64_t)(3 * C) * 2};
    cuuint32_t estr[2] = {1, 1};
    cuuint32_t boxq[2] = {64, (cuuint32_t)BQ}, boxkv[2] = {64, (cuuint32_t)BKV};   // 128-byte rows
    CUresult r1 = enc(&tq, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<__nv_bfloat16*>(qkv), gdim, gstride, boxq, estr,
                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CUresult r2 = enc(&tkv, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<__nv_bfloat16*>(qkv), gdim, gstride, boxkv, estr,
                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r1 != CUDA_SUCCESS || r2 != CUDA_SUCCESS)
      return fail(B2P_ECUDA, "cuTensorMapEncodeTiled (attention) failed: %d / %d", (int)r1, (int)r2);
  }
  dim3 grid(ceil_div(T, QW * BQ), H, B);
  B2P_LAUNCH_PDL(attention_causal_tc, grid, THREADS, SMEM_BYTES, st, tq, tkv, y, T, C);
  return 0;
}

}  // namespace attn
}  // namespace b2p
