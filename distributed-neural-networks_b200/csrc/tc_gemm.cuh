// tc_gemm.cuh -- tcgen05 / TMEM / TMA "linear" for sm_100a.
//
//   out[m,n] = act( sum_k A[m,k] * W[n,k] + bias[n] ) (+ res[m,n])
//
// A [M,K] and W [N,K] are both K-major (row-major with K contiguous), which is
// exactly how nn.Linear stores its weight ([out,in]) and how activations sit in HBM,
// so neither operand is ever transposed.  This one kernel serves every dense
// contraction of the GPT parts (c_attn, attn.c_proj, mlp.c_fc, mlp.c_proj, lm_head;
// oracle/gpt_model.py) and fc1 of the CIFAR part 1 (cifar_model_parts.py:55).
//
// Structure (one CTA per SM, persistent over output tiles, 320 threads):
//   warp 0      TMA producer : cp.async.bulk.tensor 2-D, 128B-swizzled boxes
//                              A [128 x 128B] + W [BN x 128B] per stage, mbarrier tx
//   warp 1      MMA issuer   : one lane issues tcgen05.mma.cta_group::1 (M=128, N=BN),
//                              accumulators in TMEM, 2 accumulator stages so the
//                              epilogue of tile i overlaps the MMAs of tile i+1
//   warps 2..9  epilogue     : tcgen05.ld 32x32b -> bias/activation in registers ->
//                              per-warp smem transpose -> coalesced stores (+residual);
//                              two warps per TMEM lane quarter, half the columns each
// Ragged M, N and K need no masks on the load side: TMA zero-fills out-of-bounds box
// elements; stores are guarded.
//
// KIND 0: bf16 operands (kind::f16), KIND 1: fp32 storage read as tf32 (kind::tf32),
// KIND 2: "3xTF32" -- each fp32 operand is given as two tf32-exact planes hi + lo
//         (x = hi + lo up to 2^-22 relative) and every K step issues three MMAs
//         lo*hi + hi*lo + hi*hi into the same fp32 accumulator: fp32-grade products on the
//         tensor cores (the engine's tensor-core parity mode).
#pragma once
#include <cuda_bf16.h>

#include "common.cuh"

namespace b2p {
namespace tc {

// ------------------------------------------------------------------------------------
// PTX wrappers
// ------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug becomes a trap (reported as a launch failure) instead of
// a kernel that spins forever and takes the GPU with it.
__device__ __forceinline__ uint64_t global_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const uint64_t t0 = global_ns();
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++spins & 0x3FF) == 0) {
      // A role of a ring-fed kernel may legitimately sit here for as long as the ring wait it depends on (which gives
      // up by itself, common.cuh); beyond that only a protocol bug keeps a barrier closed: trap rather than hang.
      unsigned long long limit;
      asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(limit) : "l"(&g_wait_limit_ns) : "memory");
      if (limit && global_ns() - t0 > limit + 10000000000ull) __trap();
    }
  }
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0,
                                            int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
// the same two from warp-uniform code, predicated on elect.sync (see elect_one below): the tensor-map pointer, the
// coordinates and the shared-memory addresses stay in uniform registers -- no ELECT / R2UR / BRA.U.ANY loop per load
__device__ __forceinline__ void mbar_expect_tx_p(uint32_t bar_addr, uint32_t bytes, uint32_t lead) {
  asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %2, 0;\n\t"
               "@q mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n\t}" ::"r"(bar_addr), "r"(bytes), "r"(lead) : "memory");
}
__device__ __forceinline__ void tma_load_2d_p(uint32_t smem_dst, const CUtensorMap* map, uint32_t bar_addr, int c0, int c1,
                                              uint32_t lead) {
  asm volatile(
      "{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %5, 0;\n\t"
      "@q cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n\t}"
      ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar_addr), "r"(c0), "r"(c1), "r"(lead)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {   // whole warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t ncols) {      // whole warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {   // implies fence::before_thread_sync
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
template <int KIND>
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                     uint32_t accumulate) {
  if (KIND == 0) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}
// ---- single-thread tensor-core instructions from warp-uniform code --------------------------------------------------
// tcgen05.mma / tcgen05.commit are issued by one thread.  Written as `if (lane == 0) { ... }` the compiler sees a
// divergent region, keeps every descriptor in vector registers and wraps EVERY such instruction in an
// ELECT / R2UR.BROADCAST / BRA.U.ANY loop (~14 dependent instructions): issuing an 81-cycle MMA then takes ~80 cycles
// and a barrier hand-over 150-350 (measured in part0_tc2 with clock64 stamps).  The issuing warp therefore runs its loop
// with all 32 lanes (warp index and tensor-memory base made warp-uniform with __shfl_sync(.., 0)) and predicates the
// instruction itself on elect.sync: operands stay in uniform registers, UTCHMMAs issue back to back.
__device__ __forceinline__ uint32_t elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.b32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred;
}
template <int KIND>
__device__ __forceinline__ void umma_p(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                       uint32_t accumulate, uint32_t lead) {
  if (KIND == 0) {
    asm volatile(
        "{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\tsetp.ne.b32 q, %5, 0;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(lead)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\tsetp.ne.b32 q, %5, 0;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(lead)
        : "memory");
  }
}
__device__ __forceinline__ void umma_commit_p(uint64_t* bar, uint32_t lead) {   // implies fence::before_thread_sync
  asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %1, 0;\n\t"
               "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}"
               ::"r"(smem_u32(bar)), "r"(lead) : "memory");
}
// 32 lanes x 32 consecutive 32-bit columns: thread t of the warp gets TMEM lane
// (lane_base + t), columns col .. col+31.
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major operand tile in the canonical 128B-swizzle layout: rows at a 128 B pitch,
// 8-row groups 1024 B apart (SBO), 16-byte chunk c of row r stored at chunk c^(r&7) --
// which is precisely what a TMA box of 128 B inner extent with SWIZZLE_128B writes.
// (Field layout: cute/arch/mma_sm100_desc.hpp SmemDescriptor.)
__device__ __forceinline__ uint64_t make_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);   // [0,14)  start address >> 4
  d |= static_cast<uint64_t>(1) << 16;                        // [16,30) LBO (ignored for swizzled K-major)
  d |= static_cast<uint64_t>(1024 >> 4) << 32;                // [32,46) SBO = 1024 B
  d |= static_cast<uint64_t>(1) << 46;                        // [46,48) descriptor version (Blackwell)
  d |= static_cast<uint64_t>(2) << 61;                        // [61,64) SWIZZLE_128B
  return d;
}
// Instruction descriptor (cute/arch/mma_sm100_desc.hpp InstrDescriptor): fp32 accumulate,
// A and B K-major, no negate, dense.
__host__ __device__ constexpr uint32_t make_idesc(int kind, int m, int n) {
  const uint32_t fmt = kind == 0 ? 1u /*BF16*/ : 2u /*TF32*/;
  return (1u << 4) | (fmt << 7) | (fmt << 10) | (static_cast<uint32_t>(n >> 3) << 17) |
         (static_cast<uint32_t>(m >> 4) << 24);
}

// ------------------------------------------------------------------------------------
// kernel
// ------------------------------------------------------------------------------------
enum { ACT_NONE = 0, ACT_RELU = 1, ACT_GELU = 2 };

struct EpiParams {
  const float* bias;     // [N] or null
  const float* res;      // [M, ld_res] fp32 or null: added AFTER the activation; must not overlap `out`
  void* out;             // [M, ldo] fp32 or bf16
  int64_t ldo, ld_res;
  int out_bf16;
  int act;
  int64_t split_stride;  // split-K: elements between the partial outputs of consecutive K slices
  // Greedy-id mode (LM head): when amax_val is set nothing is stored to `out`; instead every
  // epilogue warp writes the (max, first arg-max) of its rows over its columns to slot
  // (tile_n*2 + column_half) of [M][amax_slots] -- the [B,T,V] logits never touch HBM.
  float* amax_val;
  int* amax_idx;
  int amax_slots;
};

// ------------------------------------------------------------------------------------
// Epilogue of one output tile for one warp: rows [m0 + 32q, +32), columns of half `col_half` of the
// BN-wide tile whose accumulator starts at TMEM column `tmem_acc`.  Shared by the 1-CTA and the
// 2-CTA kernels.  `stg` is the warp's private [32][P] transpose buffer, `bias_s` its bias slice.
// ------------------------------------------------------------------------------------
template <int BN, int P>
__device__ __forceinline__ void epilogue_tile(const EpiParams& ep, float* stg, const float* bias_s, uint32_t tmem_acc,
                                              int m0, int n0, int M, int N, int q, int col_half, int lane, int z,
                                              int amax_slot_base, bool vec_f32, bool vec_bf16) {
  constexpr int CHUNKS = BN / 64;                 // 32-column chunks per warp
  const int row_base = m0 + q * 32;
  float* out_f32 = static_cast<float*>(ep.out) + static_cast<int64_t>(z) * ep.split_stride;
  if (ep.amax_val) {
    float best = -INFINITY;
    int bidx = 0x7fffffff;
#pragma unroll 1
    for (int cc = 0; cc < CHUNKS; ++cc) {
      const int c = col_half * CHUNKS + cc;
      const int col0 = n0 + c * 32;
      if (col0 >= N) break;
      uint32_t r[32];
      tmem_ld32(tmem_acc + c * 32 + (static_cast<uint32_t>(q * 32) << 16), r);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        float v = __uint_as_float(r[j]);
        if (ep.bias) v += bias_s[cc * 32 + j];
        if (col0 + j < N && v > best) { best = v; bidx = col0 + j; }     // ascending scan: first maximum wins
      }
    }
    const int row = row_base + lane;
    if (row < M) {
      const int64_t slot = static_cast<int64_t>(row) * ep.amax_slots + amax_slot_base + col_half;
      ep.amax_val[slot] = best;
      ep.amax_idx[slot] = bidx;
    }
    return;
  }
#pragma unroll 1
  for (int cc = 0; cc < CHUNKS; ++cc) {
    const int c = col_half * CHUNKS + cc;
    const int col0 = n0 + c * 32;
    if (col0 >= N) break;                        // warp-uniform
    uint32_t r[32];
    tmem_ld32(tmem_acc + c * 32 + (static_cast<uint32_t>(q * 32) << 16), r);
    const bool full_cols = col0 + 32 <= N;
    // Residual values are fetched NOW -- while the TMEM load is in flight and before any store
    // of this chunk.  A load issued after a may-alias store has to wait for it; leaving the
    // loads inside the store loop serialised eight ~700-cycle round trips per chunk.
    const bool res_vec = ep.res != nullptr && vec_f32 && full_cols;
    float4 rv[2][4];
    if (res_vec) {
      const int cg = (lane & 3) * 4;
#pragma unroll
      for (int sub = 0; sub < 2; ++sub)
#pragma unroll
        for (int it = 0; it < 4; ++it) {
          const int row = row_base + it * 8 + (lane >> 2);
          rv[sub][it] = row < M ? ld_produced(reinterpret_cast<const float4*>(ep.res + static_cast<int64_t>(row) * ep.ld_res +
                                                                             col0 + sub * 16 + cg))
                                : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    tmem_ld_wait();
    float v[32];
#pragma unroll
    for (int j4 = 0; j4 < 8; ++j4) {
#pragma unroll
      for (int e = 0; e < 4; ++e) v[4 * j4 + e] = __uint_as_float(r[4 * j4 + e]);
      if (ep.bias) {
        const float4 b4 = *reinterpret_cast<const float4*>(&bias_s[cc * 32 + 4 * j4]);     // smem broadcast
        v[4 * j4] += b4.x; v[4 * j4 + 1] += b4.y; v[4 * j4 + 2] += b4.z; v[4 * j4 + 3] += b4.w;
      }
    }
    if (ep.act == ACT_RELU) {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = fmax_nan(v[j], 0.f);
    } else if (ep.act == ACT_GELU) {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = gelu_erf_fast(v[j]);
    }
#pragma unroll
    for (int sub = 0; sub < 2; ++sub) {
      const int cs = col0 + sub * 16;
#pragma unroll
      for (int j4 = 0; j4 < 4; ++j4)
        *reinterpret_cast<float4*>(&stg[lane * P + 4 * j4]) =
            make_float4(v[sub * 16 + 4 * j4], v[sub * 16 + 4 * j4 + 1], v[sub * 16 + 4 * j4 + 2], v[sub * 16 + 4 * j4 + 3]);
      __syncwarp();
      if (vec_f32 && full_cols) {
        // 4 lanes x 16 B = one 64-byte row segment; 8 rows per instruction
        const int cg = (lane & 3) * 4;
#pragma unroll
        for (int it = 0; it < 4; ++it) {
          const int rr = it * 8 + (lane >> 2), row = row_base + rr;
          if (row < M) {
            float4 o = *reinterpret_cast<const float4*>(&stg[rr * P + cg]);
            if (res_vec) { o.x += rv[sub][it].x; o.y += rv[sub][it].y; o.z += rv[sub][it].z; o.w += rv[sub][it].w; }
            *reinterpret_cast<float4*>(out_f32 + static_cast<int64_t>(row) * ep.ldo + cs + cg) = o;
          }
        }
      } else if (vec_bf16 && full_cols) {
        // 2 lanes x 16 B (8 bf16) = one 32-byte row segment; 16 rows per instruction
        __nv_bfloat16* out = static_cast<__nv_bfloat16*>(ep.out);
        const int cg = (lane & 1) * 8;
#pragma unroll
        for (int it = 0; it < 2; ++it) {
          const int rr = it * 16 + (lane >> 1), row = row_base + rr;
          if (row < M) {
            const float4 x0 = *reinterpret_cast<const float4*>(&stg[rr * P + cg]);
            const float4 x1 = *reinterpret_cast<const float4*>(&stg[rr * P + cg + 4]);
            __nv_bfloat162 p0 = __floats2bfloat162_rn(x0.x, x0.y), p1 = __floats2bfloat162_rn(x0.z, x0.w);
            __nv_bfloat162 p2 = __floats2bfloat162_rn(x1.x, x1.y), p3 = __floats2bfloat162_rn(x1.z, x1.w);
            uint4 pk;
            pk.x = *reinterpret_cast<uint32_t*>(&p0); pk.y = *reinterpret_cast<uint32_t*>(&p1);
            pk.z = *reinterpret_cast<uint32_t*>(&p2); pk.w = *reinterpret_cast<uint32_t*>(&p3);
            *reinterpret_cast<uint4*>(out + static_cast<int64_t>(row) * ep.ldo + cs + cg) = pk;
          }
        }
      } else {
        // generic path (ragged N, unaligned rows, bf16 + residual): one element per lane-iteration
        const int cj = lane & 15, col = cs + cj;
#pragma unroll 4
        for (int it = 0; it < 16; ++it) {
          const int rr = it * 2 + (lane >> 4), row = row_base + rr;
          if (row < M && col < N) {
            float o = stg[rr * P + cj];
            if (ep.res) o += ep.res[static_cast<int64_t>(row) * ep.ld_res + col];
            if (ep.out_bf16) static_cast<__nv_bfloat16*>(ep.out)[static_cast<int64_t>(row) * ep.ldo + col] = __float2bfloat16_rn(o);
            else out_f32[static_cast<int64_t>(row) * ep.ldo + col] = o;
          }
        }
      }
      __syncwarp();
    }
  }
}

template <int KIND, int BN>
struct GemmCfg {
  static constexpr int BM = 128;
  static constexpr bool SPLIT = KIND == 2;     // hi/lo planes
  static constexpr int ELEM = KIND == 0 ? 2 : 4;
  static constexpr int BK = 128 / ELEM;        // one 128-byte swizzle row of K per stage
  static constexpr int UMMA_K = 32 / ELEM;     // 16 (bf16) or 8 (tf32)
  static constexpr int A_PLANE = BM * 128, B_PLANE = BN * 128;
  static constexpr int A_BYTES = A_PLANE * (SPLIT ? 2 : 1);
  static constexpr int B_BYTES = B_PLANE * (SPLIT ? 2 : 1);
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = SPLIT ? (BN <= 128 ? 3 : 2) : (BN <= 128 ? 5 : 4);
  static constexpr int EPI_PITCH = 20;       // floats: 16 columns + pad; 16-byte aligned rows, conflict-free float4 stores
  static constexpr int EPI_WARP_FLOATS = 32 * EPI_PITCH + BN / 2;   // transpose buffer + this warp's bias slice
  static constexpr int EPI_BYTES = 8 * EPI_WARP_FLOATS * 4;
  static constexpr int BAR_BYTES = 256;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + BAR_BYTES + 1024 /*alignment slack*/;
  static constexpr int TMEM_COLS = 2 * BN;     // two accumulator stages (power of two >= 32)
  static constexpr int THREADS = 320;     // TMA warp, MMA warp, 8 epilogue warps
};

template <int KIND, int BN>
__global__ void __launch_bounds__(320, 1)
linear_tc(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
          const __grid_constant__ CUtensorMap tmap_a_lo, const __grid_constant__ CUtensorMap tmap_w_lo,
          EpiParams ep, int M, int N, int K, int splits, int kb_per_split) {
  using Cfg = GemmCfg<KIND, BN>;
  // SWIZZLE_128B tiles need 1024-byte alignment (the XOR pattern is a function of the address).
  // The fix-up is an offset on the __shared__ array (not uintptr_t arithmetic) so the compiler
  // keeps the shared address space: LDS/STS instead of generic LD/ST in the epilogue.
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  unsigned char* tiles = smem;
  float* epi_smem = reinterpret_cast<float*>(smem + Cfg::STAGES * Cfg::STAGE_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::STAGES * Cfg::STAGE_BYTES + Cfg::EPI_BYTES);
  uint64_t* full = bars;                         // [STAGES]  TMA -> MMA
  uint64_t* empty = bars + Cfg::STAGES;          // [STAGES]  MMA -> TMA
  uint64_t* tmem_full = bars + 2 * Cfg::STAGES;  // [2]       MMA -> epilogue
  uint64_t* tmem_empty = tmem_full + 2;          // [2]       epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;     // warp index: uniform for the compiler
  const int tiles_m = (M + Cfg::BM - 1) / Cfg::BM, tiles_n = (N + BN - 1) / BN;
  const int num_tiles = tiles_m * tiles_n * splits;        // work items: (output tile, K slice)
  const int total_kb = (K + Cfg::BK - 1) / Cfg::BK;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_a)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_w)) : "memory");
    for (int i = 0; i < Cfg::STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 8); }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    tmem_relinquish();
  }
  pdl_trigger();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();        // everything above overlapped the previous kernel's tail; its outputs are visible from here

  if (warp == 0) {
    // ===================== TMA producer (uniform control flow, loads predicated on elect.sync) =====================
    const uint32_t lead = elect_one();
    const uint32_t tiles_u = __shfl_sync(0xffffffffu, smem_u32(tiles), 0);
    const uint32_t full_u = __shfl_sync(0xffffffffu, smem_u32(full), 0);
    uint32_t stage = 0, phase = 0;
    for (int item = blockIdx.x; item < num_tiles; item += gridDim.x) {
      const int z = item % splits, tile = item / splits;
      const int m0 = (tile % tiles_m) * Cfg::BM, n0 = (tile / tiles_m) * BN;
      const int kb0 = z * kb_per_split, kb1 = min(total_kb, kb0 + kb_per_split);
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&empty[stage], phase ^ 1);
        const uint32_t sa = tiles_u + stage * Cfg::STAGE_BYTES, fb = full_u + stage * 8;
        mbar_expect_tx_p(fb, Cfg::STAGE_BYTES, lead);
        tma_load_2d_p(sa, &tmap_a, fb, kb * Cfg::BK, m0, lead);
        tma_load_2d_p(sa + Cfg::A_BYTES, &tmap_w, fb, kb * Cfg::BK, n0, lead);
        if (Cfg::SPLIT) {
          tma_load_2d_p(sa + Cfg::A_PLANE, &tmap_a_lo, fb, kb * Cfg::BK, m0, lead);
          tma_load_2d_p(sa + Cfg::A_BYTES + Cfg::B_PLANE, &tmap_w_lo, fb, kb * Cfg::BK, n0, lead);
        }
        if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // all 32 lanes in uniform control flow, the tensor-core instructions predicated on elect.sync (see elect_one above)
    constexpr uint32_t idesc = make_idesc(KIND, Cfg::BM, BN);
    const uint32_t lead = elect_one();
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t tiles_u = __shfl_sync(0xffffffffu, smem_u32(tiles), 0);
    uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
    for (int item = blockIdx.x; item < num_tiles; item += gridDim.x) {
      const int z = item % splits;
      const int kb0 = z * kb_per_split, num_kb = min(total_kb, kb0 + kb_per_split) - kb0;
      mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_u + acc * BN;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        const uint32_t a_addr = tiles_u + stage * Cfg::STAGE_BYTES;
        const uint64_t da = make_desc_sw128(a_addr), db = make_desc_sw128(a_addr + Cfg::A_BYTES);
#pragma unroll
        for (int k = 0; k < Cfg::BK / Cfg::UMMA_K; ++k) {  // +32 B along K inside the swizzle row = +2 (>>4)
          if (Cfg::SPLIT) {
            const uint64_t da_lo = make_desc_sw128(a_addr + Cfg::A_PLANE);
            const uint64_t db_lo = make_desc_sw128(a_addr + Cfg::A_BYTES + Cfg::B_PLANE);
            umma_p<KIND>(tmem_d, da_lo + 2 * k, db + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u, lead);   // small terms first
            umma_p<KIND>(tmem_d, da + 2 * k, db_lo + 2 * k, idesc, 1u, lead);
            umma_p<KIND>(tmem_d, da + 2 * k, db + 2 * k, idesc, 1u, lead);
          } else {
            umma_p<KIND>(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u, lead);
          }
        }
        umma_commit_p(&empty[stage], lead);                  // smem slot free once these MMAs retire
        if (kb == num_kb - 1) umma_commit_p(&tmem_full[acc], lead);
        if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else {
    // ===================== epilogue (warps 2..9) =====================
    // Eight warps: two per TMEM lane quarter, each taking half of the tile's columns, so a
    // heavy epilogue (erf-GELU, residual add) keeps up with short-K main loops.  Data path per
    // 32-column chunk: tcgen05.ld -> registers (bias, activation) -> per-warp [32][20] smem
    // transpose in two 16-column halves -> row-contiguous 16-byte stores.
    constexpr int P = Cfg::EPI_PITCH;
    const int q = warp & 3;                         // TMEM lane quarter this warp may read
    const int ew = warp - 2;                        // 0..7
    const int col_half = ew >> 2;                   // which half of the tile's columns
    float* stg = epi_smem + ew * Cfg::EPI_WARP_FLOATS;
    float* bias_s = stg + 32 * P;                   // [BN/2] bias of this warp's columns for the current tile
    const bool vec_f32 = !ep.out_bf16 && (ep.ldo & 3) == 0 && (reinterpret_cast<uintptr_t>(ep.out) & 15) == 0 &&
                         (ep.split_stride & 3) == 0 &&
                         (!ep.res || ((ep.ld_res & 3) == 0 && (reinterpret_cast<uintptr_t>(ep.res) & 15) == 0));
    const bool vec_bf16 = ep.out_bf16 && (ep.ldo & 7) == 0 && (reinterpret_cast<uintptr_t>(ep.out) & 15) == 0 && !ep.res;
    uint32_t acc = 0, acc_phase = 0;
    for (int item = blockIdx.x; item < num_tiles; item += gridDim.x) {
      const int z = item % splits, tile = item / splits;
      const int m0 = (tile % tiles_m) * Cfg::BM, n0 = (tile / tiles_m) * BN;
      // This warp's bias slice goes to shared memory BEFORE waiting for the accumulator: the
      // global-load latency hides behind the MMA wait instead of sitting in front of every chunk.
      if (ep.bias) {
        const int cbase = n0 + col_half * (BN / 2);
#pragma unroll
        for (int i = lane; i < BN / 2; i += 32) bias_s[i] = (cbase + i < N) ? __ldg(ep.bias + cbase + i) : 0.f;
        __syncwarp();
      }
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      epilogue_tile<BN, P>(ep, stg, bias_s, tmem_base + acc * BN, m0, n0, M, N, q, col_half, lane, z, (tile / tiles_m) * 2,
                           vec_f32, vec_bf16);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

// ------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// 2-D K-major tensor [rows, k] with row pitch ld (elements); box = [128 bytes of K] x [box_rows].
inline int make_tmap_uncached(CUtensorMap* map, int kind, const void* ptr, int64_t rows, int64_t k, int64_t ld, int box_rows) {
  if (kind == 2) kind = 1;     // tf32x3 planes are plain fp32 arrays
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc) return fail(B2P_ECUDA, "cuTensorMapEncodeTiled is not available from the driver");
  const int elem = kind == 0 ? 2 : 4;
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || ((ld * elem) & 15))
    return fail(B2P_EINVAL, "TMA operand must be 16-byte aligned with a 16-byte multiple row pitch (ld=%lld)",
                (long long)ld);
  cuuint64_t gdim[2] = {(cuuint64_t)k, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)(ld * elem)};
  cuuint32_t box[2] = {(cuuint32_t)(128 / elem), (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, kind == 0 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                   const_cast<void*>(ptr), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(B2P_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}

// cuTensorMapEncodeTiled costs a few microseconds of host time; weights and workspace buffers
// recur with identical geometry on every call, so encoded maps are kept in a small per-thread
// direct-mapped cache keyed by everything that goes into the encoding.
inline int make_tmap(CUtensorMap* map, int kind, const void* ptr, int64_t rows, int64_t k, int64_t ld, int box_rows) {
  struct Entry { const void* ptr; int64_t rows, k, ld; int kind, box; bool valid; CUtensorMap map; };
  static thread_local Entry cache[256] = {};
  const uint64_t h = (reinterpret_cast<uintptr_t>(ptr) >> 8) * 0x9E3779B97F4A7C15ull + (uint64_t)rows * 31 + (uint64_t)k * 17 +
                     (uint64_t)box_rows * 7 + (uint64_t)kind;
  Entry& e = cache[(h >> 32) & 255];
  if (e.valid && e.ptr == ptr && e.rows == rows && e.k == k && e.ld == ld && e.kind == kind && e.box == box_rows) {
    *map = e.map;
    return 0;
  }
  B2P_TRY(make_tmap_uncached(map, kind, ptr, rows, k, ld, box_rows));
  e.ptr = ptr; e.rows = rows; e.k = k; e.ld = ld; e.kind = kind; e.box = box_rows; e.map = *map; e.valid = true;
  return 0;
}

template <int KIND, int BN>
int launch_linear_tc(const void* A, const void* A_lo, int64_t lda, const void* W, const void* W_lo, int64_t ldw,
                     const EpiParams& ep, int M, int N, int K, int splits, int sms, cudaStream_t s) {
  using Cfg = GemmCfg<KIND, BN>;
  static bool configured[64] = {};
  int dev = 0;
  B2P_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !configured[dev]) {
    B2P_CUDA(cudaFuncSetAttribute(linear_tc<KIND, BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  Cfg::SMEM_BYTES));
    configured[dev] = true;
  }
  CUtensorMap ta, tw, ta_lo, tw_lo;
  B2P_TRY(make_tmap(&ta, KIND, A, M, K, lda, Cfg::BM));
  B2P_TRY(make_tmap(&tw, KIND, W, N, K, ldw, BN));
  if (Cfg::SPLIT) {
    if (!A_lo || !W_lo) return fail(B2P_EINVAL, "linear: tf32x3 needs the lo planes of both operands");
    B2P_TRY(make_tmap(&ta_lo, KIND, A_lo, M, K, lda, Cfg::BM));
    B2P_TRY(make_tmap(&tw_lo, KIND, W_lo, N, K, ldw, BN));
  } else {
    ta_lo = ta;
    tw_lo = tw;
  }
  const int total_kb = ceil_div(K, Cfg::BK);
  splits = std::max(1, std::min(splits, total_kb));
  const int kb_per_split = ceil_div(total_kb, splits);
  splits = ceil_div(total_kb, kb_per_split);            // no empty K slice
  const int items = ceil_div(M, Cfg::BM) * ceil_div(N, BN) * splits;
  const int grid = std::min(items, sms);
  B2P_LAUNCH_PDL((linear_tc<KIND, BN>), grid, Cfg::THREADS, Cfg::SMEM_BYTES, s, ta, tw, ta_lo, tw_lo, ep, M, N, K,
                 splits, kb_per_split);
  return splits;
}

// kind: 0 bf16 operands, 1 fp32 operands multiplied as tf32.  splits > 1: split-K, partial
// sums of slice z go to out + z*split_stride (fp32, no bias/act/res).  Returns the number of
// K slices actually used (>= 1) or a negative error.  bn: 0 = pick, else 128 or 256.
inline int linear_tc_dispatch(int kind, const void* A, const void* A_lo, int64_t lda, const void* W,
                              const void* W_lo, int64_t ldw, const EpiParams& ep, int M, int N, int K, int splits,
                              int bn, cudaStream_t s) {
  if (M <= 0 || N <= 0 || K <= 0) return fail(B2P_EINVAL, "linear: M, N, K must be positive");
  if (ep.out_bf16 && ((N & 1) || (ep.ldo & 1))) return fail(B2P_EINVAL, "linear: bf16 output needs even N and ldo");
  if (splits > 1 && (ep.out_bf16 || ep.bias || ep.res || ep.act))
    return fail(B2P_EINVAL, "linear: split-K writes raw fp32 partial sums (no bias/activation/residual)");
  int dev = 0, sms = 148;
  B2P_CUDA(cudaGetDevice(&dev));
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (kind == 2) bn = 128;
  if (bn == 0) {
    // 256-wide tiles halve the shared-memory traffic per MMA (the BN=128 tile sits exactly at
    // the 128 B/clk smem limit) but only pay off when they still fill the machine
    const int64_t tiles256 = (int64_t)ceil_div(M, 128) * ceil_div(N, 256) * std::max(1, splits);
    bn = (N >= 256 && tiles256 >= sms) ? 256 : 128;
  }
#define B2P_TC_CASE(KD, TN) \
  if (kind == KD && bn == TN) return launch_linear_tc<KD, TN>(A, A_lo, lda, W, W_lo, ldw, ep, M, N, K, splits, sms, s)
  B2P_TC_CASE(0, 128);
  B2P_TC_CASE(0, 256);
  B2P_TC_CASE(1, 128);
  B2P_TC_CASE(1, 256);
  B2P_TC_CASE(2, 128);
#undef B2P_TC_CASE
  return fail(B2P_EINVAL, "linear: unknown operand kind %d / tile width %d", kind, bn);
}

}  // namespace tc
}  // namespace b2p
