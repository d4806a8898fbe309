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
// Structure (one CTA per SM, persistent over output tiles, 192 threads):
//   warp 0      TMA producer : cp.async.bulk.tensor 2-D, 128B-swizzled boxes
//                              A [128 x 128B] + W [BN x 128B] per stage, mbarrier tx
//   warp 1      MMA issuer   : one lane issues tcgen05.mma.cta_group::1 (M=128, N=BN),
//                              accumulators in TMEM, 2 accumulator stages so the
//                              epilogue of tile i overlaps the MMAs of tile i+1
//   warps 2..5  epilogue     : tcgen05.ld 32x32b -> bias/activation in registers ->
//                              per-warp smem transpose -> coalesced stores (+residual)
// Ragged M, N and K need no masks on the load side: TMA zero-fills out-of-bounds box
// elements; stores are guarded.
//
// KIND 0: bf16 operands (kind::f16), KIND 1: fp32 storage read as tf32 (kind::tf32).
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
    if ((++spins & 0x3FF) == 0 && global_ns() - t0 > 4000000000ull) __trap();   // 4 s: no legitimate wait is this long
  }
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0,
                                            int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
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
  const float* res;      // [M, ld_res] fp32 or null: added AFTER the activation
  void* out;             // [M, ldo] fp32 or bf16
  int64_t ldo, ld_res;
  int out_bf16;
  int act;
};

template <int KIND, int BN>
struct GemmCfg {
  static constexpr int BM = 128;
  static constexpr int ELEM = KIND == 0 ? 2 : 4;
  static constexpr int BK = 128 / ELEM;        // one 128-byte swizzle row of K per stage
  static constexpr int UMMA_K = 32 / ELEM;     // 16 (bf16) or 8 (tf32)
  static constexpr int A_BYTES = BM * 128;
  static constexpr int B_BYTES = BN * 128;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = BN <= 128 ? 5 : 4;
  static constexpr int EPI_BYTES = 4 * 32 * 33 * 4;
  static constexpr int BAR_BYTES = 256;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + BAR_BYTES + 1024 /*alignment slack*/;
  static constexpr int TMEM_COLS = 2 * BN;     // two accumulator stages (power of two >= 32)
  static constexpr int THREADS = 192;
};

template <int KIND, int BN>
__global__ void __launch_bounds__(192, 1)
linear_tc(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
          EpiParams ep, int M, int N, int K) {
  using Cfg = GemmCfg<KIND, BN>;
  extern __shared__ unsigned char smem_raw[];
  // SWIZZLE_128B tiles need 1024-byte alignment (the XOR pattern is a function of the address)
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                                         ~static_cast<uintptr_t>(1023));
  unsigned char* tiles = smem;
  float* epi_smem = reinterpret_cast<float*>(smem + Cfg::STAGES * Cfg::STAGE_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::STAGES * Cfg::STAGE_BYTES + Cfg::EPI_BYTES);
  uint64_t* full = bars;                         // [STAGES]  TMA -> MMA
  uint64_t* empty = bars + Cfg::STAGES;          // [STAGES]  MMA -> TMA
  uint64_t* tmem_full = bars + 2 * Cfg::STAGES;  // [2]       MMA -> epilogue
  uint64_t* tmem_empty = tmem_full + 2;          // [2]       epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_m = (M + Cfg::BM - 1) / Cfg::BM, tiles_n = (N + BN - 1) / BN;
  const int num_tiles = tiles_m * tiles_n;
  const int num_kb = (K + Cfg::BK - 1) / Cfg::BK;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_a)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_w)) : "memory");
    for (int i = 0; i < Cfg::STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 4); }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    uint32_t stage = 0, phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m0 = (tile % tiles_m) * Cfg::BM, n0 = (tile / tiles_m) * BN;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&empty[stage], phase ^ 1);
        if (lane == 0) {
          unsigned char* sa = tiles + stage * Cfg::STAGE_BYTES;
          mbar_expect_tx(&full[stage], Cfg::STAGE_BYTES);
          tma_load_2d(sa, &tmap_a, &full[stage], kb * Cfg::BK, m0);
          tma_load_2d(sa + Cfg::A_BYTES, &tmap_w, &full[stage], kb * Cfg::BK, n0);
        }
        __syncwarp();
        if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = make_idesc(KIND, Cfg::BM, BN);
    uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * BN;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t a_addr = smem_u32(tiles + stage * Cfg::STAGE_BYTES);
          const uint64_t da = make_desc_sw128(a_addr), db = make_desc_sw128(a_addr + Cfg::A_BYTES);
#pragma unroll
          for (int k = 0; k < Cfg::BK / Cfg::UMMA_K; ++k)   // +32 B along K inside the swizzle row = +2 (>>4)
            umma<KIND>(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
          umma_commit(&empty[stage]);                        // smem slot free once these MMAs retire
          if (kb == num_kb - 1) umma_commit(&tmem_full[acc]);
        }
        __syncwarp();
        if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int q = warp & 3;                         // TMEM lane quarter this warp may read
    float* stg = epi_smem + (warp - 2) * 32 * 33;   // private [32][33] transpose buffer
    uint32_t acc = 0, acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m0 = (tile % tiles_m) * Cfg::BM, n0 = (tile / tiles_m) * BN;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const int row_base = m0 + q * 32;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        const int col0 = n0 + c * 32;
        if (col0 >= N) break;                        // warp-uniform
        uint32_t r[32];
        tmem_ld32(tmem_base + acc * BN + c * 32 + (static_cast<uint32_t>(q * 32) << 16), r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          float v = __uint_as_float(r[j]);
          if (ep.bias && col0 + j < N) v += __ldg(ep.bias + col0 + j);
          if (ep.act == ACT_RELU) v = fmaxf(v, 0.f);
          else if (ep.act == ACT_GELU) v = gelu_erf(v);
          stg[lane * 33 + j] = v;
        }
        __syncwarp();
        if (!ep.out_bf16) {
          float* out = static_cast<float*>(ep.out);
          const int col = col0 + lane;
#pragma unroll 4
          for (int rr = 0; rr < 32; ++rr) {
            const int row = row_base + rr;
            if (row < M && col < N) {
              float v = stg[rr * 33 + lane];
              if (ep.res) v += ep.res[static_cast<int64_t>(row) * ep.ld_res + col];
              out[static_cast<int64_t>(row) * ep.ldo + col] = v;
            }
          }
        } else {
          __nv_bfloat16* out = static_cast<__nv_bfloat16*>(ep.out);
          const int half = lane >> 4, cj = (lane & 15) * 2, col = col0 + cj;
#pragma unroll 4
          for (int it = 0; it < 16; ++it) {
            const int rr = it * 2 + half, row = row_base + rr;
            if (row < M && col < N) {                // N is even for bf16 outputs (checked on the host)
              float v0 = stg[rr * 33 + cj], v1 = stg[rr * 33 + cj + 1];
              if (ep.res) {
                const float* rp = ep.res + static_cast<int64_t>(row) * ep.ld_res + col;
                v0 += rp[0]; v1 += rp[1];
              }
              *reinterpret_cast<__nv_bfloat162*>(out + static_cast<int64_t>(row) * ep.ldo + col) =
                  __floats2bfloat162_rn(v0, v1);
            }
          }
        }
        __syncwarp();
      }
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
inline int make_tmap(CUtensorMap* map, int kind, const void* ptr, int64_t rows, int64_t k, int64_t ld, int box_rows) {
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

template <int KIND, int BN>
int launch_linear_tc(const void* A, int64_t lda, const void* W, int64_t ldw, const EpiParams& ep, int M, int N,
                     int K, int sms, cudaStream_t s) {
  using Cfg = GemmCfg<KIND, BN>;
  static bool configured[64] = {};
  int dev = 0;
  B2P_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !configured[dev]) {
    B2P_CUDA(cudaFuncSetAttribute(linear_tc<KIND, BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  Cfg::SMEM_BYTES));
    configured[dev] = true;
  }
  CUtensorMap ta, tw;
  B2P_TRY(make_tmap(&ta, KIND, A, M, K, lda, Cfg::BM));
  B2P_TRY(make_tmap(&tw, KIND, W, N, K, ldw, BN));
  const int num_tiles = ceil_div(M, Cfg::BM) * ceil_div(N, BN);
  const int grid = std::min(num_tiles, sms);
  B2P_LAUNCH((linear_tc<KIND, BN>), grid, Cfg::THREADS, Cfg::SMEM_BYTES, s, ta, tw, ep, M, N, K);
  return 0;
}

// kind: 0 bf16 operands, 1 fp32 operands multiplied as tf32
inline int linear_tc_dispatch(int kind, const void* A, int64_t lda, const void* W, int64_t ldw,
                              const EpiParams& ep, int M, int N, int K, cudaStream_t s) {
  if (M <= 0 || N <= 0 || K <= 0) return fail(B2P_EINVAL, "linear: M, N, K must be positive");
  if (ep.out_bf16 && ((N & 1) || (ep.ldo & 1))) return fail(B2P_EINVAL, "linear: bf16 output needs even N and ldo");
  int dev = 0, sms = 148;
  B2P_CUDA(cudaGetDevice(&dev));
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (kind == 0) return launch_linear_tc<0, 128>(A, lda, W, ldw, ep, M, N, K, sms, s);
  if (kind == 1) return launch_linear_tc<1, 128>(A, lda, W, ldw, ep, M, N, K, sms, s);
  return fail(B2P_EINVAL, "linear: unknown operand kind %d", kind);
}

}  // namespace tc
}  // namespace b2p
