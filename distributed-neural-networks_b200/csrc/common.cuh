// common.cuh -- error plumbing, launch accounting and small device helpers shared
// by every translation unit of libb200pipe.so.
#pragma once

#include <cuda_runtime.h>
#include <cuda.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <stdlib.h>
#include <atomic>
#include <string>

#include "../../include/b200pipe.h"

namespace b2p {

// ---- thread-local error string -------------------------------------------------
inline std::string& err_slot() {
  static thread_local std::string s;
  return s;
}

inline int fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  err_slot() = buf;
  return code;
}

#define B2P_CUDA(expr)                                                              \
  do {                                                                              \
    cudaError_t _e = (expr);                                                        \
    if (_e != cudaSuccess)                                                          \
      return ::b2p::fail(B2P_ECUDA, "%s failed: %s (%s:%d)", #expr,                 \
                         cudaGetErrorString(_e), __FILE__, __LINE__);               \
  } while (0)

#define B2P_TRY(expr)                                                               \
  do {                                                                              \
    int _rc = (expr);                                                               \
    if (_rc < 0) return _rc;                                                        \
  } while (0)

// ---- launch accounting (bench.py reports it as gpu_launches) ---------------------
inline std::atomic<int64_t>& launch_counter() {
  static std::atomic<int64_t> c{0};
  return c;
}

// Launch + count + check.  Usage: B2P_LAUNCH(kernel, grid, block, smem, stream, args...)
#define B2P_LAUNCH(kernel, grid, block, smem, stream, ...)                          \
  do {                                                                              \
    kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);                     \
    ::b2p::launch_counter().fetch_add(1, std::memory_order_relaxed);                \
    cudaError_t _e = cudaGetLastError();                                            \
    if (_e != cudaSuccess)                                                          \
      return ::b2p::fail(B2P_ECUDA, "launch of %s failed: %s (%s:%d)", #kernel,     \
                         cudaGetErrorString(_e), __FILE__, __LINE__);               \
  } while (0)

// ---- programmatic dependent launch (PDL) --------------------------------------------------
// The stages are chains of short kernels (a few microseconds each), so the ~2.5 us it takes to
// launch a grid and run its prologue (barrier init, TMEM allocation, weight staging) is a large
// fraction of every step.  Kernels launched through B2P_LAUNCH_PDL may start while their
// predecessor in the stream is still draining; they run their prologue and then block in
// pdl_wait() until the predecessor has completed and its writes are visible.  Everything a
// kernel does BEFORE pdl_wait() must not touch memory written by earlier kernels of the stream
// (constant weights are fine), and data produced by earlier kernels must be read with PLAIN loads after
// it: __ldg / ld.global.nc promises immutability, so the compiler may hoist such loads above the wait
// (observed: the fc tail read split-K partials the GEMM had not finished writing).
// pdl_trigger() lets the successor be scheduled early.
// B2P_NO_PDL=1 in the environment, or b2p_set_pdl(0), falls back to plain stream-ordered launches
// (wanted when two stages share one GPU on two streams: an early-started grid that sits in
// pdl_wait() holds its SMs' shared memory and keeps the other stream's kernels off them).
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// Loads of data written by an earlier kernel of the stream.  Volatile asm with a memory clobber keeps
// them behind pdl_wait(); a C++ load through a `const __restrict__` pointer may be turned into
// ld.global.nc and scheduled ahead of it.
__device__ __forceinline__ float4 ld_produced(const float4* p) {
  float4 v;
  asm volatile("ld.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ int ld_produced(const int* p) {
  int v;
  asm volatile("ld.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float ld_produced(const float* p) {
  float v;
  asm volatile("ld.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

inline std::atomic<int>& pdl_switch() {
  static std::atomic<int> on{[] {
    const char* e = getenv("B2P_NO_PDL");
    return (e && e[0] && e[0] != '0') ? 0 : 3;
  }()};
  return on;
}
// Per-call override (b2p_stage_set_pdl): every stage entry point installs its stage's mask for the duration of the
// call on the calling thread, so two pipelines in one process -- or two threads -- never see each other's setting.
// -1: no override, the process-wide default above applies.
inline int& pdl_call_mask() {
  static thread_local int m = -1;
  return m;
}
struct PdlScope {
  int prev;
  explicit PdlScope(int mask) : prev(pdl_call_mask()) { if (mask >= 0) pdl_call_mask() = mask & 7; }
  ~PdlScope() { pdl_call_mask() = prev; }
};
// bit 0: the light kernel chains (GEMMs, LayerNorm, attention, fc tail, splits); bit 1: the CIFAR part-0
// kernel, whose CTAs each take a whole SM's shared memory while they wait; bit 2 (per stage only, b2p_stage_set_pdl):
// the CIFAR part-0 kernel in ring mode does not wait for the preceding kernel of its stream at all (b200pipe.h).
inline bool pdl_enabled(int bit = 1) {
  const int call = pdl_call_mask();
  return ((call >= 0 ? call : pdl_switch().load(std::memory_order_relaxed)) & bit) != 0;
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_bit(int bit, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                                  Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = pdl_enabled(bit) ? 1 : 0;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

#define B2P_LAUNCH_PDL(kernel, grid, block, smem, stream, ...) \
  B2P_LAUNCH_PDL_BIT(1, kernel, grid, block, smem, stream, __VA_ARGS__)
#define B2P_LAUNCH_PDL_BIT(bit, kernel, grid, block, smem, stream, ...)                         \
  do {                                                                                          \
    cudaError_t _e = ::b2p::launch_pdl_bit(bit, kernel, dim3(grid), dim3(block), (smem), (stream), __VA_ARGS__); \
    ::b2p::launch_counter().fetch_add(1, std::memory_order_relaxed);                            \
    if (_e != cudaSuccess)                                                                      \
      return ::b2p::fail(B2P_ECUDA, "launch of %s failed: %s (%s:%d)", #kernel,                 \
                         cudaGetErrorString(_e), __FILE__, __LINE__);                           \
  } while (0)

// Scoped "make this device current, restore on exit".
struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
    if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

inline int ceil_div(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

// ---- device-side wait policy -----------------------------------------------------------------
// Every wait a kernel of this library performs on something ANOTHER kernel or GPU has to do (a handoff-ring flag,
// ring_hooks.cuh) gives up after g_wait_limit_ns: it records what it was waiting for in g_ring_error, returns, and
// from then on every ring wait on this GPU returns at once, so the stage kernels drain (with meaningless results)
// instead of hanging or killing the context.  The host reads and clears the record with b2p_ring_check().
// 0 = wait for ever (a serving process whose front door may sit idle).  Waits on the kernel's OWN pipeline
// (mbarriers) can only fail through a protocol bug; they trap, 10 s after the ring limit, and never when it is 0.
__device__ unsigned long long g_wait_limit_ns = 20000000000ull;
__device__ unsigned long long g_ring_error[4];   // [0] 0 = clean, else 1 + wait kind; [1] flag address; [2] target; [3] last value seen

// ---- device helpers ----------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// torch's ReLU and MaxPool2d propagate NaN (relu(NaN) = NaN); fmaxf would return the other operand
__device__ __forceinline__ float fmax_nan(float a, float b) {
  float r;
  asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}

// exact (erf) GELU, the nn.GELU() default the oracle pins (oracle/gpt_model.py)
__device__ __forceinline__ float gelu_erf(float x) {
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}

// Branch-free erf GELU for the tensor-core epilogues: erf by Abramowitz-Stegun 7.1.26
// (|error| <= 1.5e-7, far inside every tolerance in the test-suite) -- 2 MUFU + ~12 FP32 ops per
// element instead of erff's ~25 with divergent branches.  A GEMM epilogue evaluates one GELU per
// output element, so with short K (c_fc: K = C) the epilogue, not the MMA loop, was the limiter.
__device__ __forceinline__ float gelu_erf_fast(float x) {
  const float z = fabsf(x) * 0.70710678118654752440f;
  float t;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, z, 1.0f)));
  float poly = fmaf(1.061405429f, t, -1.453152027f);
  poly = fmaf(poly, t, 1.421413741f);
  poly = fmaf(poly, t, -0.284496736f);
  poly = fmaf(poly, t, 0.254829592f);
  poly *= t;
  float e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-z * z * 1.4426950408889634f));
  const float erf_abs = fmaf(-poly, e, 1.0f);           // erf(|x|/sqrt2)
  return 0.5f * x + 0.5f * fabsf(x) * erf_abs;          // 0.5 x (1 + sign(x) erf_abs)
}

}  // namespace b2p
