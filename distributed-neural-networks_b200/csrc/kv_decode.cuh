// kv_decode.cuh -- key/value cache for greedy decoding through the GPT parts.
//
// No counterpart in the reference: gpt_model_parts.py:13-50 only ever runs one full-sequence forward.  The math is
// the block's attention (public nanoGPT semantics, see oracle/gpt_model.py): for the new position p of a sequence,
//     y_p = softmax(q_p . K_{0..p}^T / sqrt(64)) V_{0..p}
// with K, V of the earlier positions kept per block instead of recomputed.  Head size 64 like the other kernels.
//
// Cache layout per block: K and V as [B][H][Tmax][64] fp32 (rows of 256 B), Tmax = block_size.  The qkv buffer the
// c_attn linear writes is [M][3C] in fp32 (fp32 / tf32 / tf32x3 modes) or bf16 (bf16 mode); values are widened to
// fp32 on the way into the cache, so one cache serves every math mode.
#pragma once
#include <cuda_bf16.h>

#include "common.cuh"

namespace b2p {
namespace kv {

__device__ __forceinline__ float to_f32(float v) { return v; }
__device__ __forceinline__ float to_f32(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ void from_f32(float* p, float v) { *p = v; }
__device__ __forceinline__ void from_f32(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

// k, v of rows (b, t) -> cache position pos0 + t.  One thread per (row, head, 4 dims).
template <typename T>
__global__ void __launch_bounds__(256)
kv_append(const T* qkv, float* kc, float* vc,      // no __restrict__: qkv comes from the previous kernel (PDL), plain loads stay behind the wait
          int B, int Tn, int C, int H,
          int Tmax, int pos0, const int* pos_dev) {      // pos_dev != null: the position comes from device memory (CUDA-graph decoding)
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t total = (int64_t)B * Tn * H * 16;
  pdl_trigger();
  pdl_wait();
  if (pos_dev) pos0 = ld_produced(pos_dev);
  if (i >= total || pos0 < 0 || pos0 + Tn > Tmax) return;      // (a position beyond the cache: nothing is stored)
  const int d4 = (int)(i % 16), h = (int)((i / 16) % H);
  const int64_t row = i / (16 * H);
  const int b = (int)(row / Tn), t = (int)(row % Tn);
  const T* src = qkv + row * 3 * C + h * 64 + d4 * 4;
  const int64_t dst = (((int64_t)b * H + h) * Tmax + pos0 + t) * 64 + d4 * 4;
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    kc[dst + e] = to_f32(src[C + e]);
    vc[dst + e] = to_f32(src[2 * C + e]);
  }
}

// One CTA of 512 threads per (sequence, head); the single query sits at position `pos` (keys 0..pos are in the cache).
//   pass 1: thread t scores keys t, t+512, ... (64-term dot products, float4 loads of the 256-byte key rows)
//   pass 2: soft-max over the scores in shared memory (max, exp, sum: fp32, like the causal kernels)
//   pass 3: threads (slice, d): an eighth of the keys each, one output dim -- coalesced 256-byte reads of the value rows
// (With 128 threads -- two key slices in pass 3 -- a call took 31 us for 512 cached keys, 12 calls per decode step: each
//  thread walked 256 value rows one dependent FMA after the other.)
constexpr int DEC_THREADS = 512, DEC_SLICES = DEC_THREADS / 64;
template <typename T>
__global__ void __launch_bounds__(DEC_THREADS)
attention_decode(const T* qkv, const float* kc, const float* vc, T* y, int C, int H, int Tmax, int pos, const int* pos_dev) {
  extern __shared__ float sc[];                       // [Tmax] scores, then 64 q values, then scratch
  float* q = sc + Tmax;
  float* red = q + 64;                                 // [DEC_THREADS]
  const int h = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
  pdl_trigger();
  pdl_wait();
  if (pos_dev) pos = ld_produced(pos_dev);             // CUDA-graph decoding: the position lives in device memory
  const int n = min(max(pos, 0), Tmax - 1) + 1;
  if (tid < 64) q[tid] = to_f32(qkv[(int64_t)b * 3 * C + h * 64 + tid]) * 0.125f;       // 1/sqrt(64), exact
  __syncthreads();
  const float* kb = kc + ((int64_t)b * H + h) * Tmax * 64;
  const float* vb = vc + ((int64_t)b * H + h) * Tmax * 64;
  float mx = -INFINITY;
  for (int j = tid; j < n; j += DEC_THREADS) {
    const float4* kr = reinterpret_cast<const float4*>(kb + (int64_t)j * 64);
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const float4 k4 = kr[i];
      s = fmaf(q[4 * i], k4.x, s); s = fmaf(q[4 * i + 1], k4.y, s);
      s = fmaf(q[4 * i + 2], k4.z, s); s = fmaf(q[4 * i + 3], k4.w, s);
    }
    sc[j] = s;
    mx = fmaxf(mx, s);
  }
  red[tid] = mx;
  __syncthreads();
  for (int o = DEC_THREADS / 2; o > 0; o >>= 1) {
    if (tid < o) red[tid] = fmaxf(red[tid], red[tid + o]);
    __syncthreads();
  }
  mx = red[0];
  __syncthreads();
  float sum = 0.f;
  for (int j = tid; j < n; j += DEC_THREADS) {
    const float p = expf(sc[j] - mx);
    sc[j] = p;
    sum += p;
  }
  red[tid] = sum;
  __syncthreads();
  for (int o = DEC_THREADS / 2; o > 0; o >>= 1) {
    if (tid < o) red[tid] += red[tid + o];
    __syncthreads();
  }
  const float inv = 1.0f / red[0];
  __syncthreads();
  const int d = tid & 63, slice = tid >> 6;
  float acc = 0.f;
  for (int j = slice; j < n; j += DEC_SLICES) acc = fmaf(sc[j], vb[(int64_t)j * 64 + d], acc);
  red[tid] = acc;
  __syncthreads();
  if (tid < 64) {
    float o = red[tid];
#pragma unroll
    for (int sl = 1; sl < DEC_SLICES; ++sl) o += red[tid + 64 * sl];
    from_f32(y + (int64_t)b * C + h * 64 + tid, o * inv);
  }
}

}  // namespace kv
}  // namespace b2p
