// p2p_bulk_probe.cu -- stand-alone NVLink store-rate probe (developer aid, not part of libb200pipe.so).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o p2p_bulk_probe p2p_bulk_probe.cu && ./p2p_bulk_probe
//
// The stage-0 kernel sends every finished sample (16 KB) from shared memory into the next GPU's ring slot with one
// cp.async.bulk.  This probe asks what the link gives for exactly that pattern: 148 CTAs, each sending 16 KB tiles
// back to back from shared memory to (a) its own GPU, (b) the peer, with 1..4 copies in flight per CTA, and the same
// bytes through plain 16-byte stores from registers.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int INFLIGHT>
__global__ void bulk_store(float* dst, int tiles_per_cta, int tile_pitch_ctas) {
  extern __shared__ __align__(128) unsigned char sm[];
  for (int i = threadIdx.x; i < 4 * 4096; i += blockDim.x) reinterpret_cast<float*>(sm)[i] = (float)i;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int t = 0; t < tiles_per_cta; ++t) {
      float* d = dst + ((size_t)t * tile_pitch_ctas + blockIdx.x) * 4096;
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(d), "r"(smem_u32(sm + (t & 3) * 16384)), "r"(16384) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group %0;" ::"n"(INFLIGHT - 1) : "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

__global__ void reg_store(float4* dst, int tiles_per_cta, int tile_pitch_ctas) {
  const float4 v = make_float4(1.f, 2.f, 3.f, (float)threadIdx.x);
  for (int t = 0; t < tiles_per_cta; ++t) {
    float4* d = dst + ((size_t)t * tile_pitch_ctas + blockIdx.x) * 1024;
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) d[i] = v;
  }
}

template <typename F>
static float time_ms(F&& launch, int reps) {
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  launch();
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(a));
  for (int i = 0; i < reps; ++i) launch();
  CK(cudaEventRecord(b));
  CK(cudaDeviceSynchronize());
  float ms;
  CK(cudaEventElapsedTime(&ms, a, b));
  return ms / reps;
}

int main() {
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  const int ctas = 148, tiles = 64;                       // 148 x 64 x 16 KB = 155 MB per launch
  const size_t bytes = (size_t)ctas * tiles * 16384;
  float *local = nullptr, *peer = nullptr;
  CK(cudaSetDevice(0));
  CK(cudaMalloc(&local, bytes));
  if (ndev > 1) {
    CK(cudaSetDevice(1));
    CK(cudaMalloc(&peer, bytes));
    CK(cudaSetDevice(0));
    CK(cudaDeviceEnablePeerAccess(1, 0));
  }
  CK(cudaFuncSetAttribute(bulk_store<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
  CK(cudaFuncSetAttribute(bulk_store<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
  CK(cudaFuncSetAttribute(bulk_store<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
  for (int where = 0; where < (peer ? 2 : 1); ++where) {
    float* dst = where ? peer : local;
    const char* name = where ? "peer (NVLink)" : "own GPU";
    float ms;
    ms = time_ms([&] { bulk_store<1><<<ctas, 128, 65536>>>(dst, tiles, ctas); }, 5);
    printf("%-14s bulk 16 KB, 1 in flight per CTA: %7.1f GB/s\n", name, bytes / ms * 1e-6);
    ms = time_ms([&] { bulk_store<2><<<ctas, 128, 65536>>>(dst, tiles, ctas); }, 5);
    printf("%-14s bulk 16 KB, 2 in flight per CTA: %7.1f GB/s\n", name, bytes / ms * 1e-6);
    ms = time_ms([&] { bulk_store<4><<<ctas, 128, 65536>>>(dst, tiles, ctas); }, 5);
    printf("%-14s bulk 16 KB, 4 in flight per CTA: %7.1f GB/s\n", name, bytes / ms * 1e-6);
    ms = time_ms([&] { reg_store<<<ctas, 512>>>(reinterpret_cast<float4*>(dst), tiles, ctas); }, 5);
    printf("%-14s 16-byte stores, 512 threads per CTA: %7.1f GB/s\n", name, bytes / ms * 1e-6);
    ms = time_ms([&] { reg_store<<<ctas * 4, 512>>>(reinterpret_cast<float4*>(dst), tiles / 4, ctas * 4); }, 5);
    printf("%-14s 16-byte stores, 4 x 512 threads per SM: %7.1f GB/s\n", name, bytes / ms * 1e-6);
  }
  return 0;
}
