// ld_probe.cu -- stand-alone read-out of the tcgen05.ld fragment layouts (developer aid, not part of libb200pipe.so).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o ld_probe ld_probe.cu && ./ld_probe
//
// Tensor memory is filled through the plain 32x32b shape with value = lane * 256 + column; the same cells are then
// read back through .16x256b (x1, x2 and from lane offset 16) and every thread reports which (lane, column) each of
// its registers came from.  The part-0 kernel's conv2 drain wants rows r and r+8 of a 16-lane group in ONE thread.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void probe(uint32_t* out) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(&slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = slot;
  const uint32_t row = base + ((uint32_t)(warp * 32) << 16);
  for (int c0 = 0; c0 < 64; c0 += 8) {
    uint32_t v[8];
    for (int j = 0; j < 8; ++j) v[j] = (uint32_t)(threadIdx.x * 256 + c0 + j);
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(row + c0), "r"(v[0]),
                 "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
  }
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  uint32_t* o = out + threadIdx.x * 32;
  {  // 16x256b.x1 from the warp's first lane
    uint32_t r[4];
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(row) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int j = 0; j < 4; ++j) o[j] = r[j];
  }
  {  // 16x256b.x1 from lane offset 16
    uint32_t r[4];
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(row + (16u << 16)) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int j = 0; j < 4; ++j) o[4 + j] = r[j];
  }
  {  // 16x256b.x2 from column 16
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(row + 16) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int j = 0; j < 8; ++j) o[8 + j] = r[j];
  }
  {  // 16x128b.x2 from column 0
    uint32_t r[4];
    asm volatile("tcgen05.ld.sync.aligned.16x128b.x2.b32 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(row) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int j = 0; j < 4; ++j) o[16 + j] = r[j];
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(base) : "memory");
}

static void show(const char* what, const uint32_t* h, int first, int count) {
  printf("%s\n", what);
  for (int w = 0; w < 4; w += 3)
    for (int t = 0; t < 32; ++t) {
      printf("  warp %d thread %2d:", w, t);
      for (int j = 0; j < count; ++j) {
        const uint32_t v = h[(w * 32 + t) * 32 + first + j];
        printf("  r%d=(lane %3u, col %2u)", j, v >> 8, v & 255);
      }
      printf("\n");
    }
}

int main() {
  uint32_t* d;
  CK(cudaMalloc(&d, 128 * 32 * 4));
  CK(cudaMemset(d, 0xff, 128 * 32 * 4));
  probe<<<1, 128>>>(d);
  CK(cudaDeviceSynchronize());
  uint32_t h[128 * 32];
  CK(cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost));
  show("16x256b.x1 at the warp's lane 0, column 0", h, 0, 4);
  show("16x256b.x1 at the warp's lane 16, column 0", h, 4, 4);
  show("16x256b.x2 at the warp's lane 0, column 16", h, 8, 8);
  show("16x128b.x2 at the warp's lane 0, column 0", h, 16, 4);
  return 0;
}
