// Fixed cost of a kernel launch as a function of what the kernel asks for (dynamic shared
// memory carve-out, TMEM allocation), measured with CUDA events around ONE launch that follows
// a long-running kernel (so the host is ahead of the device, as in a CUDA graph or a busy stream).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o launch_floor launch_floor.cu && ./launch_floor
#include <cstdio>
#include <cstdint>
#include <algorithm>
#include <vector>
#include <cuda_runtime.h>

__global__ void k_fill(float4* p, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    p[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}
__global__ void k_empty(int* out) { if (out && threadIdx.x == 9999) *out = 1; }
__global__ void k_smem(int* out) {
  extern __shared__ int s[];
  if (out && threadIdx.x == 9999) *out = s[threadIdx.x];
}
__global__ void k_tmem(int* out, int cols) {
  __shared__ uint32_t slot;
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(&slot)), "r"(cols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(slot), "r"(cols));
  if (out && threadIdx.x == 9999) *out = 1;
}

template <class F>
static float med(F launch, float4* buf, size_t n, bool flush) {
  std::vector<float> ts;
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  for (int it = 0; it < 25; ++it) {
    if (flush) k_fill<<<296, 1024>>>(buf, n);
    cudaEventRecord(a);
    launch();
    cudaEventRecord(b);
    cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, a, b);
    if (it >= 5) ts.push_back(ms * 1e3f);
  }
  std::sort(ts.begin(), ts.end());
  return ts[ts.size() / 2];
}

int main() {
  size_t n = (256u << 20) / 16;
  float4* buf; cudaMalloc(&buf, n * 16);
  cudaFuncSetAttribute(k_smem, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int flush = 0; flush < 2; ++flush) {
    printf("--- %s\n", flush ? "after a 256 MB fill kernel (dirty L2, host ahead)" : "idle device");
    printf("two events back to back      : %6.2f us\n", med([] {}, buf, n, flush));
    for (int grid : {32, 128, 148}) {
      printf("grid %3d empty, 320 thr      : %6.2f us\n", grid, med([&] { k_empty<<<grid, 320>>>(nullptr); }, buf, n, flush));
      printf("grid %3d 200KB dyn smem      : %6.2f us\n", grid, med([&] { k_smem<<<grid, 320, 200 * 1024>>>(nullptr); }, buf, n, flush));
      printf("grid %3d 48KB dyn smem       : %6.2f us\n", grid, med([&] { k_smem<<<grid, 320, 48 * 1024>>>(nullptr); }, buf, n, flush));
      printf("grid %3d tmem alloc 128 cols : %6.2f us\n", grid, med([&] { k_tmem<<<grid, 320>>>(nullptr, 128); }, buf, n, flush));
    }
    printf("2x (200KB smem, then empty)  : %6.2f us\n", med([&] { k_smem<<<128, 320, 200 * 1024>>>(nullptr); k_empty<<<128, 320>>>(nullptr);
                                                                  k_smem<<<128, 320, 200 * 1024>>>(nullptr); k_empty<<<128, 320>>>(nullptr); }, buf, n, flush));
    printf("4x empty                     : %6.2f us\n", med([&] { for (int i = 0; i < 4; ++i) k_empty<<<128, 320>>>(nullptr); }, buf, n, flush));
    printf("4x 200KB smem                : %6.2f us\n", med([&] { for (int i = 0; i < 4; ++i) k_smem<<<128, 320, 200 * 1024>>>(nullptr); }, buf, n, flush));
  }
  return 0;
}
