// Shim for the reference arm: grid-stride loop macro, block size constant and atomic add
// under the names the reference kernels use.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
namespace paddle {
namespace platform {
constexpr int PADDLE_CUDA_NUM_THREADS = 512;
#define CUDA_KERNEL_LOOP(i, n) \
  for (auto i = blockIdx.x * blockDim.x + threadIdx.x; i < (n); i += blockDim.x * gridDim.x)
// (plain overloads, not a template: the reference passes a double expression for a float target)
__device__ __forceinline__ float CudaAtomicAdd(float* addr, float val) { return atomicAdd(addr, val); }
__device__ __forceinline__ double CudaAtomicAdd(double* addr, double val) { return atomicAdd(addr, val); }
__device__ __forceinline__ int CudaAtomicAdd(int* addr, int val) { return atomicAdd(addr, val); }
__device__ __forceinline__ unsigned int CudaAtomicAdd(unsigned int* addr, unsigned int val) { return atomicAdd(addr, val); }
__device__ __forceinline__ unsigned long long CudaAtomicAdd(unsigned long long* addr, unsigned long long val) {
  return atomicAdd(addr, val);
}
}  // namespace platform
}  // namespace paddle
