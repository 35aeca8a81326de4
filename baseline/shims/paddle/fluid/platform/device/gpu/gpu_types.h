// Shim for the reference arm: CUDA spellings of the gpu* type aliases.
#pragma once
#include <cuda_runtime.h>
using gpuStream_t = cudaStream_t;
using gpuError_t = cudaError_t;
using gpuEvent_t = cudaEvent_t;
namespace paddle {
using gpuStream_t = cudaStream_t;
}
