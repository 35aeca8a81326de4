// Shim for the reference arm: memory::Alloc (caching, see allocator.h) and memory::Copy.
#pragma once
#include <cuda_runtime.h>
#include "paddle/fluid/memory/allocation/allocator.h"
namespace paddle {
namespace memory {
AllocationPtr Alloc(const platform::Place& place, size_t size);
inline void Copy(const platform::Place& dst_place, void* dst, const platform::Place& src_place, const void* src,
                 size_t n, cudaStream_t stream) {
  (void)dst_place; (void)src_place;
  if (stream) cudaMemcpyAsync(dst, src, n, cudaMemcpyDefault, stream);
  else cudaMemcpy(dst, src, n, cudaMemcpyDefault);
}
}  // namespace memory
}  // namespace paddle
