// Shim for the reference arm: memory::Allocation / AllocationPtr backed by a per-device
// caching pool (baseline/harness/ref_shim.cc), so that — as with Paddle's auto-growth
// allocator — steady-state memory::Alloc calls do not reach cudaMalloc.
#pragma once
#include <cstddef>
#include <memory>
#include "paddle/fluid/platform/place.h"
namespace paddle {
namespace memory {
namespace allocation {
class Allocation {
 public:
  Allocation(void* p, size_t size, size_t bucket, int dev) : ptr_(p), size_(size), bucket_(bucket), dev_(dev) {}
  void* ptr() const { return ptr_; }
  size_t size() const { return size_; }
  size_t bucket() const { return bucket_; }
  int dev() const { return dev_; }
 private:
  void* ptr_;
  size_t size_, bucket_;
  int dev_;
};
struct AllocationDeleter {
  void operator()(Allocation* a) const;
};
using AllocationPtr = std::unique_ptr<Allocation, AllocationDeleter>;
}  // namespace allocation
using allocation::Allocation;
using allocation::AllocationPtr;
}  // namespace memory
}  // namespace paddle
