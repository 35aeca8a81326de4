// Shim for the reference arm: scoped cudaSetDevice.
#pragma once
#include <cuda_runtime.h>
#include "paddle/fluid/platform/place.h"
namespace paddle {
namespace platform {
class CUDADeviceGuard {
 public:
  explicit CUDADeviceGuard(int dev) { Set(dev); }
  explicit CUDADeviceGuard(const CUDAPlace& p) { Set(p.GetDeviceId()); }
  CUDADeviceGuard() {}
  ~CUDADeviceGuard() {
    if (prev_ >= 0) cudaSetDevice(prev_);
  }
  void SetDeviceIndex(int dev) { Set(dev); }
  void SetDevice(const CUDAPlace& p) { Set(p.GetDeviceId()); }
  CUDADeviceGuard(const CUDADeviceGuard&) = delete;
  CUDADeviceGuard& operator=(const CUDADeviceGuard&) = delete;
 private:
  void Set(int dev) {
    int cur = -1;
    cudaGetDevice(&cur);
    if (cur != dev) {
      if (prev_ < 0) prev_ = cur;
      cudaSetDevice(dev);
    }
  }
  int prev_{-1};
};
}  // namespace platform
}  // namespace paddle
