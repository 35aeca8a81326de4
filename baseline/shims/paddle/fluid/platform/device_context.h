// Shim for the reference arm: DeviceContextPool hands out one GPUContext per device whose
// stream() is the compute stream the harness registered for that device (the stream PyTorch
// runs the dense part on), which is the role Paddle's per-place compute stream plays.
#pragma once
#include <cuda_runtime.h>
#include <glog/logging.h>
#include <cassert>
#include "paddle/fluid/memory/memory.h"
#include "paddle/fluid/platform/enforce.h"
#include "paddle/fluid/platform/place.h"
namespace paddle {
namespace distributed {
class ValueAccessor;  // only ever named as a pointer type without PADDLE_WITH_PSCORE
}
namespace platform {
class DeviceContext {
 public:
  virtual ~DeviceContext() {}
};
}  // namespace platform
}  // namespace paddle
namespace phi {
class GPUContext : public paddle::platform::DeviceContext {
 public:
  cudaStream_t stream() const { return stream_; }
  void set_stream(cudaStream_t s) { stream_ = s; }
 private:
  cudaStream_t stream_{nullptr};
};
}  // namespace phi
namespace paddle {
namespace platform {
class DeviceContextPool {
 public:
  static DeviceContextPool& Instance();
  DeviceContext* Get(const Place& place);
 private:
  phi::GPUContext ctx_[64];
};
}  // namespace platform
}  // namespace paddle
