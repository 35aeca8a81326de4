// Definitions behind the shim headers in baseline/shims (reference arm only; nothing of the
// product links this): flag globals, the caching device allocator, the device-context pool.
#include <cuda_runtime.h>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>

#include "paddle/fluid/memory/memory.h"
#include "paddle/fluid/platform/device_context.h"
#include "paddle/fluid/platform/enforce.h"

// ---- flags: the reference's defaults (paddle/fluid/platform/flags.cc:983-1006); like gflags,
// an environment variable FLAGS_<name> overrides the default.
static double env_d(const char* n, double d) { const char* e = getenv(n); return e ? atof(e) : d; }
static long long env_i(const char* n, long long d) { const char* e = getenv(n); return e ? atoll(e) : d; }
static bool env_b(const char* n, bool d) {
  const char* e = getenv(n);
  if (!e) return d;
  return !(strcmp(e, "0") == 0 || strcasecmp(e, "false") == 0);
}
double FLAGS_gpugraph_hbm_table_load_factor = env_d("FLAGS_gpugraph_hbm_table_load_factor", 0.75);
bool FLAGS_gpugraph_enable_gpu_direct_access = env_b("FLAGS_gpugraph_enable_gpu_direct_access", false);
bool FLAGS_gpugraph_enable_segment_merge_grads = env_b("FLAGS_gpugraph_enable_segment_merge_grads", false);
uint64_t FLAGS_gpugraph_merge_grads_segment_size = (uint64_t)env_i("FLAGS_gpugraph_merge_grads_segment_size", 128);
int32_t FLAGS_gpugraph_dedup_pull_push_mode = (int32_t)env_i("FLAGS_gpugraph_dedup_pull_push_mode", 0);
bool FLAGS_gpugraph_enable_hbm_table_collision_stat = env_b("FLAGS_gpugraph_enable_hbm_table_collision_stat", false);

namespace paddle {
namespace memory {

// Best-fit cache of freed blocks per device.  A request takes the smallest cached block that is
// at least as large and at most twice as large; otherwise it goes to cudaMalloc.  Sizes are
// rounded up so that the slightly different lengths of successive batches share blocks.
namespace {
struct Pool {
  std::mutex mu;
  std::multimap<size_t, void*> free_blocks;
};
Pool g_pools[64];
size_t round_size(size_t n) {
  if (n < 4096) return (n + 255) / 256 * 256;
  size_t g = 4096;
  while (g * 16 < n) g <<= 1;     // granularity = 1/16 .. 1/8 of the size
  return (n + g - 1) / g * g;
}
}  // namespace

AllocationPtr Alloc(const platform::Place& place, size_t size) {
  const int dev = place.GetDeviceId();
  const size_t want = round_size(size ? size : 1);
  Pool& p = g_pools[dev];
  {
    std::lock_guard<std::mutex> lk(p.mu);
    auto it = p.free_blocks.lower_bound(want);
    if (it != p.free_blocks.end() && it->first <= want * 2) {
      void* ptr = it->second;
      size_t b = it->first;
      p.free_blocks.erase(it);
      return AllocationPtr(new Allocation(ptr, size, b, dev));
    }
  }
  int cur = -1;
  cudaGetDevice(&cur);
  if (cur != dev) cudaSetDevice(dev);
  void* ptr = nullptr;
  cudaError_t e = cudaMalloc(&ptr, want);
  if (cur != dev) cudaSetDevice(cur);
  if (e != cudaSuccess) LOG(FATAL) << "memory::Alloc(" << size << ") on device " << dev << ": " << cudaGetErrorString(e);
  return AllocationPtr(new Allocation(ptr, size, want, dev));
}

namespace allocation {
void AllocationDeleter::operator()(Allocation* a) const {
  if (!a) return;
  Pool& p = g_pools[a->dev()];
  {
    std::lock_guard<std::mutex> lk(p.mu);
    p.free_blocks.emplace(a->bucket(), a->ptr());
  }
  delete a;
}
}  // namespace allocation
}  // namespace memory

namespace platform {
DeviceContextPool& DeviceContextPool::Instance() {
  static DeviceContextPool pool;
  return pool;
}
DeviceContext* DeviceContextPool::Get(const Place& place) { return &ctx_[place.GetDeviceId()]; }
}  // namespace platform
}  // namespace paddle

extern "C" void ref_set_compute_stream(int dev, void* stream) {
  auto* c = dynamic_cast<phi::GPUContext*>(
      paddle::platform::DeviceContextPool::Instance().Get(paddle::platform::CUDAPlace(dev)));
  c->set_stream((cudaStream_t)stream);
}
