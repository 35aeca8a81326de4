// Mini-batch construction on the device (W1 of SURVEY.md §8).
// The host packs a batch record-major into ONE pinned staging buffer (keys of record 0,
// record 1, ... plus per-record slot offsets) and issues ONE H2D copy; this kernel turns
// it into the slot-major layout every sparse op consumes: keys[N] ordered by (slot, ins)
// and lod[S*B+1].  No D2H round trip, no per-slot tensors.
// Reference: MiniBatchGpuPack (6-9 H2D copies + sync, data_feed.cc:4896-4940) +
// FillSlotValueOffsetPadBoxKernel (serial per-slot scan) + D2H of the offsets +
// CopyForTensorPadBoxKernel (data_feed.cu:1208-1384, data_feed.cc:3400-3522).
#include "common.cuh"
#include <cooperative_groups.h>
#include <mutex>
namespace cg = cooperative_groups;

namespace {

constexpr int BB_THREADS = 1024;

__device__ __forceinline__ int bag_len(const int32_t* ro, int S, int B, int g) {
  const int s = g / B, b = g % B;
  const int32_t* r = ro + (int64_t)b * (S + 1) + s;
  return r[1] - r[0];
}

// block-wide exclusive scan of one int per thread; returns exclusive prefix, *total = block sum
__device__ __forceinline__ int block_exscan(int v, int* total) {
  __shared__ int warp_sums[BB_THREADS / 32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int inc = v;
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) warp_sums[w] = inc;
  __syncthreads();
  if (w == 0) {
    int ws = lane < BB_THREADS / 32 ? warp_sums[lane] : 0;
    int winc = ws;
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += t;
    }
    if (lane < BB_THREADS / 32) warp_sums[lane] = winc - ws;  // exclusive warp offsets
    if (lane == BB_THREADS / 32 - 1) *total = winc;
  }
  __syncthreads();
  int res = inc - v + warp_sums[w];
  __syncthreads();
  return res;
}

__global__ void __launch_bounds__(BB_THREADS)
k_batch_build(const uint64_t* __restrict__ rec_keys, const int32_t* __restrict__ rec_off, int B, int S,
              uint64_t* __restrict__ keys_out, int32_t* __restrict__ lod_out, int32_t* __restrict__ block_sums,
              const float* __restrict__ dense_src, float* __restrict__ dense_dst, int64_t n_dense) {
  cg::grid_group grid = cg::this_grid();
  // optional: the batch's dense block (labels + dense features) moves with the same launch
  for (int64_t i = blockIdx.x * (int64_t)BB_THREADS + threadIdx.x; i < n_dense; i += (int64_t)gridDim.x * BB_THREADS)
    dense_dst[i] = dense_src[i];
  const int G = S * B;
  const int per_block = (G + gridDim.x - 1) / gridDim.x;
  const int g0 = blockIdx.x * per_block, g1 = min(G, g0 + per_block);
  __shared__ int s_total;
  __shared__ int s_carry;

  // phase 1: per-block total length.  When a block's share fits one bag per thread (the usual
  // case: G <= blocks * 1024) the lengths and their in-block prefix stay in registers and phase 3
  // neither re-reads the offsets nor scans again.
  const bool single = per_block <= BB_THREADS;
  int len0 = 0, ex0 = 0;
  if (single) {
    const int g = g0 + threadIdx.x;
    len0 = g < g1 ? bag_len(rec_off, S, B, g) : 0;
    ex0 = block_exscan(len0, &s_total);
  } else {
    int local = 0;
    for (int g = g0 + threadIdx.x; g < g1; g += BB_THREADS) local += bag_len(rec_off, S, B, g);
    int dummy = block_exscan(local, &s_total);
    (void)dummy;
  }
  const int my_total = s_total;
  if (threadIdx.x == 0) block_sums[blockIdx.x] = my_total;
  grid.sync();

  // phase 2: every block derives its base from the (few) block sums
  int base = 0;
  for (int i = threadIdx.x; i < (int)blockIdx.x; i += BB_THREADS) base += block_sums[i];
  int unused = block_exscan(base, &s_total);
  (void)unused;
  if (threadIdx.x == 0) s_carry = s_total;
  __syncthreads();

  // phase 3: lod + key scatter
  if (single) {
    const int g = g0 + threadIdx.x;
    const int start = s_carry + ex0;
    if (g < g1) {
      lod_out[g] = start;
      const int s = g / B, b = g % B;
      const int32_t src = rec_off[(int64_t)b * (S + 1) + s];
      for (int j = 0; j < len0; ++j) keys_out[start + j] = rec_keys[src + j];
    }
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) lod_out[G] = s_carry + my_total;
    return;
  }
  for (int c0 = g0; c0 < g1; c0 += BB_THREADS) {
    const int g = c0 + threadIdx.x;
    const int len = g < g1 ? bag_len(rec_off, S, B, g) : 0;
    const int ex = block_exscan(len, &s_total);
    const int start = s_carry + ex;
    if (g < g1) {
      lod_out[g] = start;
      const int s = g / B, b = g % B;
      const int32_t src = rec_off[(int64_t)b * (S + 1) + s];
      for (int j = 0; j < len; ++j) keys_out[start + j] = rec_keys[src + j];
    }
    __syncthreads();
    if (threadIdx.x == 0) s_carry += s_total;
    __syncthreads();
  }
  if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) lod_out[G] = s_carry;
}

}  // namespace

extern "C" {

// d_rec_off: [B, S+1] global offsets into d_rec_keys (record-major CSR).
// d_scratch: >= 1024 int32 of scratch for the cross-block scan.
int pbx_batch_build_dense(const uint64_t* d_rec_keys, const int32_t* d_rec_off, int B, int S, uint64_t* d_keys_out,
                          int32_t* d_lod_out, int32_t* d_scratch, const float* d_dense_src, float* d_dense_dst,
                          int64_t n_dense, cudaStream_t st);
int pbx_batch_build(const uint64_t* d_rec_keys, const int32_t* d_rec_off, int B, int S, uint64_t* d_keys_out,
                    int32_t* d_lod_out, int32_t* d_scratch, cudaStream_t st) {
  return pbx_batch_build_dense(d_rec_keys, d_rec_off, B, S, d_keys_out, d_lod_out, d_scratch, nullptr, nullptr, 0, st);
}

// The same launch also copies n_dense floats d_dense_src -> d_dense_dst (the batch's label + dense block).
int pbx_batch_build_dense(const uint64_t* d_rec_keys, const int32_t* d_rec_off, int B, int S, uint64_t* d_keys_out,
                          int32_t* d_lod_out, int32_t* d_scratch, const float* d_dense_src, float* d_dense_dst,
                          int64_t n_dense, cudaStream_t st) {
  if (B <= 0 || S <= 0) return 0;
  int G = S * B;
  int blocks = (G + BB_THREADS - 1) / BB_THREADS;   // one bag per thread while the SMs last
  int sms = pbx_num_sms();
  if (blocks > sms) blocks = sms;
  if (blocks < 1) blocks = 1;
  if (!d_dense_src || !d_dense_dst) n_dense = 0;
  void* args[] = {(void*)&d_rec_keys, (void*)&d_rec_off, (void*)&B, (void*)&S, (void*)&d_keys_out,
                  (void*)&d_lod_out, (void*)&d_scratch, (void*)&d_dense_src, (void*)&d_dense_dst, (void*)&n_dense};
  PBX_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)k_batch_build, dim3(blocks), dim3(BB_THREADS), args, 0, st));
  pbx_note_launch(1);
  return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------
// Allocation-free variant.
//
// pbx_batch_build() needs a small device scratch array (one partial sum per thread block)
// that its caller has to allocate.  Some callers are not in a position to allocate device
// memory at all:
//   * a data-feed thread that prepares the next batch while training kernels are in flight —
//     cudaMalloc / cudaFree synchronise the whole device, so one call there stalls every
//     stream of the training step;
//   * a process whose HBM is already fully committed (the sparse table and the symmetric
//     buffers are sized to fill the card), where a late allocation can simply fail;
//   * a host application that embeds this library under a fixed memory budget and forbids
//     allocator calls from the code it loads.
// For them the scratch lives in the library image itself: a file-scope __device__ array that
// every GPU which loads libpbx_cuda.so carries (2 KiB per GPU) until the process exits.
//
// The array's name is only a host-side handle in host code — handing it to a kernel as if it
// were a pointer would make the kernel fault (or, where the GPU can reach host memory, scribble
// somewhere unrelated).  The real device address is obtained once per device with
// cudaGetSymbolAddress() while that device is current, and that address is what the kernel gets.
#define PBX_BB_STATIC_SCRATCH_INTS 512                       /* 2 KiB */
#define PBX_BB_STATIC_BAGS_PER_BLOCK 2048                    /* bags per thread block (before the SM-count cap) */
#define PBX_BB_STATIC_MAX_BAGS ((long long)PBX_BB_STATIC_SCRATCH_INTS * PBX_BB_STATIC_BAGS_PER_BLOCK)

static __device__ __align__(256) int32_t g_bb_static_scratch[PBX_BB_STATIC_SCRATCH_INTS];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_bb_static_mu;
int32_t* g_bb_static_ptr[kMaxDevices] = {nullptr};

// Resolves (once per device) the device address of g_bb_static_scratch for the current device.
int bb_static_scratch_for_current_device(int32_t** out) {
  int dev = -1;
  PBX_CUDA_CHECK(cudaGetDevice(&dev));
  PBX_REQUIRE(dev >= 0 && dev < kMaxDevices, "pbx_batch_build_static: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_bb_static_mu);
  if (g_bb_static_ptr[dev] == nullptr) {
    int32_t* p = nullptr;
    PBX_CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_bb_static_scratch));
    PBX_REQUIRE(p != nullptr, "pbx_batch_build_static: cudaGetSymbolAddress returned a null address");
    g_bb_static_ptr[dev] = p;
  }
  *out = g_bb_static_ptr[dev];
  return 0;
}
}  // namespace

extern "C" {

// Same result as pbx_batch_build(), without a caller-provided scratch buffer and without any
// allocation (no cudaMalloc / cudaFree / allocating library routine) inside the call.
//
//   * Limit: B * S (instances x sparse slots) must not exceed 1,048,576
//     (= 512 scratch words x 2048 bags per block); larger batches are rejected with an error
//     and must use pbx_batch_build() with their own scratch.  The number of keys is not limited.
//   * First call on each device: looks up the device address of the built-in scratch array
//     (cudaGetSymbolAddress with that device current).  Make that first call ahead of time,
//     before any stream capture is started on that device.
//   * The scratch is shared by every call on the device and keeps whatever the previous call
//     left in it.  Nothing in it has to start at zero: each thread block overwrites its own
//     word before any block reads it, and words beyond the current grid are never read.
//   * Because the array is shared, two calls that can overlap on the same GPU (different
//     streams, or different host threads without ordering between their streams) race on it
//     and corrupt each other's offsets.  Order such calls on one stream or with events.
int pbx_batch_build_static(const uint64_t* d_rec_keys, const int32_t* d_rec_off, int B, int S,
                           uint64_t* d_keys_out, int32_t* d_lod_out, cudaStream_t st) {
  if (B <= 0 || S <= 0) return 0;
  const long long G = (long long)S * B;
  PBX_REQUIRE(G <= PBX_BB_STATIC_MAX_BAGS,
              "pbx_batch_build_static: B*S = %lld exceeds the built-in scratch limit of %lld bags "
              "(use pbx_batch_build with caller-provided scratch)", G, PBX_BB_STATIC_MAX_BAGS);
  int32_t* scratch = nullptr;
  int rc = bb_static_scratch_for_current_device(&scratch);
  if (rc) return rc;
  int blocks = (int)((G + PBX_BB_STATIC_BAGS_PER_BLOCK - 1) / PBX_BB_STATIC_BAGS_PER_BLOCK);
  int sms = pbx_num_sms();
  if (blocks > sms) blocks = sms;
  if (blocks < 1) blocks = 1;   // blocks <= 512 = PBX_BB_STATIC_SCRATCH_INTS by the check above
  const float* no_src = nullptr;
  float* no_dst = nullptr;
  int64_t no_n = 0;
  void* args[] = {(void*)&d_rec_keys, (void*)&d_rec_off, (void*)&B, (void*)&S, (void*)&d_keys_out,
                  (void*)&d_lod_out, (void*)&scratch, (void*)&no_src, (void*)&no_dst, (void*)&no_n};
  PBX_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)k_batch_build, dim3(blocks), dim3(BB_THREADS), args, 0, st));
  pbx_note_launch(1);
  return 0;
}

}  // extern "C"
