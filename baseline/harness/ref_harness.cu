// Driver of the reference arm (bench.py --impl reference): a C ABI over the reference's own
// open sparse engine, HeterPS, compiled unmodified from baseline/_ref/src (see
// baseline/build_ref.py).  Nothing here comes from, or links, the product libraries.
//
// What this file does is what PSGPUWrapper does around HeterPs in the reference — the thin host
// orchestration of fleet/ps_gpu_wrapper.cc, which itself cannot be compiled standalone because
// its header drags in the dataset, gloo, brpc and protobuf:
//   * set-up        : InitializeGPU (ps_gpu_wrapper.h:221-292) + BuildGPUTask (ps_gpu_wrapper.cc:684-883):
//                     HeterPsResource + enable_p2p, HeterPsBase::get_instance, set_sparse_sgd /
//                     set_embedx_sgd, one HBMMemoryPool per GPU, build_ps in 500000-key chunks on 2 streams
//   * pull per batch: PSGPUWrapper::PullSparse, default FLAGS_gpugraph_dedup_pull_push_mode == 0 branch
//                     (ps_gpu_wrapper.cc:1186-1250): CopyKeys -> HeterPs::pull_sparse -> accessor CopyForPull
//   * push per batch: PSGPUWrapper::PushSparseGrad, same branch (ps_gpu_wrapper.cc:1484-1512):
//                     accessor CopyForPush -> HeterPs::push_sparse
// The engine calls (HeterPs / HeterComm / HashTable / accessor kernels / optimizer functors) are
// the reference's code; only the table CONTENT is synthetic (keys = the bench's key function,
// values initialised on the device instead of being fetched from a CPU parameter server).
#include <cuda_runtime.h>
#include <nccl.h>

#include <memory>
#include <mutex>
#include <numeric>
#include <string>
#include <unordered_map>
#include <vector>

#include "paddle/fluid/framework/fleet/heter_ps/heter_ps_base.h"
#include "paddle/fluid/framework/fleet/heter_ps/mem_pool.h"
#include "paddle/fluid/memory/memory.h"

using namespace paddle::framework;  // NOLINT
namespace platform = paddle::platform;
namespace memory = paddle::memory;

#define REF_CUDA(x)                                                                        \
  do {                                                                                     \
    cudaError_t e__ = (x);                                                                 \
    if (e__ != cudaSuccess) {                                                              \
      snprintf(g_err, sizeof(g_err), "%s:%d %s: %s", __FILE__, __LINE__, #x, cudaGetErrorString(e__)); \
      return -1;                                                                           \
    }                                                                                      \
  } while (0)

static char g_err[1024];

namespace {

struct DevState {
  memory::AllocationPtr total_keys;   // keys_tensor[dev] of PSGPUWrapper: kept from pull to push
  int64_t total_length = 0;
  HBMMemoryPool* hbm_pool = nullptr;
  int64_t n_keys = 0;
};

struct RefPs {
  int ndev = 0;
  int mf_dim = 0;
  std::vector<int> dev_ids;
  std::shared_ptr<HeterPsResource> resource;
  HeterPsBase* heter_ps = nullptr;
  OptimizerConfig opt;
  std::vector<DevState> dev;
  std::vector<ncclComm_t> comms;
  std::vector<int> slot_vector, slot_mf_dim_vector, slot_dim;
};

__device__ __forceinline__ uint64_t mix64(uint64_t k) {
  k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
  return k;
}
// the bench's synthetic key of index i (same function as the product arm uses, so that both
// arms look up the same key set): mix64(i * golden + seed) >> 1
__device__ __forceinline__ uint64_t synth_key(uint64_t i, uint64_t seed) {
  return mix64(i * 0x9E3779B97F4A7C15ull + seed) >> 1;
}

// keys of the synthetic table that live on shard `shard` (reference sharding: key % ngpu)
__global__ void k_collect_keys(uint64_t n_total, uint64_t seed, int shard, int nshard, uint64_t* out,
                               unsigned long long* counter, unsigned long long cap) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n_total;
       i += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t k = synth_key(i, seed);
    if ((int)(k % (uint64_t)nshard) == shard) {
      const unsigned long long p = atomicAdd(counter, 1ull);
      if (p < cap) out[p] = k;
    }
  }
}

// initial feature values in the HBM pool (stands in for BuildFill from the CPU table): a key
// that has been seen (show 20, click 4), with its embedx created
__global__ void k_init_values(char* pool, size_t fv_size, uint64_t n, const uint64_t* keys,
                              CommonFeatureValueAccessor acc, int mf_dim, int create_mf) {
  auto& fv = acc.common_feature_value;
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    float* v = reinterpret_cast<float*>(pool + i * fv_size);
    const uint64_t h = mix64(keys[i] ^ 0x5bd1e995ull);
    *reinterpret_cast<uint64_t*>(v + fv.CpuPtrIndex()) = 0;
    v[fv.DeltaScoreIndex()] = 0.f;
    v[fv.ShowIndex()] = 20.f;
    v[fv.ClickIndex()] = 4.f;
    v[fv.EmbedWIndex()] = ((float)(h & 0xffff) / 65536.f - 0.5f) * 0.02f;
    for (int j = 0; j < fv.EmbedDim(); ++j) v[fv.EmbedG2SumIndex() + j] = 0.f;
    v[fv.SlotIndex()] = 0.f;
    v[fv.MfDimIndex()] = (float)mf_dim;
    const int mfs = fv.MFSize(mf_dim) / sizeof(float);
    v[fv.MfSizeIndex()] = create_mf ? (float)mfs : 0.f;
    for (int j = 0; j < mfs; ++j) v[fv.EmbedxG2SumIndex() + j] = 0.f;
    if (create_mf)
      for (int j = 0; j < mf_dim; ++j)
        v[fv.EmbedxWIndex() + j] = ((float)((mix64(h + j) >> 40) & 0xffff) / 65536.f - 0.5f) * 2e-4f;
  }
}

// CopyKeysKernel of fleet/ps_gpu_wrapper.cu:34-53 restated (that file cannot be compiled standalone):
// gathers the per-slot key tensors into one flat array, slot found by binary search on the lod
__global__ void k_copy_keys(uint64_t** src_keys, uint64_t* dest_total_keys, const int64_t* len, int slot_num,
                            int total_len) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total_len; i += blockDim.x * gridDim.x) {
    int low = 0, high = slot_num - 1;
    while (low < high) {
      int mid = (low + high) / 2;
      if (i < len[mid]) high = mid; else low = mid + 1;
    }
    int y = i - (low ? len[low - 1] : 0);
    dest_total_keys[i] = src_keys[low][y];
  }
}

}  // namespace

extern "C" {

const char* ref_last_error() { return g_err; }

// ndev GPUs 0..ndev-1 driven by this one process (the reference's process model,
// boxps_trainer.cc:58-79); optimizer_type 1 = SparseAdagrad (heter_ps.cu:36-38)
void* ref_create(int ndev, int mf_dim, int n_slots, int optimizer_type) {
  RefPs* r = new RefPs();
  r->ndev = ndev;
  r->mf_dim = mf_dim;
  for (int i = 0; i < ndev; ++i) r->dev_ids.push_back(i);
  std::unordered_map<std::string, float> cfg;
  cfg["optimizer_type"] = (float)optimizer_type;
  cfg["embedx_dim"] = (float)mf_dim;
  GlobalAccessorFactory::GetInstance().Init("CtrDymfAccessor");
  GlobalAccessorFactory::GetInstance().GetAccessorWrapper()->Configure(cfg);
  r->resource = std::make_shared<HeterPsResource>(r->dev_ids);
  r->resource->enable_p2p();
  r->resource->set_multi_mf(1, mf_dim);
  r->dev.resize(ndev);
  // reference defaults, heter_ps/optimizer_conf.h:20-45 (also what the product arm's table uses)
  r->opt.set_sparse_sgd(0.1f, 1.0f, -10.f, 10.f, 0.05f, 3.0f, 1e-4f, 0.9f, 0.999f, 1e-8f);
  r->opt.set_embedx_sgd(10.f, 0.05f, 3.0f, 1e-4f, -10.f, 10.f, 0.9f, 0.999f, 1e-8f, 9008.f, 0.05f);
  for (int s = 0; s < n_slots; ++s) {
    r->slot_vector.push_back(s + 1);
    r->slot_mf_dim_vector.push_back(mf_dim);
    r->slot_dim.push_back(mf_dim + 3);
  }
  return r;
}

// Builds the sharded table for the synthetic key set {synth_key(i, seed) : i < n_total}.
int ref_build_synthetic(void* h, uint64_t n_total, uint64_t key_seed, int create_mf, int64_t* n_per_dev_out) {
  RefPs* r = (RefPs*)h;
  auto* acc_wrap = GlobalAccessorFactory::GetInstance().GetAccessorWrapper();
  CommonFeatureValueAccessor* acc = ((AccessorWrapper<CommonFeatureValueAccessor>*)acc_wrap)->AccessorPtr();
  int mf = r->mf_dim;
  const size_t fv_size = acc_wrap->GetFeatureValueSize(mf);
  std::vector<std::vector<uint64_t>> host_keys(r->ndev);
  size_t size_max = 0;
  for (int i = 0; i < r->ndev; ++i) {
    platform::CUDADeviceGuard guard(r->dev_ids[i]);
    const unsigned long long cap = n_total / r->ndev + n_total / (r->ndev * 50) + 65536;
    uint64_t* d_keys = nullptr;
    unsigned long long* d_cnt = nullptr;
    REF_CUDA(cudaMalloc(&d_keys, cap * sizeof(uint64_t)));
    REF_CUDA(cudaMalloc(&d_cnt, sizeof(unsigned long long)));
    REF_CUDA(cudaMemset(d_cnt, 0, sizeof(unsigned long long)));
    k_collect_keys<<<148 * 8, 512>>>(n_total, key_seed, i, r->ndev, d_keys, d_cnt, cap);
    unsigned long long n = 0;
    REF_CUDA(cudaMemcpy(&n, d_cnt, sizeof(n), cudaMemcpyDeviceToHost));
    if (n > cap) {
      snprintf(g_err, sizeof(g_err), "shard %d holds %llu keys, expected at most %llu", i, n, cap);
      return -1;
    }
    host_keys[i].resize(n);
    REF_CUDA(cudaMemcpy(host_keys[i].data(), d_keys, n * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    // the HBM pool of this GPU, values initialised in place (row k belongs to host_keys[i][k])
    r->dev[i].hbm_pool = new HBMMemoryPool(n, fv_size);
    r->dev[i].hbm_pool->reset(n);
    k_init_values<<<148 * 8, 256>>>(r->dev[i].hbm_pool->mem(), fv_size, n, d_keys, *acc, mf, create_mf);
    REF_CUDA(cudaDeviceSynchronize());
    REF_CUDA(cudaFree(d_keys));
    REF_CUDA(cudaFree(d_cnt));
    r->dev[i].n_keys = (int64_t)n;
    if (n_per_dev_out) n_per_dev_out[i] = (int64_t)n;
    size_max = std::max(size_max, (size_t)n);
  }
  std::unordered_map<std::string, float> cfg;
  r->heter_ps = HeterPsBase::get_instance(size_max, r->resource, cfg, "CtrDymfAccessor", 1);
  r->heter_ps->set_sparse_sgd(r->opt);
  r->heter_ps->set_embedx_sgd(r->opt);
  r->heter_ps->set_multi_mf_dim(1, mf);
  for (int i = 0; i < r->ndev; ++i) {
    platform::CUDADeviceGuard guard(r->dev_ids[i]);
    r->heter_ps->build_ps(i, host_keys[i].data(), r->dev[i].hbm_pool->mem(), host_keys[i].size(), fv_size, 500000, 2);
    std::vector<uint64_t>().swap(host_keys[i]);
  }
  return 0;
}

int ref_pull_value_floats(void* h) {
  RefPs* r = (RefPs*)h;
  int mf = r->mf_dim;
  return (int)(GlobalAccessorFactory::GetInstance().GetAccessorWrapper()->GetPullValueSize(mf) / sizeof(float));
}

// PSGPUWrapper::PullSparse (ps_gpu_wrapper.cc:1186-1250).  keys[s] / values[s]: device pointers of
// slot s ([slot_lengths[s]] uint64 keys -> [slot_lengths[s], hidden] floats); called by the worker
// thread of GPU `dev` with that device current.
int ref_pull_sparse(void* h, int dev, const uint64_t* const* keys, float* const* values, const int64_t* slot_lengths,
                    int n_slots, int hidden_size) {
  RefPs* r = (RefPs*)h;
  platform::CUDAPlace place(r->dev_ids[dev]);
  auto* acc_wrap = GlobalAccessorFactory::GetInstance().GetAccessorWrapper();
  int mf = r->mf_dim;
  const size_t feature_value_size = acc_wrap->GetPullValueSize(mf);
  auto stream = dynamic_cast<phi::GPUContext*>(platform::DeviceContextPool::Instance().Get(place))->stream();
  const int devid_2_index = r->heter_ps->get_index_by_devid(r->dev_ids[dev]);
  const size_t total_length = std::accumulate(slot_lengths, slot_lengths + n_slots, (size_t)0);
  auto buf = memory::Alloc(place, total_length * feature_value_size);
  float* total_values_gpu = reinterpret_cast<float*>(buf->ptr());
  DevState& ds = r->dev[dev];
  ds.total_keys = memory::Alloc(place, total_length * sizeof(uint64_t));
  ds.total_length = (int64_t)total_length;
  uint64_t* total_keys = reinterpret_cast<uint64_t*>(ds.total_keys->ptr());
  std::vector<int64_t> slot_lengths_lod(slot_lengths, slot_lengths + n_slots);
  for (size_t i = 1; i < slot_lengths_lod.size(); i++) slot_lengths_lod[i] += slot_lengths_lod[i - 1];
  auto buf_key = memory::Alloc(place, n_slots * sizeof(uint64_t*));
  auto buf_length = memory::Alloc(place, n_slots * sizeof(int64_t));
  uint64_t** gpu_keys = reinterpret_cast<uint64_t**>(buf_key->ptr());
  int64_t* gpu_len = reinterpret_cast<int64_t*>(buf_length->ptr());
  REF_CUDA(cudaMemcpy(gpu_keys, keys, n_slots * sizeof(uint64_t*), cudaMemcpyHostToDevice));
  REF_CUDA(cudaMemcpy(gpu_len, slot_lengths_lod.data(), n_slots * sizeof(int64_t), cudaMemcpyHostToDevice));
  auto buf_dim = memory::Alloc(place, n_slots * sizeof(int));
  int* gpu_dim = reinterpret_cast<int*>(buf_dim->ptr());
  REF_CUDA(cudaMemcpy(gpu_dim, r->slot_dim.data(), n_slots * sizeof(int), cudaMemcpyHostToDevice));
  // CopyKeys (ps_gpu_wrapper.cu:85-97): kernel + stream sync
  k_copy_keys<<<((int)total_length + 1024 - 1) / 1024, 1024, 0, stream>>>(gpu_keys, total_keys, gpu_len, n_slots,
                                                                         (int)total_length);
  REF_CUDA(cudaStreamSynchronize(stream));
  r->heter_ps->pull_sparse(devid_2_index, total_keys, total_values_gpu, total_length);
  std::vector<float*> vals(values, values + n_slots);
  acc_wrap->CopyForPull(place, gpu_keys, vals, total_values_gpu, gpu_len, n_slots, hidden_size, (int64_t)total_length,
                        gpu_dim, (int)feature_value_size);
  return 0;
}

// PSGPUWrapper::PushSparseGrad (ps_gpu_wrapper.cc:1484-1512): grad_values[s] = [slot_lengths[s], hidden]
// with show / click in columns 0 / 1 (what fused_seqpool_cvm's backward produces)
int ref_push_sparse(void* h, int dev, const float* const* grad_values, const int64_t* slot_lengths, int n_slots,
                    int hidden_size, int batch_size) {
  RefPs* r = (RefPs*)h;
  platform::CUDAPlace place(r->dev_ids[dev]);
  auto* acc_wrap = GlobalAccessorFactory::GetInstance().GetAccessorWrapper();
  int mf = r->mf_dim;
  const size_t grad_value_size = acc_wrap->GetPushValueSize(mf);
  const int devid_2_index = r->heter_ps->get_index_by_devid(r->dev_ids[dev]);
  DevState& ds = r->dev[dev];
  const int64_t total_length = std::accumulate(slot_lengths, slot_lengths + n_slots, (int64_t)0);
  if (total_length != ds.total_length) {
    snprintf(g_err, sizeof(g_err), "push of %lld keys after a pull of %lld", (long long)total_length,
             (long long)ds.total_length);
    return -1;
  }
  auto buf = memory::Alloc(place, total_length * grad_value_size);
  float* total_grad_values_gpu = reinterpret_cast<float*>(buf->ptr());
  uint64_t* total_keys = reinterpret_cast<uint64_t*>(ds.total_keys->ptr());
  std::vector<const float*> gv(grad_values, grad_values + n_slots);
  std::vector<int64_t> sl(slot_lengths, slot_lengths + n_slots);
  acc_wrap->CopyForPush(place, gv, total_grad_values_gpu, sl, (uint64_t)total_length, batch_size, grad_value_size,
                        r->slot_vector, r->slot_mf_dim_vector);
  r->heter_ps->push_sparse(devid_2_index, total_keys, total_grad_values_gpu, (size_t)total_length);
  return 0;
}

// Dense data parallelism of the reference: one NCCL communicator per GPU from ncclCommInitAll
// (platform/collective_helper.cc:120-131), every worker thread calls ncclAllReduce on its own
// communicator and compute stream (c_allreduce_sum; boxps_worker.cc:1235-1236).
int ref_nccl_init(void* h) {
  RefPs* r = (RefPs*)h;
  r->comms.resize(r->ndev);
  ncclResult_t e = ncclCommInitAll(r->comms.data(), r->ndev, r->dev_ids.data());
  if (e != ncclSuccess) {
    snprintf(g_err, sizeof(g_err), "ncclCommInitAll: %s", ncclGetErrorString(e));
    return -1;
  }
  return 0;
}

int ref_allreduce_sum(void* h, int dev, float* buf, int64_t count, void* stream) {
  RefPs* r = (RefPs*)h;
  ncclResult_t e = ncclAllReduce(buf, buf, (size_t)count, ncclFloat, ncclSum, r->comms[dev], (cudaStream_t)stream);
  if (e != ncclSuccess) {
    snprintf(g_err, sizeof(g_err), "ncclAllReduce: %s", ncclGetErrorString(e));
    return -1;
  }
  return 0;
}

int ref_destroy(void* h) {
  RefPs* r = (RefPs*)h;
  if (!r) return 0;
  for (auto c : r->comms) ncclCommDestroy(c);
  for (auto& d : r->dev) {
    d.total_keys.reset();
    delete d.hbm_pool;
  }
  delete r->heter_ps;
  delete r;
  return 0;
}

}  // extern "C"
