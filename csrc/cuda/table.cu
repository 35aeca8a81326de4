// HBM sparse-table shard: create / bulk insert (BeginPass) / lookup / update /
// dump (EndPass, save) / shrink (decay + evict) / synthetic fill for benches.
// Replaces reference kernels insert_kernel, dy_mf_search_kernel, dy_mf_update_kernel,
// dump_to_cpu (heter_ps/hashtable_kernel.cu:34-140, :247-301).
#include "table.cuh"
#include <stdarg.h>
#include <stdlib.h>
#include <mutex>
#include <string>

static thread_local std::string g_err;
extern "C" const char* pbx_last_error() { return g_err.c_str(); }
void pbx_set_error(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
}

#include <atomic>
static std::atomic<long long> g_launches{0};
extern "C" long long pbx_launch_count() { return g_launches.load(); }
void pbx_note_launch(int n) { g_launches += n; }

namespace {

__global__ void k_fill_keys(uint64_t* keys, uint64_t n) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    keys[i] = PBX_EMPTY_KEY;
}

// one thread per key; optional full rows to load (checkpoint / host tier -> HBM)
__global__ void k_insert(PbxTableView t, const uint64_t* __restrict__ keys, int64_t n,
                         const float* __restrict__ rows, int* status) {
  const int S = t.L.stride;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint64_t key = keys[i];
    if (key == PBX_EMPTY_KEY) continue;
    bool fresh;
    uint32_t s = pbx_dev_insert(t, key, &fresh);
    if (s == PBX_NO_ROW) { atomicExch(status, 2); continue; }
    float* row = t.vals + (uint64_t)s * S;
    if (rows) {
      if (fresh) {
        const float* src = rows + i * S;
        for (int j = 0; j < S - 1; ++j) row[j] = src[j];
        row[S - 1] = 0.f;   // the lock / requester-count word is per-pass state, never loaded
      }
    } else if (fresh) {
      pbx_row_init(t, row, key);
    }
  }
}

// synthetic fill: key(i) = mix(i + base) | 1 ... owned keys only.  Rows get a
// created embedx so that benches exercise the full-width pull/push path.
__global__ void k_fill_synth(PbxTableView t, uint64_t n_total, uint64_t key_seed, int rank, int nranks,
                             int create_mf, int* status) {
  const int S = t.L.stride;
  const int D = t.L.embedx_dim;
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n_total;
       i += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t key = pbx_mix64(i * 0x9E3779B97F4A7C15ull + key_seed) >> 1;  // never EMPTY
    if (pbx_owner_of(key, nranks) != rank) continue;
    bool fresh;
    uint32_t s = pbx_dev_insert(t, key, &fresh);
    if (s == PBX_NO_ROW) { atomicExch(status, 2); continue; }
    if (!fresh) continue;
    float* row = t.vals + (uint64_t)s * S;
    pbx_row_init(t, row, key);
    row[0] = 20.f + (float)(key & 15);
    row[1] = (float)(key & 3);
    row[2] = (pbx_uniform01(key, 0xfffeu, t.seed) - 0.5f) * 0.02f;
    if (create_mf) {
      row[t.L.off_mf] = (float)D;
      for (int j = 0; j < D; ++j) row[3 + j] = (pbx_uniform01(key, j, t.seed) - 0.5f) * 0.02f;
      if (t.L.optimizer == PBX_OPT_ADAM) {
        row[t.L.off_xs + 2 * D] = t.cfg.mf_beta1_decay_rate;
        row[t.L.off_xs + 2 * D + 1] = t.cfg.mf_beta2_decay_rate;
      } else if (t.L.optimizer == PBX_OPT_SHARED_ADAM) {
        row[t.L.off_xs + 2] = t.cfg.mf_beta1_decay_rate;
        row[t.L.off_xs + 3] = t.cfg.mf_beta2_decay_rate;
      }
    }
  }
}

__global__ void k_count(PbxTableView t, unsigned long long* out) {
  unsigned long long local = 0;
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < t.capacity;
       i += (uint64_t)gridDim.x * blockDim.x)
    local += (t.keys[i] != PBX_EMPTY_KEY);
  for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
  if ((threadIdx.x & 31) == 0 && local) atomicAdd(out, local);
}

// compaction dump: mode 0 = all, 1 = dirty only (delta); clears dirty flags when asked
__global__ void k_dump(PbxTableView t, uint64_t* out_keys, float* out_rows, unsigned long long max_n,
                       unsigned long long* counter, int mode, int clear_dirty) {
  const int S = t.L.stride;
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < t.capacity;
       i += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t k = t.keys[i];
    bool take = k != PBX_EMPTY_KEY && (mode == 0 || t.dirty[i]);
    unsigned mask = __ballot_sync(__activemask(), take);
    if (!take) continue;
    // warp-aggregated slot claim
    int lane = threadIdx.x & 31;
    int leader = __ffs(mask) - 1;
    unsigned long long base = 0;
    if (lane == leader) base = atomicAdd(counter, (unsigned long long)__popc(mask));
    base = __shfl_sync(mask, base, leader);
    unsigned long long pos = base + __popc(mask & ((1u << lane) - 1));
    if (pos < max_n) {
      if (out_keys) out_keys[pos] = k;
      if (out_rows) {
        const float* row = t.vals + i * S;
        float* dst = out_rows + pos * S;
        for (int j = 0; j < S - 1; ++j) dst[j] = row[j];
        dst[S - 1] = 0.f;   // lock / requester-count word: not part of the value
      }
    }
    if (clear_dirty) t.dirty[i] = 0;
  }
}

__global__ void k_lookup(PbxTableView t, const uint64_t* __restrict__ keys, int64_t n, float* __restrict__ out,
                         int out_stride, uint32_t* __restrict__ rows_out) {
  const int H = 3 + t.L.embedx_dim;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint64_t key = keys[i];
    uint32_t s = key == 0 ? PBX_NO_ROW : pbx_dev_find(t, key);
    if (rows_out) rows_out[i] = s;
    float* o = out + i * out_stride;
    if (s == PBX_NO_ROW) {
      for (int j = 0; j < out_stride; ++j) o[j] = 0.f;
    } else {
      const float* row = t.vals + (uint64_t)s * t.L.stride;
      for (int j = 0; j < H; ++j) o[j] = row[j];
      for (int j = H; j < out_stride; ++j) o[j] = 0.f;
    }
  }
}

// full value rows by key, one warp per key (coalesced row copy); absent keys give a zero row
__global__ void k_fetch_rows(PbxTableView t, const uint64_t* __restrict__ keys, int64_t n, float* __restrict__ out,
                             uint8_t* __restrict__ found) {
  const int S = t.L.stride;
  const int lane = threadIdx.x & 31;
  const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t i = warp; i < n; i += nwarps) {
    uint32_t s = PBX_NO_ROW;
    if (lane == 0) {
      const uint64_t key = keys[i];
      s = (key == 0 || key == PBX_EMPTY_KEY) ? PBX_NO_ROW : pbx_dev_find(t, key);
      if (found) found[i] = s != PBX_NO_ROW;
    }
    s = __shfl_sync(0xffffffffu, s, 0);
    const float* row = t.vals + (uint64_t)s * S;
    float* o = out + i * S;
    for (int j = lane; j < S; j += 32) o[j] = (s == PBX_NO_ROW || j == S - 1) ? 0.f : row[j];
  }
}

// unique keys only (caller merges duplicates first)
__global__ void k_update(PbxTableView t, const uint64_t* __restrict__ keys, const float* __restrict__ push,
                         int push_stride, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint64_t key = keys[i];
    if (key == 0) continue;
    uint32_t s = pbx_dev_find(t, key);
    if (s == PBX_NO_ROW) continue;
    pbx_apply_push(t, s, key, push + i * push_stride);
  }
}

// decay show/click, age unseen_days, mark rows to evict (CtrDymfAccessor::Shrink,
// reference paddle/fluid/distributed/ps/table/ctr_dymf_accessor.cc:68-85)
__global__ void k_decay_mark(PbxTableView t, float decay, float delete_threshold, float delete_after_days,
                             uint8_t* evict) {
  const int S = t.L.stride;
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < t.capacity;
       i += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t k = t.keys[i];
    if (k == PBX_EMPTY_KEY) { evict[i] = 0; continue; }
    float* row = t.vals + i * S;
    row[0] *= decay;
    row[1] *= decay;
    float unseen = row[t.L.off_tail + 2] + 1.f;
    row[t.L.off_tail + 2] = unseen;
    float score = t.cfg.nonclk_coeff * (row[0] - row[1]) + t.cfg.clk_coeff * row[1];
    evict[i] = (k != 0 && (score < delete_threshold || unseen > delete_after_days)) ? 1 : 0;
  }
}

__global__ void k_collect_survivors(PbxTableView t, const uint8_t* evict, uint64_t* out_keys, float* out_rows,
                                    uint8_t* out_dirty, unsigned long long* counter) {
  const int S = t.L.stride;
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < t.capacity;
       i += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t k = t.keys[i];
    if (k == PBX_EMPTY_KEY || evict[i]) continue;
    unsigned long long pos = atomicAdd(counter, 1ull);
    out_keys[pos] = k;
    out_dirty[pos] = t.dirty[i];
    const float* row = t.vals + i * S;
    float* dst = out_rows + pos * S;
    for (int j = 0; j < S; ++j) dst[j] = row[j];
  }
}

__global__ void k_reinsert(PbxTableView t, const uint64_t* keys, const float* rows, const uint8_t* dirty,
                           unsigned long long n) {
  const int S = t.L.stride;
  for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    bool fresh;
    uint32_t s = pbx_dev_insert(t, keys[i], &fresh);
    if (s == PBX_NO_ROW) continue;
    float* row = t.vals + (uint64_t)s * S;
    const float* src = rows + i * S;
    for (int j = 0; j < S; ++j) row[j] = src[j];
    t.dirty[s] = dirty[i];
  }
}

}  // namespace

extern "C" {

int pbx_table_create(void** out, uint64_t capacity, int embedx_dim, int optimizer, const PbxSparseOptConfig* cfg,
                     uint64_t seed) {
  PBX_REQUIRE(capacity >= 16 && capacity < 0xFFFFFFF0ull, "table capacity %llu out of range",
              (unsigned long long)capacity);
  PBX_REQUIRE(embedx_dim >= 0 && embedx_dim <= 1024, "bad embedx_dim %d", embedx_dim);
  PbxTable* t = new PbxTable();
  memset(t, 0, sizeof(*t));
  cudaGetDevice(&t->device);
  t->view.capacity = capacity;
  t->view.seed = seed;
  t->view.L = pbx_make_layout(embedx_dim, optimizer);
  t->view.cfg = *cfg;
  PBX_CUDA_CHECK(cudaMalloc(&t->view.keys, capacity * sizeof(uint64_t)));
  PBX_CUDA_CHECK(cudaMalloc(&t->view.vals, capacity * (uint64_t)t->view.L.stride * sizeof(float)));
  PBX_CUDA_CHECK(cudaMalloc(&t->view.dirty, capacity));
  PBX_CUDA_CHECK(cudaMalloc(&t->d_counter, 2 * sizeof(unsigned long long)));
  PBX_CUDA_CHECK(cudaMemset(t->view.dirty, 0, capacity));
  PBX_CUDA_CHECK(cudaMemset(t->d_counter, 0, 2 * sizeof(unsigned long long)));
  k_fill_keys<<<pbx_grid_for((long long)capacity, 256), 256>>>(t->view.keys, capacity);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  PBX_CUDA_CHECK(cudaDeviceSynchronize());
  *out = t;
  return 0;
}

int pbx_table_destroy(void* h) {
  PbxTable* t = (PbxTable*)h;
  if (!t) return 0;
  cudaFree(t->view.keys);
  cudaFree(t->view.vals);
  cudaFree(t->view.dirty);
  cudaFree(t->d_counter);
  delete t;
  return 0;
}

int pbx_table_layout(void* h, PbxRowLayout* out) {
  *out = ((PbxTable*)h)->view.L;
  return 0;
}

int pbx_table_set_config(void* h, const PbxSparseOptConfig* cfg) {
  ((PbxTable*)h)->view.cfg = *cfg;
  return 0;
}

int pbx_table_set_test_mode(void* h, int on) {
  ((PbxTable*)h)->view.test_mode = on;
  return 0;
}

int pbx_table_clear(void* h, cudaStream_t st) {
  PbxTable* t = (PbxTable*)h;
  k_fill_keys<<<pbx_grid_for((long long)t->view.capacity, 256), 256, 0, st>>>(t->view.keys, t->view.capacity);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  PBX_CUDA_CHECK(cudaMemsetAsync(t->view.dirty, 0, t->view.capacity, st));
  return 0;
}

// d_rows may be NULL (fresh rows get the optimizer's initial state).
int pbx_table_insert(void* h, const uint64_t* d_keys, int64_t n, const float* d_rows, cudaStream_t st) {
  PbxTable* t = (PbxTable*)h;
  if (n <= 0) return 0;
  int* status = (int*)(t->d_counter + 1);
  PBX_CUDA_CHECK(cudaMemsetAsync(status, 0, sizeof(int), st));
  k_insert<<<pbx_grid_for(n, 256), 256, 0, st>>>(t->view, d_keys, n, d_rows, status);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  int hs = 0;
  PBX_CUDA_CHECK(cudaMemcpyAsync(&hs, status, sizeof(int), cudaMemcpyDeviceToHost, st));
  PBX_CUDA_CHECK(cudaStreamSynchronize(st));
  PBX_REQUIRE(hs == 0, "sparse table shard is full (capacity %llu)", (unsigned long long)t->view.capacity);
  return 0;
}

int pbx_table_fill_synthetic(void* h, uint64_t n_total, uint64_t key_seed, int rank, int nranks, int create_mf,
                             cudaStream_t st) {
  PbxTable* t = (PbxTable*)h;
  int* status = (int*)(t->d_counter + 1);
  PBX_CUDA_CHECK(cudaMemsetAsync(status, 0, sizeof(int), st));
  k_fill_synth<<<pbx_grid_for((long long)n_total, 256, 16), 256, 0, st>>>(t->view, n_total, key_seed, rank, nranks,
                                                                            create_mf, status);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  int hs = 0;
  PBX_CUDA_CHECK(cudaMemcpyAsync(&hs, status, sizeof(int), cudaMemcpyDeviceToHost, st));
  PBX_CUDA_CHECK(cudaStreamSynchronize(st));
  PBX_REQUIRE(hs == 0, "sparse table shard is full during synthetic fill");
  return 0;
}

int pbx_table_count(void* h, int64_t* out, cudaStream_t st) {
  PbxTable* t = (PbxTable*)h;
  PBX_CUDA_CHECK(cudaMemsetAsync(t->d_counter, 0, sizeof(unsigned long long), st));
  k_count<<<pbx_grid_for((long long)t->view.capacity, 256), 256, 0, st>>>(t->view, t->d_counter);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  unsigned long long v = 0;
  PBX_CUDA_CHECK(cudaMemcpyAsync(&v, t->d_counter, sizeof(v), cudaMemcpyDeviceToHost, st));
  PBX_CUDA_CHECK(cudaStreamSynchronize(st));
  *out = (int64_t)v;
  return 0;
}

// mode 0: every live row; mode 1: rows touched since the last delta dump.
// Either output pointer may be NULL (count only).  *n_out gets the number found.
int pbx_table_dump(void* h, uint64_t* d_keys_out, float* d_rows_out, int64_t max_n, int mode, int clear_dirty,
                   int64_t* n_out, cudaStream_t st) {
  PbxTable* t = (PbxTable*)h;
  PBX_CUDA_CHECK(cudaMemsetAsync(t->d_counter, 0, sizeof(unsigned long long), st));
  k_dump<<<pbx_grid_for((long long)t->view.capacity, 256), 256, 0, st>>>(
      t->view, d_keys_out, d_rows_out, (unsigned long long)max_n, t->d_counter, mode, clear_dirty);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  unsigned long long v = 0;
  PBX_CUDA_CHECK(cudaMemcpyAsync(&v, t->d_counter, sizeof(v), cudaMemcpyDeviceToHost, st));
  PBX_CUDA_CHECK(cudaStreamSynchronize(st));
  *n_out = (int64_t)v;
  return 0;
}

int pbx_table_lookup(void* h, const uint64_t* d_keys, int64_t n, float* d_out, int out_stride, uint32_t* d_rows_out,
                     cudaStream_t st) {
  PbxTable* t = (PbxTable*)h;
  if (n <= 0) return 0;
  PBX_REQUIRE(out_stride >= 3 + t->view.L.embedx_dim, "lookup out_stride %d < hidden %d", out_stride,
              3 + t->view.L.embedx_dim);
  k_lookup<<<pbx_grid_for(n, 256), 256, 0, st>>>(t->view, d_keys, n, d_out, out_stride, d_rows_out);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

// Full value rows [n, stride] of the given keys (zero rows, found = 0, for absent keys): the
// by-key counterpart of pbx_table_dump — EndPass write-back of a pass's key list in bounded chunks
// (reference dump_to_cpu, hashtable_kernel.cu:247-301), and the oracle side of bench.py's self-check.
int pbx_table_fetch_rows(void* h, const uint64_t* d_keys, int64_t n, float* d_rows_out, uint8_t* d_found,
                         cudaStream_t st) {
  PbxTable* t = (PbxTable*)h;
  if (n <= 0) return 0;
  k_fetch_rows<<<pbx_grid_for(n * 32, 256), 256, 0, st>>>(t->view, d_keys, n, d_rows_out, d_found);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_table_update(void* h, const uint64_t* d_keys, const float* d_push, int push_stride, int64_t n,
                     cudaStream_t st) {
  PbxTable* t = (PbxTable*)h;
  if (n <= 0 || t->view.test_mode) return 0;
  PBX_REQUIRE(push_stride >= 4 + t->view.L.embedx_dim, "push_stride %d too small", push_stride);
  k_update<<<pbx_grid_for(n, 128), 128, 0, st>>>(t->view, d_keys, d_push, push_stride, n);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_table_shrink(void* h, float decay, float delete_threshold, float delete_after_unseen_days, int64_t* kept,
                     cudaStream_t st) {
  PbxTable* t = (PbxTable*)h;
  const uint64_t cap = t->view.capacity;
  uint8_t* evict = nullptr;
  PBX_CUDA_CHECK(cudaMalloc(&evict, cap));
  int grid = pbx_grid_for((long long)cap, 256);
  k_decay_mark<<<grid, 256, 0, st>>>(t->view, decay, delete_threshold, delete_after_unseen_days, evict);
  int64_t live = 0;
  int rc = pbx_table_count(h, &live, st);
  if (rc) { cudaFree(evict); return rc; }
  uint64_t* tk = nullptr;
  float* tr = nullptr;
  uint8_t* td = nullptr;
  size_t nl = (size_t)(live > 0 ? live : 1);
  PBX_CUDA_CHECK(cudaMalloc(&tk, nl * sizeof(uint64_t)));
  PBX_CUDA_CHECK(cudaMalloc(&tr, nl * (size_t)t->view.L.stride * sizeof(float)));
  PBX_CUDA_CHECK(cudaMalloc(&td, nl));
  PBX_CUDA_CHECK(cudaMemsetAsync(t->d_counter, 0, sizeof(unsigned long long), st));
  k_collect_survivors<<<grid, 256, 0, st>>>(t->view, evict, tk, tr, td, t->d_counter);
  unsigned long long n = 0;
  PBX_CUDA_CHECK(cudaMemcpyAsync(&n, t->d_counter, sizeof(n), cudaMemcpyDeviceToHost, st));
  PBX_CUDA_CHECK(cudaStreamSynchronize(st));
  pbx_table_clear(h, st);
  if (n) k_reinsert<<<pbx_grid_for((long long)n, 256), 256, 0, st>>>(t->view, tk, tr, td, n);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  PBX_CUDA_CHECK(cudaStreamSynchronize(st));
  cudaFree(evict);
  cudaFree(tk);
  cudaFree(tr);
  cudaFree(td);
  if (kept) *kept = (int64_t)n;
  return 0;
}

}  // extern "C"
