// Symmetric (peer-mapped) device memory + the dense-gradient collective kernels that run
// over it (W8 of SURVEY.md §8).  Every rank allocates the same-sized region, exchanges
// CUDA IPC handles once, and from then on kernels read and write peer HBM directly.
//
//   allreduce : ONE cooperative kernel = barrier -> two-shot reduce (each rank sums its
//               slice from all peers with 16-byte P2P loads, multiplies by `scale`, and
//               stores the result into every peer's buffer) -> barrier.  The 1/nranks
//               gradient scale — a separate `scale` op after c_allreduce_sum in the
//               reference (transpiler/collective.py:240-315, boxps_worker.cc:1235-1239)
//               — is fused into the reduction.
//   reduce_scatter / all_gather : the two halves, for ZeRO-1 style sharding
//               (boxps_worker.cc:1220-1233).
//   broadcast : root's slice to everyone (c_broadcast in sharding mode).
// NCCL remains available in parallel/ as the labelled baseline and for multi-node.
#include "common.cuh"
#include <cooperative_groups.h>
#include <vector>
namespace cg = cooperative_groups;

#include "symm.cuh"

namespace {

template <bool EMU>
__device__ __forceinline__ void symm_barrier(cg::grid_group& grid, const PbxSymmDev& c, uint64_t target,
                                             bool leader_block) {
  grid.sync();   // gpu-scope ordering of all threads' stores; the leader's fence below is system scope
  if (!EMU && c.nranks > 1) {
    if (leader_block && threadIdx.x < c.nranks) {
      int p = threadIdx.x;
      __threadfence_system();
      pbx_st_release_sys(c.flags[p] + c.rank, target);
      uint64_t t0 = pbx_globaltimer();
      while (pbx_ld_acquire_sys(c.flags[c.rank] + p) < target) {
        if (pbx_globaltimer() - t0 > 20000000000ull) { atomicExch(c.status, 2); break; }
      }
    }
    grid.sync();
  }
}

__device__ __forceinline__ float4 ld_f4_cg(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }

// op: 0 allreduce, 1 reduce_scatter (result only in own slice), 2 all_gather (own slice -> all),
//     3 broadcast from `root`, 4 barrier only
template <bool EMU>
__global__ void __launch_bounds__(512, 1)
k_symm_collective(const PbxSymmDev* __restrict__ ctxs, int nvr, int64_t off, int64_t n, float scale, int op,
                  int root) {
  cg::grid_group grid = cg::this_grid();
  const int bpr = gridDim.x / nvr;
  const int vr = blockIdx.x / bpr;
  const int lb = blockIdx.x % bpr;
  const PbxSymmDev& c = ctxs[vr];
  const int R = c.nranks;
  const uint64_t e0 = *c.epoch;
  const int64_t gtid = (int64_t)lb * blockDim.x + threadIdx.x;
  const int64_t gstride = (int64_t)bpr * blockDim.x;

  symm_barrier<EMU>(grid, c, e0 + 1, lb == 0);  // everybody's input is in place

  // slice r = [r*chunk, min(n, (r+1)*chunk)), chunk a multiple of 4 floats
  const int64_t chunk = ((n + R - 1) / R + 3) / 4 * 4;
  if (op == 0 || op == 1) {
    const int64_t b = min(n, (int64_t)c.rank * chunk), e = min(n, b + chunk);
    const int64_t nv = (e - b) / 4;
    for (int64_t i = gtid; i < nv; i += gstride) {
      const int64_t idx = off + b + i * 4;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int p = 0; p < PBX_MAX_RANKS; ++p) {
        if (p < R) {
          float4 v = ld_f4_cg(c.data[(c.rank + p) % R] + idx);
          acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
      }
      acc.x *= scale; acc.y *= scale; acc.z *= scale; acc.w *= scale;
      if (op == 0) {
#pragma unroll
        for (int p = 0; p < PBX_MAX_RANKS; ++p)
          if (p < R) *reinterpret_cast<float4*>(c.data[(c.rank + p) % R] + idx) = acc;
      } else {
        *reinterpret_cast<float4*>(c.data[c.rank] + idx) = acc;
      }
    }
    // tail (n not a multiple of 4): last slice owner handles scalars
    for (int64_t i = b + nv * 4 + gtid; i < e; i += gstride) {
      float acc = 0.f;
      for (int p = 0; p < R; ++p) acc += __ldcg(c.data[p] + off + i);
      acc *= scale;
      if (op == 0) for (int p = 0; p < R; ++p) c.data[p][off + i] = acc;
      else c.data[c.rank][off + i] = acc;
    }
  } else if (op == 2) {
    const int64_t b = min(n, (int64_t)c.rank * chunk), e = min(n, b + chunk);
    for (int64_t i = b + gtid; i < e; i += gstride) {
      float v = c.data[c.rank][off + i];
      for (int p = 0; p < R; ++p)
        if (p != c.rank) c.data[p][off + i] = v;
    }
  } else if (op == 3) {
    if (c.rank == root)
      for (int64_t i = gtid; i < n; i += gstride) {
        float v = c.data[root][off + i];
        for (int p = 0; p < R; ++p)
          if (p != root) c.data[p][off + i] = v;
      }
  }
  symm_barrier<EMU>(grid, c, e0 + 2, lb == 0);  // every peer's stores have landed
  if (lb == 0 && threadIdx.x == 0) *c.epoch = e0 + 2;
}

// data_norm(sync_stats): the three batch-statistics vectors of every rank sit in the symmetric
// region at [off, off + 3*C); every rank sums them over all peers with P2P loads and applies the
// decayed summary update to its own (plain, local) summary vectors — barrier, reduce, update,
// barrier in ONE launch.  The reference does 3 ncclAllReduce calls and a separate KernelUpdateParam
// (data_norm_op.cu:203-231, :90-110).
template <bool EMU>
__global__ void __launch_bounds__(256, 1)
k_symm_stat_update(const PbxSymmDev* __restrict__ ctxs, int nvr, int64_t off, int C, float scale, float decay,
                   float* const* __restrict__ summaries /* [nvr][3] */) {
  cg::grid_group grid = cg::this_grid();
  const int bpr = gridDim.x / nvr;
  const int vr = blockIdx.x / bpr;
  const int lb = blockIdx.x % bpr;
  const PbxSymmDev& c = ctxs[vr];
  const int R = c.nranks;
  const uint64_t e0 = *c.epoch;
  symm_barrier<EMU>(grid, c, e0 + 1, lb == 0);
  for (int64_t i = (int64_t)lb * blockDim.x + threadIdx.x; i < 3 * (int64_t)C; i += (int64_t)bpr * blockDim.x) {
    float acc = 0.f;
    for (int p = 0; p < R; ++p) acc += __ldcg(c.data[(c.rank + p) % R] + off + i);
    float* dst = summaries[vr * 3 + (int)(i / C)] + (i % C);
    *dst = *dst * decay + acc * scale;
  }
  symm_barrier<EMU>(grid, c, e0 + 2, lb == 0);   // nobody refills its statistics while a peer still reads them
  if (lb == 0 && threadIdx.x == 0) *c.epoch = e0 + 2;
}

void symm_bind(PbxSymm* s, int p, void* base) {
  s->peer_base[p] = base;
  s->h.flags[p] = (uint64_t*)base;
  s->h.data[p] = (float*)((char*)base + s->data_off);
  s->dirty = true;
}

}  // namespace

extern "C" {

int pbx_symm_create(void** out, int rank, int nranks, uint64_t bytes) {
  PBX_REQUIRE(nranks >= 1 && nranks <= PBX_MAX_RANKS && rank >= 0 && rank < nranks, "bad rank/nranks");
  PbxSymm* s = new PbxSymm();
  memset(s, 0, sizeof(*s));
  s->h.rank = rank;
  s->h.nranks = nranks;
  s->data_off = 1024;
  s->bytes = s->data_off + (bytes + 255) / 256 * 256;
  PBX_CUDA_CHECK(cudaMalloc(&s->base, s->bytes));
  PBX_CUDA_CHECK(cudaMemset(s->base, 0, s->bytes));
  PBX_CUDA_CHECK(cudaMalloc(&s->h.epoch, 256));
  PBX_CUDA_CHECK(cudaMemset(s->h.epoch, 0, 256));
  s->h.status = (int*)(s->h.epoch + 8);
  PBX_CUDA_CHECK(cudaMalloc(&s->d, sizeof(PbxSymmDev)));
  symm_bind(s, rank, s->base);
  PBX_CUDA_CHECK(cudaDeviceSynchronize());
  *out = s;
  return 0;
}

int pbx_symm_destroy(void* h) {
  PbxSymm* s = (PbxSymm*)h;
  if (!s) return 0;
  for (int p = 0; p < s->h.nranks; ++p)
    if (s->opened[p]) cudaIpcCloseMemHandle(s->peer_base[p]);
  cudaFree(s->base);
  cudaFree(s->h.epoch);
  cudaFree(s->d);
  cudaFree(s->stat_ptrs);
  delete s;
  return 0;
}

void* pbx_symm_ptr(void* h) { return ((PbxSymm*)h)->h.data[((PbxSymm*)h)->h.rank]; }

int pbx_symm_ipc_handle(void* h, void* out64) {
  cudaIpcMemHandle_t mh;
  PBX_CUDA_CHECK(cudaIpcGetMemHandle(&mh, ((PbxSymm*)h)->base));
  memcpy(out64, &mh, sizeof(mh));
  return 0;
}

int pbx_symm_connect(void* h, const void* handles) {
  PbxSymm* s = (PbxSymm*)h;
  for (int p = 0; p < s->h.nranks; ++p) {
    if (p == s->h.rank) continue;
    cudaIpcMemHandle_t mh;
    memcpy(&mh, (const char*)handles + (size_t)p * sizeof(mh), sizeof(mh));
    void* base = nullptr;
    PBX_CUDA_CHECK(cudaIpcOpenMemHandle(&base, mh, cudaIpcMemLazyEnablePeerAccess));
    s->opened[p] = true;
    symm_bind(s, p, base);
  }
  return 0;
}

int pbx_symm_connect_local(void* h, void* const* peers) {
  PbxSymm* s = (PbxSymm*)h;
  for (int p = 0; p < s->h.nranks; ++p)
    if (p != s->h.rank) symm_bind(s, p, ((PbxSymm*)peers[p])->base);
  return 0;
}

int pbx_symm_status(void* h, int* status, cudaStream_t st) {
  PbxSymm* s = (PbxSymm*)h;
  PBX_CUDA_CHECK(cudaMemcpyAsync(status, s->h.status, sizeof(int), cudaMemcpyDeviceToHost, st));
  PBX_CUDA_CHECK(cudaStreamSynchronize(st));
  return 0;
}

// op: 0 allreduce(sum)*scale, 1 reduce_scatter, 2 all_gather, 3 broadcast(root), 4 barrier.
// `off`/`n` in fp32 elements relative to pbx_symm_ptr(); off must be a multiple of 4.
int pbx_symm_collective(void* h, int64_t off, int64_t n, float scale, int op, int root, int max_blocks,
                        cudaStream_t st) {
  PbxSymm* s = (PbxSymm*)h;
  PBX_REQUIRE((off & 3) == 0, "collective offset must be a multiple of 4 elements");
  PBX_REQUIRE((uint64_t)(off + n) * 4 <= s->bytes - s->data_off, "collective range exceeds the symmetric region");
  if (s->dirty) {
    PBX_CUDA_CHECK(cudaMemcpyAsync(s->d, &s->h, sizeof(PbxSymmDev), cudaMemcpyHostToDevice, st));
    PBX_CUDA_CHECK(cudaStreamSynchronize(st));
    s->dirty = false;
  }
  int blocks = (int)((n / 4 / s->h.nranks + 511) / 512);
  int cap = max_blocks > 0 ? max_blocks : 64;  // leave SMs to the compute stream when overlapped
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  const PbxSymmDev* d = s->d;
  int nvr = 1;
  void* args[] = {(void*)&d, (void*)&nvr, (void*)&off, (void*)&n, (void*)&scale, (void*)&op, (void*)&root};
  PBX_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)k_symm_collective<false>, dim3(blocks), dim3(512), args, 0, st));
  pbx_note_launch(1);
  return 0;
}

// summaries = summaries * decay + scale * sum over ranks of the statistics at [off, off + 3*C) of the
// symmetric region; d_size / d_sum / d_sqsum are this rank's (ordinary) summary vectors of length C.
int pbx_symm_stat_update(void* h, int64_t off, int C, float scale, float decay, float* d_size, float* d_sum,
                         float* d_sqsum, cudaStream_t st) {
  PbxSymm* s = (PbxSymm*)h;
  PBX_REQUIRE(C > 0 && (uint64_t)(off + 3 * (int64_t)C) * 4 <= s->bytes - s->data_off,
              "stat update range exceeds the symmetric region");
  if (s->dirty) {
    PBX_CUDA_CHECK(cudaMemcpyAsync(s->d, &s->h, sizeof(PbxSymmDev), cudaMemcpyHostToDevice, st));
    PBX_CUDA_CHECK(cudaStreamSynchronize(st));
    s->dirty = false;
  }
  if (!s->stat_ptrs) PBX_CUDA_CHECK(cudaMalloc(&s->stat_ptrs, 3 * sizeof(float*)));
  float* hp[3] = {d_size, d_sum, d_sqsum};
  if (memcmp(hp, s->stat_ptrs_host, sizeof(hp)) != 0) {
    memcpy(s->stat_ptrs_host, hp, sizeof(hp));
    PBX_CUDA_CHECK(cudaMemcpyAsync(s->stat_ptrs, s->stat_ptrs_host, sizeof(hp), cudaMemcpyHostToDevice, st));
  }
  const PbxSymmDev* d = s->d;
  int nvr = 1;
  float* const* sp = s->stat_ptrs;
  int blocks = (3 * C + 255) / 256;
  if (blocks > 16) blocks = 16;
  void* args[] = {(void*)&d, (void*)&nvr, (void*)&off, (void*)&C, (void*)&scale, (void*)&decay, (void*)&sp};
  PBX_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)k_symm_stat_update<false>, dim3(blocks), dim3(256), args, 0, st));
  pbx_note_launch(1);
  return 0;
}

// the same for all virtual ranks of a same-process group in one launch (tests): summaries[r*3 + j]
int pbx_symm_stat_update_emu(void* const* hs, int nranks, int64_t off, int C, float scale, float decay,
                             float* const* summaries, cudaStream_t st) {
  std::vector<PbxSymmDev> hv(nranks);
  for (int r = 0; r < nranks; ++r) hv[r] = ((PbxSymm*)hs[r])->h;
  PbxSymmDev* d = nullptr;
  float** sp = nullptr;
  PBX_CUDA_CHECK(cudaMalloc(&d, sizeof(PbxSymmDev) * nranks));
  PBX_CUDA_CHECK(cudaMalloc(&sp, sizeof(float*) * 3 * nranks));
  PBX_CUDA_CHECK(cudaMemcpy(d, hv.data(), sizeof(PbxSymmDev) * nranks, cudaMemcpyHostToDevice));
  PBX_CUDA_CHECK(cudaMemcpy(sp, summaries, sizeof(float*) * 3 * nranks, cudaMemcpyHostToDevice));
  const PbxSymmDev* dc = d;
  float* const* spc = sp;
  int bpr = 2;
  void* args[] = {(void*)&dc, (void*)&nranks, (void*)&off, (void*)&C, (void*)&scale, (void*)&decay, (void*)&spc};
  cudaError_t e = cudaLaunchCooperativeKernel((void*)k_symm_stat_update<true>, dim3(bpr * nranks), dim3(256), args, 0, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFree(d);
  cudaFree(sp);
  PBX_CUDA_CHECK(e);
  return 0;
}

// all virtual ranks in one launch on one GPU (tests)
int pbx_symm_collective_emu(void* const* hs, int nranks, int64_t off, int64_t n, float scale, int op, int root,
                            cudaStream_t st) {
  std::vector<PbxSymmDev> hv(nranks);
  for (int r = 0; r < nranks; ++r) hv[r] = ((PbxSymm*)hs[r])->h;
  PbxSymmDev* d = nullptr;
  PBX_CUDA_CHECK(cudaMalloc(&d, sizeof(PbxSymmDev) * nranks));
  PBX_CUDA_CHECK(cudaMemcpy(d, hv.data(), sizeof(PbxSymmDev) * nranks, cudaMemcpyHostToDevice));
  int bpr = 8;
  const PbxSymmDev* dc = d;
  void* args[] = {(void*)&dc, (void*)&nranks, (void*)&off, (void*)&n, (void*)&scale, (void*)&op, (void*)&root};
  cudaError_t e =
      cudaLaunchCooperativeKernel((void*)k_symm_collective<true>, dim3(bpr * nranks), dim3(512), args, 0, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFree(d);
  PBX_CUDA_CHECK(e);
  return 0;
}

}  // extern "C"
