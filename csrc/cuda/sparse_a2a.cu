// Sharded sparse pull / push as ONE persistent cooperative kernel each (W2 / W3 of
// SURVEY.md §8).  The all-to-all is done from inside the kernel with plain stores
// into peer HBM over NVLink/NVSwitch (CUDA-IPC-mapped symmetric buffers); ranks meet
// at flag barriers in peer memory.  No NCCL, no host round trip, no sort.
//
//   pull : dedup (batch-local hash) -> bucket by owner -> remote-store keys into the
//          owner's mailbox -> barrier -> owner probes its shard and remote-stores the
//          value rows straight into the requester's value buffer -> barrier.
//   push : merge gradients by unique key (fp32 atomics into a local buffer, fused
//          with the seqpool/CVM backward) -> remote-store merged records into the
//          owner's mailbox -> barrier -> owner applies the fused sparse optimizer.
//
// Replaces the reference's host-orchestrated sequence: cub sort + per-peer
// cudaMemcpyAsync + search/update kernels + >= 8 stream syncs
// (heter_ps/heter_comm_inl.h:1117-1171, :1477-1622, :1646-1825; box_wrapper_impl.h:25-234,
// :373-522; box_wrapper.cu:324-512).
//
// Testing on fewer GPUs than ranks: with EMU=true one launch runs all ranks as slices
// of one cooperative grid on one GPU ("virtual ranks"); the cross-rank barrier then is
// the grid barrier.  The per-rank code is identical.
#include "sparse_a2a.cuh"
#include "symm.cuh"
#include <cooperative_groups.h>
#include <vector>
namespace cg = cooperative_groups;

// pull: 1024 threads (61 registers, no spills) for latency hiding; push: 512 threads so that the
// software-pipelined optimizer phase can keep four rows per lane group in registers
#define PBX_A2A_THREADS 1024
#define PBX_PUSH_THREADS 512
#define PBX_PUSH_SMEM ((PBX_PUSH_THREADS / 32) * (32 * 36 + 32 * 24) * 4)   /* Q3 warp tiles, one-line rows */
#define PBX_PUSH_WIDE_WARP_WORDS 3200                                        /* Q3 warp tiles, wide rows */
#define PBX_PUSH_SMEM_WIDE ((PBX_PUSH_THREADS / 32) * PBX_PUSH_WIDE_WARP_WORDS * 4)

struct PbxPullArgs {
  const uint64_t* keys;
  int64_t n;
  uint32_t* restore;  // [n] out: unique-value row of every occurrence
  const int32_t* n_dev;  // optional: the key count lives on the device (lod[S*B]); lets one
                         // captured CUDA graph serve batches of different lengths
};
struct PbxPullArgsN { PbxPullArgs a[PBX_MAX_RANKS]; };

struct PbxPushArgs {
  const uint32_t* restore;  // [n] from the pull of the same batch
  int64_t n;
  // mode 0: gradients come from pooled output grads (fused seqpool/CVM backward)
  const float* dout;        // [B, S*Eo]
  const float* cvm;         // [B, cvm_offset]  (show, click labels)
  const int32_t* lod;       // [S*B + 1] bag offsets, slot-major
  int B, S, Eo, cvm_offset, dim_off;
  int ld_dout;              // words between rows of dout (>= S*Eo)
  // mode 1: explicit per-occurrence gradients [n, hidden] + slot of every occurrence
  const float* gkeys;
  const int32_t* key2slot;
  int hidden;
  int mode;
  float grad_scale;         // -batch_size (box_wrapper.cu:404-410)
  // optional: the data-parallel dense-gradient allreduce rides on this kernel's barriers.  After
  // the ship barrier every rank's dense gradients are final and visible; each rank reduces its
  // slice from all peers, scales it, stores it to all peers, and one more barrier ends the kernel.
  const PbxSymmDev* dense_ctx;
  int64_t dense_off, dense_n;
  float dense_scale;
  int q0_done;              // the record phase already ran as its own launch (k_push_records)
};
struct PbxPushArgsN { PbxPushArgs a[PBX_MAX_RANKS]; };

// phase timestamps (one thread, globaltimer) for benchmarks/micro.py --phases; slots live after the epoch word
#define PBX_TS(i) do { if (lb == 0 && tid == 0) c.epoch[16 + (i)] = pbx_globaltimer(); } while (0)
// per-CTA stamp taken when the CTA FINISHES the work of a phase, before it waits for the others
// (diagnosis of load imbalance; only when pbx_sparse_debug_stamps switched it on)
#define PBX_TSB(i) do { if (c.dbg && tid == 0) c.dbg[(size_t)lb * 16 + (i)] = pbx_globaltimer(); } while (0)

namespace {

template <bool EMU>
__device__ __forceinline__ void xbarrier(cg::grid_group& grid, const PbxSparseDev& c, uint64_t target,
                                         bool leader_block) {
  // Cross-rank barrier and local grid barrier in ONE level (real multi-rank runs):
  //   arrive : every CTA orders its (possibly remote) stores at gpu scope and bumps a counter;
  //            the LAST CTA of the rank issues one system-scope fence and publishes `target` into
  //            every peer's flag array (and its own) — by cumulativity all CTAs' stores are
  //            visible to whoever acquires the flag.  (A system fence in every thread costs an
  //            ERRBAR per thread and showed up as ~15% of the kernel in ncu.)
  //   wait   : R threads of EVERY CTA poll the local flag array; no second grid barrier, no
  //            leader hop.  The own-rank flag doubles as the local grid barrier.
  if (EMU || c.nranks == 1) {
    grid.sync();
    return;
  }
  (void)leader_block;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    unsigned long long* cnt = reinterpret_cast<unsigned long long*>(c.epoch + 2);
    const unsigned long long old = atomicAdd(cnt, 1ull);
    if (old + 1ull == (unsigned long long)gridDim.x) {
      *cnt = 0ull;                              // nobody re-arrives before seeing the flags below
      __threadfence_system();
      for (int p = 0; p < c.nranks; ++p) pbx_st_release_sys(c.flags[p] + c.rank, target);
    }
  }
  if (threadIdx.x < c.nranks) {
    const uint64_t t0 = pbx_globaltimer();
    while (pbx_ld_acquire_sys(c.flags[c.rank] + threadIdx.x) < target) {
      __nanosleep(32);   // 148 CTAs x R pollers: back off so the polls do not crowd the memory system
      if (pbx_globaltimer() - t0 > 20000000000ull) {  // 20 s: a peer died; fail loudly, do not hang
        atomicExch(c.status, 2);
        break;
      }
    }
  }
  __syncthreads();
}

template <bool EMU>
__global__ void __launch_bounds__(PBX_A2A_THREADS, 1)
k_sparse_pull(const PbxSparseDev* __restrict__ ctxs, int nvr, PbxPullArgsN args) {
  cg::grid_group grid = cg::this_grid();
  const int bpr = gridDim.x / nvr;         // blocks per (virtual) rank
  const int vr = blockIdx.x / bpr;
  const int lb = blockIdx.x % bpr;
  const PbxSparseDev& c = ctxs[vr];
  PbxPullArgs a = args.a[vr];
  if (a.n_dev) a.n = max((int64_t)0, min((int64_t)__ldg(a.n_dev), a.n));   // never beyond the host-validated bound
  const int R = c.nranks;
  const uint32_t cap = c.cap;
  const int tid = threadIdx.x;
  const uint64_t e0 = *c.epoch;
  PBX_TS(0);

  __shared__ uint32_t s_cnt[PBX_MAX_RANKS];
  __shared__ uint32_t s_base[PBX_MAX_RANKS];

  // ---- P0: reset per-batch counters and (re)write the whole dedup table as EMPTY.  Writing it here
  // rather than after its last use leaves its lines resident (dirty) in L2, so the hash-insert
  // atomics of P1 are L2 hits instead of one random HBM access each; full-line writes fetch nothing.
  if (lb == 0 && tid < R) {
    c.send_cnt[tid] = 0;
    c.send_cnt[32 + tid] = 0;     // P3 tile tickets
    if (tid == 0) c.send_cnt[48] = 0;   // length of the push's duplicate-occurrence list
  }
  {
    const int64_t stride = (int64_t)bpr * PBX_A2A_THREADS;
    ulonglong2* dk = reinterpret_cast<ulonglong2*>(c.dd_keys);
    const ulonglong2 e2 = make_ulonglong2(PBX_EMPTY_KEY, PBX_EMPTY_KEY);
    for (int64_t i = (int64_t)lb * PBX_A2A_THREADS + tid; i < (int64_t)(c.dd_cap >> 1); i += stride) dk[i] = e2;
  }
  PBX_TSB(0);
  grid.sync();
  PBX_TS(1);

  // ---- P1: dedup + claim a mailbox slot per unique key + ship the key to its owner.
  // Every thread handles KPT keys per round with their first hash-table CAS all in flight at once
  // (the CAS round trips to L2, not the arithmetic, are what this phase costs); a batch of up to
  // bpr * 1024 * KPT occurrences is one round, i.e. one set of block barriers.
  {
    constexpr int KPT = 3;
    const uint32_t mask = c.dd_cap - 1;
    const int64_t chunk = (int64_t)bpr * PBX_A2A_THREADS * KPT;
    const int64_t nround = (a.n + chunk - 1) / chunk;
    for (int64_t r = 0; r < nround; ++r) {
      const int64_t base = r * chunk + (int64_t)lb * PBX_A2A_THREADS * KPT + tid;
      if (tid < R) s_cnt[tid] = 0;
      __syncthreads();
      uint64_t key[KPT];
      uint32_t pos[KPT], lrank[KPT];
      int owner[KPT];
      bool is_rep[KPT], done[KPT];
#pragma unroll
      for (int u = 0; u < KPT; ++u) {
        const int64_t i = base + (int64_t)u * PBX_A2A_THREADS;
        is_rep[u] = false;
        done[u] = i >= a.n;
        key[u] = done[u] ? 0 : a.keys[i];
        pos[u] = (uint32_t)pbx_mix64(key[u]) & mask;
      }
      unsigned long long prev[KPT];
#pragma unroll
      for (int u = 0; u < KPT; ++u)      // first probe of every key: independent atomics, all in flight
        if (!done[u])
          prev[u] = atomicCAS(reinterpret_cast<unsigned long long*>(c.dd_keys + pos[u]),
                              (unsigned long long)PBX_EMPTY_KEY, (unsigned long long)key[u]);
#pragma unroll
      for (int u = 0; u < KPT; ++u) {
        if (done[u]) continue;
        while (true) {
          if (prev[u] == PBX_EMPTY_KEY) { is_rep[u] = true; break; }
          if (prev[u] == key[u]) break;
          pos[u] = (pos[u] + 1) & mask;
          prev[u] = atomicCAS(reinterpret_cast<unsigned long long*>(c.dd_keys + pos[u]),
                              (unsigned long long)PBX_EMPTY_KEY, (unsigned long long)key[u]);
        }
        a.restore[base + (int64_t)u * PBX_A2A_THREADS] = pos[u];
        c.occ_rep[base + (int64_t)u * PBX_A2A_THREADS] = is_rep[u] ? 1 : 0;
      }
#pragma unroll
      for (int u = 0; u < KPT; ++u) owner[u] = is_rep[u] ? pbx_owner_of(key[u], R) : -1;
      // rank of every representative among its CTA's keys for the same owner: one shared-memory
      // atomic per (warp, owner) instead of one per key (with few ranks thousands of lanes would
      // otherwise queue on the same shared word)
      {
        const unsigned lane = tid & 31;
#pragma unroll
        for (int u = 0; u < KPT; ++u) {
          for (int o = 0; o < R; ++o) {
            const unsigned m = __ballot_sync(0xffffffffu, owner[u] == o);
            if (m == 0) continue;
            const int leader = __ffs(m) - 1;
            uint32_t b0 = 0;
            if ((int)lane == leader) b0 = atomicAdd(&s_cnt[o], (uint32_t)__popc(m));
            b0 = __shfl_sync(0xffffffffu, b0, leader);
            if (owner[u] == o) lrank[u] = b0 + __popc(m & ((1u << lane) - 1u));
          }
        }
      }
      __syncthreads();
      if (tid < R) s_base[tid] = s_cnt[tid] ? atomicAdd(c.send_cnt + tid, s_cnt[tid]) : 0u;
      __syncthreads();
#pragma unroll
      for (int u = 0; u < KPT; ++u) {
        if (!is_rep[u]) continue;
        const uint32_t slot = s_base[owner[u]] + lrank[u];
        if (slot < cap - 1) {
          c.dd_uid[pos[u]] = (uint32_t)owner[u] * cap + slot;
          c.mail_keys[owner[u]][(uint64_t)c.rank * cap + slot] = key[u];  // remote store (NVLink)
        } else {
          // Mailbox overflow: the key goes to the reserved LAST row of this owner's region, which is
          // never shipped — its pulled values stay zero and its merged gradient is dropped — so no
          // other key's values or table row are touched.  status 1 makes the next step fail loudly.
          c.dd_uid[pos[u]] = (uint32_t)owner[u] * cap + (cap - 1);
          atomicExch(c.status, 1);
        }
      }
      __syncthreads();
    }
  }
  PBX_TSB(1);
  grid.sync();
  PBX_TS(2);

  // ---- P2: occurrence -> unique row, publish counts, clean the dedup table
  {
    const int64_t stride = (int64_t)bpr * PBX_A2A_THREADS;
    const uint32_t tag = (uint32_t)(e0 + 1);
    for (int64_t i = (int64_t)lb * PBX_A2A_THREADS + tid; i < a.n; i += stride) {
      const uint32_t uid = c.dd_uid[a.restore[i]];
      a.restore[i] = uid;
      // a non-representative occurrence marks its key as "merged from several occurrences": the push
      // sends single-occurrence records straight to their owner and merges only the marked ones locally
      if (!c.occ_rep[i]) c.uid_multi[uid] = tag;
    }
    if (lb == 0 && tid == 0) c.epoch[5] = tag;
    if (lb == 0 && tid < R) {
      uint32_t n_send = min(c.send_cnt[tid], cap - 1);   // row cap-1 is the overflow row (see P1)
      c.send_cnt[tid] = n_send;
      c.mail_cnt[tid][c.rank] = n_send;  // remote store
    }
  }
  PBX_TSB(2);
  PBX_TS(3);
  PBX_TSB(3);
  xbarrier<EMU>(grid, c, e0 + 1, lb == 0);
  PBX_TS(4);

  // ---- P3: owner side — probe the local shard, return rows into the requester's buffer
  {
    const int lane = tid & 31;
    const int warp = lb * (PBX_A2A_THREADS / 32) + (tid >> 5);
    const int nwarps = bpr * (PBX_A2A_THREADS / 32);
    const int PS = c.pull_stride;
    const int chunks = PS >> 2;
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    // All-to-all without incast: the warps are split into R groups, group g serves requester
    // (rank + 1 + g) % R, so at any moment every owner is writing to every requester at the same
    // rate (with "everybody serves requester 0 first" the eight owners would pile onto one GPU's
    // NVLink ingress while the other seven sat idle).
    if (lb == 0 && tid < R) c.serve_cnt[tid] = min(__ldcg(c.mail_cnt[c.rank] + tid), cap);
    const int nper = nwarps >= R ? nwarps / R : nwarps;
    const int grp = nwarps >= R ? warp % R : 0;
    const int widx = nwarps >= R ? warp / R : warp;
    // Tiles of 32 keys are handed out through one ticket counter per requester (send_cnt[32 + s],
    // zeroed in P0): a warp that drew cheap tiles takes more of them, so the phase ends when the
    // work does instead of when the unluckiest warp of a static split is done.  The next ticket is
    // drawn before the current tile is processed (its round trip hides behind the tile).
    for (int sr = 0; sr < (nwarps >= R ? 1 : R); ++sr) {
      if (nwarps >= R && widx >= nper) break;
      const int s = (c.rank + 1 + grp + sr) % R;
      const uint32_t cnt = min(__ldcg(c.mail_cnt[c.rank] + s), cap);
      const uint64_t* mk = c.mail_keys[c.rank] + (uint64_t)s * cap;
      uint32_t* mrow = c.mail_row + (uint64_t)s * cap;
      float* dst = c.pull_vals[s] + (uint64_t)c.rank * cap * PS;  // requester s, my region
      uint32_t* ticket = c.send_cnt + 32 + s;
      uint32_t tk = 0;
      if (lane == 0) tk = atomicAdd(ticket, 1u);
      tk = __shfl_sync(0xffffffffu, tk, 0);
      while ((uint64_t)tk * 32u < cnt) {
        const uint32_t j0 = tk * 32u;
        uint32_t tk_next = 0;
        if (lane == 0) tk_next = atomicAdd(ticket, 1u);
        uint32_t j = j0 + lane;
        uint32_t row = PBX_NO_ROW;
        if (j < cnt) {
          uint64_t key = __ldcg(mk + j);
          if (key != 0) row = pbx_dev_find(c.table, key);
          mrow[j] = row;
        }
        const uint32_t nvalid = min(32u, cnt - j0);
        const uint32_t items = nvalid * (uint32_t)chunks;
        // UN independent 16-byte row loads per lane are issued before the first remote store needs
        // its data (a store right behind its load would hold the warp for a full HBM round trip per
        // chunk: 128 lines in flight per SM instead of the several hundred the memory system takes)
        constexpr int UN = 5;
        for (uint32_t it0 = 0; it0 < items; it0 += 32 * UN) {
          float4 v[UN];
#pragma unroll
          for (int u = 0; u < UN; ++u) {
            const uint32_t it = it0 + u * 32 + lane;
            const uint32_t kk = min(it / (uint32_t)chunks, 31u);
            const uint32_t ch = it % (uint32_t)chunks;
            const uint32_t r = __shfl_sync(0xffffffffu, row, kk);
            v[u] = zero4;
            if (it < items && r != PBX_NO_ROW)
              v[u] = __ldcg(reinterpret_cast<const float4*>(c.table.vals + (uint64_t)r * c.table.L.stride + ch * 4));
          }
#pragma unroll
          for (int u = 0; u < UN; ++u) {
            const uint32_t it = it0 + u * 32 + lane;
            if (it < items) {
              const uint32_t kk = it / (uint32_t)chunks, ch = it % (uint32_t)chunks;
              *reinterpret_cast<float4*>(dst + (uint64_t)(j0 + kk) * PS + ch * 4) = v[u];  // remote store
            }
          }
        }
        tk = __shfl_sync(0xffffffffu, tk_next, 0);
      }
    }
  }
  PBX_TSB(4);
  xbarrier<EMU>(grid, c, e0 + 2, lb == 0);
  PBX_TS(5);
  if (lb == 0 && tid == 0) *c.epoch = e0 + 2;
}

// Q0 of the push, fast path (pooled gradients, cvm_offset 2): the representative occurrence of every
// unique key stores its record, straight into the owner's mailbox when the key occurs once in the
// batch.  `s_bag`: this warp's shared-memory tile of 16 * PD floats.  Runs either inside k_sparse_push
// (virtual-rank tests) or, in production, as its own launch k_push_records right before it: the
// phase is a chain of dependent index / gradient loads per warp, and on its own it gets 64 warps per
// SM to hide them instead of the 16 the optimizer phase of the cooperative kernel leaves.
__device__ __forceinline__ void push_q0_direct(const PbxSparseDev& c, const PbxPushArgs& a, float* s_bag,
                                               int64_t gtid, int64_t gstride, uint32_t tag) {
  const int R = c.nranks;
  const uint32_t cap = c.cap;
  const int PD = c.push_stride;
  const int pchunks = PD >> 2;
  {
    // A warp takes TB consecutive bags.  All keys of one bag push the SAME record (the bag's output
    // gradient + the instance's show / click), so the warp first builds the TB bag records in shared
    // memory with coalesced loads of the gradient rows, then hands them out: every lane owns one key
    // occurrence (unique row and representative flag are coalesced loads, its bag is found by a
    // shuffle search over the lod values the warp holds) and the records are copied ITEM-wise — item
    // = (key, 16-byte chunk), adjacent lanes = adjacent chunks of one record — from shared memory to
    // wherever the key's record lives.  No global load sits on the per-key path but two index loads.
    const int hidden = 3 + c.table.L.embedx_dim;
    const int lane = threadIdx.x & 31;
    constexpr int TB = 16;                   // bags per warp task (several tasks per warp: even split)
    const int64_t nbags = (int64_t)a.S * a.B;
    const int64_t ntask = (nbags + TB - 1) / TB;
    const int64_t wid = gtid >> 5, nw = gstride >> 5;
    for (int64_t t = wid; t < ntask; t += nw) {
      const int64_t g0 = t * TB;
      const int32_t Lb = __ldg(a.lod + min(g0 + min(lane, TB), nbags));
      const int32_t kb = __shfl_sync(0xffffffffu, Lb, 0);
      const int32_t ke = __shfl_sync(0xffffffffu, Lb, TB);
      if (ke <= kb) continue;
      __syncwarp();
      // (batches of 5 rounds: all gradient loads of a batch are issued before the first shared-memory
      //  store waits for its value — a store right behind its load would serialise one L2 round trip
      //  per round)
      for (int base = 0; base < TB * PD; base += 32 * 5) {
        float val[5];
#pragma unroll
        for (int u = 0; u < 5; ++u) {
          const int idx = base + u * 32 + lane;
          val[u] = 0.f;
          if (idx >= TB * PD) continue;
          const int l = idx / PD, w = idx - l * PD;
          // (32-bit arithmetic on purpose: a 64-bit division by a runtime value is ~100 instructions, and
          //  with one per element it, not memory, was what this phase cost)
          const uint32_t gg = (uint32_t)min(g0 + l, nbags - 1);
          const uint32_t s = gg / (uint32_t)a.B, b = gg - s * (uint32_t)a.B;
          const int col = w - 1;             // push word w <-> value column w - 1, word 0 = slot id
          if (col < 0) val[u] = (float)s;
          else if (col >= hidden) val[u] = 0.f;
          else if (col < 2) val[u] = __ldg(a.cvm + (int64_t)b * 2 + col);
          else val[u] = a.grad_scale * __ldg(a.dout + (int64_t)b * a.ld_dout + (int64_t)s * a.Eo + (col - a.dim_off));
        }
#pragma unroll
        for (int u = 0; u < 5; ++u) {
          const int idx = base + u * 32 + lane;
          if (idx < TB * PD) s_bag[idx] = val[u];
        }
      }
      __syncwarp();
      for (int32_t k0 = kb; k0 < ke; k0 += 32) {
        const int32_t k = k0 + lane;
        const bool valid = k < ke;
        const int32_t kc = valid ? k : ke - 1;
        const uint32_t uid = __ldg(a.restore + kc);
        const bool rep = c.occ_rep[kc] != 0;
        int lo = 0, hi = TB - 1;             // largest l with lod[g0 + l] <= k
#pragma unroll
        for (int step = 0; step < 4; ++step) {
          const int mid = (lo + hi + 1) >> 1;
          const int32_t Lm = __shfl_sync(0xffffffffu, Lb, mid);
          if (Lm <= kc) lo = mid; else hi = mid - 1;
        }
        if (valid && !rep) {                 // merged in Q1 from the list
          const uint32_t at = atomicAdd(c.send_cnt + 48, 1u);
          c.dup_list[2 * (uint64_t)at] = (uint32_t)k;
          c.dup_list[2 * (uint64_t)at + 1] = (uint32_t)(g0 + lo);
        }
        float* dstp = nullptr;               // destination of this lane's record (representatives only)
        if (valid && rep) {
          const uint32_t o = uid / cap;
          if ((int)o == c.rank || R == 1 || c.uid_multi[uid] == tag) {
            dstp = c.push_acc + (uint64_t)uid * PD;                  // local shard, or merged locally first
          } else {
            // the key occurs once in this batch: its record is final — it goes straight into the owner's
            // mailbox over NVLink; the transfer hides behind this phase instead of being a phase of its own
            dstp = c.push_mail[o] + ((uint64_t)c.rank * cap + (uid - o * cap)) * PD;
          }
        }
        const unsigned long long dst64 = (unsigned long long)dstp;
        const int nkeys = min(32, (int)(ke - k0));
        const int items = nkeys * pchunks;
        for (int i0 = 0; i0 < items; i0 += 32) {
          const int i = i0 + lane;
          const int q = i / pchunks;
          const int r = min(q, 31), ch = i - q * pchunks;
          const int lr = __shfl_sync(0xffffffffu, lo, r);
          const unsigned long long dr = __shfl_sync(0xffffffffu, dst64, r);
          if (i < items && dr != 0ull)
            *reinterpret_cast<float4*>(reinterpret_cast<float*>(dr) + ch * 4) =
                *reinterpret_cast<const float4*>(s_bag + lr * PD + ch * 4);
        }
      }
    }
  }
}

// stand-alone launch of the same phase (any grid; dynamic shared memory = warps per CTA * 16 * PD floats)
__global__ void __launch_bounds__(256) k_push_records(const PbxSparseDev* __restrict__ ctx, PbxPushArgs a) {
  extern __shared__ float4 s_rec[];
  const PbxSparseDev& c = *ctx;
  float* s_bag = reinterpret_cast<float*>(s_rec) + (size_t)(threadIdx.x >> 5) * 16 * c.push_stride;
  push_q0_direct(c, a, s_bag, (int64_t)blockIdx.x * blockDim.x + threadIdx.x, (int64_t)gridDim.x * blockDim.x,
                 (uint32_t)c.epoch[5]);
}

template <bool EMU>
__global__ void __launch_bounds__(PBX_PUSH_THREADS, 1)
k_sparse_push(const PbxSparseDev* __restrict__ ctxs, int nvr, PbxPushArgsN args) {
  cg::grid_group grid = cg::this_grid();
  const int bpr = gridDim.x / nvr;
  const int vr = blockIdx.x / bpr;
  const int lb = blockIdx.x % bpr;
  const PbxSparseDev& c = ctxs[vr];
  const PbxPushArgs a = args.a[vr];
  const int R = c.nranks;
  const uint32_t cap = c.cap;
  const int tid = threadIdx.x;
  const int PD = c.push_stride;
  const int pchunks = PD >> 2;
  const int64_t gtid = (int64_t)lb * PBX_PUSH_THREADS + tid;
  const int64_t gstride = (int64_t)bpr * PBX_PUSH_THREADS;
  const uint64_t e0 = *c.epoch;
  PBX_TS(8);

  // ---- Q0: initialise the merge rows in use.  Fast path (pooled gradients, cvm_offset 2): the
  // representative occurrence of every unique key STORES its record (no zero pass, no atomics);
  // other paths zero the rows and merge everything with atomics in Q1.
  // fast path: pooled gradients, cvm_offset 2, and a shared-memory tile per warp that holds 16 bag records
  const bool direct = a.mode == 0 && a.cvm_offset == 2 && c.q3_tile_rows > 0 && 16 * PD <= c.q3_warp_words;
  const uint32_t tag = (uint32_t)c.epoch[5];               // tag of the pull this push belongs to
  if (direct) {
    if (!a.q0_done) {
      extern __shared__ float4 s_q3[];
      push_q0_direct(c, a, reinterpret_cast<float*>(s_q3) + (size_t)(tid >> 5) * c.q3_warp_words, gtid, gstride, tag);
    }
  } else {
    for (int o = 0; o < R; ++o) {
      const int64_t items = (int64_t)c.send_cnt[o] * pchunks;
      float4* base = reinterpret_cast<float4*>(c.push_acc + (uint64_t)o * cap * PD);
      for (int64_t it = gtid; it < items; it += gstride) base[it] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  PBX_TSB(8);
  grid.sync();
  PBX_TS(9);

  // ---- Q1: merge by unique key (fused with the seqpool/CVM backward in mode 0)
  if (direct) {
    // In CTR batches almost every unique key occurs once, so the merge is two passes instead of one
    // atomic per word: Q0 STORED the record of every representative occurrence (the first claimer of
    // its key in the pull's dedup); here the few remaining occurrences, listed by Q0, are added with
    // atomics / 16-byte vector reductions.
    // (reference: PushMergeCopy vs PushMergeCopyAtomic chosen by total < 2 * uniq, box_wrapper.cu:1034)
    const int hidden = 3 + c.table.L.embedx_dim;
    const uint32_t ndup = c.send_cnt[48];
    const int64_t items = (int64_t)ndup * pchunks;
    for (int64_t it = gtid; it < items; it += gstride) {
      const int ch = (int)(it % pchunks);
      const int64_t e = it / pchunks;
      const uint32_t k = c.dup_list[2 * e], g = c.dup_list[2 * e + 1];
      const uint32_t s = g / (uint32_t)a.B, b = g - s * (uint32_t)a.B;
      float v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int col = ch * 4 + j - 1;  // push word w <-> value column w - 1
        if (col < 0 || col >= hidden) v[j] = 0.f;
        else if (col < 2) v[j] = __ldg(a.cvm + (int64_t)b * 2 + col);
        else v[j] = a.grad_scale * __ldg(a.dout + (int64_t)b * a.ld_dout + (int64_t)s * a.Eo + (col - a.dim_off));
      }
      float* rec = c.push_acc + (uint64_t)__ldg(a.restore + k) * PD + ch * 4;
      if (ch == 0) {
        atomicAdd(rec + 1, v[1]);
        atomicAdd(rec + 2, v[2]);
        atomicAdd(rec + 3, v[3]);
      } else {
        pbx_red_add_v4(rec, v[0], v[1], v[2], v[3]);
      }
    }
  } else if (a.mode == 0) {
    const int hidden = 3 + c.table.L.embedx_dim;
    const int64_t items = (int64_t)a.S * a.B * hidden;
    for (int64_t it = gtid; it < items; it += gstride) {
      const int col = (int)(it % hidden);
      const int64_t g = it / hidden;  // bag = slot * B + ins
      const int s = (int)(g / a.B);
      const int b = (int)(g % a.B);
      const int32_t k0 = a.lod[g], k1 = a.lod[g + 1];
      if (k0 >= k1) continue;
      float v;
      if (col < a.cvm_offset) v = a.cvm[(int64_t)b * a.cvm_offset + col];
      else v = a.grad_scale * a.dout[(int64_t)b * a.ld_dout + (int64_t)s * a.Eo + (col - a.dim_off)];
      for (int32_t k = k0; k < k1; ++k) {
        float* rec = c.push_acc + (uint64_t)a.restore[k] * PD;
        atomicAdd(rec + col + 1, v);
        if (col == 0) rec[0] = (float)s;
      }
    }
  } else {
    const int hidden = a.hidden;
    const int64_t items = a.n * hidden;
    for (int64_t it = gtid; it < items; it += gstride) {
      const int col = (int)(it % hidden);
      const int64_t i = it / hidden;
      float v = a.gkeys[it];
      if (col >= a.cvm_offset) v *= a.grad_scale;
      float* rec = c.push_acc + (uint64_t)a.restore[i] * PD;
      atomicAdd(rec + col + 1, v);
      if (col == 0) rec[0] = (float)a.key2slot[i];
    }
  }
  PBX_TSB(9);
  grid.sync();
  PBX_TS(10);

  // ---- Q2: ship merged records to their owners
  // (records for the local shard are applied straight from push_acc).  The CTAs are split over
  // the R-1 remote owners — block lb ships to owner (rank + 1 + lb % (R-1)) % R — so every pair of
  // GPUs exchanges at the same rate instead of all ranks hitting owner 0 first.
  if (R > 1) {
    const int nd = R - 1;
    const int di = bpr >= nd ? lb % nd : 0;
    const int bl = bpr >= nd ? lb / nd : lb;
    const int nb = bpr >= nd ? (bpr - di + nd - 1) / nd : bpr;
    for (int rep = 0; rep < (bpr >= nd ? 1 : nd); ++rep) {
      const int o = (c.rank + 1 + di + rep) % R;
      const int64_t items = (int64_t)c.send_cnt[o] * pchunks;
      const float4* src = reinterpret_cast<const float4*>(c.push_acc + (uint64_t)o * cap * PD);
      float4* dst = reinterpret_cast<float4*>(c.push_mail[o] + (uint64_t)c.rank * cap * PD);  // remote
      // four independent 16-byte moves per thread in flight (local L2 read -> NVLink store)
      const int64_t step = (int64_t)nb * PBX_PUSH_THREADS;
      if (direct) {
        // only the records merged from several occurrences are still here; the others went out in Q0
        const uint32_t* multi = c.uid_multi + (uint64_t)o * cap;
        for (int64_t it = (int64_t)bl * PBX_PUSH_THREADS + tid; it < items; it += step)
          if (multi[it / pchunks] == tag) dst[it] = __ldcg(src + it);
        continue;
      }
      for (int64_t it = (int64_t)bl * PBX_PUSH_THREADS + tid; it < items; it += 4 * step) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (it + u * step < items) v[u] = __ldcg(src + it + u * step);
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (it + u * step < items) dst[it + u * step] = v[u];
      }
    }
  }
  PBX_TSB(10);
  xbarrier<EMU>(grid, c, e0 + 1, lb == 0);
  PBX_TS(11);
  if (lb == 0 && tid == 0) c.send_cnt[48] = 0;   // every CTA is past Q1: the duplicate list is consumed

  // ---- folded dense allreduce (one-shot per slice over peer memory, scale fused)
  if (a.dense_ctx) {
    const PbxSymmDev& d = *a.dense_ctx;
    const int64_t n = a.dense_n, off = a.dense_off;
    const int64_t chunk = ((n + R - 1) / R + 3) / 4 * 4;
    const int64_t b = min(n, (int64_t)d.rank * chunk), e = min(n, b + chunk);
    const int64_t nv = (e - b) / 4;
    for (int64_t i = gtid; i < nv; i += gstride) {
      const int64_t idx = off + b + i * 4;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int p = 0; p < PBX_MAX_RANKS; ++p)
        if (p < R) {
          const float4 v = __ldcg(reinterpret_cast<const float4*>(d.data[(d.rank + p) % R] + idx));
          acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
      acc.x *= a.dense_scale; acc.y *= a.dense_scale; acc.z *= a.dense_scale; acc.w *= a.dense_scale;
#pragma unroll
      for (int p = 0; p < PBX_MAX_RANKS; ++p)
        if (p < R) *reinterpret_cast<float4*>(d.data[(d.rank + p) % R] + idx) = acc;
    }
    for (int64_t i = b + nv * 4 + gtid; i < e; i += gstride) {
      float acc = 0.f;
      for (int p = 0; p < R; ++p) acc += __ldcg(d.data[p] + off + i);
      acc *= a.dense_scale;
      for (int p = 0; p < R; ++p) d.data[p][off + i] = acc;
    }
  }

  // ---- Q3: owner applies the fused sparse optimizer.
  // Rows of one 128-byte line (every optimizer): WARP TILES over the flattened list of all
  // requesters' records — a warp takes 32 records of one requester, loads their rows with
  // coalesced 128-byte requests (8 lanes x 16 B per row) into shared memory, ONE lane per row
  // runs the optimizer functor on the shared copy, and the rows go back with coalesced stores.
  // A key pushed by several requesters is serialised by a try-lock in the spare last word of its
  // row (no hold-and-wait: a lane that loses the race retries after the warp has released its
  // own locks), which replaces the reference's "one GPU at a time under the table write lock"
  // (heter_comm_inl.h:1779-1804) without a grid barrier per requester.
  const PbxTableView& T = c.table;
  if (T.test_mode) {
    // pushes are ignored in test mode
  } else if (T.L.stride == 32 && PD <= 24) {
    extern __shared__ float4 s_q3[];
    constexpr int RS = 36;                       // padded row stride in words: conflict-free float4
    const int lane = tid & 31, wib = tid >> 5;
    float* s_rows = reinterpret_cast<float*>(s_q3) + wib * (32 * RS + 32 * 24);
    float* s_push = s_rows + 32 * RS;
    const int64_t warp = (int64_t)lb * (PBX_PUSH_THREADS / 32) + wib;
    const int64_t nwarps = (int64_t)bpr * (PBX_PUSH_THREADS / 32);
    const int sub = lane & 7, rq = lane >> 3;
    const bool locking = R > 1;
    int64_t tile_base = 0;                       // tiles of requester s are [tile_base, tile_base + ntiles)
    for (int s = 0; s < R; ++s) {
      const uint32_t cnt = c.serve_cnt[s];
      const int64_t ntiles = ((int64_t)cnt + 31) >> 5;
      const float* pm = s == c.rank ? c.push_acc + (uint64_t)s * cap * PD
                                    : c.push_mail[c.rank] + (uint64_t)s * cap * PD;
      const uint32_t* mrow = c.mail_row + (uint64_t)s * cap;
      // first tile of this requester owned by this warp (global round-robin over all tiles)
      int64_t t0 = (warp - tile_base % nwarps + nwarps) % nwarps;
      for (int64_t tl = t0; tl < ntiles; tl += nwarps) {
        const int64_t j0 = tl * 32;
        const int nvalid = (int)min((int64_t)32, (int64_t)cnt - j0);
        const uint32_t myrow = lane < nvalid ? __ldg(mrow + j0 + lane) : PBX_NO_ROW;
        const float4* psrc = reinterpret_cast<const float4*>(pm + j0 * PD);
        const int pvec = nvalid * (PD >> 2);
        for (int idx = lane; idx < pvec; idx += 32)
          reinterpret_cast<float4*>(s_push)[idx] = __ldcg(psrc + idx);
        bool pending = myrow != PBX_NO_ROW;
        while (__any_sync(0xffffffffu, pending)) {
          bool got = pending;
          if (locking && pending) {
            unsigned int* lk = reinterpret_cast<unsigned int*>(T.vals + (uint64_t)myrow * 32 + 31);
            got = atomicCAS(lk, 0u, 1u) == 0u;
            // (no acquire fence: the row is read with ld.cg from L2, the coherence point, in program
            //  order after the CAS; the previous owner's stores reached L2 before its unlock)
          }
          const unsigned gmask = __ballot_sync(0xffffffffu, got);
          float4 rv[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {          // all eight row loads in flight before any use
            const int rr = rq + 4 * i;
            const uint32_t r = __shfl_sync(0xffffffffu, myrow, rr);
            rv[i] = ((gmask >> rr) & 1u) ? __ldcg(reinterpret_cast<const float4*>(T.vals + (uint64_t)r * 32 + 4 * sub))
                                         : make_float4(0.f, 0.f, 0.f, 0.f);
          }
#pragma unroll
          for (int i = 0; i < 8; ++i)
            *reinterpret_cast<float4*>(s_rows + (rq + 4 * i) * RS + 4 * sub) = rv[i];
          __syncwarp();
          bool touched = false;
          if (got) touched = pbx_apply_push_ptr(T, s_rows + lane * RS, myrow, s_push + lane * PD);
          __syncwarp();
          const unsigned tmask = __ballot_sync(0xffffffffu, touched);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int rr = rq + 4 * i;
            const uint32_t r = __shfl_sync(0xffffffffu, myrow, rr);
            // (the lock word, lane 7 / word 3 of the last chunk, is written back as "held": its
            //  shared copy was loaded after the lock was taken; the release below clears it)
            if ((tmask >> rr) & 1u)
              *reinterpret_cast<float4*>(T.vals + (uint64_t)r * 32 + 4 * sub) =
                  *reinterpret_cast<const float4*>(s_rows + rr * RS + 4 * sub);
          }
          if (touched && T.dirty) T.dirty[myrow] = 1;
          if (locking) {
            __threadfence();                     // release: the row stores above before the unlock
            __syncwarp();
            if (got) atomicExch(reinterpret_cast<unsigned int*>(T.vals + (uint64_t)myrow * 32 + 31), 0u);
          }
          if (got) pending = false;
          __syncwarp();
        }
      }
      tile_base += ntiles;
    }
  } else if (c.q3_tile_rows >= 2) {
    // Rows of several 128-byte lines: the same warp-tile scheme with TR (< 32) rows per tile —
    // the row and its push record are staged with coalesced 16-byte accesses, lanes [0, TR) run
    // the optimizer functor on the shared copy, the tile is written back coalesced.
    extern __shared__ float4 s_q3[];
    const int ST = T.L.stride, RS = ST + 4, SV = ST >> 2, TR = c.q3_tile_rows;
    const int lane = tid & 31, wib = tid >> 5;
    float* s_rows = reinterpret_cast<float*>(s_q3) + (size_t)wib * c.q3_warp_words;
    float* s_push = s_rows + TR * RS;
    const int64_t warp = (int64_t)lb * (PBX_PUSH_THREADS / 32) + wib;
    const int64_t nwarps = (int64_t)bpr * (PBX_PUSH_THREADS / 32);
    const bool locking = R > 1;
    const int nvec = TR * SV;
    int64_t tile_base = 0;
    for (int s = 0; s < R; ++s) {
      const uint32_t cnt = c.serve_cnt[s];
      const int64_t ntiles = ((int64_t)cnt + TR - 1) / TR;
      const float* pm = s == c.rank ? c.push_acc + (uint64_t)s * cap * PD
                                    : c.push_mail[c.rank] + (uint64_t)s * cap * PD;
      const uint32_t* mrow = c.mail_row + (uint64_t)s * cap;
      int64_t t0 = (warp - tile_base % nwarps + nwarps) % nwarps;
      for (int64_t tl = t0; tl < ntiles; tl += nwarps) {
        const int64_t j0 = tl * TR;
        const int nvalid = (int)min((int64_t)TR, (int64_t)cnt - j0);
        const uint32_t myrow = lane < nvalid ? __ldg(mrow + j0 + lane) : PBX_NO_ROW;
        const float4* psrc = reinterpret_cast<const float4*>(pm + j0 * PD);
        const int pvec = nvalid * (PD >> 2);
        for (int idx = lane; idx < pvec; idx += 32)
          reinterpret_cast<float4*>(s_push)[idx] = __ldcg(psrc + idx);
        bool pending = myrow != PBX_NO_ROW;
        while (__any_sync(0xffffffffu, pending)) {
          bool got = pending;
          if (locking && pending) {
            unsigned int* lk = reinterpret_cast<unsigned int*>(T.vals + (uint64_t)myrow * ST + ST - 1);
            got = atomicCAS(lk, 0u, 1u) == 0u;
          }
          const unsigned gmask = __ballot_sync(0xffffffffu, got);
          for (int base = 0; base < nvec; base += 128) {   // 4 independent 16-byte loads per lane in flight
            float4 rv[4];
            int rr[4], ch[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int idx = base + i * 32 + lane;
              rr[i] = min(idx / SV, TR - 1);
              ch[i] = idx - rr[i] * SV;
              const uint32_t r = __shfl_sync(0xffffffffu, myrow, rr[i]);
              rv[i] = (idx < nvec && ((gmask >> rr[i]) & 1u))
                          ? __ldcg(reinterpret_cast<const float4*>(T.vals + (uint64_t)r * ST) + ch[i])
                          : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i)
              if (base + i * 32 + lane < nvec) *reinterpret_cast<float4*>(s_rows + rr[i] * RS + 4 * ch[i]) = rv[i];
          }
          __syncwarp();
          bool touched = false;
          if (got) touched = pbx_apply_push_ptr(T, s_rows + lane * RS, myrow, s_push + lane * PD);
          __syncwarp();
          const unsigned tmask = __ballot_sync(0xffffffffu, touched);
          for (int base = 0; base < nvec; base += 32) {
            const int idx = base + lane;
            const int rr = min(idx / SV, TR - 1);
            const int ch = idx - rr * SV;
            const uint32_t r = __shfl_sync(0xffffffffu, myrow, rr);
            if (idx < nvec && ((tmask >> rr) & 1u))
              *(reinterpret_cast<float4*>(T.vals + (uint64_t)r * ST) + ch) =
                  *reinterpret_cast<const float4*>(s_rows + rr * RS + 4 * ch);
          }
          if (touched && T.dirty) T.dirty[myrow] = 1;
          if (locking) {
            __threadfence();
            __syncwarp();
            if (got) atomicExch(reinterpret_cast<unsigned int*>(T.vals + (uint64_t)myrow * ST + ST - 1), 0u);
          }
          if (got) pending = false;
          __syncwarp();
        }
      }
      tile_base += ntiles;
    }
  } else {
    // generic path (rows too wide for a shared-memory tile): one requester at a time, separated by grid barriers
    for (int s = 0; s < R; ++s) {
      const uint32_t cnt = c.serve_cnt[s];
      const float* pm = s == c.rank ? c.push_acc + (uint64_t)s * cap * PD
                                    : c.push_mail[c.rank] + (uint64_t)s * cap * PD;
      const uint32_t* mrow = c.mail_row + (uint64_t)s * cap;
      for (int64_t j = gtid; j < (int64_t)cnt; j += gstride) {
        uint32_t row = mrow[j];
        if (row == PBX_NO_ROW) continue;
        // (when the key is needed — lazy embedx creation — it is read from the table itself: a
        // faster peer may already be overwriting this rank's key mailbox with the next batch)
        pbx_apply_push_fast(T, row, pm + j * PD);
      }
      if (s + 1 < R) grid.sync();
    }
  }
  PBX_TSB(11);
  PBX_TS(12);
  if (a.dense_ctx) {
    xbarrier<EMU>(grid, c, e0 + 2, lb == 0);   // every peer's reduced slices have landed
    if (lb == 0 && tid == 0) *c.epoch = e0 + 2;
  } else {
    if (lb == 0 && tid == 0) *c.epoch = e0 + 1;
  }
}

}  // namespace

// ---------------------------------------------------------------------------
// Host object
// ---------------------------------------------------------------------------
struct PbxSparseCtx {
  PbxSparseDev h;          // host copy of the device context
  PbxSparseDev* d;         // device copy
  bool dirty;
  int device;
  void* symm_base;         // one allocation holding every symmetric buffer
  size_t symm_bytes;
  size_t off_flags, off_cnt, off_keys, off_vals, off_push;
  void* peer_base[PBX_MAX_RANKS];
  bool peer_opened[PBX_MAX_RANKS];
  int64_t max_keys;
  int* status_host;        // pinned mirror of h.status, refreshed by a 4-byte copy after every push
  int grid_blocks;         // cooperative grid size (per real rank)
  void* table;             // PbxTable*
};

static size_t pbx_push_smem(const PbxSparseDev& h) {
  return h.q3_tile_rows == 0 ? 0 : (size_t)(PBX_PUSH_THREADS / 32) * h.q3_warp_words * 4;
}

static void pbx_ctx_bind_peer(PbxSparseCtx* c, int p, void* base) {
  char* b = (char*)base;
  c->peer_base[p] = base;
  c->h.flags[p] = (uint64_t*)(b + c->off_flags);
  c->h.mail_cnt[p] = (uint32_t*)(b + c->off_cnt);
  c->h.mail_keys[p] = (uint64_t*)(b + c->off_keys);
  c->h.pull_vals[p] = (float*)(b + c->off_vals);
  c->h.push_mail[p] = (float*)(b + c->off_push);
  c->dirty = true;
}

extern "C" {

// Creates the per-rank engine: allocates the symmetric region and local scratch.
//   max_keys : largest number of key occurrences in one batch on this rank
//   cap      : mailbox capacity per (src,dst) pair (0 = choose: all of max_keys for 1 rank,
//              else 2x the even share + slack)
int pbx_sparse_create(void** out, void* table, int rank, int nranks, int64_t max_keys, int64_t cap) {
  PBX_REQUIRE(nranks >= 1 && nranks <= PBX_MAX_RANKS, "nranks %d out of range (max %d)", nranks, PBX_MAX_RANKS);
  PBX_REQUIRE(rank >= 0 && rank < nranks, "bad rank %d", rank);
  PBX_REQUIRE(max_keys > 0, "max_keys must be positive");
  PbxTable* t = (PbxTable*)table;
  PbxSparseCtx* c = new PbxSparseCtx();
  memset(c, 0, sizeof(*c));
  cudaGetDevice(&c->device);
  c->table = table;
  c->max_keys = max_keys;
  // (+1: the last row of every owner region is the overflow row, see k_sparse_pull P1)
  if (cap <= 0) cap = nranks == 1 ? max_keys + 1 : (2 * max_keys / nranks + 1024);
  if (cap > max_keys + 1) cap = max_keys + 1;
  if (cap < 2) cap = 2;
  PBX_REQUIRE((uint64_t)cap * nranks < 0xFFFFFFF0ull, "mailbox too large");
  PbxSparseDev& h = c->h;
  h.rank = rank;
  h.nranks = nranks;
  h.cap = (uint32_t)cap;
  uint64_t dd = 1024;
  while (dd < (uint64_t)max_keys * 2) dd <<= 1;
  h.dd_cap = (uint32_t)dd;
  h.pull_stride = t->view.L.pull_stride;
  h.push_stride = t->view.L.push_stride;
  h.table = t->view;
  auto al = [](size_t x) { return (x + 255) / 256 * 256; };
  size_t off = 0;
  c->off_flags = off; off += al(64 * sizeof(uint64_t));
  c->off_cnt = off;   off += al(64 * sizeof(uint32_t));
  c->off_keys = off;  off += al((size_t)nranks * cap * sizeof(uint64_t));
  c->off_vals = off;  off += al((size_t)nranks * cap * h.pull_stride * sizeof(float));
  c->off_push = off;  off += al((size_t)nranks * cap * h.push_stride * sizeof(float));
  c->symm_bytes = off;
  PBX_CUDA_CHECK(cudaMalloc(&c->symm_base, c->symm_bytes));
  PBX_CUDA_CHECK(cudaMemset(c->symm_base, 0, c->symm_bytes));
  PBX_CUDA_CHECK(cudaMalloc(&h.dd_keys, dd * sizeof(uint64_t)));
  PBX_CUDA_CHECK(cudaMemset(h.dd_keys, 0xFF, dd * sizeof(uint64_t)));
  PBX_CUDA_CHECK(cudaMalloc(&h.dd_uid, dd * sizeof(uint32_t)));
  PBX_CUDA_CHECK(cudaMalloc(&h.send_cnt, 64 * sizeof(uint32_t)));
  PBX_CUDA_CHECK(cudaMalloc(&h.serve_cnt, 64 * sizeof(uint32_t)));
  PBX_CUDA_CHECK(cudaMemset(h.send_cnt, 0, 64 * sizeof(uint32_t)));
  PBX_CUDA_CHECK(cudaMemset(h.serve_cnt, 0, 64 * sizeof(uint32_t)));
  PBX_CUDA_CHECK(cudaMalloc(&h.mail_row, (size_t)nranks * cap * sizeof(uint32_t)));
  PBX_CUDA_CHECK(cudaMalloc(&h.push_acc, (size_t)nranks * cap * h.push_stride * sizeof(float)));
  PBX_CUDA_CHECK(cudaMalloc(&h.dup_list, (size_t)max_keys * 2 * sizeof(uint32_t)));
  PBX_CUDA_CHECK(cudaMalloc(&h.uid_multi, (size_t)nranks * cap * sizeof(uint32_t)));
  PBX_CUDA_CHECK(cudaMemset(h.uid_multi, 0, (size_t)nranks * cap * sizeof(uint32_t)));
  PBX_CUDA_CHECK(cudaMalloc(&h.occ_rep, (size_t)max_keys + 16));
  PBX_CUDA_CHECK(cudaMemset(h.occ_rep, 0, (size_t)max_keys + 16));
  PBX_CUDA_CHECK(cudaMalloc(&h.epoch, 256));
  PBX_CUDA_CHECK(cudaMemset(h.epoch, 0, 256));
  h.status = (int*)(h.epoch + 8);
  PBX_CUDA_CHECK(cudaMallocHost(&c->status_host, 64));
  memset(c->status_host, 0, 64);
  PBX_CUDA_CHECK(cudaMalloc(&c->d, sizeof(PbxSparseDev)));
  pbx_ctx_bind_peer(c, rank, c->symm_base);
  // cooperative grid: one CTA per SM
  int bps = 0;
  PBX_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_sparse_pull<false>, PBX_A2A_THREADS, 0));
  PBX_REQUIRE(bps >= 1, "sparse pull kernel does not fit on an SM");
  PBX_CUDA_CHECK(cudaFuncSetAttribute(k_sparse_push<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      PBX_PUSH_SMEM_WIDE));
  PBX_CUDA_CHECK(cudaFuncSetAttribute(k_sparse_push<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      PBX_PUSH_SMEM_WIDE));
  // push Q3 tiling: one-line rows use 32-row tiles; wider rows use the largest power-of-two
  // tile that fits the per-warp shared-memory share; rows too wide for 2 per warp fall back
  // to the per-thread path
  {
    const int st = t->view.L.stride, pd = h.push_stride;
    if (st == 32 && pd <= 24) {
      h.q3_tile_rows = 32;
      h.q3_warp_words = 32 * 36 + 32 * 24;
    } else {
      int tr = 32;
      while (tr >= 2 && tr * (st + 4 + pd) > PBX_PUSH_WIDE_WARP_WORDS) tr >>= 1;
      h.q3_tile_rows = tr >= 2 ? tr : 0;
      h.q3_warp_words = PBX_PUSH_WIDE_WARP_WORDS;
    }
  }
  c->grid_blocks = pbx_num_sms();
  PBX_CUDA_CHECK(cudaDeviceSynchronize());
  *out = c;
  return 0;
}

int pbx_sparse_destroy(void* hnd) {
  PbxSparseCtx* c = (PbxSparseCtx*)hnd;
  if (!c) return 0;
  for (int p = 0; p < c->h.nranks; ++p)
    if (c->peer_opened[p]) cudaIpcCloseMemHandle(c->peer_base[p]);
  cudaFree(c->symm_base);
  cudaFree(c->h.dd_keys);
  cudaFree(c->h.dd_uid);
  cudaFree(c->h.send_cnt);
  cudaFree(c->h.serve_cnt);
  cudaFree(c->h.mail_row);
  cudaFree(c->h.push_acc);
  cudaFree(c->h.occ_rep);
  cudaFree(c->h.uid_multi);
  cudaFree(c->h.dup_list);
  cudaFree(c->h.epoch);
  cudaFree(c->d);
  cudaFreeHost(c->status_host);
  delete c;
  return 0;
}

// 64-byte CUDA IPC handle of this rank's symmetric region (exchange through the
// torch.distributed store, then call pbx_sparse_connect on every rank).
int pbx_sparse_ipc_handle(void* hnd, void* out64) {
  PbxSparseCtx* c = (PbxSparseCtx*)hnd;
  cudaIpcMemHandle_t mh;
  PBX_CUDA_CHECK(cudaIpcGetMemHandle(&mh, c->symm_base));
  memcpy(out64, &mh, sizeof(mh));
  return 0;
}

int pbx_sparse_connect(void* hnd, const void* handles /* nranks * 64 bytes */) {
  PbxSparseCtx* c = (PbxSparseCtx*)hnd;
  for (int p = 0; p < c->h.nranks; ++p) {
    if (p == c->h.rank) continue;
    cudaIpcMemHandle_t mh;
    memcpy(&mh, (const char*)handles + (size_t)p * sizeof(mh), sizeof(mh));
    void* base = nullptr;
    PBX_CUDA_CHECK(cudaIpcOpenMemHandle(&base, mh, cudaIpcMemLazyEnablePeerAccess));
    c->peer_opened[p] = true;
    pbx_ctx_bind_peer(c, p, base);
  }
  return 0;
}

// Same-process wiring (virtual ranks on one GPU, or several GPUs driven by one process).
int pbx_sparse_connect_local(void* hnd, void* const* peers) {
  PbxSparseCtx* c = (PbxSparseCtx*)hnd;
  for (int p = 0; p < c->h.nranks; ++p) {
    if (p == c->h.rank) continue;
    PbxSparseCtx* o = (PbxSparseCtx*)peers[p];
    PBX_REQUIRE(o->h.cap == c->h.cap && o->h.nranks == c->h.nranks, "peer context geometry differs");
    pbx_ctx_bind_peer(c, p, o->symm_base);
  }
  return 0;
}

int pbx_sparse_geometry(void* hnd, int64_t* cap, int64_t* n_unique_rows, int* pull_stride, int* push_stride) {
  PbxSparseCtx* c = (PbxSparseCtx*)hnd;
  *cap = c->h.cap;
  *n_unique_rows = (int64_t)c->h.cap * c->h.nranks;
  *pull_stride = c->h.pull_stride;
  *push_stride = c->h.push_stride;
  return 0;
}

// Device address of this rank's unique-value buffer [nranks*cap, pull_stride]; rows are
// addressed by the restore index written by pbx_sparse_pull.
float* pbx_sparse_pull_values(void* hnd) {
  PbxSparseCtx* c = (PbxSparseCtx*)hnd;
  return c->h.pull_vals[c->h.rank];
}

static int pbx_ctx_sync_dev(PbxSparseCtx* c, cudaStream_t st) {
  PbxTable* t = (PbxTable*)c->table;
  if (memcmp(&c->h.table, &t->view, sizeof(PbxTableView)) != 0) {
    c->h.table = t->view;
    c->dirty = true;
  }
  if (c->dirty) {
    PBX_CUDA_CHECK(cudaMemcpyAsync(c->d, &c->h, sizeof(PbxSparseDev), cudaMemcpyHostToDevice, st));
    PBX_CUDA_CHECK(cudaStreamSynchronize(st));  // h is pageable: make the copy complete before h changes
    c->dirty = false;
  }
  return 0;
}

// Diagnosis: per-CTA phase-end stamps.  on = 1 allocates [grid, 16] uint64 and switches the stamps on;
// out (host, [grid * 16]) receives the stamps of the last pull (slots 0..4) and push (8..11).
int pbx_sparse_debug_stamps(void* hnd, int on, uint64_t* out, int* grid_out, cudaStream_t st) {
  PbxSparseCtx* c = (PbxSparseCtx*)hnd;
  if (grid_out) *grid_out = c->grid_blocks;
  if (on && !c->h.dbg) {
    PBX_CUDA_CHECK(cudaMalloc(&c->h.dbg, (size_t)c->grid_blocks * 16 * sizeof(uint64_t)));
    PBX_CUDA_CHECK(cudaMemset(c->h.dbg, 0, (size_t)c->grid_blocks * 16 * sizeof(uint64_t)));
    c->dirty = true;
  }
  if (out && c->h.dbg) {
    PBX_CUDA_CHECK(cudaMemcpyAsync(out, c->h.dbg, (size_t)c->grid_blocks * 16 * sizeof(uint64_t),
                                   cudaMemcpyDeviceToHost, st));
    PBX_CUDA_CHECK(cudaStreamSynchronize(st));
  }
  return 0;
}

// copies the 16 phase timestamps (ns, globaltimer) of the last pull [0..5] and push [8..12]
int pbx_sparse_phase_times(void* hnd, uint64_t* out16, cudaStream_t st) {
  PbxSparseCtx* c = (PbxSparseCtx*)hnd;
  PBX_CUDA_CHECK(cudaMemcpyAsync(out16, c->h.epoch + 16, 16 * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
  PBX_CUDA_CHECK(cudaStreamSynchronize(st));
  return 0;
}

// Non-blocking health check: the status word as of the last push whose 4-byte copy has completed
// (every push enqueues one into pinned memory).  0 = healthy; see pbx_sparse_status for the codes.
int pbx_sparse_status_peek(void* hnd) {
  PbxSparseCtx* c = (PbxSparseCtx*)hnd;
  return *(volatile int*)c->status_host;
}

int pbx_sparse_status(void* hnd, int* status, int reset, cudaStream_t st) {
  PbxSparseCtx* c = (PbxSparseCtx*)hnd;
  PBX_CUDA_CHECK(cudaMemcpyAsync(status, c->h.status, sizeof(int), cudaMemcpyDeviceToHost, st));
  PBX_CUDA_CHECK(cudaStreamSynchronize(st));
  if (reset && *status) {
    PBX_CUDA_CHECK(cudaMemsetAsync(c->h.status, 0, sizeof(int), st));
    *c->status_host = 0;
  }
  return 0;
}

// d_n (optional): device int32 holding the number of keys; when given, `n` is only the upper
// bound used for validation and the kernel reads the real count from the device.
int pbx_sparse_pull_ex(void* hnd, const uint64_t* d_keys, int64_t n, const int32_t* d_n, uint32_t* d_restore,
                       cudaStream_t st);
int pbx_sparse_pull(void* hnd, const uint64_t* d_keys, int64_t n, uint32_t* d_restore, cudaStream_t st) {
  return pbx_sparse_pull_ex(hnd, d_keys, n, nullptr, d_restore, st);
}
int pbx_sparse_pull_ex(void* hnd, const uint64_t* d_keys, int64_t n, const int32_t* d_n, uint32_t* d_restore,
                       cudaStream_t st) {
  PbxSparseCtx* c = (PbxSparseCtx*)hnd;
  PBX_REQUIRE(n <= c->max_keys, "batch has %lld keys, engine was sized for %lld", (long long)n,
              (long long)c->max_keys);
  int rc = pbx_ctx_sync_dev(c, st);
  if (rc) return rc;
  PbxPullArgsN args;
  memset(&args, 0, sizeof(args));
  args.a[0].keys = d_keys;
  args.a[0].n = n;
  args.a[0].restore = d_restore;
  args.a[0].n_dev = d_n;
  const PbxSparseDev* dctx = c->d;
  int nvr = 1;
  void* kargs[] = {(void*)&dctx, (void*)&nvr, (void*)&args};
  PBX_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)k_sparse_pull<false>, dim3(c->grid_blocks),
                                             dim3(PBX_A2A_THREADS), kargs, 0, st));
  pbx_note_launch(1);
  return 0;
}

static int pbx_fill_push_args(PbxPushArgs& a, const uint32_t* d_restore, int64_t n, const float* d_dout,
                              const float* d_cvm, const int32_t* d_lod, int B, int S, int Eo, int cvm_offset,
                              int dim_off, const float* d_gkeys, const int32_t* d_key2slot, int hidden, int mode,
                              float grad_scale) {
  a.restore = d_restore; a.n = n; a.dout = d_dout; a.cvm = d_cvm; a.lod = d_lod;
  a.B = B; a.S = S; a.Eo = Eo; a.cvm_offset = cvm_offset; a.dim_off = dim_off; a.ld_dout = S * Eo;
  a.gkeys = d_gkeys; a.key2slot = d_key2slot; a.hidden = hidden; a.mode = mode; a.grad_scale = grad_scale;
  return 0;
}

// mode 0: d_dout [B, S*Eo] pooled-output grads, d_cvm [B, cvm_offset], d_lod [S*B+1]
// mode 1: d_gkeys [n, hidden] per-occurrence grads, d_key2slot [n]
int pbx_sparse_push_ex(void* hnd, const uint32_t* d_restore, int64_t n, const float* d_dout, int ld_dout,
                       const float* d_cvm, const int32_t* d_lod, int B, int S, int Eo, int cvm_offset, int dim_off,
                       const float* d_gkeys, const int32_t* d_key2slot, int hidden, int mode, float grad_scale,
                       cudaStream_t st);
int pbx_sparse_push_dense(void* hnd, const uint32_t* d_restore, int64_t n, const float* d_dout, int ld_dout,
                          const float* d_cvm, const int32_t* d_lod, int B, int S, int Eo, int cvm_offset, int dim_off,
                          const float* d_gkeys, const int32_t* d_key2slot, int hidden, int mode, float grad_scale,
                          void* symm, int64_t dense_off, int64_t dense_n, float dense_scale, cudaStream_t st);
int pbx_sparse_push(void* hnd, const uint32_t* d_restore, int64_t n, const float* d_dout, const float* d_cvm,
                    const int32_t* d_lod, int B, int S, int Eo, int cvm_offset, int dim_off, const float* d_gkeys,
                    const int32_t* d_key2slot, int hidden, int mode, float grad_scale, cudaStream_t st) {
  return pbx_sparse_push_ex(hnd, d_restore, n, d_dout, S * Eo, d_cvm, d_lod, B, S, Eo, cvm_offset, dim_off, d_gkeys,
                            d_key2slot, hidden, mode, grad_scale, st);
}
// ld_dout: words between rows of d_dout (lets the pooled gradient live inside a wider tensor,
// e.g. the gradient of the concatenated tower input)
int pbx_sparse_push_ex(void* hnd, const uint32_t* d_restore, int64_t n, const float* d_dout, int ld_dout,
                       const float* d_cvm, const int32_t* d_lod, int B, int S, int Eo, int cvm_offset, int dim_off,
                       const float* d_gkeys, const int32_t* d_key2slot, int hidden, int mode, float grad_scale,
                       cudaStream_t st) {
  return pbx_sparse_push_dense(hnd, d_restore, n, d_dout, ld_dout, d_cvm, d_lod, B, S, Eo, cvm_offset, dim_off, d_gkeys,
                               d_key2slot, hidden, mode, grad_scale, nullptr, 0, 0, 1.f, st);
}

// The same with the data-parallel dense-gradient allreduce folded in: `symm` is a pbx_symm handle
// whose region holds the flat gradient buffer at [dense_off, dense_off + dense_n) floats on every
// rank (same group, same rank numbering as the sparse engine); on return of the kernel that range
// holds sum over ranks * dense_scale on every rank.  symm == NULL: plain push.
int pbx_sparse_push_dense(void* hnd, const uint32_t* d_restore, int64_t n, const float* d_dout, int ld_dout,
                          const float* d_cvm, const int32_t* d_lod, int B, int S, int Eo, int cvm_offset, int dim_off,
                          const float* d_gkeys, const int32_t* d_key2slot, int hidden, int mode, float grad_scale,
                          void* symm, int64_t dense_off, int64_t dense_n, float dense_scale, cudaStream_t st) {
  PbxSparseCtx* c = (PbxSparseCtx*)hnd;
  if (mode == 0) PBX_REQUIRE(ld_dout >= S * Eo, "ld_dout %d < S*Eo %d", ld_dout, S * Eo);
  PBX_REQUIRE(c->h.table.L.embedx_dim <= 256, "embedx_dim > 256 not supported by the push kernel");
  if (mode == 0)
    PBX_REQUIRE(Eo + dim_off == 3 + c->h.table.L.embedx_dim, "pooled width %d + %d != hidden %d", Eo, dim_off,
                3 + c->h.table.L.embedx_dim);
  else
    PBX_REQUIRE(hidden == 3 + c->h.table.L.embedx_dim, "grad width %d != hidden", hidden);
  int rc = pbx_ctx_sync_dev(c, st);
  if (rc) return rc;
  PbxPushArgsN args;
  memset(&args, 0, sizeof(args));
  pbx_fill_push_args(args.a[0], d_restore, n, d_dout, d_cvm, d_lod, B, S, Eo, cvm_offset, dim_off, d_gkeys,
                     d_key2slot, hidden, mode, grad_scale);
  if (mode == 0) args.a[0].ld_dout = ld_dout;
  if (symm && c->h.nranks > 1) {
    PbxSymm* sm = (PbxSymm*)symm;
    PBX_REQUIRE(sm->h.nranks == c->h.nranks && sm->h.rank == c->h.rank, "push: symm group differs from the sparse group");
    PBX_REQUIRE((dense_off & 3) == 0 && (uint64_t)(dense_off + dense_n) * 4 <= sm->bytes - sm->data_off,
                "push: dense range outside the symmetric region");
    if (sm->dirty) {
      PBX_CUDA_CHECK(cudaMemcpyAsync(sm->d, &sm->h, sizeof(PbxSymmDev), cudaMemcpyHostToDevice, st));
      PBX_CUDA_CHECK(cudaStreamSynchronize(st));
      sm->dirty = false;
    }
    args.a[0].dense_ctx = sm->d;
    args.a[0].dense_off = dense_off;
    args.a[0].dense_n = dense_n;
    args.a[0].dense_scale = dense_scale;
  }
  const PbxSparseDev* dctx = c->d;
  const int PDh = c->h.push_stride;
  if (mode == 0 && cvm_offset == 2 && c->h.q3_tile_rows > 0 && 16 * PDh <= c->h.q3_warp_words &&
      (int64_t)S * B < 0x7fffffffll && S * B > 0) {
    const size_t smem = (size_t)(256 / 32) * 16 * PDh * sizeof(float);
    if (smem <= 48 * 1024) {
      const int64_t ntask = ((int64_t)S * B + 15) / 16;
      int grid = (int)((ntask + 7) / 8);
      const int cap_grid = pbx_num_sms() * 8;
      if (grid > cap_grid) grid = cap_grid;
      k_push_records<<<grid, 256, smem, st>>>(dctx, args.a[0]);
      PBX_CUDA_CHECK(cudaGetLastError());
      pbx_note_launch(1);
      args.a[0].q0_done = 1;
    }
  }
  int nvr = 1;
  void* kargs[] = {(void*)&dctx, (void*)&nvr, (void*)&args};
  PBX_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)k_sparse_push<false>, dim3(c->grid_blocks),
                                             dim3(PBX_PUSH_THREADS), kargs, pbx_push_smem(c->h), st));
  pbx_note_launch(1);
  // status mirror for pbx_sparse_status_peek (a memcpy node when the step is captured in a graph)
  PBX_CUDA_CHECK(cudaMemcpyAsync(c->status_host, c->h.status, sizeof(int), cudaMemcpyDeviceToHost, st));
  return 0;
}

// ---- virtual-rank emulation: all ranks of a same-process group in ONE launch on one GPU
int pbx_sparse_pull_emu(void* const* hnds, int nranks, const uint64_t* const* d_keys, const int64_t* n,
                        uint32_t* const* d_restore, cudaStream_t st) {
  PBX_REQUIRE(nranks >= 1 && nranks <= PBX_MAX_RANKS, "bad nranks");
  std::vector<PbxSparseDev> hs(nranks);
  PbxPullArgsN args;
  memset(&args, 0, sizeof(args));
  for (int r = 0; r < nranks; ++r) {
    PbxSparseCtx* c = (PbxSparseCtx*)hnds[r];
    PBX_REQUIRE(n[r] <= c->max_keys, "batch too large for rank %d", r);
    c->h.table = ((PbxTable*)c->table)->view;
    hs[r] = c->h;
    args.a[r].keys = d_keys[r];
    args.a[r].n = n[r];
    args.a[r].restore = d_restore[r];
  }
  PbxSparseDev* d = nullptr;
  PBX_CUDA_CHECK(cudaMalloc(&d, sizeof(PbxSparseDev) * nranks));
  PBX_CUDA_CHECK(cudaMemcpy(d, hs.data(), sizeof(PbxSparseDev) * nranks, cudaMemcpyHostToDevice));
  int bpr = pbx_num_sms() / nranks;
  if (bpr < 1) bpr = 1;
  const PbxSparseDev* dctx = d;
  void* kargs[] = {(void*)&dctx, (void*)&nranks, (void*)&args};
  cudaError_t e = cudaLaunchCooperativeKernel((void*)k_sparse_pull<true>, dim3(bpr * nranks), dim3(PBX_A2A_THREADS),
                                              kargs, 0, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFree(d);
  PBX_CUDA_CHECK(e);
  return 0;
}

int pbx_sparse_push_emu(void* const* hnds, int nranks, const uint32_t* const* d_restore, const int64_t* n,
                        const float* const* d_dout, const float* const* d_cvm, const int32_t* const* d_lod, int B,
                        int S, int Eo, int cvm_offset, int dim_off, float grad_scale, cudaStream_t st) {
  PBX_REQUIRE(nranks >= 1 && nranks <= PBX_MAX_RANKS, "bad nranks");
  std::vector<PbxSparseDev> hs(nranks);
  PbxPushArgsN args;
  memset(&args, 0, sizeof(args));
  for (int r = 0; r < nranks; ++r) {
    PbxSparseCtx* c = (PbxSparseCtx*)hnds[r];
    c->h.table = ((PbxTable*)c->table)->view;
    hs[r] = c->h;
    pbx_fill_push_args(args.a[r], d_restore[r], n[r], d_dout[r], d_cvm[r], d_lod[r], B, S, Eo, cvm_offset, dim_off,
                       nullptr, nullptr, 0, 0, grad_scale);
  }
  PbxSparseDev* d = nullptr;
  PBX_CUDA_CHECK(cudaMalloc(&d, sizeof(PbxSparseDev) * nranks));
  PBX_CUDA_CHECK(cudaMemcpy(d, hs.data(), sizeof(PbxSparseDev) * nranks, cudaMemcpyHostToDevice));
  int bpr = pbx_num_sms() / nranks;
  if (bpr < 1) bpr = 1;
  const PbxSparseDev* dctx = d;
  void* kargs[] = {(void*)&dctx, (void*)&nranks, (void*)&args};
  cudaError_t e = cudaLaunchCooperativeKernel((void*)k_sparse_push<true>, dim3(bpr * nranks), dim3(PBX_PUSH_THREADS),
                                              kargs, pbx_push_smem(hs[0]), st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFree(d);
  PBX_CUDA_CHECK(e);
  return 0;
}

}  // extern "C"
