// Device-side view of one HBM shard of the sparse table plus the sparse optimizer
// functors.  Open addressing (linear probing) over a key array; the value row
// of a key lives at the same index as its key, so a hit costs one key probe
// plus one contiguous row read and no pointer chase.
//
// Behavioural oracle: reference HashTable (heter_ps/hashtable_kernel.cu:34-140)
// and optimizer functors (heter_ps/optimizer.cuh.h:30-472).
#pragma once
#include "common.cuh"
#include "../../include/pbx_types.h"

struct PbxTableView {
  uint64_t* keys;     // [capacity], PBX_EMPTY_KEY when free
  float* vals;        // [capacity * stride]
  uint8_t* dirty;     // [capacity] touched since last delta save
  uint64_t capacity;
  uint64_t seed;
  PbxRowLayout L;
  PbxSparseOptConfig cfg;
  int test_mode;      // 1: pushes are ignored (reference SetTestMode)
};

__device__ __forceinline__ uint32_t pbx_dev_find(const PbxTableView& t, uint64_t key) {
  uint64_t s = pbx_home_slot(key, t.capacity);
  for (uint64_t probe = 0; probe < t.capacity; ++probe) {
    uint64_t k = __ldg(t.keys + s);   // keys are immutable during pull/push phases
    if (k == key) return (uint32_t)s;
    if (k == PBX_EMPTY_KEY) return PBX_NO_ROW;
    if (++s == t.capacity) s = 0;
  }
  return PBX_NO_ROW;
}

// returns slot; *fresh = true when this thread created the entry
__device__ __forceinline__ uint32_t pbx_dev_insert(const PbxTableView& t, uint64_t key, bool* fresh) {
  uint64_t s = pbx_home_slot(key, t.capacity);
  *fresh = false;
  for (uint64_t probe = 0; probe < t.capacity; ++probe) {
    unsigned long long prev =
        atomicCAS(reinterpret_cast<unsigned long long*>(t.keys + s), (unsigned long long)PBX_EMPTY_KEY,
                  (unsigned long long)key);
    if (prev == PBX_EMPTY_KEY) { *fresh = true; return (uint32_t)s; }
    if (prev == key) return (uint32_t)s;
    if (++s == t.capacity) s = 0;
  }
  return PBX_NO_ROW;
}

__device__ __forceinline__ void pbx_row_init(const PbxTableView& t, float* row, uint64_t key) {
  const PbxRowLayout& L = t.L;
  for (int i = 0; i < L.stride; ++i) row[i] = 0.f;
  // embed_w ~ U(0,1) * initial_range (0 by default, optimizer_conf.h:27)
  row[2] = t.cfg.initial_range != 0.f ? pbx_uniform01(key, 0xffffu, t.seed) * t.cfg.initial_range : 0.f;
  if (L.optimizer != PBX_OPT_ADAGRAD) {
    row[L.off_es + 2] = t.cfg.beta1_decay_rate;
    row[L.off_es + 3] = t.cfg.beta2_decay_rate;
  }
}

// ---- optimizer functors -----------------------------------------------------
// g is already multiplied by -batch_size by the push merge (box_wrapper.cu:404-410);
// the update divides by the pushed show count and *adds* (optimizer.cuh.h:60-62).
__device__ __forceinline__ void pbx_adagrad(float* w, float* g2sum, const float* g, int n, float scale,
                                            float lr, float g2_0, float lo, float hi) {
  float ratio = lr * sqrtf(g2_0 / (g2_0 + *g2sum));
  float add = 0.f;
  float inv = 1.f / scale;
  for (int i = 0; i < n; ++i) {
    float sg = g[i] * inv;
    float v = w[i] + sg * ratio;
    v = fminf(fmaxf(v, lo), hi);
    w[i] = v;
    add += sg * sg;
  }
  *g2sum += add / (float)n;
}

__device__ __forceinline__ void pbx_adam(float* w, float* m, float* v, float* b1p, float* b2p,
                                         const float* g, int n, float scale, float lr, float b1, float b2,
                                         float eps, float lo, float hi, bool shared) {
  float b1pow = *b1p, b2pow = *b2p;
  float ratio = lr * sqrtf(1.f - b2pow) / (1.f - b1pow);
  float inv = 1.f / scale;
  if (!shared) {
    for (int i = 0; i < n; ++i) {
      float sg = g[i] * inv;
      float nm = b1 * m[i] + (1.f - b1) * sg;
      float nv = b2 * v[i] + (1.f - b2) * sg * sg;
      float x = w[i] + ratio * (nm / (sqrtf(nv) + eps));
      w[i] = fminf(fmaxf(x, lo), hi);
      m[i] = nm;
      v[i] = nv;
    }
  } else {
    // shared moments: one (m, v) pair for all n dims, fed by the mean gradient
    // (optimizer.cuh.h:329-472)
    float sum_m = 0.f, sum_v = 0.f;
    float m0 = *m, v0 = *v;
    for (int i = 0; i < n; ++i) {
      float sg = g[i] * inv;
      float nm = b1 * m0 + (1.f - b1) * sg;
      float nv = b2 * v0 + (1.f - b2) * sg * sg;
      float x = w[i] + ratio * (nm / (sqrtf(nv) + eps));
      w[i] = fminf(fmaxf(x, lo), hi);
      sum_m += nm;
      sum_v += nv;
    }
    *m = sum_m / (float)n;
    *v = sum_v / (float)n;
  }
  *b1p = b1pow * b1;
  *b2p = b2pow * b2;
}

// Apply one merged push record to one row held at `row` (global or shared memory).
// push = [slot, show, click, embed_g, embedx_g[D]].  Returns false when the record is empty.
// The key (needed only for lazy embedx creation) is read from the table on demand.
__device__ __forceinline__ bool pbx_apply_push_ptr(const PbxTableView& t, float* row, uint32_t slot_idx,
                                                   const float* push) {
  const PbxRowLayout& L = t.L;
  const PbxSparseOptConfig& c = t.cfg;
  float g_show = push[1], g_click = push[2];
  if (g_show <= 0.f) return false;  // nothing merged into this record
  row[L.off_tail + 1] = push[0];  // slot
  float show = row[0] + g_show;
  float click = row[1] + g_click;
  row[0] = show;
  row[1] = click;
  row[L.off_tail + 0] += c.nonclk_coeff * (g_show - g_click) + c.clk_coeff * g_click;
  row[L.off_tail + 2] = 0.f;  // seen in this pass
  float* es = row + L.off_es;
  float* xs = row + L.off_xs;
  const int D = L.embedx_dim;
  if (L.optimizer == PBX_OPT_ADAGRAD) {
    pbx_adagrad(row + 2, es, push + 3, 1, g_show, c.learning_rate, c.initial_g2sum, c.min_bound, c.max_bound);
  } else {
    pbx_adam(row + 2, es, es + 1, es + 2, es + 3, push + 3, 1, g_show, c.learning_rate, c.beta1_decay_rate,
             c.beta2_decay_rate, c.ada_epsilon, c.min_bound, c.max_bound, false);
  }
  if (row[L.off_mf] == 0.f) {
    // lazy embedx creation (optimizer.cuh.h:104-122)
    if (c.mf_create_thresholds <= c.nonclk_coeff * (show - click) + c.clk_coeff * click) {
      const uint64_t key = t.keys[slot_idx];
      row[L.off_mf] = (float)D;
      for (int i = 0; i < D; ++i) row[3 + i] = pbx_uniform01(key, (uint32_t)i, t.seed) * c.mf_initial_range;
      if (L.optimizer == PBX_OPT_ADAM) {
        xs[2 * D] = c.mf_beta1_decay_rate;
        xs[2 * D + 1] = c.mf_beta2_decay_rate;
      } else if (L.optimizer == PBX_OPT_SHARED_ADAM) {
        xs[2] = c.mf_beta1_decay_rate;
        xs[3] = c.mf_beta2_decay_rate;
      }
    }
  } else {
    if (L.optimizer == PBX_OPT_ADAGRAD) {
      pbx_adagrad(row + 3, xs, push + 4, D, g_show, c.mf_learning_rate, c.mf_initial_g2sum, c.mf_min_bound,
                  c.mf_max_bound);
    } else if (L.optimizer == PBX_OPT_ADAM) {
      pbx_adam(row + 3, xs, xs + D, xs + 2 * D, xs + 2 * D + 1, push + 4, D, g_show, c.mf_learning_rate,
               c.mf_beta1_decay_rate, c.mf_beta2_decay_rate, c.mf_ada_epsilon, c.mf_min_bound, c.mf_max_bound,
               false);
    } else {
      pbx_adam(row + 3, xs, xs + 1, xs + 2, xs + 3, push + 4, D, g_show, c.mf_learning_rate,
               c.mf_beta1_decay_rate, c.mf_beta2_decay_rate, c.mf_ada_epsilon, c.mf_min_bound, c.mf_max_bound,
               true);
    }
  }
  return true;
}

// in-place update of the table row `slot_idx` (generic path; `key` kept for API stability)
__device__ __forceinline__ void pbx_apply_push(const PbxTableView& t, uint32_t slot_idx, uint64_t key,
                                               const float* push) {
  (void)key;
  if (pbx_apply_push_ptr(t, t.vals + (uint64_t)slot_idx * t.L.stride, slot_idx, push) && t.dirty)
    t.dirty[slot_idx] = 1;
}

// Streaming Adagrad update for embedx widths that are multiples of 4: the row and the push
// record move as 16-byte vectors, but only a sliding window (one row chunk, two push chunks)
// plus the three state-carrying chunks live in registers, so the kernels that call it keep a
// small register footprint (-> 1024 threads per SM for latency hiding).  Same arithmetic and
// accumulation order as pbx_apply_push (Adagrad branch).
//
// Word map (D = embedx dim): row  [show clk w x0 | x1..x4 | ... | x(D-3) x(D-2) x(D-1) mf |
//                                  delta slot unseen g2 | xg2 pad pad pad]
//                            push [slot gshow gclk gw | g0..g3 | g4..g7 | ...]
// so embedx element e sits at row word 3+e and its gradient at push word 4+e: row chunk k
// needs push chunk k lanes 1..3 and push chunk k+1 lane 0.
__device__ __forceinline__ void pbx_apply_push_adagrad_stream(const PbxTableView& t, uint32_t slot_idx,
                                                              const float* __restrict__ push) {
  const int D = t.L.embedx_dim;
  const int q = D >> 2;                       // chunk holding x(D-3..D-1), mf
  const int S4 = t.L.stride >> 2;
  const float4* pv = reinterpret_cast<const float4*>(push);
  const float4 p0 = __ldcg(pv);               // slot, g_show, g_click, embed_g
  const float g_show = p0.y, g_click = p0.z;
  if (g_show <= 0.f) return;
  float4* rowv = reinterpret_cast<float4*>(t.vals + (uint64_t)slot_idx * (S4 * 4));
  float4 r0 = rowv[0];
  float4 rA = rowv[q];                        // x(D-3), x(D-2), x(D-1), mf_size
  float4 rB = rowv[q + 1];                    // delta_score, slot, unseen_days, embed g2sum
  float4 rC = rowv[q + 2];                    // embedx g2sum, pad
  const PbxSparseOptConfig& c = t.cfg;
  const float show = r0.x + g_show, click = r0.y + g_click;
  r0.x = show;
  r0.y = click;
  rB.x += c.nonclk_coeff * (g_show - g_click) + c.clk_coeff * g_click;
  rB.y = p0.x;
  rB.z = 0.f;
  const float inv = 1.f / g_show;
  {
    const float ratio = c.learning_rate * sqrtf(c.initial_g2sum / (c.initial_g2sum + rB.w));
    const float sg = p0.w * inv;
    r0.z = fminf(fmaxf(r0.z + sg * ratio, c.min_bound), c.max_bound);
    rB.w += sg * sg;
  }
  if (rA.w == 0.f) {
    if (c.mf_create_thresholds <= c.nonclk_coeff * (show - click) + c.clk_coeff * click) {
      // lazy embedx creation; the key is only needed here (counter-based RNG)
      const uint64_t key = t.keys[slot_idx];
      const float rg = c.mf_initial_range;
      rA.w = (float)D;
      r0.w = pbx_uniform01(key, 0u, t.seed) * rg;
      for (int k = 1; k < q; ++k)
        rowv[k] = make_float4(pbx_uniform01(key, 4 * k - 3, t.seed) * rg, pbx_uniform01(key, 4 * k - 2, t.seed) * rg,
                              pbx_uniform01(key, 4 * k - 1, t.seed) * rg, pbx_uniform01(key, 4 * k, t.seed) * rg);
      rA.x = pbx_uniform01(key, D - 3, t.seed) * rg;
      rA.y = pbx_uniform01(key, D - 2, t.seed) * rg;
      rA.z = pbx_uniform01(key, D - 1, t.seed) * rg;
    }
  } else {
    const float ratio = c.mf_learning_rate * sqrtf(c.mf_initial_g2sum / (c.mf_initial_g2sum + rC.x));
    const float lo = c.mf_min_bound, hi = c.mf_max_bound;
    float add = 0.f;
    float4 pk = __ldcg(pv + 1);               // g0..g3
    float sg = pk.x * inv;
    r0.w = fminf(fmaxf(r0.w + sg * ratio, lo), hi);
    add += sg * sg;
    for (int k = 1; k < q; ++k) {
      const float4 pn = __ldcg(pv + 1 + k);
      float4 rk = rowv[k];
      sg = pk.y * inv; rk.x = fminf(fmaxf(rk.x + sg * ratio, lo), hi); add += sg * sg;
      sg = pk.z * inv; rk.y = fminf(fmaxf(rk.y + sg * ratio, lo), hi); add += sg * sg;
      sg = pk.w * inv; rk.z = fminf(fmaxf(rk.z + sg * ratio, lo), hi); add += sg * sg;
      sg = pn.x * inv; rk.w = fminf(fmaxf(rk.w + sg * ratio, lo), hi); add += sg * sg;
      rowv[k] = rk;
      pk = pn;
    }
    sg = pk.y * inv; rA.x = fminf(fmaxf(rA.x + sg * ratio, lo), hi); add += sg * sg;
    sg = pk.z * inv; rA.y = fminf(fmaxf(rA.y + sg * ratio, lo), hi); add += sg * sg;
    sg = pk.w * inv; rA.z = fminf(fmaxf(rA.z + sg * ratio, lo), hi); add += sg * sg;
    rC.x += add / (float)D;
  }
  rowv[0] = r0;
  rowv[q] = rA;
  rowv[q + 1] = rB;
  rowv[q + 2] = rC;
  if (t.dirty) t.dirty[slot_idx] = 1;
}

// dispatch: streaming vector path for Adagrad with D % 4 == 0 (D >= 4), generic scalar otherwise
__device__ __forceinline__ void pbx_apply_push_fast(const PbxTableView& t, uint32_t slot_idx, const float* push) {
  if (t.L.optimizer == PBX_OPT_ADAGRAD && t.L.embedx_dim >= 4 && (t.L.embedx_dim & 3) == 0) {
    pbx_apply_push_adagrad_stream(t, slot_idx, push);
    return;
  }
  pbx_apply_push(t, slot_idx, t.keys[slot_idx], push);
}

// Host-side table object (opaque to callers; see include/pbx_cuda.h)
struct PbxTable {
  PbxTableView view;
  int device;
  unsigned long long* d_counter;  // scratch counter for count/dump
};
