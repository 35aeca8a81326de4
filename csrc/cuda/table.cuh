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

// Apply one merged push record to one row.  push = [slot, show, click, embed_g, embedx_g[D]].
__device__ __forceinline__ void pbx_apply_push(const PbxTableView& t, uint32_t slot_idx, uint64_t key,
                                               const float* push) {
  const PbxRowLayout& L = t.L;
  const PbxSparseOptConfig& c = t.cfg;
  float* row = t.vals + (uint64_t)slot_idx * L.stride;
  float g_show = push[1], g_click = push[2];
  if (g_show <= 0.f) return;  // nothing merged into this record
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
  if (t.dirty) t.dirty[slot_idx] = 1;
}

// Register-resident Adagrad update for a compile-time embedx width: the whole row (stride
// = ceil4(9 + D) words) and the push record are moved with 16-byte loads/stores and the math
// runs on registers.  Same arithmetic and order as pbx_apply_push (Adagrad branch).
template <int D>
__device__ __forceinline__ void pbx_apply_push_adagrad_vec(const PbxTableView& t, uint32_t slot_idx, uint64_t key,
                                                           const float* __restrict__ push) {
  constexpr int S4 = (9 + D + 3) / 4;
  constexpr int P4 = (4 + D + 3) / 4;
  constexpr int OFF_MF = 3 + D, OFF_T = 4 + D, OFF_ES = 7 + D, OFF_XS = 8 + D;
  float p[P4 * 4];
#pragma unroll
  for (int i = 0; i < P4; ++i) {
    float4 v = __ldcg(reinterpret_cast<const float4*>(push) + i);
    p[4 * i] = v.x; p[4 * i + 1] = v.y; p[4 * i + 2] = v.z; p[4 * i + 3] = v.w;
  }
  const float g_show = p[1], g_click = p[2];
  if (g_show <= 0.f) return;
  float4* rowv = reinterpret_cast<float4*>(t.vals + (uint64_t)slot_idx * (S4 * 4));
  float r[S4 * 4];
#pragma unroll
  for (int i = 0; i < S4; ++i) {
    float4 v = rowv[i];
    r[4 * i] = v.x; r[4 * i + 1] = v.y; r[4 * i + 2] = v.z; r[4 * i + 3] = v.w;
  }
  const PbxSparseOptConfig& c = t.cfg;
  r[OFF_T + 1] = p[0];
  const float show = r[0] + g_show, click = r[1] + g_click;
  r[0] = show;
  r[1] = click;
  r[OFF_T + 0] += c.nonclk_coeff * (g_show - g_click) + c.clk_coeff * g_click;
  r[OFF_T + 2] = 0.f;
  const float inv = 1.f / g_show;
  {  // embed_w
    float ratio = c.learning_rate * sqrtf(c.initial_g2sum / (c.initial_g2sum + r[OFF_ES]));
    float sg = p[3] * inv;
    r[2] = fminf(fmaxf(r[2] + sg * ratio, c.min_bound), c.max_bound);
    r[OFF_ES] += sg * sg;
  }
  if (r[OFF_MF] == 0.f) {
    if (c.mf_create_thresholds <= c.nonclk_coeff * (show - click) + c.clk_coeff * click) {
      r[OFF_MF] = (float)D;
#pragma unroll
      for (int i = 0; i < D; ++i) r[3 + i] = pbx_uniform01(key, (uint32_t)i, t.seed) * c.mf_initial_range;
    }
  } else {
    float ratio = c.mf_learning_rate * sqrtf(c.mf_initial_g2sum / (c.mf_initial_g2sum + r[OFF_XS]));
    float add = 0.f;
#pragma unroll
    for (int i = 0; i < D; ++i) {
      float sg = p[4 + i] * inv;
      r[3 + i] = fminf(fmaxf(r[3 + i] + sg * ratio, c.mf_min_bound), c.mf_max_bound);
      add += sg * sg;
    }
    r[OFF_XS] += add / (float)D;
  }
#pragma unroll
  for (int i = 0; i < S4; ++i) rowv[i] = make_float4(r[4 * i], r[4 * i + 1], r[4 * i + 2], r[4 * i + 3]);
  if (t.dirty) t.dirty[slot_idx] = 1;
}

// dispatch: vectorised Adagrad for the common widths, generic scalar path otherwise
__device__ __forceinline__ void pbx_apply_push_fast(const PbxTableView& t, uint32_t slot_idx, uint64_t key,
                                                    const float* push) {
  if (t.L.optimizer == PBX_OPT_ADAGRAD) {
    switch (t.L.embedx_dim) {
      case 8: pbx_apply_push_adagrad_vec<8>(t, slot_idx, key, push); return;
      case 16: pbx_apply_push_adagrad_vec<16>(t, slot_idx, key, push); return;
      case 32: pbx_apply_push_adagrad_vec<32>(t, slot_idx, key, push); return;
      default: break;   // wider rows would spill under the 128-register budget of 512-thread CTAs
    }
  }
  pbx_apply_push(t, slot_idx, key, push);
}

// Host-side table object (opaque to callers; see include/pbx_cuda.h)
struct PbxTable {
  PbxTableView view;
  int device;
  unsigned long long* d_counter;  // scratch counter for count/dump
};
