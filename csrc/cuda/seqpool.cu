// fused_seqpool_cvm forward / backward as ONE kernel each (W5 of SURVEY.md §8).
// The reference runs pool and CVM as two kernels through an [S*B, E] intermediate plus
// S small LoD memcpys per call (fused/fused_seqpool_cvm_op.cu:35-520, :814-1151); here
// pooling, score filter, embed-norm filter, quantisation, embedx-concat and the CVM
// transform happen in a single pass, and the input rows may be read *through the pull's
// restore index* so the per-occurrence [N, hidden] tensor never has to exist.
//
// Layouts: bags are slot-major, bag g = slot * B + ins, keys of bag g are
// [lod[g], lod[g+1]).  Output is [B, S * Eo] (instance-major), i.e. already the
// concatenation over slots that every CTR tower consumes; slot s is the column block
// [s*Eo, (s+1)*Eo).
#include "common.cuh"

#include "../../include/pbx_types.h"

namespace {

__device__ __forceinline__ float quantize(float v, int q) {
  // reference: static_cast<int>(v * q + 0.5) / static_cast<float>(q)
  return (float)((int)(v * (float)q + 0.5f)) / (float)q;
}

__device__ __forceinline__ bool keep_key(const PbxSeqpoolParams& p, const float* row, float thr) {
  if (p.need_filter) {
    float show = row[0], clk = row[1];
    if ((show - clk) * p.show_coeff + clk * p.clk_coeff < thr) return false;
  }
  if (p.embed_threshold_filter) {
    const float* e = row + p.cvm_offset;
    float sc = 0.f;
    for (int i = 1; i < p.embed_thres_size; ++i) sc += e[i] * e[i];
    sc = sqrtf(sc) + fabsf(e[0]);
    if (sc < p.embed_threshold) return false;
  }
  return true;
}

// one thread per (bag, concat block, source column)
// optional tail: dense features copied into columns [off, off + cols) of the same output rows
// (the tower input is [pooled slots | dense features] in one buffer: no separate copy launch)
struct SeqpoolTail { const float* src; int ld, cols, off; };
__device__ __forceinline__ void seqpool_copy_tail(const SeqpoolTail& t, int B, float* __restrict__ out,
                                                  int out_stride) {
  if (!t.src) return;
  const int64_t total = (int64_t)B * t.cols;
  for (int64_t it = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; it < total;
       it += (int64_t)gridDim.x * blockDim.x) {
    const int b = (int)(it / t.cols), c = (int)(it % t.cols);
    out[(int64_t)b * out_stride + t.off + c] = __ldg(t.src + (int64_t)b * t.ld + c);
  }
}

__global__ void k_seqpool_cvm_fwd(PbxSeqpoolParams p, const float* __restrict__ x,
                                  const uint32_t* __restrict__ restore, const int32_t* __restrict__ lod,
                                  float* __restrict__ out, SeqpoolTail tail) {
  seqpool_copy_tail(tail, p.B, out, p.out_stride);
  const int H = p.hidden;
  const int64_t per_bag = (int64_t)p.concat * H;
  const int64_t total = (int64_t)p.S * p.B * per_bag;
  // width of one concat block in the output
  const int blockw = p.Eo / p.concat;
  for (int64_t it = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; it < total;
       it += (int64_t)gridDim.x * blockDim.x) {
    const int col = (int)(it % H);
    const int ci = (int)((it / H) % p.concat);
    const int64_t g = it / per_bag;
    const int s = (int)(g / p.B);
    const int b = (int)(g % p.B);
    // which output column does this source column land in (or none)
    int oc;
    if (p.use_cvm) {
      if (p.variant == 0) {
        if (p.clk_filter) { if (col == 1) continue; oc = col == 0 ? 0 : col - 1; }
        else oc = col;
      } else {  // conv / credit: optional show filter drops column 0
        if (p.show_filter) { if (col == 0) continue; oc = col - 1; }
        else oc = col;
      }
    } else {
      if (col < p.cvm_offset) continue;
      oc = col - p.cvm_offset;
    }
    int32_t k0 = lod[g], k1 = lod[g + 1];
    bool filtering = p.need_filter || p.embed_threshold_filter;
    if (p.concat > 1) {
      // block ci holds the ci-th key of the bag, un-pooled (fused_seqpool_cvm_op.cu:116-123)
      k0 = k0 + ci;
      k1 = min(k0 + 1, k1);
      if (p.embed_threshold_filter && !p.concat_filter) filtering = false;  // (:352-354)
    }
    const float thr = p.slot_threshold ? p.slot_threshold[s] : p.threshold;
    float val = p.pad_value;
    float show_sum = p.pad_value, aux_sum = p.pad_value;  // for the ratio columns
    for (int32_t k = k0; k < k1; ++k) {
      const float* row = x + (uint64_t)(restore ? restore[k] : (uint32_t)k) * p.x_stride;
      if (filtering && !keep_key(p, row, thr)) continue;
      float v = row[col];
      if (col >= p.cvm_offset) {
        if (p.trade_w) v *= p.trade_w[k];
        if (p.quant_ratio > 0) v = quantize(v, p.quant_ratio);
      }
      val += v;
      if (p.use_cvm && col >= 1 && col < p.cvm_offset) {
        show_sum += row[0];
        if (col >= 2) aux_sum += row[col - 1];
      }
    }
    if (p.use_cvm && col < p.cvm_offset) {
      if (p.variant == 0) {
        // [log(show+1), log(clk+1) - log(show+1)]  (fused_seqpool_cvm_op.cu:385-391)
        val = col == 0 ? logf(val + 1.f) : (col == 1 ? logf(val + 1.f) - logf(show_sum + 1.f) : val);
      } else if (p.variant == 1) {
        // conv: [log(show+1), log(clk+1), log(conv+1) - log(clk+1)]
        val = col == 2 ? logf(val + 1.f) - logf(aux_sum + 1.f) : logf(val + 1.f);
      } else {
        val = logf(val + 1.f);  // credit: every cvm column is log(x+1)
      }
    }
    out[(int64_t)b * p.out_stride + (int64_t)s * p.Eo + (int64_t)ci * blockw + oc] = val;
  }
}

// Fast path for the common configuration (base variant, cvm_offset 2, no filter / quant /
// concat, 16-byte-aligned value rows): one thread per (bag, 4-column chunk), 128-bit gathers
// of the value rows, CVM transform applied in the chunk that holds show/clk.
__global__ void k_seqpool_cvm_fwd_fast(PbxSeqpoolParams p, const float* __restrict__ x,
                                       const uint32_t* __restrict__ restore, const int32_t* __restrict__ lod,
                                       float* __restrict__ out, SeqpoolTail tail) {
  seqpool_copy_tail(tail, p.B, out, p.out_stride);
  const int H = p.hidden;
  const int chunks = (H + 3) >> 2;
  const int64_t total = (int64_t)p.S * p.B * chunks;
  for (int64_t it = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; it < total;
       it += (int64_t)gridDim.x * blockDim.x) {
    const int ch = (int)(it % chunks);
    const int64_t g = it / chunks;
    const int s = (int)(g / p.B);
    const int b = (int)(g % p.B);
    const int32_t k0 = __ldg(lod + g), k1 = __ldg(lod + g + 1);
    float4 acc = make_float4(p.pad_value, p.pad_value, p.pad_value, p.pad_value);
    for (int32_t k = k0; k < k1; ++k) {
      const uint32_t r = restore ? __ldg(restore + k) : (uint32_t)k;
      const float4 v = *reinterpret_cast<const float4*>(x + (uint64_t)r * p.x_stride + ch * 4);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    float v[4] = {acc.x, acc.y, acc.z, acc.w};
    float* o = out + (int64_t)b * p.out_stride + (int64_t)s * p.Eo;
    if (p.use_cvm) {
      if (ch == 0) {
        const float ls = logf(v[0] + 1.f);
        v[1] = logf(v[1] + 1.f) - ls;
        v[0] = ls;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (ch * 4 + j < H) o[ch * 4 + j] = v[j];
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int col = ch * 4 + j;
        if (col >= 2 && col < H) o[col - 2] = v[j];
      }
    }
  }
}

// dX[k, col] for every occurrence k: columns < cvm_offset come from the CVM input
// (show / click labels), the others are dOut broadcast over the bag
// (fused_seqpool_cvm_op.cu:814-1010; CPU twin fused_seqpool_cvm_op.cc:394-460).
__global__ void k_seqpool_cvm_bwd(PbxSeqpoolParams p, const float* __restrict__ dout,
                                  const float* __restrict__ cvm, const int32_t* __restrict__ lod,
                                  float* __restrict__ dx, int dx_stride) {
  const int H = p.hidden;
  const int64_t total = (int64_t)p.S * p.B * H;
  const int blockw = p.Eo / p.concat;
  int dim_off;
  if (p.use_cvm) dim_off = (p.variant == 0 ? p.clk_filter : p.show_filter) ? 1 : 0;
  else dim_off = p.cvm_offset;
  for (int64_t it = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; it < total;
       it += (int64_t)gridDim.x * blockDim.x) {
    const int col = (int)(it % H);
    const int64_t g = it / H;
    const int s = (int)(g / p.B);
    const int b = (int)(g % p.B);
    const int32_t k0 = lod[g], k1 = lod[g + 1];
    for (int32_t k = k0; k < k1; ++k) {
      float v;
      if (col < p.cvm_offset) {
        v = cvm[(int64_t)b * p.cvm_offset + col];
      } else {
        int ci = min((int)(k - k0), p.concat - 1);  // remaining keys share the last block's grad
        v = dout[(int64_t)b * p.out_stride + (int64_t)s * p.Eo + (int64_t)ci * blockw + (col - dim_off)];
      }
      dx[(int64_t)k * dx_stride + col] = v;
    }
  }
}

// fan unique values out to occurrences: out[i, :hidden] = vals[restore[i], :hidden]
// (reference PullCopy*, box_wrapper.cu:74-143)
__global__ void k_pull_scatter(const float* __restrict__ vals, int vstride, const uint32_t* __restrict__ restore,
                               int64_t n, int hidden, int skip, float scale_embedx, float* __restrict__ out) {
  const int W = hidden - skip;
  const int64_t total = n * W;
  for (int64_t it = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; it < total;
       it += (int64_t)gridDim.x * blockDim.x) {
    const int col = (int)(it % W) + skip;
    const int64_t i = it / W;
    float v = vals[(uint64_t)restore[i] * vstride + col];
    if (col >= 3) v *= scale_embedx;
    out[it] = v;
  }
}

// key2slot[i] = slot of occurrence i (reference CopyKeysKernel, box_wrapper.cu:324-343)
__global__ void k_key2slot(const int32_t* __restrict__ lod, int B, int S, int64_t n, int32_t* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int lo = 0, hi = S;  // first slot whose end offset > i
    while (lo < hi) {
      int mid = (lo + hi) >> 1;
      if (lod[(int64_t)(mid + 1) * B] > i) hi = mid; else lo = mid + 1;
    }
    out[i] = lo;
  }
}

}  // namespace

extern "C" {

int pbx_seqpool_cvm_fwd_tail(const PbxSeqpoolParams* p, const float* d_x, const uint32_t* d_restore,
                             const int32_t* d_lod, float* d_out, const float* d_tail, int tail_ld, int tail_cols,
                             int tail_off, cudaStream_t st);
int pbx_seqpool_cvm_fwd(const PbxSeqpoolParams* p, const float* d_x, const uint32_t* d_restore,
                        const int32_t* d_lod, float* d_out, cudaStream_t st) {
  return pbx_seqpool_cvm_fwd_tail(p, d_x, d_restore, d_lod, d_out, nullptr, 0, 0, 0, st);
}

// The same, and d_tail [B, tail_cols] (row stride tail_ld) is copied into columns
// [tail_off, tail_off + tail_cols) of every output row by the same launch.
int pbx_seqpool_cvm_fwd_tail(const PbxSeqpoolParams* p, const float* d_x, const uint32_t* d_restore,
                             const int32_t* d_lod, float* d_out, const float* d_tail, int tail_ld, int tail_cols,
                             int tail_off, cudaStream_t st) {
  SeqpoolTail tail{d_tail, tail_ld, tail_cols, tail_off};
  if (d_tail) PBX_REQUIRE(tail_off + tail_cols <= p->out_stride, "seqpool tail exceeds the output row");
  PBX_REQUIRE(p->concat >= 1 && p->Eo % p->concat == 0, "bad embedx_concate_size");
  int64_t total = (int64_t)p->S * p->B * p->concat * p->hidden;
  if (total == 0) return 0;
  const int chunks = (p->hidden + 3) / 4;
  const bool fast = p->variant == 0 && p->cvm_offset == 2 && !p->clk_filter && !p->need_filter &&
                    !p->embed_threshold_filter && p->quant_ratio == 0 && p->concat == 1 && !p->trade_w &&
                    !p->slot_threshold && (p->x_stride & 3) == 0 && p->x_stride >= chunks * 4 &&
                    ((uintptr_t)d_x & 15) == 0;
  if (fast) {
    int64_t items = (int64_t)p->S * p->B * chunks;
    k_seqpool_cvm_fwd_fast<<<pbx_grid_for(items, 256, 16), 256, 0, st>>>(*p, d_x, d_restore, d_lod, d_out, tail);
    PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
    return 0;
  }
  k_seqpool_cvm_fwd<<<pbx_grid_for(total, 256, 16), 256, 0, st>>>(*p, d_x, d_restore, d_lod, d_out, tail);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_seqpool_cvm_bwd(const PbxSeqpoolParams* p, const float* d_dout, const float* d_cvm, const int32_t* d_lod,
                        float* d_dx, int dx_stride, cudaStream_t st) {
  int64_t total = (int64_t)p->S * p->B * p->hidden;
  if (total == 0) return 0;
  k_seqpool_cvm_bwd<<<pbx_grid_for(total, 256, 16), 256, 0, st>>>(*p, d_dout, d_cvm, d_lod, d_dx, dx_stride);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_pull_scatter(const float* d_vals, int vstride, const uint32_t* d_restore, int64_t n, int hidden, int skip,
                     float scale_embedx, float* d_out, cudaStream_t st) {
  int64_t total = n * (hidden - skip);
  if (total <= 0) return 0;
  k_pull_scatter<<<pbx_grid_for(total, 256, 16), 256, 0, st>>>(d_vals, vstride, d_restore, n, hidden, skip,
                                                                scale_embedx, d_out);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_key2slot(const int32_t* d_lod, int B, int S, int64_t n, int32_t* d_out, cudaStream_t st) {
  if (n <= 0) return 0;
  k_key2slot<<<pbx_grid_for(n, 256, 16), 256, 0, st>>>(d_lod, B, S, n, d_out);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

}  // extern "C"
