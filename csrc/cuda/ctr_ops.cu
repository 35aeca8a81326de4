// The remaining CTR operator family as single-pass sm_100a kernels (W5 policies and W10 of
// SURVEY.md §8).  The GEMM-shaped ops (rank_attention, batch_fc, scaled_fc, scaled_int8fc) run
// their products on the tcgen05 kernel of gemm_tcgen05.cu; what lives here is everything around
// those products — operand builders that gather / quantise / cast straight into the bf16 layouts
// the tensor-core kernel reads — and the memory-bound ops, one launch each:
//
//   rank_attention      expand-by-rank-offset A operand (+ transposed copy for the weight gradient)
//                       and the scatter of the input gradient     (rank_attention.cu.h:28-195)
//   scaled_int8fc       expand / clip / quantise to integer-valued bf16   (scaled_int8fc_op.cu:38-72)
//   fused_seqpool_cvm_with_pcoc   pool + filter + quant + PCOC CVM transform, and its backward
//                                                          (fused_seqpool_cvm_with_pcoc_op.cu:36-294)
//   fused_concat / fused_seqpool_concat   column slice + concat of many inputs, pooled variant,
//                       and the split backward                     (fused_concat_op.cu:34-258)
//   fused_seq_tensor    DIN sequence tensors (din / mask / side / session)   (fused_seq_tensor_op.cu:14-147)
//   cross_norm_hadamard field-pair [a, b, a*b, <a,b>] expansion and its backward
//                                                          (cross_norm_hadamard.cu.h:44-225)
#include "common.cuh"

namespace {

#define PBX_GRID_STRIDE(it, total)                                                          \
  for (int64_t it = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; it < (total);           \
       it += (int64_t)gridDim.x * blockDim.x)

// ---------------------------------------------------------------------------------------------
// rank_attention.  A_wide[i, (own_i * R + faster_k) * C + c] = x[index_k, c] for every valid k
// (own_i = rank of instance i, faster_k / index_k = rank and row of the k-th ad of its page view);
// every other element is zero.  Then out = A_wide . rank_param is ONE GEMM with K = R*R*C instead
// of the reference's N per-instance [1, R*C] x [R*C, out] products (rank_attention_op.cu:110).
__global__ void k_rank_expand(const float* __restrict__ x, int ldx, const int32_t* __restrict__ ro, int N, int C,
                              int R, __nv_bfloat16* __restrict__ out, int ldo, __nv_bfloat16* __restrict__ out_t,
                              int ldt) {
  const int K = R * R * C;
  const int W = 2 * R + 1;
  PBX_GRID_STRIDE(it, (int64_t)N * K) {
    const int i = (int)(it / K), j = (int)(it % K);
    const int blk = j / C, c = j - blk * C;
    const int own = blk / R, fj = blk - own * R;
    float v = 0.f;
    const int32_t* r = ro + (int64_t)i * W;
    if (r[0] - 1 == own) {
      for (int k = 0; k < R; ++k) {
        const int faster = r[1 + 2 * k] - 1, idx = r[2 + 2 * k];
        if (faster == fj && idx >= 0) v += __ldg(x + (int64_t)idx * ldx + c);
      }
    }
    const __nv_bfloat16 b = __float2bfloat16(v);
    if (out) out[(int64_t)i * ldo + j] = b;
    if (out_t) out_t[(int64_t)j * ldt + i] = b;
  }
}

// dx[index_k, c] += dA[i, (own_i*R + faster_k)*C + c]   (merge_input_gradient_kernel, rank_attention.cu.h:152-195)
__global__ void k_rank_dx(const float* __restrict__ dA, int lda, const int32_t* __restrict__ ro, int N, int C, int R,
                          float* __restrict__ dx, int ldx) {
  const int W = 2 * R + 1;
  PBX_GRID_STRIDE(it, (int64_t)N * R * C) {
    const int c = (int)(it % C);
    const int k = (int)((it / C) % R);
    const int i = (int)(it / ((int64_t)C * R));
    const int32_t* r = ro + (int64_t)i * W;
    const int own = r[0] - 1, faster = r[1 + 2 * k] - 1, idx = r[2 + 2 * k];
    if (own < 0 || faster < 0 || idx < 0) continue;
    atomicAdd(dx + (int64_t)idx * ldx + c, dA[(int64_t)i * lda + (own * R + faster) * C + c]);
  }
}

// ---------------------------------------------------------------------------------------------
// scaled_int8fc operand: q = trunc(clamp(v * expand, -clip, clip) / interval + 0.5), interval =
// 2*clip/range, as an integer-valued bf16 (|q| <= 128 is exact in bf16; the tensor-core product of
// two such operands accumulates exactly in fp32 while K * 128^2 < 2^24)
__global__ void k_int8_quant(const float* __restrict__ x, int ldx, int M, int K, float expand, float clip,
                             float inv_interval, __nv_bfloat16* __restrict__ out, int ldo,
                             __nv_bfloat16* __restrict__ out_t, int ldt) {
  PBX_GRID_STRIDE(it, (int64_t)M * K) {
    const int m = (int)(it / K), k = (int)(it % K);
    float v = __ldg(x + (int64_t)m * ldx + k) * expand;
    v = fminf(fmaxf(v, -clip), clip);
    float q = truncf(v * inv_interval + 0.5f);
    q = fminf(fmaxf(q, -128.f), 127.f);
    const __nv_bfloat16 b = __float2bfloat16(q);
    if (out) out[(int64_t)m * ldo + k] = b;
    if (out_t) out_t[(int64_t)k * ldt + m] = b;
  }
}

// ---------------------------------------------------------------------------------------------
// fused_seqpool_cvm_with_pcoc.  Value rows [N, H]: show, clk, show2, clk2, pclk_0..pclk_{pn-1}
// (used = 4 + pn statistics columns, padded to mx), then the embedding.  One thread per (bag,
// output column); the ratio columns pool the two statistics they need in the same pass.
struct PcocParams {
  int B, S, H, x_stride, out_stride, Eo;
  int use_cvm, used, mx, need_filter, quant_ratio;
  float pad_value, show_coeff, clk_coeff, threshold;
};

__global__ void k_seqpool_pcoc_fwd(PcocParams p, const float* __restrict__ x, const int32_t* __restrict__ lod,
                                   float* __restrict__ out) {
  const int pn = p.used - 4;
  const int emb0 = p.use_cvm ? 2 + 2 * pn : 0;        // first embedding column of the output block
  PBX_GRID_STRIDE(it, (int64_t)p.S * p.B * p.Eo) {
    const int oc = (int)(it % p.Eo);
    const int64_t g = it / p.Eo;
    const int s = (int)(g / p.B), b = (int)(g % p.B);
    // source columns: c1 is pooled into `a`; ratio columns also pool c2 into `d`
    int c1, c2 = -1;
    if (oc >= emb0) c1 = p.mx + (oc - emb0);
    else if (oc == 0) c1 = 0;
    else if (oc == 1) { c1 = 1; c2 = 0; }
    else if (oc < 2 + pn) { c1 = 4 + (oc - 2); c2 = 2; }
    else { c1 = 4 + (oc - 2 - pn); c2 = 3; }
    float a = p.pad_value, d = p.pad_value;
    for (int32_t k = lod[g]; k < lod[g + 1]; ++k) {
      const float* row = x + (int64_t)k * p.x_stride;
      if (p.need_filter && (row[0] - row[1]) * p.show_coeff + row[1] * p.clk_coeff < p.threshold) continue;
      float v = row[c1];
      if (c1 >= p.mx && p.quant_ratio > 0) v = truncf(v * p.quant_ratio + 0.5f) / p.quant_ratio;
      a += v;
      if (c2 >= 0) d += row[c2];
    }
    float r = a;
    if (oc < emb0) r = c2 >= 0 ? logf(a + 1.f) - logf(d + 1.f) : logf(a + 1.f);
    out[(int64_t)b * p.out_stride + (int64_t)s * p.Eo + oc] = r;
  }
}

// dx[k, :4] = cvm[b, :4]; dx[k, 4:used] = q_values[b, :pn] (zero without); dx[k, used:mx] = 0;
// dx[k, mx:] = dOut of the bag's embedding columns
__global__ void k_seqpool_pcoc_bwd(PcocParams p, const float* __restrict__ dout, const float* __restrict__ cvm,
                                   int cvm_ld, const float* __restrict__ q, const int32_t* __restrict__ lod,
                                   float* __restrict__ dx, int dx_stride) {
  const int pn = p.used - 4;
  const int emb0 = p.use_cvm ? 2 + 2 * pn : 0;
  PBX_GRID_STRIDE(it, (int64_t)p.S * p.B * p.H) {
    const int col = (int)(it % p.H);
    const int64_t g = it / p.H;
    const int s = (int)(g / p.B), b = (int)(g % p.B);
    float v;
    if (col < 4) v = cvm[(int64_t)b * cvm_ld + col];
    else if (col < p.used) v = q ? q[(int64_t)b * pn + (col - 4)] : 0.f;
    else if (col < p.mx) v = 0.f;
    else v = dout[(int64_t)b * p.out_stride + (int64_t)s * p.Eo + emb0 + (col - p.mx)];
    for (int32_t k = lod[g]; k < lod[g + 1]; ++k) dx[(int64_t)k * dx_stride + col] = v;
  }
}

// ---------------------------------------------------------------------------------------------
// fused_concat: out[b, i*len + c] = in_i[b, start + c]; backward splits the gradient back and
// zeroes the columns outside the slice.  Up to 64 inputs per launch (pointer table by value).
struct ConcatPtrs { const float* in[64]; float* din[64]; int ld[64]; int width[64]; };

__global__ void k_fused_concat_fwd(ConcatPtrs t, int n_in, int B, int start, int len, float* __restrict__ out,
                                   int ldo) {
  PBX_GRID_STRIDE(it, (int64_t)B * n_in * len) {
    const int c = (int)(it % len);
    const int i = (int)((it / len) % n_in);
    const int b = (int)(it / ((int64_t)len * n_in));
    out[(int64_t)b * ldo + (int64_t)i * len + c] = __ldg(t.in[i] + (int64_t)b * t.ld[i] + start + c);
  }
}

__global__ void k_fused_concat_bwd(ConcatPtrs t, int n_in, int B, int start, int len, int max_w,
                                   const float* __restrict__ dout, int ldo) {
  PBX_GRID_STRIDE(it, (int64_t)B * n_in * max_w) {
    const int c = (int)(it % max_w);
    const int i = (int)((it / max_w) % n_in);
    const int b = (int)(it / ((int64_t)max_w * n_in));
    if (c >= t.width[i]) continue;
    const bool in = c >= start && c < start + len;
    t.din[i][(int64_t)b * t.ld[i] + c] = in ? dout[(int64_t)b * ldo + (int64_t)i * len + (c - start)] : 0.f;
  }
}

// fused_seqpool_concat: two groups of per-key rows sharing one lod are sum-pooled per (slot, ins)
// and the column ranges [off_g, off_g + dim_g) are laid side by side:
//   out[b, s*(d0+d1) + c]       = sum_k x0[k, off0 + c]        (c < d0)
//   out[b, s*(d0+d1) + d0 + c]  = sum_k x1[k, off1 + c]        (c < d1)
__global__ void k_seqpool_concat_fwd(const float* __restrict__ x0, int ld0, int off0, int d0,
                                     const float* __restrict__ x1, int ld1, int off1, int d1,
                                     const int32_t* __restrict__ lod, int B, int S, float* __restrict__ out, int ldo) {
  const int W = d0 + d1;
  PBX_GRID_STRIDE(it, (int64_t)S * B * W) {
    const int c = (int)(it % W);
    const int64_t g = it / W;
    const int s = (int)(g / B), b = (int)(g % B);
    const float* x = c < d0 ? x0 : x1;
    const int ld = c < d0 ? ld0 : ld1;
    const int col = c < d0 ? off0 + c : off1 + (c - d0);
    float a = 0.f;
    for (int32_t k = lod[g]; k < lod[g + 1]; ++k) a += __ldg(x + (int64_t)k * ld + col);
    out[(int64_t)b * ldo + (int64_t)s * W + c] = a;
  }
}

__global__ void k_seqpool_concat_bwd(const float* __restrict__ dout, int ldo, const int32_t* __restrict__ lod, int B,
                                     int S, float* __restrict__ dx0, int ld0, int w0, int off0, int d0,
                                     float* __restrict__ dx1, int ld1, int w1, int off1, int d1) {
  const int W = d0 + d1;
  const int WW = w0 + w1;
  PBX_GRID_STRIDE(it, (int64_t)S * B * WW) {
    const int cc = (int)(it % WW);
    const int64_t g = it / WW;
    const int s = (int)(g / B), b = (int)(g % B);
    const bool first = cc < w0;
    const int col = first ? cc : cc - w0;
    const int off = first ? off0 : off1, d = first ? d0 : d1;
    float v = 0.f;
    if (col >= off && col < off + d)
      v = dout[(int64_t)b * ldo + (int64_t)s * W + (first ? 0 : d0) + (col - off)];
    float* dx = first ? dx0 : dx1;
    const int ld = first ? ld0 : ld1;
    for (int32_t k = lod[g]; k < lod[g + 1]; ++k) dx[(int64_t)k * ld + col] = v;
  }
}

// ---------------------------------------------------------------------------------------------
// fused_seq_tensor.  x [ins, bc, slot_num, L, D], ad [ins, bc, an, D] ->
//   din  [bc, ins, L, 4*an*D] = [seq, ad, seq - ad, seq * ad] of the ad slots
//   sess [bc, ins, L, an*D], side [bc, ins, L, (slot_num-an)*D], mask [bc, ins, L]
__global__ void k_seq_tensor(const float* __restrict__ x, const float* __restrict__ ad, int ins, int bc, int SN,
                             int L, int D, int an, int ad_off, int side_off, float* __restrict__ din,
                             float* __restrict__ sess, float* __restrict__ side, float* __restrict__ mask) {
  const int AD = an * D, SD = (SN - an) * D;
  const int64_t n_din = (int64_t)bc * ins * L * AD;
  const int64_t n_side = (int64_t)bc * ins * L * SD;
  const int64_t n_mask = (int64_t)bc * ins * L;
  PBX_GRID_STRIDE(it, n_din + n_side + n_mask) {
    if (it < n_din) {
      const int e = (int)(it % AD);
      const int t = (int)((it / AD) % L);
      const int i = (int)((it / ((int64_t)AD * L)) % ins);
      const int c = (int)(it / ((int64_t)AD * L * ins));
      const int a = e / D, d = e - a * D;
      const float sv = __ldg(x + ((((int64_t)i * bc + c) * SN + ad_off + a) * L + t) * D + d);
      const float av = __ldg(ad + (((int64_t)i * bc + c) * an + a) * D + d);
      float* o = din + (((int64_t)c * ins + i) * L + t) * (4 * AD);
      o[e] = sv;
      o[AD + e] = av;
      o[2 * AD + e] = sv - av;
      o[3 * AD + e] = sv * av;
      sess[(((int64_t)c * ins + i) * L + t) * AD + e] = sv;
    } else if (it < n_din + n_side) {
      const int64_t j = it - n_din;
      const int e = (int)(j % SD);
      const int t = (int)((j / SD) % L);
      const int i = (int)((j / ((int64_t)SD * L)) % ins);
      const int c = (int)(j / ((int64_t)SD * L * ins));
      const int a = e / D, d = e - a * D;
      side[(((int64_t)c * ins + i) * L + t) * SD + e] =
          __ldg(x + ((((int64_t)i * bc + c) * SN + side_off + a) * L + t) * D + d);
    } else {
      const int64_t j = it - n_din - n_side;
      const int t = (int)(j % L);
      const int i = (int)((j / L) % ins);
      const int c = (int)(j / ((int64_t)L * ins));
      float a = 0.f;
      for (int sl = 0; sl < SN; ++sl) {
        const float* p = x + ((((int64_t)i * bc + c) * SN + sl) * L + t) * D;
        for (int d = 0; d < D; ++d) a += __ldg(p + d);
      }
      mask[j] = fabsf(a) > 1e-8f ? 1.f : 0.f;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// cross_norm_hadamard expansion.  x [B, f, 2, d] -> y [B, f, 3d+1] = [a, b, a*b, <a,b>];
// backward: da = ga + (gh + gdot) * b, db = gb + (gh + gdot) * a.
__global__ void k_cross_hadamard_fwd(const float* __restrict__ x, int ldx, int B, int f, int d, float* __restrict__ y,
                                     int ldy) {
  const int W = 3 * d + 1;
  PBX_GRID_STRIDE(it, (int64_t)B * f * W) {
    const int c = (int)(it % W);
    const int fi = (int)((it / W) % f);
    const int b = (int)(it / ((int64_t)W * f));
    const float* a = x + (int64_t)b * ldx + (int64_t)fi * 2 * d;
    const float* bb = a + d;
    float v;
    if (c < d) v = a[c];
    else if (c < 2 * d) v = bb[c - d];
    else if (c < 3 * d) v = a[c - 2 * d] * bb[c - 2 * d];
    else {
      v = 0.f;
      for (int j = 0; j < d; ++j) v += a[j] * bb[j];
    }
    y[(int64_t)b * ldy + (int64_t)fi * W + c] = v;
  }
}

__global__ void k_cross_hadamard_bwd(const float* __restrict__ x, int ldx, const float* __restrict__ gy, int ldy,
                                     int B, int f, int d, float* __restrict__ dx, int lddx) {
  const int W = 3 * d + 1;
  PBX_GRID_STRIDE(it, (int64_t)B * f * 2 * d) {
    const int c = (int)(it % (2 * d));
    const int fi = (int)((it / (2 * d)) % f);
    const int b = (int)(it / ((int64_t)2 * d * f));
    const float* g = gy + (int64_t)b * ldy + (int64_t)fi * W;
    const float* xa = x + (int64_t)b * ldx + (int64_t)fi * 2 * d;
    const int j = c < d ? c : c - d;
    const float other = c < d ? xa[d + j] : xa[j];
    dx[(int64_t)b * lddx + (int64_t)fi * 2 * d + c] = g[c] + (g[2 * d + j] + g[3 * d]) * other;
  }
}

}  // namespace

extern "C" {

int pbx_rank_attention_expand(const float* d_x, int ldx, const int32_t* d_rank_offset, int N, int C, int R,
                              void* d_out_bf16, int ldo, void* d_out_t_bf16, int ldt, cudaStream_t st) {
  if (N <= 0) return 0;
  PBX_REQUIRE(C > 0 && R > 0 && R <= 16, "rank_attention: bad C %d / max_rank %d", C, R);
  k_rank_expand<<<pbx_grid_for((int64_t)N * R * R * C, 256, 16), 256, 0, st>>>(
      d_x, ldx, d_rank_offset, N, C, R, (__nv_bfloat16*)d_out_bf16, ldo, (__nv_bfloat16*)d_out_t_bf16, ldt);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

// d_dx [N, C] must be zero on entry (gradients of rows referenced by several instances add up)
int pbx_rank_attention_dx(const float* d_dA, int lda, const int32_t* d_rank_offset, int N, int C, int R, float* d_dx,
                          int ldx, cudaStream_t st) {
  if (N <= 0) return 0;
  k_rank_dx<<<pbx_grid_for((int64_t)N * R * C, 256, 16), 256, 0, st>>>(d_dA, lda, d_rank_offset, N, C, R, d_dx, ldx);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_int8_quant_bf16(const float* d_x, int ldx, int M, int K, float expand, float clip, float int8_range,
                        void* d_out, int ldo, void* d_out_t, int ldt, cudaStream_t st) {
  if ((int64_t)M * K <= 0) return 0;
  PBX_REQUIRE(clip > 0.f && int8_range > 0.f, "int8 quant: clip and range must be positive");
  const float inv_interval = int8_range / (2.f * clip);
  k_int8_quant<<<pbx_grid_for((int64_t)M * K, 256, 16), 256, 0, st>>>(d_x, ldx, M, K, expand, clip, inv_interval,
                                                                      (__nv_bfloat16*)d_out, ldo,
                                                                      (__nv_bfloat16*)d_out_t, ldt);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

static int pcoc_params(PcocParams& p, int B, int S, int H, int x_stride, int out_stride, int use_cvm, int used, int mx,
                       int need_filter, int quant_ratio, float pad_value, float show_coeff, float clk_coeff,
                       float threshold) {
  PBX_REQUIRE(used >= 4 && mx >= used && H > mx, "with_pcoc: need 4 <= cvm_offset %d <= max_cvm_offset %d < hidden %d",
              used, mx, H);
  p.B = B; p.S = S; p.H = H; p.x_stride = x_stride; p.out_stride = out_stride;
  p.use_cvm = use_cvm; p.used = used; p.mx = mx; p.need_filter = need_filter; p.quant_ratio = quant_ratio;
  p.pad_value = pad_value; p.show_coeff = show_coeff; p.clk_coeff = clk_coeff; p.threshold = threshold;
  p.Eo = (use_cvm ? 2 + 2 * (used - 4) : 0) + (H - mx);
  return 0;
}

int pbx_seqpool_pcoc_fwd(const float* d_x, int x_stride, const int32_t* d_lod, int B, int S, int H, int use_cvm,
                         int used, int mx, int need_filter, int quant_ratio, float pad_value, float show_coeff,
                         float clk_coeff, float threshold, float* d_out, int out_stride, cudaStream_t st) {
  PcocParams p;
  if (pcoc_params(p, B, S, H, x_stride, out_stride, use_cvm, used, mx, need_filter, quant_ratio, pad_value, show_coeff,
                  clk_coeff, threshold)) return -1;
  PBX_REQUIRE(out_stride >= S * p.Eo, "with_pcoc: out_stride %d < S*Eo %d", out_stride, S * p.Eo);
  const int64_t total = (int64_t)S * B * p.Eo;
  if (total == 0) return 0;
  k_seqpool_pcoc_fwd<<<pbx_grid_for(total, 256, 16), 256, 0, st>>>(p, d_x, d_lod, d_out);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_seqpool_pcoc_bwd(const float* d_dout, int out_stride, const float* d_cvm, int cvm_ld, const float* d_q,
                         const int32_t* d_lod, int B, int S, int H, int use_cvm, int used, int mx, float* d_dx,
                         int dx_stride, cudaStream_t st) {
  PcocParams p;
  if (pcoc_params(p, B, S, H, dx_stride, out_stride, use_cvm, used, mx, 0, 0, 0.f, 0.f, 0.f, 0.f)) return -1;
  const int64_t total = (int64_t)S * B * H;
  if (total == 0) return 0;
  k_seqpool_pcoc_bwd<<<pbx_grid_for(total, 256, 16), 256, 0, st>>>(p, d_dout, d_cvm, cvm_ld, d_q, d_lod, d_dx,
                                                                    dx_stride);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

// inputs: n_in device pointers (host array), row strides and widths; d_din NULL for the forward
int pbx_fused_concat(const float* const* d_in, float* const* d_din, const int* ld, const int* width, int n_in, int B,
                     int start, int len, float* d_out_or_dout, int ldo, int backward, cudaStream_t st) {
  PBX_REQUIRE(n_in >= 1 && n_in <= 64, "fused_concat: %d inputs (max 64 per launch)", n_in);
  PBX_REQUIRE(start >= 0 && len > 0, "fused_concat: bad slice");
  ConcatPtrs t;
  int max_w = 0;
  for (int i = 0; i < n_in; ++i) {
    PBX_REQUIRE(start + len <= width[i], "fused_concat: slice [%d, %d) exceeds input %d of width %d", start,
                start + len, i, width[i]);
    t.in[i] = d_in ? d_in[i] : nullptr;
    t.din[i] = d_din ? d_din[i] : nullptr;
    t.ld[i] = ld[i];
    t.width[i] = width[i];
    if (width[i] > max_w) max_w = width[i];
  }
  if (B <= 0) return 0;
  if (!backward)
    k_fused_concat_fwd<<<pbx_grid_for((int64_t)B * n_in * len, 256, 16), 256, 0, st>>>(t, n_in, B, start, len,
                                                                                       d_out_or_dout, ldo);
  else
    k_fused_concat_bwd<<<pbx_grid_for((int64_t)B * n_in * max_w, 256, 16), 256, 0, st>>>(t, n_in, B, start, len, max_w,
                                                                                         d_out_or_dout, ldo);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_seqpool_concat_fwd(const float* d_x0, int ld0, int off0, int d0, const float* d_x1, int ld1, int off1, int d1,
                           const int32_t* d_lod, int B, int S, float* d_out, int ldo, cudaStream_t st) {
  const int64_t total = (int64_t)S * B * (d0 + d1);
  if (total <= 0) return 0;
  k_seqpool_concat_fwd<<<pbx_grid_for(total, 256, 16), 256, 0, st>>>(d_x0, ld0, off0, d0, d_x1, ld1, off1, d1, d_lod, B,
                                                                     S, d_out, ldo);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_seqpool_concat_bwd(const float* d_dout, int ldo, const int32_t* d_lod, int B, int S, float* d_dx0, int ld0,
                           int w0, int off0, int d0, float* d_dx1, int ld1, int w1, int off1, int d1,
                           cudaStream_t st) {
  const int64_t total = (int64_t)S * B * (w0 + w1);
  if (total <= 0) return 0;
  k_seqpool_concat_bwd<<<pbx_grid_for(total, 256, 16), 256, 0, st>>>(d_dout, ldo, d_lod, B, S, d_dx0, ld0, w0, off0, d0,
                                                                     d_dx1, ld1, w1, off1, d1);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_fused_seq_tensor(const float* d_x, const float* d_ad, int ins, int bc, int slot_num, int max_length, int dim,
                         int ad_slot_num, int ad_slot_offset, float* d_din, float* d_sess, float* d_side,
                         float* d_mask, cudaStream_t st) {
  PBX_REQUIRE(ad_slot_num > 0 && ad_slot_num <= slot_num && ad_slot_offset >= 0 &&
                  ad_slot_offset + ad_slot_num <= slot_num, "fused_seq_tensor: bad ad slot range");
  const int side_off = ad_slot_offset == 0 ? ad_slot_num : 0;
  const int64_t total = (int64_t)bc * ins * max_length * ((int64_t)slot_num * dim + 1);
  if (total <= 0) return 0;
  k_seq_tensor<<<pbx_grid_for(total, 256, 16), 256, 0, st>>>(d_x, d_ad, ins, bc, slot_num, max_length, dim,
                                                             ad_slot_num, ad_slot_offset, side_off, d_din, d_sess,
                                                             d_side, d_mask);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_cross_hadamard_fwd(const float* d_x, int ldx, int B, int fields, int dim, float* d_y, int ldy,
                           cudaStream_t st) {
  const int64_t total = (int64_t)B * fields * (3 * dim + 1);
  if (total <= 0) return 0;
  k_cross_hadamard_fwd<<<pbx_grid_for(total, 256, 16), 256, 0, st>>>(d_x, ldx, B, fields, dim, d_y, ldy);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_cross_hadamard_bwd(const float* d_x, int ldx, const float* d_gy, int ldy, int B, int fields, int dim,
                           float* d_dx, int lddx, cudaStream_t st) {
  const int64_t total = (int64_t)B * fields * 2 * dim;
  if (total <= 0) return 0;
  k_cross_hadamard_bwd<<<pbx_grid_for(total, 256, 16), 256, 0, st>>>(d_x, ldx, d_gy, ldy, B, fields, dim, d_dx, lddx);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

}  // extern "C"
