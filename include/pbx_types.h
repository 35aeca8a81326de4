/* paddlebox_b200 — plain-C types shared by the CUDA library, the host library and
 * the Python ctypes bindings.
 *
 * Feature-value row layout (fp32 words), D = embedx_dim.  Semantics follow the
 * HeterPS accessor (reference: paddle/fluid/framework/fleet/heter_ps/feature_value.h:44-158)
 * but the order is ours: the pull record is a literal prefix of the row, so serving a
 * pull is a straight 16-byte-vector copy of the first pull_stride words.
 *
 *   [0] show  [1] click  [2] embed_w
 *   [3 .. 3+D)            embedx_w   (all zero until the embedx is created)
 *   [M] mf_size (0 = embedx not created yet, else D)           (M = 3 + D)
 *   [T+0] delta_score  [T+1] slot  [T+2] unseen_days           (T = 4 + D)
 *   [T+3 .. T+3+es)       embed optimizer state   (adagrad: g2sum | adam: m,v,b1pow,b2pow)
 *   [.. +xs)              embedx optimizer state  (adagrad: g2sum | adam: m[D],v[D],b1pow,b2pow |
 *                                                  shared adam: m,v,b1pow,b2pow)
 *   row stride is rounded up to a multiple of 4 words.
 *
 * Pull record : [show, click, embed_w, embedx_w[D]]            hidden = 3 + D
 *               (reference: feature_value.h:161-183, box_wrapper.cu:74-106)
 * Push record : [slot, show, click, embed_g, embedx_g[D]]      4 + D words
 *               (reference: feature_value.h:185-237, box_wrapper.cu:344-416)
 */
#ifndef PBX_TYPES_H_
#define PBX_TYPES_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum PbxOptimizer { PBX_OPT_ADAGRAD = 0, PBX_OPT_ADAM = 1, PBX_OPT_SHARED_ADAM = 2 };

/* Mirror of the reference's OptimizerConfig (heter_ps/optimizer_conf.h:20-45). */
typedef struct PbxSparseOptConfig {
  float nonclk_coeff;         /* 0.1  */
  float clk_coeff;            /* 1.0  */
  float min_bound;            /* -10  */
  float max_bound;            /* 10   */
  float learning_rate;        /* 0.05 */
  float initial_g2sum;        /* 3.0  */
  float initial_range;        /* 0    */
  float beta1_decay_rate;     /* 0.9  */
  float beta2_decay_rate;     /* 0.999*/
  float ada_epsilon;          /* 1e-8 */
  float mf_create_thresholds; /* 10   */
  float mf_learning_rate;     /* 0.05 */
  float mf_initial_g2sum;     /* 3.0  */
  float mf_initial_range;     /* 1e-4 */
  float mf_beta1_decay_rate;  /* 0.9  */
  float mf_beta2_decay_rate;  /* 0.999*/
  float mf_min_bound;         /* -10  */
  float mf_max_bound;         /* 10   */
  float mf_ada_epsilon;       /* 1e-8 */
  float reserved;
} PbxSparseOptConfig;

typedef struct PbxRowLayout {
  int embedx_dim;  /* D */
  int optimizer;   /* PbxOptimizer */
  int embed_state; /* es */
  int embedx_state;/* xs */
  int off_mf;      /* M: mf_size word */
  int off_tail;    /* T */
  int off_es;
  int off_xs;
  int stride;      /* words per row */
  int pull_stride; /* round_up(3 + D, 4) */
  int push_stride; /* round_up(4 + D, 4) */
} PbxRowLayout;

static inline PbxRowLayout pbx_make_layout(int embedx_dim, int optimizer) {
  PbxRowLayout l;
  l.embedx_dim = embedx_dim;
  l.optimizer = optimizer;
  l.embed_state = optimizer == PBX_OPT_ADAGRAD ? 1 : 4;
  l.embedx_state = optimizer == PBX_OPT_ADAGRAD ? 1
                   : optimizer == PBX_OPT_ADAM  ? 2 * embedx_dim + 2
                                                : 4;
  l.off_mf = 3 + embedx_dim;
  l.off_tail = 4 + embedx_dim;
  l.off_es = l.off_tail + 3;
  l.off_xs = l.off_es + l.embed_state;
  int words = l.off_xs + l.embedx_state;
  /* rows are padded to whole 128-byte lines: a row (and its read-modify-write by the optimizer)
   * then touches the minimum number of DRAM lines; random-line throughput, not bandwidth, bounds
   * the sparse kernels */
  l.stride = (words + 1 + 31) / 32 * 32;   /* + 1: the last word of a row is its update lock */
  l.pull_stride = (3 + embedx_dim + 3) / 4 * 4;
  l.push_stride = (4 + embedx_dim + 3) / 4 * 4;
  return l;
}

/* Attributes of the fused_seqpool_cvm family (reference attrs:
 * operators/fused/fused_seqpool_cvm_op.cc:124-146). Layout: bags slot-major, output [B, S*Eo]. */
typedef struct PbxSeqpoolParams {
  int B, S, hidden;
  int x_stride, out_stride, Eo;
  int use_cvm, cvm_offset, clk_filter;
  int need_filter, quant_ratio;
  int embed_threshold_filter, embed_thres_size;
  int concat, concat_filter;
  int variant;      /* 0 base, 1 with_conv, 2 with_credit */
  int show_filter;
  float pad_value, show_coeff, clk_coeff, threshold, embed_threshold;
  const float* slot_threshold; /* with_diff_thres: device [S] or NULL */
  const float* trade_w;        /* tradew: device [N] or NULL */
} PbxSeqpoolParams;

/* One problem of a grouped GEMM launch (pbx_gemm_bf16_tn_grouped): the arguments of
 * pbx_gemm_bf16_tn as a record.  All problems of a launch share the tile width. */
typedef struct PbxGemmProblem {
  const void* A;        /* bf16 [M, lda] */
  const void* B;        /* bf16 [N, ldb] */
  const float* bias;
  const void* mask;
  void* out_bf16;
  void* out_t;
  float* out_f32;
  float* col_out;
  int lda, ldb, M, N, K;
  int ld_mask, ld_out, ld_t, ld_f32;
  int relu, accumulate, splitk, col_split;
  float alpha;
} PbxGemmProblem;

#ifdef __cplusplus
}
#endif
#endif /* PBX_TYPES_H_ */
