/* paddlebox_b200 — C ABI of the sm_100a kernel library (libpbx_cuda.so).
 *
 * Every function returns 0 on success; on failure it returns a CUDA error code (> 0) or
 * -1 and pbx_last_error() describes the problem.  `stream` arguments are cudaStream_t
 * handles (pass torch.cuda.current_stream().cuda_stream).  Pointers prefixed d_ are
 * device pointers.  Reference citations are relative to the PaddleBox tree.
 */
#ifndef PBX_CUDA_H_
#define PBX_CUDA_H_
#include <stdint.h>
#include "pbx_types.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef void* pbx_stream_t;

const char* pbx_last_error(void);
/* number of kernel launches this library has issued in the calling process */
long long pbx_launch_count(void);

/* ---- sparse table shard (csrc/cuda/table.cu) --------------------------------
 * reference: heter_ps/hashtable.h, hashtable_kernel.cu:34-337, optimizer.cuh.h,
 * ctr_dymf_accessor.cc:68-142 (shrink). */
int pbx_table_create(void** out, uint64_t capacity, int embedx_dim, int optimizer,
                     const PbxSparseOptConfig* cfg, uint64_t seed);
int pbx_table_destroy(void* table);
int pbx_table_layout(void* table, PbxRowLayout* out);
int pbx_table_set_config(void* table, const PbxSparseOptConfig* cfg);
int pbx_table_set_test_mode(void* table, int on);
int pbx_table_clear(void* table, pbx_stream_t stream);
int pbx_table_insert(void* table, const uint64_t* d_keys, int64_t n, const float* d_rows, pbx_stream_t stream);
int pbx_table_fill_synthetic(void* table, uint64_t n_total, uint64_t key_seed, int rank, int nranks,
                             int create_mf, pbx_stream_t stream);
int pbx_table_count(void* table, int64_t* out, pbx_stream_t stream);
int pbx_table_dump(void* table, uint64_t* d_keys_out, float* d_rows_out, int64_t max_n, int mode,
                   int clear_dirty, int64_t* n_out, pbx_stream_t stream);
int pbx_table_lookup(void* table, const uint64_t* d_keys, int64_t n, float* d_out, int out_stride,
                     uint32_t* d_rows_out, pbx_stream_t stream);
int pbx_table_update(void* table, const uint64_t* d_keys, const float* d_push, int push_stride, int64_t n,
                     pbx_stream_t stream);
/* full value rows [n, stride] by key (zero row + found = 0 when absent); d_found may be NULL.
 * reference: dump_to_cpu (heter_ps/hashtable_kernel.cu:247-301) walks the whole table; this is the
 * by-key form used for chunked EndPass write-back. */
int pbx_table_fetch_rows(void* table, const uint64_t* d_keys, int64_t n, float* d_rows_out, uint8_t* d_found,
                         pbx_stream_t stream);
int pbx_table_shrink(void* table, float decay, float delete_threshold, float delete_after_unseen_days,
                     int64_t* kept, pbx_stream_t stream);

/* ---- sharded sparse pull / push over NVLink peer memory (csrc/cuda/sparse_a2a.cu) ----
 * reference: HeterComm::pull_sparse / push_sparse (heter_ps/heter_comm_inl.h:1477-1825),
 * BoxWrapper::PullSparseCaseGPU / PushSparseGradCaseGPU (fleet/box_wrapper_impl.h:25-522),
 * dedup contract heter_comm_inl.h:2231-2343. */
int pbx_sparse_create(void** out, void* table, int rank, int nranks, int64_t max_keys, int64_t cap);
int pbx_sparse_destroy(void* ctx);
int pbx_sparse_ipc_handle(void* ctx, void* out64);
int pbx_sparse_connect(void* ctx, const void* handles);
int pbx_sparse_connect_local(void* ctx, void* const* peers);
int pbx_sparse_geometry(void* ctx, int64_t* cap, int64_t* n_unique_rows, int* pull_stride, int* push_stride);
float* pbx_sparse_pull_values(void* ctx);
int pbx_sparse_status(void* ctx, int* status, int reset, pbx_stream_t stream);
/* Non-blocking: the status as of the last push whose 4-byte status copy has reached pinned host
 * memory (every push enqueues one).  Lets the trainer fail at the NEXT step after a mailbox
 * overflow / barrier timeout without synchronising. */
int pbx_sparse_status_peek(void* ctx);
/* diagnosis: per-CTA phase-end timestamps of the last pull / push (benchmarks/micro.py --what imbalance) */
int pbx_sparse_debug_stamps(void* ctx, int on, uint64_t* out, int* grid_out, pbx_stream_t stream);
/* Copies the 16 phase-boundary globaltimer stamps (ns) of the last pull (0..5) and push (8..12); synchronises. */
int pbx_sparse_phase_times(void* ctx, uint64_t* out16, pbx_stream_t stream);
int pbx_sparse_pull(void* ctx, const uint64_t* d_keys, int64_t n, uint32_t* d_restore, pbx_stream_t stream);
/* d_n: optional device int32 with the real key count (n is then the upper bound) — graph-replay friendly */
int pbx_sparse_pull_ex(void* ctx, const uint64_t* d_keys, int64_t n, const int32_t* d_n, uint32_t* d_restore,
                       pbx_stream_t stream);
int pbx_sparse_push(void* ctx, const uint32_t* d_restore, int64_t n, const float* d_dout, const float* d_cvm,
                    const int32_t* d_lod, int B, int S, int Eo, int cvm_offset, int dim_off, const float* d_gkeys,
                    const int32_t* d_key2slot, int hidden, int mode, float grad_scale, pbx_stream_t stream);
/* as pbx_sparse_push, with an explicit row stride (in words) of d_dout */
int pbx_sparse_push_ex(void* ctx, const uint32_t* d_restore, int64_t n, const float* d_dout, int ld_dout,
                       const float* d_cvm, const int32_t* d_lod, int B, int S, int Eo, int cvm_offset, int dim_off,
                       const float* d_gkeys, const int32_t* d_key2slot, int hidden, int mode, float grad_scale,
                       pbx_stream_t stream);
/* push_ex with the data-parallel dense-gradient allreduce folded into the kernel's own barriers:
 * `symm` (a pbx_symm handle of the same group) holds the flat gradient buffer at
 * [dense_off, dense_off + dense_n) floats; afterwards that range is sum over ranks * dense_scale
 * on every rank.  symm == NULL: plain push. */
int pbx_sparse_push_dense(void* ctx, const uint32_t* d_restore, int64_t n, const float* d_dout, int ld_dout,
                          const float* d_cvm, const int32_t* d_lod, int B, int S, int Eo, int cvm_offset,
                          int dim_off, const float* d_gkeys, const int32_t* d_key2slot, int hidden, int mode,
                          float grad_scale, void* symm, int64_t dense_off, int64_t dense_n, float dense_scale,
                          pbx_stream_t stream);
int pbx_sparse_pull_emu(void* const* ctxs, int nranks, const uint64_t* const* d_keys, const int64_t* n,
                        uint32_t* const* d_restore, pbx_stream_t stream);
int pbx_sparse_push_emu(void* const* ctxs, int nranks, const uint32_t* const* d_restore, const int64_t* n,
                        const float* const* d_dout, const float* const* d_cvm, const int32_t* const* d_lod, int B,
                        int S, int Eo, int cvm_offset, int dim_off, float grad_scale, pbx_stream_t stream);

/* ---- fused_seqpool_cvm family (csrc/cuda/seqpool.cu) ------------------------
 * reference: operators/fused/fused_seqpool_cvm_op.cu:34-1349 and the
 * _with_conv/_with_credit/_with_diff_thres/_tradew variants. */
int pbx_seqpool_cvm_fwd(const PbxSeqpoolParams* p, const float* d_x, const uint32_t* d_restore,
                        const int32_t* d_lod, float* d_out, pbx_stream_t stream);
/* the same launch also copies d_tail [B, tail_cols] (row stride tail_ld) into columns
 * [tail_off, tail_off + tail_cols) of every output row (dense features next to the pooled slots) */
int pbx_seqpool_cvm_fwd_tail(const PbxSeqpoolParams* p, const float* d_x, const uint32_t* d_restore,
                             const int32_t* d_lod, float* d_out, const float* d_tail, int tail_ld, int tail_cols,
                             int tail_off, pbx_stream_t stream);
int pbx_seqpool_cvm_bwd(const PbxSeqpoolParams* p, const float* d_dout, const float* d_cvm,
                        const int32_t* d_lod, float* d_dx, int dx_stride, pbx_stream_t stream);
int pbx_pull_scatter(const float* d_vals, int vstride, const uint32_t* d_restore, int64_t n, int hidden, int skip,
                     float scale_embedx, float* d_out, pbx_stream_t stream);
int pbx_key2slot(const int32_t* d_lod, int B, int S, int64_t n, int32_t* d_out, pbx_stream_t stream);

/* ---- tcgen05 / TMA GEMM (csrc/cuda/gemm_tcgen05.cu) -------------------------
 * C = epilogue(alpha * A[M,K] . B[N,K]^T); A, B bf16 K-contiguous.
 * reference call sites: scaled_fc_op.cu:214, batch_fc_op.cu:243, rank_attention_op.cu:110,
 * phi mul/matmul/fc kernels (all cuBLAS). */
/* n <= 8 independent problems in one launch (block_n in {32, 64, 128}; 0 = 64) */
int pbx_gemm_bf16_tn_grouped(const PbxGemmProblem* problems, int n, int block_n, pbx_stream_t stream);
/* debug: in-kernel timeline (8 globaltimer stamps of CTA 0) of the next tcgen05 GEMM launches */
int pbx_gemm_debug_enable(int on);
int pbx_gemm_debug_ts(unsigned long long* out8);
int pbx_gemm_bf16_tn(const void* A, int lda, const void* B, int ldb, int M, int N, int K, const float* bias,
                     const void* mask, int ld_mask, void* out_bf16, int ld_out, void* out_t, int ld_t,
                     float* out_f32, int ld_f32, int relu, int accumulate, float alpha, int splitk, int block_n,
                     int impl, float* col_out, int col_split, pbx_stream_t stream);
/* col_out/col_split (optional): output column `col_split` is written to col_out[row] instead of
 * out_f32 — with a row of ones appended to B this yields the bias gradient inside the wgrad GEMM. */

/* ---- dense helpers (csrc/cuda/dense_ops.cu) --------------------------------- */
int pbx_cast_transpose(const float* x, int ld, int M, int K, void* out_bf16, int ldo, void* out_t, int ldt,
                       pbx_stream_t stream);
/* the same for n <= 12 tensors in one launch (arrays of length n; out / out_t entries may be NULL) */
int pbx_cast_transpose_multi(int n, const float* const* x, const int* ld, const int* M, const int* K,
                             void* const* out, const int* ldo, void* const* out_t, const int* ldt,
                             pbx_stream_t stream);
int pbx_colsum_bf16(const void* x, int ld, int M, int N, float* out, pbx_stream_t stream);
/* step_dev: 2 floats {step count, scratch}; the kernel advances the step itself (no second launch) */
int pbx_adam_flat(float* p, float* g, float* m, float* v, int64_t n, float lr, float b1, float b2, float eps,
                  float weight_decay, float grad_scale, float* step_dev, int zero_grad, pbx_stream_t stream);
int pbx_sgd_flat(float* p, float* g, int64_t n, float lr, float grad_scale, int zero_grad, pbx_stream_t stream);
int pbx_sigmoid_logloss(const float* logit, const float* label, int B, float* pred, float* loss_sum,
                        float* dlogit, float gscale, pbx_stream_t stream);
/* fused CTR head: logit = tower[i*ld_tower] + extra[i] + bias[0]; pred, mean log-loss (written to
 * *loss_out) and dlogit = (pred - label)/B in one single-block pass; the backward scales dlogit by
 * *gloss and writes (accumulate: adds) its sum to *d_bias.  extra / bias / d_* may be NULL. */
int pbx_ctr_head_fwd(const float* tower, int ld_tower, const float* extra, const float* bias, const float* label,
                     int ld_label, int B, float* pred, float* loss_out, float* dlogit, pbx_stream_t stream);
/* dz_bf16 [B, ld_dz] / dzt_bf16 [>= B] (optional): bf16 copies of the tower gradient written straight
 * into the operands of the tower's last-layer backward GEMMs (saves a cast launch) */
int pbx_ctr_head_bwd(const float* dlogit, const float* gloss, int B, float* d_tower, float* d_extra, float* d_bias,
                     int accumulate, void* dz_bf16, int ld_dz, void* dzt_bf16, pbx_stream_t stream);
int pbx_fm_fwd(const float* x, int ldx, int B, int S, int E, int wcol, int vcol, int D, float* fm,
               pbx_stream_t stream);
int pbx_fm_bwd(const float* x, int ldx, const float* g, int B, int S, int E, int wcol, int vcol, int D, float* dx,
               int lddx, int accumulate, pbx_stream_t stream);
/* reference: operators/cvm_op.cu:29-75 */
int pbx_cvm_fwd(const float* x, int64_t rows, int W, int use_cvm, float* y, pbx_stream_t stream);
int pbx_cvm_bwd(const float* dy, const float* cvm, int64_t rows, int W, int use_cvm, float* dx,
                pbx_stream_t stream);
/* reference: operators/data_norm_op.cu:39-110 */
int pbx_data_norm_fwd(const float* x, int64_t rows, int C, const float* bsize, const float* bsum, const float* bsq,
                      float* y, float* means, float* scales, pbx_stream_t stream);
int pbx_data_norm_bwd(const float* x, const float* dy, int64_t rows, int C, const float* means,
                      const float* scales, float eps, float* dx, float* d_size, float* d_sum, float* d_sq,
                      pbx_stream_t stream);
/* reference: fleet/metrics.cc:115-136 (host bucketing) */
int pbx_auc_bucket(const float* pred, const float* label, const float* mask, int64_t n, int nbuckets,
                   double* table, double* sums, pbx_stream_t stream);
/* reference: fleet/box_wrapper.cu:1210-1224 */
int pbx_row_gather(const float* table, const int64_t* idx, int64_t n, int dim, float* out, pbx_stream_t stream);
/* one launch copies a mini-batch (keys, lod, dense) into the static buffers of a captured step;
 * key buffers 16-byte aligned */
int pbx_copy_batch(const void* keys_src, void* keys_dst, int64_t n_keys, const int32_t* lod_src, int32_t* lod_dst,
                   int64_t n_lod, const float* dense_src, float* dense_dst, int64_t n_dense, pbx_stream_t stream);
/* n floats device -> pinned host memory by a device-side store (no copy-engine hop) */
int pbx_store_host(const float* d_src, float* pinned_dst, int n, pbx_stream_t stream);
int pbx_l2_flush(void* buf, int64_t bytes, pbx_stream_t stream);

/* ---- device-side batch construction (csrc/cuda/batch_build.cu) ----------------
 * reference: data_feed.cu:1208-1384, data_feed.cc:3400-3522, :4896-4940 */
int pbx_batch_build(const uint64_t* d_rec_keys, const int32_t* d_rec_off, int B, int S, uint64_t* d_keys_out,
                    int32_t* d_lod_out, int32_t* d_scratch, pbx_stream_t stream);
/* the same launch also copies n_dense floats d_dense_src -> d_dense_dst (label + dense block) */
int pbx_batch_build_dense(const uint64_t* d_rec_keys, const int32_t* d_rec_off, int B, int S, uint64_t* d_keys_out,
                          int32_t* d_lod_out, int32_t* d_scratch, const float* d_dense_src, float* d_dense_dst,
                          int64_t n_dense, pbx_stream_t stream);
/* Allocation-free variant of pbx_batch_build (this entry point has no counterpart in the
 * reference).  For callers that cannot allocate device memory — a feed thread running beside
 * in-flight training kernels (cudaMalloc/cudaFree synchronise the device), a process whose HBM
 * is fully committed, or a host that forbids allocator calls from loaded code — the block-sum
 * scratch is a 2 KiB __device__ array inside the library (held by every GPU that loads it until
 * exit), so the call needs only its inputs and outputs and performs no allocation.
 * Limit: B * S <= 1,048,576; larger batches are rejected with an error.  The first call on a
 * device resolves the array's device address (cudaGetSymbolAddress) — make it before any stream
 * capture.  All calls on a device share the array (contents persist between calls; nothing in it
 * needs zeroing); calls that may overlap on the same GPU race and must be ordered by the caller. */
int pbx_batch_build_static(const uint64_t* d_rec_keys, const int32_t* d_rec_off, int B, int S,
                           uint64_t* d_keys_out, int32_t* d_lod_out, pbx_stream_t stream);

/* ---- symmetric memory + dense collectives (csrc/cuda/symm.cu) ------------------
 * reference: boxps_worker.cc:1191-1258 (SyncParam), c_allreduce_x_op.cc:114-151,
 * c_mixallgather_op.cc:221-328, transpiler/collective.py:240-315 */
int pbx_symm_create(void** out, int rank, int nranks, uint64_t bytes);
int pbx_symm_destroy(void* symm);
void* pbx_symm_ptr(void* symm);
int pbx_symm_ipc_handle(void* symm, void* out64);
int pbx_symm_connect(void* symm, const void* handles);
int pbx_symm_connect_local(void* symm, void* const* peers);
int pbx_symm_status(void* symm, int* status, pbx_stream_t stream);
int pbx_symm_collective(void* symm, int64_t off, int64_t n, float scale, int op, int root, int max_blocks,
                        pbx_stream_t stream);
int pbx_symm_collective_emu(void* const* symms, int nranks, int64_t off, int64_t n, float scale, int op,
                            int root, pbx_stream_t stream);

/* ---- remaining CTR operator family (csrc/cuda/ctr_ops.cu) ----
 * reference: rank_attention.cu.h:28-195 + rank_attention_op.cu:30-392; scaled_int8fc_op.cu:38-72;
 * fused/fused_seqpool_cvm_with_pcoc_op.cu:36-294; fused/fused_concat_op.cu:34-258;
 * fused/fused_seq_tensor_op.cu:14-147; cross_norm_hadamard.cu.h:44-225. */
int pbx_rank_attention_expand(const float* d_x, int ldx, const int32_t* d_rank_offset, int N, int C, int R,
                              void* d_out_bf16, int ldo, void* d_out_t_bf16, int ldt, pbx_stream_t stream);
int pbx_rank_attention_dx(const float* d_dA, int lda, const int32_t* d_rank_offset, int N, int C, int R, float* d_dx,
                          int ldx, pbx_stream_t stream);
int pbx_int8_quant_bf16(const float* d_x, int ldx, int M, int K, float expand, float clip, float int8_range,
                        void* d_out, int ldo, void* d_out_t, int ldt, pbx_stream_t stream);
int pbx_seqpool_pcoc_fwd(const float* d_x, int x_stride, const int32_t* d_lod, int B, int S, int H, int use_cvm,
                         int used, int mx, int need_filter, int quant_ratio, float pad_value, float show_coeff,
                         float clk_coeff, float threshold, float* d_out, int out_stride, pbx_stream_t stream);
int pbx_seqpool_pcoc_bwd(const float* d_dout, int out_stride, const float* d_cvm, int cvm_ld, const float* d_q,
                         const int32_t* d_lod, int B, int S, int H, int use_cvm, int used, int mx, float* d_dx,
                         int dx_stride, pbx_stream_t stream);
int pbx_fused_concat(const float* const* d_in, float* const* d_din, const int* ld, const int* width, int n_in, int B,
                     int start, int len, float* d_out_or_dout, int ldo, int backward, pbx_stream_t stream);
int pbx_seqpool_concat_fwd(const float* d_x0, int ld0, int off0, int d0, const float* d_x1, int ld1, int off1, int d1,
                           const int32_t* d_lod, int B, int S, float* d_out, int ldo, pbx_stream_t stream);
int pbx_seqpool_concat_bwd(const float* d_dout, int ldo, const int32_t* d_lod, int B, int S, float* d_dx0, int ld0,
                           int w0, int off0, int d0, float* d_dx1, int ld1, int w1, int off1, int d1,
                           pbx_stream_t stream);
int pbx_fused_seq_tensor(const float* d_x, const float* d_ad, int ins, int bc, int slot_num, int max_length, int dim,
                         int ad_slot_num, int ad_slot_offset, float* d_din, float* d_sess, float* d_side,
                         float* d_mask, pbx_stream_t stream);
int pbx_cross_hadamard_fwd(const float* d_x, int ldx, int B, int fields, int dim, float* d_y, int ldy,
                           pbx_stream_t stream);
int pbx_cross_hadamard_bwd(const float* d_x, int ldx, const float* d_gy, int ldy, int B, int fields, int dim,
                           float* d_dx, int lddx, pbx_stream_t stream);
/* data_norm(sync_stats): 3 stat vectors reduced over the ranks + decayed summary update in one launch
 * (reference: 3 ncclAllReduce + KernelUpdateParam, data_norm_op.cu:203-231, :90-110) */
int pbx_symm_stat_update(void* symm, int64_t off, int C, float scale, float decay, float* d_size, float* d_sum,
                         float* d_sqsum, pbx_stream_t stream);
int pbx_symm_stat_update_emu(void* const* symms, int nranks, int64_t off, int C, float scale, float decay,
                             float* const* summaries, pbx_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* PBX_CUDA_H_ */
