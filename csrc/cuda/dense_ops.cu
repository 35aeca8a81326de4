// Dense-side helper kernels of the CTR step (all plain HBM-bound SIMT — no GEMM shape):
//   cast+transpose feeding the tcgen05 GEMMs, column sums (bias grads), fused flat Adam
//   (+ bf16 weight refresh), FM interaction head, sigmoid+log-loss, CVM, data_norm,
//   AUC bucketing, replica-cache gather.
// Reference kernels replaced: cvm_op.cu:29-75, data_norm_op.cu:39-110, metrics.cc:115-136
// (host bucketing under a mutex), box_wrapper.cu:1210-1224, phi adam / log_loss /
// sigmoid / scale kernels.
#include "common.cuh"

#define PBX_CT_MAX 12

namespace {

// fp32 [M, K] (ld) -> bf16 [M, ldo] (+ optional bf16 transposed [K, ldt]); 32x32 smem tiles
__global__ void k_cast_transpose(const float* __restrict__ x, int ld, int M, int K, __nv_bfloat16* __restrict__ out,
                                 int ldo, __nv_bfloat16* __restrict__ out_t, int ldt) {
  __shared__ float tile[32][33];
  const int m0 = blockIdx.y * 32, k0 = blockIdx.x * 32;
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    int m = m0 + r, k = k0 + threadIdx.x;
    float v = (m < M && k < K) ? x[(int64_t)m * ld + k] : 0.f;
    tile[r][threadIdx.x] = v;
    if (out && m < M && k < K) out[(int64_t)m * ldo + k] = __float2bfloat16(v);
  }
  if (!out_t) return;
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    int k = k0 + r, m = m0 + threadIdx.x;
    if (k < K && m < M) out_t[(int64_t)k * ldt + m] = __float2bfloat16(tile[threadIdx.x][r]);
  }
}

// The same for up to PBX_CT_MAX tensors in ONE launch (all weights of a tower + its input): a
// training step is launch-bound (~1.7 us per dependent launch on B200), so five 4-us kernels
// cost more in gaps than in work.
struct CastTransposeBatch {
  int n;
  int tile_end[PBX_CT_MAX];        // exclusive prefix of 32x32 tiles
  const float* x[PBX_CT_MAX];
  __nv_bfloat16* out[PBX_CT_MAX];
  __nv_bfloat16* out_t[PBX_CT_MAX];
  int ld[PBX_CT_MAX], M[PBX_CT_MAX], K[PBX_CT_MAX], ldo[PBX_CT_MAX], ldt[PBX_CT_MAX];
};

__global__ void __launch_bounds__(256)
k_cast_transpose_multi(const __grid_constant__ CastTransposeBatch b) {
  // 64x64 tiles: coalesced fp32 reads into shared memory, then both outputs leave as packed
  // bf16x2 (128 bytes per warp row)
  __shared__ float tile[64][65];
  int t = 0;
  while (t < b.n - 1 && (int)blockIdx.x >= b.tile_end[t]) ++t;
  const int local = blockIdx.x - (t ? b.tile_end[t - 1] : 0);
  const int M = b.M[t], K = b.K[t], ld = b.ld[t], ldo = b.ldo[t], ldt = b.ldt[t];
  const int tk = (K + 63) / 64;
  const int m0 = (local / tk) * 64, k0 = (local % tk) * 64;
  const float* __restrict__ x = b.x[t];
  __nv_bfloat16* __restrict__ out = b.out[t];
  __nv_bfloat16* __restrict__ out_t = b.out_t[t];
  const int tx = threadIdx.x, ty = threadIdx.y;
#pragma unroll
  for (int r = ty; r < 64; r += 8) {
    const int m = m0 + r;
    const float* xr = x + (int64_t)m * ld + k0;
    tile[r][tx] = (m < M && k0 + tx < K) ? __ldg(xr + tx) : 0.f;
    tile[r][tx + 32] = (m < M && k0 + tx + 32 < K) ? __ldg(xr + tx + 32) : 0.f;
  }
  __syncthreads();
  if (out) {
    // pad columns [K, ldo) inside the tile are written as zeros (operands of a TMA-fed GEMM)
#pragma unroll
    for (int r = ty; r < 64; r += 8) {
      const int m = m0 + r, k = k0 + 2 * tx;
      if (m < M && k < ldo)
        *reinterpret_cast<__nv_bfloat162*>(out + (int64_t)m * ldo + k) =
            __floats2bfloat162_rn(tile[r][2 * tx], tile[r][2 * tx + 1]);
    }
  }
  if (out_t) {
#pragma unroll
    for (int r = ty; r < 64; r += 8) {
      const int k = k0 + r, m = m0 + 2 * tx;
      if (k < K && m < ldt)
        *reinterpret_cast<__nv_bfloat162*>(out_t + (int64_t)k * ldt + m) =
            __floats2bfloat162_rn(tile[2 * tx][r], tile[2 * tx + 1][r]);
    }
  }
}

// out[n] += sum_m x[m, n]   (x bf16 [M, ld])
__global__ void k_colsum_bf16(const __nv_bfloat16* __restrict__ x, int ld, int M, int N, float* __restrict__ out) {
  const int n = blockIdx.x * 32 + (threadIdx.x & 31);
  const int rlane = threadIdx.x >> 5;      // 8 row lanes
  const int rows_per_block = (M + gridDim.y - 1) / gridDim.y;
  const int mbeg = blockIdx.y * rows_per_block, mend = min(M, mbeg + rows_per_block);
  float acc = 0.f;
  if (n < N)
    for (int m = mbeg + rlane; m < mend; m += 8) acc += __bfloat162float(x[(int64_t)m * ld + n]);
  __shared__ float red[8][33];
  red[rlane][threadIdx.x & 31] = acc;
  __syncthreads();
  if (rlane == 0 && n < N) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += red[i][threadIdx.x & 31];
    atomicAdd(out + n, s);
  }
}

// flat Adam over every dense parameter at once (reference: one adam op per param).
// step_dev[0] holds the step count t (float), advanced by the kernel's first thread.
__global__ void k_adam_flat(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m,
                            float* __restrict__ v, int64_t n, float lr, float b1, float b2, float eps,
                            float weight_decay, float grad_scale, float* __restrict__ step_dev,
                            int zero_grad) {
  // step_dev[0] = step count, step_dev[1] = blocks finished (as float bits are avoided: counts
  // stay far below 2^24).  The last block to finish advances the step: every block has read it.
  const float t = step_dev[0] + 1.f;
  const float c1 = 1.f - powf(b1, t), c2 = 1.f - powf(b2, t);
  const float step_size = lr * sqrtf(c2) / c1;
  auto upd = [&](float& pi, float& gi_, float& mi_, float& vi_) {
    float gi = gi_ * grad_scale;
    if (weight_decay != 0.f) gi += weight_decay * pi;
    mi_ = b1 * mi_ + (1.f - b1) * gi;
    vi_ = b2 * vi_ + (1.f - b2) * gi * gi;
    pi = pi - step_size * mi_ / (sqrtf(vi_) + eps * sqrtf(c2));
    if (zero_grad) gi_ = 0.f;
  };
  const int64_t tid = blockIdx.x * (int64_t)blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
  const bool vec = ((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0);
  const int64_t n4 = vec ? n / 4 : 0;
  for (int64_t i = tid; i < n4; i += nth) {        // 16-byte accesses: the flat buffers are 16-byte aligned
    float4 P = reinterpret_cast<float4*>(p)[i], G = reinterpret_cast<float4*>(g)[i];
    float4 M = reinterpret_cast<float4*>(m)[i], V = reinterpret_cast<float4*>(v)[i];
    upd(P.x, G.x, M.x, V.x); upd(P.y, G.y, M.y, V.y); upd(P.z, G.z, M.z, V.z); upd(P.w, G.w, M.w, V.w);
    reinterpret_cast<float4*>(p)[i] = P; reinterpret_cast<float4*>(m)[i] = M; reinterpret_cast<float4*>(v)[i] = V;
    if (zero_grad) reinterpret_cast<float4*>(g)[i] = G;
  }
  for (int64_t i = n4 * 4 + tid; i < n; i += nth) {
    float pi = p[i], gi = g[i], mi = m[i], vi = v[i];
    upd(pi, gi, mi, vi);
    p[i] = pi; m[i] = mi; v[i] = vi;
    if (zero_grad) g[i] = gi;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const float done = atomicAdd(step_dev + 1, 1.f);
    if (done == (float)(gridDim.x - 1)) {
      step_dev[1] = 0.f;
      step_dev[0] = t;
    }
  }
}

__global__ void k_sgd_flat(float* __restrict__ p, float* __restrict__ g, int64_t n, float lr, float grad_scale,
                           int zero_grad) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    p[i] -= lr * grad_scale * g[i];
    if (zero_grad) g[i] = 0.f;
  }
}

// sigmoid + log-loss (mean) forward with the gradient produced in the same pass.
//   pred = sigmoid(logit); loss += -(y log p + (1-y) log(1-p)) / B; dlogit = (p - y) * gscale
__global__ void k_sigmoid_logloss(const float* __restrict__ logit, const float* __restrict__ label, int B,
                                  float* __restrict__ pred, float* __restrict__ loss_sum,
                                  float* __restrict__ dlogit, float gscale) {
  float local = 0.f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < B; i += gridDim.x * blockDim.x) {
    float z = logit[i], y = label[i];
    float p = 1.f / (1.f + __expf(-z));
    if (pred) pred[i] = p;
    // stable: max(z,0) - z*y + log(1+exp(-|z|))
    local += fmaxf(z, 0.f) - z * y + log1pf(__expf(-fabsf(z)));
    if (dlogit) dlogit[i] = (p - y) * gscale;
  }
  for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
  if ((threadIdx.x & 31) == 0 && loss_sum) atomicAdd(loss_sum, local / (float)B);
}

// CTR head in one pass: logit = tower + extra + bias; pred = sigmoid(logit); mean log-loss;
// dlogit = (pred - label) / B kept for the backward.  Replaces three adds, a copy and the loss op.
// ONE block: the loss is reduced in the block and written (not accumulated), so no zero-fill
// launch precedes it and the sum is deterministic.
__device__ __forceinline__ float pbx_block_sum_1024(float v, float* red) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
  if (threadIdx.x < 32)
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  return t;   // valid in thread 0
}

__global__ void __launch_bounds__(1024)
k_ctr_head_fwd(const float* __restrict__ tower, int ld_tower, const float* __restrict__ extra,
               const float* __restrict__ bias, const float* __restrict__ label, int ld_label, int B,
               float* __restrict__ pred, float* __restrict__ loss_out, float* __restrict__ dlogit) {
  __shared__ float red[32];
  float local = 0.f;
  const float bv = bias ? bias[0] : 0.f;
  const float invB = 1.f / (float)B;
  for (int i0 = threadIdx.x; i0 < B; i0 += 4 * blockDim.x) {
    float zt[4], ze[4], yy[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {               // all loads of four elements in flight together
      const int i = i0 + u * blockDim.x;
      const bool ok = i < B;
      zt[u] = ok ? __ldg(tower + (int64_t)i * ld_tower) : 0.f;
      ze[u] = (ok && extra) ? __ldg(extra + i) : 0.f;
      yy[u] = ok ? __ldg(label + (int64_t)i * ld_label) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * blockDim.x;
      if (i >= B) break;
      const float z = zt[u] + ze[u] + bv, y = yy[u];
      const float p = 1.f / (1.f + __expf(-z));
      pred[i] = p;
      local += fmaxf(z, 0.f) - z * y + log1pf(__expf(-fabsf(z)));
      dlogit[i] = (p - y) * invB;
    }
  }
  const float tot = pbx_block_sum_1024(local, red);
  if (threadIdx.x == 0) loss_out[0] = tot * invB;
}

// backward of the head: every logit term receives gloss * dlogit; the bias its sum (written, or
// added to an existing gradient when `accumulate`).
__global__ void __launch_bounds__(1024)
k_ctr_head_bwd(const float* __restrict__ dlogit, const float* __restrict__ gloss, int B,
               float* __restrict__ d_tower, float* __restrict__ d_extra, float* __restrict__ d_bias,
               int accumulate, __nv_bfloat16* __restrict__ dz, int ld_dz, __nv_bfloat16* __restrict__ dzt) {
  __shared__ float red[32];
  const float g = gloss[0];
  float local = 0.f;
  for (int i0 = threadIdx.x; i0 < B; i0 += 4 * blockDim.x) {
    float dv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * blockDim.x;
      dv[u] = i < B ? __ldg(dlogit + i) * g : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * blockDim.x;
      if (i >= B) break;
      if (d_tower) d_tower[i] = dv[u];
      if (d_extra) d_extra[i] = dv[u];
      if (dz) {            // bf16 operands of the tower's last-layer backward GEMMs, written in place
        const __nv_bfloat16 h = __float2bfloat16(dv[u]);
        dz[(int64_t)i * ld_dz] = h;
        dzt[i] = h;
      }
      local += dv[u];
    }
  }
  if (d_bias) {
    const float tot = pbx_block_sum_1024(local, red);
    if (threadIdx.x == 0) d_bias[0] = accumulate ? d_bias[0] + tot : tot;
  }
}

// DeepFM first/second-order head on pooled slot embeddings.
//   x: [B, ldx], slot s occupies columns [s*E, (s+1)*E); column `wcol` is the first-order
//   weight, columns [vcol, vcol+D) the factor vector.
//   fm[b] = sum_s w_s + 0.5 * sum_d ((sum_s v_sd)^2 - sum_s v_sd^2)
// One warp per row; the row (S*E floats) is read ONCE with coalesced loads into shared memory,
// per-dimension sums are formed by lanes [0, D) from the shared copy.
#define PBX_FM_WARPS 8
__global__ void __launch_bounds__(PBX_FM_WARPS * 32)
k_fm_fwd(const float* __restrict__ x, int ldx, int B, int S, int E, int wcol, int vcol, int D,
         float* __restrict__ fm) {
  extern __shared__ float s_fm[];
  const int W = S * E;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  float* srow = s_fm + wib * (W + 32);
  const int nw = blockDim.x >> 5;
  const int warps = gridDim.x * nw;
  for (int b = blockIdx.x * nw + wib; b < B; b += warps) {
    const float* row = x + (int64_t)b * ldx;
#pragma unroll 4
    for (int e = lane; e < W; e += 32) srow[e] = __ldg(row + e);
    __syncwarp();
    float acc = 0.f;
    for (int d = lane; d < D; d += 32) {
      float sv = 0.f, sq = 0.f;
      for (int s = 0; s < S; ++s) {
        const float v = srow[s * E + vcol + d];
        sv += v;
        sq += v * v;
      }
      acc += 0.5f * (sv * sv - sq);
    }
    for (int s = lane; s < S; s += 32) acc += srow[s * E + wcol];
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) fm[b] = acc;
    __syncwarp();
  }
}
// dx[b, s*E + wcol] += g[b];  dx[b, s*E+vcol+d] += g[b] * (sum_s' v_s'd - v_sd); other columns of
// the slot block get 0 (or stay, when accumulating).  Row in and gradient row out are coalesced.
__global__ void __launch_bounds__(PBX_FM_WARPS * 32)
k_fm_bwd(const float* __restrict__ x, int ldx, const float* __restrict__ g, int B, int S, int E,
         int wcol, int vcol, int D, float* __restrict__ dx, int lddx, int accumulate) {
  extern __shared__ float s_fm[];
  const int W = S * E;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  float* srow = s_fm + wib * (W + 32 + D);
  float* ssum = srow + W + 32;
  const int nw = blockDim.x >> 5;
  const int warps = gridDim.x * nw;
  for (int b = blockIdx.x * nw + wib; b < B; b += warps) {
    const float* row = x + (int64_t)b * ldx;
    float* drow = dx + (int64_t)b * lddx;
#pragma unroll 4
    for (int e = lane; e < W; e += 32) srow[e] = __ldg(row + e);
    const float gb = __ldg(g + b);
    __syncwarp();
    for (int d = lane; d < D; d += 32) {
      float sv = 0.f;
      for (int s = 0; s < S; ++s) sv += srow[s * E + vcol + d];
      ssum[d] = sv;
    }
    __syncwarp();
    for (int e0 = lane; e0 < W; e0 += 128) {          // four independent read-modify-writes in flight
      float gv[4], old[4];
      bool on[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int e = e0 + 32 * u;
        on[u] = e < W;
        gv[u] = 0.f;
        if (on[u]) {
          const int c = e % E;
          if (c == wcol) gv[u] = gb;
          else if (c >= vcol && c < vcol + D) gv[u] = gb * (ssum[c - vcol] - srow[e]);
          else if (accumulate) on[u] = false;
        }
        old[u] = (on[u] && accumulate) ? drow[e] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (on[u]) drow[e0 + 32 * u] = old[u] + gv[u];
    }
    __syncwarp();
  }
}

// standalone CVM transform (cvm_op.cu:29-75): use_cvm: y = [log(x0+1), log(x1+1)-log(x0+1), x2..]
// else y = x[2:]
__global__ void k_cvm_fwd(const float* __restrict__ x, int64_t rows, int W, int use_cvm, float* __restrict__ y) {
  const int Wo = use_cvm ? W : W - 2;
  const int64_t total = rows * Wo;
  for (int64_t it = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; it < total;
       it += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(it % Wo);
    const int64_t r = it / Wo;
    const float* xr = x + r * W;
    float v;
    if (use_cvm) v = c == 0 ? logf(xr[0] + 1.f) : (c == 1 ? logf(xr[1] + 1.f) - logf(xr[0] + 1.f) : xr[c]);
    else v = xr[c + 2];
    y[it] = v;
  }
}
// dx = [cvm0, cvm1, dy...]  (cvm_op.cu:50-75)
__global__ void k_cvm_bwd(const float* __restrict__ dy, const float* __restrict__ cvm, int64_t rows, int W,
                          int use_cvm, float* __restrict__ dx) {
  const int Wo = use_cvm ? W : W - 2;
  const int64_t total = rows * W;
  for (int64_t it = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; it < total;
       it += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(it % W);
    const int64_t r = it / W;
    dx[it] = c < 2 ? cvm[r * 2 + c] : dy[r * Wo + (use_cvm ? c : c - 2)];
  }
}

// data_norm forward: y = (x - sum/size) * sqrt(size/sqsum)   (data_norm_op.cu:39-60)
__global__ void k_data_norm_fwd(const float* __restrict__ x, int64_t rows, int C, const float* __restrict__ bsize,
                                const float* __restrict__ bsum, const float* __restrict__ bsq,
                                float* __restrict__ y, float* __restrict__ means, float* __restrict__ scales) {
  const int64_t total = rows * C;
  for (int64_t it = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; it < total;
       it += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(it % C);
    const float mean = bsum[c] / bsize[c];
    const float scale = sqrtf(bsize[c] / bsq[c]);
    y[it] = (x[it] - mean) * scale;
    if (it < C) { means[c] = mean; scales[c] = scale; }
  }
}
// backward: dx = dy * scale; per-column batch statistics for the three summaries
//   d_size = 1, d_sum = mean(x), d_sqsum = mean((x-mean)^2) + eps   (GPU kernel, data_norm_op.cu:67-88)
__global__ void k_data_norm_bwd(const float* __restrict__ x, const float* __restrict__ dy, int64_t rows, int C,
                                const float* __restrict__ means, const float* __restrict__ scales, float eps,
                                float* __restrict__ dx, float* __restrict__ d_size, float* __restrict__ d_sum,
                                float* __restrict__ d_sq) {
  // grid.x over column tiles of 32, grid.y over row chunks
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  const int rlane = threadIdx.x >> 5;
  const int64_t rpb = (rows + gridDim.y - 1) / gridDim.y;
  const int64_t r0 = blockIdx.y * rpb, r1 = min(rows, r0 + rpb);
  float s = 0.f, q = 0.f;
  if (c < C) {
    const float mean = means[c], scale = scales[c];
    for (int64_t r = r0 + rlane; r < r1; r += 8) {
      float xv = x[r * C + c];
      if (dx) dx[r * C + c] = dy[r * C + c] * scale;
      s += xv;
      q += (xv - mean) * (xv - mean);
    }
  }
  __shared__ float rs[8][33], rq[8][33];
  rs[rlane][threadIdx.x & 31] = s;
  rq[rlane][threadIdx.x & 31] = q;
  __syncthreads();
  if (rlane == 0 && c < C) {
    float ts = 0.f, tq = 0.f;
    for (int i = 0; i < 8; ++i) { ts += rs[i][threadIdx.x & 31]; tq += rq[i][threadIdx.x & 31]; }
    const float inv = 1.f / (float)rows;
    atomicAdd(d_sum + c, ts * inv);
    atomicAdd(d_sq + c, tq * inv + (blockIdx.y == 0 ? eps : 0.f));
    if (blockIdx.y == 0) d_size[c] = 1.f;
  }
}

// AUC bucketing on device (W9): table[0][bucket] += (1 - label), table[1][bucket] += label,
// plus running sums {abs_err, sq_err, pred, label, count} in fp64.
// Reference does a D2H copy and buckets on the CPU under a mutex (metrics.cc:115-136).
__global__ void k_auc_bucket(const float* __restrict__ pred, const float* __restrict__ label,
                             const float* __restrict__ mask, int64_t n, int nbuckets,
                             double* __restrict__ table /* [2, nbuckets] */, double* __restrict__ sums /* [5] */) {
  double l_abs = 0, l_sq = 0, l_pred = 0, l_lab = 0, l_cnt = 0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    if (mask && mask[i] == 0.f) continue;
    float p = pred[i], y = label[i];
    p = fminf(fmaxf(p, 0.f), 1.f);
    int bkt = min((int)(p * (float)nbuckets), nbuckets - 1);
    atomicAdd(table + bkt, (double)(1.f - y));
    atomicAdd(table + nbuckets + bkt, (double)y);
    l_abs += fabs((double)p - y);
    l_sq += ((double)p - y) * ((double)p - y);
    l_pred += p;
    l_lab += y;
    l_cnt += 1;
  }
  for (int o = 16; o > 0; o >>= 1) {
    l_abs += __shfl_xor_sync(0xffffffffu, l_abs, o);
    l_sq += __shfl_xor_sync(0xffffffffu, l_sq, o);
    l_pred += __shfl_xor_sync(0xffffffffu, l_pred, o);
    l_lab += __shfl_xor_sync(0xffffffffu, l_lab, o);
    l_cnt += __shfl_xor_sync(0xffffffffu, l_cnt, o);
  }
  if ((threadIdx.x & 31) == 0 && l_cnt > 0) {
    atomicAdd(sums + 0, l_abs);
    atomicAdd(sums + 1, l_sq);
    atomicAdd(sums + 2, l_pred);
    atomicAdd(sums + 3, l_lab);
    atomicAdd(sums + 4, l_cnt);
  }
}

// replica cache gather (box_wrapper.cu:1210-1224): out[i, :] = table[idx[i], :]
__global__ void k_row_gather(const float* __restrict__ table, const int64_t* __restrict__ idx, int64_t n, int dim,
                             float* __restrict__ out) {
  const int64_t total = n * dim;
  for (int64_t it = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; it < total;
       it += (int64_t)gridDim.x * blockDim.x)
    out[it] = table[idx[it / dim] * dim + it % dim];
}

// n floats device -> pinned (UVA-mapped) host memory by a store from the device: a step's loss
// reaches the host without a copy-engine hop between two graph launches (~10 us on this box)
__global__ void k_store_host(const float* __restrict__ src, float* __restrict__ dst_host, int n) {
  for (int i = threadIdx.x; i < n; i += blockDim.x) dst_host[i] = src[i];
  __threadfence_system();
}

// keys + lod + dense of one mini-batch into the static buffers of the captured step: ONE launch
// instead of three device-to-device copies
__global__ void k_copy_batch(const int4* __restrict__ k_src, int4* __restrict__ k_dst, int64_t k16,
                             const int* __restrict__ l_src, int* __restrict__ l_dst, int64_t nl,
                             const float* __restrict__ d_src, float* __restrict__ d_dst, int64_t nd) {
  const int64_t tid = blockIdx.x * (int64_t)blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = tid; i < (k16 >> 1); i += nth) k_dst[i] = k_src[i];      // k16 = number of 64-bit ids here
  if ((k16 & 1) && tid == 0)
    reinterpret_cast<long long*>(k_dst)[k16 - 1] = reinterpret_cast<const long long*>(k_src)[k16 - 1];
  for (int64_t i = tid; i < nl; i += nth) l_dst[i] = l_src[i];
  for (int64_t i = tid; i < nd; i += nth) d_dst[i] = d_src[i];
}

// L2 flush helper for benches: writes a buffer larger than L2
__global__ void k_fill_u32(uint32_t* p, int64_t n, uint32_t v) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    p[i] = v;
}

}  // namespace

// warps per block / grid / dynamic shared memory of the FM kernels: as many warps (<= 8) as fit
// 200 KB of per-warp row copies
static int pbx_fm_geometry(const void* fn, int words_per_warp, int B, int* nw, int* blocks, size_t* smem) {
  int w = PBX_FM_WARPS;
  while (w > 1 && (size_t)w * words_per_warp * sizeof(float) > 200 * 1024) w >>= 1;
  *smem = (size_t)w * words_per_warp * sizeof(float);
  PBX_REQUIRE(*smem <= 200 * 1024, "fm: a row of %d floats does not fit in shared memory", words_per_warp);
  if (*smem > 48 * 1024)
    PBX_CUDA_CHECK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)*smem));
  *nw = w;
  int b = (B + w - 1) / w;
  if (b > 148 * 8) b = 148 * 8;
  *blocks = b;
  return 0;
}

extern "C" {

int pbx_cast_transpose(const float* x, int ld, int M, int K, void* out_bf16, int ldo, void* out_t, int ldt,
                       cudaStream_t st) {
  if (M <= 0 || K <= 0) return 0;
  dim3 grid((K + 31) / 32, (M + 31) / 32);
  k_cast_transpose<<<grid, dim3(32, 8), 0, st>>>(x, ld, M, K, (__nv_bfloat16*)out_bf16, ldo, (__nv_bfloat16*)out_t,
                                                 ldt);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

// n <= PBX_CT_MAX tensors; arrays of length n (out / out_t entries may be null)
int pbx_cast_transpose_multi(int n, const float* const* x, const int* ld, const int* M, const int* K,
                             void* const* out, const int* ldo, void* const* out_t, const int* ldt, cudaStream_t st) {
  PBX_REQUIRE(n >= 1 && n <= PBX_CT_MAX, "cast_transpose_multi: %d tensors (max %d)", n, PBX_CT_MAX);
  CastTransposeBatch b;
  memset(&b, 0, sizeof(b));
  b.n = n;
  int tiles = 0;
  for (int i = 0; i < n; ++i) {
    PBX_REQUIRE(M[i] > 0 && K[i] > 0, "cast_transpose_multi: empty tensor %d", i);
    tiles += ((M[i] + 63) / 64) * ((K[i] + 63) / 64);
    b.tile_end[i] = tiles;
    b.x[i] = x[i]; b.out[i] = (__nv_bfloat16*)out[i]; b.out_t[i] = (__nv_bfloat16*)out_t[i];
    PBX_REQUIRE((!out[i] || (ldo[i] & 1) == 0) && (!out_t[i] || (ldt[i] & 1) == 0),
                "cast_transpose_multi: output leading dims must be even (tensor %d)", i);
    b.ld[i] = ld[i]; b.M[i] = M[i]; b.K[i] = K[i]; b.ldo[i] = ldo[i]; b.ldt[i] = ldt[i];
  }
  k_cast_transpose_multi<<<tiles, dim3(32, 8), 0, st>>>(b);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_ctr_head_fwd(const float* tower, int ld_tower, const float* extra, const float* bias, const float* label,
                     int ld_label, int B, float* pred, float* loss_sum, float* dlogit, cudaStream_t st) {
  if (B <= 0) return 0;
  k_ctr_head_fwd<<<1, 1024, 0, st>>>(tower, ld_tower, extra, bias, label, ld_label, B, pred, loss_sum, dlogit);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_ctr_head_bwd(const float* dlogit, const float* gloss, int B, float* d_tower, float* d_extra, float* d_bias,
                     int accumulate, void* dz_bf16, int ld_dz, void* dzt_bf16, cudaStream_t st) {
  if (B <= 0) return 0;
  k_ctr_head_bwd<<<1, 1024, 0, st>>>(dlogit, gloss, B, d_tower, d_extra, d_bias, accumulate,
                                     (__nv_bfloat16*)dz_bf16, ld_dz, (__nv_bfloat16*)dzt_bf16);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_colsum_bf16(const void* x, int ld, int M, int N, float* out, cudaStream_t st) {
  if (M <= 0 || N <= 0) return 0;
  int gy = (M + 63) / 64;   // 64 rows per block: short serial loops, plenty of blocks
  if (gy > 128) gy = 128;
  k_colsum_bf16<<<dim3((N + 31) / 32, gy), 256, 0, st>>>((const __nv_bfloat16*)x, ld, M, N, out);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_adam_flat(float* p, float* g, float* m, float* v, int64_t n, float lr, float b1, float b2, float eps,
                  float weight_decay, float grad_scale, float* step_dev, int zero_grad, cudaStream_t st) {
  if (n <= 0) return 0;
  // (at most two blocks per SM: every block ends with one atomic on the step word)
  k_adam_flat<<<pbx_grid_for((n + 3) / 4, 256, 2), 256, 0, st>>>(p, g, m, v, n, lr, b1, b2, eps, weight_decay, grad_scale,
                                                                 step_dev, zero_grad);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_sgd_flat(float* p, float* g, int64_t n, float lr, float grad_scale, int zero_grad, cudaStream_t st) {
  if (n <= 0) return 0;
  k_sgd_flat<<<pbx_grid_for(n, 256, 8), 256, 0, st>>>(p, g, n, lr, grad_scale, zero_grad);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_sigmoid_logloss(const float* logit, const float* label, int B, float* pred, float* loss_sum, float* dlogit,
                        float gscale, cudaStream_t st) {
  if (B <= 0) return 0;
  k_sigmoid_logloss<<<pbx_grid_for(B, 256, 1), 256, 0, st>>>(logit, label, B, pred, loss_sum, dlogit, gscale);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_fm_fwd(const float* x, int ldx, int B, int S, int E, int wcol, int vcol, int D, float* fm, cudaStream_t st) {
  if (B <= 0) return 0;
  int nw, blocks;
  size_t smem;
  int rc = pbx_fm_geometry((const void*)k_fm_fwd, S * E + 32, B, &nw, &blocks, &smem);
  if (rc) return rc;
  k_fm_fwd<<<blocks, nw * 32, smem, st>>>(x, ldx, B, S, E, wcol, vcol, D, fm);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_fm_bwd(const float* x, int ldx, const float* g, int B, int S, int E, int wcol, int vcol, int D, float* dx,
               int lddx, int accumulate, cudaStream_t st) {
  if (B <= 0) return 0;
  int nw, blocks;
  size_t smem;
  int rc = pbx_fm_geometry((const void*)k_fm_bwd, S * E + 32 + D, B, &nw, &blocks, &smem);
  if (rc) return rc;
  k_fm_bwd<<<blocks, nw * 32, smem, st>>>(x, ldx, g, B, S, E, wcol, vcol, D, dx, lddx, accumulate);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_cvm_fwd(const float* x, int64_t rows, int W, int use_cvm, float* y, cudaStream_t st) {
  int64_t total = rows * (use_cvm ? W : W - 2);
  if (total <= 0) return 0;
  k_cvm_fwd<<<pbx_grid_for(total, 256, 8), 256, 0, st>>>(x, rows, W, use_cvm, y);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_cvm_bwd(const float* dy, const float* cvm, int64_t rows, int W, int use_cvm, float* dx, cudaStream_t st) {
  if (rows * W <= 0) return 0;
  k_cvm_bwd<<<pbx_grid_for(rows * W, 256, 8), 256, 0, st>>>(dy, cvm, rows, W, use_cvm, dx);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_data_norm_fwd(const float* x, int64_t rows, int C, const float* bsize, const float* bsum, const float* bsq,
                      float* y, float* means, float* scales, cudaStream_t st) {
  if (rows * C <= 0) return 0;
  k_data_norm_fwd<<<pbx_grid_for(rows * C, 256, 8), 256, 0, st>>>(x, rows, C, bsize, bsum, bsq, y, means, scales);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

// d_size/d_sum/d_sq must be zeroed by the caller (they are accumulated with atomics)
int pbx_data_norm_bwd(const float* x, const float* dy, int64_t rows, int C, const float* means, const float* scales,
                      float eps, float* dx, float* d_size, float* d_sum, float* d_sq, cudaStream_t st) {
  if (rows * C <= 0) return 0;
  int gy = (int)((rows + 255) / 256);
  if (gy > 64) gy = 64;
  if (gy < 1) gy = 1;
  k_data_norm_bwd<<<dim3((C + 31) / 32, gy), 256, 0, st>>>(x, dy, rows, C, means, scales, eps, dx, d_size, d_sum,
                                                           d_sq);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_auc_bucket(const float* pred, const float* label, const float* mask, int64_t n, int nbuckets, double* table,
                   double* sums, cudaStream_t st) {
  if (n <= 0) return 0;
  k_auc_bucket<<<pbx_grid_for(n, 256, 2), 256, 0, st>>>(pred, label, mask, n, nbuckets, table, sums);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_row_gather(const float* table, const int64_t* idx, int64_t n, int dim, float* out, cudaStream_t st) {
  if (n * dim <= 0) return 0;
  k_row_gather<<<pbx_grid_for(n * dim, 256, 8), 256, 0, st>>>(table, idx, n, dim, out);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

// dst_host must be pinned host memory (device-accessible under unified addressing)
int pbx_store_host(const float* d_src, float* pinned_dst, int n, cudaStream_t st) {
  if (n <= 0) return 0;
  k_store_host<<<1, 32, 0, st>>>(d_src, pinned_dst, n);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

// keys: n_keys 64-bit ids (both buffers 16-byte aligned); lod: n_lod int32;
// dense: n_dense floats (src/dst may be NULL with n_dense = 0)
int pbx_copy_batch(const void* keys_src, void* keys_dst, int64_t n_keys, const int32_t* lod_src, int32_t* lod_dst,
                   int64_t n_lod, const float* dense_src, float* dense_dst, int64_t n_dense, cudaStream_t st) {
  PBX_REQUIRE((((uintptr_t)keys_src | (uintptr_t)keys_dst) & 15) == 0, "copy_batch: key buffers must be 16-byte aligned");
  const int64_t k16 = n_keys;
  int64_t work = n_keys / 2 + 1 > n_lod ? n_keys / 2 + 1 : n_lod;
  if (n_dense > work) work = n_dense;
  if (work <= 0) return 0;
  k_copy_batch<<<pbx_grid_for(work, 256, 4), 256, 0, st>>>((const int4*)keys_src, (int4*)keys_dst, k16, lod_src, lod_dst,
                                                            n_lod, dense_src, dense_dst, n_dense);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

int pbx_l2_flush(void* buf, int64_t bytes, cudaStream_t st) {
  k_fill_u32<<<pbx_grid_for(bytes / 4, 256, 8), 256, 0, st>>>((uint32_t*)buf, bytes / 4, 0u);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

}  // extern "C"
