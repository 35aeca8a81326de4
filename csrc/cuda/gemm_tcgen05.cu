// Dense-tower GEMM on the 5th-gen tensor cores (W7 of SURVEY.md §8):
//
//     C[M,N] = epilogue( alpha * A[M,K] . B[N,K]^T )        A, B bf16, K-contiguous
//
// * operands staged by TMA (cp.async.bulk.tensor, 128B swizzle) through a 4-stage
//   mbarrier ring, multiplied by tcgen05.mma (cta_group::1, M=128, N=BN, K=16) issued
//   by one elected thread, accumulated in TMEM, read back with tcgen05.ld;
// * warp roles: warp0 = TMA producer, warp1 = TMEM owner + MMA issuer, warps2-5 =
//   epilogue (one TMEM lane quadrant each);
// * fused epilogue: + bias[N], ReLU, * (mask > 0) (ReLU backward), bf16 row-major
//   store, bf16 *transposed* store (feeds the next wgrad without a transpose pass),
//   fp32 store, or fp32 atomic accumulate (split-K wgrad).
//
// All three GEMMs of a dense layer (fprop, dgrad, wgrad) are this one "TN" kernel:
// transposed bf16 copies of activations / weights are produced by the epilogues.
// Replaces the reference's cuBLAS calls (mul/fc phi kernels; scaled_fc_op.cu:214;
// batch_fc_op.cu:243) — there is no tensor-core code in the reference.
#include "common.cuh"
#include "pbx_types.h"
#include <cstdlib>
#include <cuda.h>
#include <cudaTypedefs.h>

namespace {

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = 128 bytes = one swizzle row
constexpr int GEMM_THREADS = 320;  // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue
constexpr int EPI_THREADS = 256;

struct GemmEpilogue {
  const float* bias;          // [N] or null
  const __nv_bfloat16* mask;  // [M, ld_mask] or null: out *= (mask > 0)
  __nv_bfloat16* out_bf16;    // [M, ld_out] or null
  __nv_bfloat16* out_t;       // [N, ld_t] or null  (transposed copy)
  float* out_f32;             // [M, ld_f32] or null
  int ld_mask, ld_out, ld_t, ld_f32;
  int relu;
  int accumulate;             // atomicAdd into out_f32 (split-K)
  float alpha;
  float* col_out;             // optional: output column `col_split` goes to col_out[row] instead of
  int col_split;              // out_f32 (bias gradient from a ones row appended to the B operand)
  unsigned long long* ts;     // optional debug timeline of CTA (0,0,0): 8 globaltimer stamps
};
#define GEMM_TS(i) do { if (ep.ts && (bx | by | bz) == 0) ep.ts[i] = pbx_globaltimer(); } while (0)

// ---------------- PTX wrappers ----------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t addr = smem_u32(bar);
  uint32_t done = 0;
  // bounded: a mis-programmed pipeline must trap, not hang the GPU
  for (uint32_t spin = 0; spin < (1u << 28); ++spin) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
    if (done) return;
  }
  __trap();
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)),
               "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major, 128B-swizzled smem tile descriptor (cute::UMMA::SmemDescriptor, sm_100):
//   [0,14) start>>4 | [16,30) LBO>>4 (unused for swizzled K-major) | [32,46) SBO>>4 = 1024B
//   | [46,48) version = 1 | [61,64) layout = 2 (SWIZZLE_128B)
__device__ __forceinline__ uint64_t make_smem_desc(const void* smem_tile) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_u32(smem_tile) & 0x3FFFF) >> 4);
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// instruction descriptor (cute::UMMA::InstrDescriptor): c=F32 [4,6)=1, a=BF16 [7,10)=1,
// b=BF16 [10,13)=1, a/b K-major, N>>3 at [17,23), M>>4 at [24,29)
__host__ __device__ constexpr uint32_t make_idesc(int m, int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

template <int BN, int STG = 0>
struct SmemLayout {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  // deep pipelines: the MLP-tower GEMMs have 4-8 k-blocks and start cold, so every k-block that
  // is in flight at once removes one exposed DRAM round trip from the main loop
  // STG = 0: as deep as 196 KB allows (one CTA per SM); STG > 0: that many stages (the grouped
  // kernel uses ~96 KB so that two CTAs share an SM and one's epilogue overlaps the other's loads)
  static constexpr int STAGES = STG > 0 ? STG : ((196 * 1024 / STAGE_BYTES) > 8 ? 8 : (196 * 1024 / STAGE_BYTES));
  static constexpr int BAR_BYTES = 256;
  static constexpr int TOTAL = STAGES * STAGE_BYTES + 1024 /*align slack*/ + BAR_BYTES + BN * 4 /*bias*/;
  // the epilogue reuses the stage memory for a row-major and a transposed bf16 staging tile
  static_assert(BM * (BN + 8) * 2 + BN * (BM + 32) * 2 <= STAGES * STAGE_BYTES, "staging tiles do not fit");
};

// One output tile (bx, by) of split bz of one GEMM problem; the whole CTA participates.
template <int BN, int STG = 0>
__device__ __forceinline__ void gemm_tile(const CUtensorMap* pmap_a, const CUtensorMap* pmap_b, int M, int N, int K,
                                          int kb_per_split, const GemmEpilogue& ep, int bx, int by, int bz) {
  extern __shared__ uint8_t smem_raw[];
  using L = SmemLayout<BN, STG>;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  constexpr int STAGES = L::STAGES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * L::STAGE_BYTES);
  float* s_bias = reinterpret_cast<float*>(smem + STAGES * L::STAGE_BYTES + L::BAR_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int m0 = bx * BM;
  const int n0 = by * BN;
  const int num_kb_total = (K + BK - 1) / BK;
  const int kb_begin = bz * kb_per_split;
  const int kb_end = min(kb_begin + kb_per_split, num_kb_total);
  const int num_kb = kb_end - kb_begin;
  if (threadIdx.x == 0) GEMM_TS(0);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(pmap_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(pmap_b) : "memory");
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar + s, 1);
      mbar_init(empty_bar + s, 1);
    }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(tmem_slot, BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) GEMM_TS(1);

  if (num_kb > 0) {
    if (warp == 0) {
      // ===== TMA producer =====
      if (lane == 0) {
        for (int i = 0; i < num_kb; ++i) {
          const int s = i % STAGES;
          const uint32_t ph = (uint32_t)(i / STAGES) & 1u;
          mbar_wait(empty_bar + s, ph ^ 1u);
          uint8_t* sa = smem + s * L::STAGE_BYTES;
          uint8_t* sb = sa + L::A_BYTES;
          mbar_expect_tx(full_bar + s, (uint32_t)L::STAGE_BYTES);
          const int kc = (kb_begin + i) * BK;
          tma_load_2d(sa, pmap_a, full_bar + s, kc, m0);
          tma_load_2d(sb, pmap_b, full_bar + s, kc, n0);
        }
      }
    } else if (warp == 1) {
      // ===== MMA issuer =====
      if (lane == 0) {
        constexpr uint32_t idesc = make_idesc(BM, BN);
        for (int i = 0; i < num_kb; ++i) {
          const int s = i % STAGES;
          const uint32_t ph = (uint32_t)(i / STAGES) & 1u;
          mbar_wait(full_bar + s, ph);
          tc_fence_after();
          if (i == 0) GEMM_TS(2);
          uint8_t* sa = smem + s * L::STAGE_BYTES;
          uint8_t* sb = sa + L::A_BYTES;
          const uint64_t da = make_smem_desc(sa);
          const uint64_t db = make_smem_desc(sb);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            // advance 16 bf16 = 32 bytes along K inside the swizzle atom: +2 in the (>>4) address field
            umma_bf16(tmem_base, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (i | k) ? 1u : 0u);
          }
          umma_commit(empty_bar + s);  // frees the smem stage when these MMAs retire
        }
        umma_commit(tmem_full_bar);    // accumulator complete
        GEMM_TS(3);
      }
      __syncwarp();
    } else {
      // ===== epilogue: warps 2..9.  TMEM lane quadrant = warp % 4 (hardware rule); the two warps
      // of a quadrant split the tile's columns in halves.  One warp per scheduler is latency-bound
      // (the in-kernel timeline showed the 4-warp epilogue at 2x the main loop), hence eight. =====
      // bf16 outputs are staged through shared memory (the pipeline stages are free once the
      // accumulator is complete) so that both the row-major tile and its transpose leave the SM
      // as 16-byte, fully coalesced stores.  (Writing the transpose straight from registers
      // costs one 2-byte store instruction per element: ncu showed the epilogue at ~10x the
      // tensor-pipe time.)
      const int q = warp & 3;
      const int rl = q * 32 + lane;          // row inside the tile
      const int row = m0 + rl;
      const int half = (warp - 2) >> 2;      // which half of the columns
      constexpr int NCH = BN >= 64 ? BN / 64 : 1;            // 32-column chunks per warp
      const bool active = BN >= 64 || half == 0;
      constexpr int LDC = BN + 8, LDT = BM + 32;             // LDT/2 = 16 (mod 32): paired-row stores of
                                                             // neighbouring columns hit disjoint banks
      __nv_bfloat16* sC = reinterpret_cast<__nv_bfloat16*>(smem);
      __nv_bfloat16* sT = sC + BM * LDC;
      const bool stage = ep.out_bf16 != nullptr || ep.out_t != nullptr;
      const int et = threadIdx.x - 64;       // 0..255 over the eight epilogue warps
      // While the main loop runs these warps are idle: fetch everything the epilogue reads from
      // global memory NOW (bias -> shared, ReLU mask of this thread's row -> registers), so no
      // cold DRAM round trip sits between the accumulator and the stores.
      if (ep.bias)
        for (int i = et; i < BN; i += EPI_THREADS) s_bias[i] = (n0 + i < N) ? __ldg(ep.bias + n0 + i) : 0.f;
      constexpr bool kMaskRegs = BN <= 128;
      uint4 mk[kMaskRegs ? NCH * 4 : 1];
      const int cbase = BN >= 64 ? half * (BN / 2) : 0;
      const bool mask_pre = kMaskRegs && ep.mask && row < M && n0 + BN <= N && (ep.ld_mask & 7) == 0 &&
                            ((uintptr_t)ep.mask & 15) == 0 && (n0 & 7) == 0;
      if (kMaskRegs && mask_pre) {
        const uint4* mrow4 = reinterpret_cast<const uint4*>(ep.mask + (int64_t)row * ep.ld_mask + n0 + cbase);
#pragma unroll
        for (int j = 0; j < NCH * 4; ++j) mk[j] = __ldg(mrow4 + j);
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");   // s_bias visible to the eight epilogue warps
      mbar_wait(tmem_full_bar, 0);
      tc_fence_after();
      if (threadIdx.x == 64) GEMM_TS(4);
#pragma unroll
      for (int cc = 0; cc < NCH; ++cc) {
        const int c0 = cbase + cc * 32;
        if (!active) break;
        uint32_t r[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, r);
        if (n0 + c0 >= N) continue;
        float v[32];
        if (ep.bias) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 bq = *reinterpret_cast<const float4*>(s_bias + c0 + 4 * j);   // warp-uniform: broadcast
            v[4 * j] = __uint_as_float(r[4 * j]) * ep.alpha + bq.x;
            v[4 * j + 1] = __uint_as_float(r[4 * j + 1]) * ep.alpha + bq.y;
            v[4 * j + 2] = __uint_as_float(r[4 * j + 2]) * ep.alpha + bq.z;
            v[4 * j + 3] = __uint_as_float(r[4 * j + 3]) * ep.alpha + bq.w;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * ep.alpha;
        }
        if (ep.relu) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
        }
        if (kMaskRegs && mask_pre) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const __nv_bfloat16* mb = reinterpret_cast<const __nv_bfloat16*>(&mk[kMaskRegs ? cc * 4 + j : 0]);
#pragma unroll
            for (int e = 0; e < 8; ++e)
              if (!(__bfloat162float(mb[e]) > 0.f)) v[8 * j + e] = 0.f;
          }
        } else if (ep.mask && row < M) {
          const __nv_bfloat16* mrow = ep.mask + (int64_t)row * ep.ld_mask + n0 + c0;
          if (n0 + c0 + 32 <= N && (ep.ld_mask & 7) == 0) {
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              const uint4 mq = __ldg(reinterpret_cast<const uint4*>(mrow + j));
              const __nv_bfloat16* mb = reinterpret_cast<const __nv_bfloat16*>(&mq);
#pragma unroll
              for (int e = 0; e < 8; ++e)
                if (!(__bfloat162float(mb[e]) > 0.f)) v[j + e] = 0.f;
            }
          } else {
            for (int j = 0; j < 32; ++j)
              if (n0 + c0 + j < N && !(__bfloat162float(mrow[j]) > 0.f)) v[j] = 0.f;
          }
        }
        if (stage) {
          if (ep.out_bf16) {
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              __nv_bfloat162 p0 = __floats2bfloat162_rn(v[j], v[j + 1]);
              __nv_bfloat162 p1 = __floats2bfloat162_rn(v[j + 2], v[j + 3]);
              __nv_bfloat162 p2 = __floats2bfloat162_rn(v[j + 4], v[j + 5]);
              __nv_bfloat162 p3 = __floats2bfloat162_rn(v[j + 6], v[j + 7]);
              uint4 pk;
              pk.x = *reinterpret_cast<uint32_t*>(&p0);
              pk.y = *reinterpret_cast<uint32_t*>(&p1);
              pk.z = *reinterpret_cast<uint32_t*>(&p2);
              pk.w = *reinterpret_cast<uint32_t*>(&p3);
              *reinterpret_cast<uint4*>(sC + rl * LDC + c0 + j) = pk;
            }
          }
          if (ep.out_t) {
            // neighbouring lanes (rows rl, rl^1) swap one column of each column pair, so every lane
            // writes one packed (row, row+1) bf16x2 instead of two 2-byte stores
            const bool odd = lane & 1;
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
              const float mine = odd ? v[j + 1] : v[j];
              const float give = odd ? v[j] : v[j + 1];
              const float got = __shfl_xor_sync(0xffffffffu, give, 1);
              const __nv_bfloat162 pr = odd ? __floats2bfloat162_rn(got, mine) : __floats2bfloat162_rn(mine, got);
              *reinterpret_cast<__nv_bfloat162*>(sT + (c0 + j + (odd ? 1 : 0)) * LDT + (rl & ~1)) = pr;
            }
          }
        }
        if (ep.out_f32 && row < M) {
          float* o = ep.out_f32 + (int64_t)row * ep.ld_f32 + n0 + c0;
          const bool full = n0 + c0 + 32 <= N && (ep.ld_f32 & 3) == 0 &&
                            !(ep.col_out && ep.col_split >= n0 + c0 && ep.col_split < n0 + c0 + 32);
          if (ep.accumulate) {
            if (full) {
#pragma unroll
              for (int j = 0; j < 32; j += 4) pbx_red_add_v4(o + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
            } else {
              for (int j = 0; j < 32; ++j) {
                const int n = n0 + c0 + j;
                if (n >= N) break;
                if (ep.col_out && n == ep.col_split) atomicAdd(ep.col_out + row, v[j]);
                else atomicAdd(o + j, v[j]);
              }
            }
          } else if (full) {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              *reinterpret_cast<float4*>(o + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
          } else {
            for (int j = 0; j < 32; ++j) {
              const int n = n0 + c0 + j;
              if (n >= N) break;
              if (ep.col_out && n == ep.col_split) ep.col_out[row] = v[j];
              else o[j] = v[j];
            }
          }
        }
      }
      if (threadIdx.x == 64) GEMM_TS(5);
      if (stage) {
        asm volatile("bar.sync 1, 256;" ::: "memory");   // the eight epilogue warps only
        if (ep.out_bf16) {
          constexpr int VPR = BN / 8;
          const bool vec_ok = (ep.ld_out & 7) == 0;
          for (int base = 0; base < BM * VPR; base += EPI_THREADS * 4) {   // 4 independent 16-byte moves per thread
            uint4 val[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int idx = base + u * EPI_THREADS + et;
              if (idx < BM * VPR) val[u] = *reinterpret_cast<const uint4*>(sC + (idx / VPR) * LDC + (idx % VPR) * 8);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int idx = base + u * EPI_THREADS + et;
              const int gr = m0 + idx / VPR, gc = n0 + (idx % VPR) * 8;
              if (idx >= BM * VPR || gr >= M || gc >= N) continue;
              __nv_bfloat16* o = ep.out_bf16 + (int64_t)gr * ep.ld_out + gc;
              if (gc + 8 <= N && vec_ok) {
                *reinterpret_cast<uint4*>(o) = val[u];
              } else {
                const __nv_bfloat16* e8 = reinterpret_cast<const __nv_bfloat16*>(&val[u]);
                for (int e = 0; e < 8 && gc + e < N; ++e) o[e] = e8[e];
              }
            }
          }
        }
        if (ep.out_t) {
          constexpr int VPR = BM / 8;
          const bool vec_ok = (ep.ld_t & 7) == 0;
          for (int base = 0; base < BN * VPR; base += EPI_THREADS * 4) {
            uint4 val[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int idx = base + u * EPI_THREADS + et;
              if (idx < BN * VPR) val[u] = *reinterpret_cast<const uint4*>(sT + (idx / VPR) * LDT + (idx % VPR) * 8);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int idx = base + u * EPI_THREADS + et;
              const int gn = n0 + idx / VPR, gm = m0 + (idx % VPR) * 8;
              if (idx >= BN * VPR || gn >= N || gm >= M) continue;
              __nv_bfloat16* o = ep.out_t + (int64_t)gn * ep.ld_t + gm;
              if (gm + 8 <= M && vec_ok) {
                *reinterpret_cast<uint4*>(o) = val[u];
              } else {
                const __nv_bfloat16* e8 = reinterpret_cast<const __nv_bfloat16*>(&val[u]);
                for (int e = 0; e < 8 && gm + e < M; ++e) o[e] = e8[e];
              }
            }
          }
        }
      }
    }
  }
  if (threadIdx.x == 64) GEMM_TS(6);
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, BN);
  if (threadIdx.x == 0) GEMM_TS(7);
}

template <int BN, int STG>
__global__ void __launch_bounds__(GEMM_THREADS, STG > 0 ? 2 : 1)
k_gemm_tcgen05(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, int M, int N,
               int K, int kb_per_split, GemmEpilogue ep) {
  gemm_tile<BN, STG>(&map_a, &map_b, M, N, K, kb_per_split, ep, blockIdx.x, blockIdx.y, blockIdx.z);
}

// Several independent GEMM problems in ONE launch (all weight gradients of a tower; the experts of
// an MMoE layer): a dependent launch costs ~1.7 us plus a cold pipeline fill per kernel, which is
// comparable to the work of one tower-sized GEMM.  CTAs are numbered problem after problem.
constexpr int GEMM_GROUP_MAX = 8;
struct GroupedGemm {
  CUtensorMap map_a[GEMM_GROUP_MAX];
  CUtensorMap map_b[GEMM_GROUP_MAX];
  GemmEpilogue ep[GEMM_GROUP_MAX];
  int M[GEMM_GROUP_MAX], N[GEMM_GROUP_MAX], K[GEMM_GROUP_MAX], kb_per[GEMM_GROUP_MAX];
  int gx[GEMM_GROUP_MAX], gy[GEMM_GROUP_MAX];
  int cta_end[GEMM_GROUP_MAX];
  int n;
};

template <int BN>
struct GroupedStages { static constexpr int value = BN >= 128 ? 3 : 4; };

template <int BN>
__global__ void __launch_bounds__(GEMM_THREADS, 2)
k_gemm_tcgen05_grouped(const __grid_constant__ GroupedGemm g) {
  int p = 0;
  while (p < g.n - 1 && (int)blockIdx.x >= g.cta_end[p]) ++p;
  const int local = blockIdx.x - (p ? g.cta_end[p - 1] : 0);
  const int bx = local % g.gx[p];
  const int by = (local / g.gx[p]) % g.gy[p];
  const int bz = local / (g.gx[p] * g.gy[p]);
  gemm_tile<BN, GroupedStages<BN>::value>(&g.map_a[p], &g.map_b[p], g.M[p], g.N[p], g.K[p], g.kb_per[p], g.ep[p], bx, by,
                                      bz);
}

// --------- plain SIMT kernel with the same contract: numerics cross-check / fallback ---------
__global__ void k_gemm_simt(const __nv_bfloat16* __restrict__ A, int lda, const __nv_bfloat16* __restrict__ B,
                            int ldb, int M, int N, int K, GemmEpilogue ep) {
  __shared__ float sa[16][17], sb[16][17];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int row = blockIdx.y * 16 + ty, col = blockIdx.x * 16 + tx;
  float acc = 0.f;
  for (int k0 = 0; k0 < K; k0 += 16) {
    sa[ty][tx] = (row < M && k0 + tx < K) ? __bfloat162float(A[(int64_t)row * lda + k0 + tx]) : 0.f;
    const int brow = blockIdx.x * 16 + ty;
    sb[ty][tx] = (brow < N && k0 + tx < K) ? __bfloat162float(B[(int64_t)brow * ldb + k0 + tx]) : 0.f;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) acc += sa[ty][k] * sb[tx][k];
    __syncthreads();
  }
  if (row >= M || col >= N) return;
  float x = acc * ep.alpha;
  if (ep.bias) x += ep.bias[col];
  if (ep.relu) x = fmaxf(x, 0.f);
  if (ep.mask && !(__bfloat162float(ep.mask[(int64_t)row * ep.ld_mask + col]) > 0.f)) x = 0.f;
  if (ep.out_bf16) ep.out_bf16[(int64_t)row * ep.ld_out + col] = __float2bfloat16(x);
  if (ep.out_t) ep.out_t[(int64_t)col * ep.ld_t + row] = __float2bfloat16(x);
  if (ep.out_f32) {
    float* dst = (ep.col_out && col == ep.col_split) ? ep.col_out + row : ep.out_f32 + (int64_t)row * ep.ld_f32 + col;
    if (ep.accumulate) atomicAdd(dst, x);
    else *dst = x;
  }
}

PFN_cuTensorMapEncodeTiled_v12000 get_encode_fn() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (PFN_cuTensorMapEncodeTiled_v12000)p;
  }
  return fn;
}

int make_map(CUtensorMap* map, const void* base, int rows, int cols, int ld, int box_rows) {
  auto fn = get_encode_fn();
  PBX_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  PBX_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d): base %p rows %d cols %d ld %d", (int)r, base,
              rows, cols, ld);
  return 0;
}

// STG = 0: deep pipeline, one CTA per SM (grids of at most one wave: latency matters);
// STG > 0: ~96 KB pipeline, two CTAs per SM (multi-wave grids: one CTA's epilogue overlaps the other's loads)
template <int BN, int STG>
int launch_tc(const __nv_bfloat16* A, int lda, const __nv_bfloat16* B, int ldb, int M, int N, int K, int splitk,
              const GemmEpilogue& ep, cudaStream_t st) {
  CUtensorMap ma, mb;
  int rc = make_map(&ma, A, M, K, lda, BM);
  if (rc) return rc;
  rc = make_map(&mb, B, N, K, ldb, BN);
  if (rc) return rc;
  const int num_kb = (K + BK - 1) / BK;
  if (splitk < 1) splitk = 1;
  if (splitk > num_kb) splitk = num_kb;
  const int kb_per = (num_kb + splitk - 1) / splitk;
  splitk = (num_kb + kb_per - 1) / kb_per;
  // (the attribute is per DEVICE: one flag per device ordinal, for processes that drive several GPUs)
  static bool attr_set[64] = {false};
  int dev_ord = 0;
  cudaGetDevice(&dev_ord);
  if (dev_ord < 0 || dev_ord >= 64 || !attr_set[dev_ord]) {
    PBX_CUDA_CHECK(cudaFuncSetAttribute(k_gemm_tcgen05<BN, STG>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        SmemLayout<BN, STG>::TOTAL));
    if (dev_ord >= 0 && dev_ord < 64) attr_set[dev_ord] = true;
  }
  dim3 grid((M + BM - 1) / BM, (N + BN - 1) / BN, splitk);
  k_gemm_tcgen05<BN, STG><<<grid, GEMM_THREADS, SmemLayout<BN, STG>::TOTAL, st>>>(ma, mb, M, N, K, kb_per, ep);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

template <int BN>
int launch_grouped(const PbxGemmProblem* pr, int n, cudaStream_t st) {
  static thread_local GroupedGemm g;     // ~3 KB per calling thread: filled on the host, passed by value
  memset(&g, 0, sizeof(g));
  g.n = n;
  int ctas = 0;
  for (int i = 0; i < n; ++i) {
    const PbxGemmProblem& q = pr[i];
    PBX_REQUIRE(q.M > 0 && q.N > 0 && q.K > 0, "grouped gemm: empty problem %d", i);
    PBX_REQUIRE(((uintptr_t)q.A & 15) == 0 && ((uintptr_t)q.B & 15) == 0 && (q.lda & 7) == 0 && (q.ldb & 7) == 0,
                "grouped gemm: problem %d needs 16B-aligned operands and leading dims that are multiples of 8", i);
    int rc = make_map(&g.map_a[i], q.A, q.M, q.K, q.lda, BM);
    if (rc) return rc;
    rc = make_map(&g.map_b[i], q.B, q.N, q.K, q.ldb, BN);
    if (rc) return rc;
    const int num_kb = (q.K + BK - 1) / BK;
    int splitk = q.splitk < 1 ? 1 : q.splitk;
    if (splitk > num_kb) splitk = num_kb;
    const int kb_per = (num_kb + splitk - 1) / splitk;
    splitk = (num_kb + kb_per - 1) / kb_per;
    if (splitk > 1)
      PBX_REQUIRE(q.accumulate && q.out_f32 && !q.out_bf16 && !q.out_t && !q.relu && !q.bias && !q.mask,
                  "grouped gemm: split-K needs a pure fp32 accumulate epilogue (problem %d)", i);
    GemmEpilogue& ep = g.ep[i];
    ep.bias = q.bias; ep.mask = (const __nv_bfloat16*)q.mask;
    ep.out_bf16 = (__nv_bfloat16*)q.out_bf16; ep.out_t = (__nv_bfloat16*)q.out_t; ep.out_f32 = q.out_f32;
    ep.ld_mask = q.ld_mask; ep.ld_out = q.ld_out; ep.ld_t = q.ld_t; ep.ld_f32 = q.ld_f32;
    ep.relu = q.relu; ep.accumulate = q.accumulate; ep.alpha = q.alpha;
    ep.col_out = q.col_out; ep.col_split = q.col_split; ep.ts = nullptr;
    g.M[i] = q.M; g.N[i] = q.N; g.K[i] = q.K; g.kb_per[i] = kb_per;
    g.gx[i] = (q.M + BM - 1) / BM;
    g.gy[i] = (q.N + BN - 1) / BN;
    ctas += g.gx[i] * g.gy[i] * splitk;
    g.cta_end[i] = ctas;
  }
  static bool attr_set[64] = {false};
  int dev_ord = 0;
  cudaGetDevice(&dev_ord);
  if (dev_ord < 0 || dev_ord >= 64 || !attr_set[dev_ord]) {
    PBX_CUDA_CHECK(cudaFuncSetAttribute(k_gemm_tcgen05_grouped<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        SmemLayout<BN, GroupedStages<BN>::value>::TOTAL));
    if (dev_ord >= 0 && dev_ord < 64) attr_set[dev_ord] = true;
  }
  k_gemm_tcgen05_grouped<BN><<<ctas, GEMM_THREADS, SmemLayout<BN, GroupedStages<BN>::value>::TOTAL, st>>>(g);
  PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
  return 0;
}

unsigned long long* g_gemm_ts = nullptr;   // device buffer of 8 stamps when the debug timeline is on

}  // namespace

extern "C" {

int pbx_gemm_bf16_tn_grouped(const PbxGemmProblem* problems, int n, int block_n, cudaStream_t st) {
  PBX_REQUIRE(n >= 1 && n <= GEMM_GROUP_MAX, "grouped gemm: %d problems (max %d)", n, GEMM_GROUP_MAX);
  switch (block_n == 0 ? 64 : block_n) {
    case 32: return launch_grouped<32>(problems, n, st);
    case 64: return launch_grouped<64>(problems, n, st);
    case 128: return launch_grouped<128>(problems, n, st);
    default: PBX_REQUIRE(false, "grouped gemm: unsupported block_n %d", block_n);
  }
  return 0;
}

// Debug timeline of CTA (0,0,0) of the next tcgen05 GEMM launches: kernel entry, setup done, first
// stage landed, last MMA issued, accumulator complete, epilogue registers done, stores issued, exit.
int pbx_gemm_debug_enable(int on) {
  if (on && !g_gemm_ts) {
    PBX_CUDA_CHECK(cudaMalloc(&g_gemm_ts, 8 * sizeof(unsigned long long)));
    PBX_CUDA_CHECK(cudaMemset(g_gemm_ts, 0, 8 * sizeof(unsigned long long)));
  } else if (!on && g_gemm_ts) {
    cudaFree(g_gemm_ts);
    g_gemm_ts = nullptr;
  }
  return 0;
}

int pbx_gemm_debug_ts(unsigned long long* out8) {
  PBX_REQUIRE(g_gemm_ts != nullptr, "gemm debug timeline is off");
  PBX_CUDA_CHECK(cudaDeviceSynchronize());
  PBX_CUDA_CHECK(cudaMemcpy(out8, g_gemm_ts, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  return 0;
}

// impl: 0 = tcgen05 (default), 1 = SIMT cross-check.  block_n in {32, 64, 128, 256} (0 = auto).
// Requirements for the tcgen05 path: A/B 16-byte aligned, lda/ldb multiples of 8 elements.
int pbx_gemm_bf16_tn(const void* A, int lda, const void* B, int ldb, int M, int N, int K, const float* bias,
                     const void* mask, int ld_mask, void* out_bf16, int ld_out, void* out_t, int ld_t,
                     float* out_f32, int ld_f32, int relu, int accumulate, float alpha, int splitk, int block_n,
                     int impl, float* col_out, int col_split, cudaStream_t st) {
  if (M <= 0 || N <= 0) return 0;
  PBX_REQUIRE(K > 0, "gemm: K must be positive");
  GemmEpilogue ep;
  ep.bias = bias;
  ep.mask = (const __nv_bfloat16*)mask;
  ep.out_bf16 = (__nv_bfloat16*)out_bf16;
  ep.out_t = (__nv_bfloat16*)out_t;
  ep.out_f32 = out_f32;
  ep.ld_mask = ld_mask; ep.ld_out = ld_out; ep.ld_t = ld_t; ep.ld_f32 = ld_f32;
  ep.relu = relu;
  ep.accumulate = accumulate;
  ep.alpha = alpha;
  ep.col_out = col_out;
  ep.col_split = col_split;
  ep.ts = g_gemm_ts;
  if (splitk > 1) PBX_REQUIRE(accumulate && out_f32 && !out_bf16 && !out_t && !relu && !bias && !mask,
                              "gemm: split-K needs a pure fp32 accumulate epilogue");
  if (impl == 1) {
    dim3 grid((N + 15) / 16, (M + 15) / 16);
    k_gemm_simt<<<grid, dim3(16, 16), 0, st>>>((const __nv_bfloat16*)A, lda, (const __nv_bfloat16*)B, ldb, M, N, K,
                                               ep);
    PBX_CUDA_CHECK(cudaGetLastError()); pbx_note_launch(1);
    return 0;
  }
  PBX_REQUIRE(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0 && (lda & 7) == 0 && (ldb & 7) == 0,
              "gemm: tcgen05 path needs 16B-aligned operands and leading dims that are multiples of 8");
  if (block_n == 0) {
    // enough CTAs to cover the 148 SMs when the problem allows it
    long long tiles128 = (long long)((M + BM - 1) / BM) * ((N + 127) / 128) * (splitk > 0 ? splitk : 1);
    // up to one wave of 128-wide tiles: 64-wide tiles instead (twice the CTAs, two per SM with the
    // shallow pipeline, so one CTA's epilogue overlaps the other's loads: 12.3 vs 14.3 us on the
    // 4096x512x507 layer)
    block_n = (N <= 32) ? 32 : (N <= 64 || tiles128 <= 148) ? 64 : 128;
    // large problems: 128x256 tiles halve the A-operand traffic per FLOP (measured 0.80 of the
    // cuBLAS-measured peak at 8192^3 vs 0.51 with 128x128) once there are >= 2 waves of them
    long long tiles256 = (long long)((M + BM - 1) / BM) * ((N + 255) / 256) * (splitk > 0 ? splitk : 1);
    if (N >= 256 && tiles256 >= 2 * 148) block_n = 256;
  }
  const __nv_bfloat16* a = (const __nv_bfloat16*)A;
  const __nv_bfloat16* b = (const __nv_bfloat16*)B;
  // more CTAs than SMs: the shallow two-CTAs-per-SM variant (PBX_GEMM_PIPE=deep|shallow overrides)
  static const char* pipe_env = getenv("PBX_GEMM_PIPE");
  const long long ctas = (long long)((M + BM - 1) / BM) * ((N + block_n - 1) / block_n) * (splitk > 0 ? splitk : 1);
  bool shallow = ctas > 148;
  if (pipe_env) shallow = pipe_env[0] == 's';
  switch (block_n) {
    case 32: return shallow ? launch_tc<32, 4>(a, lda, b, ldb, M, N, K, splitk, ep, st)
                            : launch_tc<32, 0>(a, lda, b, ldb, M, N, K, splitk, ep, st);
    case 64: return shallow ? launch_tc<64, 4>(a, lda, b, ldb, M, N, K, splitk, ep, st)
                            : launch_tc<64, 0>(a, lda, b, ldb, M, N, K, splitk, ep, st);
    case 128: return shallow ? launch_tc<128, 3>(a, lda, b, ldb, M, N, K, splitk, ep, st)
                             : launch_tc<128, 0>(a, lda, b, ldb, M, N, K, splitk, ep, st);
    case 256: return launch_tc<256, 0>(a, lda, b, ldb, M, N, K, splitk, ep, st);
    default: PBX_REQUIRE(false, "gemm: unsupported block_n %d", block_n);
  }
  return 0;
}

}  // extern "C"
