// Shared device/host helpers for the sm_100a kernels of paddlebox_b200.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#define PBX_MAX_RANKS 8
#define PBX_EMPTY_KEY 0xFFFFFFFFFFFFFFFFull
#define PBX_NO_ROW 0xFFFFFFFFu

// ---- error plumbing (host) ------------------------------------------------
extern "C" const char* pbx_last_error();
void pbx_set_error(const char* fmt, ...);

#define PBX_CUDA_CHECK(expr)                                                     \
  do {                                                                           \
    cudaError_t _e = (expr);                                                     \
    if (_e != cudaSuccess) {                                                     \
      pbx_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr,                 \
                    cudaGetErrorString(_e));                                     \
      return (int)_e;                                                            \
    }                                                                            \
  } while (0)

// every kernel launch made by this library is counted (bench.py reports it as gpu_launches)
extern "C" long long pbx_launch_count();
void pbx_note_launch(int n);

#define PBX_REQUIRE(cond, ...)                                                   \
  do {                                                                           \
    if (!(cond)) {                                                               \
      pbx_set_error(__VA_ARGS__);                                                \
      return -1;                                                                 \
    }                                                                            \
  } while (0)

static inline int pbx_num_sms() {
  static int sms[64] = {0};     // same bound as kMaxDevices of the allocation-free entry point
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (sms[dev] == 0) {
    cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev);
    if (sms[dev] <= 0) sms[dev] = 148;
  }
  return sms[dev];
}

static inline int pbx_grid_for(long long n, int block, int max_waves = 8) {
  long long g = (n + block - 1) / block;
  long long cap = (long long)pbx_num_sms() * max_waves;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

// ---- device helpers ---------------------------------------------------------
__host__ __device__ __forceinline__ uint64_t pbx_mix64(uint64_t k) {
  // murmur3 finalizer: full-avalanche 64-bit mix
  k ^= k >> 33;
  k *= 0xff51afd7ed558ccdull;
  k ^= k >> 33;
  k *= 0xc4ceb9fe1a85ec53ull;
  k ^= k >> 33;
  return k;
}

__host__ __device__ __forceinline__ int pbx_owner_of(uint64_t key, int nranks) {
  // Shard by key hash (reference: key % ngpu, heter_comm_kernel.cu:85-93).
  return (int)((pbx_mix64(key) >> 40) % (uint64_t)nranks);
}

__device__ __forceinline__ uint64_t pbx_home_slot(uint64_t key, uint64_t capacity) {
  return __umul64hi(pbx_mix64(key ^ 0x9e3779b97f4a7c15ull), capacity);
}

// counter-based uniform in (0,1]: keyed by (key, lane index, seed) so that lazy
// embedx creation is reproducible (the reference seeds curand with clock64()).
__host__ __device__ __forceinline__ float pbx_uniform01(uint64_t key, uint32_t idx, uint64_t seed) {
  uint64_t x = pbx_mix64(key + 0x632be59bd9b4e019ull * (uint64_t)(idx + 1) + seed);
  return ((float)((x >> 40) + 1)) * (1.0f / 16777216.0f);
}

__device__ __forceinline__ void pbx_st_release_sys(uint64_t* p, uint64_t v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t pbx_ld_acquire_sys(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t pbx_globaltimer() {
  uint64_t t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ float4 pbx_ldg_f4(const float* p) {
  return __ldg(reinterpret_cast<const float4*>(p));
}
__device__ __forceinline__ void pbx_red_add_v4(float* p, float a, float b, float c, float d) {
  // sm_90+: vectorised fp32 reduction (one L2 atomic op for 16 bytes)
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c),
               "f"(d)
               : "memory");
}
