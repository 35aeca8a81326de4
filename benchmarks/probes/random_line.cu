// Probe: how many random 128-byte lines per nanosecond can this GPU read (and read-modify-write)
// from a table far larger than L2?  This is the ceiling the owner-side phases of the sparse
// kernels (k_sparse_pull P3: key probe + row read; k_sparse_push Q3: row read + row write) run
// against, so it is measured rather than assumed.
//
//   pattern "row"  : 8 lanes read one whole 128-B line with 16-B loads (how the kernels read rows)
//   pattern "word" : every lane reads 8 B of its own random line (how a key probe touches a line)
//   pattern "rmw"  : "row" + the line is written back
// U = independent random lines in flight per lane group before any is consumed.
//
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o random_line random_line.cu && ./random_line [GiB]
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

__device__ __forceinline__ uint64_t mix64(uint64_t k) {
  k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
  return k;
}

template <int U, int MODE, int TPB>   // MODE 0 row, 1 word, 2 rmw
__global__ void __launch_bounds__(TPB, 1) k_probe(float4* __restrict__ tab, uint64_t n_lines, int iters, float* sink) {
  const uint64_t gtid = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  const int sub = threadIdx.x & 7;
  const uint64_t group = MODE == 1 ? gtid : gtid >> 3;
  float acc = 0.f;
  for (int it = 0; it < iters; ++it) {
    float4 v[U];
    uint64_t line[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      line[u] = __umul64hi(mix64(group * 0x9E3779B97F4A7C15ull + (uint64_t)(it * U + u) * 0xD1B54A32D192ED03ull + 12345), n_lines);
      if (MODE == 1) {
        const double2* p = reinterpret_cast<const double2*>(tab + line[u] * 8);
        double2 d = __ldcg(p);
        v[u].x = (float)d.x;
      } else {
        v[u] = __ldcg(tab + line[u] * 8 + sub);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      acc += v[u].x;
      if (MODE == 2) {
        v[u].x += 1.f;
        tab[line[u] * 8 + sub] = v[u];
      }
    }
  }
  if (acc == 1234.5678f) *sink = acc;
}

template <int U, int MODE, int TPB = 1024>
double run(float4* tab, uint64_t n_lines, float* sink, int grid) {
  const int iters = 64 / U < 1 ? 1 : 64 / U;
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  k_probe<U, MODE, TPB><<<grid, TPB>>>(tab, n_lines, iters, sink);    // warm-up
  cudaDeviceSynchronize();
  cudaEventRecord(a);
  k_probe<U, MODE, TPB><<<grid, TPB>>>(tab, n_lines, iters, sink);
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms = 0;
  cudaEventElapsedTime(&ms, a, b);
  const double groups = (double)grid * TPB / (MODE == 1 ? 1 : 8);
  const double lines = groups * iters * U;
  return lines / (ms * 1e6);   // lines per ns
}

int main(int argc, char** argv) {
  const double gib = argc > 1 ? atof(argv[1]) : 24.0;
  const uint64_t n_lines = (uint64_t)(gib * 1024 * 1024 * 1024 / 128);
  float4* tab = nullptr;
  float* sink = nullptr;
  if (cudaMalloc(&tab, n_lines * 128) != cudaSuccess) { printf("cudaMalloc of %.1f GiB failed\n", gib); return 1; }
  cudaMalloc(&sink, 4);
  cudaMemset(tab, 0, n_lines * 128);
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  printf("{\"table_gib\": %.1f, \"sms\": %d, \"unit\": \"random 128-B lines per ns (x128 = GB/s)\",\n", gib, sms);
  const char* names[3] = {"row_read_8lanes_x16B", "word_read_8B_per_lane", "row_read_modify_write"};
  for (int mode = 0; mode < 3; ++mode) {
    double r[5];
    for (int g = 1; g <= 2; ++g) {
      const int grid = sms * g;   // g CTAs of 1024 threads per SM (2 only fit while the kernel stays under 32 registers)
      if (mode == 0) { r[0] = run<1, 0>(tab, n_lines, sink, grid); r[1] = run<2, 0>(tab, n_lines, sink, grid); r[2] = run<4, 0>(tab, n_lines, sink, grid); r[3] = run<8, 0>(tab, n_lines, sink, grid); r[4] = run<16, 0>(tab, n_lines, sink, grid); }
      if (mode == 1) { r[0] = run<1, 1>(tab, n_lines, sink, grid); r[1] = run<2, 1>(tab, n_lines, sink, grid); r[2] = run<4, 1>(tab, n_lines, sink, grid); r[3] = run<8, 1>(tab, n_lines, sink, grid); r[4] = run<16, 1>(tab, n_lines, sink, grid); }
      if (mode == 2) { r[0] = run<1, 2>(tab, n_lines, sink, grid); r[1] = run<2, 2>(tab, n_lines, sink, grid); r[2] = run<4, 2>(tab, n_lines, sink, grid); r[3] = run<8, 2>(tab, n_lines, sink, grid); r[4] = run<16, 2>(tab, n_lines, sink, grid); }
      printf(" \"%s_ctas_per_sm_%d\": {\"U1\": %.2f, \"U2\": %.2f, \"U4\": %.2f, \"U8\": %.2f, \"U16\": %.2f},\n", names[mode], g,
             r[0], r[1], r[2], r[3], r[4]);
    }
  }
  // 256-thread CTAs, 8 per SM (2048 threads per SM whatever the register count up to 32; fewer resident otherwise)
  {
    double r[5];
    const int grid = sms * 8;
    r[0] = run<1, 0, 256>(tab, n_lines, sink, grid); r[1] = run<2, 0, 256>(tab, n_lines, sink, grid); r[2] = run<4, 0, 256>(tab, n_lines, sink, grid); r[3] = run<8, 0, 256>(tab, n_lines, sink, grid); r[4] = run<16, 0, 256>(tab, n_lines, sink, grid);
    printf(" \"row_read_256thr_x8\": {\"U1\": %.2f, \"U2\": %.2f, \"U4\": %.2f, \"U8\": %.2f, \"U16\": %.2f},\n", r[0], r[1], r[2], r[3], r[4]);
    r[0] = run<1, 2, 256>(tab, n_lines, sink, grid); r[1] = run<2, 2, 256>(tab, n_lines, sink, grid); r[2] = run<4, 2, 256>(tab, n_lines, sink, grid); r[3] = run<8, 2, 256>(tab, n_lines, sink, grid); r[4] = run<16, 2, 256>(tab, n_lines, sink, grid);
    printf(" \"rmw_256thr_x8\": {\"U1\": %.2f, \"U2\": %.2f, \"U4\": %.2f, \"U8\": %.2f, \"U16\": %.2f}\n", r[0], r[1], r[2], r[3], r[4]);
  }
  printf("}\n");
  return 0;
}
