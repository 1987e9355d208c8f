// peaks.cu -- on-box microbenchmarks for the two on-chip rooflines of the fold kernels.
//
// The fold DP is integer min-plus over irregular gathers: it is bounded by shared-memory load
// throughput and by INT32 ALU issue, not by HBM or the tensor cores.
// Neither peak is in MEASURED_PEAKS.json, so bench.py measures both on the device it runs on:
//   * conflict-free ld.shared.b32 stream            -> 4-byte shared loads per second
//   * independent add+min chains (DPX VIADDMNMX)    -> min-plus candidate updates per second
#include <cuda_runtime.h>
#include <stdint.h>

#include "fold_kernels.h"

namespace rnafold {

namespace {

constexpr int kPeakThreads = 256;
constexpr int kPeakIters = 512;
constexpr int kPeakUnroll = 16;

__global__ void __launch_bounds__(kPeakThreads) lds_stream_kernel(int *out, int iters, unsigned rot) {
  __shared__ int buf[kPeakThreads * kPeakUnroll];
  for (int k = threadIdx.x; k < kPeakThreads * kPeakUnroll; k += kPeakThreads) buf[k] = k * 7 + blockIdx.x;
  __syncthreads();
  const unsigned base = (unsigned)__cvta_generic_to_shared(buf) + 4u * threadIdx.x;
  int acc[kPeakUnroll];
#pragma unroll
  for (int u = 0; u < kPeakUnroll; ++u) acc[u] = 0;
  // Consecutive lanes -> consecutive words (conflict free).  Each of the 16 loads of an iteration reads its own
  // row; the rows move by `rot` (a run-time value) per iteration and the loads are ld.volatile, so ptxas can
  // neither prove two loads equal nor reuse one across iterations: 16 LDS execute per iteration, as counted
  // (profiles/r2_peaks_sass.txt).  Round 1's version rotated the rows by a compile-time 1 without volatile: ptxas
  // shared each load between two iterations and the kernel executed half the loads it was credited with.
  unsigned row = 0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < kPeakUnroll; ++u) {
      int v;
      asm volatile("ld.volatile.shared.b32 %0, [%1];"
                   : "=r"(v)
                   : "r"(base + 4u * kPeakThreads * ((row + (unsigned)u) & (kPeakUnroll - 1)))
                   : "memory");
      acc[u] += v;
    }
    row += rot;
  }
  int r = 0;
#pragma unroll
  for (int u = 0; u < kPeakUnroll; ++u) r ^= acc[u];
  if (r == 0x7fffffff) out[0] = r;
}

__global__ void __launch_bounds__(kPeakThreads) minplus_stream_kernel(int *out, int iters, int seed) {
  int acc[kPeakUnroll];
  int x = seed + threadIdx.x;
#pragma unroll
  for (int u = 0; u < kPeakUnroll; ++u) acc[u] = 1000000 + u;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < kPeakUnroll; ++u) acc[u] = __viaddmin_s32(acc[u], x, acc[(u + 1) % kPeakUnroll]);
  }
  int r = 0;
#pragma unroll
  for (int u = 0; u < kPeakUnroll; ++u) r ^= acc[u];
  if (r == 0x7fffffff) out[0] = r;
}

}  // namespace

// Returns per-second rates over the whole device (all SMs busy, 8 CTAs of 256 threads per SM).
cudaError_t measure_onchip_peaks(int num_sms, cudaStream_t st, double *lds_per_s, double *minplus_per_s) {
  // 256 threads x 16 KB static shared memory: 8 CTAs per SM stay resident (128 KB)
  int *d_out = nullptr;
  cudaError_t e = cudaMalloc((void **)&d_out, 64);
  if (e != cudaSuccess) return e;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int grid = num_sms * 8;
  float best_lds = 1e30f, best_alu = 1e30f;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(e0, st);
    lds_stream_kernel<<<grid, kPeakThreads, 0, st>>>(d_out, kPeakIters, 5u);
    cudaEventRecord(e1, st);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep > 0 && ms < best_lds) best_lds = ms;
    cudaEventRecord(e0, st);
    minplus_stream_kernel<<<grid, kPeakThreads, 0, st>>>(d_out, kPeakIters * 4, rep);
    cudaEventRecord(e1, st);
    cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep > 0 && ms < best_alu) best_alu = ms;
  }
  e = cudaGetLastError();
  const double n_lds = (double)grid * kPeakThreads * (double)kPeakIters * kPeakUnroll;
  const double n_alu = (double)grid * kPeakThreads * (double)kPeakIters * 4 * kPeakUnroll;
  *lds_per_s = n_lds / (best_lds * 1e-3);
  *minplus_per_s = n_alu / (best_alu * 1e-3);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d_out);
  return e;
}

}  // namespace rnafold
