// peaks.cu -- live FP64 peak measurements for the roofline denominators (bench.py).
// MEASURED_PEAKS.json carries an HBM copy figure and a bf16 GEMM figure but no FP64 figure, and the hot
// kernels of this path are FP64: a dependent-free DFMA loop (FP64 pipe) and an mma.sync m8n8k4 f64 loop
// (FP64 tensor pipe) are timed with CUDA events, best of several launches.
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdint>

namespace {

__global__ void k_dfma_peak(double* out, int iters, double seed) {
  double a[8];
#pragma unroll
  for (int j = 0; j < 8; j++) a[j] = seed + j + threadIdx.x * 1e-9;
  double b = 1.0000001, c = 1e-7;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int j = 0; j < 8; j++) a[j] = fma(a[j], b, c);
  }
  double s = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) s += a[j];
  if (s == 12345.678) out[0] = s;
}

__global__ void k_dmma_peak(double* out, int iters, double seed) {
  double d[8][2];
#pragma unroll
  for (int j = 0; j < 8; j++) d[j][0] = d[j][1] = 0.0;
  double a = seed + (threadIdx.x & 31) * 1e-3, b = 1.0 + (threadIdx.x & 31) * 1e-4;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int j = 0; j < 8; j++)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(d[j][0]), "+d"(d[j][1]) : "d"(a), "d"(b));
  }
  double s = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) s += d[j][0] + d[j][1];
  if (s == 12345.678) out[0] = s;
}

// DMMA and DFMA issued together: do the FP64 tensor sub-pipe and the FP64 pipe add up, or share their multipliers?
__global__ void k_mixed_peak(double* out, int iters, double seed) {
  double d[4][2], a8[8];
#pragma unroll
  for (int j = 0; j < 4; j++) d[j][0] = d[j][1] = 0.0;
#pragma unroll
  for (int j = 0; j < 8; j++) a8[j] = seed + j + threadIdx.x * 1e-9;
  double a = seed + (threadIdx.x & 31) * 1e-3, b = 1.0 + (threadIdx.x & 31) * 1e-4, bb = 1.0000001, c = 1e-7;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int j = 0; j < 4; j++)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(d[j][0]), "+d"(d[j][1]) : "d"(a), "d"(b));
#pragma unroll
    for (int r = 0; r < 4; r++)
#pragma unroll
      for (int j = 0; j < 8; j++) a8[j] = fma(a8[j], bb, c);
  }
  double s = 0;
#pragma unroll
  for (int j = 0; j < 4; j++) s += d[j][0] + d[j][1];
#pragma unroll
  for (int j = 0; j < 8; j++) s += a8[j];
  if (s == 12345.678) out[0] = s;
}

// 8 independent DFMA chains with NI independent integer multiply-adds per trip: how much of the FP64 pipe is left when the
// instruction stream is mixed?  (If a DFMA holds the scheduler's dispatch port for two cycles, every other instruction costs a
// cycle on top of 2 x the FP64 count: the model behind the family kernels' instruction budgets in DESIGN.md.)
template <int NI>
__global__ void k_dfma_int_mix(double* out, int iters, double seed) {
  double a[8];
  unsigned x[NI > 0 ? NI : 1];
#pragma unroll
  for (int j = 0; j < 8; j++) a[j] = seed + j + threadIdx.x * 1e-9;
#pragma unroll
  for (int j = 0; j < NI; j++) x[j] = threadIdx.x * 7u + j;
  double b = 1.0000001, c = 1e-7;
  unsigned m = (unsigned)(seed * 3.0) | 1u;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int j = 0; j < 8; j++) a[j] = fma(a[j], b, c);
#pragma unroll
    for (int j = 0; j < NI; j++) x[j] = x[j] * m + (unsigned)i;
  }
  double s = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) s += a[j];
  unsigned t = 0;
#pragma unroll
  for (int j = 0; j < NI; j++) t ^= x[j];
  if (s == 12345.678 || t == 0x12345u) out[0] = s + t;
}

}  // namespace

// DFMA TFLOP/s with 0, 4, 8 and 16 integer instructions per 8 DFMAs in the loop
extern "C" int nimb200_measure_fp64_issue_mix(int device, double* tflops4) {
  if (cudaSetDevice(device) != cudaSuccess) return 1;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return 1;
  double* out = nullptr;
  if (cudaMalloc(&out, 8) != cudaSuccess) return 1;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int threads = 256, blocks = prop.multiProcessorCount * 8, iters = 4096;
  for (int v = 0; v < 4; v++) {
    double best = 0;
    for (int rep = 0; rep < 4; rep++) {
      float ms = 0;
      cudaEventRecord(e0);
      if (v == 0) k_dfma_int_mix<0><<<blocks, threads>>>(out, iters, 1.0);
      else if (v == 1) k_dfma_int_mix<4><<<blocks, threads>>>(out, iters, 1.0);
      else if (v == 2) k_dfma_int_mix<8><<<blocks, threads>>>(out, iters, 1.0);
      else k_dfma_int_mix<16><<<blocks, threads>>>(out, iters, 1.0);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      cudaEventElapsedTime(&ms, e0, e1);
      const double flops = 2.0 * 8 * iters * (double)threads * blocks;
      if (rep > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
    }
    tflops4[v] = best;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(out);
  return cudaGetLastError() == cudaSuccess ? 0 : 1;
}

// flops of a loop that issues 4 DMMA (2048 flop/warp) and 32 DFMA (2048 flop/warp) per iteration
extern "C" int nimb200_measure_fp64_mixed(int device, double* tflops) {
  if (cudaSetDevice(device) != cudaSuccess) return 1;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return 1;
  double* out = nullptr;
  if (cudaMalloc(&out, 8) != cudaSuccess) return 1;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int threads = 256, blocks = prop.multiProcessorCount * 8, iters = 4096;
  double best = 0;
  for (int rep = 0; rep < 5; rep++) {
    float ms = 0;
    cudaEventRecord(e0);
    k_mixed_peak<<<blocks, threads>>>(out, iters, 1.0);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    double flops = (4.0 * 2 * 256 + 32.0 * 2 * 32) * iters * (double)(threads / 32) * blocks;
    if (rep > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(out);
  *tflops = best;
  return cudaGetLastError() == cudaSuccess ? 0 : 1;
}

extern "C" int nimb200_measure_fp64_peaks(int device, double* dfma_tflops, double* dmma_tflops) {
  if (cudaSetDevice(device) != cudaSuccess) return 1;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return 1;
  double* out = nullptr;
  if (cudaMalloc(&out, 8) != cudaSuccess) return 1;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int threads = 256, blocks = prop.multiProcessorCount * 8, iters = 4096;
  double best_f = 0, best_m = 0;
  for (int rep = 0; rep < 6; rep++) {
    float ms = 0;
    cudaEventRecord(e0);
    k_dfma_peak<<<blocks, threads>>>(out, iters, 1.0);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    double flops = 2.0 * 8 * iters * (double)threads * blocks;
    if (rep > 0) best_f = std::max(best_f, flops / (ms * 1e-3) / 1e12);
    cudaEventRecord(e0);
    k_dmma_peak<<<blocks, threads>>>(out, iters, 1.0);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    double mflops = 2.0 * 8 * 8 * 4 * 8 * iters * (double)(threads / 32) * blocks;
    if (rep > 0) best_m = std::max(best_m, mflops / (ms * 1e-3) / 1e12);
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(out);
  if (cudaGetLastError() != cudaSuccess) return 1;
  *dfma_tflops = best_f;
  *dmma_tflops = best_m;
  return 0;
}
