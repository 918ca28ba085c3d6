// pbgpu_peak.cu -- FP64 DFMA peak microbenchmark (the roofline denominator of this path).
//
// MEASURED_PEAKS.json (driver-written) holds the HBM copy bandwidth and the bf16 GEMM rate only; the
// batched DAE path is bound by the FP64 pipe, so bench.py measures the DFMA
// rate of the box it runs on with this kernel: 8 independent FMA chains per thread, all SMs full.
#include <cuda_runtime.h>

#include <cstdio>

#include "pbgpu_peak.h"

__global__ void __launch_bounds__(256) pb_dfma_kernel(double* out, int iters, double a, double b) {
  double x0 = threadIdx.x * 1e-3, x1 = x0 + 1.0, x2 = x0 + 2.0, x3 = x0 + 3.0;
  double x4 = x0 + 4.0, x5 = x0 + 5.0, x6 = x0 + 6.0, x7 = x0 + 7.0;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int u = 0; u < 8; u++) {
      x0 = fma(x0, a, b);
      x1 = fma(x1, a, b);
      x2 = fma(x2, a, b);
      x3 = fma(x3, a, b);
      x4 = fma(x4, a, b);
      x5 = fma(x5, a, b);
      x6 = fma(x6, a, b);
      x7 = fma(x7, a, b);
    }
  }
  const double s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
  if (s == 123.456) out[0] = s;  // never true: keeps the chains alive
}

extern "C" int pbgpu_measure_fp64_peak(int device, int iters, double* tflops, float* ms_out) {
  if (cudaSetDevice(device) != cudaSuccess) return 1;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return 1;
  double* out = nullptr;
  if (cudaMalloc(&out, sizeof(double)) != cudaSuccess) return 1;
  const int blocks = prop.multiProcessorCount * 8, threads = 256;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  pb_dfma_kernel<<<blocks, threads>>>(out, 64, 0.999999, 1e-9);  // warm-up
  float best = 1e30f;
  for (int rep = 0; rep < 5; rep++) {
    cudaEventRecord(e0);
    pb_dfma_kernel<<<blocks, threads>>>(out, iters, 0.999999, 1e-9);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) return 1;
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  const double flops = 2.0 * 64.0 * (double)iters * (double)blocks * (double)threads;
  *tflops = flops / (best * 1e-3) / 1e12;
  if (ms_out) *ms_out = best;
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(out);
  return 0;
}
