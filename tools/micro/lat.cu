// Latency microbenchmarks (one warp): dependent DFMA / DADD / DMUL chains, LDS.64 pointer chase, SHFL+DADD butterfly.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(double* out, long long* cyc, int iters) {
  __shared__ double sm[1024];
  __shared__ int idx[1024];
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) { sm[i] = 1.0 + i * 1e-9; idx[i] = (i * 37 + 11) & 1023; }
  __syncthreads();
  double x = threadIdx.x * 1e-3 + 1.0, a = 0.999999, b = 1e-9;
  long long t0 = clock64();
  for (int i = 0; i < iters; i++) { x = fma(x, a, b); x = fma(x, a, b); x = fma(x, a, b); x = fma(x, a, b); }
  long long t1 = clock64();
  for (int i = 0; i < iters; i++) { x = x + b; x = x + a; x = x + b; x = x + a; }
  long long t2 = clock64();
  for (int i = 0; i < iters; i++) { x = x * a; x = x * a; x = x * a; x = x * a; }
  long long t3 = clock64();
  int j = threadIdx.x;
  for (int i = 0; i < iters; i++) { j = idx[j]; j = idx[j]; j = idx[j]; j = idx[j]; }
  long long t4 = clock64();
  double y = x;
  for (int i = 0; i < iters; i++) { int q = (int)(__double_as_longlong(y) & 1023); y = sm[q]; q = (int)(__double_as_longlong(y) & 1023); y += sm[q]; }
  long long t5 = clock64();
  for (int i = 0; i < iters; i++) { y += __shfl_xor_sync(0xffffffffu, y, 1); y += __shfl_xor_sync(0xffffffffu, y, 2); }
  long long t6 = clock64();
  for (int i = 0; i < iters; i++) { __syncwarp(); y += 1.0; __syncwarp(); y += 1.0; }
  long long t7 = clock64();
  for (int i = 0; i < iters; i++) { __syncthreads(); y += 1.0; __syncthreads(); y += 1.0; }
  long long t8 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    cyc[0] = t1 - t0; cyc[1] = t2 - t1; cyc[2] = t3 - t2; cyc[3] = t4 - t3; cyc[4] = t5 - t4; cyc[5] = t6 - t5; cyc[6] = t7 - t6; cyc[7] = t8 - t7;
  }
  out[threadIdx.x + blockIdx.x * blockDim.x] = x + y + j;
}
int main() {
  double* out; long long* cyc; cudaMalloc(&out, 8 * 1024 * 256); cudaMalloc(&cyc, 64);
  const int iters = 2000;
  for (int threads : {32, 128, 256}) {
    k<<<1, threads>>>(out, cyc, iters); cudaDeviceSynchronize();
    k<<<1, threads>>>(out, cyc, iters); cudaDeviceSynchronize();
    long long h[8]; cudaMemcpy(h, cyc, 64, cudaMemcpyDeviceToHost);
    printf("threads %d: DFMA %.1f DADD %.1f DMUL %.1f LDS.32chase %.1f LDS.64+decode(+DADD every 2nd) %.1f SHFL64+DADD %.1f syncwarp+DADD %.1f syncthreads+DADD %.1f cycles/op\n",
           threads, h[0] / (4.0 * iters), h[1] / (4.0 * iters), h[2] / (4.0 * iters), h[3] / (4.0 * iters), h[4] / (2.0 * iters), h[5] / (2.0 * iters), h[6] / (2.0 * iters), h[7] / (2.0 * iters));
  }
  return 0;
}
