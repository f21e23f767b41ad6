// Microbenchmark: do warp shuffles compete with shared-memory loads for the same pipe?
// 8 warps per SM, one CTA per SM; cycles per iteration of 24 LDS.64, 24 64-bit shuffles
// (2 x SHFL.32 each), and both interleaved.
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void __launch_bounds__(256, 1) k(double *out, long long *cyc, int iters)
{
  extern __shared__ double sm[];
  const int tid = threadIdx.x;
  for (int i = tid; i < 8192; i += 256) sm[i] = i * 0.5;
  __syncthreads();
  double acc = tid * 1e-3, v = tid + 0.25;
  long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
    if (MODE == 0 || MODE == 2) {
#pragma unroll
      for (int j = 0; j < 24; j++) acc += sm[(tid + it + j * 37) & 8191];
    }
    if (MODE == 1 || MODE == 2) {
#pragma unroll
      for (int j = 0; j < 24; j++) acc += __shfl_sync(0xffffffffu, v + j, (tid + 1 + (j % 3)) & 31);
    }
    v += 1.0;
  }
  long long t1 = clock64();
  out[blockIdx.x * 256 + tid] = acc;
  if (tid == 0) cyc[blockIdx.x] = t1 - t0;
}
int main()
{
  double *out; long long *cyc;
  cudaMalloc(&out, 148 * 256 * 8); cudaMalloc(&cyc, 148 * 8);
  const int iters = 2000;
  cudaFuncSetAttribute(k<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  const char *names[] = {"24 LDS.64", "24 x 64-bit shuffle (48 SHFL.32)", "24 LDS.64 + 24 x 64-bit shuffle"};
  for (int m = 0; m < 3; m++) {
    for (int rep = 0; rep < 2; rep++) {
      if (m == 0) k<0><<<148, 256, 65536>>>(out, cyc, iters);
      if (m == 1) k<1><<<148, 256, 65536>>>(out, cyc, iters);
      if (m == 2) k<2><<<148, 256, 65536>>>(out, cyc, iters);
    }
    cudaError_t e = cudaGetLastError(); if (e == cudaSuccess) e = cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    printf("%-36s: %8.1f cycles/iter per SM (8 warps) %s\n", names[m], (double)h[0] / iters, cudaGetErrorString(e));
  }
  return 0;
}
