// Microbenchmark: is TMEM usable as per-thread scratch next to shared memory?
// Measures cycles per warp-wide tcgen05.ld/.st (32x32b.x2 = one double per thread)
// alone and interleaved with LDS.64, 8 warps per SM, one CTA per SM.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void tm_st2(unsigned taddr, double v)
{
  unsigned lo = __double2loint(v), hi = __double2hiint(v);
  asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};" ::"r"(taddr), "r"(lo), "r"(hi) : "memory");
}
__device__ __forceinline__ void tm_ld2_issue(unsigned taddr, unsigned &lo, unsigned &hi)
{ asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(lo), "=r"(hi) : "r"(taddr) : "memory"); }
__device__ __forceinline__ void tm_ld8_issue(unsigned taddr, unsigned (&r)[8])
{ asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(taddr) : "memory"); }
__device__ __forceinline__ void tm_ld16_issue(unsigned taddr, unsigned (&r)[16])
{ asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]) : "r"(taddr) : "memory"); }
__device__ __forceinline__ void tm_st8(unsigned taddr, const unsigned (&r)[8])
{ asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" :: "r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory"); }
__device__ __forceinline__ void tm_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tm_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

template <int MODE>
__global__ void __launch_bounds__(256, 1) k(double *out, long long *cyc, int iters)
{
  extern __shared__ double sm[];
  __shared__ unsigned tbase;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 8192; i += 256) sm[i] = i * 0.5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tbase)), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  // this warp's lanes: 32*(warp%4); warps 4..7 use the upper 256 columns
  const unsigned t0 = tbase + ((unsigned)((warp & 3) * 32) << 16) + (unsigned)((warp >> 2) * 256);
  for (int c = 0; c < 256; c += 2) tm_st2(t0 + c, (double)(tid * 1000 + c));
  tm_wait_st();
  __syncthreads();
  double acc = 0.0;
  long long t_start = clock64();
  for (int it = 0; it < iters; it++) {
    if (MODE == 0) {            // 24 TMEM loads, one wait
      unsigned lo[24], hi[24];
#pragma unroll
      for (int j = 0; j < 24; j++) tm_ld2_issue(t0 + ((it + j * 2) & 127) * 2, lo[j], hi[j]);
      tm_wait_ld();
#pragma unroll
      for (int j = 0; j < 24; j++) acc += __hiloint2double(hi[j], lo[j]);
    } else if (MODE == 1) {     // 24 LDS.64
#pragma unroll
      for (int j = 0; j < 24; j++) acc += sm[(tid + it + j * 37) & 8191];
    } else if (MODE == 2) {     // both interleaved
      unsigned lo[24], hi[24];
#pragma unroll
      for (int j = 0; j < 24; j++) tm_ld2_issue(t0 + ((it + j * 2) & 127) * 2, lo[j], hi[j]);
#pragma unroll
      for (int j = 0; j < 24; j++) acc += sm[(tid + it + j * 37) & 8191];
      tm_wait_ld();
#pragma unroll
      for (int j = 0; j < 24; j++) acc += __hiloint2double(hi[j], lo[j]);
    } else if (MODE == 3) {     // 24 TMEM stores
#pragma unroll
      for (int j = 0; j < 24; j++) tm_st2(t0 + ((it + j * 2) & 127) * 2, acc + j);
      tm_wait_st();
    } else if (MODE == 5) {     // 6 x ld.x8 (same 24 doubles), one wait
      unsigned r[6][8];
#pragma unroll
      for (int j = 0; j < 6; j++) tm_ld8_issue(t0 + ((it + j * 4) & 31) * 8, r[j]);
      tm_wait_ld();
#pragma unroll
      for (int j = 0; j < 6; j++)
#pragma unroll
        for (int q = 0; q < 4; q++) acc += __hiloint2double(r[j][2 * q + 1], r[j][2 * q]);
    } else if (MODE == 6) {     // 3 x ld.x16
      unsigned r[3][16];
#pragma unroll
      for (int j = 0; j < 3; j++) tm_ld16_issue(t0 + ((it + j * 4) & 15) * 16, r[j]);
      tm_wait_ld();
#pragma unroll
      for (int j = 0; j < 3; j++)
#pragma unroll
        for (int q = 0; q < 8; q++) acc += __hiloint2double(r[j][2 * q + 1], r[j][2 * q]);
    } else if (MODE == 7) {     // 6 x st.x8
#pragma unroll
      for (int j = 0; j < 6; j++) {
        unsigned r[8];
#pragma unroll
        for (int q = 0; q < 8; q++) r[q] = __double2loint(acc) + q + j;
        tm_st8(t0 + ((it + j * 4) & 31) * 8, r);
      }
      tm_wait_st();
    } else if (MODE == 4) {     // load-wait-use one at a time (latency)
#pragma unroll
      for (int j = 0; j < 24; j++) {
        unsigned lo, hi;
        tm_ld2_issue(t0 + ((it + j * 2) & 127) * 2, lo, hi);
        tm_wait_ld();
        acc += __hiloint2double(hi, lo);
      }
    }
  }
  long long t_end = clock64();
  out[blockIdx.x * 256 + tid] = acc;
  if (tid == 0) cyc[blockIdx.x] = t_end - t_start;
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tbase), "n"(512));
}

int main()
{
  double *out; long long *cyc;
  cudaMalloc(&out, 148 * 256 * 8); cudaMalloc(&cyc, 148 * 8);
  const int iters = 2000;
  cudaFuncSetAttribute(k<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<7>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  const char *names[] = {"24 tcgen05.ld.x2 + 1 wait", "24 LDS.64", "24 tcgen05.ld.x2 + 24 LDS.64", "24 tcgen05.st.x2 + wait", "24 x (ld.x2, wait, use)", "6 tcgen05.ld.x8 + wait (24 dbl)", "3 tcgen05.ld.x16 + wait (24 dbl)", "6 tcgen05.st.x8 + wait (24 dbl)"};
  for (int m = 0; m < 8; m++) {
    for (int rep = 0; rep < 2; rep++) {
      if (m == 0) k<0><<<148, 256, 65536>>>(out, cyc, iters);
      if (m == 1) k<1><<<148, 256, 65536>>>(out, cyc, iters);
      if (m == 2) k<2><<<148, 256, 65536>>>(out, cyc, iters);
      if (m == 3) k<3><<<148, 256, 65536>>>(out, cyc, iters);
      if (m == 4) k<4><<<148, 256, 65536>>>(out, cyc, iters);
      if (m == 5) k<5><<<148, 256, 65536>>>(out, cyc, iters);
      if (m == 6) k<6><<<148, 256, 65536>>>(out, cyc, iters);
      if (m == 7) k<7><<<148, 256, 65536>>>(out, cyc, iters);
    }
    cudaError_t e = cudaGetLastError(); if (e == cudaSuccess) e = cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double d0; cudaMemcpy(&d0, out + 5, 8, cudaMemcpyDeviceToHost);
    printf("%-34s: %8.1f cycles/iter per SM (8 warps), err=%s chk=%g\n", names[m], (double)h[0] / iters, cudaGetErrorString(e), d0);
  }
  return 0;
}
