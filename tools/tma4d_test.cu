// Stand-alone check of the 4-D TMA boxes rhs_dly uses: halo box {38,14,1,V} at aligned x and the
// interior box {32,8,1,3} at x0+3 (24-byte offset) of a (mx,my,mz,mvar) double tensor.
#include <cstdio>
#include <cuda.h>
#include <cuda_runtime.h>
#include <vector>
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__global__ void k(const __grid_constant__ CUtensorMap tm, int x, int y, int z, int v, int nbytes, double *out, int n)
{
  extern __shared__ __align__(128) unsigned char raw[];
  double *sm = reinterpret_cast<double *>(raw);
  __shared__ unsigned long long bar;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(nbytes) : "memory");
    asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
                 ::"r"(smem_u32(sm)), "l"(&tm), "r"(x), "r"(y), "r"(z), "r"(v), "r"(smem_u32(&bar)) : "memory");
  }
  unsigned ok = 0;
  for (int it = 0; it < (1 << 22) && !ok; it++)
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(ok) : "r"(smem_u32(&bar)) : "memory");
  if (threadIdx.x == 0) out[n] = ok ? 1.0 : -1.0;
  if (ok) for (int i = threadIdx.x; i < n; i += blockDim.x) out[i] = sm[i];
}
int main()
{
  const int mx = 38, my = 38, mz = 38, mvar = 4;
  const size_t mw = (size_t)mx * my * mz;
  std::vector<double> h(mw * mvar);
  for (size_t i = 0; i < h.size(); i++) h[i] = (double)i;
  double *d, *out; cudaMalloc(&d, h.size() * 8); cudaMalloc(&out, 65536 * 8);
  cudaMemcpy(d, h.data(), h.size() * 8, cudaMemcpyHostToDevice);
  typedef CUresult (*encode_t)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                               const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void *fn = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  encode_t encode = (encode_t)fn;
  cuuint64_t dims[4] = {mx, my, mz, mvar};
  cuuint64_t strides[3] = {(cuuint64_t)mx * 8, (cuuint64_t)mx * my * 8, (cuuint64_t)mw * 8};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  struct Case { cuuint32_t box[4]; int x, y, z, v; const char *name; } cases[] = {
    {{38, 14, 1, 4}, 0, 8, 5, 0, "halo box, all variables"},
    {{34, 8, 1, 3}, 2, 3, 5, 0, "box {34,8,1,3} at x0+2 (16-byte aligned), variables 0-2"},
    {{34, 8, 1, 1}, 2, 3, 5, 3, "box {34,8,1,1} at x0+2, variable 3"},
    {{32, 8, 1, 3}, 3, 3, 5, 0, "box {32,8,1,3} at x0+3 (24-byte offset: faults)"},
  };
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  for (auto &cs : cases) {
    CUtensorMap tm;
    CUresult r = encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 4, d, dims, strides, cs.box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    const int n = cs.box[0] * cs.box[1] * cs.box[2] * cs.box[3];
    k<<<1, 128, 65536>>>(tm, cs.x, cs.y, cs.z, cs.v, n * 8, out, n);
    cudaError_t e = cudaDeviceSynchronize();
    std::vector<double> o(n + 1, 0.0);
    if (e == cudaSuccess) cudaMemcpy(o.data(), out, (n + 1) * 8, cudaMemcpyDeviceToHost);
    long bad = 0;
    for (int i = 0; i < n && e == cudaSuccess; i++) {
      const int ix = i % cs.box[0], iy = (i / cs.box[0]) % cs.box[1], iv = i / (cs.box[0] * cs.box[1]);
      const double want = (double)((size_t)(cs.v + iv) * mw + (size_t)cs.z * mx * my + (size_t)(cs.y + iy) * mx + cs.x + ix);
      if (o[i] != want) bad++;
    }
    printf("%-55s encode=%d run=%s landed=%g mismatches=%ld\n", cs.name, (int)r, cudaGetErrorString(e), o[n], bad);
    if (e != cudaSuccess) return 1;
  }
  return 0;
}
