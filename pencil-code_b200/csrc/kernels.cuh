// kernels.cuh -- CUDA kernels of the solver core (sm_100a).
//
//  rhs_march<MAG,ENT,AUXM>   fused RHS + 2N-RK3 update, z-marching: each CTA
//      owns a 32x8 (x,y) tile and marches along z; ghost-padded z-planes are
//      staged in a shared-memory ring with cp.async (8 slots: 7 live + 1 in
//      flight), the thread's own z-column lives in a register sliding window.
//  rhs_global<MAG,ENT,AUXM>  same physics straight from global memory (L1/L2
//      cached); used for thin boundary slabs and as an in-library cross-check.
//  bc_*                      ghost-zone fills (boundcond.f90 semantics).
//  axpy_update               f_out = f_in + c*w   (substep 1 when dt comes from
//      the CFL reduction of the same substep, timestep.f90:148-163).
//  reduce_partials           deterministic second pass of the dt/diagnostic
//      reductions (no floating-point atomics).
#pragma once
#include "physics.cuh"
#include <cuda_runtime.h>

__constant__ DevPar c_par;

// reduction slots
enum { RD_DT1 = 0, RD_U2MAX, RD_B2MAX, RD_J2MAX, RD_DTC, RD_DTU, RD_DTNU, RD_DTCHI, NRMAX,
       RS_U2 = NRMAX, RS_RHO, RS_SS, RS_B2, RS_J2, RS_AB, RS_JB, RS_OU, NRED };

#define TX 32
#define TY 8
#define NTHR (TX * TY)
#define HX (TX + 2 * NG)
#define HY (TY + 2 * NG)
#define PE (HX * HY)          // elements of one staged plane tile
#define NSLOT 8

struct SubstepArgs {
  const double *f_in;
  double *w;
  double *f_out;
  double alpha;       // w = alpha*w + rhs   (alpha==0: w is not read)
  double beta;        // f_out = f_in + beta*dt*w  (timestep.f90:152,162)
  const double *dt_dev;   // dt lives on the device: no host round trip inside a step
  int write_w;        // 0 on the last substep
  int write_f;        // 0 on substep 1 when dt is not yet known
  double *partials;   // [nblocks][NRED] when AUXM>0
  int z_lo, z_hi;     // range of output planes (interior-relative, [z_lo,z_hi))
  int zchunk;
  // tile sub-range of a launch (x/y face exchange overlapped with interior tiles): the CTA's tile is
  // (tx0 + blockIdx.x*txs, ty0 + blockIdx.y*tys); ntx/nty = grid extent (0 = all tiles)
  int tx0, txs, ntx, ty0, tys, nty;
};

// ------------------------------------------------------------------------
// warp/block reduction of the auxiliary quantities (deterministic order)
// red[NRED] per thread; first NRMAX are maxima, the rest sums.
// AUXM==1: only RD_DT1 is reduced.
// ------------------------------------------------------------------------
template <int AUXM>
__device__ __forceinline__ void block_reduce_store(const double *red, int tid, int nthreads,
                                                   double *sm, double *partials, int block_linear)
{
  const int nq = AUXM == 1 ? 1 : NRED;
  const int lane = tid & 31, warp = tid >> 5;
  const int nwarp = (nthreads + 31) >> 5;
#pragma unroll
  for (int q = 0; q < NRED; q++) {
    if (q >= nq) break;
    double v = red[q];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double t = __shfl_xor_sync(0xffffffffu, v, o);
      v = q < NRMAX ? fmax(v, t) : v + t;
    }
    if (lane == 0) sm[warp * NRED + q] = v;
  }
  __syncthreads();
  if (tid < nq) {
    double v = sm[tid];
    for (int wv = 1; wv < nwarp; wv++) v = tid < NRMAX ? fmax(v, sm[wv * NRED + tid]) : v + sm[wv * NRED + tid];
    partials[(size_t)block_linear * NRED + tid] = v;
  }
}

__global__ void reduce_partials(const double *__restrict__ partials, int nblocks, double *__restrict__ out)
{
  // one block of NRED*32 threads: warp q reduces slot q in a fixed order
  const int q = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (q >= NRED) return;
  double v = 0.0;
  for (int b = lane; b < nblocks; b += 32) {
    const double t = partials[(size_t)b * NRED + q];
    v = q < NRMAX ? fmax(v, t) : v + t;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double t = __shfl_xor_sync(0xffffffffu, v, o);
    v = q < NRMAX ? fmax(v, t) : v + t;
  }
  if (lane == 0) out[q] = v;
}

// ------------------------------------------------------------------------
// accessor over global memory
// ------------------------------------------------------------------------
struct GlobalAcc {
  static constexpr bool PS = true, UPW = true, NOZMIX = false;
  const double *p[NROLE];
  long long sy, sz;
  bool bd;
  PC_DEV bool bidiag() const { return bd; }
  template <int R> PC_DEV double c() const { return __ldg(p[R]); }
  template <int R, int K> PC_DEV double x() const { return __ldg(p[R] + K); }
  template <int R, int K> PC_DEV double y() const { return __ldg(p[R] + K * sy); }
  template <int R, int K> PC_DEV double z() const { return __ldg(p[R] + K * sz); }
  template <int R, int KX, int KY> PC_DEV double xy() const { return __ldg(p[R] + KX + KY * sy); }
  template <int R, int KX, int KZ> PC_DEV double xz() const { return __ldg(p[R] + KX + KZ * sz); }
  template <int R, int KY, int KZ> PC_DEV double yz() const { return __ldg(p[R] + KY * sy + KZ * sz); }
};

template <bool MAG, bool ENT>
__device__ __forceinline__ void store_cell(const SubstepArgs &a, long long idx, const double *rhs)
{
  const DevPar &c = c_par;
#pragma unroll
  for (int r = 0; r < NROLE; r++) {
    if (!MAG && (r == R_AX || r == R_AY || r == R_AZ)) continue;
    if (!ENT && r == R_SS) continue;
    const long long i = idx + (long long)c.slot[r] * c.mw;
    double wv = rhs[r];
    if (a.alpha != 0.0) wv += a.alpha * a.w[i];
    if (a.write_w) a.w[i] = wv;
    if (a.write_f) a.f_out[i] = a.f_in[i] + (a.beta * a.dt_dev[0]) * wv;
  }
}

template <bool MAG, bool ENT, int AUXM>
__global__ void __launch_bounds__(256) rhs_global(SubstepArgs a)
{
  const DevPar &c = c_par;
  __shared__ double sm[8 * NRED];
  const int l = blockIdx.x * blockDim.x + threadIdx.x;
  const int m = blockIdx.y * blockDim.y + threadIdx.y;
  const int n = a.z_lo + blockIdx.z * blockDim.z + threadIdx.z;
  double red[NRED];
#pragma unroll
  for (int q = 0; q < NRED; q++) red[q] = 0.0;
  if (l < c.nx && m < c.ny && n < a.z_hi) {
    const long long idx = (long long)(n + NG) * c.sz + (long long)(m + NG) * c.sy + (l + NG);
    GlobalAcc acc;
    acc.sy = c.sy; acc.sz = c.sz; acc.bd = c.lbidiag != 0;
#pragma unroll
    for (int r = 0; r < NROLE; r++) acc.p[r] = c.slot[r] >= 0 ? a.f_in + (long long)c.slot[r] * c.mw + idx : a.f_in;
    double rhs[NROLE];
    CellAux aux;
    rhs_cell<MAG, ENT, (AUXM > 0)>(acc, c, n, rhs, aux);
    store_cell<MAG, ENT>(a, idx, rhs);
    if (AUXM > 0) {
      red[RD_DT1] = aux.dt1;
      if (AUXM > 1) {
        red[RD_U2MAX] = aux.u2; red[RD_B2MAX] = aux.b2; red[RD_J2MAX] = aux.j2;
        red[RD_DTC] = aux.dtc; red[RD_DTU] = aux.dtu; red[RD_DTNU] = aux.dtnu; red[RD_DTCHI] = aux.dtchi;
        red[RS_U2] = aux.u2; red[RS_RHO] = aux.rho; red[RS_SS] = aux.ss; red[RS_B2] = aux.b2; red[RS_J2] = aux.j2;
        red[RS_AB] = aux.ab; red[RS_JB] = aux.jb; red[RS_OU] = aux.ou;
      }
    }
  }
  if (AUXM > 0) {
    const int tid = threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
    const int bl = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
    block_reduce_store<AUXM>(red, tid, blockDim.x * blockDim.y * blockDim.z, sm, a.partials, bl);
  }
}

// ------------------------------------------------------------------------
// z-marching kernel with a shared-memory plane ring
// ------------------------------------------------------------------------
__device__ __forceinline__ void cp_async8(double *smem_dst, const double *gsrc)
{
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(d), "l"(gsrc));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }

// Roles R_UX..R_AZ (0..5) are "ring" fields: they need x/y neighbours on the
// planes z+-k (bidiagonal xz/yz mixed derivatives), so 7 planes are live.
// R_LNRHO,R_SS are "plain" fields: only the centre plane is staged (2 slots).
// shared-memory accessor: byte offsets in 32-bit registers, role/neighbour offsets as immediates
PC_DEV double lds_at(const double *base, unsigned byteoff, int imm_elems)
{ return *reinterpret_cast<const double *>(reinterpret_cast<const char *>(base) + byteoff + imm_elems * 8); }

template <bool MAG, bool ENT>
struct MarchAcc {
  static constexpr bool PS = true, UPW = true, NOZMIX = false;
  const double *smem0;  // dynamic shared memory: ring [ring role][NSLOT][PE], then plain [plain role][2][PE]
  unsigned sb[7];       // byte offset of this thread's cell in the ring slot of plane z+k, k=-3..3
  unsigned pb;          // byte offset of this thread's cell in the plain-field slot of plane z
  double w[NROLE][7];   // own column, planes z-3..z+3
  PC_DEV constexpr bool bidiag() const { return true; }
  template <int R, int OFF> PC_DEV double cenv() const
  { return R < 6 ? lds_at(smem0, sb[3], (R * NSLOT) * PE + OFF) : lds_at(smem0, pb, ((R - 6) * 2) * PE + OFF); }
  template <int R> PC_DEV double c() const { return w[R][3]; }
  template <int R, int K> PC_DEV double x() const { return cenv<R, K>(); }
  template <int R, int K> PC_DEV double y() const { return cenv<R, K * HX>(); }
  template <int R, int K> PC_DEV double z() const { return w[R][3 + K]; }
  template <int R, int KX, int KY> PC_DEV double xy() const { return cenv<R, KX + KY * HX>(); }
  template <int R, int KX, int KZ> PC_DEV double xz() const { return lds_at(smem0, sb[3 + KZ], (R * NSLOT) * PE + KX); }
  template <int R, int KY, int KZ> PC_DEV double yz() const { return lds_at(smem0, sb[3 + KZ], (R * NSLOT) * PE + KY * HX); }
};

template <bool MAG, bool ENT> struct MarchCfg {
  static constexpr int NRF = MAG ? 6 : 3;
  static constexpr int NPF = ENT ? 2 : 1;
  static constexpr size_t SMEM = sizeof(double) * ((size_t)NRF * NSLOT * PE + (size_t)NPF * 2 * PE + 8 * NRED);
};

template <bool MAG, bool ENT> __device__ __forceinline__ bool role_present(int r)
{
  if (r >= R_AX && r <= R_AZ) return MAG;
  if (r == R_SS) return ENT;
  return true;
}

template <bool MAG, bool ENT, int AUXM>
__global__ void __launch_bounds__(NTHR, 1) rhs_march(SubstepArgs a)
{
  const DevPar &c = c_par;
  extern __shared__ double smem[];
  double *ring = smem;
  double *plain = smem + (size_t)MarchCfg<MAG, ENT>::NRF * NSLOT * PE;
  double *redsm = plain + (size_t)MarchCfg<MAG, ENT>::NPF * 2 * PE;

  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * TX + tx;
  const int x0 = blockIdx.x * TX, y0 = blockIdx.y * TY;           // tile origin (interior coords == padded coords of the halo origin)
  const int zb = a.z_lo + blockIdx.z * a.zchunk;
  const int ze = min(zb + a.zchunk, a.z_hi);
  const bool active = (x0 + tx < c.nx) && (y0 + ty < c.ny);

  // per-thread share of a plane tile: up to 3 elements
  int eoff[3]; long long goff[3]; bool eok[3];
#pragma unroll
  for (int i = 0; i < 3; i++) {
    const int e = tid + i * NTHR;
    const int hy = e / HX, hx = e - hy * HX;
    eoff[i] = e;
    eok[i] = e < PE && (x0 + hx) < c.mx && (y0 + hy) < c.my;
    goff[i] = (long long)(y0 + hy) * c.sy + (x0 + hx);
  }
  // column offset of this thread in a plane
  const long long col = (long long)(y0 + ty + NG) * c.sy + (x0 + tx + NG);

  MarchAcc<MAG, ENT> acc;
  acc.smem0 = smem;
  const int tb = (ty + NG) * HX + tx + NG;
  constexpr int PLAIN0 = MarchCfg<MAG, ENT>::NRF * NSLOT * PE;

  // prologue: ring planes zb-3..zb+3 (padded zb..zb+6) into slots 0..6, plain plane zb into slot 0
#pragma unroll 1
  for (int q = 0; q < 7; q++) {
    const long long pz = (long long)(zb + q) * c.sz;
#pragma unroll
    for (int r = 0; r < 6; r++) {
      if (!role_present<MAG, ENT>(r)) continue;
      const double *src = a.f_in + c.foff[r] + pz;
#pragma unroll
      for (int i = 0; i < 3; i++) if (eok[i]) cp_async8(ring + ((size_t)r * NSLOT + q) * PE + eoff[i], src + goff[i]);
    }
  }
#pragma unroll
  for (int r = 6; r < NROLE; r++) {
    if (!role_present<MAG, ENT>(r)) continue;
    const double *src = a.f_in + c.foff[r] + (long long)(zb + NG) * c.sz;
#pragma unroll
    for (int i = 0; i < 3; i++) if (eok[i]) cp_async8(plain + ((size_t)(r - 6) * 2) * PE + eoff[i], src + goff[i]);
  }
  cp_async_commit();
  // register window: planes zb-3..zb+3 of the own column
  double wnext[NROLE];
#pragma unroll
  for (int r = 0; r < NROLE; r++) {
    if (!role_present<MAG, ENT>(r)) continue;
    const double *src = a.f_in + c.foff[r] + col;
#pragma unroll
    for (int k = 0; k < 7; k++) acc.w[r][k] = active ? __ldg(src + (long long)(zb + k) * c.sz) : 0.0;
  }
  double red[NRED];
#pragma unroll
  for (int q = 0; q < NRED; q++) red[q] = 0.0;
  const double beta_dt = a.write_f ? a.beta * a.dt_dev[0] : 0.0;
  cp_async_wait_all();
  __syncthreads();

#pragma unroll 1
  for (int n = zb; n < ze; n++) {
    const int q0 = n - zb;                       // ring plane index of z-3
    // 1. stage the next planes (ring: plane q0+7 == padded n+7 ; plain: padded n+NG+1)
    if (n + 1 < ze) {
      const int sl = (q0 + 7) & (NSLOT - 1);
      const long long pz = (long long)(n + 7) * c.sz;
#pragma unroll
      for (int r = 0; r < 6; r++) {
        if (!role_present<MAG, ENT>(r)) continue;
        const double *src = a.f_in + c.foff[r] + pz;
#pragma unroll
        for (int i = 0; i < 3; i++) if (eok[i]) cp_async8(ring + ((size_t)r * NSLOT + sl) * PE + eoff[i], src + goff[i]);
      }
      const int ps = (q0 + 1) & 1;
      const long long pz1 = (long long)(n + NG + 1) * c.sz;
#pragma unroll
      for (int r = 6; r < NROLE; r++) {
        if (!role_present<MAG, ENT>(r)) continue;
        const double *src = a.f_in + c.foff[r] + pz1;
#pragma unroll
        for (int i = 0; i < 3; i++) if (eok[i]) cp_async8(plain + ((size_t)(r - 6) * 2 + ps) * PE + eoff[i], src + goff[i]);
      }
      cp_async_commit();
#pragma unroll
      for (int r = 6; r < NROLE; r++) {
        if (!role_present<MAG, ENT>(r)) continue;
        wnext[r] = active ? __ldg(a.f_in + c.foff[r] + col + pz) : 0.0;
      }
    }
    // 2. compute the cell of plane n
#pragma unroll
    for (int k = 0; k < 7; k++) acc.sb[k] = (unsigned)((((q0 + k) & (NSLOT - 1)) * PE + tb) * 8);
    acc.pb = (unsigned)((PLAIN0 + (q0 & 1) * PE + tb) * 8);
    if (active) {
      const long long idx = (long long)(n + NG) * c.sz + col;
      // old 2N register: issued before the RHS so that its latency hides behind the stencils
      double wold[NROLE];
#pragma unroll
      for (int r = 0; r < NROLE; r++) {
        if (!role_present<MAG, ENT>(r)) continue;
        wold[r] = a.alpha != 0.0 ? __ldcs(a.w + idx + c.foff[r]) : 0.0;
      }
      double rhs[NROLE];
      CellAux aux;
      rhs_cell<MAG, ENT, (AUXM > 0)>(acc, c, n, rhs, aux);
      // 3. 2N-RK3 update (timestep.f90:119-122,162-163); f_in(cell) is the window centre
#pragma unroll
      for (int r = 0; r < NROLE; r++) {
        if (!role_present<MAG, ENT>(r)) continue;
        const long long i = idx + c.foff[r];
        const double wv = rhs[r] + a.alpha * wold[r];
        if (a.write_w) __stcs(a.w + i, wv);
        if (a.write_f) __stcs(a.f_out + i, acc.w[r][3] + beta_dt * wv);
      }
      if (AUXM > 0) {
        red[RD_DT1] = fmax(red[RD_DT1], aux.dt1);
        if (AUXM > 1) {
          red[RD_U2MAX] = fmax(red[RD_U2MAX], aux.u2); red[RD_B2MAX] = fmax(red[RD_B2MAX], aux.b2);
          red[RD_J2MAX] = fmax(red[RD_J2MAX], aux.j2);
          red[RD_DTC] = fmax(red[RD_DTC], aux.dtc); red[RD_DTU] = fmax(red[RD_DTU], aux.dtu);
          red[RD_DTNU] = fmax(red[RD_DTNU], aux.dtnu); red[RD_DTCHI] = fmax(red[RD_DTCHI], aux.dtchi);
          red[RS_U2] += aux.u2; red[RS_RHO] += aux.rho; red[RS_SS] += aux.ss; red[RS_B2] += aux.b2; red[RS_J2] += aux.j2;
          red[RS_AB] += aux.ab; red[RS_JB] += aux.jb; red[RS_OU] += aux.ou;
        }
      }
    }
    // 4. retire the oldest plane, admit the new one
    cp_async_wait_all();
    __syncthreads();
#pragma unroll
    for (int r = 0; r < NROLE; r++) {
      if (!role_present<MAG, ENT>(r)) continue;
#pragma unroll
      for (int k = 0; k < 6; k++) acc.w[r][k] = acc.w[r][k + 1];
      if (r < 6) acc.w[r][6] = ring[((size_t)r * NSLOT + ((q0 + 7) & (NSLOT - 1))) * PE + tb];   // plane z+4 has just landed
      else acc.w[r][6] = wnext[r];
    }
  }
  if (AUXM > 0) {
    const int bl = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
    block_reduce_store<AUXM>(red, tid, NTHR, redsm, a.partials, bl);
  }
}

// ------------------------------------------------------------------------
// streaming update for substep 1 when dt comes out of the same substep:
// f_out = f_in + c*w on the interior (timestep.f90:162-163)
// ------------------------------------------------------------------------
__global__ void axpy_update(const double *__restrict__ f_in, const double *__restrict__ w, double *__restrict__ f_out,
                            const double *__restrict__ dt_dev, double beta)
{
  const DevPar &c = c_par;
  const int l = blockIdx.x * blockDim.x + threadIdx.x;
  const int m = blockIdx.y;
  const int nv = blockIdx.z;               // n + nz*var
  const int n = nv % c.nz, k = nv / c.nz;
  if (l >= c.nx) return;
  const long long i = (long long)k * c.mw + (long long)(n + NG) * c.sz + (long long)(m + NG) * c.sy + (l + NG);
  f_out[i] = f_in[i] + (beta * dt_dev[0]) * w[i];
}

// dt = 1/max(dt1, 1/dtmax) (equ.f90:280-290, sub.f90:3911) from the reduced partials
// extra: optional second contribution to dt1_max (the particle CFL terms)
__global__ void finish_dt(const double *__restrict__ red, double dt1_floor, double *__restrict__ dt_dev, const double *__restrict__ extra)
{
  if (threadIdx.x == 0) {
    double d = fmax(red[RD_DT1], dt1_floor);
    if (extra) d = fmax(d, extra[0]);
    dt_dev[1] = d; dt_dev[0] = 1.0 / d;
  }
}
__global__ void set_dt_from_dt1(double *__restrict__ dt_dev)   // after the cross-rank max of dt_dev[1]
{
  if (threadIdx.x == 0) dt_dev[0] = 1.0 / dt_dev[1];
}

// ------------------------------------------------------------------------
// helical forcing kick (forcing.f90:1968-2110 with default switches)
// ------------------------------------------------------------------------
struct ForcingArgs { double coef1[3], coef2[3], fda[3], kk[3], phase, fact, x0, y0, z0, dx, dy, dz; };
__global__ void forcing_kick(double *__restrict__ f, ForcingArgs g)
{
  const DevPar &c = c_par;
  const int l = blockIdx.x * blockDim.x + threadIdx.x, m = blockIdx.y, n = blockIdx.z;
  if (l >= c.nx) return;
  double sx, cx, sy, cy, sz, cz;
  sincos(g.kk[0] * (g.x0 + l * g.dx) + g.phase, &sx, &cx);
  sincos(g.kk[1] * (g.y0 + m * g.dy), &sy, &cy);
  sincos(g.kk[2] * (g.z0 + n * g.dz), &sz, &cz);
  const double fxr = cx * g.fact, fxi = sx * g.fact;
  const double pr = fxr * cy - fxi * sy, pi = fxr * sy + fxi * cy;
  const double qr = pr * cz - pi * sz, qi = pr * sz + pi * cz;
  const long long idx = (long long)(n + NG) * c.sz + (long long)(m + NG) * c.sy + (l + NG);
#pragma unroll
  for (int j = 0; j < 3; j++) {
    const long long i = idx + (long long)c.slot[R_UX + j] * c.mw;
    f[i] += g.fda[j] * (g.coef1[j] * qr - g.coef2[j] * qi);
  }
}

// ------------------------------------------------------------------------
// boundary conditions: one thread per ghost line, all variables and both
// faces in the reference order (bot: var 1..mvar, top: var 1..mvar;
// boundcond.f90:1303-1316).  axis 0/1/2 = x/y/z; lines span the FULL extent
// of the two other axes.  do_face[f]==0 skips a face (interior process
// boundary: ghosts come from the neighbour rank instead).
// ------------------------------------------------------------------------
struct BcArgs {
  double *f;
  int axis;
  int do_face[2];
  int code[2][PCB200_MAXVAR];
  int ilnrho, iss;             // storage slots
  double cp, cv, cv1, gamma_m1, cs20, lnrho0, cs2top, Fbot, FbotKbot, chi;
  int chiconst;
  double dz2_bound[7];
};
__global__ void bc_fill(BcArgs b)
{
  const DevPar &c = c_par;
  const int mm[3] = {c.mx, c.my, c.mz};
  const int nn[3] = {c.nx, c.ny, c.nz};
  const long long str[3] = {1, c.sy, c.sz};
  const int ax = b.axis, bx = ax == 0 ? 1 : 0, cx = ax == 2 ? 1 : 2;   // b = fastest remaining axis
  const int ib = blockIdx.x * blockDim.x + threadIdx.x, ic = blockIdx.y;
  if (ib >= mm[bx] || ic >= mm[cx]) return;
  const long long sa = str[ax];
  const int n1 = NG, n2 = NG + nn[ax] - 1;
  const long long line = ib * str[bx] + ic * str[cx];
  for (int face = 0; face < 2; face++) {
    if (!b.do_face[face]) continue;
    for (int k = 0; k < c.mvar; k++) {
      double *p = b.f + (long long)k * c.mw + line;
      const int code = b.code[face][k];
      if (code == PCB200_BC_P) {
        if (face == 0) { for (int i = 0; i < NG; i++) p[i * sa] = p[(n2 - NG + 1 + i) * sa]; }
        else           { for (int i = 0; i < NG; i++) p[(n2 + 1 + i) * sa] = p[(n1 + i) * sa]; }
      } else if (code == PCB200_BC_S || code == PCB200_BC_A) {
        const double sgn = code == PCB200_BC_S ? 1.0 : -1.0;
        if (face == 0) { for (int i = 1; i <= NG; i++) p[(n1 - i) * sa] = sgn * p[(n1 + i) * sa]; if (sgn < 0) p[n1 * sa] = 0.0; }
        else           { for (int i = 1; i <= NG; i++) p[(n2 + i) * sa] = sgn * p[(n2 - i) * sa]; if (sgn < 0) p[n2 * sa] = 0.0; }
      } else if (code == PCB200_BC_A2) {
        if (face == 0) { for (int i = 1; i <= NG; i++) p[(n1 - i) * sa] = 2.0 * p[n1 * sa] - p[(n1 + i) * sa]; }
        else           { for (int i = 1; i <= NG; i++) p[(n2 + i) * sa] = p[n2 * sa] + (p[n2 * sa] - p[(n2 - i) * sa]); }
      } else if (code == PCB200_BC_C1 && ax == 2 && face == 0) {
        const double *lnr = b.f + (long long)b.ilnrho * c.mw + line;
        const double rho_xy = exp(lnr[n1 * sa]);
        const double cs2_xy = b.cs20 * exp(b.gamma_m1 * (lnr[n1 * sa] - b.lnrho0) + b.cv1 * p[n1 * sa]);
        const double tmp_xy = b.chiconst ? b.Fbot / (rho_xy * b.chi * cs2_xy) : b.FbotKbot / cs2_xy;
        for (int i = 1; i <= NG; i++) {
          const double dlnrho = lnr[(n1 + i) * sa] - lnr[(n1 - i) * sa];
          p[(n1 - i) * sa] = p[(n1 + i) * sa] + b.cp * (b.cp - b.cv) * (dlnrho + b.dz2_bound[3 - i] * tmp_xy);
        }
      } else if (code == PCB200_BC_CT && ax == 2 && face == 1) {
        const double *lnr = b.f + (long long)b.ilnrho * c.mw + line;
        const double tmp = 2.0 * b.cv * log(b.cs2top / b.cs20);
        p[n2 * sa] = 0.5 * tmp - (b.cp - b.cv) * (lnr[n2 * sa] - b.lnrho0);
        for (int i = 1; i <= NG; i++)
          p[(n2 + i) * sa] = -p[(n2 - i) * sa] + tmp - (b.cp - b.cv) * (lnr[(n2 - i) * sa] + lnr[(n2 + i) * sa] - 2.0 * b.lnrho0);
      }
    }
  }
}

// Face packing for the x and y halo exchange (isendrcv_bdry_x, mpicomm.f90:2049-2101, and the
// y buffers lbufyo/ubufyo, :1592-1690): NG planes of one axis over the FULL ghost-inclusive
// extent of the other two, so that the sequential x -> y -> z exchange carries edges and
// corners without separate corner messages.  buf layout [var][c][b][NG planes] (axis 0) or
// [var][iz][NG][mx] (axis 1).  unpack=1 scatters buf back into f.
struct FaceArgs { double *f; double *buf; int axis, start, unpack, nplanes; };   // unpack: 0 pack, 1 copy back, 2 add back
__global__ void face_copy(FaceArgs a)
{
  const DevPar &c = c_par;
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int np = a.nplanes;
  const long long per_var = a.axis == 0 ? (long long)np * c.my * c.mz : (a.axis == 1 ? (long long)np * c.mx * c.mz : (long long)np * c.mx * c.my);
  if (t >= per_var) return;
  int ix, iy, iz;
  if (a.axis == 0) { ix = a.start + (int)(t % np); const long long r = t / np; iy = (int)(r % c.my); iz = (int)(r / c.my); }
  else if (a.axis == 1) { ix = (int)(t % c.mx); const long long r = t / c.mx; iy = a.start + (int)(r % np); iz = (int)(r / np); }
  else { ix = (int)(t % c.mx); const long long r = t / c.mx; iy = (int)(r % c.my); iz = a.start + (int)(r / c.my); }
  double *p = a.f + (long long)blockIdx.y * c.mw + (long long)iz * c.sz + (long long)iy * c.sy + ix;
  double *q = a.buf + (long long)blockIdx.y * per_var + t;
  if (a.unpack == 0) *q = *p; else if (a.unpack == 1) *p = *q; else *p += *q;
}
// Both faces of one axis in one launch (x or y halo exchange): blockIdx.y = variable, threads over the lower
// face first, then the upper one; 32-bit index arithmetic (a face has < 2^31 cells; the 64-bit divisions of
// face_copy cost more than the copy).  Same buffer layouts as face_copy.  pack: lower/upper INTERIOR planes
// (start NG and n) -> buf_lo/buf_hi; unpack: buf_lo -> lower ghosts (start 0), buf_hi -> upper ghosts (start NG+n).
struct Face2Args { double *f; double *buf_lo; double *buf_hi; int axis, n, unpack, do_lo, do_hi; };
__global__ void face_copy2(Face2Args a)
{
  const DevPar &c = c_par;
  const unsigned mx = c.mx, my = c.my, mz = c.mz;
  const unsigned per_var = a.axis == 0 ? NG * my * mz : NG * mx * mz;
  unsigned t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= 2 * per_var) return;
  const bool hi = t >= per_var;
  if (hi) t -= per_var;
  if (hi ? !a.do_hi : !a.do_lo) return;
  const unsigned start = a.unpack ? (hi ? NG + a.n : 0) : (hi ? a.n : NG);
  unsigned ix, iy, iz;
  if (a.axis == 0) { const unsigned r = t / NG; ix = start + (t - r * NG); iz = r / my; iy = r - iz * my; }
  else { const unsigned r = t / mx; ix = t - r * mx; iz = r / NG; iy = start + (r - iz * NG); }
  double *p = a.f + (long long)blockIdx.y * c.mw + (long long)iz * c.sz + (long long)iy * c.sy + ix;
  double *q = (hi ? a.buf_hi : a.buf_lo) + (size_t)blockIdx.y * per_var + t;
  if (a.unpack) *p = *q; else *q = *p;
}
// zero one plane (all blockIdx.y fields) -- the folded ghost layer (ghostfold_mpicomm.f90:69-70)
__global__ void plane_zero(double *f, int axis, int plane)
{
  const DevPar &c = c_par;
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long per = axis == 0 ? (long long)c.my * c.mz : (axis == 1 ? (long long)c.mx * c.mz : (long long)c.mx * c.my);
  if (t >= per) return;
  int ix, iy, iz;
  if (axis == 0) { ix = plane; iy = (int)(t % c.my); iz = (int)(t / c.my); }
  else if (axis == 1) { ix = (int)(t % c.mx); iy = plane; iz = (int)(t / c.mx); }
  else { ix = (int)(t % c.mx); iy = (int)(t / c.mx); iz = plane; }
  f[(long long)blockIdx.y * c.mw + (long long)iz * c.sz + (long long)iy * c.sy + ix] = 0.0;
}

// All locally periodic directions in one launch (bc_per_x/y/z, boundcond.f90:1813-1961).
// A ghost cell takes the value of the interior cell at the wrapped coordinates, which is what
// the sequential x -> y -> z sweeps produce at edges and corners.  Sources are interior in
// every wrapped direction and destinations are ghost in at least one, so there is no race.
struct BcPerArgs { double *f; int wrap[3]; long long n0, n1, n2; };
__global__ void bc_periodic(BcPerArgs b)
{
  const DevPar &c = c_par;
  // 32-bit index arithmetic (the three ghost shells have < 2^31 cells each; 64-bit divisions cost more than the copy)
  unsigned t = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned n0 = (unsigned)b.n0, n1 = (unsigned)b.n1, n2 = (unsigned)b.n2, mx = c.mx, my = c.my;
  int ix, iy, iz;
  if (t < n0) {                                     // x ghosts: 6 x my x mz
    const unsigned gx = t % 6u, r = t / 6u;
    iy = (int)(r % my); iz = (int)(r / my); ix = gx < NG ? (int)gx : c.nx + (int)gx;
  } else if ((t -= n0) < n1) {                      // y ghosts: mx x 6 x mz
    ix = (int)(t % mx); const unsigned r = t / mx;
    const unsigned gy = r % 6u; iz = (int)(r / 6u); iy = gy < NG ? (int)gy : c.ny + (int)gy;
  } else if ((t -= n1) < n2) {                      // z ghosts: mx x my x 6
    ix = (int)(t % mx); const unsigned r = t / mx;
    iy = (int)(r % my); const unsigned gz = r / my; iz = gz < NG ? (int)gz : c.nz + (int)gz;
  } else return;
  const int sx = !b.wrap[0] ? ix : (ix < NG ? ix + c.nx : (ix >= NG + c.nx ? ix - c.nx : ix));
  const int sy = !b.wrap[1] ? iy : (iy < NG ? iy + c.ny : (iy >= NG + c.ny ? iy - c.ny : iy));
  const int sz = !b.wrap[2] ? iz : (iz < NG ? iz + c.nz : (iz >= NG + c.nz ? iz - c.nz : iz));
  double *p = b.f + (long long)blockIdx.y * c.mw;
  p[(long long)iz * c.sz + (long long)iy * c.sy + ix] = p[(long long)sz * c.sz + (long long)sy * c.sy + sx];
}


// ------------------------------------------------------------------------
// rhs_tma: the production z-marching kernel.  Same scheme as rhs_march, but
//  * plane tiles are fetched by the TMA unit (cp.async.bulk.tensor.3d, one
//    box {HX,HY,1} per field and plane, completion on an mbarrier) instead of
//    8-byte cp.async issued by every thread: frees the LSU queue (no
//    lg_throttle), zero-fills out-of-range halo at the domain edge;
//  * whether the 2N register w is read is a template parameter (RW), so its
//    prefetch is an unconditional load that the compiler can keep in flight;
//  * edge tiles clamp their column instead of predicating the whole body.
// The global tensor is f viewed as 3-D (mx, my, mz*mvar).
// ------------------------------------------------------------------------
#include <cuda.h>

#ifndef PCB200_WPF
#define PCB200_WPF (MAG && ENT)
#endif
#define PEP 544                 // plane tile padded to a multiple of 128 B (TMA smem alignment)

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count)
{ asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count)); }
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes)
{ asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory"); }
__device__ __forceinline__ bool mbar_try_wait(unsigned long long *bar, unsigned parity)
{
  unsigned ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ bool mbar_test_wait(unsigned long long *bar, unsigned parity)   // non-blocking probe
{
  unsigned ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity)
{
  // bounded spin: a protocol error traps instead of hanging the GPU (not unrolled: the copies of the
  // poll made up a quarter of the z-loop body of rhs_dly, which has to stay in the instruction cache)
#pragma unroll 1
  for (unsigned it = 0; it < (1u << 28); it++) if (mbar_try_wait(bar, parity)) return;
  __trap();
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *bar)
{ asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory"); }
__device__ __forceinline__ void tma_load_3d(void *dst, const CUtensorMap *tm, int x, int y, int z, unsigned long long *bar)
{
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
               ::"r"(smem_u32(dst)), "l"(tm), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar)) : "memory");
}

template <bool MAG, bool ENT, bool BD = true>
struct TmaAcc {
  static constexpr bool UPW = true, NOZMIX = false;
  static constexpr bool PS = !(MAG && ENT);     // see physics.cuh: the 8-variable kernel spills more with the pre-scaled form
  PC_DEV constexpr bool bidiag() const { return BD; }
  const double *smem0;  // ring [ring role][NSLOT][PEP], then plain [plain role][2][PEP]
  unsigned sb[7];       // byte offset of this thread's cell in the ring slot of plane z+k, k=-3..3
  unsigned pb;          // byte offset of this thread's cell in the plain-field slot of plane z
  double w[NROLE][7];   // own column, planes z-3..z+3
  template <int R, int OFF> PC_DEV double cenv() const
  { return R < 6 ? lds_at(smem0, sb[3], (R * NSLOT) * PEP + OFF) : lds_at(smem0, pb, ((R - 6) * 2) * PEP + OFF); }
  template <int R> PC_DEV double c() const { return w[R][3]; }
  template <int R, int K> PC_DEV double x() const { return cenv<R, K>(); }
  template <int R, int K> PC_DEV double y() const { return cenv<R, K * HX>(); }
  template <int R, int K> PC_DEV double z() const { return w[R][3 + K]; }
  template <int R, int KX, int KY> PC_DEV double xy() const { return cenv<R, KX + KY * HX>(); }
  template <int R, int KX, int KZ> PC_DEV double xz() const { return lds_at(smem0, sb[3 + KZ], (R * NSLOT) * PEP + KX); }
  template <int R, int KY, int KZ> PC_DEV double yz() const { return lds_at(smem0, sb[3 + KZ], (R * NSLOT) * PEP + KY * HX); }
};

template <bool MAG, bool ENT> struct TmaCfg {
  static constexpr int NRF = MAG ? 6 : 3;
  static constexpr int NPF = ENT ? 2 : 1;
  static constexpr int NDBL = NRF * NSLOT * PEP + NPF * 2 * PEP;          // field tiles
  static constexpr size_t SMEM = sizeof(double) * ((size_t)NDBL + 8 * NRED + 2) + 128;   // + reduction scratch + mbarrier + alignment slack
};

template <bool MAG, bool ENT, int AUXM, bool RW, bool BD = true>
__global__ void __launch_bounds__(NTHR, 1) rhs_tma(SubstepArgs a, const __grid_constant__ CUtensorMap tmap)
{
  const DevPar &c = c_par;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  // align the tile area to 128 B (TMA destination requirement)
  // (offset arithmetic on the __shared__ symbol itself, so that the accesses stay LDS rather than generic LD)
  double *smem = reinterpret_cast<double *>(smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u));
  constexpr int NRF = TmaCfg<MAG, ENT>::NRF, NPF = TmaCfg<MAG, ENT>::NPF;
  double *ring = smem;
  double *plain = smem + (size_t)NRF * NSLOT * PEP;
  double *redsm = plain + (size_t)NPF * 2 * PEP;
  unsigned long long *bar = reinterpret_cast<unsigned long long *>(redsm + 8 * NRED);
  constexpr unsigned BOX_BYTES = PE * 8;
  constexpr bool WPF = PCB200_WPF;     // old 2N register: L2 prefetch + late load instead of an early load held in registers

  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * TX + tx;
  const int x0 = (a.tx0 + blockIdx.x * a.txs) * TX, y0 = (a.ty0 + blockIdx.y * a.tys) * TY;
  const int zb = a.z_lo + blockIdx.z * a.zchunk;
  const int ze = min(zb + a.zchunk, a.z_hi);
  const bool active = (x0 + tx < c.nx) && (y0 + ty < c.ny);
  // edge tiles: inactive threads shadow the last valid cell (loads stay in range, stores are masked)
  const int lx = min(x0 + tx, c.nx - 1), ly = min(y0 + ty, c.ny - 1);
  const long long col = (long long)(ly + NG) * c.sy + (lx + NG);
  const int tb = (ty + NG) * HX + tx + NG;
  const bool ww = a.write_w != 0 && active, wf = a.write_f != 0 && active;

  // bar[0] = "full": phase 0 is the prologue load, phase q+1 the planes issued in step q.
  // bar[1] = "empty": phase q completes when all NTHR threads are done with step q.
  // No __syncthreads in the z loop: warps drift up to one plane apart.
  unsigned long long *full = bar, *empty = bar + 1;
  if (tid == 0) {
    mbar_init(full, 1);
    mbar_init(empty, NTHR);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  // prologue: ring planes zb..zb+6 (padded) -> slots 0..6, plain plane zb+NG -> slot 0
  if (tid == 0) {
    mbar_expect_tx(full, (unsigned)((NRF * 7 + NPF) * BOX_BYTES));
#pragma unroll 1
    for (int q = 0; q < 7; q++)
#pragma unroll
      for (int r = 0; r < 6; r++) {
        if (!role_present<MAG, ENT>(r)) continue;
        tma_load_3d(ring + ((size_t)r * NSLOT + q) * PEP, &tmap, x0, y0, c.slot[r] * c.mz + zb + q, full);
      }
#pragma unroll
    for (int r = 6; r < NROLE; r++) {
      if (!role_present<MAG, ENT>(r)) continue;
      tma_load_3d(plain + ((size_t)(r - 6) * 2) * PEP, &tmap, x0, y0, c.slot[r] * c.mz + zb + NG, full);
    }
  }
  TmaAcc<MAG, ENT, BD> acc;
  acc.smem0 = smem;
  double wnext[NROLE];
#pragma unroll
  for (int r = 0; r < NROLE; r++) {
    if (!role_present<MAG, ENT>(r)) continue;
    const double *src = a.f_in + c.foff[r] + col;
#pragma unroll
    for (int k = 0; k < 7; k++) acc.w[r][k] = __ldg(src + (long long)(zb + k) * c.sz);
  }
  double red[NRED];
#pragma unroll
  for (int q = 0; q < NRED; q++) red[q] = 0.0;
  const double beta_dt = a.write_f ? a.beta * a.dt_dev[0] : 0.0;
  mbar_wait(full, 0);
  // stage the planes of step q0+1 (ring: padded plane n+7, plain: padded plane n+NG+1); they
  // overwrite the slots of step q0-1, so the elected thread first waits for "empty" phase q0-1
  auto issue = [&](int n, int q0) {
    mbar_expect_tx(full, (unsigned)((NRF + NPF) * BOX_BYTES));
    const int sl = (q0 + 7) & (NSLOT - 1);
#pragma unroll
    for (int r = 0; r < 6; r++) {
      if (!role_present<MAG, ENT>(r)) continue;
      tma_load_3d(ring + ((size_t)r * NSLOT + sl) * PEP, &tmap, x0, y0, c.slot[r] * c.mz + n + 7, full);
    }
    const int ps = (q0 + 1) & 1;
#pragma unroll
    for (int r = 6; r < NROLE; r++) {
      if (!role_present<MAG, ENT>(r)) continue;
      tma_load_3d(plain + ((size_t)(r - 6) * 2 + ps) * PEP, &tmap, x0, y0, c.slot[r] * c.mz + n + NG + 1, full);
    }
  };

#pragma unroll 1
  for (int n = zb; n < ze; n++) {
    const int q0 = n - zb;
    const bool more = n + 1 < ze;
    bool issued = true;
    if (more) {
      if (tid == 0) {
        // try once without blocking; if the slowest warp is still in step q0-1, retry after the stencils
        issued = q0 == 0 || mbar_test_wait(empty, (unsigned)((q0 - 1) & 1));
        if (issued) issue(n, q0);
      }
#pragma unroll
      for (int r = 6; r < NROLE; r++) {
        if (!role_present<MAG, ENT>(r)) continue;
        wnext[r] = __ldg(a.f_in + c.foff[r] + col + (long long)(n + 7) * c.sz);
      }
    }
    const long long idx = (long long)(n + NG) * c.sz + col;
    // old 2N register: unconditional loads, consumed only after the stencils
    double wold[NROLE];
    if (RW) {
#pragma unroll
      for (int r = 0; r < NROLE; r++) {
        if (!role_present<MAG, ENT>(r)) continue;
        if (WPF) asm volatile("prefetch.global.L2 [%0];" ::"l"(a.w + idx + c.foff[r]));
        else wold[r] = __ldcs(a.w + idx + c.foff[r]);
      }
    }
    // 2. the cell of plane n
#pragma unroll
    for (int k = 0; k < 7; k++) acc.sb[k] = (unsigned)((((q0 + k) & (NSLOT - 1)) * PEP + tb) * 8);
    acc.pb = (unsigned)((NRF * NSLOT * PEP + (q0 & 1) * PEP + tb) * 8);
    double rhs[NROLE];
    CellAux aux;
    rhs_cell<MAG, ENT, (AUXM > 0)>(acc, c, n, rhs, aux);
    if (tid == 0 && !issued) { mbar_wait(empty, (unsigned)((q0 - 1) & 1)); issue(n, q0); }
    if (RW && WPF) {      // the lines were prefetched into L2 before the stencils: no registers held across them
#pragma unroll
      for (int r = 0; r < NROLE; r++) {
        if (!role_present<MAG, ENT>(r)) continue;
        wold[r] = __ldcs(a.w + idx + c.foff[r]);
      }
    }
    // 3. 2N-RK3 update (timestep.f90:119-122,162-163); f_in(cell) is the window centre
#pragma unroll
    for (int r = 0; r < NROLE; r++) {
      if (!role_present<MAG, ENT>(r)) continue;
      const long long i = idx + c.foff[r];
      const double wv = RW ? rhs[r] + a.alpha * wold[r] : rhs[r];
      if (ww) __stcs(a.w + i, wv);
      if (wf) __stcs(a.f_out + i, acc.w[r][3] + beta_dt * wv);
    }
    if (AUXM > 0 && active) {
      red[RD_DT1] = fmax(red[RD_DT1], aux.dt1);
      if (AUXM > 1) {
        red[RD_U2MAX] = fmax(red[RD_U2MAX], aux.u2); red[RD_B2MAX] = fmax(red[RD_B2MAX], aux.b2);
        red[RD_J2MAX] = fmax(red[RD_J2MAX], aux.j2);
        red[RD_DTC] = fmax(red[RD_DTC], aux.dtc); red[RD_DTU] = fmax(red[RD_DTU], aux.dtu);
        red[RD_DTNU] = fmax(red[RD_DTNU], aux.dtnu); red[RD_DTCHI] = fmax(red[RD_DTCHI], aux.dtchi);
        red[RS_U2] += aux.u2; red[RS_RHO] += aux.rho; red[RS_SS] += aux.ss; red[RS_B2] += aux.b2; red[RS_J2] += aux.j2;
        red[RS_AB] += aux.ab; red[RS_JB] += aux.jb; red[RS_OU] += aux.ou;
      }
    }
    // 4. next planes have landed -> shift the z window, then release step q0's slots
    if (!more) break;
    mbar_wait(full, (unsigned)((q0 + 1) & 1));
#pragma unroll
    for (int r = 0; r < NROLE; r++) {
      if (!role_present<MAG, ENT>(r)) continue;
#pragma unroll
      for (int k = 0; k < 6; k++) acc.w[r][k] = acc.w[r][k + 1];
      if (r < 6) acc.w[r][6] = ring[((size_t)r * NSLOT + ((q0 + 7) & (NSLOT - 1))) * PEP + tb];
      else acc.w[r][6] = wnext[r];
    }
    mbar_arrive(empty);
  }
  if (AUXM > 0) {
    const int bl = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
    block_reduce_store<AUXM>(red, tid, NTHR, redsm, a.partials, bl);
  }
}

// ------------------------------------------------------------------------
// cfl_prepass: dt1_max of the state at the beginning of a time step (isothermal runs).
// The reference reduces dt1_max over the RHS evaluation of substep 1 and only then updates f
// (timestep.f90:141-163), which costs a separate axpy pass over f and w.  Without entropy the
// CFL terms (equ.f90:1086-1151) need only |u|, cs2, rho1 and B = curl A (hydro.f90:3810,
// noentropy.f90:367, magnetic.f90:5438, viscosity/resistivity constants), so a light z-marching
// kernel computes them first (56 B/cell read, 6 derivatives) and substep 1 applies the update
// itself.  Same per-cell formula as hydro_pre/hydro_finish (AUX part); partials[block][RD_DT1].
// ------------------------------------------------------------------------
struct CflAcc {
  static constexpr bool PS = true, UPW = true, NOZMIX = false;
  const double *sm;     // [3 A fields][2 slots][PE]
  int cb;               // element offset of this thread's cell in the centre slot
  double wx[7], wy[7];  // own column of Ax, Ay, planes z-3..z+3
  template <int R, int K> PC_DEV double x() const { return sm[(R - R_AX) * 2 * PE + cb + K]; }
  template <int R, int K> PC_DEV double y() const { return sm[(R - R_AX) * 2 * PE + cb + K * HX]; }
  template <int R, int K> PC_DEV double z() const { return R == R_AX ? wx[3 + K] : wy[3 + K]; }
};

__device__ __forceinline__ void cp_async16(double *smem_dst, const double *gsrc)
{
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gsrc));
}

template <bool MAG>
__global__ void __launch_bounds__(NTHR, 3) cfl_prepass(const double *__restrict__ f, double *__restrict__ partials, int z_lo, int z_hi, int zchunk)
{
  const DevPar &c = c_par;
  __shared__ __align__(16) double sm[MAG ? 3 * 2 * PE : 2];
  __shared__ double redsm[8 * NRED];
  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * TX + tx;
  const int x0 = blockIdx.x * TX, y0 = blockIdx.y * TY;
  const int zb = z_lo + blockIdx.z * zchunk;
  const int ze = min(zb + zchunk, z_hi);
  const bool active = (x0 + tx < c.nx) && (y0 + ty < c.ny);
  const int lx = min(x0 + tx, c.nx - 1), ly = min(y0 + ty, c.ny - 1);
  const long long col = (long long)(ly + NG) * c.sy + (lx + NG);
  // this thread's share of a halo tile: pairs of doubles (16-byte copies) when every row start is
  // 16-byte aligned (mx even; x0 and HX are even), single doubles otherwise
  const bool pairs = (c.mx & 1) == 0;
  const int per = pairs ? 2 : 1;
  int eoff[3]; long long goff[3]; bool eok[3];
#pragma unroll
  for (int i = 0; i < 3; i++) {
    const int e = (tid + i * NTHR) * per;
    const int hy = e / HX, hx = e - hy * HX;
    eoff[i] = e;
    eok[i] = e < PE && (x0 + hx + per - 1) < c.mx && (y0 + hy) < c.my;
    goff[i] = (long long)(y0 + hy) * c.sy + (x0 + hx);
  }
  auto stage = [&](int n) {       // centre-plane tiles of interior plane n into slot n&1
    if (MAG) {
      const long long pz = (long long)(n + NG) * c.sz;
#pragma unroll
      for (int r = 0; r < 3; r++)
#pragma unroll
        for (int i = 0; i < 3; i++) {
          if (!eok[i]) continue;
          double *dst = sm + (r * 2 + (n & 1)) * PE + eoff[i];
          const double *src = f + c.foff[R_AX + r] + pz + goff[i];
          if (pairs) cp_async16(dst, src); else cp_async8(dst, src);
        }
    }
    cp_async_commit();
  };
  CflAcc acc;
  acc.sm = sm;
  const double *pu = f + col;
  if (MAG) {
#pragma unroll
    for (int k = 0; k < 7; k++) {
      acc.wx[k] = __ldg(pu + c.foff[R_AX] + (long long)(zb + k) * c.sz);
      acc.wy[k] = __ldg(pu + c.foff[R_AY] + (long long)(zb + k) * c.sz);
    }
  }
  stage(zb);
  double dt1max = 0.0;
  const double dt1_diffus = (MAG ? fmax(c.nu * c.dxyz_2, c.eta * c.dxyz_2) : c.nu * c.dxyz_2) / c.cdtv;
  const double advec_cs2 = c.cs20 * c.dxyz_2;
  // own-cell values are loaded one plane ahead
  long long idx = (long long)(zb + NG) * c.sz;
  double ux = __ldg(pu + c.foff[R_UX] + idx), uy = __ldg(pu + c.foff[R_UY] + idx), uz = __ldg(pu + c.foff[R_UZ] + idx);
  double lnrho = MAG ? __ldg(pu + c.foff[R_LNRHO] + idx) : 0.0;
  cp_async_wait_all();
  __syncthreads();
#pragma unroll 1
  for (int n = zb; n < ze; n++) {
    const bool more = n + 1 < ze;
    double nux = 0.0, nuy = 0.0, nuz = 0.0, nln = 0.0, nwx = 0.0, nwy = 0.0;
    if (more) {
      stage(n + 1);
      idx += c.sz;
      nux = __ldg(pu + c.foff[R_UX] + idx); nuy = __ldg(pu + c.foff[R_UY] + idx); nuz = __ldg(pu + c.foff[R_UZ] + idx);
      if (MAG) {
        nln = __ldg(pu + c.foff[R_LNRHO] + idx);
        nwx = __ldg(pu + c.foff[R_AX] + idx + 3 * c.sz); nwy = __ldg(pu + c.foff[R_AY] + idx + 3 * c.sz);
      }
    }
    const double advec_uu = fabs(ux) * c.dx_1 + fabs(uy) * c.dy_1 + fabs(uz) * c.dz_1;     // hydro.f90:3810
    double advec_va2 = 0.0;
    if (MAG) {
      acc.cb = (n & 1) * PE + (ty + NG) * HX + tx + NG;
      const double axy = ddy<R_AX>(acc, c), axz = ddz<R_AX>(acc, c);
      const double ayx = ddx<R_AY>(acc, c), ayz = ddz<R_AY>(acc, c);
      const double azx = ddx<R_AZ>(acc, c), azy = ddy<R_AZ>(acc, c);
      const double bx = azy - ayz, by = axz - azx, bz = ayx - axy;
      const double rho1 = exp(-lnrho);
      const double t0 = bx * c.dx_1, t1 = by * c.dy_1, t2 = bz * c.dz_1;
      advec_va2 = (t0 * t0 + t1 * t1 + t2 * t2) * c.mu01 * rho1;                            // magnetic.f90:5438
    }
    // dt1 = sqrt((maxadvec/cdt)^2 + dt1_diffus^2) (equ.f90:1149) is a monotonic function of maxadvec (every
    // step of it is monotonic in floating point too), so the maximum is taken over maxadvec and the
    // function applied once per thread: same bits, one sqrt and one division less per cell
    const double maxadvec = advec_uu + sqrt(advec_va2 + advec_cs2);
    if (active) dt1max = fmax(dt1max, maxadvec);
    if (!more) break;
    cp_async_wait_all();
    __syncthreads();
    ux = nux; uy = nuy; uz = nuz; lnrho = nln;
    if (MAG) {
#pragma unroll
      for (int k = 0; k < 6; k++) { acc.wx[k] = acc.wx[k + 1]; acc.wy[k] = acc.wy[k + 1]; }
      acc.wx[6] = nwx; acc.wy[6] = nwy;
    }
  }
  double red[NRED];
#pragma unroll
  for (int q = 0; q < NRED; q++) red[q] = 0.0;
  {
    const double dt1_advec = dt1max / c.cdt;
    red[RD_DT1] = sqrt(dt1_advec * dt1_advec + dt1_diffus * dt1_diffus);
  }
  const int bl = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
  block_reduce_store<1>(red, tid, NTHR, redsm, partials, bl);
}
