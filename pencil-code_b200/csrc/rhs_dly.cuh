// rhs_dly.cuh -- z-marching kernel with delayed z-mixed derivatives (sm_100a).
//
// rhs_tma spends 96 of its 234 shared-memory loads per cell on the xz/yz
// bidiagonal mixed derivatives (derij_main, deriv.f90:1385-1423):
//     d2g/dxdz (z) = fzx * sum_k c_k [ D_k(z+k) - D_k(z-k) ],   D_k(z') = g(x+k,y,z') - g(x-k,y,z')
// D_k(z') is exactly the difference the thread forms anyway for dg/dx on the
// plane z'.  This kernel therefore never loads a neighbour outside the centre
// plane: it keeps the (pre-scaled) D_k of the last six planes in a per-thread
// history in TENSOR MEMORY (tcgen05.st / tcgen05.ld, 32x32b: lane = thread,
// columns = private scratch, 128 doubles per thread) and completes the terms
// that need the z-mixed derivatives three planes late.  Those terms enter the
// right-hand side linearly:
//     du/dt  += nu/3 * (z-mixed part of grad div u)              (viscosity.f90:1421-1448)
//     du/dt  += rho1 * (Jm x B),  Jm = z-mixed part of J         (magnetic.f90:3787,4360)
// so step p computes everything else of plane p ("phase A", u-part parked in
// TMEM together with rho1*B) and finishes du/dt of plane p-3 ("phase B").
// Shared memory holds 6 slots per field (centre .. centre+4 for the thread's
// own z-column, +1 in flight) for all fields instead of 8 for six of them.
//
// Used for lbidiagonal_derij=T, AUXM<2 (the full 0-D diagnostics need |J|^2 at the centre plane:
// those substeps use rhs_tma).
//
// Two forms of the bookkeeping in tensor memory (DlyCfg::ACCF):
//  * history form (hydro, hydro+entropy, isothermal MHD): the D_k of the last planes are kept (5+6+7
//    groups) and plane s-3 gathers its six terms at step s: 3 groups stored, 5 loaded per step;
//  * accumulator form (magnetic+entropy): every pending plane owns ONE running sum; step s scatters
//    +-D_k(s) into the sums of planes s-+k (7 groups live, 5 loaded and 5 stored per step).  That
//    leaves room for the larger pending state of Ohmic heating eta*mu0*|Jp+Jm|^2/(rho T)
//    (magnetic.f90:5122-5128), which is quadratic in J: the entropy equation of plane s-3 is finished
//    in phase B too, from the parked partial sum, eta*mu0*rho1/T and Jp.  8 fields need 34 KB per ring
//    slot: the six slots leave room for ONE staging buffer of the old 2N register, refilled for the next
//    step right after all threads have read it (measured dead ends at 512^3: five ring slots with three w
//    buffers 14.2 ms per substep -- the plane that enters the z window at the end of a step had only that
//    step to arrive; six slots with the w tiles requested at the start of their own step 17.0 ms).
#pragma once
#include "kernels.cuh"
#include <type_traits>

#ifndef PCB200_DLY_ACC
#define PCB200_DLY_ACC 0       // 1: accumulator form for every variant (A/B timing)
#endif
#ifndef PCB200_DLY_WPROD
#define PCB200_DLY_WPROD 0     // thread that requests the w tiles: 0 = the thread that requests the planes does both; 32 (warp 1, another scheduler) measured 1.9 % slower
#endif
#ifndef PCB200_DLY_ZS
#define PCB200_DLY_ZS 0        // 1: z+1..z+3 from the ring for every variant (A/B timing)
#endif

// ---- tensor memory as per-thread scratch ----------------------------------
PC_DEV void tm_st2(unsigned ta, double a)
{ asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};" ::"r"(ta), "r"(__double2loint(a)), "r"(__double2hiint(a))); }
PC_DEV void tm_st4(unsigned ta, double a, double b)
{ asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(ta), "r"(__double2loint(a)), "r"(__double2hiint(a)),
               "r"(__double2loint(b)), "r"(__double2hiint(b))); }
PC_DEV void tm_st8(unsigned ta, double a, double b, double c, double d)
{ asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(ta),
               "r"(__double2loint(a)), "r"(__double2hiint(a)), "r"(__double2loint(b)), "r"(__double2hiint(b)),
               "r"(__double2loint(c)), "r"(__double2hiint(c)), "r"(__double2loint(d)), "r"(__double2hiint(d))); }
PC_DEV void tm_ld2(unsigned ta, unsigned *r)
{ asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(ta)); }
PC_DEV void tm_ld4(unsigned ta, unsigned *r)
{ asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(ta)); }
PC_DEV void tm_ld8(unsigned ta, unsigned *r)
{ asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(ta)); }
PC_DEV void tm_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
PC_DEV void tm_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// Group g (0..NGRP-1) of NC doubles (NC = 3 or 6) of this thread.  Every access is naturally
// aligned: the first 2/3 of a group lives in region A (units of UA columns), the rest in
// region B (units of UA/2 columns) behind it.
#define DLY_NGRP 21            // history D1: 5 slots (groups 0-4), D2: 6 (5-10), D3: 7 (11-17); pending: 3 (18-20)
template <int NC> PC_DEV void tm_st_group(unsigned t0, int g, const double *v)
{
  if (NC == 6) { tm_st8(t0 + g * 8, v[0], v[1], v[2], v[3]); tm_st4(t0 + DLY_NGRP * 8 + g * 4, v[4], v[5]); }
  else         { tm_st4(t0 + g * 4, v[0], v[1]); tm_st2(t0 + DLY_NGRP * 4 + g * 2, v[2]); }
}
template <int NC> PC_DEV void tm_ld_group(unsigned t0, int g, unsigned *r)   // r[2*NC]; valid after tm_wait_ld()
{
  if (NC == 6) { tm_ld8(t0 + g * 8, r); tm_ld4(t0 + DLY_NGRP * 8 + g * 4, r + 8); }
  else         { tm_ld4(t0 + g * 4, r); tm_ld2(t0 + DLY_NGRP * 4 + g * 2, r + 4); }
}
PC_DEV double tm_dbl(const unsigned *r, int i) { return __hiloint2double((int)r[2 * i + 1], (int)r[2 * i]); }

// Position of a role inside a ring slot = its storage slot in f for the standard registration
// order of the reference (register.f90:124-135: ux,uy,uz,lnrho,[ss],ax,ay,az), so that ONE 4-D TMA
// box {HX,HY,1,mvar} brings a whole plane.  The host checks the order and mvar before it selects
// this kernel (other layouts run rhs_tma).
template <bool MAG, bool ENT> PC_DEV constexpr int dly_fidx(int R)
{ return R < 3 ? R : (R == R_LNRHO ? 3 : (R == R_SS ? 4 : (R - R_AX) + (ENT ? 5 : 4))); }
__device__ __forceinline__ void tma_load_4d(void *dst, const CUtensorMap *tm, int x, int y, int z, int v, unsigned long long *bar)
{
  asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
               ::"r"(smem_u32(dst)), "l"(tm), "r"(x), "r"(y), "r"(z), "r"(v), "r"(smem_u32(bar)) : "memory");
}

template <bool MAG, bool ENT> struct DlyCfg {
  static constexpr int NF = 4 + (MAG ? 3 : 0) + (ENT ? 1 : 0);
  static constexpr int NC = MAG ? 6 : 3;                      // z-mixed combinations kept per plane and distance k
  static constexpr bool ACCF = (MAG && ENT) || PCB200_DLY_ACC;   // accumulator form (see the head of this file)
  static constexpr int PG = (MAG && ENT) ? 2 : 1;             // pending groups per plane
  // ring slots per field: centre .. centre+4 landed, centre+5 in flight
  static constexpr int NSD = 6;
  static constexpr int NP = NSD - 1;                          // planes of the prologue
  static constexpr int SLD = (NF * PE + 15) / 16 * 16;        // doubles per ring slot [field][HY][HX], padded to 128 B
  static constexpr int NDBL = NSD * SLD;
  // old-2N-register staging [3 steps][field][TY][WTX]: a TMA box must start at a 16-byte aligned
  // global address (measured: x0+3 faults, tools/tma4d_test.cu), so the tile starts at x0+2
  static constexpr int WTX = TX + 2, WTE = WTX * TY;
  // three buffers filled two steps ahead; 8 fields (34 KB per ring slot): one buffer, refilled for step s+1 as
  // soon as every thread has read the tiles of step s (its own "wempty" barrier): almost a whole step ahead
  static constexpr int NWBUF = (MAG && ENT) ? 1 : 3, WD = (MAG && ENT) ? 1 : 2;
  static constexpr int NWB = NWBUF * NF * WTE;
  static constexpr size_t SMEM = sizeof(double) * ((size_t)NDBL + NWB + 8 * NRED + 12) + 128;
  // tensor-memory groups per thread (NC doubles each), see DLY_NGRP
  static_assert(DLY_NGRP * NC * 2 <= 256, "tensor-memory budget per thread");
  static_assert(!ACCF || 7 + 3 * PG <= DLY_NGRP, "tensor-memory groups of the accumulator form");
};

// centre-plane accessor: x/y neighbours from the ring slot of the centre plane, z from the
// register window; the z-mixed accessors return 0 (those terms are completed in phase B)
// ZS: the planes z+1..z+3 of the thread's own column are read from the ring slots (they are there anyway)
// instead of being carried in the register window: 8 fields x 3 planes = 48 registers less for the
// magnetic+entropy variant, whose spills went to L2 (223 KB of shared memory leave almost no L1).
template <bool MAG, bool ENT, bool UPW_>
struct DlyAcc {
  static constexpr bool ZS = (MAG && ENT) || PCB200_DLY_ZS;
  // UPW: upwinding compiled in (lupw_uu/lupw_lnrho/lupw_ss are run parameters; behind a run-time branch the
  // compiler evaluated the 104 FP64 instructions of the der6 stencils in every cell, used or not)
  static constexpr bool PS = true, UPW = UPW_, NOZMIX = true;
  PC_DEV constexpr bool bidiag() const { return true; }
  const double *smem0;
  unsigned cb;            // byte offset of this thread's cell in the ring slot of the centre plane
  unsigned zo[3];         // ZS: the same for the slots of planes z+1..z+3
  double w[NROLE][ZS ? 4 : 7];     // own column, planes z-3..z+3 (ZS: z-3..z)
  template <int R, int OFF> PC_DEV double cenv() const { return lds_at(smem0, cb, dly_fidx<MAG, ENT>(R) * PE + OFF); }
  template <int R> PC_DEV double c() const { return w[R][3]; }
  template <int R, int K> PC_DEV double x() const { return cenv<R, K>(); }
  template <int R, int K> PC_DEV double y() const { return cenv<R, K * HX>(); }
  template <int R, int K> PC_DEV double z() const
  { if (ZS && K > 0) return lds_at(smem0, zo[K > 0 ? K - 1 : 0], dly_fidx<MAG, ENT>(R) * PE); else return w[R][K > 0 && ZS ? 3 : 3 + K]; }
  template <int R, int KX, int KY> PC_DEV double xy() const { return cenv<R, KX + KY * HX>(); }
  template <int R, int KX, int KZ> PC_DEV double xz() const { return 0.0; }
  template <int R, int KY, int KZ> PC_DEV double yz() const { return 0.0; }
  template <int R, int K> PC_DEV double dx() const { return cenv<R, K>() - cenv<R, -K>(); }
  template <int R, int K> PC_DEV double dy() const { return cenv<R, K * HX>() - cenv<R, -K * HX>(); }
};

// the NC pre-scaled combinations of distance K on the centre plane
template <bool MAG, bool ENT, bool UPW, int K>
PC_DEV void dly_combos(const DlyAcc<MAG, ENT, UPW> &a, const DevPar &c, double *__restrict__ d)
{
  const double *su = c.dlyu[K - 1], *sa = c.dlya[K - 1];      // nu/3*c_K*{fzx,fyz}, mu01*c_K*{fzx,fyz}
  d[0] = su[0] * a.template dx<R_UZ, K>();                                           // nu/3 dxz(uz)
  d[1] = su[1] * a.template dy<R_UZ, K>();                                           // nu/3 dyz(uz)
  d[2] = fma(su[0], a.template dx<R_UX, K>(), su[1] * a.template dy<R_UY, K>());     // nu/3 (dxz(ux) + dyz(uy))
  if (MAG) {
    d[3] = sa[0] * a.template dx<R_AZ, K>();                                         // Jm_x = mu01 dxz(Az)
    d[4] = sa[1] * a.template dy<R_AZ, K>();                                         // Jm_y = mu01 dyz(Az)
    d[5] = fma(sa[0], a.template dx<R_AX, K>(), sa[1] * a.template dy<R_AY, K>());   // Jm_z
  }
}

template <bool MAG, bool ENT, int AUXM, bool RW, bool UPW>
__global__ void __launch_bounds__(NTHR, 1) rhs_dly(SubstepArgs a, const __grid_constant__ CUtensorMap tmap,
                                                   const __grid_constant__ CUtensorMap tmap_wu, const __grid_constant__ CUtensorMap tmap_wr)
{
  // tmap: f_in as (mx,my,mz,mvar), box {HX,HY,1,mvar}.  tmap_wu / tmap_wr: the 2N register w, boxes
  // {TX+2,TY,1,3} (u) and {TX+2,TY,1,mvar-3} (the other fields) -- own cells only.  Magnetic+entropy:
  // tmap_wr has a one-variable box (lnrho at plane n, ss at plane n-3) and A comes through tmap_wu.
  static_assert(AUXM < 2, "full diagnostics run through rhs_tma");
  typedef DlyCfg<MAG, ENT> Cfg;
  constexpr int NC = Cfg::NC, NF = Cfg::NF, NSD = Cfg::NSD, NP = Cfg::NP, PG = Cfg::PG, NWBUF = Cfg::NWBUF;
  constexpr bool ACCF = Cfg::ACCF, ME = MAG && ENT;
  const DevPar &c = c_par;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double *smem = reinterpret_cast<double *>(smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u));
  double *ring = smem;
  double *wbuf = smem + (size_t)Cfg::NDBL;
  double *redsm = wbuf + Cfg::NWB;
  unsigned long long *bar = reinterpret_cast<unsigned long long *>(redsm + 8 * NRED);
  unsigned *tbase_sm = reinterpret_cast<unsigned *>(bar + 8);
  constexpr unsigned BOX_BYTES = NF * PE * 8;       // one plane of every field

  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * TX + tx, warp = tid >> 5;
  const int x0 = (a.tx0 + blockIdx.x * a.txs) * TX, y0 = (a.ty0 + blockIdx.y * a.tys) * TY;
  const int zb = a.z_lo + blockIdx.z * a.zchunk;
  const int ze = min(zb + a.zchunk, a.z_hi);
  const int nzc = ze - zb;
  const int nstep = nzc + 6;                  // centre planes zb-3 .. ze+2 (ring planes 0 .. nzc+5)
  const bool active = (x0 + tx < c.nx) && (y0 + ty < c.ny);
  const int lx = min(x0 + tx, c.nx - 1), ly = min(y0 + ty, c.ny - 1);
  const long long col = (long long)(ly + NG) * c.sy + (lx + NG);
  const int tb = (ty + NG) * HX + tx + NG;
  const bool ww = a.write_w != 0 && active, wf = a.write_f != 0 && active;

  // full[b]: the NP planes of the prologue complete phase 0 of full[0]; ring plane q >= NP is transaction
  // j = q-NP+1 and completes phase j/2 of full[j&1]; two barriers because plane q+1 is armed while plane q
  // may still be in flight.
  // empty[s&1]: phase s/2 completes when all NTHR threads are done with step s (warps drift up to
  // two steps apart, so consecutive steps must not share a barrier).
  // wfull[s%NWBUF]: phase s/NWBUF = the w tiles of step s have landed (issued WD steps ahead).
  // wempty (single w buffer only): phase s completes when all threads have read the w tiles of step s.
  unsigned long long *full = bar, *empty = bar + 2, *wfull = bar + 4, *wempty = bar + 7;
  if (tid == 0) {
    mbar_init(full, 1);
    mbar_init(full + 1, 1);
    mbar_init(empty, NTHR);
    mbar_init(empty + 1, NTHR);
    mbar_init(wfull, 1);
    mbar_init(wfull + 1, 1);
    mbar_init(wfull + 2, 1);
    mbar_init(wempty, NTHR);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tbase_sm)), "n"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  // this warp's lane quarter; warps 4..7 use the upper 256 columns
  const unsigned t0 = *tbase_sm + ((unsigned)((warp & 3) * 32) << 16) + (unsigned)((warp >> 2) * 256);

  auto issue_plane = [&](int q, int sl) __attribute__((always_inline)) {     // ring plane q (padded z = zb+q) of every field into slot sl = q % NSD
    tma_load_4d(ring + (size_t)sl * Cfg::SLD, &tmap, x0, y0, zb + q, 0, full + (q >= NP ? ((q - NP + 1) & 1) : 0));
  };
  // The old 2N register w of step s is fetched by TMA two steps ahead into wbuf[s%3][field][TY][WTX]
  // (own cells only; no per-thread loads, no registers held across the stencils).  u belongs to plane n-3.
  auto issue_w = [&](int s, int s3) __attribute__((always_inline)) {   // s3 = s % 3
    if (!RW) return;
    double *dst = wbuf + (size_t)s3 * NF * Cfg::WTE;
    mbar_expect_tx(wfull + s3, (unsigned)(NF * Cfg::WTE * 8));
    const int zl = zb + s - (s >= 6 ? 3 : 0);                            // fields finished in phase B: plane n-3
    tma_load_4d(dst, &tmap_wu, x0 + NG - 1, y0 + NG, zl, 0, wfull + s3);
    if (ME) {
      tma_load_4d(dst + 3 * Cfg::WTE, &tmap_wr, x0 + NG - 1, y0 + NG, zb + s, 3, wfull + s3);     // lnrho
      tma_load_4d(dst + 4 * Cfg::WTE, &tmap_wr, x0 + NG - 1, y0 + NG, zl, 4, wfull + s3);         // ss
      tma_load_4d(dst + 5 * Cfg::WTE, &tmap_wu, x0 + NG - 1, y0 + NG, zb + s, 5, wfull + s3);     // A
    } else {
      tma_load_4d(dst + 3 * Cfg::WTE, &tmap_wr, x0 + NG - 1, y0 + NG, zb + s, 3, wfull + s3);
    }
  };
  if (tid == 0) {
    mbar_expect_tx(full, (unsigned)(NP * BOX_BYTES));
#pragma unroll 1
    for (int q = 0; q < NP; q++) issue_plane(q, q);
    issue_w(0, 0);
    if (NWBUF == 3) issue_w(1, 1);
  }
  // zero the history and the pending slots (defined values in the first steps)
  {
    double z6[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
#pragma unroll 1
    for (int g = 0; g < DLY_NGRP; g++) tm_st_group<NC>(t0, g, z6);
  }
  DlyAcc<MAG, ENT, UPW> acc;
  acc.smem0 = smem;
  constexpr bool ZS = DlyAcc<MAG, ENT, UPW>::ZS;
  constexpr int NW = ZS ? 4 : 7;
#pragma unroll
  for (int r = 0; r < NROLE; r++)
#pragma unroll
    for (int k = 0; k < NW; k++) acc.w[r][k] = 0.0;
  double red[NRED];
#pragma unroll
  for (int q = 0; q < NRED; q++) red[q] = 0.0;
  const double beta_dt = a.write_f ? a.beta * a.dt_dev[0] : 0.0;
  // store addresses as byte pointers: one pointer into w advanced by a plane per step, the role's offset
  // added from the constant bank, f_out at a fixed distance from w (64-bit element indices scaled by 8 cost
  // five integer instructions per store)
  char *pw = reinterpret_cast<char *>(a.w) + ((long long)zb * c.sz + col) * 8;        // centre plane of step 0 is zb-NG (interior-relative)
  const long long f_minus_w = reinterpret_cast<char *>(a.f_out) - reinterpret_cast<char *>(a.w);
  const long long sz8 = c.sz * 8;
  mbar_wait(full, 0);
  // window of step 0: centre = ring plane 0, planes 0..3 are in the ring
#pragma unroll
  for (int r = 0; r < NROLE; r++) {
    if (!role_present<MAG, ENT>(r)) continue;
#pragma unroll
    for (int k = 0; k < NW - 3; k++) acc.w[r][3 + k] = ring[(size_t)k * Cfg::SLD + dly_fidx<MAG, ENT>(r) * PE + tb];
  }

  int c5 = 0, c6 = 0, c7 = 0, c3 = 0;         // s % 5, s % 6, s % 7, s % 3
  int cr = 0;                                 // s % NSD: ring slot of the centre plane
  unsigned p3 = 0;                            // (s / 3) & 1
  // one step; FULL: centre plane and plane n-3 both inside the chunk (no masks)
  auto step = [&](const int s, auto full_tag) __attribute__((always_inline)) {
    constexpr bool FULL = decltype(full_tag)::value;
    const int n = zb - NG + s;                // centre plane, interior-relative
    const bool more = s + NP < nstep;         // ring plane s+NP exists: issue it into the slot of plane s-1
    const bool morew = RW && NWBUF == 3 && s + 2 < nstep;   // w tiles of step s+2 into the buffer of step s-1
    // both overwrite what step s-1 used: wait for "empty" of step s-1 (probe now, block after the stencils).
    // One producer by default (PCB200_DLY_WPROD=0).  Splitting the ~80 producer instructions over two warps on two
    // schedulers (thread 0 the plane, thread 32 the w tiles) measured 7.53 against 7.39 ms per 512^3 substep.
    const bool prodP = more && tid == 0, prodW = morew && tid == PCB200_DLY_WPROD;
    auto issue_mine = [&]() __attribute__((always_inline)) {
      if (prodP) { mbar_expect_tx(full + ((s + 1) & 1), BOX_BYTES); issue_plane(s + NP, cr == 0 ? NSD - 1 : cr - 1); }
      if (prodW) issue_w(s + 2, c3 == 0 ? 2 : c3 - 1);
    };
    bool issued = true;
    if (prodP || prodW) {
      issued = s == 0 || mbar_test_wait(empty + ((s - 1) & 1), (unsigned)(((s - 1) >> 1) & 1));
      if (issued) issue_mine();
    }
    const bool doA = FULL || (s >= 3 && s < nzc + 3);   // phase A produces output (centre plane inside the chunk)
    const bool doB = FULL || s >= 6;                    // phase B finishes plane n-3
    // ---- phase A: everything of plane n except the z-mixed terms ----
    acc.cb = (unsigned)((cr * Cfg::SLD + tb) * 8);
    if (ZS) {
#pragma unroll
      for (int k = 0; k < 3; k++) { const int sl = cr + k + 1 >= NSD ? cr + k + 1 - NSD : cr + k + 1; acc.zo[k] = (unsigned)((sl * Cfg::SLD + tb) * 8); }
    }
    // the pre-scaled x/y differences of this plane (history form: to their slots at once, none is read in this step)
    double d1[6], d2[6], d3[6];
    dly_combos<MAG, ENT, UPW, 1>(acc, c, d1);
    if (!ACCF) tm_st_group<NC>(t0, c5, d1);
    dly_combos<MAG, ENT, UPW, 2>(acc, c, d2);
    if (!ACCF) tm_st_group<NC>(t0, 5 + c6, d2);
    dly_combos<MAG, ENT, UPW, 3>(acc, c, d3);
    if (!ACCF) tm_st_group<NC>(t0, 11 + c7, d3);
    double rhs[NROLE];
    CellAux aux;
    MagOut m;
    m.fx = m.fy = m.fz = m.j2 = m.advec_va2 = m.b2 = m.ab = m.jb = 0.0;
    m.bx = m.by = m.bz = m.jx = m.jy = m.jz = 0.0;
    const double rho1 = pc_exp(-acc.template c<R_LNRHO>());
    if (MAG) mag_part<(AUXM > 0)>(acc, c, acc.template c<R_UX>(), acc.template c<R_UY>(), acc.template c<R_UZ>(), rho1, rhs + R_AX, m);
    HydroPend hp;
    hydro_pre<MAG, ENT, (AUXM > 0)>(acc, c, FULL ? n : min(max(n, 0), c.nz - 1), rhs, aux, hp);
    hydro_finish<MAG, ENT, (AUXM > 0), false>(c, hp, m, rhs, aux);     // Ohmic heating needs the whole J: phase B
    if (!issued) {
      mbar_wait(empty + ((s - 1) & 1), (unsigned)(((s - 1) >> 1) & 1));
      issue_mine();
    }
    // ---- phase B: finish plane n-3 (du/dt; with magnetic+entropy also ds/dt) ----
    // ga..ge: the five groups loaded per step.  History form: D3(s-6), D2(s-1), D2(s-5), D1(s-2), D1(s-4).
    // Accumulator form: the running sums of planes s-3 (complete after + D3(s)), s-2, s-1, s+1, s+2.
    unsigned ga[2 * NC], gb[2 * NC], gc[2 * NC], gd[2 * NC], ge[2 * NC], pn[2 * NC], pn2[ME ? 2 * NC : 2];
    int sb = 0, sc = 0, sd = 0, se = 0, sf = 0;                     // accumulator form: groups of planes s-2, s-1, s+1, s+2, s+3
    tm_wait_st();                                                   // the tensor-memory stores of step s-1 are visible
    if (!ACCF) {
      tm_ld_group<NC>(t0, 11 + (c7 == 6 ? 0 : c7 + 1), ga);         // D3(s-6)
      tm_ld_group<NC>(t0, 5 + (c6 == 0 ? 5 : c6 - 1), gb);          // D2(s-1)
      tm_ld_group<NC>(t0, 5 + (c6 == 5 ? 0 : c6 + 1), gc);          // D2(s-5)
      tm_ld_group<NC>(t0, c5 >= 2 ? c5 - 2 : c5 + 3, gd);           // D1(s-2)
      tm_ld_group<NC>(t0, c5 == 4 ? 0 : c5 + 1, ge);                // D1(s-4)
    } else {
      sb = c7 >= 2 ? c7 - 2 : c7 + 5; sc = c7 >= 1 ? c7 - 1 : 6; sd = c7 == 6 ? 0 : c7 + 1; se = c7 >= 5 ? c7 - 5 : c7 + 2;
      sf = c7 >= 4 ? c7 - 4 : c7 + 3;                               // plane s+3 takes over the group of plane s-4
      tm_ld_group<NC>(t0, c7 >= 3 ? c7 - 3 : c7 + 4, ga);
      tm_ld_group<NC>(t0, sb, gb);
      tm_ld_group<NC>(t0, sc, gc);
      tm_ld_group<NC>(t0, sd, gd);
      tm_ld_group<NC>(t0, se, ge);
    }
    const int pg = ACCF ? 7 + c3 * PG : 18 + c3;                    // pending group(s) of plane s-3 (and of plane s)
    tm_ld_group<NC>(t0, pg, pn);
    if (ME) tm_ld_group<NC>(t0, pg + 1, pn2);
    double wold[NROLE];
    if (RW) {
      const int cw = NWBUF == 1 ? 0 : c3;
      mbar_wait(wfull + cw, NWBUF == 1 ? (unsigned)(s & 1) : p3);   // this step's w tiles have landed
      const double *src = wbuf + (size_t)cw * NF * Cfg::WTE + ty * Cfg::WTX + tx + 1;
#pragma unroll
      for (int r = 0; r < NROLE; r++) {
        if (!role_present<MAG, ENT>(r)) continue;
        wold[r] = src[dly_fidx<MAG, ENT>(r) * Cfg::WTE];
      }
      if (NWBUF == 1) mbar_arrive(wempty);
    }
    // fields without delayed terms leave while the tensor-memory loads are in flight
#pragma unroll
    for (int r = 3; r < NROLE; r++) {
      if (!role_present<MAG, ENT>(r) || (ME && r == R_SS)) continue;
      char *const q = pw + c.foff8[r];
      const double wv = RW ? rhs[r] + a.alpha * wold[r] : rhs[r];
      if (ww && doA) __stcs(reinterpret_cast<double *>(q), wv);
      if (wf && doA) __stcs(reinterpret_cast<double *>(q + f_minus_w), acc.w[r][3] + beta_dt * wv);
    }
    if (AUXM > 0 && active && doA) red[RD_DT1] = fmax(red[RD_DT1], aux.dt1);
    tm_wait_ld();
    {
      double pv[6];
      pv[0] = rhs[R_UX]; pv[1] = rhs[R_UY]; pv[2] = rhs[R_UZ];
      if (MAG) { pv[3] = rho1 * m.bx; pv[4] = rho1 * m.by; pv[5] = rho1 * m.bz; }
      tm_st_group<NC>(t0, pg, pv);
      if (ME) {
        double pw2[6];
        pw2[0] = rhs[R_SS]; pw2[1] = hp.ohm; pw2[2] = m.jx; pw2[3] = m.jy; pw2[4] = m.jz; pw2[5] = 0.0;
        tm_st_group<NC>(t0, pg + 1, pw2);
      }
    }
    double mz[6];
    if (!ACCF) {
#pragma unroll
      for (int q = 0; q < NC; q++)
        mz[q] = (d3[q] - tm_dbl(ga, q)) + (tm_dbl(gb, q) - tm_dbl(gc, q)) + (tm_dbl(gd, q) - tm_dbl(ge, q));
    } else {
      double ub[6], uc[6], ud[6], ue[6], uf[6];
#pragma unroll
      for (int q = 0; q < NC; q++) {
        mz[q] = tm_dbl(ga, q) + d3[q];
        ub[q] = tm_dbl(gb, q) + d2[q]; uc[q] = tm_dbl(gc, q) + d1[q];
        ud[q] = tm_dbl(gd, q) - d1[q]; ue[q] = tm_dbl(ge, q) - d2[q];
        uf[q] = -d3[q];
      }
      tm_st_group<NC>(t0, sb, ub); tm_st_group<NC>(t0, sc, uc); tm_st_group<NC>(t0, sd, ud); tm_st_group<NC>(t0, se, ue);
      tm_st_group<NC>(t0, sf, uf);
    }
    double ru[3];
    ru[0] = tm_dbl(pn, 0) + mz[0]; ru[1] = tm_dbl(pn, 1) + mz[1]; ru[2] = tm_dbl(pn, 2) + mz[2];
    if (MAG) {
      const double qx = tm_dbl(pn, 3), qy = tm_dbl(pn, 4), qz = tm_dbl(pn, 5);      // rho1*B of plane n-3
      ru[0] += mz[4] * qz - mz[5] * qy;
      ru[1] += mz[5] * qx - mz[3] * qz;
      ru[2] += mz[3] * qy - mz[4] * qx;
    }
    char *const pwB = pw - 3 * sz8;
#pragma unroll
    for (int r = 0; r < 3; r++) {
      char *const q = pwB + c.foff8[r];
      const double wv = RW ? ru[r] + a.alpha * wold[r] : ru[r];
      if (ww && doB) __stcs(reinterpret_cast<double *>(q), wv);
      if (wf && doB) __stcs(reinterpret_cast<double *>(q + f_minus_w), acc.w[r][0] + beta_dt * wv);
    }
    if (ME) {
      // ds/dt of plane n-3: parked partial sum + eta*mu0*rho1/T * |Jp + Jm|^2 (magnetic.f90:5122-5128)
      const double jx = tm_dbl(pn2, 2) + mz[3], jy = tm_dbl(pn2, 3) + mz[4], jz = tm_dbl(pn2, 4) + mz[5];
      const double rs = fma(tm_dbl(pn2, 1), jx * jx + jy * jy + jz * jz, tm_dbl(pn2, 0));
      char *const q = pwB + c.foff8[R_SS];
      const double wv = RW ? rs + a.alpha * wold[R_SS] : rs;
      if (ww && doB) __stcs(reinterpret_cast<double *>(q), wv);
      if (wf && doB) __stcs(reinterpret_cast<double *>(q + f_minus_w), acc.w[R_SS][0] + beta_dt * wv);
    }

    // ---- next plane: shift the z window, release the centre slot ----
    if (s + 1 == nstep) return;
    if (RW && NWBUF == 1 && tid == 0) {                               // refill the single w buffer for step s+1
      mbar_wait(wempty, (unsigned)(s & 1));
      issue_w(s + 1, 0);
    }
    pw += sz8;
    const bool have4 = s + 4 < nstep;
    if (have4) { const int j = s + 5 - NP; mbar_wait(full + (j & 1), (unsigned)((j >> 1) & 1)); }   // ring plane s+4 = transaction j
    const int nl = cr + 4 >= NSD ? cr + 4 - NSD : cr + 4;             // slot of ring plane s+4
#pragma unroll
    for (int r = 0; r < NROLE; r++) {
      if (!role_present<MAG, ENT>(r)) continue;
#pragma unroll
      for (int k = 0; k < NW - 1; k++) asm volatile("mov.b64 %0, %1;" : "=d"(acc.w[r][k]) : "d"(acc.w[r][k + 1]));   // one move each, in this order
      if (ZS) acc.w[r][3] = lds_at(smem, acc.zo[0], dly_fidx<MAG, ENT>(r) * PE);       // plane s+1 becomes the centre
      else acc.w[r][6] = have4 ? ring[(size_t)nl * Cfg::SLD + dly_fidx<MAG, ENT>(r) * PE + tb] : 0.0;
    }
    mbar_arrive(empty + (s & 1));
    c5 = c5 == 4 ? 0 : c5 + 1;
    c6 = c6 == 5 ? 0 : c6 + 1;
    c7 = c7 == 6 ? 0 : c7 + 1;
    cr = cr == NSD - 1 ? 0 : cr + 1;
    if (c3 == 2) { c3 = 0; p3 ^= 1u; } else c3++;
  };
  {
    int s = 0;
    const int s_main_end = nzc + 3;
#pragma unroll 1
    for (; s < 6; s++) step(s, std::false_type());
#pragma unroll 1
    for (; s < s_main_end; s++) step(s, std::true_type());
#pragma unroll 1
    for (; s < nstep; s++) step(s, std::false_type());
  }
  if (AUXM > 0) {
    const int bl = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
    block_reduce_store<AUXM>(red, tid, NTHR, redsm, a.partials, bl);
  }
  tm_wait_st();
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(*tbase_sm), "n"(512) : "memory");
}
