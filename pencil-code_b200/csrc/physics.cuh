// physics.cuh -- the fused right-hand side of one grid cell.
//
// Replaces, for one (l,m,n), everything the reference evaluates per pencil in
// rhs_cpu (equ.f90:816-1239): calc_pencils_{hydro,density,eos,energy,viscosity,
// magnetic} and duu_dt / dlnrho_dt / denergy_dt / daa_dt / duu_dt_grav plus the
// CFL combination (equ.f90:1086-1151).  The stencils are the 6th-order ones of
// deriv.f90 (der :89, der2 :474, derij bidiagonal :1373, der6 upwind :938).
// The data access is abstracted by an accessor `A` so that the same physics
// serves the z-marching shared-memory kernel and the plain global-memory
// kernel.  Evaluation order is free here (FMA contraction on): parity with the
// Fortran order is a 1e-10 tolerance, checked against oracle/pcref.c.
#pragma once
#include "devpar.h"

#define PC_DEV __device__ __forceinline__

// ---- exp for the density pencils ---------------------------------------
// rho1 = exp(-lnrho) is evaluated in every cell.  The library routine materialises its eleven polynomial
// coefficients as 22 immediates per call (measured in the z loop of rhs_dly: 60 instructions, 18 of them FP64);
// here they are constant-bank operands of the FMAs.  Cody-Waite reduction x = n ln2 + r, |r| <= ln2/2, Taylor
// polynomial of degree 13 (truncation 4e-18 relative), scaling by 2^n through the exponent field.  Arguments
// beyond +-700 (never a physical ln rho; exp leaves the normal range at 709) saturate to inf / 0, NaN stays NaN.
// No fallback call: a call inside the z loop cost 48 more shared-memory loads per cell (registers saved around it).
__constant__ double c_expc[12] = {0.5, 0.16666666666666666, 0.041666666666666664, 0.008333333333333333, 0.001388888888888889, 0.0001984126984126984, 2.48015873015873e-05, 2.7557319223985893e-06, 2.755731922398589e-07, 2.505210838544172e-08, 2.08767569878681e-09, 1.6059043836821613e-10};   // 1/2! .. 1/13!
PC_DEV double pc_exp(double x)
{
  const double t = fma(x, 1.4426950408889634, 6755399441055744.0);      // x*log2(e) + 1.5*2^52: n in the low word
  const double tn = t - 6755399441055744.0;
  double r = fma(tn, -6.93147180369123816490e-01, x);                  // ln2 high part (trailing zeros: n*hi is exact)
  r = fma(tn, -1.90821492927058770002e-10, r);                         // ln2 low part
  double q = c_expc[11];
#pragma unroll
  for (int k = 10; k >= 0; k--) q = fma(q, r, c_expc[k]);
  q = fma(q, r, 1.0);
  q = fma(q, r, 1.0);
  const int n = __double2loint(t);
  const double e = q * __hiloint2double((n + 1023) << 20, 0);
  return fabs(x) < 700.0 ? e : (x != x ? x : (x > 0.0 ? __longlong_as_double(0x7ff0000000000000LL) : 0.0));
}

// ---- stencils ---------------------------------------------------------
// Two formulations of the same stencils.  A::PS (pre-scaled): coefficients f*c_k come from the
// constant bank (DevPar::k1x ...), one multiply-add per tap and merged Laplacian pairs: ~10 % fewer
// FP64 instructions.  !A::PS: factor * (integer-coefficient sum), kept for the MHD+entropy variant
// of rhs_tma, whose register allocation spills more with the pre-scaled form (13.4 vs 12.4 ms).
#define PC_D1(ACC, K, F) (A::PS ? fma(K[2], a.template ACC<R, 3>() - a.template ACC<R, -3>(), fma(K[1], a.template ACC<R, 2>() - a.template ACC<R, -2>(), \
                                  K[0] * (a.template ACC<R, 1>() - a.template ACC<R, -1>()))) \
                                : F * (45.0 * (a.template ACC<R, 1>() - a.template ACC<R, -1>()) - 9.0 * (a.template ACC<R, 2>() - a.template ACC<R, -2>()) \
                                       + (a.template ACC<R, 3>() - a.template ACC<R, -3>())))
template <int R, class A> PC_DEV double ddx(const A &a, const DevPar &c) { return PC_D1(x, c.k1x, c.fx1); }
template <int R, class A> PC_DEV double ddy(const A &a, const DevPar &c) { return PC_D1(y, c.k1y, c.fy1); }
template <int R, class A> PC_DEV double ddz(const A &a, const DevPar &c) { return PC_D1(z, c.k1z, c.fz1); }
#undef PC_D1

#define PC_C0 (-490.0 / 180.0)
#define PC_C1 (270.0 / 180.0)
#define PC_C2 (-27.0 / 180.0)
#define PC_C3 (2.0 / 180.0)
// taps of one axis added to acc (pre-scaled form)
#define PC_D2T(ACC, K, acc) fma(K[3], a.template ACC<R, 3>() + a.template ACC<R, -3>(), fma(K[2], a.template ACC<R, 2>() + a.template ACC<R, -2>(), \
                            fma(K[1], a.template ACC<R, 1>() + a.template ACC<R, -1>(), acc)))
#define PC_D2F(ACC, F) (F * (PC_C0 * a.template c<R>() + PC_C1 * (a.template ACC<R, 1>() + a.template ACC<R, -1>()) \
                             + PC_C2 * (a.template ACC<R, 2>() + a.template ACC<R, -2>()) + PC_C3 * (a.template ACC<R, 3>() + a.template ACC<R, -3>())))
template <int R, class A> PC_DEV double d2x(const A &a, const DevPar &c) { return A::PS ? PC_D2T(x, c.k2x, c.k2x[0] * a.template c<R>()) : PC_D2F(x, c.fx2); }
template <int R, class A> PC_DEV double d2y(const A &a, const DevPar &c) { return A::PS ? PC_D2T(y, c.k2y, c.k2y[0] * a.template c<R>()) : PC_D2F(y, c.fy2); }
template <int R, class A> PC_DEV double d2z(const A &a, const DevPar &c) { return A::PS ? PC_D2T(z, c.k2z, c.k2z[0] * a.template c<R>()) : PC_D2F(z, c.fz2); }
// pair sums d2y+d2z, d2z+d2x, d2x+d2y and the Laplacian, with a common centre coefficient
template <int R, class A> PC_DEV double p2yz(const A &a, const DevPar &c)
{ if (!A::PS) return d2y<R>(a, c) + d2z<R>(a, c); const double t = PC_D2T(y, c.k2y, c.k2yz0 * a.template c<R>()); return PC_D2T(z, c.k2z, t); }
template <int R, class A> PC_DEV double p2zx(const A &a, const DevPar &c)
{ if (!A::PS) return d2x<R>(a, c) + d2z<R>(a, c); const double t = PC_D2T(z, c.k2z, c.k2zx0 * a.template c<R>()); return PC_D2T(x, c.k2x, t); }
template <int R, class A> PC_DEV double p2xy(const A &a, const DevPar &c)
{ if (!A::PS) return d2x<R>(a, c) + d2y<R>(a, c); const double t = PC_D2T(x, c.k2x, c.k2xy0 * a.template c<R>()); return PC_D2T(y, c.k2y, t); }
template <int R, class A> PC_DEV double del2(const A &a, const DevPar &c)
{
  if (!A::PS) return d2x<R>(a, c) + d2y<R>(a, c) + d2z<R>(a, c);
  const double t = PC_D2T(x, c.k2x, c.k2xyz0 * a.template c<R>()); const double u = PC_D2T(y, c.k2y, t); return PC_D2T(z, c.k2z, u);
}
#undef PC_D2T
#undef PC_D2F

// upwind correction D6 * dx^5/60 (der6 with upwind=T)
template <int R, class A> PC_DEV double d6x(const A &a, double f1)
{ return f1 * (-20.0 * a.template c<R>() + 15.0 * (a.template x<R, 1>() + a.template x<R, -1>()) - 6.0 * (a.template x<R, 2>() + a.template x<R, -2>()) + (a.template x<R, 3>() + a.template x<R, -3>())); }
template <int R, class A> PC_DEV double d6y(const A &a, double f1)
{ return f1 * (-20.0 * a.template c<R>() + 15.0 * (a.template y<R, 1>() + a.template y<R, -1>()) - 6.0 * (a.template y<R, 2>() + a.template y<R, -2>()) + (a.template y<R, 3>() + a.template y<R, -3>())); }
template <int R, class A> PC_DEV double d6z(const A &a, double f1)
{ return f1 * (-20.0 * a.template c<R>() + 15.0 * (a.template z<R, 1>() + a.template z<R, -1>()) - 6.0 * (a.template z<R, 2>() + a.template z<R, -2>()) + (a.template z<R, 3>() + a.template z<R, -3>())); }

// mixed derivatives (derij_main, deriv.f90:1343-1580).  a.bidiag(): lbidiagonal_derij=T, the
// 12-point scheme (:1385-1423); otherwise the 36-point product of two first-derivative
// stencils (:1430-1557), inner differences along x (xy, zx) or y (yz), outer along y or z.
// fij = d_i d_j/720 in both cases (the product stencil has 1/3600 = 0.2/720).
#define PC_DROW(ACC, K2) (45.0 * (a.template ACC<R, 1, K2>() - a.template ACC<R, -1, K2>()) \
                          - 9.0 * (a.template ACC<R, 2, K2>() - a.template ACC<R, -2, K2>()) \
                          + (a.template ACC<R, 3, K2>() - a.template ACC<R, -3, K2>()))
#define PC_DFULL(ACC) (0.2 * fij) * (45.0 * (PC_DROW(ACC, 1) - PC_DROW(ACC, -1)) - 9.0 * (PC_DROW(ACC, 2) - PC_DROW(ACC, -2)) \
                                     + (PC_DROW(ACC, 3) - PC_DROW(ACC, -3)))
template <int R, class A> PC_DEV double dxy(const A &a, const double *k, double fij)
{
  if (!a.bidiag()) return PC_DFULL(xy);
  if (!A::PS)
    return fij * (270.0 * (a.template xy<R, 1, 1>() - a.template xy<R, -1, 1>() + a.template xy<R, -1, -1>() - a.template xy<R, 1, -1>())
                  - 27.0 * (a.template xy<R, 2, 2>() - a.template xy<R, -2, 2>() + a.template xy<R, -2, -2>() - a.template xy<R, 2, -2>())
                  + 2.0 * (a.template xy<R, 3, 3>() - a.template xy<R, -3, 3>() + a.template xy<R, -3, -3>() - a.template xy<R, 3, -3>()));
  return fma(k[2], a.template xy<R, 3, 3>() - a.template xy<R, -3, 3>() + a.template xy<R, -3, -3>() - a.template xy<R, 3, -3>(),
         fma(k[1], a.template xy<R, 2, 2>() - a.template xy<R, -2, 2>() + a.template xy<R, -2, -2>() - a.template xy<R, 2, -2>(),
             k[0] * (a.template xy<R, 1, 1>() - a.template xy<R, -1, 1>() + a.template xy<R, -1, -1>() - a.template xy<R, 1, -1>())));
}
template <int R, class A> PC_DEV double dxz(const A &a, const double *k, double fij)
{
  if (A::NOZMIX) return 0.0;
  if (!a.bidiag()) return PC_DFULL(xz);
  if (!A::PS)
    return fij * (270.0 * (a.template xz<R, 1, 1>() - a.template xz<R, -1, 1>() + a.template xz<R, -1, -1>() - a.template xz<R, 1, -1>())
                  - 27.0 * (a.template xz<R, 2, 2>() - a.template xz<R, -2, 2>() + a.template xz<R, -2, -2>() - a.template xz<R, 2, -2>())
                  + 2.0 * (a.template xz<R, 3, 3>() - a.template xz<R, -3, 3>() + a.template xz<R, -3, -3>() - a.template xz<R, 3, -3>()));
  return fma(k[2], a.template xz<R, 3, 3>() - a.template xz<R, -3, 3>() + a.template xz<R, -3, -3>() - a.template xz<R, 3, -3>(),
         fma(k[1], a.template xz<R, 2, 2>() - a.template xz<R, -2, 2>() + a.template xz<R, -2, -2>() - a.template xz<R, 2, -2>(),
             k[0] * (a.template xz<R, 1, 1>() - a.template xz<R, -1, 1>() + a.template xz<R, -1, -1>() - a.template xz<R, 1, -1>())));
}
template <int R, class A> PC_DEV double dyz(const A &a, const double *k, double fij)
{
  if (A::NOZMIX) return 0.0;
  if (!a.bidiag()) return PC_DFULL(yz);
  if (!A::PS)
    return fij * (270.0 * (a.template yz<R, 1, 1>() - a.template yz<R, -1, 1>() + a.template yz<R, -1, -1>() - a.template yz<R, 1, -1>())
                  - 27.0 * (a.template yz<R, 2, 2>() - a.template yz<R, -2, 2>() + a.template yz<R, -2, -2>() - a.template yz<R, 2, -2>())
                  + 2.0 * (a.template yz<R, 3, 3>() - a.template yz<R, -3, 3>() + a.template yz<R, -3, -3>() - a.template yz<R, 3, -3>()));
  return fma(k[2], a.template yz<R, 3, 3>() - a.template yz<R, -3, 3>() + a.template yz<R, -3, -3>() - a.template yz<R, 3, -3>(),
         fma(k[1], a.template yz<R, 2, 2>() - a.template yz<R, -2, 2>() + a.template yz<R, -2, -2>() - a.template yz<R, 2, -2>(),
             k[0] * (a.template yz<R, 1, 1>() - a.template yz<R, -1, 1>() + a.template yz<R, -1, -1>() - a.template yz<R, 1, -1>())));
}
#undef PC_DFULL
#undef PC_DROW

// per-cell by-products for the CFL step and the 0-D diagnostics
struct CellAux {
  double dt1;        // local inverse time step (equ.f90:1149)
  double u2, rho, ss, b2, j2, ab, jb, ou;
  double dtc, dtu, dtnu, dtchi;
};

// what the induction part hands to the momentum/entropy part
struct MagOut {
  double fx, fy, fz;     // Lorentz force J x B / rho (magnetic.f90:4360)
  double j2;             // for Ohmic heating (magnetic.f90:5127)
  double advec_va2;      // Alfven CFL term (magnetic.f90:5438)
  double b2, ab, jb;     // diagnostics
  double bx, by, bz;     // B itself (rhs_dly finishes the Lorentz force three planes late)
  double jx, jy, jz;     // J (rhs_dly: its centre-plane part; Ohmic heating is finished three planes late)
};

// ---------------- magnetic (magnetic.f90:3587-3990, 4229-5589) ----------------
// rhsA[0..2] = dA/dt = u x B + eta del2 A (magnetic.f90:4399,5238,5521)
template <bool AUX, class A>
PC_DEV void mag_part(const A &a, const DevPar &c, double ux, double uy, double uz, double rho1,
                     double *__restrict__ rhsA, MagOut &m)
{
  const double axy = ddy<R_AX>(a, c), axz = ddz<R_AX>(a, c);
  const double ayx = ddx<R_AY>(a, c), ayz = ddz<R_AY>(a, c);
  const double azx = ddx<R_AZ>(a, c), azy = ddy<R_AZ>(a, c);
  const double bx = azy - ayz, by = axz - azx, bz = ayx - axy;   // curl_mn
  // second derivatives (gij_etc, sub.f90:2732): del2 A and J = curl curl A / mu0.  jj(1)=bij(3,2)-bij(2,3)
  // etc. with bij(i,l)=eps d2A (sub.f90:2773-2781, magnetic.f90:3787) = mixed terms minus the
  // pair sums d2y+d2z (Ax), d2z+d2x (Ay), d2x+d2y (Az), which del2 A needs too
  const double axxx = d2x<R_AX>(a, c), ax_yz = p2yz<R_AX>(a, c);
  const double ayyy = d2y<R_AY>(a, c), ay_zx = p2zx<R_AY>(a, c);
  const double azzz = d2z<R_AZ>(a, c), az_xy = p2xy<R_AZ>(a, c);
  // A::NOZMIX (rhs_dly): the z-mixed derivatives are completed three planes late, not here
  const double jx = c.mu01 * (A::NOZMIX ? dxy<R_AY>(a, c.kxy, c.fxy) - ax_yz : (dxy<R_AY>(a, c.kxy, c.fxy) + dxz<R_AZ>(a, c.kzx, c.fzx)) - ax_yz);
  const double jy = c.mu01 * (A::NOZMIX ? dxy<R_AX>(a, c.kxy, c.fxy) - ay_zx : (dyz<R_AZ>(a, c.kyz, c.fyz) + dxy<R_AX>(a, c.kxy, c.fxy)) - ay_zx);
  const double jz = c.mu01 * (A::NOZMIX ? -az_xy : (dxz<R_AX>(a, c.kzx, c.fzx) + dyz<R_AY>(a, c.kyz, c.fyz)) - az_xy);
  m.bx = bx; m.by = by; m.bz = bz;
  m.jx = jx; m.jy = jy; m.jz = jz;
  m.j2 = jx * jx + jy * jy + jz * jz;
  m.fx = rho1 * (jy * bz - jz * by);
  m.fy = rho1 * (jz * bx - jx * bz);
  m.fz = rho1 * (jx * by - jy * bx);
  rhsA[0] = (uy * bz - uz * by) + c.eta * (axxx + ax_yz);
  rhsA[1] = (uz * bx - ux * bz) + c.eta * (ayyy + ay_zx);
  rhsA[2] = (ux * by - uy * bx) + c.eta * (azzz + az_xy);
  if (AUX) {
    const double t0 = bx * c.dx_1, t1 = by * c.dy_1, t2 = bz * c.dz_1;
    m.advec_va2 = (t0 * t0 + t1 * t1 + t2 * t2) * c.mu01 * rho1;
    m.b2 = bx * bx + by * by + bz * bz;
    const double ax = a.template c<R_AX>(), ay = a.template c<R_AY>(), az = a.template c<R_AZ>();
    m.ab = ax * bx + ay * by + az * bz;
    m.jb = jx * bx + jy * by + jz * bz;
  }
}

// ---------------- momentum, continuity, entropy + CFL ----------------
// rhs[R_UX..R_UZ], rhs[R_LNRHO], rhs[R_SS]; m = output of mag_part (zeros without magnetic)
// iz = z index of the cell relative to n1 (for the 1-D z profiles).
// quantities the hydro part keeps until the induction part's results are in
struct HydroPend { double ohm, advec_uu, advec_cs2, dt1_diffus, hm; };

template <bool MAG, bool ENT, bool AUX, class A>
PC_DEV void hydro_pre(const A &a, const DevPar &c, int iz, double *__restrict__ rhs, CellAux &aux, HydroPend &p)
{
  // ---------------- hydro pencils (hydro.f90:2913-3207) ----------------
  const double ux = a.template c<R_UX>(), uy = a.template c<R_UY>(), uz = a.template c<R_UZ>();
  const double uxx = ddx<R_UX>(a, c), uxy = ddy<R_UX>(a, c), uxz = ddz<R_UX>(a, c);
  const double uyx = ddx<R_UY>(a, c), uyy = ddy<R_UY>(a, c), uyz = ddz<R_UY>(a, c);
  const double uzx = ddx<R_UZ>(a, c), uzy = ddy<R_UZ>(a, c), uzz = ddz<R_UZ>(a, c);
  const double divu = uxx + uyy + uzz;
  // traceless rate of strain (sub.f90:8314)
  const double third = 1.0 / 3.0;
  const double sxx = uxx - third * divu, syy = uyy - third * divu, szz = uzz - third * divu;
  const double sxy = 0.5 * (uxy + uyx), sxz = 0.5 * (uxz + uzx), syz = 0.5 * (uyz + uzy);
  // u.grad u (sub.f90:3227), optional upwinding (sub.f90:6941)
  double ugux = ux * uxx + uy * uxy + uz * uxz;
  double uguy = ux * uyx + uy * uyy + uz * uyz;
  double uguz = ux * uzx + uy * uzy + uz * uzz;
  const double aux_ = fabs(ux), auy_ = fabs(uy), auz_ = fabs(uz);
  if (A::UPW && c.lupw_uu) {
    ugux -= aux_ * d6x<R_UX>(a, c.fx1) + auy_ * d6y<R_UX>(a, c.fy1) + auz_ * d6z<R_UX>(a, c.fz1);
    uguy -= aux_ * d6x<R_UY>(a, c.fx1) + auy_ * d6y<R_UY>(a, c.fy1) + auz_ * d6z<R_UY>(a, c.fz1);
    uguz -= aux_ * d6x<R_UZ>(a, c.fx1) + auy_ * d6y<R_UZ>(a, c.fy1) + auz_ * d6z<R_UZ>(a, c.fz1);
  }
  // del2 u and grad div u (del2v_etc, sub.f90:2397)
  const double uxxx = d2x<R_UX>(a, c), uyyy = d2y<R_UY>(a, c), uzzz = d2z<R_UZ>(a, c);
  const double del2ux = uxxx + p2yz<R_UX>(a, c);
  const double del2uy = uyyy + p2zx<R_UY>(a, c);
  const double del2uz = uzzz + p2xy<R_UZ>(a, c);
  const double gdx = A::NOZMIX ? uxxx + dxy<R_UY>(a, c.kxy, c.fxy) : uxxx + dxy<R_UY>(a, c.kxy, c.fxy) + dxz<R_UZ>(a, c.kzx, c.fzx);
  const double gdy = A::NOZMIX ? dxy<R_UX>(a, c.kxy, c.fxy) + uyyy : dxy<R_UX>(a, c.kxy, c.fxy) + uyyy + dyz<R_UZ>(a, c.kyz, c.fyz);
  const double gdz = A::NOZMIX ? uzzz : dxz<R_UX>(a, c.kzx, c.fzx) + dyz<R_UY>(a, c.kyz, c.fyz) + uzzz;

  // ---------------- density pencils (density.f90:2341) ----------------
  const double lnrho = a.template c<R_LNRHO>();
  const double rho1 = pc_exp(-lnrho);
  const double glx = ddx<R_LNRHO>(a, c), gly = ddy<R_LNRHO>(a, c), glz = ddz<R_LNRHO>(a, c);
  double uglnrho = ux * glx + uy * gly + uz * glz;
  if (A::UPW && c.lupw_lnrho)
    uglnrho -= aux_ * d6x<R_LNRHO>(a, c.fx1) + auy_ * d6y<R_LNRHO>(a, c.fy1) + auz_ * d6z<R_LNRHO>(a, c.fz1);
  const double sglx = sxx * glx + sxy * gly + sxz * glz;
  const double sgly = sxy * glx + syy * gly + syz * glz;
  const double sglz = sxz * glx + syz * gly + szz * glz;

  // ---------------- EOS + energy pencils ----------------
  double cs2 = c.cs20, TT1 = 0.0, ee = 1.0, ss = 0.0;
  double gTx = 0.0, gTy = 0.0, gTz = 0.0, del2lnTT = 0.0, ugss = 0.0;
  double fpx, fpy, fpz;
  if (ENT) {
    // eos_idealgas.f90:843-868
    ss = a.template c<R_SS>();
    const double gsx = ddx<R_SS>(a, c), gsy = ddy<R_SS>(a, c), gsz = ddz<R_SS>(a, c);
    const double del2ss = del2<R_SS>(a, c);
    const double del2lnrho = del2<R_LNRHO>(a, c);
    const double th = c.cv1 * ss + c.gamma_m1 * (lnrho - c.lnrho0);
    const double eth = exp(th);
    cs2 = c.cs20 * eth;
    const double lnTT = c.lnTT0 + th;
    const double TT = exp(lnTT);
    ee = c.cv * TT;
    TT1 = 1.0 / TT;
    gTx = c.gamma_m1 * glx + c.cv1 * gsx; gTy = c.gamma_m1 * gly + c.cv1 * gsy; gTz = c.gamma_m1 * glz + c.cv1 * gsz;
    del2lnTT = c.gamma_m1 * del2lnrho + c.cv1 * del2ss;
    ugss = ux * gsx + uy * gsy + uz * gsz;                         // entropy.f90:3189
    if (A::UPW && c.lupw_ss)
      ugss -= aux_ * d6x<R_SS>(a, c.fx1) + auy_ * d6y<R_SS>(a, c.fy1) + auz_ * d6z<R_SS>(a, c.fz1);
    const double pf = -cs2 * c.gamma1;                             // entropy.f90:3209
    fpx = pf * (glx + gTx); fpy = pf * (gly + gTy); fpz = pf * (glz + gTz);
  } else {
    // noentropy with gamma /= 1 selects the isentropic eos (noentropy.f90:113-116, leos_isentropic branch of
    // calc_pencils_eos, eos_idealgas.f90:809-815): cs2 = cs20*exp(gamma_m1*(lnrho-lnrho0)), glnTT = gamma_m1*glnrho,
    // fpres = -cs2*(glnrho+glnTT)/gamma.  Compiled in with A::UPW only (the general variant of rhs_dly).
    if (A::UPW && c.gamma_m1 != 0.0) {
      cs2 = c.cs20 * pc_exp(c.gamma_m1 * (lnrho - c.lnrho0));
      const double pf = -cs2 * c.gamma1;
      fpx = pf * (glx + c.gamma_m1 * glx); fpy = pf * (gly + c.gamma_m1 * gly); fpz = pf * (glz + c.gamma_m1 * glz);
    } else {
      fpx = -cs2 * glx; fpy = -cs2 * gly; fpz = -cs2 * glz;        // noentropy.f90:291
    }
  }

  // ---------------- momentum (hydro.f90:3704, viscosity.f90:1421-1448,2571) ----------------
  double dux = -ugux + c.Omega2 * uy + c.nu * (del2ux + 2.0 * sglx + third * gdx) + fpx;
  double duy = -uguy - c.Omega2 * ux + c.nu * (del2uy + 2.0 * sgly + third * gdy) + fpy;
  double duz = -uguz + c.nu * (del2uz + 2.0 * sglz + third * gdz) + fpz;
  if (c.lgravz) duz += c.gravz;                                    // gravity_simple.f90:914
  rhs[R_UX] = dux; rhs[R_UY] = duy; rhs[R_UZ] = duz;

  // ---------------- continuity (density.f90:2542-2623) ----------------
  rhs[R_LNRHO] = -divu - uglnrho;

  // ---------------- entropy equation (entropy.f90:3242-3452) ----------------
  double Hmax = 0.0, diffus_chi = 0.0;
  if (ENT) {
    const double sij2 = sxx * sxx + syy * syy + szz * szz + 2.0 * (sxy * sxy + sxz * sxz + syz * syz);
    double dss = -ugss;
    const double visc_heat = 2.0 * c.nu * sij2;                    // viscosity.f90:1447
    if (c.lviscosity_heat) { dss += TT1 * visc_heat; Hmax += visc_heat; }
    if (c.iheatcond == PCB200_HCOND_KPROFILE) {                    // calc_heatcond :5333
      const double hcond = c.zprof[PCB200_PROF_HCOND][iz];
      const double glhc = c.zprof[PCB200_PROF_DLNHCOND][iz];
      const double g2 = gTx * gTx + gTy * gTy + gTz * (gTz + glhc);
      dss += rho1 * hcond * (del2lnTT + g2);
      diffus_chi += c.gamma * (rho1 * hcond * c.cp1) * c.dxyz_2;
    } else if (c.iheatcond == PCB200_HCOND_CHICONST) {             // calc_heatcond_constchi :4266
      const double g2 = (glx + gTx) * gTx + (gly + gTy) * gTy + (glz + gTz) * gTz;
      dss += c.chi * c.cp * (del2lnTT + g2);
      diffus_chi += (c.gamma * c.chi) * c.dxyz_2;
    }
    if (c.lcooling) {                                              // get_heat_cool_gravz :6196
      const double heat = -c.cool * c.zprof[PCB200_PROF_COOL][iz] * (cs2 - c.cs2cool) / c.cs2cool;
      dss += TT1 * rho1 * heat;
      Hmax += heat * rho1;
    }
    p.ohm = (MAG && c.lohmic_heat) ? c.eta * c.mu0 * rho1 * TT1 : 0.0;   // magnetic.f90:5127
    rhs[R_SS] = dss;
  }

  // ---------------- time-step constraint, part 1 (equ.f90:1086-1151) ----------------
  if (AUX) {
    p.advec_uu = aux_ * c.dx_1 + auy_ * c.dy_1 + auz_ * c.dz_1;                  // hydro.f90:3810
    p.advec_cs2 = cs2 * c.dxyz_2;
    double maxdiffus = c.nu * c.dxyz_2;
    if (MAG) maxdiffus = fmax(maxdiffus, c.eta * c.dxyz_2);
    if (ENT) maxdiffus = fmax(maxdiffus, diffus_chi);
    p.dt1_diffus = maxdiffus / c.cdtv;
    p.hm = ENT ? fabs(Hmax / ee / c.cdts) : 0.0;                                 // entropy.f90:3444
    aux.u2 = ux * ux + uy * uy + uz * uz;
    aux.rho = 1.0 / rho1;
    aux.ss = ss;
    aux.ou = (uzy - uyz) * ux + (uxz - uzx) * uy + (uyx - uxy) * uz;
    aux.dtc = sqrt(p.advec_cs2) / c.cdt;
    aux.dtu = p.advec_uu / c.cdt;
    aux.dtnu = c.nu * c.dxyz_2 / c.cdtv;
    aux.dtchi = diffus_chi / c.cdtv;
  }
}

// adds what depends on the induction part: Lorentz force, Ohmic heating, Alfven CFL term
// OHM: Ohmic heating added here (rhs_dly adds it once the whole J is known)
template <bool MAG, bool ENT, bool AUX, bool OHM = true>
PC_DEV void hydro_finish(const DevPar &c, const HydroPend &p, const MagOut &m, double *__restrict__ rhs, CellAux &aux)
{
  if (MAG) {
    rhs[R_UX] += m.fx; rhs[R_UY] += m.fy; rhs[R_UZ] += m.fz;
    if (ENT && OHM) rhs[R_SS] += p.ohm * m.j2;
  }
  if (AUX) {
    const double maxadvec = p.advec_uu + sqrt((MAG ? m.advec_va2 : 0.0) + p.advec_cs2);
    const double dt1_advec = maxadvec / c.cdt;
    double dt1 = sqrt(dt1_advec * dt1_advec + p.dt1_diffus * p.dt1_diffus);
    if (ENT) dt1 = fmax(dt1, p.hm);
    aux.dt1 = dt1;
    aux.b2 = MAG ? m.b2 : 0.0; aux.j2 = MAG ? m.j2 : 0.0; aux.ab = MAG ? m.ab : 0.0; aux.jb = MAG ? m.jb : 0.0;
  }
}

// one thread does both parts (global-memory kernel, non-specialised marching kernel)
template <bool MAG, bool ENT, bool AUX, class A>
PC_DEV void rhs_cell(const A &a, const DevPar &c, int iz, double *__restrict__ rhs, CellAux &aux)
{
  MagOut m;
  m.fx = m.fy = m.fz = m.j2 = m.advec_va2 = m.b2 = m.ab = m.jb = 0.0;
  if (MAG) {
    const double rho1 = pc_exp(-a.template c<R_LNRHO>());
    mag_part<AUX>(a, c, a.template c<R_UX>(), a.template c<R_UY>(), a.template c<R_UZ>(), rho1, rhs + R_AX, m);
  }
  HydroPend p;
  hydro_pre<MAG, ENT, AUX>(a, c, iz, rhs, aux, p);
  hydro_finish<MAG, ENT, AUX>(c, p, m, rhs, aux);
}
