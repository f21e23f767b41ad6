/* pcref_pencil.c -- PENCIL-WISE evaluation of the same right-hand side, for TIMING the CPU arm only.
 *
 * Test infrastructure like everything under oracle/ (never shipped, never the checker: the checker is
 * pcref.c, which evaluates one grid point at a time in the reference's order of floating-point
 * operations).  The reference itself works on x-pencils: every derivative of a variable is formed for a
 * whole row f(l1:l2,m,n) at once (deriv.f90:89-171 der_main etc. operate on (nx) arrays inside the
 * mn-loop of rhs_cpu, equ.f90:816-1239), which is what lets a Fortran compiler vectorise it.  A CPU
 * baseline that re-computes index arithmetic per point undersells the reference, so bench.py times THIS
 * routine where it applies:
 *      HYDRO + DENSITY [+ MAGNETIC], no entropy module, gamma = 1, lbidiagonal_derij = T, no upwinding
 * (BASELINE configs 2 and 3, the configurations the headline metric is quoted on); everything else falls
 * back to pcref_timestep.  Same stencils, same physics, same RK3/2N scheme; only the association of a few
 * sums differs from pcref.c (checked to 1e-12 against it by tests/test_oracle_pencil.py).
 */
#include "pcref.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define NG PCB200_NGHOST

/* one row of a first derivative: stride s in elements, fac = d_1/60 */
#define ROW_D1(out, p, s, fac) \
  for (int l = 0; l < nx; l++) (out)[l] = (fac) * (45.0 * ((p)[l + (s)] - (p)[l - (s)]) - 9.0 * ((p)[l + 2 * (s)] - (p)[l - 2 * (s)]) \
                                                   + ((p)[l + 3 * (s)] - (p)[l - 3 * (s)]));
/* second derivative: fac = d_1^2 */
#define ROW_D2(out, p, s, fac) \
  for (int l = 0; l < nx; l++) (out)[l] = (fac) * (c0 * (p)[l] + c1 * ((p)[l + (s)] + (p)[l - (s)]) + c2 * ((p)[l + 2 * (s)] + (p)[l - 2 * (s)]) \
                                                   + c3 * ((p)[l + 3 * (s)] + (p)[l - 3 * (s)]));
/* bidiagonal mixed derivative (deriv.f90:1385-1423): strides a, b; fac = da_1*db_1/720 */
#define ROW_DIJ(out, p, a, b, fac) \
  for (int l = 0; l < nx; l++) (out)[l] = (fac) * (270.0 * ((p)[l + (a) + (b)] - (p)[l - (a) + (b)] + (p)[l - (a) - (b)] - (p)[l + (a) - (b)]) \
                                                   - 27.0 * ((p)[l + 2 * (a) + 2 * (b)] - (p)[l - 2 * (a) + 2 * (b)] + (p)[l - 2 * (a) - 2 * (b)] - (p)[l + 2 * (a) - 2 * (b)]) \
                                                   + 2.0 * ((p)[l + 3 * (a) + 3 * (b)] - (p)[l - 3 * (a) + 3 * (b)] + (p)[l - 3 * (a) - 3 * (b)] - (p)[l + 3 * (a) - 3 * (b)]));

int pcref_pencil_applies(const pcb200_params *par)
{
  return par->lhydro && par->ldensity && !par->lentropy && par->gamma == 1.0 && par->lbidiagonal_derij &&
         !par->lupw_uu && !par->lupw_lnrho && par->nx > 1 && par->ny > 1 && par->nz > 1;
}

/* df = alpha*df + F(f) over the interior, pencil by pencil; returns max dt1 when ldt1 */
static double rhs_pencils(const double *f, double *df, const pcb200_params *par, double alpha, int ldt1)
{
  const int nx = par->nx, ny = par->ny, nz = par->nz;
  const int mx = nx + 2 * NG, my = ny + 2 * NG, mz = nz + 2 * NG;
  const long sy = mx, sz = (long)mx * my, mw = sz * mz;
  const int M = par->lmagnetic;
  const int iu = par->iux - 1, ilr = par->ilnrho - 1, ia = par->iax - 1;
  const double c0 = -490. / 180., c1 = 270. / 180., c2 = -27. / 180., c3 = 2. / 180.;
  const double dx_1 = par->dx_1, dy_1 = par->dy_1, dz_1 = par->dz_1;
  const double f1[3] = {dx_1 / 60.0, dy_1 / 60.0, dz_1 / 60.0};
  const double f2[3] = {dx_1 * dx_1, dy_1 * dy_1, dz_1 * dz_1};
  const double fxy = dx_1 * dy_1 / 720.0, fyz = dy_1 * dz_1 / 720.0, fzx = dz_1 * dx_1 / 720.0;
  const long st[3] = {1, sy, sz};
  const double dxyz_2 = dx_1 * dx_1 + dy_1 * dy_1 + dz_1 * dz_1;
  const double nu = par->nu, eta = par->eta, mu01 = 1.0 / par->mu0, cs2 = par->cs20, third = 1.0 / 3.0;
  const double Om2 = 2.0 * par->Omega;
  double dt1 = 0.0;
#pragma omp parallel reduction(max : dt1)
  {
    /* pencils of this thread: 9 du, 9 d2u, 6 mixed u, 3 glnrho; the same for A */
    double *buf = (double *)malloc(sizeof(double) * (size_t)nx * 64);
    double *uij[3][3], *u2d[3][3], *umx[6], *gl[3], *aij[3][3], *a2d[3][3], *amx[6], *r1;
    {
      double *q = buf;
      for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { uij[i][j] = q; q += nx; u2d[i][j] = q; q += nx; aij[i][j] = q; q += nx; a2d[i][j] = q; q += nx; }
      for (int k = 0; k < 6; k++) { umx[k] = q; q += nx; amx[k] = q; q += nx; }
      for (int j = 0; j < 3; j++) { gl[j] = q; q += nx; }
      r1 = q;
    }
#pragma omp for collapse(2) schedule(static)
    for (int n = 0; n < nz; n++) {
      for (int m = 0; m < ny; m++) {
        const long row = ((long)(n + NG) * my + (m + NG)) * mx + NG;
        const double *pu[3] = {f + (long)iu * mw + row, f + (long)(iu + 1) * mw + row, f + (long)(iu + 2) * mw + row};
        const double *pl = f + (long)ilr * mw + row;
        for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
          ROW_D1(uij[i][j], pu[i], st[j], f1[j])
          ROW_D2(u2d[i][j], pu[i], st[j], f2[j])
        }
        /* grad div u needs d2 u_j / dx_j dx_i: (x: uy_xy, uz_xz), (y: ux_xy, uz_yz), (z: ux_xz, uy_yz) */
        ROW_DIJ(umx[0], pu[1], 1, sy, fxy) ROW_DIJ(umx[1], pu[2], 1, sz, fzx)
        ROW_DIJ(umx[2], pu[0], 1, sy, fxy) ROW_DIJ(umx[3], pu[2], sy, sz, fyz)
        ROW_DIJ(umx[4], pu[0], 1, sz, fzx) ROW_DIJ(umx[5], pu[1], sy, sz, fyz)
        for (int j = 0; j < 3; j++) { ROW_D1(gl[j], pl, st[j], f1[j]) }
        const double *pa[3] = {0, 0, 0};
        if (M) {
          for (int i = 0; i < 3; i++) pa[i] = f + (long)(ia + i) * mw + row;
          for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
            if (i != j) { ROW_D1(aij[i][j], pa[i], st[j], f1[j]) }
            ROW_D2(a2d[i][j], pa[i], st[j], f2[j])
          }
          /* J = curl curl A / mu0: jx = (Ay_xy + Az_xz) - (Ax_yy + Ax_zz), ... */
          ROW_DIJ(amx[0], pa[1], 1, sy, fxy) ROW_DIJ(amx[1], pa[2], 1, sz, fzx)
          ROW_DIJ(amx[2], pa[0], 1, sy, fxy) ROW_DIJ(amx[3], pa[2], sy, sz, fyz)
          ROW_DIJ(amx[4], pa[0], 1, sz, fzx) ROW_DIJ(amx[5], pa[1], sy, sz, fyz)
          for (int l = 0; l < nx; l++) r1[l] = exp(-pl[l]);
        }
        double *du[3] = {df + (long)iu * mw + row, df + (long)(iu + 1) * mw + row, df + (long)(iu + 2) * mw + row};
        double *dl = df + (long)ilr * mw + row;
        double *da[3] = {0, 0, 0};
        if (M) for (int i = 0; i < 3; i++) da[i] = df + (long)(ia + i) * mw + row;
        double rowmax = 0.0;
        for (int l = 0; l < nx; l++) {
          const double ux = pu[0][l], uy = pu[1][l], uz = pu[2][l];
          const double uxx = uij[0][0][l], uxy = uij[0][1][l], uxz = uij[0][2][l];
          const double uyx = uij[1][0][l], uyy = uij[1][1][l], uyz = uij[1][2][l];
          const double uzx = uij[2][0][l], uzy = uij[2][1][l], uzz = uij[2][2][l];
          const double divu = uxx + uyy + uzz;
          const double sxx = uxx - third * divu, syy = uyy - third * divu, szz = uzz - third * divu;
          const double sxy = 0.5 * (uxy + uyx), sxz = 0.5 * (uxz + uzx), syz = 0.5 * (uyz + uzy);
          const double glx = gl[0][l], gly = gl[1][l], glz = gl[2][l];
          const double del2x = u2d[0][0][l] + u2d[0][1][l] + u2d[0][2][l];
          const double del2y = u2d[1][0][l] + u2d[1][1][l] + u2d[1][2][l];
          const double del2z = u2d[2][0][l] + u2d[2][1][l] + u2d[2][2][l];
          const double gdx = u2d[0][0][l] + umx[0][l] + umx[1][l];
          const double gdy = umx[2][l] + u2d[1][1][l] + umx[3][l];
          const double gdz = umx[4][l] + umx[5][l] + u2d[2][2][l];
          double rx = -(ux * uxx + uy * uxy + uz * uxz) + Om2 * uy + nu * (del2x + 2.0 * (sxx * glx + sxy * gly + sxz * glz) + third * gdx) - cs2 * glx;
          double ry = -(ux * uyx + uy * uyy + uz * uyz) - Om2 * ux + nu * (del2y + 2.0 * (sxy * glx + syy * gly + syz * glz) + third * gdy) - cs2 * gly;
          double rz = -(ux * uzx + uy * uzy + uz * uzz) + nu * (del2z + 2.0 * (sxz * glx + syz * gly + szz * glz) + third * gdz) - cs2 * glz;
          if (par->lgravz) rz += par->gravz;
          const double rl = -divu - (ux * glx + uy * gly + uz * glz);
          double va2 = 0.0;
          if (M) {
            const double bx = aij[2][1][l] - aij[1][2][l], by = aij[0][2][l] - aij[2][0][l], bz = aij[1][0][l] - aij[0][1][l];
            const double jx = mu01 * ((amx[0][l] + amx[1][l]) - (a2d[0][1][l] + a2d[0][2][l]));
            const double jy = mu01 * ((amx[2][l] + amx[3][l]) - (a2d[1][0][l] + a2d[1][2][l]));
            const double jz = mu01 * ((amx[4][l] + amx[5][l]) - (a2d[2][0][l] + a2d[2][1][l]));
            const double rho1 = r1[l];
            rx += rho1 * (jy * bz - jz * by);
            ry += rho1 * (jz * bx - jx * bz);
            rz += rho1 * (jx * by - jy * bx);
            const double rax = (uy * bz - uz * by) + eta * (a2d[0][0][l] + a2d[0][1][l] + a2d[0][2][l]);
            const double ray = (uz * bx - ux * bz) + eta * (a2d[1][0][l] + a2d[1][1][l] + a2d[1][2][l]);
            const double raz = (ux * by - uy * bx) + eta * (a2d[2][0][l] + a2d[2][1][l] + a2d[2][2][l]);
            da[0][l] = alpha * da[0][l] + rax; da[1][l] = alpha * da[1][l] + ray; da[2][l] = alpha * da[2][l] + raz;
            if (ldt1) { const double t0 = bx * dx_1, t1 = by * dy_1, t2 = bz * dz_1; va2 = (t0 * t0 + t1 * t1 + t2 * t2) * mu01 * rho1; }
          }
          du[0][l] = alpha * du[0][l] + rx; du[1][l] = alpha * du[1][l] + ry; du[2][l] = alpha * du[2][l] + rz;
          dl[l] = alpha * dl[l] + rl;
          if (ldt1) {
            const double maxadvec = (fabs(ux) * dx_1 + fabs(uy) * dy_1 + fabs(uz) * dz_1) + sqrt(va2 + cs2 * dxyz_2);
            if (maxadvec > rowmax) rowmax = maxadvec;
          }
        }
        if (ldt1) {   /* equ.f90:1149, monotonic in maxadvec: applied once per row */
          const double maxdiffus = M ? fmax(nu * dxyz_2, eta * dxyz_2) : nu * dxyz_2;
          const double a = rowmax / par->cdt, d = maxdiffus / par->cdtv;
          const double v = sqrt(a * a + d * d);
          if (v > dt1) dt1 = v;
        }
      }
    }
    free(buf);
  }
  return dt1;
}

/* same contract as pcref_timestep (diag must be NULL here: the diagnostics stay with the checker) */
void pcref_timestep_pencil(double *f, double *df, const pcb200_params *par,
                           const pcb200_forcing *frc, double *dt, double *diag)
{
  if (diag || !pcref_pencil_applies(par)) { pcref_timestep(f, df, par, frc, dt, diag); return; }
  const int mx = par->nx + 2 * NG, my = par->ny + 2 * NG, mz = par->nz + 2 * NG;
  const long mw = (long)mx * my * mz;
  for (int itsub = 0; itsub < 3; itsub++) {
    const int lfirst = itsub == 0;
    pcref_boundconds(f, par);
    const double alpha = lfirst ? 0.0 : par->alpha_ts[itsub];
    if (lfirst) memset(df, 0, sizeof(double) * (size_t)mw * par->mvar);       /* alpha = 0: old df must not leak NaNs */
    double dt1 = rhs_pencils(f, df, par, alpha, lfirst && par->ldt);
    if (lfirst && par->ldt) { const double fl = par->dtmax != 0.0 ? 1.0 / par->dtmax : 0.0; *dt = 1.0 / fmax(dt1, fl); }
    if (lfirst && !par->ldt) *dt = par->dt_fixed;
    const double dt_beta = *dt * par->beta_ts[itsub];
#pragma omp parallel for collapse(2) schedule(static)
    for (int k = 0; k < par->mvar; k++) for (int n = NG; n < NG + par->nz; n++)
      for (int m = NG; m < NG + par->ny; m++) {
        const long o = (long)k * mw + ((long)n * my + m) * mx + NG;
        for (int l = 0; l < par->nx; l++) f[o + l] = f[o + l] + dt_beta * df[o + l];
      }
  }
  if (frc) pcref_addforce(f, par, frc);
}
