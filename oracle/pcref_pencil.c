/* pcref_pencil.c -- PENCIL-WISE evaluation of the same right-hand side, for TIMING the CPU arm only.
 *
 * Test infrastructure like everything under oracle/ (never shipped, never the checker: the checker is
 * pcref.c, which evaluates one grid point at a time in the reference's order of floating-point
 * operations).  The reference itself works on x-pencils: every derivative of a variable is formed for a
 * whole row f(l1:l2,m,n) at once (deriv.f90:89-171 der_main etc. operate on (nx) arrays inside the
 * mn-loop of rhs_cpu, equ.f90:816-1239), which is what lets a Fortran compiler vectorise it.  A CPU
 * baseline that re-computes index arithmetic per point undersells the reference, so bench.py times THIS
 * routine where it applies:
 *      HYDRO + DENSITY [+ MAGNETIC] [+ ENTROPY with iheatcond none/K-profile/chi-const, cooling layer,
 *      viscous and Ohmic heating], upwinding, gravz, Coriolis, lbidiagonal_derij = T
 * (BASELINE configs 1-4); runs without the entropy module and gamma /= 1, one-point directions and
 * lbidiagonal_derij = F fall back to pcref_timestep.  Same stencils, same physics, same RK3/2N scheme; only the association of a few
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
/* upwind correction, der6 with upwind=T (deriv.f90:938-1048): fac = d_1/60 */
#define ROW_D6U(out, p, s, fac) \
  for (int l = 0; l < nx; l++) (out)[l] = (fac) * (-20.0 * (p)[l] + 15.0 * ((p)[l + (s)] + (p)[l - (s)]) - 6.0 * ((p)[l + 2 * (s)] + (p)[l - 2 * (s)]) \
                                                   + ((p)[l + 3 * (s)] + (p)[l - 3 * (s)]));
#define ROW_DIJ(out, p, a, b, fac) \
  for (int l = 0; l < nx; l++) (out)[l] = (fac) * (270.0 * ((p)[l + (a) + (b)] - (p)[l - (a) + (b)] + (p)[l - (a) - (b)] - (p)[l + (a) - (b)]) \
                                                   - 27.0 * ((p)[l + 2 * (a) + 2 * (b)] - (p)[l - 2 * (a) + 2 * (b)] + (p)[l - 2 * (a) - 2 * (b)] - (p)[l + 2 * (a) - 2 * (b)]) \
                                                   + 2.0 * ((p)[l + 3 * (a) + 3 * (b)] - (p)[l - 3 * (a) + 3 * (b)] + (p)[l - 3 * (a) - 3 * (b)] - (p)[l + 3 * (a) - 3 * (b)]));

int pcref_pencil_applies(const pcb200_params *par)
{
  return par->lhydro && par->ldensity && (par->lentropy || par->gamma == 1.0) && par->lbidiagonal_derij &&
         par->nx > 1 && par->ny > 1 && par->nz > 1;
}

/* df = alpha*df + F(f) over the interior, pencil by pencil; returns max dt1 when ldt1 */
static double rhs_pencils(const double *f, double *df, const pcb200_params *par, double alpha, int ldt1)
{
  const int nx = par->nx, ny = par->ny, nz = par->nz;
  const int mx = nx + 2 * NG, my = ny + 2 * NG, mz = nz + 2 * NG;
  const long sy = mx, sz = (long)mx * my, mw = sz * mz;
  const int M = par->lmagnetic, E = par->lentropy;
  const int is = par->iss - 1;
  const double gamma = par->gamma, gamma1 = 1.0 / gamma, gamma_m1 = gamma - 1.0, cp = par->cp, cp1 = 1.0 / cp;
  const double cv = gamma1 * cp, cv1 = gamma * cp1;
  const double lnTT0 = gamma_m1 != 0.0 ? log(par->cs20 / (cp * gamma_m1)) : log(par->cs20 / cp);
  const double *zp_hc = pcref_get_zprofile(PCB200_PROF_HCOND), *zp_dl = pcref_get_zprofile(PCB200_PROF_DLNHCOND),
               *zp_cl = pcref_get_zprofile(PCB200_PROF_COOL);
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
    double *buf = (double *)malloc(sizeof(double) * (size_t)nx * 96);
    double *uij[3][3], *u2d[3][3], *umx[6], *gl[3], *aij[3][3], *a2d[3][3], *amx[6], *r1;
    double *gs[3], *d2s, *d2l, *u6[5][3], *tmp;        /* entropy: grad ss, del2 ss, del2 lnrho; upwind rows of ux,uy,uz,lnrho,ss */
    {
      double *q = buf;
      for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { uij[i][j] = q; q += nx; u2d[i][j] = q; q += nx; aij[i][j] = q; q += nx; a2d[i][j] = q; q += nx; }
      for (int k = 0; k < 6; k++) { umx[k] = q; q += nx; amx[k] = q; q += nx; }
      for (int j = 0; j < 3; j++) { gl[j] = q; q += nx; gs[j] = q; q += nx; }
      r1 = q; q += nx; d2s = q; q += nx; d2l = q; q += nx; tmp = q; q += nx;
      for (int k = 0; k < 5; k++) for (int j = 0; j < 3; j++) { u6[k][j] = q; q += nx; }
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
        }
        if (M || E) for (int l = 0; l < nx; l++) r1[l] = exp(-pl[l]);
        const double *ps = E ? f + (long)is * mw + row : 0;
        if (E) {
          for (int j = 0; j < 3; j++) { ROW_D1(gs[j], ps, st[j], f1[j]) }
          ROW_D2(d2s, ps, st[0], f2[0]) ROW_D2(tmp, ps, st[1], f2[1]) for (int l = 0; l < nx; l++) d2s[l] += tmp[l];
          ROW_D2(tmp, ps, st[2], f2[2]) for (int l = 0; l < nx; l++) d2s[l] += tmp[l];
          ROW_D2(d2l, pl, st[0], f2[0]) ROW_D2(tmp, pl, st[1], f2[1]) for (int l = 0; l < nx; l++) d2l[l] += tmp[l];
          ROW_D2(tmp, pl, st[2], f2[2]) for (int l = 0; l < nx; l++) d2l[l] += tmp[l];
        }
        if (par->lupw_uu) for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { ROW_D6U(u6[i][j], pu[i], st[j], f1[j]) }
        if (par->lupw_lnrho) for (int j = 0; j < 3; j++) { ROW_D6U(u6[3][j], pl, st[j], f1[j]) }
        if (E && par->lupw_ss) for (int j = 0; j < 3; j++) { ROW_D6U(u6[4][j], ps, st[j], f1[j]) }
        double *ds = E ? df + (long)is * mw + row : 0;
        const double hcond = (E && par->iheatcond == PCB200_HCOND_KPROFILE) ? zp_hc[n] : 0.0;
        const double glhc = (E && par->iheatcond == PCB200_HCOND_KPROFILE) ? zp_dl[n] : 0.0;
        const double coolp = (E && par->lcooling) ? zp_cl[n] : 0.0;
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
          double ugx = ux * uxx + uy * uxy + uz * uxz, ugy = ux * uyx + uy * uyy + uz * uyz, ugz = ux * uzx + uy * uzy + uz * uzz;
          double ugl = ux * glx + uy * gly + uz * glz;
          const double aux = fabs(ux), auy = fabs(uy), auz = fabs(uz);
          if (par->lupw_uu) {
            ugx -= aux * u6[0][0][l] + auy * u6[0][1][l] + auz * u6[0][2][l];
            ugy -= aux * u6[1][0][l] + auy * u6[1][1][l] + auz * u6[1][2][l];
            ugz -= aux * u6[2][0][l] + auy * u6[2][1][l] + auz * u6[2][2][l];
          }
          if (par->lupw_lnrho) ugl -= aux * u6[3][0][l] + auy * u6[3][1][l] + auz * u6[3][2][l];
          /* equation of state and pressure gradient (eos_idealgas.f90:843-868, entropy.f90:3209 / noentropy.f90:291) */
          double cs2l = cs2, TT1 = 0.0, ee = 1.0, gTx = 0.0, gTy = 0.0, gTz = 0.0, fpx, fpy, fpz;
          if (E) {
            const double th = cv1 * ps[l] + gamma_m1 * (pl[l] - par->lnrho0);
            cs2l = par->cs20 * exp(th);
            const double TT = exp(lnTT0 + th);
            ee = cv * TT; TT1 = 1.0 / TT;
            gTx = gamma_m1 * glx + cv1 * gs[0][l]; gTy = gamma_m1 * gly + cv1 * gs[1][l]; gTz = gamma_m1 * glz + cv1 * gs[2][l];
            const double pf = -cs2l * gamma1;
            fpx = pf * (glx + gTx); fpy = pf * (gly + gTy); fpz = pf * (glz + gTz);
          } else { fpx = -cs2 * glx; fpy = -cs2 * gly; fpz = -cs2 * glz; }
          double rx = -ugx + Om2 * uy + nu * (del2x + 2.0 * (sxx * glx + sxy * gly + sxz * glz) + third * gdx) + fpx;
          double ry = -ugy - Om2 * ux + nu * (del2y + 2.0 * (sxy * glx + syy * gly + syz * glz) + third * gdy) + fpy;
          double rz = -ugz + nu * (del2z + 2.0 * (sxz * glx + syz * gly + szz * glz) + third * gdz) + fpz;
          if (par->lgravz) rz += par->gravz;
          const double rl = -divu - ugl;
          double va2 = 0.0, j2 = 0.0;
          if (M) {
            const double bx = aij[2][1][l] - aij[1][2][l], by = aij[0][2][l] - aij[2][0][l], bz = aij[1][0][l] - aij[0][1][l];
            const double jx = mu01 * ((amx[0][l] + amx[1][l]) - (a2d[0][1][l] + a2d[0][2][l]));
            const double jy = mu01 * ((amx[2][l] + amx[3][l]) - (a2d[1][0][l] + a2d[1][2][l]));
            const double jz = mu01 * ((amx[4][l] + amx[5][l]) - (a2d[2][0][l] + a2d[2][1][l]));
            const double rho1 = r1[l];
            j2 = jx * jx + jy * jy + jz * jz;
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
          double diffus_chi = 0.0, hm = 0.0;
          if (E) {                                   /* denergy_dt, entropy.f90:3242-3452 */
            const double rho1 = r1[l];
            double ugs = ux * gs[0][l] + uy * gs[1][l] + uz * gs[2][l];
            if (par->lupw_ss) ugs -= aux * u6[4][0][l] + auy * u6[4][1][l] + auz * u6[4][2][l];
            double rs = -ugs, Hmax = 0.0;
            if (par->lviscosity_heat) {
              const double vh = 2.0 * nu * (sxx * sxx + syy * syy + szz * szz + 2.0 * (sxy * sxy + sxz * sxz + syz * syz));
              rs += TT1 * vh; Hmax += vh;
            }
            const double del2lnTT = gamma_m1 * d2l[l] + cv1 * d2s[l];
            if (par->iheatcond == PCB200_HCOND_KPROFILE) {
              rs += rho1 * hcond * (del2lnTT + (gTx * gTx + gTy * gTy + gTz * (gTz + glhc)));
              diffus_chi = gamma * (rho1 * hcond * cp1) * dxyz_2;
            } else if (par->iheatcond == PCB200_HCOND_CHICONST) {
              rs += par->chi * cp * (del2lnTT + ((glx + gTx) * gTx + (gly + gTy) * gTy + (glz + gTz) * gTz));
              diffus_chi = gamma * par->chi * dxyz_2;
            }
            if (par->lcooling) {
              const double heat = -par->cool * coolp * (cs2l - par->cs2cool) / par->cs2cool;
              rs += TT1 * rho1 * heat; Hmax += heat * rho1;
            }
            if (M && par->lohmic_heat) rs += eta * par->mu0 * j2 * rho1 * TT1;
            ds[l] = alpha * ds[l] + rs;
            hm = fabs(Hmax / ee / par->cdts);
          }
          if (ldt1) {
            const double maxadvec = (aux * dx_1 + auy * dy_1 + auz * dz_1) + sqrt(va2 + cs2l * dxyz_2);
            if (!E) { if (maxadvec > rowmax) rowmax = maxadvec; }
            else {      /* the diffusive and heating limits vary from point to point: equ.f90:1149 per point */
              double maxdiffus = fmax(nu * dxyz_2, diffus_chi);
              if (M) maxdiffus = fmax(maxdiffus, eta * dxyz_2);
              const double a = maxadvec / par->cdt, d = maxdiffus / par->cdtv;
              const double v = fmax(sqrt(a * a + d * d), hm);
              if (v > dt1) dt1 = v;
            }
          }
        }
        if (ldt1 && !E) {   /* equ.f90:1149, monotonic in maxadvec: applied once per row */
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
