/* pcref_particles.c -- CPU restatement of the dust-particle part of the hot path.
 *
 * TEST INFRASTRUCTURE ONLY (see pcref.c).  Parity: PINNED IN PART.  The particle
 * integration (initial state from the nr_f90 streams of 4 processes, boundconds,
 * nearest grid point, pencil sort, drag acceleration, drag time step, RK3 update)
 * reproduces every printed digit of samples/ParticlesCaustics/reference.out
 * (xp2m, yp2m, zp2m, vprms, dt over 8 steps; tests/test_oracle_particles_caustics.py)
 * with the two settings of that sample this file carries for the purpose
 * (pcref_particles_fixture: CIC interpolation, pure Stokes drag).  The TSC
 * assignment, the quadratic-spline interpolation and the back-reaction on the gas
 * are NOT pinned by a reference file (samples/convective-overstability-particles and
 * streaming_instability need shear, shock, self-gravity, sink particles); they are
 * checked by known-answer tests (drag relaxation, momentum conservation, TSC
 * partition of unity, index arithmetic) in tests/test_oracle_particles.py.
 *
 * Follows, routine by routine (all under /root/reference/src):
 *   boundconds_particles   particles_sub.f90:135-330      periodic wrap
 *   real_to_index/inverse_grid  grid.f90:2225-2343        index space, NINT, nearest()
 *   map_nearest_grid       particles_map.f90:895-984
 *   setup_mm_nn            general.f90:230-360            pencil order imn_array
 *   particle_pencil_index  particles_map.f90:1920-1948    npar_imn, k1_imn, k2_imn
 *   sort_particles_imn     particles_map.f90:986-1163     stable (counting sort, case 3)
 *   map_xxp_grid           particles_map.f90:1295-1742    np, TSC rhop, fold_f
 *   interpolate_quadratic_spline  particles_map.f90:581-726
 *   dxxp_dt_pencil, dvvp_dt_pencil  particles_dust.f90:3918-3977, 3979-4911
 *   dxxp_dt, dvvp_dt       particles_dust.f90:3120-3197, 3199-3250
 *   fold_df, fold_f        ghostfold_mpicomm.f90:28-153, 155-280 (nproc=1 branch)
 *   particles_timestep_first/second  particles_main.f90:486-551
 *
 * Layout: fp[6][mpar_loc] (= Fortran fp(mpar_loc,mparray)), dfp likewise,
 * ipar[mpar_loc], ineargrid[3][mpar_loc] holding the Fortran 1-based l,m,n.
 * Single rank only (the oracle is always run on the undecomposed domain).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "pcref.h"

#define NG 3
enum { IXP = 0, IYP, IZP, IVPX, IVPY, IVPZ, MPVAR };

static inline long cell(const pcb200_params *p, int l, int m, int n)     /* 0-based incl. ghosts */
{
  const long mx = p->nx + 2 * NG, my = p->ny + 2 * NG;
  return ((long)n * my + m) * mx + l;
}
static inline double gx(const pcb200_params *p, int axis, int i0)        /* grid coordinate, 0-based index */
{
  const double x0[3] = {p->x0_loc, p->y0_loc, p->z0_loc}, d[3] = {p->dx, p->dy, p->dz};
  return x0[axis] + (i0 - NG) * d[axis];
}

/* boundconds_particles, bcp='p' (particles_sub.f90:170-215) */
void pcref_particles_boundconds(double *fp, const pcb200_params *par, const pcb200_particle_params *pp)
{
  const long M = pp->mpar_loc;
  for (int ax = 0; ax < 3; ax++) {
    const double x0 = par->xyz0[ax], x1 = par->xyz0[ax] + par->Lxyz[ax], L = par->Lxyz[ax];
    double *q = fp + (long)ax * M;
    for (long k = 0; k < pp->npar_loc; k++) {
      if (q[k] < x0) q[k] = q[k] + L;
      if (q[k] >= x1) q[k] = q[k] - L;
    }
  }
}

/* inverse_grid(local=.true.) + NINT (grid.f90:2310-2343, particles_map.f90:935) */
static int nearest_index(const pcb200_params *par, int ax, double x)
{
  const int nloc[3] = {par->nx, par->ny, par->nz}, ip[3] = {par->ipx, par->ipy, par->ipz};
  const double h[3] = {par->dx, par->dy, par->dz};
  double xi = (x - par->xyz0[ax]) / h[ax];
  if (par->lperi[ax]) xi = xi + (double)NG + 0.5; else xi = xi + (double)(NG + 1);
  xi = xi - (double)(nloc[ax] * ip[ax]);
  const int inear = (int)round(xi);                      /* NINT: half away from zero */
  const int inear_max = NG + nloc[ax];
  const double xyz1 = par->xyz0[ax] + par->Lxyz[ax];
  if ((x < xyz1 && inear > inear_max) || (x < par->xyz0[ax] && inear > NG)) xi = nextafter(xi, -INFINITY);
  return (int)round(xi);
}

void pcref_particles_map_nearest_grid(const double *fp, int32_t *ineargrid, const pcb200_params *par,
                                      const pcb200_particle_params *pp)
{
  const long M = pp->mpar_loc;
  const int nloc[3] = {par->nx, par->ny, par->nz};
  for (int ax = 0; ax < 3; ax++)
    for (long k = 0; k < pp->npar_loc; k++) {
      int i = nearest_index(par, ax, fp[(long)ax * M + k]);
      if (pp->lcheck_exact_frontier) { if (i < NG + 1) i = NG + 1; if (i > NG + nloc[ax]) i = NG + nloc[ax]; }
      ineargrid[(long)ax * M + k] = i;
    }
}

/* setup_mm_nn (general.f90:230-335): imn_array[(n-n1)*ny + (m-m1)] = imn (1-based) */
void pcref_setup_imn_array(const pcb200_params *par, int32_t *imn_array)
{
  const int ny = par->ny, nz = par->nz;
  const int ncpus = par->nprocx * par->nprocy * par->nprocz;
  const int m1 = NG + 1, m2 = NG + ny, n1 = NG + 1, n2 = NG + nz;       /* Fortran 1-based */
  const int m1i = m1 + NG - 1, m2i = m2 - NG + 1, n1i = n1 + NG - 1, n2i = n2 - NG + 1;
#define IMN(m, n) imn_array[(long)((n) - n1) * ny + ((m) - m1)]
  for (long i = 0; i < (long)ny * nz; i++) imn_array[i] = 0;
  int imn = 1;
  if (ncpus == 1) {
    for (int n = n1; n <= n2; n++) for (int m = m1; m <= m2; m++) IMN(m, n) = imn++;
  } else {
#define VISIT(m, n) do { if (IMN(m, n) == 0) IMN(m, n) = imn++; } while (0)
    for (int n = n1i + 2; n <= n2i - 2; n++) for (int m = m1i + 2; m <= m2i - 2; m++) VISIT(m, n);
    { const int a = n2i - 1 > n1 + 1 ? n2i - 1 : n1 + 1;
      for (int n = a; n <= n2; n++) for (int m = m1i + 2; m <= m2i - 2; m++) VISIT(m, n); }
    { const int b = n1i + 1 < n2 ? n1i + 1 : n2;
      for (int n = n1; n <= b; n++) for (int m = m1i + 2; m <= m2i - 2; m++) VISIT(m, n); }
    const int min_m1i_m2 = m1i + 1 < m2 ? m1i + 1 : m2;
    const int max_m2i_m1 = m2i - 1 > min_m1i_m2 + 1 ? m2i - 1 : min_m1i_m2 + 1;
    for (int n = n1; n <= n2; n++) {
      for (int m = m1; m <= min_m1i_m2; m++) VISIT(m, n);
      for (int m = max_m2i_m1; m <= m2; m++) VISIT(m, n);
    }
#undef VISIT
  }
#undef IMN
}

/* particle_pencil_index (particles_map.f90:1920-1948); arrays of ny*nz, 1-based particle numbers */
static void pencil_index(const int32_t *ineargrid, const int32_t *imn_array, const pcb200_params *par,
                         const pcb200_particle_params *pp, int32_t *npar_imn, int32_t *k1_imn, int32_t *k2_imn)
{
  const long M = pp->mpar_loc;
  const int ny = par->ny, nimn = par->ny * par->nz;
  for (int i = 0; i < nimn; i++) npar_imn[i] = 0;
  for (long k = 0; k < pp->npar_loc; k++) {
    const int iy0 = ineargrid[M + k], iz0 = ineargrid[2 * M + k];
    npar_imn[imn_array[(long)(iz0 - NG - 1) * ny + (iy0 - NG - 1)] - 1]++;
  }
  int k = 0;
  for (int i = 0; i < nimn; i++) {
    if (npar_imn[i] != 0) { k1_imn[i] = k + 1; k2_imn[i] = k1_imn[i] + npar_imn[i] - 1; k += npar_imn[i]; }
    else { k1_imn[i] = 0; k2_imn[i] = 0; }
  }
}

/* sort_particles_imn, counting sort (particles_map.f90:1109-1131): stable in the current order */
void pcref_particles_sort(double *fp, double *dfp, int32_t *ipar, int32_t *ineargrid, int32_t *k1_imn, int32_t *k2_imn,
                          const pcb200_params *par, const pcb200_particle_params *pp)
{
  const long M = pp->mpar_loc, N = pp->npar_loc;
  const int ny = par->ny, nimn = par->ny * par->nz;
  int32_t *imn_array = (int32_t *)malloc(sizeof(int32_t) * nimn), *npar_imn = (int32_t *)malloc(sizeof(int32_t) * nimn);
  int32_t *kk = (int32_t *)malloc(sizeof(int32_t) * nimn);
  long *sorted = (long *)malloc(sizeof(long) * (N > 0 ? N : 1));
  pcref_setup_imn_array(par, imn_array);
  pencil_index(ineargrid, imn_array, par, pp, npar_imn, k1_imn, k2_imn);
  for (int i = 0; i < nimn; i++) kk[i] = k1_imn[i];
  for (long k = 0; k < N; k++) {
    const int iy0 = ineargrid[M + k], iz0 = ineargrid[2 * M + k];
    const int imn = imn_array[(long)(iz0 - NG - 1) * ny + (iy0 - NG - 1)] - 1;
    sorted[kk[imn] - 1] = k;
    kk[imn]++;
  }
  double *tmp = (double *)malloc(sizeof(double) * (N > 0 ? N : 1));
  int32_t *itmp = (int32_t *)malloc(sizeof(int32_t) * (N > 0 ? N : 1));
  for (int j = 0; j < MPVAR; j++) {
    for (long k = 0; k < N; k++) tmp[k] = fp[j * M + sorted[k]];
    memcpy(fp + j * M, tmp, sizeof(double) * N);
    if (dfp) { for (long k = 0; k < N; k++) tmp[k] = dfp[j * M + sorted[k]]; memcpy(dfp + j * M, tmp, sizeof(double) * N); }
  }
  for (int j = 0; j < 3; j++) { for (long k = 0; k < N; k++) itmp[k] = ineargrid[j * M + sorted[k]]; memcpy(ineargrid + j * M, itmp, sizeof(int32_t) * N); }
  for (long k = 0; k < N; k++) itmp[k] = ipar[sorted[k]];
  memcpy(ipar, itmp, sizeof(int32_t) * N);
  free(tmp); free(itmp); free(sorted); free(kk); free(npar_imn); free(imn_array);
}

/* TSC assignment weight of grid point ixx (0-based) for a particle at xp with nearest point ix0
 * (particles_dust.f90:4467-4490, particles_map.f90:1607-1630) */
static inline double tsc_weight(const pcb200_params *par, int ax, double xp, int ixx, int ix0)
{
  const double d_1[3] = {par->dx_1, par->dy_1, par->dz_1};
  if (ixx - ix0 == -1 || ixx - ix0 == 1) {
    const double a = fabs(xp - gx(par, ax, ixx)) * d_1[ax];
    return 1.125 - 1.5 * a + 0.5 * (a * a);
  }
  const double b = (xp - gx(par, ax, ixx)) * d_1[ax];
  return 0.75 - b * b;
}

/* fold the first ghost layer into the opposite interior edge, z then y then x
 * (fold_df/fold_f, nproc=1 branches, ghostfold_mpicomm.f90:46-52,86-92,120-124) */
static void fold_field(double *a, const pcb200_params *p)
{
  const int nx = p->nx, ny = p->ny, nz = p->nz;
  const int l1 = NG, l2 = NG + nx - 1, m1 = NG, m2 = NG + ny - 1, n1 = NG, n2 = NG + nz - 1;    /* 0-based */
  for (int m = m1 - 1; m <= m2 + 1; m++) for (int l = l1 - 1; l <= l2 + 1; l++) {
    a[cell(p, l, m, n1)] = a[cell(p, l, m, n1)] + a[cell(p, l, m, n2 + 1)];
    a[cell(p, l, m, n2)] = a[cell(p, l, m, n2)] + a[cell(p, l, m, n1 - 1)];
    a[cell(p, l, m, n1 - 1)] = 0.0; a[cell(p, l, m, n2 + 1)] = 0.0;
  }
  for (int n = n1; n <= n2; n++) for (int l = l1 - 1; l <= l2 + 1; l++) {
    a[cell(p, l, m1, n)] = a[cell(p, l, m1, n)] + a[cell(p, l, m2 + 1, n)];
    a[cell(p, l, m2, n)] = a[cell(p, l, m2, n)] + a[cell(p, l, m1 - 1, n)];
    a[cell(p, l, m1 - 1, n)] = 0.0; a[cell(p, l, m2 + 1, n)] = 0.0;
  }
  for (int n = n1; n <= n2; n++) for (int m = m1; m <= m2; m++) {
    a[cell(p, l1, m, n)] = a[cell(p, l1, m, n)] + a[cell(p, l2 + 1, m, n)];
    a[cell(p, l2, m, n)] = a[cell(p, l2, m, n)] + a[cell(p, l1 - 1, m, n)];
    a[cell(p, l1 - 1, m, n)] = 0.0; a[cell(p, l2 + 1, m, n)] = 0.0;
  }
}

/* map_xxp_grid (particles_map.f90:1295-1742): np (count at the nearest point) and TSC rhop,
 * folded, times rhop_swarm.  np, rhop: full (mx,my,mz) arrays. */
void pcref_particles_map_xxp_grid(const double *fp, const int32_t *ineargrid, double *np, double *rhop,
                                  const pcb200_params *par, const pcb200_particle_params *pp)
{
  const long M = pp->mpar_loc;
  const long mw = (long)(par->nx + 2 * NG) * (par->ny + 2 * NG) * (par->nz + 2 * NG);
  memset(np, 0, sizeof(double) * mw);
  memset(rhop, 0, sizeof(double) * mw);
  for (long k = 0; k < pp->npar_loc; k++) {
    const int ix0 = ineargrid[k] - 1, iy0 = ineargrid[M + k] - 1, iz0 = ineargrid[2 * M + k] - 1;
    np[cell(par, ix0, iy0, iz0)] += 1.0;
  }
  for (long k = 0; k < pp->npar_loc; k++) {
    const int ix0 = ineargrid[k] - 1, iy0 = ineargrid[M + k] - 1, iz0 = ineargrid[2 * M + k] - 1;
    for (int izz = iz0 - 1; izz <= iz0 + 1; izz++) for (int iyy = iy0 - 1; iyy <= iy0 + 1; iyy++)
      for (int ixx = ix0 - 1; ixx <= ix0 + 1; ixx++) {
        const double wx = tsc_weight(par, 0, fp[k], ixx, ix0);
        const double wy = tsc_weight(par, 1, fp[M + k], iyy, iy0);
        const double wz = tsc_weight(par, 2, fp[2 * M + k], izz, iz0);
        double w = 1.0;
        w = w * wx; w = w * wy; w = w * wz;
        rhop[cell(par, ixx, iyy, izz)] += w;
      }
  }
  fold_field(rhop, par);
  for (int n = NG; n < NG + par->nz; n++) for (int m = NG; m < NG + par->ny; m++) for (int l = NG; l < NG + par->nx; l++)
    rhop[cell(par, l, m, n)] = pp->rhop_swarm * rhop[cell(par, l, m, n)];
}

/* interpolate_quadratic_spline, dimensionality 3 (particles_map.f90:686-722) */
static void interp_qspline(const double *f, const pcb200_params *par, int ivar1, const double *xxp, const int *inear, double *gp)
{
  const long mw = (long)(par->nx + 2 * NG) * (par->ny + 2 * NG) * (par->nz + 2 * NG);
  const int ix0 = inear[0], iy0 = inear[1], iz0 = inear[2];          /* 0-based */
  const double dxp0 = (xxp[0] - gx(par, 0, ix0)) * par->dx_1;
  const double dyp0 = (xxp[1] - gx(par, 1, iy0)) * par->dy_1;
  const double dzp0 = (xxp[2] - gx(par, 2, iz0)) * par->dz_1;
  const double fxm = 0.5 * ((0.5 - dxp0) * (0.5 - dxp0)), fx0 = 0.75 - dxp0 * dxp0, fxp = 0.5 * ((0.5 + dxp0) * (0.5 + dxp0));
  const double fym = 0.5 * ((0.5 - dyp0) * (0.5 - dyp0)), fy0 = 0.75 - dyp0 * dyp0, fyp = 0.5 * ((0.5 + dyp0) * (0.5 + dyp0));
  const double fzm = 0.5 * ((0.5 - dzp0) * (0.5 - dzp0)), fz0 = 0.75 - dzp0 * dzp0, fzp = 0.5 * ((0.5 + dzp0) * (0.5 + dzp0));
  for (int j = 0; j < 3; j++) {
    const double *g = f + (long)(ivar1 + j) * mw;
#define F(a, b, c) g[cell(par, ix0 + (a), iy0 + (b), iz0 + (c))]
    gp[j] = fx0 * fy0 * fz0 * F(0, 0, 0) +
            fx0 * fy0 * (F(0, 0, 1) * fzp + F(0, 0, -1) * fzm) +
            fx0 * fz0 * (F(0, 1, 0) * fyp + F(0, -1, 0) * fym) +
            fy0 * fz0 * (F(1, 0, 0) * fxp + F(-1, 0, 0) * fxm) +
            fxp * fyp * (F(1, 1, 1) * fzp + F(1, 1, -1) * fzm) +
            fxp * fym * (F(1, -1, 1) * fzp + F(1, -1, -1) * fzm) +
            fxm * fyp * (F(-1, 1, 1) * fzp + F(-1, 1, -1) * fzm) +
            fxm * fym * (F(-1, -1, 1) * fzp + F(-1, -1, -1) * fzm) +
            fx0 * fyp * (F(0, 1, 1) * fzp + F(0, 1, -1) * fzm) +
            fx0 * fym * (F(0, -1, 1) * fzp + F(0, -1, -1) * fzm) +
            fy0 * fzp * (F(1, 0, 1) * fxp + F(-1, 0, 1) * fxm) +
            fy0 * fzm * (F(1, 0, -1) * fxp + F(-1, 0, -1) * fxm) +
            fz0 * fxp * (F(1, 1, 0) * fyp + F(1, -1, 0) * fym) +
            fz0 * fxm * (F(-1, 1, 0) * fyp + F(-1, -1, 0) * fym);
#undef F
  }
}

/* ---- options of the reference fixture samples/ParticlesCaustics (checker only; the shipped library has the
 * quadratic spline and a constant friction time): interp_pol_uu='cic' and ldraglaw_purestokes with the radii of
 * particles_radius.  Set once by the test that replays the fixture, off by default. ---- */
static int g_fix_cic = 0;
static const double *g_fix_ap = 0;         /* radius by particle identity ipar (1-based) */
static long g_fix_nap = 0;
static double g_fix_rhopmat = 0.0;
static const int32_t *g_fix_ipar = 0;      /* ipar of the state pcref_particles_pde is called on */
void pcref_particles_fixture(int interp_cic, const double *ap_by_ipar, long nap, double rhopmat, const int32_t *ipar)
{
  g_fix_cic = interp_cic; g_fix_ap = ap_by_ipar; g_fix_nap = ap_by_ipar ? nap : 0; g_fix_rhopmat = rhopmat; g_fix_ipar = ipar;
}

/* interpolate_linear_range (particles_map.f90:112-287), equidistant grid, all three directions present */
static void interp_linear(const double *f, const pcb200_params *par, int ivar1, int nvar, const double *xxp, const int *inear, double *gp)
{
  const long mw = (long)(par->nx + 2 * NG) * (par->ny + 2 * NG) * (par->nz + 2 * NG);
  int ix0 = inear[0], iy0 = inear[1], iz0 = inear[2];
  if (gx(par, 0, ix0) > xxp[0]) ix0 = ix0 - 1;
  if (gx(par, 1, iy0) > xxp[1]) iy0 = iy0 - 1;
  if (gx(par, 2, iz0) > xxp[2]) iz0 = iz0 - 1;
  const double xp0 = xxp[0] - gx(par, 0, ix0), yp0 = xxp[1] - gx(par, 1, iy0), zp0 = xxp[2] - gx(par, 2, iz0);
  const double dx1 = par->dx_1, dy1 = par->dy_1, dz1 = par->dz_1;
  const double dxdy1 = dx1 * dy1, dxdz1 = dx1 * dz1, dydz1 = dy1 * dz1, dxdydz1 = dx1 * dy1 * dz1;
  for (int j = 0; j < nvar; j++) {
    const double *g = f + (long)(ivar1 + j) * mw;
    const double g1 = g[cell(par, ix0, iy0, iz0)], g2 = g[cell(par, ix0 + 1, iy0, iz0)];
    const double g3 = g[cell(par, ix0, iy0 + 1, iz0)], g4 = g[cell(par, ix0 + 1, iy0 + 1, iz0)];
    const double g5 = g[cell(par, ix0, iy0, iz0 + 1)], g6 = g[cell(par, ix0 + 1, iy0, iz0 + 1)];
    const double g7 = g[cell(par, ix0, iy0 + 1, iz0 + 1)], g8 = g[cell(par, ix0 + 1, iy0 + 1, iz0 + 1)];
    gp[j] = g1 + xp0 * dx1 * (-g1 + g2) + yp0 * dy1 * (-g1 + g3) + zp0 * dz1 * (-g1 + g5) +
            xp0 * yp0 * dxdy1 * (g1 - g2 - g3 + g4) + xp0 * zp0 * dxdz1 * (g1 - g2 - g5 + g6) +
            yp0 * zp0 * dydz1 * (g1 - g3 - g5 + g7) +
            xp0 * yp0 * zp0 * dxdydz1 * (-g1 + g2 + g3 - g4 + g5 - g6 - g7 + g8);
  }
}

/* dxxp_dt_pencil + dvvp_dt_pencil over all pencils (particles_dust.f90:3918-3977, 3979-4911),
 * then dxxp_dt, dvvp_dt (:3120-3197, :3199-3250) and fold_df(df,iux,iuz) (equ.f90:407-413).
 * f: state with valid ghost zones; df: gas RHS register (mx,my,mz,mvar), receives the
 * back-reaction; dfp updated; *dt1_max updated when lfirst&&ldt. */
void pcref_particles_pde(const double *f, double *df, const double *fp, double *dfp, const int32_t *ineargrid,
                         const pcb200_params *par, const pcb200_particle_params *pp, int lfirst, double *dt1_max)
{
  const long M = pp->mpar_loc, N = pp->npar_loc;
  const long mw = (long)(par->nx + 2 * NG) * (par->ny + 2 * NG) * (par->nz + 2 * NG);
  double tausp1 = pp->tausp != 0.0 ? 1.0 / pp->tausp : 0.0;
  const double *lnrho = f + (long)(par->ilnrho - 1) * mw;
  const int iux = par->iux - 1;
  const int ldt = lfirst && par->ldt;
  double *dd = 0, *dg = 0;
  if (ldt) { dd = (double *)calloc(mw, sizeof(double)); dg = (double *)calloc(mw, sizeof(double)); }
  for (long k = 0; k < N; k++) {
    const int inear[3] = {ineargrid[k] - 1, ineargrid[M + k] - 1, ineargrid[2 * M + k] - 1};
    const int ix0 = inear[0], iy0 = inear[1], iz0 = inear[2];
    const double xxp[3] = {fp[IXP * M + k], fp[IYP * M + k], fp[IZP * M + k]};
    const double vvp[3] = {fp[IVPX * M + k], fp[IVPY * M + k], fp[IVPZ * M + k]};
    /* dxxp_dt_pencil :3936-3968 */
    if (ldt && pp->ldt_adv_par) {
      const double ax = fabs(vvp[0]) * par->dx_1, ay = fabs(vvp[1]) * par->dy_1, az = fabs(vvp[2]) * par->dz_1;
      const double v = sqrt(ax * ax + ay * ay + az * az) / pp->cdtp;
      if (v > *dt1_max) *dt1_max = v;
    }
    if (pp->ldragforce_dust_par) {
      double uup[3];
      if (g_fix_cic) interp_linear(f, par, iux, 3, xxp, inear, uup);
      else interp_qspline(f, par, iux, xxp, inear, uup);                  /* :4226 */
      if (g_fix_nap) {                                                    /* calc_draglaw_purestokes :6314-6362 */
        double lr;
        interp_linear(f, par, par->ilnrho - 1, 1, xxp, inear, &lr);       /* interpolate_quantities, particles_map.f90:2151-2158 */
        const double interp_rho = exp(lr), dia = 2.0 * g_fix_ap[g_fix_ipar[k] - 1];
        tausp1 = 18.0 * par->nu / ((g_fix_rhopmat / interp_rho) * (dia * dia));
      }
      double dragforce[3];
      for (int j = 0; j < 3; j++) dragforce[j] = -tausp1 * (vvp[j] - uup[j]);      /* :4264 */
      for (int j = 0; j < 3; j++) dfp[(IVPX + j) * M + k] = dfp[(IVPX + j) * M + k] + dragforce[j];   /* :4275 */
      if (pp->ldragforce_gas_par) {                                       /* TSC back-reaction :4443-4536 */
        for (int izz = iz0 - 1; izz <= iz0 + 1; izz++) for (int iyy = iy0 - 1; iyy <= iy0 + 1; iyy++)
          for (int ixx = ix0 - 1; ixx <= ix0 + 1; ixx++) {
            const double wx = tsc_weight(par, 0, xxp[0], ixx, ix0);
            const double wy = tsc_weight(par, 1, xxp[1], iyy, iy0);
            const double wz = tsc_weight(par, 2, xxp[2], izz, iz0);
            double weight = 1.0;
            weight = weight * wx; weight = weight * wy; weight = weight * wz;
            const long c = cell(par, ixx, iyy, izz);
            const double rho1_point = 1.0 / exp(lnrho[c]);                /* get_gas_density */
            const double mp_vcell = pp->rhop_swarm;                       /* get_rhopswarm, equidistant Cartesian */
            for (int j = 0; j < 3; j++)
              df[(long)(iux + j) * mw + c] = df[(long)(iux + j) * mw + c] - mp_vcell * rho1_point * dragforce[j] * weight;
          }
      }
      if (ldt) {                                                          /* :4843-4847 */
        const long c0 = cell(par, ix0, iy0, iz0);
        if (tausp1 > dd[c0]) dd[c0] = tausp1;
        if (pp->ldragforce_gas_par) dg[c0] = dg[c0] + pp->rhop_swarm * exp(-lnrho[c0]) * tausp1;
      }
    }
  }
  if (ldt && pp->ldragforce_dust_par) {                                   /* :4901-4910 */
    for (long c = 0; c < mw; c++) {
      double d = pp->ldragforce_gas_par ? dd[c] + dg[c] : dd[c];
      d = d / pp->cdtp_drag;
      if (d > *dt1_max) *dt1_max = d;
    }
  }
  if (dd) { free(dd); free(dg); }
  /* dxxp_dt :3155-3161 */
  for (int j = 0; j < 3; j++) for (long k = 0; k < N; k++) dfp[(IXP + j) * M + k] = dfp[(IXP + j) * M + k] + fp[(IVPX + j) * M + k];
  /* dvvp_dt Coriolis :3226-3233 */
  if (par->Omega != 0.0 && pp->lcoriolis_force_par) {
    const double Omega2 = 2 * par->Omega;
    for (long k = 0; k < N; k++) {
      dfp[IVPX * M + k] = dfp[IVPX * M + k] + Omega2 * fp[IVPY * M + k];
      dfp[IVPY * M + k] = dfp[IVPY * M + k] - Omega2 * fp[IVPX * M + k];
    }
  }
  /* fold_df(df,iux,iuz), equ.f90:407-413 (lfold_df set in particles_dust.f90:734) */
  if (pp->ldragforce_dust_par && pp->ldragforce_gas_par)
    for (int j = 0; j < 3; j++) fold_field(df + (long)(iux + j) * mw, par);
}

/* time_step with particles (timestep.f90:113-192): per substep
 * particles_timestep_first -> pde {particles_boundconds, BCs, gas RHS, particle pencils,
 * fold_df} -> set_dt -> gas update -> particles_timestep_second.
 * np, rhop: (mx,my,mz) outputs of map_xxp_grid of the LAST substep; k1_imn,k2_imn: ny*nz. */
void pcref_timestep_particles(double *f, double *df, const pcb200_params *par, const pcb200_forcing *frc,
                              double *dt, double *diag, const pcb200_particle_params *pp,
                              double *fp, double *dfp, int32_t *ipar, int32_t *ineargrid,
                              double *np, double *rhop, int32_t *k1_imn, int32_t *k2_imn)
{
  const int mx = par->nx + 2 * NG, my = par->ny + 2 * NG, mz = par->nz + 2 * NG;
  const long mw = (long)mx * my * mz, ntot = mw * par->mvar;
  const long M = pp->mpar_loc, N = pp->npar_loc;
  for (int itsub = 0; itsub < 3; itsub++) {
    const int lfirst = itsub == 0;
    if (lfirst) memset(df, 0, sizeof(double) * (size_t)ntot);
    else { const double a = par->alpha_ts[itsub]; for (long i = 0; i < ntot; i++) df[i] = a * df[i]; }
    /* particles_timestep_first, particles_main.f90:494-498 */
    for (int j = 0; j < MPVAR; j++) for (long k = 0; k < N; k++)
      dfp[j * M + k] = lfirst ? 0.0 : par->alpha_ts[itsub] * dfp[j * M + k];
    /* particles_boundconds, particles_main.f90:625-664 (equ.f90:169) */
    pcref_particles_boundconds(fp, par, pp);
    pcref_particles_map_nearest_grid(fp, ineargrid, par, pp);
    pcref_particles_sort(fp, dfp, ipar, ineargrid, k1_imn, k2_imn, par, pp);
    pcref_particles_map_xxp_grid(fp, ineargrid, np, rhop, par, pp);
    pcref_boundconds(f, par);
    double dt1_max = 0.0;
    if (lfirst && par->ldt) dt1_max = par->dtmax != 0.0 ? 1. / par->dtmax : 0.0;
    pcref_rhs(f, df, par, lfirst, &dt1_max, lfirst ? diag : 0);
    pcref_particles_pde(f, df, fp, dfp, ineargrid, par, pp, lfirst, &dt1_max);
    if (lfirst && par->ldt) *dt = 1.0 / dt1_max;
    if (lfirst && !par->ldt) *dt = par->dt_fixed;
    const double dt_beta = *dt * par->beta_ts[itsub];
    for (int k = 0; k < par->mvar; k++) for (int n = NG; n < NG + par->nz; n++)
      for (int m = NG; m < NG + par->ny; m++) for (int l = NG; l < NG + par->nx; l++) {
        const long i = (long)k * mw + ((long)n * my + m) * mx + l;
        f[i] = f[i] + dt_beta * df[i];
      }
    /* particles_timestep_second, particles_main.f90:521-522 */
    for (int j = 0; j < MPVAR; j++) for (long k = 0; k < N; k++) fp[j * M + k] = fp[j * M + k] + dt_beta * dfp[j * M + k];
  }
  if (frc) pcref_addforce(f, par, frc);
}
