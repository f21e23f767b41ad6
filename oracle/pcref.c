/* pcref.c -- CPU ORACLE (test infrastructure, NOT product code; see pcref.h).
 *
 * Restates, in the Fortran's own order of floating-point operations, the
 * reference hot path: deriv.f90 stencils, sub.f90 pencil algebra, the module
 * RHS terms, boundary conditions, the CFL time step and the 2N-RK3 update.
 * Build: gcc -O2 -ffp-contract=off -fopenmp (see oracle/Makefile); FMA
 * contraction stays off because the reference build (gfortran -O, no -march,
 * config/compilers/separate/GNU-gfortran.conf:6-8) emits none on x86-64.
 *
 * Index convention: all (l,m,n) are 0-based and include the 3 ghost cells,
 * f[((k*mz+n)*my+m)*mx+l] == Fortran f(l+1,m+1,n+1,k+1).
 */
#include "pcref.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define NG PCB200_NGHOST

static const double *g_zprof[PCB200_NPROF] = {0, 0, 0};
static int g_zprof_n[PCB200_NPROF] = {0, 0, 0};
static double *g_zprof_own[PCB200_NPROF] = {0, 0, 0};

/* z profiles (hcond_prof, dlnhcond_prof, cooling prof), nz values */
void pcref_set_zprofile(int which, const double *prof, int nz)
{
  if (which < 0 || which >= PCB200_NPROF) return;
  free(g_zprof_own[which]);
  g_zprof_own[which] = (double *)malloc(sizeof(double) * (size_t)nz);
  memcpy(g_zprof_own[which], prof, sizeof(double) * (size_t)nz);
  g_zprof[which] = g_zprof_own[which];
  g_zprof_n[which] = nz;
}

/* bench.py --impl reference under torchrun inherits OMP_NUM_THREADS=1: the CPU arm sets the count itself */
/* the profile set by pcref_set_zprofile (pcref_pencil.c reads the same arrays) */
const double *pcref_get_zprofile(int which)
{
  return (which >= 0 && which < PCB200_NPROF) ? g_zprof[which] : 0;
}

void pcref_set_num_threads(int n)
{
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

int pcref_num_threads(void)
{
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* stride of axis j (1=x,2=y,3=z) in a (mx,my,mz) block */
static inline long axis_stride(int mx, int my, int j)
{
  return j == 1 ? 1 : (j == 2 ? (long)mx : (long)mx * my);
}

/* der_main, deriv.f90:89-171: fac*(45*(f+1 - f-1) - 9*(f+2 - f-2) + (f+3 - f-3)),
 * fac = (1/60)*dx_1 */
double pcref_der(const double *f, int mx, int my, int l, int m, int n, int j, double d_1)
{
  const long s = axis_stride(mx, my, j);
  const double *p = f + ((long)n * my + m) * mx + l;
  const double a = 1.0 / 60.0;
  const double fac = a * d_1;
  return fac * (+45.0 * (p[s] - p[-s]) - 9.0 * (p[2 * s] - p[-2 * s]) + (p[3 * s] - p[-3 * s]));
}

/* der2_main, deriv.f90:474-564, coefficients deriv.f90:37-40 ('standard') */
double pcref_der2(const double *f, int mx, int my, int l, int m, int n, int j, double d_1)
{
  const long s = axis_stride(mx, my, j);
  const double *p = f + ((long)n * my + m) * mx + l;
  const double c0 = -490. / 180., c1 = 270. / 180., c2 = -27. / 180., c3 = 2. / 180.;
  const double fac = d_1 * d_1;
  return fac * (c0 * p[0] + c1 * (p[s] + p[-s]) + c2 * (p[2 * s] + p[-2 * s]) + c3 * (p[3 * s] + p[-3 * s]));
}

/* der6_main with upwind=T, deriv.f90:938-1048: fac=(1/60)*dx_1 */
double pcref_der6_upwind(const double *f, int mx, int my, int l, int m, int n, int j, double d_1)
{
  const long s = axis_stride(mx, my, j);
  const double *p = f + ((long)n * my + m) * mx + l;
  const double fac = (1.0 / 60) * d_1;
  return fac * (-20.0 * p[0] + 15.0 * (p[s] + p[-s]) - 6.0 * (p[2 * s] + p[-2 * s]) + (p[3 * s] + p[-3 * s]));
}

/* derij_main, bidiagonal scheme (lbidiagonal_derij=T), deriv.f90:1373-1423.
 * The Fortran writes fac=(1./720.)*dxi_1*dxj_1 with (i,j) ordered as
 * (x,y), (y,z), (z,x); the first listed axis carries the inner sign flip. */
double pcref_derij(const double *f, int mx, int my, int l, int m, int n, int i, int j,
                   double di_1, double dj_1)
{
  const double *p = f + ((long)n * my + m) * mx + l;
  long sa, sb;     /* sa: axis of the inner +/- difference, sb: outer */
  double fac;
  if (i + j == 3) {          /* x,y: fac=(1/720)*dx_1*dy_1 ; inner x, outer y */
    const double dx_1 = (i == 1) ? di_1 : dj_1, dy_1 = (i == 1) ? dj_1 : di_1;
    fac = (1. / 720.) * dx_1 * dy_1; sa = 1; sb = mx;
  } else if (i + j == 5) {   /* y,z: facs=(1/720)*dy_1*dz_1 ; terms f(m+k,n+k)-f(m+k,n-k)+f(m-k,n-k)-f(m-k,n+k) */
    const double dy_1 = (i == 2) ? di_1 : dj_1, dz_1 = (i == 2) ? dj_1 : di_1;
    fac = (1. / 720.) * dy_1 * dz_1;
    const long sy = mx, sz = (long)mx * my;
    return fac * (270. * (p[sy + sz] - p[sy - sz] + p[-sy - sz] - p[-sy + sz])
                  - 27. * (p[2 * sy + 2 * sz] - p[2 * sy - 2 * sz] + p[-2 * sy - 2 * sz] - p[-2 * sy + 2 * sz])
                  + 2. * (p[3 * sy + 3 * sz] - p[3 * sy - 3 * sz] + p[-3 * sy - 3 * sz] - p[-3 * sy + 3 * sz]));
  } else {                   /* z,x: fac=(1/720)*dz_1*dx_1 ; inner x, outer z */
    const double dx_1 = (i == 1) ? di_1 : dj_1, dz_1 = (i == 1) ? dj_1 : di_1;
    fac = (1. / 720.) * dz_1 * dx_1; sa = 1; sb = (long)mx * my;
  }
  return fac * (270. * (p[sa + sb] - p[-sa + sb] + p[-sa - sb] - p[sa - sb])
                - 27. * (p[2 * sa + 2 * sb] - p[-2 * sa + 2 * sb] + p[-2 * sa - 2 * sb] - p[2 * sa - 2 * sb])
                + 2. * (p[3 * sa + 3 * sb] - p[-3 * sa + 3 * sb] + p[-3 * sa - 3 * sb] - p[3 * sa - 3 * sb]));
}

/* derij_main, lbidiagonal_derij=F: "the old, straight-forward scheme", product of two
 * 1-D 6th-order first-derivative stencils, 36 points (deriv.f90:1430-1557).  Inner
 * differences along a (x for (x,y) and (z,x); y for (y,z)), outer along b. */
double pcref_derij_full(const double *f, int mx, int my, int l, int m, int n, int i, int j,
                        double di_1, double dj_1)
{
  const double *p = f + ((long)n * my + m) * mx + l;
  const long sx = 1, sy = mx, sz = (long)mx * my;
  long sa, sb;
  double fac;
  if (i + j == 3) { const double dx_1 = (i == 1) ? di_1 : dj_1, dy_1 = (i == 1) ? dj_1 : di_1;
                    fac = (1. / (60. * 60.)) * dx_1 * dy_1; sa = sx; sb = sy; }
  else if (i + j == 5) { const double dy_1 = (i == 2) ? di_1 : dj_1, dz_1 = (i == 2) ? dj_1 : di_1;
                    fac = (1. / (60. * 60.)) * dy_1 * dz_1; sa = sy; sb = sz; }
  else { const double dx_1 = (i == 1) ? di_1 : dj_1, dz_1 = (i == 1) ? dj_1 : di_1;
                    fac = (1. / (60. * 60.)) * dz_1 * dx_1; sa = sx; sb = sz; }
#define DROW(q) (45. * ((q)[sa] - (q)[-sa]) - 9. * ((q)[2 * sa] - (q)[-2 * sa]) + ((q)[3 * sa] - (q)[-3 * sa]))
  return fac * (45. * (DROW(p + sb) - DROW(p - sb)) - 9. * (DROW(p + 2 * sb) - DROW(p - 2 * sb))
                + (DROW(p + 3 * sb) - DROW(p - 3 * sb)));
#undef DROW
}

void pcref_derij_full_field(const double *f, int nx, int ny, int nz, int i, int j,
                            double di_1, double dj_1, double *out)
{
  const int mx = nx + 2 * NG, my = ny + 2 * NG;
  for (int n = 0; n < nz; n++) for (int m = 0; m < ny; m++) for (int l = 0; l < nx; l++)
    out[((long)n * ny + m) * nx + l] = pcref_derij_full(f, mx, my, l + NG, m + NG, n + NG, i, j, di_1, dj_1);
}

void pcref_der_field(const double *f, int nx, int ny, int nz, int j, double d_1, double *out)
{
  const int mx = nx + 2 * NG, my = ny + 2 * NG;
  for (int n = 0; n < nz; n++) for (int m = 0; m < ny; m++) for (int l = 0; l < nx; l++)
    out[((long)n * ny + m) * nx + l] = pcref_der(f, mx, my, l + NG, m + NG, n + NG, j, d_1);
}
void pcref_der2_field(const double *f, int nx, int ny, int nz, int j, double d_1, double *out)
{
  const int mx = nx + 2 * NG, my = ny + 2 * NG;
  for (int n = 0; n < nz; n++) for (int m = 0; m < ny; m++) for (int l = 0; l < nx; l++)
    out[((long)n * ny + m) * nx + l] = pcref_der2(f, mx, my, l + NG, m + NG, n + NG, j, d_1);
}
void pcref_der6_field(const double *f, int nx, int ny, int nz, int j, double d_1, double *out)
{
  const int mx = nx + 2 * NG, my = ny + 2 * NG;
  for (int n = 0; n < nz; n++) for (int m = 0; m < ny; m++) for (int l = 0; l < nx; l++)
    out[((long)n * ny + m) * nx + l] = pcref_der6_upwind(f, mx, my, l + NG, m + NG, n + NG, j, d_1);
}
void pcref_derij_field(const double *f, int nx, int ny, int nz, int i, int j,
                       double di_1, double dj_1, double *out)
{
  const int mx = nx + 2 * NG, my = ny + 2 * NG;
  for (int n = 0; n < nz; n++) for (int m = 0; m < ny; m++) for (int l = 0; l < nx; l++)
    out[((long)n * ny + m) * nx + l] = pcref_derij(f, mx, my, l + NG, m + NG, n + NG, i, j, di_1, dj_1);
}

/* ------------------------------------------------------------------------
 * Boundary conditions: boundconds_x, boundconds_y, boundconds_z
 * (boundcond.f90:659,1018,1283), applied in this order over the FULL extent
 * of the other two axes (so edges/corners get the reference's values).
 * Per axis: loop over faces (bot, top), inside over variables (:1303-1316).
 * bc_per_* boundcond.f90:1813-1961; bc_sym_* :3160-3222 ('s','a','a2');
 * 'c1' bc_ss_flux eos_idealgas.f90:2245-2347; 'cT' bc_ss_temp_z :3434-3563.
 * ---------------------------------------------------------------------- */
static void bc_axis(double *f, const pcb200_params *par, int axis)
{
  const int mx = par->nx + 2 * NG, my = par->ny + 2 * NG, mz = par->nz + 2 * NG;
  const int mm[3] = {mx, my, mz};
  const int nn[3] = {par->nx, par->ny, par->nz};
  const long str[3] = {1, mx, (long)mx * my};
  const int a = axis, b = (axis + 1) % 3, c = (axis + 2) % 3;
  const long sa = str[a];
  const int n1 = NG, n2 = NG + nn[a] - 1;    /* first/last interior index on this axis */
  const long mw = (long)mx * my * mz;
  const int32_t (*bc)[PCB200_MAXVAR] = a == 0 ? par->bcx : (a == 1 ? par->bcy : par->bcz);
  if (nn[a] == 1) return;                     /* degenerate direction: no boundary */
  /* units_eos, eos_idealgas.f90:127-128,236-238 */
  const double cp = par->cp, gamma1 = 1 / par->gamma, cv = gamma1 * cp;
  const double gamma_m1 = par->gamma - 1.0, cv1 = par->gamma * (1 / cp);
  for (int face = 0; face < 2; face++) {
    for (int k = 0; k < par->mvar; k++) {
      double *fk = f + k * mw;
      const int code = bc[face][k];
      for (int ic = 0; ic < mm[c]; ic++) for (int ib = 0; ib < mm[b]; ib++) {
        double *p = fk + ib * str[b] + ic * str[c];        /* line along axis a */
        if (code == PCB200_BC_P) {
          if (face == 0) { for (int i = 0; i < NG; i++) p[i * sa] = p[(n2 - NG + 1 + i) * sa]; }
          else           { for (int i = 0; i < NG; i++) p[(n2 + 1 + i) * sa] = p[(n1 + i) * sa]; }
        } else if (code == PCB200_BC_S || code == PCB200_BC_A) {
          const double sgn = code == PCB200_BC_S ? 1.0 : -1.0;
          if (face == 0) {
            for (int i = 1; i <= NG; i++) p[(n1 - i) * sa] = sgn * p[(n1 + i) * sa];
            if (sgn < 0) p[n1 * sa] = 0.;
          } else {
            for (int i = 1; i <= NG; i++) p[(n2 + i) * sa] = sgn * p[(n2 - i) * sa];
            if (sgn < 0) p[n2 * sa] = 0.;
          }
        } else if (code == PCB200_BC_A2) {
          if (face == 0) { for (int i = 1; i <= NG; i++) p[(n1 - i) * sa] = 2 * p[n1 * sa] + (-1.0) * p[(n1 + i) * sa]; }
          else           { for (int i = 1; i <= NG; i++) p[(n2 + i) * sa] = p[n2 * sa] + (p[n2 * sa] + (-1.0) * p[(n2 - i) * sa]); }
        } else if (code == PCB200_BC_C1 && a == 2 && face == 0) {
          /* bc_ss_flux 'bot', not chi-const/kramers branch unless iheatcond says so */
          const double *lnr = f + (long)(par->ilnrho - 1) * mw + ib * str[b] + ic * str[c];
          const double rho_xy = exp(lnr[n1 * sa]);                       /* getrho */
          double cs2_xy = p[n1 * sa];
          cs2_xy = par->cs20 * exp(gamma_m1 * (lnr[n1 * sa] - par->lnrho0) + cv1 * cs2_xy);
          double tmp_xy;
          if (par->iheatcond == PCB200_HCOND_CHICONST) tmp_xy = par->Fbot / (rho_xy * par->chi * cs2_xy);
          else { const double FbotKbot = par->Fbot / (par->hcond0 * par->hcond1); tmp_xy = FbotKbot / cs2_xy; }
          for (int i = 1; i <= NG; i++) {
            const double dlnrho = lnr[(n1 + i) * sa] - lnr[(n1 - i) * sa];   /* getdlnrho_z */
            p[(n1 - i) * sa] = p[(n1 + i) * sa] + cp * (cp - cv) * (dlnrho + par->dz2_bound[3 - i] * tmp_xy);
          }
        } else if (code == PCB200_BC_CT && a == 2 && face == 1) {
          /* bc_ss_temp_z 'top' */
          const double *lnr = f + (long)(par->ilnrho - 1) * mw + ib * str[b] + ic * str[c];
          const double tmp = 2 * cv * log(par->cs2top / par->cs20);
          p[n2 * sa] = 0.5 * tmp - (cp - cv) * (lnr[n2 * sa] - par->lnrho0);
          for (int i = 1; i <= NG; i++)
            p[(n2 + i) * sa] = -p[(n2 - i) * sa] + tmp
                               - (cp - cv) * (lnr[(n2 - i) * sa] + lnr[(n2 + i) * sa] - 2 * par->lnrho0);
        }
      }
    }
  }
}

void pcref_boundconds(double *f, const pcb200_params *par)
{
  bc_axis(f, par, 0);
  bc_axis(f, par, 1);
  bc_axis(f, par, 2);
}

/* ------------------------------------------------------------------------
 * rhs_cpu mn-loop, equ.f90:816-1239, evaluated point by point (every pencil
 * operation of the reference is elementwise in x, so a pencil of nx points
 * and nx scalar evaluations perform identical arithmetic).
 * ---------------------------------------------------------------------- */
typedef struct { double dt1, d[PCB200_NDIAG]; } pt_out;

static inline double dot3(const double *a, const double *b)
{ /* dot_mn, sub.f90:585-616 */
  double c = a[0] * b[0]; c = c + a[1] * b[1]; c = c + a[2] * b[2]; return c;
}
static inline double dot2v(const double *a)
{ /* dot2_mn, sub.f90:841-885 */
  return a[0] * a[0] + a[1] * a[1] + a[2] * a[2];
}
static inline void cross3(const double *a, const double *b, double *c)
{ /* cross_mn, sub.f90:1368-1382 */
  c[0] = a[1] * b[2] - a[2] * b[1];
  c[1] = a[2] * b[0] - a[0] * b[2];
  c[2] = a[0] * b[1] - a[1] * b[0];
}

static void rhs_point(const double *f, double *df, const pcb200_params *par,
                      int l, int m, int n, int lfirst, pt_out *out)
{
  const int mx = par->nx + 2 * NG, my = par->ny + 2 * NG, mz = par->nz + 2 * NG;
  const long mw = (long)mx * my * mz;
  const long idx = ((long)n * my + m) * mx + l;
  const double d1[3] = {par->nx > 1 ? par->dx_1 : 0., par->ny > 1 ? par->dy_1 : 0., par->nz > 1 ? par->dz_1 : 0.};
  /* grid.f90:2567-2570: dxyz_2 = dline_1(:,1)**2+dline_1(:,2)**2+dline_1(:,3)**2 */
  const double dxyz_2 = d1[0] * d1[0] + d1[1] * d1[1] + d1[2] * d1[2];
  const int ldt1 = lfirst && par->ldt;
#define FV(k) (f + (long)(k) * mw)
#define DER(k, j)  (d1[(j)-1] == 0. ? 0. : pcref_der (FV(k), mx, my, l, m, n, (j), d1[(j)-1]))
#define DER2(k, j) (d1[(j)-1] == 0. ? 0. : pcref_der2(FV(k), mx, my, l, m, n, (j), d1[(j)-1]))
#define DER6U(k, j) (d1[(j)-1] == 0. ? 0. : pcref_der6_upwind(FV(k), mx, my, l, m, n, (j), d1[(j)-1]))
#define DERIJ(k, i, j) ((d1[(i)-1] == 0. || d1[(j)-1] == 0.) ? 0. : \
        (par->lbidiagonal_derij ? pcref_derij(FV(k), mx, my, l, m, n, (i), (j), d1[(i)-1], d1[(j)-1]) \
                                : pcref_derij_full(FV(k), mx, my, l, m, n, (i), (j), d1[(i)-1], d1[(j)-1])))
  const int iu = par->iux - 1, ilr = par->ilnrho - 1, is = par->iss - 1, ia = par->iax - 1;

  double maxadvec = 0., advec2 = 0., advec_cs2 = 0., maxdiffus = 0.;
  double advec_uu = 0., diffus_nu = 0., diffus_chi = 0.;
  double dt1_max_pt = 0.;

  /* ---- calc_pencils_hydro_nonlinear, hydro.f90:2913-3207 ---- */
  double uu[3] = {0, 0, 0}, u2 = 0, uij[3][3], divu = 0, sij[3][3], sij2 = 0, oo[3] = {0, 0, 0};
  double ugu[3], del2u[3], graddivu[3];
  if (par->lhydro) {
    for (int i = 0; i < 3; i++) uu[i] = FV(iu + i)[idx];
    u2 = dot2v(uu);
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) uij[i][j] = DER(iu + i, j + 1);     /* gij sub.f90:1417 */
    divu = uij[0][0] + uij[1][1] + uij[2][2];                                                    /* div_mn :1767 */
    for (int j = 0; j < 3; j++) {                                                                /* traceless_strain :8314 */
      sij[j][j] = uij[j][j] - (1. / 3.) * divu;
      for (int i = j + 1; i < 3; i++) { sij[i][j] = .5 * (uij[i][j] + uij[j][i]); sij[j][i] = sij[i][j]; }
    }
    sij2 = sij[0][0] * sij[0][0];                                                                /* multm2_sym_mn :1171 */
    for (int i = 1; i < 3; i++) {
      sij2 = sij2 + sij[i][i] * sij[i][i];
      for (int j = 0; j < i; j++) sij2 = sij2 + 2 * (sij[i][j] * sij[i][j]);
    }
    oo[0] = uij[2][1] - uij[1][2]; oo[1] = uij[0][2] - uij[2][0]; oo[2] = uij[1][0] - uij[0][1]; /* curl_mn :1829 */
    for (int j = 0; j < 3; j++) {                                                                /* u_dot_grad_vec :3227 */
      ugu[j] = dot3(uu, uij[j]);
      if (par->lupw_uu) {                                                                        /* doupwind :6941 */
        double d6[3];
        for (int q = 0; q < 3; q++) d6[q] = fabs(uu[q]) * DER6U(iu + j, q + 1);
        ugu[j] = ugu[j] - ((d6[0] + d6[1]) + d6[2]);
      }
    }
    /* del2v_etc, sub.f90:2397-2536: fijj(i,j)=der2(f_i,j); fjji(i,j)=derij(f_j,j,i),
     * diagonal i=j keeps the der2 value left in tmp (:2444-2452) */
    double fijj[3][3], fjji[3][3];
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
      fijj[i][j] = DER2(iu + i, j + 1);
      fjji[i][j] = (i == j) ? fijj[i][j] : DERIJ(iu + j, j + 1, i + 1);
    }
    for (int i = 0; i < 3; i++) {
      del2u[i] = fijj[i][0] + fijj[i][1] + fijj[i][2];
      graddivu[i] = fjji[i][0] + fjji[i][1] + fjji[i][2];
    }
  }

  /* ---- calc_pencils_log_density_pnc, density.f90:2341-2421 ---- */
  double lnrho = 0, rho1 = 1, rho = 1, glnrho[3] = {0, 0, 0}, uglnrho = 0, sglnrho[3] = {0, 0, 0}, del2lnrho = 0;
  if (par->ldensity) {
    lnrho = FV(ilr)[idx];
    rho1 = exp(-lnrho);
    rho = 1.0 / rho1;
    for (int j = 0; j < 3; j++) glnrho[j] = DER(ilr, j + 1);                                     /* grad sub.f90:1488 */
    uglnrho = dot3(uu, glnrho);
    if (par->lupw_lnrho) {
      double d6[3];
      for (int q = 0; q < 3; q++) d6[q] = fabs(uu[q]) * DER6U(ilr, q + 1);
      uglnrho = uglnrho - ((d6[0] + d6[1]) + d6[2]);
    }
    if (par->lentropy) del2lnrho = (DER2(ilr, 1) + DER2(ilr, 2)) + DER2(ilr, 3);                 /* del2 sub.f90:2091 */
    if (par->lhydro) for (int i = 0; i < 3; i++) {                                               /* multmv_mn :1192 */
      double t = sij[i][0] * glnrho[0]; t = t + sij[i][1] * glnrho[1]; t = t + sij[i][2] * glnrho[2];
      sglnrho[i] = t;
    }
  }

  /* ---- calc_pencils_eos_pencpar, eos_idealgas.f90:771-870 (ilnrho_ss) ---- */
  const double cp = par->cp, cp1 = 1 / cp, gamma = par->gamma, gamma1 = 1 / gamma, gamma_m1 = gamma - 1.0;
  const double cv = gamma1 * cp, cv1 = gamma * cp1;
  const double lnTT0 = gamma_m1 != 0.0 ? log(par->cs20 / (cp * gamma_m1)) : log(par->cs20 / cp);
  double cs2 = par->cs20, ss = 0, gss[3] = {0, 0, 0}, del2ss = 0, lnTT = 0, TT1 = 0, glnTT[3] = {0, 0, 0}, del2lnTT = 0, ee = 0;
  if (par->lentropy) {
    ss = FV(is)[idx];
    for (int j = 0; j < 3; j++) gss[j] = DER(is, j + 1);
    del2ss = (DER2(is, 1) + DER2(is, 2)) + DER2(is, 3);
    cs2 = par->cs20 * exp(cv1 * ss + gamma_m1 * (lnrho - par->lnrho0));
    lnTT = lnTT0 + cv1 * ss + gamma_m1 * (lnrho - par->lnrho0);
    ee = cv * exp(lnTT);
    TT1 = exp(-lnTT);
    for (int j = 0; j < 3; j++) glnTT[j] = gamma_m1 * glnrho[j] + cv1 * gss[j];
    del2lnTT = gamma_m1 * del2lnrho + cv1 * del2ss;
  } else if (gamma_m1 != 0.0) {
    /* noentropy with gamma /= 1 selects the ISENTROPIC eos (noentropy.f90:113-116: select_eos_variable('ss',-1));
     * leos_isentropic branch of calc_pencils_eos, eos_idealgas.f90:809-815 */
    cs2 = par->cs20 * exp(gamma_m1 * (lnrho - par->lnrho0));
  }

  /* ---- calc_pencils_energy: entropy.f90:3165-3240 / noentropy.f90:255-313 ---- */
  double ugss = 0, fpres[3] = {0, 0, 0};
  if (par->lentropy) {
    ugss = dot3(uu, gss);
    if (par->lupw_ss) {
      double d6[3];
      for (int q = 0; q < 3; q++) d6[q] = fabs(uu[q]) * DER6U(is, q + 1);
      ugss = ugss - ((d6[0] + d6[1]) + d6[2]);
    }
    for (int j = 0; j < 3; j++) fpres[j] = -cs2 * (glnrho[j] + glnTT[j]) * gamma1;
  } else if (par->ldensity) {
    for (int j = 0; j < 3; j++) fpres[j] = -cs2 * glnrho[j];
  }

  /* ---- calc_pencils_viscosity, nu-const, viscosity.f90:1421-1448 ---- */
  double fvisc[3] = {0, 0, 0}, visc_heat = 0;
  if (par->lhydro) {
    for (int j = 0; j < 3; j++) {
      if (par->ldensity) fvisc[j] = 0. + par->nu * (del2u[j] + 2. * sglnrho[j] + 1. / 3. * graddivu[j]);
      else               fvisc[j] = 0. + par->nu * (del2u[j] + 1.0 / 3.0 * graddivu[j]);
    }
    visc_heat = 0. + 2 * par->nu * sij2;
  }

  /* ---- calc_pencils_magnetic_pencpar, magnetic.f90:3587-3990 ---- */
  const double mu01 = 1. / par->mu0;
  double aa[3], aij[3][3], bb[3] = {0, 0, 0}, b2 = 0, uxb[3], del2a[3], jj[3] = {0, 0, 0}, j2 = 0, jxb[3], jxbr[3], ab = 0, jb = 0;
  if (par->lmagnetic) {
    for (int i = 0; i < 3; i++) aa[i] = FV(ia + i)[idx];
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) aij[i][j] = DER(ia + i, j + 1);
    bb[0] = aij[2][1] - aij[1][2]; bb[1] = aij[0][2] - aij[2][0]; bb[2] = aij[1][0] - aij[0][1];
    b2 = dot2v(bb);
    ab = dot3(aa, bb);
    cross3(uu, bb, uxb);
    /* gij_etc, sub.f90:2732-2900: d2A(j,k,i) = d^2 A_i / dx_j dx_k */
    double d2A[3][3][3];
    for (int i = 0; i < 3; i++) {
      for (int j = 0; j < 3; j++) d2A[j][j][i] = DER2(ia + i, j + 1);
      d2A[1][2][i] = DERIJ(ia + i, 2, 3); d2A[2][1][i] = d2A[1][2][i];
      d2A[2][0][i] = DERIJ(ia + i, 3, 1); d2A[0][2][i] = d2A[2][0][i];
      d2A[0][1][i] = DERIJ(ia + i, 1, 2); d2A[1][0][i] = d2A[0][1][i];
    }
    double bij[3][3];
    for (int q = 0; q < 3; q++) {
      bij[0][q] = d2A[1][q][2] - d2A[2][q][1];
      bij[1][q] = d2A[2][q][0] - d2A[0][q][2];
      bij[2][q] = d2A[0][q][1] - d2A[1][q][0];
    }
    for (int i = 0; i < 3; i++) del2a[i] = d2A[0][0][i] + d2A[1][1][i] + d2A[2][2][i];
    jj[0] = bij[2][1] - bij[1][2]; jj[1] = bij[0][2] - bij[2][0]; jj[2] = bij[1][0] - bij[0][1];
    for (int i = 0; i < 3; i++) jj[i] = mu01 * jj[i];                                            /* :3847 */
    j2 = dot2v(jj);
    jb = dot3(jj, bb);
    cross3(jj, bb, jxb);
    for (int i = 0; i < 3; i++) jxbr[i] = rho1 * jxb[i];                                         /* multsv_mn */
  }

  /* ================= duu_dt, hydro.f90:3613-3922 ================= */
  if (par->lhydro) {
    for (int j = 0; j < 3; j++) df[(long)(iu + j) * mw + idx] = df[(long)(iu + j) * mw + idx] - ugu[j];   /* :3704 */
    if (par->Omega != 0.) {                                                                      /* coriolis_cartesian :5262 */
      const double c2 = 2 * par->Omega;
      df[(long)(iu) * mw + idx] = df[(long)(iu) * mw + idx] + c2 * uu[1];
      df[(long)(iu + 1) * mw + idx] = df[(long)(iu + 1) * mw + idx] - c2 * uu[0];
    }
    for (int j = 0; j < 3; j++) df[(long)(iu + j) * mw + idx] = df[(long)(iu + j) * mw + idx] + fvisc[j];  /* calc_viscous_force viscosity.f90:2571 */
    if (ldt1) {
      diffus_nu = par->nu * dxyz_2;
      maxdiffus = fmax(maxdiffus, diffus_nu);
      advec_uu = (fabs(uu[0]) * d1[0] + fabs(uu[1]) * d1[1]) + fabs(uu[2]) * d1[2];              /* hydro.f90:3810 */
      maxadvec = maxadvec + advec_uu;
    }
  }
  /* ================= dlnrho_dt, density.f90:2474-2872 ================= */
  if (par->ldensity) {
    double density_rhs = -divu;
    density_rhs = density_rhs - uglnrho;
    df[(long)ilr * mw + idx] = df[(long)ilr * mw + idx] + density_rhs;
  }
  /* ================= denergy_dt ================= */
  double Hmax = 0.;
  if (par->ldensity && par->lhydro && ldt1) advec_cs2 = cs2 * dxyz_2;                            /* noentropy.f90:367 / entropy.f90:3277 */
  if (par->lhydro && par->ldensity)
    for (int j = 0; j < 3; j++) df[(long)(iu + j) * mw + idx] = df[(long)(iu + j) * mw + idx] + fpres[j];
  if (par->lentropy) {
    double *dss = df + (long)is * mw + idx;
    *dss = *dss - ugss;                                                                          /* entropy.f90:3348 */
    if (par->lviscosity_heat && par->lhydro) {                                                   /* calc_viscous_heat viscosity.f90:2524 */
      *dss = *dss + TT1 * visc_heat;
      if (ldt1) Hmax = Hmax + visc_heat;
    }
    if (par->iheatcond == PCB200_HCOND_KPROFILE) {                                               /* calc_heatcond entropy.f90:5333 */
      const double hcond = g_zprof[PCB200_PROF_HCOND][n - NG];
      const double glhc[3] = {0., 0., g_zprof[PCB200_PROF_DLNHCOND][n - NG]};
      const double chix = rho1 * hcond * cp1;
      double glnThcond[3];
      for (int j = 0; j < 3; j++) glnThcond[j] = glnTT[j] + glhc[j];
      const double g2 = dot3(glnTT, glnThcond);
      const double thdiff = rho1 * hcond * (del2lnTT + g2);
      *dss = *dss + thdiff;
      if (ldt1) diffus_chi = diffus_chi + gamma * chix * dxyz_2;
    } else if (par->iheatcond == PCB200_HCOND_CHICONST) {                                        /* calc_heatcond_constchi :4266 */
      double gsum[3];
      for (int j = 0; j < 3; j++) gsum[j] = glnrho[j] + glnTT[j];
      const double g2 = dot3(gsum, glnTT);
      const double thdiff = par->chi * cp * (del2lnTT + g2);
      *dss = *dss + thdiff;
      if (ldt1) diffus_chi = diffus_chi + (gamma * par->chi) * dxyz_2;
    }
    if (ldt1) maxdiffus = fmax(maxdiffus, diffus_chi);
    if (par->lcooling) {                                                                         /* calc_heat_cool :5830, get_heat_cool_gravz :6196 */
      double heat = 0.0;
      const double prof = g_zprof[PCB200_PROF_COOL][n - NG];
      heat = heat - par->cool * prof * (cs2 - par->cs2cool) / par->cs2cool;
      *dss = *dss + TT1 * rho1 * heat;
      if (ldt1) Hmax = Hmax + heat * rho1;
    }
    if (ldt1) dt1_max_pt = fmax(dt1_max_pt, fabs(Hmax / ee / par->cdts));                        /* :3444 */
  }
  /* ================= daa_dt, magnetic.f90:4229-5589 ================= */
  if (par->lmagnetic) {
    if (par->lhydro)
      for (int j = 0; j < 3; j++) df[(long)(iu + j) * mw + idx] = df[(long)(iu + j) * mw + idx] + jxbr[j];   /* :4360 */
    double dAdt[3];
    for (int j = 0; j < 3; j++) {
      const double fres = 0. + par->eta * del2a[j];                                              /* :4399 */
      dAdt[j] = 0. + uxb[j] + fres;                                                              /* :5238 */
    }
    if (par->lentropy && par->lohmic_heat)                                                       /* :5127 */
      df[(long)is * mw + idx] = df[(long)is * mw + idx] + par->eta * par->mu0 * j2 * rho1 * TT1;
    if (ldt1) {
      double advec_va2 = 0.;
      if (par->lhydro) {
        const double t0 = bb[0] * d1[0], t1 = bb[1] * d1[1], t2 = bb[2] * d1[2];
        advec_va2 = ((t0 * t0 + t1 * t1) + t2 * t2) * mu01 * rho1;                              /* :5438 */
      }
      advec2 = advec2 + advec_va2;
    }
    for (int j = 0; j < 3; j++) df[(long)(ia + j) * mw + idx] = df[(long)(ia + j) * mw + idx] + dAdt[j];   /* :5521 */
    if (ldt1) maxdiffus = fmax(maxdiffus, par->eta * dxyz_2);                                    /* :5551,5568 */
  }
  /* ================= duu_dt_grav, gravity_simple.f90:914 ================= */
  if (par->lgravz && par->lhydro) df[(long)(iu + 2) * mw + idx] = df[(long)(iu + 2) * mw + idx] + par->gravz;

  /* ================= time-step combination, equ.f90:1086-1151 ================= */
  if (ldt1) {
    advec2 = advec2 + advec_cs2;
    if (par->lentropy || par->ldensity || par->lmagnetic) maxadvec = maxadvec + sqrt(advec2);
    maxadvec = maxadvec + sqrt(0.0);                  /* advec2_hypermesh = 0 */
    const double dt1_advec = maxadvec / par->cdt;
    const double dt1_diffus = maxdiffus / par->cdtv;  /* + maxdiffus2/cdtv2 + maxdiffus3/cdtv3, both 0 */
    const double dt1_src = 0.;
    dt1_max_pt = fmax(dt1_max_pt, sqrt(dt1_advec * dt1_advec + dt1_diffus * dt1_diffus + dt1_src * dt1_src));
  }
  out->dt1 = dt1_max_pt;
  /* diagnostics (calc_0d_diagnostics_*: hydro.f90:3924, density.f90:2966, entropy.f90:3454, magnetic.f90:5676) */
  out->d[PCB200_DIAG_U2SUM] = u2;        out->d[PCB200_DIAG_U2MAX] = u2;
  out->d[PCB200_DIAG_RHOSUM] = par->ldensity ? rho : 0.;
  out->d[PCB200_DIAG_SSSUM] = ss;
  out->d[PCB200_DIAG_B2SUM] = b2;        out->d[PCB200_DIAG_B2MAX] = b2;
  out->d[PCB200_DIAG_J2SUM] = j2;        out->d[PCB200_DIAG_J2MAX] = j2;
  out->d[PCB200_DIAG_DTC] = sqrt(advec_cs2) / par->cdt;
  out->d[PCB200_DIAG_DTU] = advec_uu / par->cdt;
  out->d[PCB200_DIAG_DTNU] = diffus_nu / par->cdtv;
  out->d[PCB200_DIAG_DTCHI] = diffus_chi / par->cdtv;
  out->d[PCB200_DIAG_DT1MAX] = dt1_max_pt;
  out->d[PCB200_DIAG_ABSUM] = ab;
  out->d[PCB200_DIAG_JBSUM] = jb;
  out->d[PCB200_DIAG_OUSUM] = dot3(oo, uu);
#undef FV
#undef DER
#undef DER2
#undef DER6U
#undef DERIJ
}

static const int diag_is_max[PCB200_NDIAG] = {0, 1, 0, 0, 0, 1, 0, 1, 1, 1, 1, 1, 1, 0, 0, 0};

void pcref_rhs(const double *f, double *df, const pcb200_params *par,
               int lfirst, double *dt1_max, double *diag)
{
  const int nx = par->nx, ny = par->ny, nz = par->nz;
  double dt1 = dt1_max ? *dt1_max : 0.;
  /* per-pencil partial results so that the reduction order does not depend on
   * the thread count (sum_mn_name adds pencil sums in mn order,
   * diagnostics.f90:1751-1886) */
  double *pen = diag ? (double *)calloc((size_t)ny * nz * PCB200_NDIAG, sizeof(double)) : 0;
#pragma omp parallel for collapse(2) schedule(static) reduction(max : dt1)
  for (int n = 0; n < nz; n++) {
    for (int m = 0; m < ny; m++) {
      double acc[PCB200_NDIAG];
      for (int q = 0; q < PCB200_NDIAG; q++) acc[q] = 0.;
      for (int l = 0; l < nx; l++) {
        pt_out o;
        rhs_point(f, df, par, l + NG, m + NG, n + NG, lfirst, &o);
        if (o.dt1 > dt1) dt1 = o.dt1;
        if (pen) for (int q = 0; q < PCB200_NDIAG; q++) {
          if (diag_is_max[q]) acc[q] = fmax(acc[q], o.d[q]); else acc[q] += o.d[q];
        }
      }
      if (pen) memcpy(pen + ((size_t)n * ny + m) * PCB200_NDIAG, acc, sizeof(acc));
    }
  }
  if (dt1_max) *dt1_max = dt1;
  if (diag) {
    for (size_t p = 0; p < (size_t)ny * nz; p++) for (int q = 0; q < PCB200_NDIAG; q++) {
      const double v = pen[p * PCB200_NDIAG + q];
      if (diag_is_max[q]) diag[q] = fmax(diag[q], v); else diag[q] += v;
    }
    free(pen);
  }
}

/* forcing_hel application, forcing.f90:1968-2110 with the default switches
 * (force_ampl=1, profx_hel=profyz_hel=1, omega_ff=0, force1_scl=1):
 * f_u_j += fda_j * Re[(coef1_j + i coef2_j) * fx(l) fy(m) fz(n)],
 * fx,fy,fz from fxyz_coefs_hel (:1788-1849). */
void pcref_addforce(double *f, const pcb200_params *par, const pcb200_forcing *frc)
{
  const int mx = par->nx + 2 * NG, my = par->ny + 2 * NG, mz = par->nz + 2 * NG;
  const long mw = (long)mx * my * mz;
  const int iu = par->iux - 1;
#pragma omp parallel for collapse(2) schedule(static)
  for (int n = NG; n < NG + par->nz; n++) for (int m = NG; m < NG + par->ny; m++) {
    const double zz = par->z0_loc + (n - NG) * par->dz, yy = par->y0_loc + (m - NG) * par->dy;
    const double fyr = cos(frc->kk[1] * yy), fyi = sin(frc->kk[1] * yy);
    const double fzr = cos(frc->kk[2] * zz), fzi = sin(frc->kk[2] * zz);
    for (int l = NG; l < NG + par->nx; l++) {
      const double xx = par->x0_loc + (l - NG) * par->dx;
      const double arg = frc->kk[0] * xx + frc->phase;
      const double fxr = cos(arg) * frc->fact, fxi = sin(arg) * frc->fact;
      const double pr = fxr * fyr - fxi * fyi, pi_ = fxr * fyi + fxi * fyr;      /* fx*fy */
      const double qr = pr * fzr - pi_ * fzi, qi = pr * fzi + pi_ * fzr;         /* (fx*fy)*fz */
      const long idx = ((long)n * my + m) * mx + l;
      for (int j = 0; j < 3; j++) {
        const double re = frc->coef1[j] * qr - frc->coef2[j] * qi;
        f[(long)(iu + j) * mw + idx] = f[(long)(iu + j) * mw + idx] + frc->fda[j] * re;
      }
    }
  }
}

/* time_step, timestep.f90:65-198 (itorder=3), around pde, equ.f90:24-668 */
void pcref_timestep(double *f, double *df, const pcb200_params *par,
                    const pcb200_forcing *frc, double *dt, double *diag)
{
  const int mx = par->nx + 2 * NG, my = par->ny + 2 * NG, mz = par->nz + 2 * NG;
  const long mw = (long)mx * my * mz;
  const long ntot = mw * par->mvar;
  for (int itsub = 0; itsub < 3; itsub++) {
    const int lfirst = itsub == 0;
    if (lfirst) memset(df, 0, sizeof(double) * (size_t)ntot);                    /* timestep.f90:119 */
    else {
      const double a = par->alpha_ts[itsub];
#pragma omp parallel for schedule(static)
      for (long i = 0; i < ntot; i++) df[i] = a * df[i];                         /* :122 */
    }
    pcref_boundconds(f, par);                                                    /* equ.f90:214-229 */
    double dt1_max = 0.0;
    if (lfirst && par->ldt) dt1_max = par->dtmax != 0.0 ? 1. / par->dtmax : 0.0; /* equ.f90:280-290 */
    pcref_rhs(f, df, par, lfirst, &dt1_max, lfirst ? diag : 0);
    if (lfirst && par->ldt) *dt = 1.0 / dt1_max;                                 /* set_dt sub.f90:3896-3915 */
    if (lfirst && !par->ldt) *dt = par->dt_fixed;
    const double dt_beta = *dt * par->beta_ts[itsub];                            /* timestep.f90:152 */
#pragma omp parallel for collapse(2) schedule(static)
    for (int k = 0; k < par->mvar; k++) for (int n = NG; n < NG + par->nz; n++)
      for (int m = NG; m < NG + par->ny; m++) for (int l = NG; l < NG + par->nx; l++) {
        const long i = (long)k * mw + ((long)n * my + m) * mx + l;
        f[i] = f[i] + dt_beta * df[i];                                           /* :162-163 */
      }
  }
  if (frc) pcref_addforce(f, par, frc);                                          /* run.f90:729 */
}
