/* pcref.h -- CPU ORACLE (test infrastructure, NOT product code).
 *
 * A restatement in C of the Pencil Code's pencil-wise RHS / 2N-RK3 hot path
 * (reference: /root/reference/src/{deriv,sub,hydro,density,noentropy,entropy,
 * eos_idealgas,viscosity,magnetic,gravity_simple,boundcond,equ,timestep}.f90).
 * Each function cites the file:line it follows and keeps the Fortran order of
 * floating-point operations (compile with -ffp-contract=off).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.  The product (libpcb200.so)
 * never links or calls it.
 *
 * PARITY PIN: the reference is Fortran+MPI and cannot be compiled in this
 * image (no Fortran compiler).  The oracle is pinned against the reference's
 * own golden time series samples/conv-slab/reference.out.double through
 * tests/test_oracle_convslab.py (initial condition restated in
 * oracle/initcond.py); see DESIGN.md "Oracle".
 */
#ifndef PCREF_H
#define PCREF_H
#include "../include/pcb200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* deriv.f90 stencils on one point; f points at the variable's (mx,my,mz) block,
 * (l,m,n) are 0-based indices including ghosts; j = 1,2,3 axis. */
double pcref_der (const double *f, int mx, int my, int l, int m, int n, int j, double d_1);
double pcref_der2(const double *f, int mx, int my, int l, int m, int n, int j, double d_1);
double pcref_der6_upwind(const double *f, int mx, int my, int l, int m, int n, int j, double d_1);
double pcref_derij(const double *f, int mx, int my, int l, int m, int n, int i, int j,
                   double di_1, double dj_1);

/* lbidiagonal_derij=F: 36-point product stencil (deriv.f90:1430-1557) */
double pcref_derij_full(const double *f, int mx, int my, int l, int m, int n, int i, int j,
                        double di_1, double dj_1);

void pcref_derij_full_field(const double *f, int nx, int ny, int nz, int i, int j,
                            double di_1, double dj_1, double *out);

/* whole-array versions used by the unit tests (interior only, out has nx*ny*nz) */
void pcref_der_field (const double *f, int nx, int ny, int nz, int j, double d_1, double *out);
void pcref_der2_field(const double *f, int nx, int ny, int nz, int j, double d_1, double *out);
void pcref_der6_field(const double *f, int nx, int ny, int nz, int j, double d_1, double *out);
void pcref_derij_field(const double *f, int nx, int ny, int nz, int i, int j,
                       double di_1, double dj_1, double *out);

/* boundconds_x/y/z on the full ghost-padded array (boundcond.f90:659,1018,1283),
 * single domain (the oracle always runs the global grid on one process). */
void pcref_boundconds(double *f, const pcb200_params *par);

/* rhs_cpu mn-loop (equ.f90:816-1239): df += RHS(f).  If lfirst, *dt1_max is
 * updated (must be initialised by the caller, equ.f90:280-290); if diag!=NULL
 * (PCB200_NDIAG doubles, zero-initialised by the caller) the 0-D diagnostics
 * are accumulated. */
void pcref_rhs(const double *f, double *df, const pcb200_params *par,
               int lfirst, double *dt1_max, double *diag);

/* time_step (timestep.f90:65-198): 3 substeps incl. boundary conditions,
 * set_dt (sub.f90:3896) and the optional forcing kick (run.f90:729).
 * df is scratch (mx*my*mz*mvar). *dt in/out. */
void pcref_timestep(double *f, double *df, const pcb200_params *par,
                    const pcb200_forcing *frc, double *dt, double *diag);
/* pcref_pencil.c: the same time step evaluated pencil-wise (timing arm of bench.py only; falls back to
 * pcref_timestep where it does not apply or when diagnostics are requested) */
int pcref_pencil_applies(const pcb200_params *par);
const double *pcref_get_zprofile(int which);
void pcref_timestep_pencil(double *f, double *df, const pcb200_params *par,
                           const pcb200_forcing *frc, double *dt, double *diag);

/* forcing_hel application (forcing.f90:1968-2110) */
void pcref_addforce(double *f, const pcb200_params *par, const pcb200_forcing *frc);

int pcref_num_threads(void);
void pcref_set_num_threads(int n);

/* ---- dust particles (pcref_particles.c; parity unpinned, see its header) ---- */
void pcref_particles_boundconds(double *fp, const pcb200_params *par, const pcb200_particle_params *pp);
void pcref_particles_map_nearest_grid(const double *fp, int32_t *ineargrid, const pcb200_params *par,
                                      const pcb200_particle_params *pp);
void pcref_setup_imn_array(const pcb200_params *par, int32_t *imn_array);
void pcref_particles_sort(double *fp, double *dfp, int32_t *ipar, int32_t *ineargrid, int32_t *k1_imn, int32_t *k2_imn,
                          const pcb200_params *par, const pcb200_particle_params *pp);
void pcref_particles_map_xxp_grid(const double *fp, const int32_t *ineargrid, double *np, double *rhop,
                                  const pcb200_params *par, const pcb200_particle_params *pp);
void pcref_particles_fixture(int interp_cic, const double *ap_by_ipar, long nap, double rhopmat, const int32_t *ipar);
void pcref_particles_pde(const double *f, double *df, const double *fp, double *dfp, const int32_t *ineargrid,
                         const pcb200_params *par, const pcb200_particle_params *pp, int lfirst, double *dt1_max);
void pcref_timestep_particles(double *f, double *df, const pcb200_params *par, const pcb200_forcing *frc,
                              double *dt, double *diag, const pcb200_particle_params *pp,
                              double *fp, double *dfp, int32_t *ipar, int32_t *ineargrid,
                              double *np, double *rhop, int32_t *k1_imn, int32_t *k2_imn);

#ifdef __cplusplus
}
#endif
#endif
