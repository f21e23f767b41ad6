/* pcb200.h -- C ABI of the B200-native solver core for the Pencil Code.
 *
 * This is the drop-in boundary: the Fortran module `GPU` (public list in the
 * reference's src/gpu.h:1 -- gpu_init, register_GPU, initialize_GPU,
 * finalize_GPU, rhs_GPU, copy_farray_from_GPU) is re-implemented by
 * pencil-code_b200/fortran/gpu_b200.f90 on top of the entry points below
 * through ISO_C_BINDING.  Every entry point names the reference interface it
 * replaces.  Plain pointers and sizes only; all reals are FP64
 * (REAL_PRECISION=double), all integers 32 bit (FINT, src/headers_c.h:15).
 *
 * Every function returns 0 on success and a negative PCB200_E* code on
 * failure; pcb200_last_error() gives the message.  There is NO CPU fallback:
 * without a CUDA device pcb200_init() fails.
 */
#ifndef PCB200_H
#define PCB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PCB200_ABI_VERSION 1
#define PCB200_MAXVAR 8      /* ux uy uz lnrho ss ax ay az */
#define PCB200_NGHOST 3      /* src/deriv.f90:7 / cparam.f90 nghost */

/* error codes */
#define PCB200_OK          0
#define PCB200_ENODEV     -1  /* no CUDA device / driver            */
#define PCB200_EUNSUPP    -2  /* unsupported module / flag combo    */
#define PCB200_EARG       -3  /* bad argument                       */
#define PCB200_ECUDA      -4  /* CUDA runtime error                 */
#define PCB200_ESTATE     -5  /* call out of order                  */
#define PCB200_ENCCL      -6  /* NCCL error                         */

/* boundary-condition codes, one per variable and face.
 * Reference strings (run.in bcx/bcy/bcz, src/boundcond.f90:1283-1460):
 *  'p' periodic, 's' symmetric, 'a' antisymmetric, 'a2' antisym. rel. to
 *  boundary value, 'c1' constant heat flux (ss, bottom; bc_ss_flux
 *  eos_idealgas.f90:2245), 'cT' constant temperature (ss; bc_ss_temp_z
 *  eos_idealgas.f90:3434). */
enum {
  PCB200_BC_P  = 0,
  PCB200_BC_S  = 1,
  PCB200_BC_A  = 2,
  PCB200_BC_A2 = 3,
  PCB200_BC_C1 = 4,
  PCB200_BC_CT = 5
};

/* heat conduction choice (entropy.f90 iheatcond) */
enum {
  PCB200_HCOND_NONE     = 0,
  PCB200_HCOND_KPROFILE = 1,  /* 'K-profile', calc_heatcond entropy.f90:5333      */
  PCB200_HCOND_CHICONST = 2   /* 'chi-const', calc_heatcond_constchi :4266        */
};

/* Flat parameter block filled by the Fortran wrapper from cdata/cparam and the
 * module run_pars (replaces the generated cparam_c.h/cdata_c.h headers and the
 * pushpars2c pointer tables, gpu_astaroth_ansi.c:31-35, astaroth/gpu_astaroth.cc:242-317).
 * Variable slots are 1-based Fortran indices (iux..iaz in cdata), 0 = absent. */
typedef struct pcb200_params {
  int32_t abi_version;            /* must be PCB200_ABI_VERSION              */
  /* local (per rank) interior sizes and process layout (cparam.f90:18-55)  */
  int32_t nx, ny, nz;
  int32_t nprocx, nprocy, nprocz;
  int32_t ipx, ipy, ipz;
  int32_t lperi[3];               /* cdata lperi                             */
  int32_t mvar, mfarray;
  int32_t iux, iuy, iuz, ilnrho, iss, iax, iay, iaz;
  /* module switches (Makefile.local choices -> cparam l* constants)        */
  int32_t lhydro, ldensity, lentropy, lmagnetic, lgravz, lforcing;
  int32_t lupw_uu, lupw_lnrho, lupw_ss;   /* hydro/density/entropy run_pars  */
  int32_t lbidiagonal_derij;      /* cdata.f90:568; only 1 is supported      */
  int32_t iheatcond;              /* PCB200_HCOND_*                          */
  int32_t lcooling;               /* entropy cool/=0 (get_heat_cool_gravz)   */
  int32_t lviscosity_heat;        /* entropy.f90:132                         */
  int32_t lohmic_heat;            /* magnetic lohmic_heat                    */
  int32_t ldt;                    /* cdata ldt: dt from CFL (1) or fixed (0) */
  int32_t reserved_i[7];
  /* BC codes [face 0=lower,1=upper][variable slot-1]                       */
  int32_t bcx[2][PCB200_MAXVAR];
  int32_t bcy[2][PCB200_MAXVAR];
  int32_t bcz[2][PCB200_MAXVAR];
  /* equidistant Cartesian grid of this rank (grid.f90 construct_grid)      */
  double dx, dy, dz;              /* cdata dx,dy,dz                          */
  double dx_1, dy_1, dz_1;        /* cdata dx_1(l1),dy_1(m1),dz_1(n1): the
                                     host's own 1/dx (grid.f90:218), not 1/dx
                                     recomputed                              */
  double dz2_bound[7];            /* grid.f90:2670-2672, index i+3, i=-3..3  */
  double xyz0[3];                 /* lower corner of the GLOBAL box          */
  double Lxyz[3];
  double x0_loc, y0_loc, z0_loc;  /* x(l1), y(m1), z(n1) of this rank        */
  /* time step control (cdata.f90:145-153)                                  */
  double cdt, cdtv, cdts, dtmax, dt_fixed;
  double alpha_ts[3], beta_ts[3]; /* timestep.f90:40-41                      */
  /* physics                                                                */
  double nu;                      /* viscosity_run_pars, ivisc='nu-const'    */
  double eta;                     /* magnetic_run_pars, iresistivity='eta-const' */
  double mu0;                     /* cdata.f90:252                           */
  double cs20, gamma, cp, lnrho0; /* eos (units_eos eos_idealgas.f90:123)    */
  double Omega;                   /* hydro_run_pars, theta=0                 */
  double gravz;                   /* gravity_simple gravz_profile='const'    */
  double chi;                     /* entropy chi (chi-const)                 */
  double hcond0, hcond1, hcond2;  /* entropy K-profile, lmultilayer          */
  double z1, z2, widthss;         /* gravity z1,z2 ; entropy widthss         */
  double cool, wcool, cs2cool;    /* entropy cooling layer                   */
  double ztop;                    /* xyz0(3)+Lxyz(3)                         */
  double Fbot;                    /* entropy Fbot (bc 'c1')                  */
  double cs2top;                  /* entropy cs2top (bc 'cT')                */
  double reserved_d[8];
} pcb200_params;

/* helical forcing coefficients for one time step; the random wave vector and
 * phase stay on the host (forcing.f90:1554-1849 fconst_coefs_hel/fxyz_coefs_hel;
 * consumed like forcing_pars_hel, astaroth/forcing.h:19).
 * f_u_j += Re[(coef1_j + i coef2_j) * fx(l) fy(m) fz(n)] * fda_j  (forcing.f90:1968-2110) */
typedef struct pcb200_forcing {
  double coef1[3], coef2[3], fda[3];
  double kk[3];                   /* wave vector * k1_ff                     */
  double phase;                   /* random phase                            */
  double fact;                    /* force/ffnorm*sqrt(dt)  (forcing.f90:1741) */
} pcb200_forcing;

/* indices into the diagnostics vector returned by pcb200_get_diagnostics
 * (replaces calc_0d_diagnostics_* accumulation via sum_mn_name/max_mn_name,
 * diagnostics.f90:1686-1886; raw sums/maxima of this rank -- the caller
 * finishes them exactly like diagnostic(), diagnostics.f90:643-716). */
enum {
  PCB200_DIAG_U2SUM = 0,  /* sum u^2            -> urms = sqrt(sum/nw)       */
  PCB200_DIAG_U2MAX,      /* max u^2            -> umax = sqrt(max)          */
  PCB200_DIAG_RHOSUM,     /* sum rho            -> rhom                      */
  PCB200_DIAG_SSSUM,      /* sum ss             -> ssm                       */
  PCB200_DIAG_B2SUM,      /* sum b^2            -> brms                      */
  PCB200_DIAG_B2MAX,      /* max b^2            -> bmax                      */
  PCB200_DIAG_J2SUM,      /* sum j^2            -> jrms                      */
  PCB200_DIAG_J2MAX,      /* max j^2            -> jmax                      */
  PCB200_DIAG_DTC,        /* max sqrt(advec_cs2)/cdt      (x dt -> dtc)      */
  PCB200_DIAG_DTU,        /* max advec_uu/cdt             (x dt -> dtu)      */
  PCB200_DIAG_DTNU,       /* max nu*dxyz_2/cdtv           (x dt -> dtnu)     */
  PCB200_DIAG_DTCHI,      /* max diffus_chi/cdtv          (x dt -> dtchi)    */
  PCB200_DIAG_DT1MAX,     /* max dt1 of this rank (1/dt before allreduce)    */
  PCB200_DIAG_ABSUM,      /* sum a.b            -> abm                       */
  PCB200_DIAG_JBSUM,      /* sum j.b            -> jbm                       */
  PCB200_DIAG_OUSUM,      /* sum o.u            -> oum                       */
  PCB200_NDIAG = 16
};

/* gpu_init (run.f90:125; init_gpu_c_ gpu_astaroth_ansi.c:46): pick the
 * device for this rank (local_rank) and create streams. */
int pcb200_init(int local_rank);

/* initialize_GPU (register.f90:375; initialize_gpu_c_ :37): upload
 * parameters, (re)allocate device buffers. Idempotent (RELOAD, run.f90:571). */
int pcb200_initialize(const pcb200_params *par);

/* register_GPU (run.f90:223; register_gpu_c_ :54): remember the address of
 * the host array f(mx,my,mz,mfarray) (aliased, never freed). */
int pcb200_register(double *f_host);

/* rhs_GPU (equ.f90:343; rhs_gpu_c_ :89): one RK substep (BCs, halo exchange,
 * RHS, CFL dt on substep 1, 2N-RK3 update) fully on the device.
 * isubstep in {1,2,3}; full!=0 uploads the registered host f first
 * (lvery_first, gpu_astaroth.f90:107,124); early_finalize is accepted for
 * interface parity and IGNORED (the device always completes its halos itself).
 * With ldt=0, *dt (or dt_fixed) must be positive: PCB200_EARG otherwise.
 * ldiagnos!=0 also accumulates the 0-D diagnostics on substep 1.
 * *dt (in/out): on substep 1 with ldt the new time step is returned. */
int pcb200_substep(int isubstep, int full, int early_finalize, int ldiagnos, double *dt);

/* One full time step = 3 x pcb200_substep + optional forcing kick, for
 * callers that do not need per-substep control (bench, tests). */
int pcb200_timestep(int full, int ldiagnos, double *dt);

/* copy_farray_from_GPU (equ.f90:345, snapshot.f90:286; copy_farray_c_ :130):
 * download all mvar variables incl. ghosts into f_host (NULL = the registered
 * array). */
int pcb200_copy_farray(double *f_host);

/* wsnap + output_snap (snapshot.f90:286-302, io_dist.f90:110-167) from the device, without stalling the time loop.
 * The reference copies all of f to the host synchronously (copy_farray_from_GPU, snapshot.f90:286) and then writes
 * record 1 = f(mx,my,mz,mvar) incl. ghost zones, record 2 = t,x(mx),y(my),z(mz),dx,dy,dz.  pcb200_snapshot_begin
 * costs the caller one device-to-device copy of f into a staging array; a worker thread then pulls it through two
 * page-locked chunk buffers (PCB200_SNAP_CHUNK_MB, default 64) on its own stream and writes `path` in exactly that
 * layout (Fortran sequential unformatted, 4-byte markers, gfortran subrecords above 2147483639 bytes), while the
 * caller goes on with pcb200_substep.  One snapshot in flight: the next begin, pcb200_snapshot_wait and
 * pcb200_finalize wait for it.  x,y,z: this rank's mesh incl. ghost points (cdata x,y,z). */
int pcb200_snapshot_begin(const char *path, double t, const double *x, const double *y, const double *z,
                          double dx, double dy, double dz);
int pcb200_snapshot_wait(void);

/* upload f from the host (NULL = registered array) -- `full` path stand-alone */
int pcb200_load_farray(const double *f_host);

/* addforce/forcing_hel (run.f90:729; forcing.f90:1851): coefficients for the
 * kick applied after substep 3 of the current step.  Pass NULL to disable. */
int pcb200_set_forcing(const pcb200_forcing *frc);

/* 1-D z profiles the entropy module precomputes on the host (nz values for
 * n=n1..n2 of this rank): hcond_prof/dlnhcond_prof (get_gravz_heatcond,
 * entropy.f90:7243-7262, used by get_prof_pencil :7379) and the cooling
 * profile prof(z) of get_heat_cool_gravz (entropy.f90:6244-6272). */
enum { PCB200_PROF_HCOND = 0, PCB200_PROF_DLNHCOND = 1, PCB200_PROF_COOL = 2, PCB200_NPROF = 3 };
int pcb200_set_zprofile(int which, const double *prof, int nz);

/* 0-D diagnostics of the last substep-1 call made with ldiagnos!=0.  */
int pcb200_get_diagnostics(double *out, int n);

/* Multi-GPU: NCCL communicator bootstrap.  Rank 0 calls get_unique_id and
 * the host broadcasts the 128 bytes (MPI_Bcast in the Fortran wrapper,
 * torch.distributed in bench.py); then every rank calls comm_init.
 * Replaces mpicomm_init's Cartesian setup (mpicomm.f90:367-605). */
int pcb200_nccl_unique_id(char id128[128]);
int pcb200_comm_init(const char id128[128], int rank, int nranks);

/* finalize_GPU (register.f90:455; finalize_gpu_c_ :62) */
int pcb200_finalize(void);

/* timing helpers on the library's own compute stream (bench.py; the host
 * prints 'Wall clock time/timestep/meshpoint', run.f90:939-941) */
int pcb200_timer_start(void);
int pcb200_timer_stop(float *ms);
int pcb200_sync(void);
/* number of kernels launched by the library since the last call (resets) */
int64_t pcb200_launch_count(void);
/* average duration [ms] and count of the fused RHS+RK kernel launches since
 * the last call (CUDA events on the compute stream; resets) */
int pcb200_rhs_kernel_time(float *avg_ms, int *count);
int pcb200_enable_kernel_timing(int on);

/* ---------------------------------------------------------------------
 * Dust particles (particles_dust with TSC particle-mesh).
 * The reference's GPU slot has no particle entry points (gpu_astaroth.f90:32-71
 * refuses lparticles); these are what gpu_b200.f90 calls from rhs_GPU when
 * lparticles is on.  Layouts are the reference's own (particles_cdata.f90):
 *   fp(mpar_loc,mparray), dfp(mpar_loc,mpvar)  column-major, particle index
 *   fastest; columns ixp,iyp,izp,ivpx,ivpy,ivpz = 1..6;
 *   ipar(mpar_loc) int32; ineargrid(mpar_loc,3) int32 (1-based l,m,n incl. ghosts).
 * Per substep the device does what pde does around the mn-loop:
 *   particles_timestep_first  (particles_main.f90:486)  dfp = alpha*dfp
 *   particles_boundconds      (:625)  wrap -> map_nearest_grid -> stable sort by
 *                                      pencil index imn -> np, TSC rhop
 *   dxxp_dt_pencil/dvvp_dt_pencil (particles_dust.f90:3918,3979): interpolation
 *       of u (quadratic spline), drag, TSC back-reaction into df_u, CFL terms
 *   dxxp_dt/dvvp_dt (:3120,:3199), fold_df (ghostfold_mpicomm.f90:28)
 *   particles_timestep_second (particles_main.f90:507)  fp += dt*beta*dfp      */
typedef struct pcb200_particle_params {
  int32_t abi_version;
  int32_t npar_loc;              /* live particles on this rank */
  int32_t mpar_loc;              /* capacity = leading dimension of fp, dfp, ipar, ineargrid */
  int32_t ldragforce_dust_par;   /* drag on the particles (particles_dust.f90:4264) */
  int32_t ldragforce_gas_par;    /* back-reaction on the gas (:4443-4536) */
  int32_t lcoriolis_force_par;   /* dvvp_dt Coriolis term, uses pcb200_params.Omega (:3230) */
  int32_t ldt_adv_par;           /* particle advection CFL (dxxp_dt_pencil :3936) */
  int32_t lcheck_exact_frontier; /* particles_map.f90:940-943.  The device ALWAYS clamps ineargrid into the interior
                                    (the behaviour of lcheck_exact_frontier=T); with F the reference stops with
                                    fatal_error when round-off puts a particle one cell outside -- the device does
                                    not reproduce that stop */
  int32_t reserved_i[8];
  double tausp;                  /* constant friction time (ldraglaw_epstein, get_frictiontime :5410) */
  double rhop_swarm;             /* mass density of one superparticle in a cell (get_rhopswarm, particles_sub.f90:950) */
  double cdtp, cdtp_drag;        /* particles_dust.f90:59 */
  double reserved_d[8];
} pcb200_particle_params;

/* parameters + device buffers; call after pcb200_initialize. npar_loc=0 switches particles off. */
int pcb200_particles_initialize(const pcb200_particle_params *pp);
/* upload particle state from the host arrays (leading dimension mpar_loc) */
int pcb200_particles_load(const double *fp, const int32_t *ipar);
/* download particle state; any pointer may be NULL. k1_imn/k2_imn: ny*nz entries
 * (1-based first/last particle of each pencil in imn order, 0 if empty, particles_map.f90:1920) */
int pcb200_particles_store(double *fp, double *dfp, int32_t *ipar, int32_t *ineargrid,
                           int32_t *k1_imn, int32_t *k2_imn);
/* npar_loc now (it changes when particles migrate between ranks, particles_mpicomm.f90:267) */
int pcb200_particles_count(int32_t *npar_loc);
/* aux fields of map_xxp_grid on the full (mx,my,mz) grid: np and rhop */
int pcb200_particles_get_aux(double *np, double *rhop);

/* name of the fused RHS+RK kernel of the last substep launch ("rhs_dly", "rhs_tma", "rhs_march",
 * "rhs_global"; "" before the first launch).  Selection: PCB200_KERNEL={dly,tma,cpasync,global};
 * default rhs_dly where it applies (lbidiagonal_derij=T, not magnetic+entropy, no full-diagnostics substep). */
const char *pcb200_kernel_name(void);
const char *pcb200_last_error(void);
int pcb200_abi_version(void);

/* No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * CFL time step of a device-resident state -- the same cfl_prepass kernel that substep 1 of an
 * isothermal run launches -- with the per-block partial maxima in a file-scope device array instead
 * of allocated scratch.  Calls no cudaMalloc/cudaFree or allocating library routine.
 *   f_dev       device pointer (16-byte aligned) to f(mx,my,mz,mvar) of the grid given to
 *               pcb200_initialize, ghost zones already valid; read only
 *   dt_out_dev  device pointer to two doubles: [0] = dt, [1] = 1/dt = max(dt1_max, 1/dtmax) of THIS
 *               rank (no cross-rank maximum, no particle terms)
 * Asynchronous on the library's compute stream (pcb200_sync() waits).  Limit: at most
 * PCB200_NOALLOC_MAX_BLOCKS blocks of 32x8x64 cells (512^3 cells per rank; the array is 1 MiB per GPU,
 * held until exit); larger grids and runs with the entropy module are refused with PCB200_EARG /
 * PCB200_EUNSUPP.  All calls share the array: calls that overlap on the GPU race. */
#define PCB200_NOALLOC_MAX_BLOCKS 8192
int pcb200_cfl_dt_noalloc(const double *f_dev, double *dt_out_dev);

#ifdef __cplusplus
}
#endif
#endif /* PCB200_H */
