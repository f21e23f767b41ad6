/* abi_driver.c -- replays the call sequence of the Fortran host through the C ABI, with no Python in between:
 *   pcb200_init -> pcb200_initialize -> pcb200_register -> [pcb200_substep x 3] x N -> pcb200_copy_farray -> pcb200_finalize
 * i.e. gpu_init (run.f90:125), initialize_GPU (register.f90:375), register_GPU (run.f90:223), rhs_GPU per substep
 * (equ.f90:343, lvery_first -> full upload), copy_farray_from_GPU (snapshot.f90:286), finalize_GPU (register.f90:455).
 * The library is dlopen'ed, so this file builds with plain gcc (no CUDA toolkit):
 *   gcc -O1 -I include tests/abi_driver.c -o abi_driver -ldl -lm
 *   ./abi_driver pencil-code_b200/libpcb200.so 32 4 out.bin
 * Set-up: periodic isothermal MHD n^3 on [-pi,pi)^3, the parameters of samples/helical-MHDturb/run.in
 * (cdt=0.4, cdtv=0.3, nu=eta=5e-3), an analytic smooth initial field evaluated on the ghost-padded grid.
 * out.bin: int32 n, nsteps, mvar | double dt[nsteps] | double f0[mvar*m^3] | double f[mvar*m^3]  (m = n+6).
 * tests/test_gpu_parity.py::test_c_driver_replays_host_sequence runs it and replays the same input through the
 * Python mirror of the Fortran module: results must be bit-identical. */
#include <dlfcn.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "pcb200.h"

#define SYM(T, name) T name = (T)dlsym(h, #name); if (!name) { fprintf(stderr, "missing symbol %s\n", #name); return 2; }
#define CK(call) do { int rc_ = (call); if (rc_) { fprintf(stderr, "%s -> %d: %s\n", #call, rc_, pcb200_last_error()); return 3; } } while (0)

typedef int (*init_t)(int);
typedef int (*initialize_t)(const pcb200_params *);
typedef int (*reg_t)(double *);
typedef int (*substep_t)(int, int, int, int, double *);
typedef int (*copy_t)(double *);
typedef int (*fin_t)(void);
typedef const char *(*err_t)(void);

int main(int argc, char **argv)
{
  if (argc < 5) { fprintf(stderr, "usage: %s <libpcb200.so> <n> <nsteps> <out.bin>\n", argv[0]); return 1; }
  void *h = dlopen(argv[1], RTLD_NOW | RTLD_GLOBAL);
  if (!h) { fprintf(stderr, "dlopen: %s\n", dlerror()); return 2; }
  const int n = atoi(argv[2]), nsteps = atoi(argv[3]);
  SYM(init_t, pcb200_init) SYM(initialize_t, pcb200_initialize) SYM(reg_t, pcb200_register)
  SYM(substep_t, pcb200_substep) SYM(copy_t, pcb200_copy_farray) SYM(fin_t, pcb200_finalize) SYM(err_t, pcb200_last_error)

  const double pi = 3.14159265358979323846;
  pcb200_params P;
  memset(&P, 0, sizeof(P));
  P.abi_version = PCB200_ABI_VERSION;
  P.nx = P.ny = P.nz = n;
  P.nprocx = P.nprocy = P.nprocz = 1;
  P.lperi[0] = P.lperi[1] = P.lperi[2] = 1;
  P.mvar = P.mfarray = 7;
  P.iux = 1; P.iuy = 2; P.iuz = 3; P.ilnrho = 4; P.iax = 5; P.iay = 6; P.iaz = 7;     /* register.f90:124-135 */
  P.lhydro = P.ldensity = P.lmagnetic = 1;
  P.lbidiagonal_derij = 1; P.lviscosity_heat = 1; P.lohmic_heat = 1; P.ldt = 1;
  /* bcx/bcy/bcz: all 'p' = PCB200_BC_P = 0 (memset) */
  const double L = 2 * pi, dx = L / n;
  P.dx = P.dy = P.dz = dx;
  P.dx_1 = P.dy_1 = P.dz_1 = 1.0 / dx;
  for (int i = 0; i < 3; i++) { P.xyz0[i] = -pi; P.Lxyz[i] = L; }
  P.x0_loc = P.y0_loc = P.z0_loc = -pi + 0.5 * dx;                                      /* grid.f90:118-222, periodic */
  for (int i = 1; i <= 3; i++) { P.dz2_bound[3 - i] = 2.0 * i * dx; P.dz2_bound[3 + i] = 2.0 * i * dx; }
  P.cdt = 0.4; P.cdtv = 0.3; P.cdts = 1.0; P.dtmax = 1e37;
  P.alpha_ts[0] = 0.0; P.alpha_ts[1] = -5.0 / 9.0; P.alpha_ts[2] = -153.0 / 128.0;      /* timestep.f90:40-41 */
  P.beta_ts[0] = 1.0 / 3.0; P.beta_ts[1] = 15.0 / 16.0; P.beta_ts[2] = 8.0 / 15.0;
  P.nu = 5e-3; P.eta = 5e-3; P.mu0 = 1.0;
  P.cs20 = 1.0; P.gamma = 1.0; P.cp = 1.0; P.lnrho0 = 0.0;
  P.ztop = -pi + L;

  const int m = n + 6;
  const size_t mw = (size_t)m * m * m, ne = mw * 7;
  double *f = (double *)malloc(ne * sizeof(double)), *f0 = (double *)malloc(ne * sizeof(double));
  double *dts = (double *)calloc(nsteps > 0 ? nsteps : 1, sizeof(double));
  for (int k = 0; k < 7; k++)
    for (int iz = 0; iz < m; iz++)
      for (int iy = 0; iy < m; iy++)
        for (int ix = 0; ix < m; ix++) {
          const double x = P.x0_loc + (ix - 3) * dx, y = P.y0_loc + (iy - 3) * dx, z = P.z0_loc + (iz - 3) * dx;
          double v;
          switch (k) {                       /* smooth, periodic on the box, all derivatives non-trivial */
            case 0: v = 0.10 * sin(y + 0.3) * cos(2 * z) + 0.02 * sin(3 * x + y); break;
            case 1: v = 0.08 * sin(z - 0.2) * cos(x) + 0.03 * cos(2 * y - z); break;
            case 2: v = 0.09 * sin(x + 0.7) * cos(2 * y) + 0.02 * sin(x - 2 * z); break;
            case 3: v = 0.05 * cos(x + y) * sin(z) + 0.02 * cos(2 * x); break;
            case 4: v = 0.010 * sin(2 * y) * cos(z + 0.1) + 0.004 * sin(x + 3 * z); break;
            case 5: v = 0.012 * cos(2 * z) * sin(x - 0.4) + 0.003 * cos(3 * y + x); break;
            default: v = 0.011 * sin(2 * x) * cos(y + 0.5) + 0.004 * sin(2 * y - z); break;
          }
          f[(size_t)k * mw + ((size_t)iz * m + iy) * m + ix] = v;
        }
  memcpy(f0, f, ne * sizeof(double));

  CK(pcb200_init(0));
  CK(pcb200_initialize(&P));
  CK(pcb200_register(f));
  double dt = 0.0;
  int lvery_first = 1;                                   /* gpu_astaroth.f90:107: first call uploads the registered f */
  for (int it = 0; it < nsteps; it++) {
    for (int isub = 1; isub <= 3; isub++) {
      CK(pcb200_substep(isub, lvery_first, 0, 0, &dt));
      lvery_first = 0;
    }
    dts[it] = dt;
    printf("it=%d dt=%.17g\n", it, dt);
  }
  CK(pcb200_copy_farray(f));
  CK(pcb200_finalize());

  FILE *fp = fopen(argv[4], "wb");
  if (!fp) { perror("fopen"); return 4; }
  int32_t hdr[3] = {n, nsteps, 7};
  fwrite(hdr, sizeof(hdr), 1, fp);
  fwrite(dts, sizeof(double), nsteps, fp);
  fwrite(f0, sizeof(double), ne, fp);
  fwrite(f, sizeof(double), ne, fp);
  fclose(fp);
  free(f); free(f0); free(dts);
  return 0;
}
