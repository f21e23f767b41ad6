// devpar.h -- device-side parameter block (constant memory) and field roles.
#pragma once
#include <stdint.h>
#include "../../include/pcb200.h"

#define NG PCB200_NGHOST

// Field roles: compile-time identity of a variable inside the kernels.  The
// storage slot (position in f(mx,my,mz,mvar), i.e. iux-1 ...) is run time.
enum Role { R_UX = 0, R_UY = 1, R_UZ = 2, R_AX = 3, R_AY = 4, R_AZ = 5, R_LNRHO = 6, R_SS = 7, NROLE = 8 };

struct DevPar {
  int nx, ny, nz, mx, my, mz;
  long long sy, sz, mw;          // strides (elements): y, z, variable
  int mvar;
  int slot[NROLE];               // role -> storage slot, -1 = absent
  long long foff[NROLE];         // slot*mw: element offset of the role's variable
  long long foff8[NROLE];        // the same in bytes
  int lupw_uu, lupw_lnrho, lupw_ss;
  int lbidiag;                 // lbidiagonal_derij (cdata.f90:568)
  int iheatcond, lcooling, lviscosity_heat, lohmic_heat, lgravz, ldensity, lhydro;
  double dx_1, dy_1, dz_1;       // 0 for degenerate directions
  double dxyz_2;                 // grid.f90:2567-2570
  double cdt, cdtv, cdts;
  double nu, eta, mu0, mu01;
  double cs20, gamma, gamma1, gamma_m1, cp, cp1, cv, cv1, lnrho0, lnTT0;
  double Omega2;                 // 2*Omega
  double gravz;
  double chi, cool, cs2cool;
  // derived stencil factors
  double fx1, fy1, fz1;          // dx_1/60 ...
  double fx2, fy2, fz2;          // dx_1^2 ...
  double fxy, fyz, fzx;          // dx_1*dy_1/720 ...
  // pre-scaled stencil coefficients (constant-bank operands: one FMA per tap, no separate scaling)
  double k1x[3], k1y[3], k1z[3]; // f1*{45,-9,1}
  double k2x[4], k2y[4], k2z[4]; // f2*{-490,270,-27,2}/180
  double k2yz0, k2zx0, k2xy0;    // centre coefficients of the pair sums d2y+d2z, d2z+d2x, d2x+d2y
  double k2xyz0;                 // ... and of the full Laplacian
  double kxy[3], kyz[3], kzx[3]; // fij*{270,-27,2}
  double dlyu[3][2], dlya[3][2]; // rhs_dly history scales: nu/3*c_k*{fzx,fyz} and mu01*c_k*{fzx,fyz}
  const double *zprof[PCB200_NPROF];   // device pointers, nz values each
};
