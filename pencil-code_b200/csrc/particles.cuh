// particles.cuh -- dust particles on the device.
//
// What pde does around the mn-loop when lparticles is on, one kernel per stage:
//   part_first        particles_timestep_first   (particles_main.f90:486-498)
//   part_bc_map       boundconds_particles 'p'   (particles_sub.f90:170-215) + map_nearest_grid
//                     (particles_map.f90:895-984; inverse_grid, grid.f90:2310-2343) + sort key
//   [stable radix sort of (key, index) by the pencil index imn, particles_map.f90:986-1163]
//   part_gather       fp, dfp, ipar, ineargrid permuted into pencil order
//   part_pencil_bounds  k1_imn, k2_imn             (particles_map.f90:1920-1948)
//   part_cellkey      cell id of every particle + count per cell (= np, map_xxp_grid :1322-1327)
//   [second stable sort by cell id -> cell bins; exclusive scan of the counts -> bin offsets]
//   part_drag         dxxp_dt_pencil/dvvp_dt_pencil (particles_dust.f90:3918-4911): spline
//                     interpolation of u, drag, CFL; dxxp_dt, dvvp_dt (:3120,:3199)
//   part_dt_cells     dt1_drag per cell             (particles_dust.f90:4843-4910)
//   part_deposit      TSC back-reaction (:4443-4536) and TSC rhop (particles_map.f90:1560-1640)
//                     as a GATHER: one thread per grid cell walks the particles of its 27
//                     neighbour bins.  No atomics, fixed summation order => deterministic.
//                     The result joins the 2N register / new state of u (equ.f90:1046).
//   part_second       particles_timestep_second  (particles_main.f90:521)
// Layout = the reference's: fp[6][mpar] (particle index fastest), ineargrid[3][mpar] 1-based.
// The gather reads a neighbour bin through the periodic wrap, which is fold_df/fold_f
// (ghostfold_mpicomm.f90:46-124) without the intermediate ghost layer.
#pragma once

struct PartPar {
  int npar, mpar;
  int ldrag_dust, ldrag_gas, lcoriolis, ldt_adv;
  int wrap[3];                 // direction wraps inside this rank (fold == periodic index wrap)
  int lperi[3], shift[3];      // shift = n*ip (grid.f90:2291-2300)
  double xyz0[3], Lxyz[3], h[3], h_1[3], x0_loc[3];
  double tausp1, rhop_swarm, cdtp, cdtp_drag, Omega2;
  int iux, ilnrho;             // storage slots (0-based)
};

__global__ void part_first(double *__restrict__ dfp, int npar, int mpar, double alpha)
{
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= npar) return;
#pragma unroll
  for (int j = 0; j < 6; j++) dfp[(size_t)j * mpar + k] = alpha == 0.0 ? 0.0 : alpha * dfp[(size_t)j * mpar + k];
}

// NINT = round half away from zero = round() (not rint)
__device__ __forceinline__ int part_nearest(const PartPar &q, int ax, int nloc, double x)
{
  double xi = (x - q.xyz0[ax]) / q.h[ax];
  xi = q.lperi[ax] ? xi + (double)NG + 0.5 : xi + (double)(NG + 1);
  xi = xi - (double)q.shift[ax];
  const int inear = (int)round(xi);
  const double xyz1 = q.xyz0[ax] + q.Lxyz[ax];
  if ((x < xyz1 && inear > NG + nloc) || (x < q.xyz0[ax] && inear > NG)) xi = ::nextafter(xi, -1.0e308);
  int i = (int)round(xi);
  return min(max(i, NG + 1), NG + nloc);           // lcheck_exact_frontier (particles_map.f90:940-943)
}

__global__ void part_bc_map(PartPar q, double *__restrict__ fp, int *__restrict__ inear, const int *__restrict__ imn_array,
                            unsigned *__restrict__ keys, unsigned *__restrict__ vals)
{
  const DevPar &c = c_par;
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= q.npar) return;
  const int nloc[3] = {c.nx, c.ny, c.nz};
  int idx[3];
#pragma unroll
  for (int ax = 0; ax < 3; ax++) {
    double x = fp[(size_t)ax * q.mpar + k];
    // bcpx=bcpy=bcpz='p' (the particle boundary condition, independent of the gas one)
    if (x < q.xyz0[ax]) x = x + q.Lxyz[ax];
    if (x >= q.xyz0[ax] + q.Lxyz[ax]) x = x - q.Lxyz[ax];
    fp[(size_t)ax * q.mpar + k] = x;
    idx[ax] = part_nearest(q, ax, nloc[ax], x);
    inear[(size_t)ax * q.mpar + k] = idx[ax];
  }
  keys[k] = (unsigned)(imn_array[(idx[2] - NG - 1) * c.ny + (idx[1] - NG - 1)] - 1);
  vals[k] = (unsigned)k;
}

__global__ void part_gather(int npar, int mpar, const unsigned *__restrict__ perm,
                            const double *__restrict__ fp_in, double *__restrict__ fp_out,
                            const double *__restrict__ dfp_in, double *__restrict__ dfp_out,
                            const int *__restrict__ ipar_in, int *__restrict__ ipar_out,
                            const int *__restrict__ in_in, int *__restrict__ in_out)
{
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= npar) return;
  const unsigned s = perm[k];
#pragma unroll
  for (int j = 0; j < 6; j++) {
    fp_out[(size_t)j * mpar + k] = fp_in[(size_t)j * mpar + s];
    dfp_out[(size_t)j * mpar + k] = dfp_in[(size_t)j * mpar + s];
  }
#pragma unroll
  for (int j = 0; j < 3; j++) in_out[(size_t)j * mpar + k] = in_in[(size_t)j * mpar + s];
  ipar_out[k] = ipar_in[s];
}

__global__ void part_pencil_bounds(int npar, const unsigned *__restrict__ keys, int *__restrict__ k1, int *__restrict__ k2)
{
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= npar) return;
  const unsigned key = keys[k];
  if (k == 0 || keys[k - 1] != key) k1[key] = k + 1;
  if (k == npar - 1 || keys[k + 1] != key) k2[key] = k + 1;
}

// TSC weight of grid point ixx (0-based) for a particle at xp with nearest point ix0
// (particles_dust.f90:4467-4490)
__device__ __forceinline__ double part_tsc(const PartPar &q, int ax, double xp, int ixx, int ix0)
{
  const double xg = q.x0_loc[ax] + (ixx - NG) * q.h[ax];
  if (ixx != ix0) { const double a = fabs(xp - xg) * q.h_1[ax]; return 1.125 - 1.5 * a + 0.5 * (a * a); }
  const double b = (xp - xg) * q.h_1[ax];
  return 0.75 - b * b;
}
// interior-relative cell id of the nearest grid point; ccount = particles per cell (exact)
__global__ void part_cellkey(int npar, int mpar, const int *__restrict__ inear, unsigned *__restrict__ ckeys,
                             unsigned *__restrict__ cvals, int *__restrict__ ccount)
{
  const DevPar &c = c_par;
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= npar) return;
  const int l = inear[k] - 1 - NG, m = inear[(size_t)mpar + k] - 1 - NG, n = inear[(size_t)2 * mpar + k] - 1 - NG;
  const unsigned id = (unsigned)((n * c.ny + m) * c.nx + l);
  ckeys[k] = id; cvals[k] = (unsigned)k;
  atomicAdd(&ccount[id], 1);
}

// rank[k] = position of particle k in the cell-sorted list
__global__ void part_invperm(int npar, const unsigned *__restrict__ perm, unsigned *__restrict__ rank)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < npar) rank[perm[i]] = (unsigned)i;
}

__device__ __forceinline__ void atomic_max_nonneg(double *addr, double v)   // exact for v >= 0
{ atomicMax(reinterpret_cast<unsigned long long *>(addr), (unsigned long long)__double_as_longlong(v)); }

// One thread per particle.  f = state of this substep with valid ghost zones; csort[6][mpar]
// receives position and drag force in CELL-SORTED order (slot rank[k]) for part_deposit;
// dt1 (one double) takes the advective particle CFL term when ldt.
__global__ void __launch_bounds__(128) part_drag(PartPar q, const double *__restrict__ f, const double *__restrict__ fp,
                                                  double *__restrict__ dfp, const int *__restrict__ inear,
                                                  const unsigned *__restrict__ rank, double *__restrict__ csort,
                                                  double *__restrict__ dt1, int ldt)
{
  const DevPar &c = c_par;
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  double v_adv = 0.0;
  if (k < q.npar) {
    const size_t M = q.mpar;
    const int i0[3] = {inear[k] - 1, inear[M + k] - 1, inear[2 * M + k] - 1};
    const double xp[3] = {fp[k], fp[M + k], fp[2 * M + k]};
    const double vp[3] = {fp[3 * M + k], fp[4 * M + k], fp[5 * M + k]};
    if (ldt && q.ldt_adv) {                                         // dxxp_dt_pencil :3936-3968
      const double ax = fabs(vp[0]) * q.h_1[0], ay = fabs(vp[1]) * q.h_1[1], az = fabs(vp[2]) * q.h_1[2];
      v_adv = sqrt(ax * ax + ay * ay + az * az) / q.cdtp;
    }
    double dvp[3] = {0.0, 0.0, 0.0};
    if (q.ldrag_dust) {
      // quadratic spline interpolation of u (particles_map.f90:686-722)
      double sx[3], sy[3], sz[3];
      {
        const double d0 = (xp[0] - (q.x0_loc[0] + (i0[0] - NG) * q.h[0])) * q.h_1[0];
        const double d1 = (xp[1] - (q.x0_loc[1] + (i0[1] - NG) * q.h[1])) * q.h_1[1];
        const double d2 = (xp[2] - (q.x0_loc[2] + (i0[2] - NG) * q.h[2])) * q.h_1[2];
        sx[0] = 0.5 * ((0.5 - d0) * (0.5 - d0)); sx[1] = 0.75 - d0 * d0; sx[2] = 0.5 * ((0.5 + d0) * (0.5 + d0));
        sy[0] = 0.5 * ((0.5 - d1) * (0.5 - d1)); sy[1] = 0.75 - d1 * d1; sy[2] = 0.5 * ((0.5 + d1) * (0.5 + d1));
        sz[0] = 0.5 * ((0.5 - d2) * (0.5 - d2)); sz[1] = 0.75 - d2 * d2; sz[2] = 0.5 * ((0.5 + d2) * (0.5 + d2));
      }
      const long long c0 = (long long)i0[2] * c.sz + (long long)i0[1] * c.sy + i0[0];
      double uup[3] = {0.0, 0.0, 0.0};
#pragma unroll
      for (int dz = 0; dz < 3; dz++)
#pragma unroll
        for (int dy = 0; dy < 3; dy++) {
          const double wyz = sy[dy] * sz[dz];
          const long long o = c0 + (long long)(dz - 1) * c.sz + (long long)(dy - 1) * c.sy;
#pragma unroll
          for (int dx = 0; dx < 3; dx++) {
            const double w = sx[dx] * wyz;
#pragma unroll
            for (int j = 0; j < 3; j++) uup[j] += w * __ldg(f + (long long)(q.iux + j) * c.mw + o + dx - 1);
          }
        }
      double drag[3];
#pragma unroll
      for (int j = 0; j < 3; j++) { drag[j] = -q.tausp1 * (vp[j] - uup[j]); dvp[j] = drag[j]; }   // :4264
      if (q.ldrag_gas) {                                            // consumed by part_deposit
        const unsigned r = rank[k];
#pragma unroll
        for (int j = 0; j < 3; j++) csort[(size_t)(3 + j) * M + r] = drag[j];
      }
    }
    {
      const unsigned r = rank[k];
#pragma unroll
      for (int j = 0; j < 3; j++) csort[(size_t)j * M + r] = xp[j];
    }
    if (q.lcoriolis && q.Omega2 != 0.0) { dvp[0] += q.Omega2 * vp[1]; dvp[1] -= q.Omega2 * vp[0]; }   // dvvp_dt :3226-3233
#pragma unroll
    for (int j = 0; j < 3; j++) {
      dfp[(size_t)j * M + k] += vp[j];                              // dxxp_dt :3155-3161
      dfp[(size_t)(3 + j) * M + k] += dvp[j];
    }
  }
  if (ldt && q.ldt_adv) {
    for (int o = 16; o > 0; o >>= 1) v_adv = fmax(v_adv, __shfl_xor_sync(0xffffffffu, v_adv, o));
    if ((threadIdx.x & 31) == 0 && v_adv > 0.0) atomic_max_nonneg(dt1, v_adv);
  }
}

// dt1_drag = (max_k tausp1 + sum_k mp_vcell*rho1*tausp1)/cdtp_drag per cell with particles
// (particles_dust.f90:4843-4847, 4901-4910); constant tausp => the sum is np * (...)
__global__ void part_dt_cells(PartPar q, const double *__restrict__ f, const int *__restrict__ ccount, double *__restrict__ dt1)
{
  const DevPar &c = c_par;
  const int l = blockIdx.x * blockDim.x + threadIdx.x, m = blockIdx.y, n = blockIdx.z;
  double d = 0.0;
  if (l < c.nx) {
    const int npc = ccount[(n * c.ny + m) * c.nx + l];
    if (npc > 0) {
      const long long i = (long long)(n + NG) * c.sz + (long long)(m + NG) * c.sy + (l + NG);
      d = q.tausp1;
      if (q.ldrag_gas) d += npc * (q.rhop_swarm * exp(-f[(long long)q.ilnrho * c.mw + i]) * q.tausp1);
      d /= q.cdtp_drag;
    }
  }
  for (int o = 16; o > 0; o >>= 1) d = fmax(d, __shfl_xor_sync(0xffffffffu, d, o));
  if ((threadIdx.x & 31) == 0 && d > 0.0) atomic_max_nonneg(dt1, d);
}

// One thread per interior cell (l,m,n): gathers from the particles whose nearest grid point is
// one of the 27 neighbours.  csort[6][mpar] = x,y,z,drag of the particles in cell-sorted order,
// cs = exclusive scan of the per-cell counts, so the three x-neighbour bins of one (y,z) row
// are ONE contiguous range (plus one wrapped bin at the x edges).  The TSC weight of this grid
// point for a particle at distance a (in cells) is 0.75-a^2 if it is the particle's nearest
// point (|a|<=1/2) and 1.125-1.5|a|+a^2/2 otherwise (particles_dust.f90:4467-4490; the two
// forms agree at |a|=1/2).  In y,z (and the wrapped x bin) the distance is taken to the image
// of this grid point next to the bin, i.e. through the periodic fold (ghostfold_mpicomm.f90:46-124).
//   rhop[c]  = sum of weights                          (particles_map.f90:1631, times rhop_swarm on output)
//   dfu_j    = -rhop_swarm/rho(c) * sum_k w_k drag_kj   (particles_dust.f90:4527)
//   w_u += dfu ; f_out_u += beta*dt*dfu                (equ.f90:1046 + timestep.f90:162; f_out may be null)
__device__ __forceinline__ double part_w(double a)
{ const double b = fabs(a); return b <= 0.5 ? 0.75 - a * a : 1.125 - 1.5 * b + 0.5 * (a * a); }

// GHOST=false: single rank, every direction folds locally; result fused into w / f_out.
// GHOST=true : some directions are split between ranks; along those the bins of the neighbour
//   rank are not here, so (as the reference does) the first GHOST LAYER is filled as well and
//   fold_df/fold_f ships it to the neighbour (ghostfold_mpicomm.f90:46-153).  Threads cover the
//   interior plus that layer; acc[0] = sum of weights, acc[1..3] = sum of weight*drag (padded grid).
template <bool GHOST>
__global__ void __launch_bounds__(128) part_deposit(PartPar q, const double *__restrict__ f, const double *__restrict__ csort,
                                                     const int *__restrict__ cs, double *__restrict__ rhop,
                                                     double *__restrict__ w, double *__restrict__ f_out,
                                                     const double *__restrict__ dt_dev, double beta, double *__restrict__ acc)
{
  const DevPar &c = c_par;
  const int e = GHOST ? 1 : 0;
  const int l = blockIdx.x * blockDim.x + threadIdx.x - e, m = (int)blockIdx.y - e, n = (int)blockIdx.z - e;
  if (l >= c.nx + e) return;
  if (GHOST && ((q.wrap[0] && (l < 0 || l >= c.nx)) || (q.wrap[1] && (m < 0 || m >= c.ny)) || (q.wrap[2] && (n < 0 || n >= c.nz)))) return;
  const size_t M = q.mpar;
  const double xg = q.x0_loc[0] + l * q.h[0];
  // x bins of the main range and of the wrapped bin (if x folds locally and this cell sits at an x edge)
  const int la = l > 0 ? l - 1 : 0, lb = l < c.nx - 1 ? l + 1 : c.nx - 1;
  const int lw = (GHOST && !q.wrap[0]) ? -1 : (l == 0 ? c.nx - 1 : (l == c.nx - 1 ? 0 : -1));
  const double xgw = l == 0 ? xg + q.Lxyz[0] : xg - q.Lxyz[0];      // image of this point seen from the wrapped bin
  int r0[9], r1[9], w0[9], w1[9];
#pragma unroll
  for (int j = 0; j < 9; j++) {
    const int oz = j / 3 - 1, oy = j % 3 - 1;
    int nn = n + oz, mm = m + oy;
    bool ok = true;
    if (nn < 0 || nn >= c.nz) { if (GHOST && !q.wrap[2]) ok = false; else nn = nn < 0 ? nn + c.nz : nn - c.nz; }
    if (mm < 0 || mm >= c.ny) { if (GHOST && !q.wrap[1]) ok = false; else mm = mm < 0 ? mm + c.ny : mm - c.ny; }
    const int row = ok ? (nn * c.ny + mm) * c.nx : 0;
    r0[j] = ok ? __ldg(cs + row + la) : 0; r1[j] = ok ? __ldg(cs + row + lb + 1) : 0;
    w0[j] = ok && lw >= 0 ? __ldg(cs + row + lw) : 0; w1[j] = ok && lw >= 0 ? __ldg(cs + row + lw + 1) : 0;
  }
  double v0 = 0.0, v1 = 0.0, v2 = 0.0, ws = 0.0;
#pragma unroll
  for (int j = 0; j < 9; j++) {
    const int oz = j / 3 - 1, oy = j % 3 - 1;
    int nn = n + oz; nn = nn < 0 ? nn + c.nz : (nn >= c.nz ? nn - c.nz : nn);
    int mm = m + oy; mm = mm < 0 ? mm + c.ny : (mm >= c.ny ? mm - c.ny : mm);
    const double zg = q.x0_loc[2] + (nn - oz) * q.h[2], yg = q.x0_loc[1] + (mm - oy) * q.h[1];
#pragma unroll 1
    for (int pass = 0; pass < 2; pass++) {
      const int i0 = pass ? w0[j] : r0[j], i1 = pass ? w1[j] : r1[j];
      const double xgi = pass ? xgw : xg;
      for (int i = i0; i < i1; i++) {
        const double wt = (part_w((__ldg(csort + i) - xgi) * q.h_1[0]) * part_w((__ldg(csort + M + i) - yg) * q.h_1[1]))
                          * part_w((__ldg(csort + 2 * M + i) - zg) * q.h_1[2]);
        ws += wt;
        if (q.ldrag_gas) { v0 += wt * __ldg(csort + 3 * M + i); v1 += wt * __ldg(csort + 4 * M + i); v2 += wt * __ldg(csort + 5 * M + i); }
      }
    }
  }
  const long long i = (long long)(n + NG) * c.sz + (long long)(m + NG) * c.sy + (l + NG);
  if (GHOST) { acc[i] = ws; acc[c.mw + i] = v0; acc[2 * c.mw + i] = v1; acc[3 * c.mw + i] = v2; return; }
  rhop[i] = ws;
  if (q.ldrag_gas) {
    const double fac = -q.rhop_swarm * exp(-f[(long long)q.ilnrho * c.mw + i]);
    const double bdt = f_out ? beta * dt_dev[0] : 0.0;
    const double d[3] = {fac * v0, fac * v1, fac * v2};
#pragma unroll
    for (int j = 0; j < 3; j++) {
      const long long o = (long long)(q.iux + j) * c.mw + i;
      w[o] += d[j];
      if (f_out) f_out[o] += bdt * d[j];
    }
  }
}

// after the fold: rhop and the back-reaction from the accumulated sums (interior cells)
__global__ void part_backreact_acc(PartPar q, const double *__restrict__ f, const double *__restrict__ acc, double *__restrict__ rhop,
                                   double *__restrict__ w, double *__restrict__ f_out, const double *__restrict__ dt_dev, double beta)
{
  const DevPar &c = c_par;
  const int l = blockIdx.x * blockDim.x + threadIdx.x, m = blockIdx.y, n = blockIdx.z;
  if (l >= c.nx) return;
  const long long i = (long long)(n + NG) * c.sz + (long long)(m + NG) * c.sy + (l + NG);
  rhop[i] = acc[i];
  if (!q.ldrag_gas) return;
  const double fac = -q.rhop_swarm * exp(-f[(long long)q.ilnrho * c.mw + i]);
  const double bdt = f_out ? beta * dt_dev[0] : 0.0;
#pragma unroll
  for (int j = 0; j < 3; j++) {
    const double d = fac * acc[(long long)(1 + j) * c.mw + i];
    const long long o = (long long)(q.iux + j) * c.mw + i;
    w[o] += d;
    if (f_out) f_out[o] += bdt * d;
  }
}

__global__ void part_second(double *__restrict__ fp, const double *__restrict__ dfp, int npar, int mpar,
                            const double *__restrict__ dt_dev, double beta)
{
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= npar) return;
  const double dtb = dt_dev[0] * beta;
#pragma unroll
  for (int j = 0; j < 6; j++) fp[(size_t)j * mpar + k] = fp[(size_t)j * mpar + k] + dtb * dfp[(size_t)j * mpar + k];
}

// aux fields for the host on the padded grid: np (count at the nearest point) and
// rhop = rhop_swarm * sum of TSC weights (particles_map.f90:1733-1737)
__global__ void part_aux_out(const int *__restrict__ ccount, const double *__restrict__ rhop, double rhop_swarm,
                             double *__restrict__ np_out, double *__restrict__ rhop_out)
{
  const DevPar &c = c_par;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= c.mw) return;
  const int l = (int)(i % c.mx) - NG, m = (int)((i / c.mx) % c.my) - NG, n = (int)(i / c.sz) - NG;
  const bool in = l >= 0 && l < c.nx && m >= 0 && m < c.ny && n >= 0 && n < c.nz;
  np_out[i] = in ? (double)ccount[(n * c.ny + m) * c.nx + l] : 0.0;
  rhop_out[i] = in ? rhop_swarm * rhop[i] : 0.0;
}

// ------------------------------------------------------------------------
// migrate_particles (particles_mpicomm.f90:267-696), device version.
// Same bookkeeping order as the reference so that the per-rank storage order evolves the same
// way: particles are scanned from the LAST to the first; a leaver's slot is refilled with the
// current last particle (fp(k,:)=fp(npar_loc,:)); leavers are grouped by receiving rank in scan
// order; arrivals are appended sender by sender in ascending rank order.
// The sequential refill is equivalent to: with N' = N - nleave, the i-th largest hole below N'
// receives the i-th largest survivor at or above N'.
struct MigPar {
  int np[3], ip[3], nloc[3];
  int nslots, cap;            // distinct neighbour ranks; capacity of one message (particles)
  int iproc;
};
enum { MIG_STAY = 31, MIG_NVAL = 13 };   // payload: fp(6), dfp(6), ipar

// counters[0..26] leavers per slot, [27] nleave, [28] N', [29] error flags, [30] new npar, [32..58] slot offsets
__global__ void part_mig_classify(PartPar q, MigPar g, const double *__restrict__ fp, const int *__restrict__ slot_of_rank,
                                  unsigned *__restrict__ rkeys, unsigned *__restrict__ rvals, int *__restrict__ leave,
                                  int *__restrict__ counters)
{
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= q.npar) return;
  int rec[3];
#pragma unroll
  for (int a = 0; a < 3; a++) {
    const double x = fp[(size_t)a * q.mpar + k], xx = x - q.xyz0[a];
    const double w = q.Lxyz[a] / g.np[a];                    // procx_bounds(j) = xyz0 + j*Lx/nprocx (grid.f90:835, linear grid)
    const double x0l = g.ip[a] * w, x1l = (g.ip[a] + 1) * w; // xx0_loc, xx1_loc (:285-286)
    int r = g.ip[a];
    if (xx >= x1l && g.ip[a] < g.np[a] - 1) {
      for (int j = g.ip[a] + 1; j <= g.np[a] - 1; j++) if (x < q.xyz0[a] + (j + 1) * w) { r = j; break; }
    } else if (xx < x0l && g.ip[a] > 0) {
      for (int j = g.ip[a] - 1; j >= 0; j--) if (x >= q.xyz0[a] + j * w) { r = j; break; }
    }
    rec[a] = r;
  }
  const int iproc_rec = rec[0] + g.np[0] * (rec[1] + g.np[1] * rec[2]);
  unsigned key = MIG_STAY;
  int lv = 0;
  if (iproc_rec != g.iproc) {
    const int s = slot_of_rank[iproc_rec];
    if (s < 0) atomicOr(&counters[29], 1);                   // "wants to migrate to a processor that is not a neighbour" (:399)
    else { key = (unsigned)s; lv = 1; atomicAdd(&counters[s], 1); }
  }
  leave[k] = lv;
  const int r = q.npar - 1 - k;                              // scan order of the reference: k = npar_loc ... 1
  rkeys[r] = key; rvals[r] = (unsigned)k;
}

__global__ void part_mig_offsets(MigPar g, int npar, int *__restrict__ counters)
{
  if (threadIdx.x || blockIdx.x) return;
  int tot = 0;
  for (int s = 0; s < g.nslots; s++) { counters[32 + s] = tot; tot += counters[s]; if (counters[s] > g.cap) atomicOr(&counters[29], 2); }
  counters[27] = tot; counters[28] = npar - tot;
}

// sendbuf: [nslots][MIG_NVAL*cap + 1], last entry = count
__global__ void part_mig_pack(MigPar g, int npar, int mpar, const unsigned *__restrict__ skeys, const unsigned *__restrict__ svals,
                              const double *__restrict__ fp, const double *__restrict__ dfp, const int *__restrict__ ipar,
                              const int *__restrict__ counters, double *__restrict__ sendbuf)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)MIG_NVAL * g.cap + 1;
  if (i < g.nslots) sendbuf[(size_t)i * stride + (size_t)MIG_NVAL * g.cap] = (double)counters[i];
  if (i >= counters[27] || i >= npar) return;
  const unsigned s = skeys[i], k = svals[i];
  const int j = i - counters[32 + s];
  if (j >= g.cap) return;
  double *b = sendbuf + (size_t)s * stride;
#pragma unroll
  for (int v = 0; v < 6; v++) { b[(size_t)v * g.cap + j] = fp[(size_t)v * mpar + k]; b[(size_t)(6 + v) * g.cap + j] = dfp[(size_t)v * mpar + k]; }
  b[(size_t)12 * g.cap + j] = (double)ipar[k];
}

// pre = exclusive scan of leave.  holes[i] / tails[i]: see the header comment.
__global__ void part_mig_holes(int npar, const int *__restrict__ leave, const int *__restrict__ pre, const int *__restrict__ counters,
                               int *__restrict__ holes, int *__restrict__ tails)
{
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= npar) return;
  const int nleave = counters[27], nstay = counters[28];
  if (leave[k]) {
    if (k < nstay) { const int pre_n = nstay < npar ? pre[nstay] : nleave; holes[pre_n - pre[k] - 1] = k; }
  } else if (k >= nstay) {
    tails[(npar - 1 - k) - (nleave - pre[k])] = k;
  }
}
__global__ void part_mig_fill(int npar, int mpar, const int *__restrict__ pre, const int *__restrict__ counters,
                              const int *__restrict__ holes, const int *__restrict__ tails,
                              double *__restrict__ fp, double *__restrict__ dfp, int *__restrict__ ipar)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int nstay = counters[28], nleave = counters[27];
  const int nholes = nstay < npar ? pre[nstay] : nleave;
  if (i >= nholes) return;
  const int d = holes[i], s = tails[i];
#pragma unroll
  for (int v = 0; v < 6; v++) { fp[(size_t)v * mpar + d] = fp[(size_t)v * mpar + s]; dfp[(size_t)v * mpar + d] = dfp[(size_t)v * mpar + s]; }
  ipar[d] = ipar[s];
}

// arrivals: sender slots are in ascending rank order (":do i=0,ncpus-1" receive loop, :613)
__global__ void part_mig_append(MigPar g, int mpar, const double *__restrict__ recvbuf, int *__restrict__ counters,
                                double *__restrict__ fp, double *__restrict__ dfp, int *__restrict__ ipar)
{
  const int s = blockIdx.y, j = blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)MIG_NVAL * g.cap + 1;
  int off = counters[28], tot = counters[28];
  for (int t = 0; t < g.nslots; t++) {
    const int c = (int)recvbuf[(size_t)t * stride + (size_t)MIG_NVAL * g.cap];
    if (t < s) off += c;
    tot += c;
  }
  if (s == 0 && j == 0) { counters[30] = tot; if (tot > mpar) atomicOr(&counters[29], 4); }
  const int cnt = (int)recvbuf[(size_t)s * stride + (size_t)MIG_NVAL * g.cap];
  if (j >= cnt || off + j >= mpar) return;
  const double *b = recvbuf + (size_t)s * stride;
  const int d = off + j;
#pragma unroll
  for (int v = 0; v < 6; v++) { fp[(size_t)v * mpar + d] = b[(size_t)v * g.cap + j]; dfp[(size_t)v * mpar + d] = b[(size_t)(6 + v) * g.cap + j]; }
  ipar[d] = (int)b[(size_t)12 * g.cap + j];
}

// periodic wrap only (boundconds_particles runs before migrate_particles, particles_sub.f90:135)
__global__ void part_wrap(PartPar q, double *__restrict__ fp)
{
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= q.npar) return;
#pragma unroll
  for (int ax = 0; ax < 3; ax++) {
    double x = fp[(size_t)ax * q.mpar + k];
    if (x < q.xyz0[ax]) x = x + q.Lxyz[ax];
    if (x >= q.xyz0[ax] + q.Lxyz[ax]) x = x - q.Lxyz[ax];
    fp[(size_t)ax * q.mpar + k] = x;
  }
}
