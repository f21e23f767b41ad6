"""Run set-ups for the hot path: what start.in/run.in + cparam.local resolve to
for the BASELINE configs, as a Params block, plus the
deterministic synthetic initial fields of SURVEY.md section 8(d).

The numbers mirror the reference's shipped inputs:
  conv-slab        samples/conv-slab/{start.in,run.in,src/cparam.local}
  helical-MHDturb  samples/helical-MHDturb/{start.in,run.in}
Defaults: cdata.f90:145-153 (cdt,cdtv,cdts,dtmax), timestep.f90:40-41.
"""
import math

import numpy as np

from . import abi

NG = abi.NGHOST


def grid_axis(n, x0, L, periodic):
    """Equidistant grid incl. ghosts (grid.f90:118-222): periodic xi=0.5..N-0.5,
    dx=L/N; non-periodic xi=0..N-1, dx=L/(N-1)."""
    dx = L / n if periodic else L / max(n - 1, 1)
    xi = np.arange(-NG, n + NG, dtype=np.float64) + (0.5 if periodic else 0.0)
    return x0 + dx * xi, dx


def make_params(nxgrid, nygrid, nzgrid, *, magnetic=False, entropy=False, xyz0=(-math.pi,) * 3,
                Lxyz=(2 * math.pi,) * 3, lperi=(1, 1, 1), nprocz=1, ipz=0, nproc=None, iproc=None, **kw):
    """Build (Params, zprofiles, grid) for one block of the global grid.

    Decomposition as the reference (cparam.f90:18, mpicomm.f90:367-369): nproc=(nprocx,nprocy,nprocz),
    rank iproc = ipx + nprocx*(ipy + nprocy*ipz); `nprocz=,ipz=` is the z-slab shorthand."""
    P = abi.Params()
    P.abi_version = abi.ABI_VERSION
    if nproc is None:
        nproc, ip = (1, 1, nprocz), (0, 0, ipz)
    else:
        iproc = 0 if iproc is None else iproc
        ip = (iproc % nproc[0], (iproc // nproc[0]) % nproc[1], iproc // (nproc[0] * nproc[1]))
    assert nxgrid % nproc[0] == 0 and nygrid % nproc[1] == 0 and nzgrid % nproc[2] == 0
    nx, ny, nz = nxgrid // nproc[0], nygrid // nproc[1], nzgrid // nproc[2]
    ipz = ip[2]
    P.nx, P.ny, P.nz = nx, ny, nz
    P.nprocx, P.nprocy, P.nprocz = nproc
    P.ipx, P.ipy, P.ipz = ip
    for i in range(3):
        P.lperi[i] = int(lperi[i])
        P.xyz0[i] = xyz0[i]
        P.Lxyz[i] = Lxyz[i]
    # variable slots in registration order (register.f90:124-135)
    P.iux, P.iuy, P.iuz, P.ilnrho = 1, 2, 3, 4
    nv = 4
    if entropy:
        nv += 1
        P.iss = nv
    if magnetic:
        P.iax, P.iay, P.iaz = nv + 1, nv + 2, nv + 3
        nv += 3
    P.mvar = P.mfarray = nv
    P.lhydro = P.ldensity = 1
    P.lentropy, P.lmagnetic = int(entropy), int(magnetic)
    P.lbidiagonal_derij = 1
    P.lviscosity_heat = 1
    P.lohmic_heat = 1
    P.ldt = 1
    xg, dx = grid_axis(nxgrid, xyz0[0], Lxyz[0], lperi[0])
    yg, dy = grid_axis(nygrid, xyz0[1], Lxyz[1], lperi[1])
    zg, dz = grid_axis(nzgrid, xyz0[2], Lxyz[2], lperi[2])
    x = xg[ip[0] * nx: ip[0] * nx + nx + 2 * NG].copy()
    y = yg[ip[1] * ny: ip[1] * ny + ny + 2 * NG].copy()
    z = zg[ipz * nz: ipz * nz + nz + 2 * NG].copy()
    P.dx, P.dy, P.dz = dx, dy, dz
    P.dx_1, P.dy_1, P.dz_1 = 1.0 / dx, 1.0 / dy, 1.0 / dz
    for i in range(1, NG + 1):          # grid.f90:2670-2672
        P.dz2_bound[3 - i] = 2.0 * (z[NG + i] - z[NG])
        P.dz2_bound[3 + i] = 2.0 * (z[NG + nz - 1] - z[NG + nz - 1 - i])
    P.x0_loc, P.y0_loc, P.z0_loc = x[NG], y[NG], z[NG]
    P.cdt, P.cdtv, P.cdts, P.dtmax = 0.9, 0.25, 1.0, 1e37
    P.alpha_ts[0], P.alpha_ts[1], P.alpha_ts[2] = 0.0, -5 / 9.0, -153 / 128.0
    P.beta_ts[0], P.beta_ts[1], P.beta_ts[2] = 1 / 3.0, 15 / 16.0, 8 / 15.0
    P.mu0 = 1.0
    P.cs20, P.gamma, P.cp, P.lnrho0 = 1.0, (5.0 / 3.0 if entropy else 1.0), 1.0, 0.0
    P.ztop = xyz0[2] + Lxyz[2]
    for name, val in kw.items():
        if name in ("bcx", "bcy", "bcz"):
            continue
        if not hasattr(P, name):
            raise KeyError(name)
        setattr(P, name, val)
    for axis in ("bcx", "bcy", "bcz"):
        spec = kw.get(axis)
        arr = getattr(P, axis)
        for k in range(abi.MAXVAR):
            lo = hi = "p"
            if spec is not None and k < len(spec):
                s = spec[k]
                lo, hi = s.split(":") if ":" in s else (s, s)
            arr[0][k] = abi.BC[lo]
            arr[1][k] = abi.BC[hi]
    grid = dict(x=x, y=y, z=z, zglobal=zg, dx=dx, dy=dy, dz=dz)
    zprof = {}
    if entropy:
        zprof = entropy_zprofiles(P, z[NG:NG + nz])
    return P, zprof, grid


def _step(x, x0, width):          # sub.f90:4544 step_vector
    tini = np.finfo(np.float64).tiny * 5
    return 0.5 * (1 + np.tanh((x - x0) / (width + tini)))


def _der_step(x, x0, width):      # sub.f90 der_step
    tini = np.finfo(np.float64).tiny * 5
    arg = np.abs((x - x0) / (width + tini))
    return np.where(arg >= 8.0, 2.0 / width * np.exp(-2.0 * arg), 0.5 / (width * np.cosh(np.minimum(arg, 8.0)) ** 2))


def entropy_zprofiles(P, z):
    """hcond_prof, dlnhcond_prof (get_gravz_heatcond, entropy.f90:7243-7262) and the
    'gaussian' cooling profile of get_heat_cool_gravz (entropy.f90:6244)."""
    out = {}
    if P.iheatcond == abi.HCOND["K-profile"]:
        prof = 1.0 + (P.hcond1 - 1.0) * _step(z, P.z1, -P.widthss) + (P.hcond2 - 1.0) * _step(z, P.z2, P.widthss)
        dln = ((P.hcond1 - 1.0) * _der_step(z, P.z1, -P.widthss) + (P.hcond2 - 1.0) * _der_step(z, P.z2, P.widthss)) / prof
        out[abi.PROF_HCOND] = P.hcond0 * prof
        out[abi.PROF_DLNHCOND] = dln
    if P.lcooling:
        out[abi.PROF_COOL] = np.exp(-0.5 * ((P.ztop - z) / P.wcool) ** 2)
    return out


# --------------------------------------------------------------------------
# synthetic smooth fields, seed 12345
# --------------------------------------------------------------------------
def _mode_sum(rng, grid, nmodes, kmax, ampl, Lxyz, xyz0):
    """sum_i a_i cos(k_i.(x-x0) + phi_i), k_i integer multiples of 2pi/L: periodic
    on the box, evaluated on the ghost-padded grid (ghosts consistent)."""
    x, y, z = grid["x"], grid["y"], grid["z"]
    acc = np.zeros((z.size, y.size, x.size))
    k = rng.integers(-kmax, kmax + 1, size=(nmodes, 3))
    k[(k == 0).all(axis=1)] = (1, 0, 0)
    a = rng.standard_normal(nmodes) / math.sqrt(nmodes)
    phi = rng.uniform(0, 2 * math.pi, nmodes)
    ex = np.exp(1j * np.outer(k[:, 0] * 2 * math.pi / Lxyz[0], x - xyz0[0]))
    ey = np.exp(1j * np.outer(k[:, 1] * 2 * math.pi / Lxyz[1], y - xyz0[1]))
    ez = np.exp(1j * np.outer(k[:, 2] * 2 * math.pi / Lxyz[2], z - xyz0[2])) * (a * np.exp(1j * phi))[:, None]
    exy = (ey[:, :, None] * ex[:, None, :]).reshape(nmodes, -1)        # [i, y*x]
    acc = (ez.T.real @ exy.real - ez.T.imag @ exy.imag).reshape(z.size, y.size, x.size)
    return ampl * acc


def synthetic_fields(P, grid, seed=12345, ampl_u=0.1, ampl_lnrho=0.1, ampl_a=1e-2, ampl_ss=0.05):
    """f(mvar,mz,my,mx) C-ordered == Fortran f(mx,my,mz,mvar)."""
    rng = np.random.default_rng(seed)
    L = [P.Lxyz[i] for i in range(3)]
    x0 = [P.xyz0[i] for i in range(3)]
    mx, my, mz = P.nx + 2 * NG, P.ny + 2 * NG, P.nz + 2 * NG
    f = np.zeros((P.mvar, mz, my, mx))
    # the random draws are made in a fixed order so that every rank of a
    # decomposed run generates the same global field
    for j in range(3):
        f[P.iux - 1 + j] = _mode_sum(rng, grid, 16, 4, ampl_u, L, x0)
    f[P.ilnrho - 1] = _mode_sum(rng, grid, 8, 3, ampl_lnrho, L, x0)
    if P.lentropy:
        f[P.iss - 1] = _mode_sum(rng, grid, 8, 3, ampl_ss, L, x0)
    if P.lmagnetic:
        for j in range(3):
            f[P.iax - 1 + j] = _mode_sum(rng, grid, 16, 8, ampl_a, L, x0)
    return f


def config_hydro(n=64, **kw):
    """BASELINE config 2: periodic isothermal hydro (forcing optional)."""
    d = dict(nu=5e-3, cdt=0.4, cdtv=0.3)
    d.update(kw)
    return make_params(n, n, n, **d)


def config_mhd(n=64, nprocz=1, ipz=0, **kw):
    """BASELINE config 3: isothermal helical MHD (samples/helical-MHDturb/run.in:
    cdt=0.4, cdtv=0.3, eta=nu=5e-3)."""
    d = dict(nu=5e-3, eta=5e-3, cdt=0.4, cdtv=0.3)
    d.update(kw)
    return make_params(n, n, n, magnetic=True, nprocz=nprocz, ipz=ipz, **d)


def config_mhd_entropy(n=64, nprocz=1, ipz=0, **kw):
    """BASELINE config 4: MHD + entropy, iheatcond='chi-const' (SURVEY A.5)."""
    d = dict(nu=5e-3, eta=5e-3, cdt=0.4, cdtv=0.3, chi=1e-4, iheatcond=abi.HCOND["chi-const"])
    d.update(kw)
    return make_params(n, n, n, magnetic=True, entropy=True, nprocz=nprocz, ipz=ipz, **d)


def config_convslab(nxgrid=32, nygrid=32, nzgrid=32, nprocz=1, ipz=0, **kw):
    """BASELINE config 1: samples/conv-slab (start.in, run.in)."""
    gamma = 1.6666666
    cs0 = 0.57735
    hcond0, mpoly0, mpoly1, mpoly2, gravz = 8e-3, 1.0, 3.0, 0.0, -1.0
    d = dict(
        xyz0=(-0.5, -0.5, -0.68), Lxyz=(1.0, 1.0, 2.0), lperi=(1, 1, 0),
        nu=4e-3, cdt=0.4, cdtv=0.3, Omega=0.1, lgravz=1, gravz=gravz,
        lupw_lnrho=1, lupw_ss=1, cs20=cs0 ** 2, gamma=gamma,
        iheatcond=abi.HCOND["K-profile"], hcond0=hcond0,
        hcond1=(mpoly1 + 1.0) / (mpoly0 + 1.0), hcond2=(mpoly2 + 1.0) / (mpoly0 + 1.0),
        z1=0.0, z2=1.0, widthss=0.05, lcooling=1, cool=15.0, wcool=0.2, cs2cool=0.333333,
        cs2top=0.333333,
        Fbot=-gamma / (gamma - 1) * hcond0 * gravz / (mpoly0 + 1),          # entropy.f90:716
        bcz=["s", "s", "a", "a2", "c1:cT"],
    )
    d.update(kw)
    return make_params(nxgrid, nygrid, nzgrid, entropy=True, nprocz=nprocz, ipz=ipz, **d)


# ---------------------------------------------------------------------------
# dust particles (BASELINE config 5)
class ParticleState:
    """Host-side particle arrays in the reference's layout (particles_cdata.f90):
    fp(mpar_loc,mparray), dfp(mpar_loc,mpvar) -> [6][mpar_loc]; ipar; ineargrid [3][mpar_loc];
    np/rhop aux fields (mz,my,mx); k1_imn/k2_imn (ny*nz)."""

    def __init__(self, P, npar, mpar=None):
        mpar = npar if mpar is None else mpar
        self.npar, self.mpar = npar, mpar
        self.fp = np.zeros((6, mpar))
        self.dfp = np.zeros((6, mpar))
        self.ipar = np.zeros(mpar, dtype=np.int32)
        self.ineargrid = np.zeros((3, mpar), dtype=np.int32)
        shape = (P.nz + 2 * NG, P.ny + 2 * NG, P.nx + 2 * NG)
        self.np_ = np.zeros(shape)
        self.rhop = np.zeros(shape)
        self.k1_imn = np.zeros(P.ny * P.nz, dtype=np.int32)
        self.k2_imn = np.zeros(P.ny * P.nz, dtype=np.int32)

    def copy(self):
        import copy
        return copy.deepcopy(self)


def make_particle_params(P, npar, mpar=None, *, tausp=0.1, eps_dtog=0.5, backreaction=True, **kw):
    """particles_run_pars: constant friction time (ldraglaw_epstein with tausp), TSC mesh,
    rhop_swarm from eps_dtog as particles_dust.f90:830-850 does for a uniform gas of density
    rho0=exp(lnrho0): rhop_swarm = eps_dtog*rho0/(npar/nwgrid)."""
    pp = abi.ParticleParams()
    pp.abi_version = abi.ABI_VERSION
    pp.npar_loc = npar
    pp.mpar_loc = npar if mpar is None else mpar
    pp.ldragforce_dust_par = 1
    pp.ldragforce_gas_par = int(backreaction)
    pp.lcoriolis_force_par = 1
    pp.ldt_adv_par = 1
    pp.lcheck_exact_frontier = 1
    pp.tausp = tausp
    nw = P.nx * P.nprocx * P.ny * P.nprocy * P.nz * P.nprocz
    npar_glob = npar * P.nprocx * P.nprocy * P.nprocz
    pp.rhop_swarm = eps_dtog * math.exp(P.lnrho0) / (npar_glob / nw)
    pp.cdtp, pp.cdtp_drag = 0.2, 0.2
    for name, val in kw.items():
        if not hasattr(pp, name):
            raise KeyError(name)
        setattr(pp, name, val)
    return pp


def synthetic_particles(P, pp, seed=4321, vampl=0.05, block_only=False):
    """Uniform random positions in the global box (or, block_only, in this rank's block),
    small random velocities; ipar = 1..npar (offset by iproc*npar when block_only)."""
    iproc = P.ipx + P.nprocx * (P.ipy + P.nprocy * P.ipz)
    rng = np.random.default_rng(seed + (iproc if block_only else 0))
    st = ParticleState(P, pp.npar_loc, pp.mpar_loc)
    n = pp.npar_loc
    npr, ip = (P.nprocx, P.nprocy, P.nprocz), (P.ipx, P.ipy, P.ipz)
    for j in range(3):
        lo, L = P.xyz0[j], P.Lxyz[j]
        if block_only:
            L = P.Lxyz[j] / npr[j]
            lo = P.xyz0[j] + ip[j] * L
        st.fp[j, :n] = lo + L * rng.random(n)
        st.fp[3 + j, :n] = vampl * rng.standard_normal(n)
    st.ipar[:n] = np.arange(1, n + 1, dtype=np.int32) + (iproc * n if block_only else 0)
    return st


# ---------------------------------------------------------------------------
# helical forcing coefficients (host side of forcing_hel; the random draws stay on the host)
def forcing_shell(kmin=4.5, kmax=5.5):
    """integer wave vectors with kmin <= |k| < kmax: what samples/helical-MHDturb/k.dat lists for kav=5"""
    r = int(math.ceil(kmax))
    ks = [(kx, ky, kz) for kx in range(-r, r + 1) for ky in range(-r, r + 1) for kz in range(-r, r + 1)
          if kmin <= math.sqrt(kx * kx + ky * ky + kz * kz) < kmax]
    return np.array(ks, dtype=np.float64)


def helical_forcing_coefs(kvec, phase, phi, dt, *, force=0.07, relhel=1.0, kav=5.0, cs0eff=1.0, slope_ff=0.0):
    """fconst_coefs_hel (forcing.f90:1554-1760) for one wave vector, phase and e-vector angle phi (the three random
    numbers of a step), default switches (no shear, no k scaling, lhelical_test=F): returns abi.Forcing.
    f_u += Re[(coef1 + i coef2) exp(i(k.x + phase))] * fact,  fact = force/ffnorm*sqrt(dt)  (forcing.f90:1741)."""
    kx, ky, kz = (float(v) for v in kvec)
    k2 = kx * kx + ky * ky + kz * kz
    k = math.sqrt(k2)
    ee = np.array([0.0, 1.0, 0.0]) if (ky == 0 and kz == 0) else np.array([1.0, 0.0, 0.0])   # forcing.f90:1651-1655
    kk = np.array([kx, ky, kz])
    e1 = np.cross(kk, ee); e1 /= math.sqrt(e1 @ e1)
    e2 = np.cross(kk, e1); e2 /= math.sqrt(e2 @ e2)
    ee = math.cos(phi) * e1 + math.sin(phi) * e2                                            # :1667-1675
    kde = float(kk @ ee)
    ke = np.cross(kk, ee)
    kke = np.cross(kk, ke)
    ffnorm = math.sqrt(1.0 + relhel ** 2) * k * math.sqrt(k2 - kde ** 2) / math.sqrt(kav * cs0eff ** 3) * (k / kav) ** slope_ff
    frc = abi.Forcing()
    for j in range(3):
        frc.coef1[j] = k * ke[j]
        frc.coef2[j] = relhel * kke[j]
        frc.fda[j] = 1.0
        frc.kk[j] = kk[j]
    frc.phase = phase
    frc.fact = force / ffnorm * math.sqrt(dt)
    return frc


def forcing_sequence(nsteps, seed=777):
    """a recorded (k, phase, phi) sequence for bench.py --forcing (numpy generator, fixed seed: synthetic, not the
    reference's nr_f90 stream)"""
    rng = np.random.default_rng(seed)
    shell = forcing_shell()
    return [(shell[int(rng.integers(len(shell)))], math.pi * (2 * rng.random() - 1), 2 * math.pi * rng.random()) for _ in range(nsteps)]
