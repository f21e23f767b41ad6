"""PIN of the DUST-PARTICLE path of the checker against a golden vector the reference holds:
samples/ParticlesCaustics/reference.out -- 30 000 particles_dust particles in helically forced isothermal hydro at
32^3 (the flow of samples/helical-MHDturb without the magnetic field: HYDRO+DENSITY+FORCING, gamma=1, nu=5e-3,
cdt=0.4, cdtv=0.3), run by the reference on 4 processes (nprocz=4), random_gen='nr_f90'.  No shear, no self-gravity:
the one particle fixture of the reference that lies inside the hot path.  Printed every 2 steps:
t, dt, umax, urms, orms, omax, oum, mass, xp2m, yp2m, zp2m, vprms (blowupm is an uninitialised value in the file,
trsigmapm belongs to particles_caustics, a diagnostics module outside the path).

What the file pins, routine by routine of oracle/pcref_particles.c:
  * the initial particle state: every process draws, from its own stream (seed -(10+iproc), start.f90:383), first the
    radius class of its npar/ncpus particles (set_particle_radius 'constant' with 5 radii, particles_radius.f90:256-265,
    called before init_particles: particles_main.f90:282,289), then x, y, z per particle ('random-box',
    particles_dust.f90:1328-1356) -> xp2m, yp2m, zp2m to all four printed digits at it=0;
  * the time step limited by the drag: dt = cdtp_drag/max(tausp1) = 1.42e-3 (particles_dust.f90:4843-4847,4901-4910);
  * particles_timestep_first/second, boundconds_particles, map_nearest_grid, sort_particles_imn, the drag acceleration
    -(v - u(x_p))/tau (dvvp_dt_pencil :4264-4275), dxxp_dt: vprms to four digits over 8 steps, i.e. 24 substeps in
    which the particles are driven by the interpolated, forced gas velocity.
Two settings of the sample differ from what libpcb200.so implements and exist in the CHECKER ONLY
(pcref_particles_fixture): interp_pol_uu='cic' (interpolate_linear, particles_map.f90:112-287) instead of the
quadratic spline, and ldraglaw_purestokes (tausp1 = 18 nu/((rhopmat/rho(x_p)) (2 a_p)^2), particles_dust.f90:6314-6362)
instead of a constant friction time.  The TSC assignment, the quadratic-spline interpolation and the back-reaction on
the gas stay covered by known-answer tests only (tests/test_oracle_particles.py); the CUDA path is tied to this
checker by test_particles_match_oracle in tests/test_gpu_parity.py.

vprms as the reference prints it is sqrt(sum_k v_k^2)/npar (itype ilabel_sum_sqrt_par, diagnostics.f90:769-770), not
sqrt(sum/npar); reproduced as such.  Tolerance: 1.5 units of the last printed digit (the reference's harness)."""
import math
import os

import numpy as np

from tests.test_oracle_helical_mhdturb import printed_tolerance

HERE = os.path.dirname(os.path.abspath(__file__))
COLS = {"t": 1, "dt": 2, "umax": 3, "urms": 4, "oum": 7, "mass": 8, "xp2m": 9, "yp2m": 10, "zp2m": 11, "vprms": 12}
NPAR, NCPUS = 30000, 4
AP0 = (4e-4, 6e-4, 8e-4, 10e-4, 12e-4)
RHOPMAT = 1e3


def initial_particles(pk, P):
    from oracle import nr_f90
    st = pk.setups.ParticleState(P, NPAR)
    ap = np.zeros(NPAR)
    nloc = NPAR // NCPUS
    for iproc in range(NCPUS):
        g = nr_f90.start_stream(iproc)
        sl = slice(iproc * nloc, (iproc + 1) * nloc)
        ap[sl] = [AP0[int(math.ceil(len(AP0) * r)) - 1] for r in g.draws(nloc)]
        r = g.draws(3 * nloc).reshape(nloc, 3)
        for j in range(3):
            st.fp[j, sl] = -3.14 + r[:, j] * 6.28                  # xp0 + r*Lx0 in the whole box (start.in)
    st.ipar[:] = np.arange(1, NPAR + 1)
    return st, ap


def test_oracle_reproduces_particles_caustics_reference_out(pk, orc):
    from oracle import nr_f90
    P, zprof, grid = pk.setups.make_params(32, 32, 32, magnetic=False, nu=5e-3, cdt=0.4, cdtv=0.3)
    pp = pk.setups.make_particle_params(P, NPAR, tausp=0.0, eps_dtog=1.0, backreaction=False, cdtp=0.9, cdtp_drag=0.2,
                                        lcoriolis_force_par=0)
    f = np.zeros((P.mvar, P.nz + 6, P.ny + 6, P.nx + 6))
    df = np.zeros_like(f)
    st, ap = initial_particles(pk, P)
    nk, kav, kkx, kky, kkz = nr_f90.read_kdat(os.path.join(HERE, "golden", "helical-MHDturb.k.dat"))   # the sample's k.dat is this file
    stream = nr_f90.run_stream()
    gold = [ln.split() for ln in open(os.path.join(HERE, "golden", "ParticlesCaustics.reference.out")) if not ln.startswith("#")]
    nw, vol = P.nx * P.ny * P.nz, (2 * math.pi) ** 3
    rows, t, dt = [], 0.0, 0.0
    try:
        for it in range(9):
            d = np.zeros(16)
            trial = st.copy()
            orc.particles_fixture(trial, ap, RHOPMAT, interp_cic=True)
            dt_step = orc.timestep_particles(f.copy(), df.copy(), P, pp, trial, dt, None, d)   # dt of this step, diagnostics
            if it % 2 == 0:
                n = dict(zip(pk.abi.DIAG_NAMES, d))
                fp = st.fp
                rows.append({"t": t, "dt": dt_step, "umax": math.sqrt(n["u2max"]), "urms": math.sqrt(n["u2sum"] / nw),
                             "oum": n["ousum"] / nw, "mass": n["rhosum"] / nw * vol,
                             "xp2m": (fp[0] ** 2).mean(), "yp2m": (fp[1] ** 2).mean(), "zp2m": (fp[2] ** 2).mean(),
                             "vprms": math.sqrt((fp[3:6] ** 2).sum()) / NPAR})
            phase, ik, phi = nr_f90.helical_draw(stream, nk)
            frc = pk.setups.helical_forcing_coefs((kkx[ik - 1], kky[ik - 1], kkz[ik - 1]), phase, phi, dt_step,
                                                  force=0.07, relhel=1.0, kav=kav)
            orc.particles_fixture(st, ap, RHOPMAT, interp_cic=True)
            dt = orc.timestep_particles(f, df, P, pp, st, dt, frc, None)
            assert dt == dt_step
            t += dt
    finally:
        orc.particles_fixture()                                    # the options are process-wide in the checker: off again
    assert len(rows) == len(gold) == 5
    worst = 0.0
    for got, ref in zip(rows, gold):
        for name, j in COLS.items():
            tol = printed_tolerance(ref[j])
            err = abs(got[name] - float(ref[j]))
            worst = max(worst, err / tol)
            assert err <= tol, (ref[0], name, got[name], ref[j])
    assert rows[-1]["vprms"] > 1e-5                                # the particles did pick up the gas motion
    print("oracle vs ParticlesCaustics reference.out: worst error / tolerance", worst)


def test_fixture_options_are_off_by_default(pk, orc):
    """with the options off the checker is the path the library implements: constant tausp and the quadratic spline"""
    P, zprof, grid = pk.setups.make_params(8, 8, 8, magnetic=False, nu=5e-3)
    pp = pk.setups.make_particle_params(P, 64, tausp=0.1, backreaction=False)
    st = pk.setups.synthetic_particles(P, pp)
    f = pk.setups.synthetic_fields(P, grid)
    a, b = st.copy(), st.copy()
    orc.timestep_particles(f.copy(), np.zeros_like(f), P, pp, a, 0.0)
    ap = np.full(64, 5e-4)
    orc.particles_fixture(b, ap, RHOPMAT, interp_cic=True)
    try:
        orc.timestep_particles(f.copy(), np.zeros_like(f), P, pp, b, 0.0)
    finally:
        orc.particles_fixture()
    assert not np.array_equal(a.fp, b.fp)
    c = st.copy()
    orc.timestep_particles(f.copy(), np.zeros_like(f), P, pp, c, 0.0)
    assert np.array_equal(a.fp, c.fp)
