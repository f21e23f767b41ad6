"""Known-answer tests of the particle oracle (oracle/pcref_particles.c).

The reference has no double-precision golden file for a particle run inside the
path (parity unpinned, see the oracle header), so the restatement is checked
against properties that follow from the reference's formulas:
index arithmetic (grid.f90:2310-2343), stable pencil sort (particles_map.f90:986),
TSC partition of unity (particles_map.f90:1607-1640), exactness of the quadratic
spline for linear fields (particles_map.f90:686-722), momentum conservation of the
drag back-reaction (particles_dust.f90:4443-4536) and the drag relaxation time.
"""
import numpy as np
import pytest

NG = 3


def _setup(pk, n=16, npar=4000, **kw):
    P, zprof, grid = pk.setups.make_params(n, n, n, nu=1e-3, cdt=0.4, cdtv=0.3)
    pp = pk.setups.make_particle_params(P, npar, **kw)
    st = pk.setups.synthetic_particles(P, pp)
    return P, grid, pp, st


def test_nearest_grid_and_wrap(pk, orc):
    P, grid, pp, st = _setup(pk)
    # push some particles outside the box: wrap must bring them back (particles_sub.f90:170-215)
    st.fp[0, :50] += P.Lxyz[0] * 0.75
    st.fp[1, 50:100] -= P.Lxyz[1] * 0.75
    orc.particles_boundconds(st, P, pp)
    for j in range(3):
        assert (st.fp[j] >= P.xyz0[j]).all() and (st.fp[j] < P.xyz0[j] + P.Lxyz[j]).all()
    orc.particles_map_nearest_grid(st, P, pp)
    for j, ax in enumerate("xyz"):
        i0 = st.ineargrid[j] - 1                     # 1-based Fortran index -> 0-based
        assert i0.min() >= NG and i0.max() <= NG + 16 - 1
        d = np.abs(st.fp[j] - grid[ax][i0]) / grid["d" + ax]
        assert d.max() <= 0.5 + 1e-12                # really the nearest point


def test_imn_array_orders(pk, orc):
    P, _, _ = pk.setups.make_params(16, 12, 10)
    a = orc.imn_array(P)
    assert (a.ravel() == np.arange(1, 12 * 10 + 1)).all()        # ncpus=1: plain (m,n) order, general.f90:250-261
    P2, _, _ = pk.setups.make_params(16, 24, 20, nproc=(1, 2, 2), iproc=0)
    b = orc.imn_array(P2)
    assert sorted(b.ravel().tolist()) == list(range(1, 12 * 10 + 1))   # a permutation: every pencil once
    assert b[5, 5] < b[0, 0] and b[5, 5] < b[9, 11]              # inner rectangle first (general.f90:267-277)


def test_sort_is_stable_counting_sort(pk, orc):
    P, grid, pp, st = _setup(pk)
    st.dfp[:] = np.random.default_rng(1).random(st.dfp.shape)
    orc.particles_map_nearest_grid(st, P, pp)
    key = orc.imn_array(P)[st.ineargrid[2] - NG - 1, st.ineargrid[1] - NG - 1]
    order = np.argsort(key, kind="stable")
    want_fp, want_dfp, want_ipar = st.fp[:, order].copy(), st.dfp[:, order].copy(), st.ipar[order].copy()
    orc.particles_sort(st, P, pp)
    assert (st.fp == want_fp).all() and (st.dfp == want_dfp).all() and (st.ipar == want_ipar).all()
    skey = np.sort(key)
    for imn in (1, 7, 100, 256):
        idx = np.nonzero(skey == imn)[0]
        if len(idx):
            assert st.k1_imn[imn - 1] == idx[0] + 1 and st.k2_imn[imn - 1] == idx[-1] + 1
        else:
            assert st.k1_imn[imn - 1] == 0 and st.k2_imn[imn - 1] == 0


def test_tsc_partition_of_unity(pk, orc):
    P, grid, pp, st = _setup(pk)
    orc.particles_map_nearest_grid(st, P, pp)
    orc.particles_sort(st, P, pp)
    orc.particles_map_xxp_grid(st, P, pp)
    inner = (slice(NG, -NG),) * 3
    assert st.np_[inner].sum() == pp.npar_loc
    assert abs(st.rhop[inner].sum() / (pp.rhop_swarm * pp.npar_loc) - 1) < 1e-12
    # ghost layer was folded back and zeroed (ghostfold_mpicomm.f90:69-70)
    assert st.rhop[NG - 1].max() == 0.0 and st.rhop[:, :, -NG].max() == 0.0


def test_drag_conserves_momentum_and_spline_is_exact_for_linear_fields(pk, orc):
    P, grid, pp, st = _setup(pk, npar=3000)
    mz = my = mx = 16 + 2 * NG
    f = np.zeros((P.mvar, mz, my, mx))
    X = grid["x"][None, None, :] + 0 * grid["y"][None, :, None] + 0 * grid["z"][:, None, None]
    Y = 0 * X + grid["y"][None, :, None]
    Z = 0 * X + grid["z"][:, None, None]
    f[P.iux - 1] = 0.3 + 0.1 * X - 0.2 * Y + 0.05 * Z           # linear: the spline reproduces it exactly
    f[P.iuy - 1] = -0.1
    f[P.iuz - 1] = 0.2 * Z
    orc.particles_map_nearest_grid(st, P, pp)
    orc.particles_sort(st, P, pp)
    df = np.zeros_like(f)
    st.dfp[:] = 0.0
    orc.particles_pde(f, df, st, P, pp, lfirst=False)
    xp, yp, zp = st.fp[0], st.fp[1], st.fp[2]
    uup = np.stack([0.3 + 0.1 * xp - 0.2 * yp + 0.05 * zp, -0.1 + 0 * xp, 0.2 * zp])
    drag = -(st.fp[3:6] - uup) / pp.tausp
    assert np.abs(st.dfp[3:6] - drag).max() < 1e-12
    assert np.abs(st.dfp[0:3] - st.fp[3:6]).max() == 0.0         # dxxp_dt: dx/dt = v
    # rho = 1: sum over the (folded) grid of df_u = -rhop_swarm * sum of drag forces
    inner = (slice(NG, -NG),) * 3
    for j in range(3):
        tot = df[P.iux - 1 + j][inner].sum()
        assert abs(tot + pp.rhop_swarm * drag[j].sum()) < 1e-10 * max(1.0, abs(tot))
        assert df[P.iux - 1 + j][NG - 1].max() == 0.0            # ghost layer folded away


def test_drag_relaxation_time(pk, orc):
    """dv/dt = -(v-u)/tausp in a uniform stream: v(t) = u + (v0-u) exp(-t/tausp)."""
    P, grid, pp, st = _setup(pk, npar=64, backreaction=False, tausp=0.2)
    P.ldt, P.dt_fixed = 0, 2e-3
    mz = 16 + 2 * NG
    f = np.zeros((P.mvar, mz, mz, mz))
    f[P.iux - 1] = 0.25
    st.fp[3:6] = 0.0
    st.fp[3] = -0.1
    df = np.zeros_like(f)
    t, dt = 0.0, 0.0
    for _ in range(50):
        dt = orc.timestep_particles(f, df, P, pp, st, dt)
        t += dt
    # the gas stays uniform (no forces), so the exact solution applies; RK3 error ~ (dt/tau)^3
    assert np.abs(f[P.iux - 1][NG:-NG, NG:-NG, NG:-NG] - 0.25).max() < 1e-13
    want = 0.25 + (-0.1 - 0.25) * np.exp(-t / pp.tausp)
    assert np.abs(st.fp[3, :64] - want).max() < 1e-7
    assert np.abs(st.fp[4, :64]).max() < 1e-15
