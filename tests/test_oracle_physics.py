"""CPU checks of the oracle's physics against analytic statements."""
import numpy as np


def test_rhs_of_uniform_state_is_zero(pk, orc):
    P, zprof, grid = pk.setups.config_mhd(16)
    f = np.zeros((P.mvar, 22, 22, 22))
    f[P.ilnrho - 1] = 0.3
    f[P.iux - 1] = 0.0
    df = np.zeros_like(f)
    dt1 = orc.rhs(f, df, P, True, 0.0)
    assert np.abs(df).max() == 0.0
    # sound-speed + viscous limit: dt1 = sqrt((cs0*sqrt(dxyz_2)/cdt)^2 + (nu*dxyz_2/cdtv)^2)
    dxyz_2 = 3 * P.dx_1 ** 2
    want = np.hypot(np.sqrt(P.cs20 * dxyz_2) / P.cdt, max(P.nu, P.eta) * dxyz_2 / P.cdtv)
    assert abs(dt1 - want) < 1e-13 * want


def test_beltrami_field_is_force_free_and_decays(pk, orc):
    """A = (cos z, sin z... ) ABC-type Beltrami field: curl A = A  =>  J x B = 0 and dA/dt = -eta A for u=0."""
    n = 32
    P, zprof, grid = pk.setups.config_mhd(n, nu=0.0)
    x, y, z = grid["x"], grid["y"], grid["z"]
    Z, Y, X = np.meshgrid(z, y, x, indexing="ij")
    f = np.zeros((P.mvar, n + 6, n + 6, n + 6))
    f[P.iax - 1] = np.sin(Z) + np.cos(Y)
    f[P.iay - 1] = np.sin(X) + np.cos(Z)
    f[P.iaz - 1] = np.sin(Y) + np.cos(X)
    df = np.zeros_like(f)
    orc.rhs(f, df, P, True, 0.0)
    s = (slice(3, -3),) * 3
    for k in range(3):
        assert np.abs(df[P.iux - 1 + k][s]).max() < 2e-6          # J x B = O(dx^6)
        a = f[P.iax - 1 + k][s]
        assert np.abs(df[P.iax - 1 + k][s] + P.eta * a).max() < 1e-8


def test_periodic_bc_wraps(pk, orc):
    P, zprof, grid = pk.setups.config_hydro(8)
    rng = np.random.default_rng(0)
    f = rng.standard_normal((P.mvar, 14, 14, 14))
    orc.boundconds(f, P)
    assert np.array_equal(f[:, :, :, :3], f[:, :, :, 8:11])
    assert np.array_equal(f[:, :, 11:, :], f[:, :, 3:6, :])
    assert np.array_equal(f[:, :3, :, :], f[:, 8:11, :, :])


def test_sym_antisym_bcs(pk, orc):
    P, zprof, grid = pk.setups.config_convslab(8, 8, 8)
    rng = np.random.default_rng(1)
    f = rng.standard_normal((P.mvar, 14, 14, 14)) * 0.1
    orc.boundconds(f, P)
    n1, n2 = 3, 10
    for i in (1, 2, 3):
        assert np.array_equal(f[0, n1 - i], f[0, n1 + i])             # 's'
        assert np.array_equal(f[2, n1 - i], -f[2, n1 + i])            # 'a'
        assert np.allclose(f[3, n2 + i], 2 * f[3, n2] - f[3, n2 - i])  # 'a2'
    assert np.all(f[2, n1] == 0) and np.all(f[2, n2] == 0)
    # 'cT' top: cs2(n2) == cs2top
    cv1 = P.gamma / P.cp
    cs2 = P.cs20 * np.exp(cv1 * f[4, n2] + (P.gamma - 1) * (f[3, n2] - P.lnrho0))
    assert np.allclose(cs2, P.cs2top, rtol=1e-13)


def test_rk3_third_order_convergence(pk, orc):
    """fixed dt halved -> error of the 2N-RK3 (Williamson) drops ~8x"""
    errs = []
    for nsub in (1, 2, 4):
        P, zprof, grid = pk.setups.config_hydro(16, ldt=0, dt_fixed=0.02 / nsub, nu=1e-2)
        f = pk.setups.synthetic_fields(P, grid, ampl_u=0.05, ampl_lnrho=0.05)
        df = np.zeros_like(f)
        for _ in range(4 * nsub):
            orc.timestep(f, df, P, 0.0)
        errs.append(f[:, 3:-3, 3:-3, 3:-3].copy())
    P, zprof, grid = pk.setups.config_hydro(16, ldt=0, dt_fixed=0.02 / 16, nu=1e-2)
    f = pk.setups.synthetic_fields(P, grid, ampl_u=0.05, ampl_lnrho=0.05)
    df = np.zeros_like(f)
    for _ in range(64):
        orc.timestep(f, df, P, 0.0)
    ref = f[:, 3:-3, 3:-3, 3:-3]
    e = [np.abs(a - ref).max() for a in errs]
    assert e[0] / e[1] > 6.0 and e[1] / e[2] > 6.0, e
