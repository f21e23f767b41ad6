"""The pencil-wise evaluation used to TIME the CPU arm (oracle/pcref_pencil.c) against the checker (oracle/pcref.c):
same stencils and physics, rows of nx points at a time like the reference's mn-loop; fields to 1e-12, dt to 1e-13."""
import numpy as np
import pytest


@pytest.mark.parametrize("case", ["hydro", "mhd", "mhd_ragged", "hydro_upwind", "mhd_entropy", "mhd_entropy_upwind", "convslab"])
def test_pencil_timestep_matches_the_checker(pk, orc, case):
    if case == "hydro":
        P, zprof, grid = pk.setups.config_hydro(24, Omega=0.1)
    elif case == "mhd":
        P, zprof, grid = pk.setups.config_mhd(24)
    elif case == "mhd_ragged":
        P, zprof, grid = pk.setups.make_params(20, 12, 8, magnetic=True, nu=5e-3, eta=4e-3, cdt=0.4, cdtv=0.3)
    elif case == "hydro_upwind":
        P, zprof, grid = pk.setups.config_hydro(20, lupw_lnrho=1, lupw_uu=1)
    elif case == "mhd_entropy":
        P, zprof, grid = pk.setups.config_mhd_entropy(20)
    elif case == "mhd_entropy_upwind":
        P, zprof, grid = pk.setups.config_mhd_entropy(16, lupw_ss=1, lupw_uu=1, lupw_lnrho=1)
    else:
        P, zprof, grid = pk.setups.config_convslab(16, 16, 16)
    assert orc.pencil_applies(P)
    orc.set_zprofiles(zprof)
    f0 = pk.setups.synthetic_fields(P, grid)
    if case == "convslab":
        f0[P.ilnrho - 1] *= 0.1
    frc = pk.setups.helical_forcing_coefs((3.0, -1.0, 2.0), 0.3, 1.1, 1e-2)
    fa, fb = f0.copy(), f0.copy()
    dfa, dfb = np.zeros_like(f0), np.full_like(f0, np.nan)     # the pencil version must not read stale df in substep 1
    dta = dtb = 0.0
    for _ in range(3):
        dta = orc.timestep(fa, dfa, P, dta, frc)
        dtb = orc.timestep_pencil(fb, dfb, P, dtb, frc)
        assert abs(dta - dtb) <= 1e-13 * dta
    inner = (slice(None), slice(3, -3), slice(3, -3), slice(3, -3))
    for k in range(P.mvar):
        den = np.abs(fa[k][inner[1:]]).max()
        assert np.abs(fa[k][inner[1:]] - fb[k][inner[1:]]).max() <= 1e-12 * den


def test_pencil_version_falls_back_where_it_does_not_apply(pk, orc):
    P, zprof, grid = pk.setups.config_mhd_entropy(16, lbidiagonal_derij=0)
    assert not orc.pencil_applies(P)
    orc.set_zprofiles(zprof)
    f0 = pk.setups.synthetic_fields(P, grid)
    fa, fb = f0.copy(), f0.copy()
    dfa, dfb = np.zeros_like(f0), np.zeros_like(f0)
    assert orc.timestep(fa, dfa, P, 0.0) == orc.timestep_pencil(fb, dfb, P, 0.0)
    assert np.array_equal(fa, fb)
