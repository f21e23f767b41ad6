"""PIN of the entropy path with iheatcond='chi-const', viscous heating, the x/y wall boundary conditions 'a'/'s' and
degenerate directions against four more golden vectors of the reference: samples/1d-tests/sod_10, sod_100, sod_1000,
sod_10y/reference.out -- Sod shock tubes on 400 points (HYDRO+DENSITY+ENTROPY, nu-const + chi-const, no shock
viscosity), printed every 100 time steps with 4-5 digits.  start.in: initlnrho/initss='xjump' ('yjump'):
f = left + (right-left)*(1+tanh(x/width))/2 (initcond.f90:1357-1387), lnrho_left=log(rho_left) (density.f90:1019);
run.in: bcx (bcy) = 'a','s','s','s','s', cdt=0.4, cdtv=0.25, chi, nu.  Directions with one grid point have dy_1=0.

Columns checked: t, dt, urms, rhom, ssm for every row of each file (up to 1000 time steps); tolerance: 1.5 units of
the last printed digit (the reference harness).  The other columns (ekin, ethm, ruxm, ...) are not among the 16
diagnostics of the hot path."""
import math
import os
import re

import numpy as np
import pytest

from tests.test_oracle_helical_mhdturb import printed_tolerance

HERE = os.path.dirname(os.path.abspath(__file__))

#        name        axis  x0    rho_left ss_left  chi    nu
CASES = [("sod_10", 0, -4.0, 10.0, -0.615, 5.0e-4, 0.02),
         ("sod_100", 0, -3.7, 100.0, -1.536, 1.0e-3, 0.04),
         ("sod_1000", 0, -3.7, 1.0e3, -2.46, 5.0e-4, 0.08),
         ("sod_10y", 1, -3.7, 10.0, -0.615, 5.0e-4, 0.02)]


def read_table(name):
    lines = open(os.path.join(HERE, "golden", name + ".reference.out")).read().splitlines()
    head = [h for h in re.split(r"-+", lines[0].lstrip("#")) if h]
    rows = []
    for ln in lines[1:]:
        if ln.strip():
            # columns can touch when a value is negative ("3.739E+01-3.761E-01")
            rows.append(re.findall(r"[-+]?\d+\.?\d*(?:[eE][-+]?\d+)?", ln))
    return head, rows


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_oracle_reproduces_sod_reference_out(pk, orc, case):
    name, axis, x0, rho_left, ss_left, chi, nu = case
    n = [1, 1, 1]; n[axis] = 400
    xyz0 = [0.0, 0.0, 0.0]; xyz0[axis] = x0
    L = [1.0, 1.0, 1.0]; L[axis] = 10.0
    lperi = [1, 1, 1]; lperi[axis] = 0
    bc = {("bcx", "bcy", "bcz")[axis]: ["a", "s", "s", "s", "s"]}
    P, zprof, grid = pk.setups.make_params(n[0], n[1], n[2], entropy=True, xyz0=tuple(xyz0), Lxyz=tuple(L), lperi=tuple(lperi),
                                           nu=nu, chi=chi, iheatcond=pk.abi.HCOND["chi-const"], cdt=0.4, cdtv=0.25,
                                           gamma=5.0 / 3.0, **bc)
    if axis != 0: P.dx_1 = 0.0          # one grid point: no derivative along that direction (grid.f90 construct_grid)
    if axis != 1: P.dy_1 = 0.0
    P.dz_1 = 0.0
    x = grid["xyz"[axis]]
    f = np.zeros((P.mvar, P.nz + 6, P.ny + 6, P.nx + 6))
    shape = [1, 1, 1]; shape[2 - axis] = x.size
    prof = 0.5 * (1.0 + np.tanh(x / 0.03)).reshape(shape)
    f[P.ilnrho - 1] += math.log(rho_left) + (0.0 - math.log(rho_left)) * prof          # rho_right = 1
    f[P.iss - 1] += ss_left + (0.306 - ss_left) * prof
    orc.set_zprofiles(zprof)
    head, gold = read_table(name)
    col = {h: j for j, h in enumerate(head)}
    df = np.zeros_like(f)
    nw, t, dt = 400, 0.0, 0.0
    last = int(gold[-1][0])
    k = 0
    for it in range(last + 1):
        d = np.zeros(16) if it % 100 == 0 else None
        dt = orc.timestep(f, df, P, dt, None, d)
        if d is not None:
            v = dict(zip(pk.abi.DIAG_NAMES, d))
            got = {"t": t, "dt": dt, "urms": math.sqrt(v["u2sum"] / nw), "rhom": v["rhosum"] / nw, "ssm": v["sssum"] / nw}
            ref = gold[k]
            assert int(ref[0]) == it
            for nm, val in got.items():
                tok = ref[col[nm]]
                assert abs(val - float(tok)) <= printed_tolerance(tok), (name, it, nm, val, tok)
            k += 1
        t += dt
    assert k == len(gold)
