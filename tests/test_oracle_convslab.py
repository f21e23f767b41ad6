"""PIN of the oracle against the reference's own golden vector:
samples/conv-slab/reference.out.double (13-14 digit time series of the real
Fortran/MPI build; copied to tests/golden/).  The initial condition is restated
in oracle/initcond.py, the run is the oracle's pcref_timestep with the
parameters of samples/conv-slab/{start.in,run.in}.  Columns compared:
t, dt, urms, umax, rhom, ssm, dtc, dtu, dtnu, dtchi at it=0,10,20,30,40."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def read_golden():
    rows = []
    for line in open(os.path.join(HERE, "golden", "conv-slab.reference.out.double")):
        if line.startswith("#") or not line.strip():
            continue
        rows.append([float(v) for v in line.split()])
    return np.array(rows)


def oracle_time_series(pk, orc, nsteps=41):
    from oracle import initcond
    P, zprof, grid = pk.setups.config_convslab(32, 32, 32)
    f = initcond.convslab_initial_condition(P, grid)
    orc.set_zprofiles(zprof)
    df = np.zeros_like(f)
    nw = P.nx * P.ny * P.nz
    out, t, dt = [], 0.0, 0.0
    for it in range(nsteps):
        d = np.zeros(16)
        dt = orc.timestep(f, df, P, dt, None, d)
        n = dict(zip(pk.abi.DIAG_NAMES, d))
        if it % 10 == 0:
            out.append([it, t, dt, np.sqrt(n["u2sum"] / nw), np.sqrt(n["u2max"]), n["rhosum"] / nw, n["sssum"] / nw,
                        dt * n["dtc"], dt * n["dtu"], dt * n["dtnu"], dt * n["dtchi"]])
        t += dt
    return np.array(out)


def test_oracle_reproduces_reference_out_double(pk, orc):
    gold = read_golden()
    got = oracle_time_series(pk, orc)
    assert got.shape == gold.shape
    names = ["it", "t", "dt", "urms", "umax", "rhom", "ssm", "dtc", "dtu", "dtnu", "dtchi"]
    worst = 0.0
    for j, nm in enumerate(names):
        for i in range(gold.shape[0]):
            ref, val = gold[i, j], got[i, j]
            rel = abs(val - ref) / max(abs(ref), 1e-300) if ref != 0 else abs(val)
            worst = max(worst, rel)
            # the reference harness accepts 1.5 units of the last printed digit
            # (lib/perl/Test/NumericFileComparator.pm:152,595-615); 13 digits are printed
            assert rel < 2e-11, (nm, i, val, ref, rel)
    print("worst relative deviation from reference.out.double:", worst)
