"""RESTART from a snapshot the reference itself wrote: samples/backwards-compatible ships data/proc0/var.dat (4^3 forced
helical MHD at t = 2.625, single precision, with the persistent block that holds the nr_f90 random-number state) and the
reference.out of the 10 steps run.x makes from it (HYDRO+DENSITY+noentropy+MAGNETIC+FORCING, gamma=1, nu=eta=0.5,
cdt=0.4, iforce='helical', force=0.07; the sample exists "to make sure we notice when a new version of the code can no
longer restart from old snapshot").  Nothing of the initial state is restated here: the fields, the time and the
generator state all come out of the reference's file through pencil-code_b200/snapshot.py, so this pins
  * the snapshot reader incl. the persistent block (io_dist.f90:535-568, :772-846; persist.f90:156-160),
  * the diagnostics of a state the reference produced (row it=0 needs no time step at all),
  * 8 forced MHD steps in the diffusion-dominated regime (dt from sqrt((advec/cdt)^2 + (diffus/cdtv)^2), equ.f90:1149),
    with the forcing wave vectors drawn from the restored stream,
on a grid where every ghost zone wraps onto 3 of the 4 interior planes.
Columns: t, dt, umax, urms, brms, jrms, jmax, abm, jbm, oum, rhom, all five rows; tolerance 1.5 units of the last
printed digit.  The file is single precision, the checker double: the printed 3-4 digits do not see the difference."""
import math
import os

import numpy as np

from tests.test_oracle_helical_mhdturb import printed_tolerance

HERE = os.path.dirname(os.path.abspath(__file__))
COLS = {"t": 1, "dt": 2, "umax": 3, "urms": 4, "brms": 5, "jrms": 6, "jmax": 7, "abm": 11, "jbm": 12, "oum": 13, "rhom": 14}


def test_oracle_continues_the_reference_run_from_its_own_var_dat(pk, orc):
    from oracle import nr_f90
    P, zprof, grid = pk.setups.make_params(4, 4, 4, magnetic=True, nu=0.5, eta=0.5, cdt=0.4, cdtv=0.25)
    path = os.path.join(HERE, "golden", "backwards-compatible.var.dat")
    snap = pk.snapshot.read_var(path, 10, 10, 10, 7, dtype=np.float32)
    assert abs(snap["t"] - 2.62457) < 1e-5 and abs(snap["dx"] - math.pi / 2) < 1e-6
    seed = np.frombuffer(pk.snapshot.read_persist(path)[pk.snapshot.ID_RECORD_RANDOM_SEEDS], dtype=np.int32)
    assert seed.size == 2 and seed[1] != 0
    # file order = the reference's registration order ux,uy,uz,lnrho,ax,ay,az (data/proc0/dim.dat: mvar=7)
    src = {P.iux - 1: 0, P.iux: 1, P.iux + 1: 2, P.ilnrho - 1: 3, P.iax - 1: 4, P.iax: 5, P.iax + 1: 6}
    f = np.zeros((P.mvar, 10, 10, 10))
    for slot, k in src.items():
        f[slot] = snap["f"][k]
    f = np.ascontiguousarray(np.pad(f[:, 3:-3, 3:-3, 3:-3], ((0, 0), (3, 3), (3, 3), (3, 3)), mode="wrap"))   # update_ghosts
    df = np.zeros_like(f)
    stream = nr_f90.MarsRan()
    stream.put([int(seed[0]), int(seed[1])])                      # a saved state restores the sequence (general.f90:571-597)
    nk, kav, kkx, kky, kkz = nr_f90.read_kdat(os.path.join(HERE, "golden", "helical-MHDturb.k.dat"))   # the sample's k.dat is this file
    gold = [ln.split() for ln in open(os.path.join(HERE, "golden", "backwards-compatible.reference.out")) if not ln.startswith("#")]
    nw, rows, t, dt = P.nx * P.ny * P.nz, [], snap["t"], 0.0
    for it in range(9):
        d = np.zeros(16)
        dt_step = orc.timestep(f.copy(), df.copy(), P, dt, None, d)
        if it % 2 == 0:
            n = dict(zip(pk.abi.DIAG_NAMES, d))
            rows.append({"t": t, "dt": dt_step, "umax": math.sqrt(n["u2max"]), "urms": math.sqrt(n["u2sum"] / nw),
                         "brms": math.sqrt(n["b2sum"] / nw), "jrms": math.sqrt(n["j2sum"] / nw), "jmax": math.sqrt(n["j2max"]),
                         "abm": n["absum"] / nw, "jbm": n["jbsum"] / nw, "oum": n["ousum"] / nw, "rhom": n["rhosum"] / nw})
        phase, ik, phi = nr_f90.helical_draw(stream, nk)
        frc = pk.setups.helical_forcing_coefs((kkx[ik - 1], kky[ik - 1], kkz[ik - 1]), phase, phi, dt_step,
                                              force=0.07, relhel=1.0, kav=kav)
        dt = orc.timestep(f, df, P, dt, frc, None)
        t += dt
    assert len(rows) == len(gold) == 5
    for got, ref in zip(rows, gold):
        for name, j in COLS.items():
            assert abs(got[name] - float(ref[j])) <= printed_tolerance(ref[j]), (ref[0], name, got[name], ref[j])
