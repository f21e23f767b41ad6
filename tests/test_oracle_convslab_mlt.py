"""A second convection fixture of the reference: samples/2d-tests/conv-slab-MLT/reference.out -- 2-D (x,z) compressible
convection at 64x1x64 (HYDRO+DENSITY+ENTROPY+gravity_simple, single precision, 50 steps, printed every 10):
  start.in  initss='mixinglength' (a mixing-length stratification integrated downwards from the top and iterated until
            rho(z1)=rho0; it re-defines rho0, which run.x reads back), inituu='gaussian-noise' (default random_gen
            'min_std'); the 'blob' listed as initss(2) has amplitude ampl_ss(2)=0 (start.in sets ampl_ss(1) only) and
            adds nothing;
  run.in    iheatcond='K-profile' with hcond0 from mixinglength_flux (entropy.f90:669-675: three layers, hcond1=4,
            hcond2=0.1), cooling_profile='cubic_step' (entropy.f90:6250), cool=10, wcool=0.1, upwinding of lnrho and ss,
            nu=5e-3, cdt=0.25, bcz = s,s,a,a2,c1:cT.
Everything the kernels need beyond samples/conv-slab: a one-point y direction together with the entropy module, a
different cooling profile (profiles reach the device as arrays, so this is host side), lnrho0 /= 0 in the equation of
state, three conductivity layers with a strongly sub-adiabatic top.  Columns: t, dt, urms, rhom, ssm, dtc, dtnu, dtchi
of all five rows; tolerance 1.5 units of the last printed digit."""
import math
import os

import numpy as np

from tests.test_oracle_helical_mhdturb import printed_tolerance

HERE = os.path.dirname(os.path.abspath(__file__))
COLS = {"t": 1, "dt": 2, "urms": 3, "rhom": 4, "ssm": 5, "dtc": 6, "dtnu": 7, "dtchi": 8}


def test_oracle_reproduces_convslab_mlt_reference_out(pk, orc):
    from oracle import initcond, nr_f90
    abi = pk.abi
    gamma, cs20 = 5.0 / 3.0, 0.57735 ** 2
    flux, mpoly0, mpoly1, mpoly2, gravz = 1e-2, 0.0, 3.0, -0.9, -1.0
    nxg, nzg, z0, Lz, z1, z2 = 64, 64, -2.0, 3.2, 0.0, 1.0
    lnrho, ss, lnrho0 = initcond.mixinglength_stratification(flux, nzgrid=nzg, z0=z0, Lz=Lz, z1=z1, z2=z2, gravz=gravz,
                                                             gamma=gamma, cs20=cs20, mpoly0=mpoly0, mpoly1=mpoly1)
    P, zprof, grid = pk.setups.make_params(
        nxg, 1, nzg, entropy=True, xyz0=(-2.0, 0.0, z0), Lxyz=(4.0, 1.0, Lz), lperi=(1, 1, 0), nu=5e-3, cdt=0.25, cdtv=0.25,
        lgravz=1, gravz=gravz, lupw_lnrho=1, lupw_ss=1, cs20=cs20, gamma=gamma, lnrho0=lnrho0,
        iheatcond=abi.HCOND["K-profile"], hcond0=-flux * (mpoly0 + 1.0) * (gamma - 1.0) / gamma / gravz,       # entropy.f90:671
        hcond1=(mpoly1 + 1.0) / (mpoly0 + 1.0), hcond2=(mpoly2 + 1.0) / (mpoly0 + 1.0), z1=z1, z2=z2, widthss=0.05,
        lcooling=1, cool=10.0, wcool=0.1, cs2cool=cs20, cs2top=cs20, Fbot=flux, bcz=["s", "s", "a", "a2", "c1:cT"])
    P.dy_1 = 0.0                                                   # nygrid=1 (grid.f90:481-489)
    xi = np.clip((grid["z"][3:-3] - z2) / (P.wcool + initcond.TINI), -1.0, 1.0)
    zprof[abi.PROF_COOL] = 0.5 + xi * (0.75 - xi ** 2 * 0.25)     # cubic_step (sub.f90:4646-4671)
    f = np.zeros((P.mvar, nzg + 6, 7, nxg + 6))
    nr_f90.gaunoise_vect(nr_f90.start_stream_min_std(0), f, 3e-4, P.iux - 1, P.iuz - 1)
    f[P.ilnrho - 1, 3:-3] = lnrho[:, None, None]
    f[P.iss - 1, 3:-3] = ss[:, None, None]
    g = np.zeros_like(f)
    g[:, 3:-3, 3:-3, 3:-3] = f[:, 3:-3, 3:-3, 3:-3]               # run.x sets the ghost zones
    f = g
    orc.set_zprofiles(zprof)
    gold = [ln.split() for ln in open(os.path.join(HERE, "golden", "conv-slab-MLT.reference.out")) if not ln.startswith("#")]
    df = np.zeros_like(f)
    nw, t, dt, rows = nxg * nzg, 0.0, 0.0, []
    for it in range(41):
        d = np.zeros(16) if it % 10 == 0 else None
        dt = orc.timestep(f, df, P, dt, None, d)
        if d is not None:
            n = dict(zip(abi.DIAG_NAMES, d))
            rows.append({"t": t, "dt": dt, "urms": math.sqrt(n["u2sum"] / nw), "rhom": n["rhosum"] / nw, "ssm": n["sssum"] / nw,
                         "dtc": dt * n["dtc"], "dtnu": dt * n["dtnu"], "dtchi": dt * n["dtchi"]})
        t += dt
    assert len(rows) == len(gold) == 5
    for got, ref in zip(rows, gold):
        for name, j in COLS.items():
            assert abs(got[name] - float(ref[j])) <= printed_tolerance(ref[j]), (ref[0], name, got[name], ref[j])
