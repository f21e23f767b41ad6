"""Initial condition of samples/conv-slab, restated from the reference's start.x
modules (TEST INFRASTRUCTURE: used to pin the oracle against the reference's own
golden time series samples/conv-slab/reference.out.double).

  lnrho: 'piecew-poly'  density.f90:1315-1345 + polytropic_lnrho_z (density.f90, same file)
  ss   : 'piecew-poly'  entropy.f90:1628-1650 + polytropic_ss_z (entropy.f90:1849-1925)
  uu   : 'up-down'      hydro.f90:2189-2203
Parameters: samples/conv-slab/start.in.

Also: initss='mixinglength' of samples/2d-tests/conv-slab-MLT (entropy.f90:2108-2218 mixinglength, :7678-7741 strat_MLT,
sub.f90:6092-6132 interp1), see mixinglength_stratification below.
"""
import math

import numpy as np

EPSI = 5 * np.finfo(np.float64).eps      # cparam.f90:118
TINI = 5 * np.finfo(np.float64).tiny


def _step(x, x0, width):                 # sub.f90:4544 step_vector
    return 0.5 * (1 + np.tanh((x - x0) / (width + TINI)))


def convslab_initial_condition(P, grid, *, cs0=0.57735, gamma=1.6666666, rho0=1.0, gravz=-1.0, z1=0.0, z2=1.0,
                               zref=1.32, mpoly0=1.0, mpoly1=3.0, mpoly2=0.0, isothtop=1, widthlnrho=0.05,
                               widthss=0.05, ampluu=0.1, widthuu=0.1, fac_cs=1.0, isothmid=0):
    x, y, z = grid["x"], grid["y"], grid["zglobal"] if "zglobal" in grid and P.nprocz == 1 else grid["z"]
    mx, my, mz = x.size, y.size, z.size
    x0, y0, z0 = P.xyz0[0], P.xyz0[1], P.xyz0[2]
    Lx, Ly, Lz = P.Lxyz[0], P.Lxyz[1], P.Lxyz[2]
    ztop = z0 + Lz
    gamma_m1 = gamma - 1.0
    cp1 = 1.0 / P.cp
    lnrho0 = math.log(rho0)
    f = np.zeros((P.mvar, mz, my, mx))

    # ---- lnrho: density.f90:1315-1345 ------------------------------------
    lnr = np.full(mz, lnrho0)
    st = {"cs2int": cs0 ** 2, "lnrhoint": lnrho0}

    def poly_lnrho(mpoly, zint, zbot, zblend, isoth, fac=None):
        stp = _step(z, zblend, widthlnrho)
        cs2int, lnrhoint = st["cs2int"], st["lnrhoint"]
        if isoth != 0:
            beta1 = 0.0
            tmp = lnrhoint + gamma * gravz / cs2int * (z - zint)
        else:
            beta1 = gamma * gravz / (mpoly + 1)
            tmp = 1.0 + beta1 * (z - zint) / cs2int
            tmp = np.maximum(tmp, EPSI)
            tmp = lnrhoint + mpoly * np.log(tmp)
        lnr[:] = stp * lnr + (1 - stp) * tmp
        if isoth != 0:
            st["lnrhoint"] = lnrhoint + gamma * gravz / cs2int * (zbot - zint)
        else:
            st["lnrhoint"] = lnrhoint + mpoly * math.log(1 + beta1 * (zbot - zint) / cs2int)
        if isoth != 0 and fac is not None:
            st["cs2int"] = fac ** 2 * cs2int
        else:
            st["cs2int"] = cs2int + beta1 * (zbot - zint)

    poly_lnrho(mpoly2, zref, z2, ztop + Lz, isothtop, fac_cs)      # top region
    poly_lnrho(mpoly0, z2, z1, z2, isothmid)                       # unstable layer
    poly_lnrho(mpoly1, z1, z0, z1, 0)                              # stable layer

    # ---- ss: entropy.f90:1628-1650 -----------------------------------------
    ssz = np.zeros(mz)
    se = {"cs2int": cs0 ** 2, "ssint": 0.0}

    def poly_ss(mpoly, zint, zbot, zblend, isoth, fac=None):
        stp = _step(z, zblend, widthss)
        cs2int, ssint = se["cs2int"], se["ssint"]
        if isoth != 0:
            beta1 = 0.0
            tmp = ssint - gamma_m1 * gravz * (z - zint) / cs2int / cp1
        else:
            beta1 = gamma * gravz / (mpoly + 1)
            tmp = 1.0 + beta1 * (z - zint) / cs2int
            tmp = np.maximum(tmp, EPSI)
            tmp = ssint + (1 - mpoly * gamma_m1) / gamma * np.log(tmp) / cp1
        ssz[:] = stp * ssz + (1 - stp) * tmp
        if isoth != 0:
            if fac is not None:
                se["ssint"] = ssint - gamma_m1 * gravz * (zbot - zint) / cs2int / cp1 + 2 * math.log(fac) / gamma / cp1
            else:
                se["ssint"] = ssint - gamma_m1 * gravz * (zbot - zint) / cs2int / cp1
        else:
            se["ssint"] = ssint + (1 - mpoly * gamma_m1) / gamma * math.log(1 + beta1 * (zbot - zint) / cs2int) / cp1
        if isoth != 0 and fac is not None:
            se["cs2int"] = fac ** 2 * cs2int
        else:
            se["cs2int"] = cs2int + beta1 * (zbot - zint)

    poly_ss(mpoly2, zref, z2, z0 + 2 * Lz, isothtop, fac_cs)
    poly_ss(mpoly0, z2, z1, z2, isothmid)
    poly_ss(mpoly1, z1, z0, z1, 0)

    f[P.ilnrho - 1] = lnr[:, None, None]
    f[P.iss - 1] = ssz[:, None, None]

    # ---- uz: 'up-down', hydro.f90:2189-2203 ---------------------------------
    prof = ampluu * np.exp(-0.5 * (z - z1) ** 2 / widthuu ** 2)
    X, Y = x[None, :], y[:, None]
    t1 = np.sqrt((X - (x0 + 0.3 * Lx)) ** 2 + (Y - (y0 + 0.3 * Ly)) ** 2)
    t2 = np.sqrt((X - (x0 + 0.5 * Lx)) ** 2 + (Y - (y0 + 0.8 * Ly)) ** 2)
    uz = prof[:, None, None] * np.exp(-0.5 * t1 ** 2 / widthuu ** 2)[None]
    uz = uz - 0.7 * prof[:, None, None] * np.exp(-0.5 * t2 ** 2 / widthuu ** 2)[None]
    f[P.iuz - 1] = uz
    # start.x fills only the interior; ghost zones are set by the boundary conditions in run.x
    g = np.zeros_like(f)
    g[:, 3:-3, 3:-3, 3:-3] = f[:, 3:-3, 3:-3, 3:-3]
    return g


# ---------------------------------------------------------------------------
# samples/interlocked-fluxrings: initial_condition/fluxrings.f90
# ---------------------------------------------------------------------------
def _norm_ring(xx1, yy1, zz1, fring, Iring, R0, width):
    """norm_ring, profile 'tanh' (initial_condition/fluxrings.f90:318-371): vector potential of a
    canonical ring lying in the x-y plane."""
    tmp = np.sqrt(xx1 ** 2 + yy1 ** 2) - R0
    v3 = -fring * 0.5 * (1 + np.tanh(tmp / width)) * 0.5 / width / np.cosh(zz1 / width) ** 2
    tmp = width - np.sqrt(tmp ** 2 + zz1 ** 2)
    tmp = Iring * 0.5 * (1 + np.tanh(tmp / width))
    phi = np.arctan2(yy1, xx1)
    return -tmp * np.sin(phi), tmp * np.cos(phi), v3


def fluxrings_initial_condition(P, grid, *, amplaa=1.0,
                                rings=((1.0, 0.0, 1.0, 0.3, (0.0, 0.0, 1.0), (0.0, -0.5, 0.0)),
                                       (1.0, 0.0, 1.0, 0.3, (1.0, 0.0, 0.0), (0.0, 0.5, 0.0)))):
    """`fluxrings` (initial_condition/fluxrings.f90:205-298) with the parameters of
    samples/interlocked-fluxrings/start.in: AA(x) = D AA0(D^-1 (x - disp)), D = rotation by phi about z
    followed by theta about y.  uu=0, lnrho=0 (start.in sets no init for hydro/density).
    rings: (fring, Iring, Rring, wr, axis, disp) per ring."""
    x, y, z = grid["x"], grid["y"], grid["z"]
    f = np.zeros((P.mvar, z.size, y.size, x.size))
    X, Y, Z = x[None, None, :], y[None, :, None], z[:, None, None]
    ia = P.iax - 1
    for fring, Iring, R0, width, axis, disp in rings:
        phi = math.atan2(axis[1], axis[0] + EPSI)
        theta = math.atan2(math.sqrt(axis[0] ** 2 + axis[1] ** 2) + EPSI, axis[2])
        ct, st, cp, sp = math.cos(theta), math.sin(theta), math.cos(phi), math.sin(phi)
        xx1 = ct * cp * (X - disp[0]) + ct * sp * (Y - disp[1]) - st * (Z - disp[2])
        yy1 = -sp * (X - disp[0]) + cp * (Y - disp[1]) + 0.0 * Z
        zz1 = st * cp * (X - disp[0]) + st * sp * (Y - disp[1]) + ct * (Z - disp[2])
        v1, v2, v3 = _norm_ring(xx1, yy1, zz1, fring, Iring, R0, width)
        f[ia] += amplaa * (ct * cp * v1 - sp * v2 + st * cp * v3)
        f[ia + 1] += amplaa * (ct * sp * v1 + cp * v2 + st * sp * v3)
        f[ia + 2] += amplaa * (-st * v1 + ct * v3)
    g = np.zeros_like(f)                     # start.x fills the interior only
    g[:, 3:-3, 3:-3, 3:-3] = f[:, 3:-3, 3:-3, 3:-3]
    return g


# ---------------------------------------------------------------------------
# samples/2d-tests/conv-slab-MLT: initss='mixinglength'
# ---------------------------------------------------------------------------
def _interp1_descending(r, fr, r0):
    """interp1(r,fr,nr,r0,ldescending=.true.) (sub.f90:6092-6132): linear interpolation in the reversed arrays"""
    r1, fr1 = r[::-1], fr[::-1]
    if r0 == r1[0]:
        return fr1[0]
    if r0 > r1[-1]:
        return fr1[-1]
    i = 0
    while True:                      # do while (istop /= 1): if (r1(i) >= r0) istop=1; i=i+1
        stop = r1[i] >= r0
        i += 1
        if stop:
            break
    i1, i2 = i - 2, i - 1            # Fortran i-2, i-1 with the 1-based i one past the hit
    return (fr1[i1] * (r1[i2] - r0) + fr1[i2] * (r0 - r1[i1])) / (r1[i2] - r1[i1])


def strat_MLT(rhotop, flux, *, nzgrid, ztop, dz, z1, z2, gravz, gamma, cs2top, mpoly0, mpoly1, alpha_MLT=1.5):
    """entropy.f90:7678-7741: integrate temperature and ln rho downwards from the top with the mixing-length gradient
    del = delad + alpha_MLT*(Fconv/(rho cs^3))^(2/3) in z1 <= z <= z2, the polytrope 1/(mpoly1+1) below, isothermal above.
    -> lnrhom, tempm (index 0 = top), rho at z1"""
    gamma_m1 = gamma - 1.0
    lnrhom, tempm, zz = np.zeros(nzgrid), np.zeros(nzgrid), np.zeros(nzgrid)
    lnrhom[0], tempm[0], zz[0] = math.log(rhotop), cs2top / gamma_m1, ztop
    polyad, delad = 1.0 / gamma_m1, 1.0 - 1.0 / gamma
    fc = (1.0 - delad * (mpoly0 + 1.0)) * flux
    for iz in range(1, nzgrid):
        zm = ztop - iz * dz
        zz[iz] = zm
        if zm < z1:
            dl = 1.0 / (mpoly1 + 1.0)
        elif zm <= z2:
            dl = delad + alpha_MLT * (fc / (math.exp(lnrhom[iz - 1]) * (gamma_m1 * tempm[iz - 1]) ** 1.5)) ** (2.0 / 3.0)
        else:
            dl = 0.0
        dtemp = gamma * polyad * gravz * dl
        dlnrho = gamma * polyad * gravz * (1.0 - dl) / tempm[iz - 1]
        tempm[iz] = tempm[iz - 1] - dtemp * dz
        lnrhom[iz] = lnrhom[iz - 1] - dlnrho * dz
    return lnrhom, tempm, math.exp(_interp1_descending(zz, lnrhom, z1))


def mixinglength_stratification(flux, *, nzgrid, z0, Lz, z1, z2, gravz, gamma, cs20, mpoly0, mpoly1, rho0=1.0, cp=1.0):
    """mixinglength (entropy.f90:2108-2218): secant iteration on rho at the top until rho(z1) = rho0, then rho0 := rhotop
    (the value run.x reads back from param.nml, written after the initial condition, start.f90:562).
    -> lnrho(z), ss(z) bottom to top, the new lnrho0"""
    kw = dict(nzgrid=nzgrid, ztop=z0 + Lz, dz=Lz / (nzgrid - 1), z1=z1, z2=z2, gravz=gravz, gamma=gamma, cs2top=cs20,
              mpoly0=mpoly0, mpoly1=mpoly1)
    rbot, rt_old, rt_new = rho0, 0.1 * rho0, 0.12 * rho0
    rb_old = strat_MLT(rt_old, flux, **kw)[2]
    lnrhom, tempm, rb_new = strat_MLT(rt_new, flux, **kw)
    for _ in range(10):
        rhotop = rt_old + (rt_new - rt_old) / (rb_new - rb_old) * (rbot - rb_old)
        if abs(rhotop / rt_new - 1.0) <= 1e-4:
            break
        lnrhom, tempm, rhobot = strat_MLT(rhotop, flux, **kw)
        rt_old, rb_old, rt_new, rb_new = rt_new, rb_new, rhotop, rhobot
    lnrho0 = math.log(rhotop)
    gamma_m1, cv = gamma - 1.0, cp / gamma
    lnTT0 = math.log(cs20 / (cp * gamma_m1))
    lnrho = lnrhom[::-1].copy()                                    # f(n) = lnrhom(nzgrid-iz+1)
    ss = cv * (np.log(tempm[::-1]) - lnTT0 - gamma_m1 * (lnrho - lnrho0))       # eoscalc(ilnrho_lnTT), eos_idealgas.f90
    return lnrho, ss, lnrho0
