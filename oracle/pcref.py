"""ctypes binding of the CPU ORACLE libpcref.so (test infrastructure, NOT product
code: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs import this)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "libpcref.so")


def build(force=False):
    srcs = [os.path.join(_HERE, n) for n in ("pcref.c", "pcref_particles.c", "pcref_pencil.c", "pcref.h")]
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < max(os.path.getmtime(q) for q in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return LIB


FAST_LIB = os.path.join(_HERE, "libpcref_fast.so")


def build_fast():
    """-O3 -march=native build for TIMING the CPU arm (bench.py); compiled on the machine that runs it.
    Returns None when it cannot be built (the caller then times the checker build and says so)."""
    try:
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "fast"])
        return FAST_LIB
    except Exception:
        return None


class Oracle:
    def __init__(self, abi, lib_path=None):
        """abi: the product's ctypes module (only for the shared struct layout).
        lib_path: another build of the same sources (bench.py's timing build)."""
        build()
        self.abi = abi
        self.lib_path = lib_path or LIB
        L = self.lib = C.CDLL(self.lib_path)
        d, i, p = C.c_double, C.c_int, C.c_void_p
        for name in ("pcref_der_field", "pcref_der2_field", "pcref_der6_field"):
            getattr(L, name).argtypes = [p, i, i, i, i, d, p]
            getattr(L, name).restype = None
        L.pcref_derij_field.argtypes = [p, i, i, i, i, i, d, d, p]
        L.pcref_derij_field.restype = None
        L.pcref_derij_full_field.argtypes = [p, i, i, i, i, i, d, d, p]
        L.pcref_derij_full_field.restype = None
        L.pcref_boundconds.argtypes = [p, C.POINTER(abi.Params)]
        L.pcref_boundconds.restype = None
        L.pcref_rhs.argtypes = [p, p, C.POINTER(abi.Params), i, C.POINTER(d), p]
        L.pcref_rhs.restype = None
        L.pcref_timestep.argtypes = [p, p, C.POINTER(abi.Params), C.POINTER(abi.Forcing), C.POINTER(d), p]
        L.pcref_timestep.restype = None
        L.pcref_timestep_pencil.argtypes = [p, p, C.POINTER(abi.Params), C.POINTER(abi.Forcing), C.POINTER(d), p]
        L.pcref_timestep_pencil.restype = None
        L.pcref_pencil_applies.argtypes = [C.POINTER(abi.Params)]
        L.pcref_pencil_applies.restype = i
        L.pcref_addforce.argtypes = [p, C.POINTER(abi.Params), C.POINTER(abi.Forcing)]
        L.pcref_addforce.restype = None
        L.pcref_set_zprofile.argtypes = [i, p, i]
        L.pcref_set_zprofile.restype = None
        L.pcref_num_threads.restype = i
        L.pcref_set_num_threads.argtypes = [i]
        L.pcref_set_num_threads.restype = None
        PP, PA = C.POINTER(abi.ParticleParams), C.POINTER(abi.Params)
        L.pcref_particles_boundconds.argtypes = [p, PA, PP]
        L.pcref_particles_map_nearest_grid.argtypes = [p, p, PA, PP]
        L.pcref_setup_imn_array.argtypes = [PA, p]
        L.pcref_particles_sort.argtypes = [p, p, p, p, p, p, PA, PP]
        L.pcref_particles_map_xxp_grid.argtypes = [p, p, p, p, PA, PP]
        L.pcref_particles_pde.argtypes = [p, p, p, p, p, PA, PP, i, C.POINTER(d)]
        L.pcref_timestep_particles.argtypes = [p, p, PA, C.POINTER(abi.Forcing), C.POINTER(d), p, PP, p, p, p, p, p, p, p, p]
        L.pcref_particles_fixture.argtypes = [i, p, C.c_long, d, p]
        L.pcref_particles_fixture.restype = None
        for name in ("pcref_particles_boundconds", "pcref_particles_map_nearest_grid", "pcref_setup_imn_array",
                     "pcref_particles_sort", "pcref_particles_map_xxp_grid", "pcref_particles_pde",
                     "pcref_timestep_particles"):
            getattr(L, name).restype = None

    def set_zprofiles(self, zprof):
        for which, arr in (zprof or {}).items():
            a = np.ascontiguousarray(arr, dtype=np.float64)
            self.lib.pcref_set_zprofile(int(which), a.ctypes.data, a.size)

    def _field(self, fn, f, nx, ny, nz, *args):
        out = np.empty((nz, ny, nx))
        fn(f.ctypes.data, nx, ny, nz, *args, out.ctypes.data)
        return out

    def der(self, f, j, d_1):
        nz, ny, nx = (s - 6 for s in f.shape)
        return self._field(self.lib.pcref_der_field, f, nx, ny, nz, j, d_1)

    def der2(self, f, j, d_1):
        nz, ny, nx = (s - 6 for s in f.shape)
        return self._field(self.lib.pcref_der2_field, f, nx, ny, nz, j, d_1)

    def der6(self, f, j, d_1):
        nz, ny, nx = (s - 6 for s in f.shape)
        return self._field(self.lib.pcref_der6_field, f, nx, ny, nz, j, d_1)

    def derij(self, f, i, j, di_1, dj_1):
        nz, ny, nx = (s - 6 for s in f.shape)
        return self._field(self.lib.pcref_derij_field, f, nx, ny, nz, i, j, di_1, dj_1)

    def derij_full(self, f, i, j, di_1, dj_1):
        """lbidiagonal_derij=F (36-point product stencil)."""
        mz, my, mx = f.shape
        return self._field(self.lib.pcref_derij_full_field, f, mx - 6, my - 6, mz - 6, i, j, di_1, dj_1)

    def boundconds(self, f, par):
        self.lib.pcref_boundconds(f.ctypes.data, C.byref(par))

    def rhs(self, f, df, par, lfirst=True, dt1_max=0.0, diag=None):
        d = C.c_double(dt1_max)
        self.lib.pcref_rhs(f.ctypes.data, df.ctypes.data, C.byref(par), int(lfirst), C.byref(d),
                           diag.ctypes.data if diag is not None else None)
        return d.value

    def timestep(self, f, df, par, dt=0.0, frc=None, diag=None):
        d = C.c_double(dt)
        self.lib.pcref_timestep(f.ctypes.data, df.ctypes.data, C.byref(par),
                                C.byref(frc) if frc is not None else None, C.byref(d),
                                diag.ctypes.data if diag is not None else None)
        return d.value

    def timestep_pencil(self, f, df, par, dt=0.0, frc=None):
        """the same time step evaluated pencil-wise like the reference's mn-loop (pcref_pencil.c): for TIMING the CPU arm;
        falls back to timestep() where it does not apply"""
        d = C.c_double(dt)
        self.lib.pcref_timestep_pencil(f.ctypes.data, df.ctypes.data, C.byref(par),
                                       C.byref(frc) if frc is not None else None, C.byref(d), None)
        return d.value

    def pencil_applies(self, par):
        return bool(self.lib.pcref_pencil_applies(C.byref(par)))

    def num_threads(self):
        return self.lib.pcref_num_threads()

    def set_num_threads(self, n):
        self.lib.pcref_set_num_threads(int(n))

    # ---- dust particles (pcref_particles.c) ----------------------------------
    def imn_array(self, par):
        out = np.zeros(par.ny * par.nz, dtype=np.int32)
        self.lib.pcref_setup_imn_array(C.byref(par), out.ctypes.data)
        return out.reshape(par.nz, par.ny)

    def particles_boundconds(self, st, par, pp):
        self.lib.pcref_particles_boundconds(st.fp.ctypes.data, C.byref(par), C.byref(pp))

    def particles_map_nearest_grid(self, st, par, pp):
        self.lib.pcref_particles_map_nearest_grid(st.fp.ctypes.data, st.ineargrid.ctypes.data, C.byref(par), C.byref(pp))

    def particles_sort(self, st, par, pp):
        self.lib.pcref_particles_sort(st.fp.ctypes.data, st.dfp.ctypes.data, st.ipar.ctypes.data, st.ineargrid.ctypes.data,
                                      st.k1_imn.ctypes.data, st.k2_imn.ctypes.data, C.byref(par), C.byref(pp))

    def particles_map_xxp_grid(self, st, par, pp):
        self.lib.pcref_particles_map_xxp_grid(st.fp.ctypes.data, st.ineargrid.ctypes.data, st.np_.ctypes.data,
                                              st.rhop.ctypes.data, C.byref(par), C.byref(pp))

    def particles_pde(self, f, df, st, par, pp, lfirst=True, dt1_max=0.0):
        d = C.c_double(dt1_max)
        self.lib.pcref_particles_pde(f.ctypes.data, df.ctypes.data, st.fp.ctypes.data, st.dfp.ctypes.data,
                                     st.ineargrid.ctypes.data, C.byref(par), C.byref(pp), int(lfirst), C.byref(d))
        return d.value

    def particles_fixture(self, st=None, ap_by_ipar=None, rhopmat=0.0, interp_cic=False):
        """options of the reference fixture samples/ParticlesCaustics (CIC interpolation, pure Stokes drag with the radii
        of particles_radius, looked up by st.ipar); call without arguments to switch them off again.  The arrays must
        stay alive while the options are on."""
        if st is None:
            self.lib.pcref_particles_fixture(0, None, 0, 0.0, None)
        else:
            self.lib.pcref_particles_fixture(int(interp_cic), ap_by_ipar.ctypes.data if ap_by_ipar is not None else None,
                                             0 if ap_by_ipar is None else ap_by_ipar.size, float(rhopmat), st.ipar.ctypes.data)

    def timestep_particles(self, f, df, par, pp, st, dt=0.0, frc=None, diag=None):
        d = C.c_double(dt)
        self.lib.pcref_timestep_particles(f.ctypes.data, df.ctypes.data, C.byref(par),
                                          C.byref(frc) if frc is not None else None, C.byref(d),
                                          diag.ctypes.data if diag is not None else None, C.byref(pp),
                                          st.fp.ctypes.data, st.dfp.ctypes.data, st.ipar.ctypes.data,
                                          st.ineargrid.ctypes.data, st.np_.ctypes.data, st.rhop.ctypes.data,
                                          st.k1_imn.ctypes.data, st.k2_imn.ctypes.data)
        return d.value
