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
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(os.path.join(_HERE, "pcref.c")):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return LIB


class Oracle:
    def __init__(self, abi):
        """abi: the product's ctypes module (only for the shared struct layout)."""
        build()
        self.abi = abi
        L = self.lib = C.CDLL(LIB)
        d, i, p = C.c_double, C.c_int, C.c_void_p
        for name in ("pcref_der_field", "pcref_der2_field", "pcref_der6_field"):
            getattr(L, name).argtypes = [p, i, i, i, i, d, p]
            getattr(L, name).restype = None
        L.pcref_derij_field.argtypes = [p, i, i, i, i, i, d, d, p]
        L.pcref_derij_field.restype = None
        L.pcref_boundconds.argtypes = [p, C.POINTER(abi.Params)]
        L.pcref_boundconds.restype = None
        L.pcref_rhs.argtypes = [p, p, C.POINTER(abi.Params), i, C.POINTER(d), p]
        L.pcref_rhs.restype = None
        L.pcref_timestep.argtypes = [p, p, C.POINTER(abi.Params), C.POINTER(abi.Forcing), C.POINTER(d), p]
        L.pcref_timestep.restype = None
        L.pcref_addforce.argtypes = [p, C.POINTER(abi.Params), C.POINTER(abi.Forcing)]
        L.pcref_addforce.restype = None
        L.pcref_set_zprofile.argtypes = [i, p, i]
        L.pcref_set_zprofile.restype = None
        L.pcref_num_threads.restype = i

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

    def num_threads(self):
        return self.lib.pcref_num_threads()
