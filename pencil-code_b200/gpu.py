"""Host-side mirror of the reference's Fortran module `GPU` (src/gpu.h:1) on top
of libpcb200.so.

Same public list and call moments as the reference
(gpu_init run.f90:125, register_GPU run.f90:223, initialize_GPU register.f90:375,
rhs_GPU equ.f90:343, copy_farray_from_GPU equ.f90:345/snapshot.f90:286,
finalize_GPU register.f90:455).  The production host is the Fortran wrapper
fortran/gpu_b200.f90; this Python class exists because this image has no
Fortran compiler, and drives the very same C entry points (tests, bench).

There is no CPU fallback: if libpcb200.so is missing or no CUDA device is
present every call raises.
"""
import ctypes as C
import os

import numpy as np

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
# PCB200_LIB: another build of the same library (A/B timing of kernel variants); the default is the in-tree build
LIB_PATH = os.environ.get("PCB200_LIB") or os.path.join(_HERE, "libpcb200.so")


class PCB200Error(RuntimeError):
    pass


def load_library(path=LIB_PATH):
    if not os.path.exists(path):
        raise PCB200Error(
            f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(path, mode=C.RTLD_GLOBAL)
    return abi.bind(lib)


class GPU:
    """module GPU: gpu_init, register_GPU, initialize_GPU, finalize_GPU, rhs_GPU,
    copy_farray_from_GPU (+ the extras the C ABI offers)."""

    def __init__(self, lib=None):
        self.lib = lib if lib is not None else load_library()
        self._f = None
        self._lvery_first = True          # gpu_astaroth.f90:107
        self.dt = 0.0
        self.par = None

    def _ck(self, rc):
        if rc != 0:
            raise PCB200Error(f"pcb200 error {rc}: {self.lib.pcb200_last_error().decode()}")

    # -- reference interface -------------------------------------------------
    def gpu_init(self, local_rank=0):
        self._ck(self.lib.pcb200_init(int(local_rank)))

    def comm_init(self, id128, rank, nranks):
        self._ck(self.lib.pcb200_comm_init(id128, int(rank), int(nranks)))

    def nccl_unique_id(self):
        buf = C.create_string_buffer(128)
        self._ck(self.lib.pcb200_nccl_unique_id(buf))
        return buf.raw

    def initialize_GPU(self, par, zprofiles=None):
        self.par = par
        self._ck(self.lib.pcb200_initialize(C.byref(par)))
        if zprofiles:
            for which, arr in zprofiles.items():
                a = np.ascontiguousarray(arr, dtype=np.float64)
                self._ck(self.lib.pcb200_set_zprofile(int(which), a.ctypes.data, a.size))
        self._lvery_first = True
        self.dt = 0.0

    def register_GPU(self, f):
        assert f.dtype == np.float64 and f.flags["C_CONTIGUOUS"]
        self._f = f                        # aliased, like registerGPU
        self._ck(self.lib.pcb200_register(f.ctypes.data))

    def rhs_GPU(self, f, isubstep, early_finalize=False, ldiagnos=False, dt=None):
        if f is not None and f is not self._f:
            self.register_GPU(f)
        d = C.c_double(self.dt if dt is None else dt)
        self._ck(self.lib.pcb200_substep(int(isubstep), int(self._lvery_first), int(early_finalize),
                                         int(ldiagnos), C.byref(d)))
        self._lvery_first = False
        self.dt = d.value
        return self.dt

    def copy_farray_from_GPU(self, f=None):
        tgt = self._f if f is None else f
        self._ck(self.lib.pcb200_copy_farray(tgt.ctypes.data))
        return tgt

    def finalize_GPU(self):
        self._ck(self.lib.pcb200_finalize())

    # -- extras ----------------------------------------------------------------
    def load_farray(self, f=None):
        src = self._f if f is None else f
        self._ck(self.lib.pcb200_load_farray(src.ctypes.data))
        self._lvery_first = False

    def time_step(self, ldiagnos=False, dt=None):
        """time_step (timestep.f90:65): 3 substeps; returns dt."""
        d = C.c_double(self.dt if dt is None else dt)
        self._ck(self.lib.pcb200_timestep(int(self._lvery_first), int(ldiagnos), C.byref(d)))
        self._lvery_first = False
        self.dt = d.value
        return self.dt

    def set_forcing(self, frc):
        self._ck(self.lib.pcb200_set_forcing(C.byref(frc) if frc is not None else None))

    def diagnostics(self):
        out = (C.c_double * abi.NDIAG)()
        self._ck(self.lib.pcb200_get_diagnostics(out, abi.NDIAG))
        return dict(zip(abi.DIAG_NAMES, list(out)))

    def sync(self):
        self._ck(self.lib.pcb200_sync())

    def timer_start(self):
        self._ck(self.lib.pcb200_timer_start())

    def timer_stop(self):
        ms = C.c_float()
        self._ck(self.lib.pcb200_timer_stop(C.byref(ms)))
        return ms.value

    def launch_count(self):
        return int(self.lib.pcb200_launch_count())

    def kernel_name(self):
        """fused RHS+RK kernel of the last substep launch (rhs_dly / rhs_tma / rhs_march / rhs_global)"""
        return self.lib.pcb200_kernel_name().decode()

    def cfl_dt_noalloc(self, f_dev_ptr, dt_out_dev_ptr):
        """pcb200_cfl_dt_noalloc: CFL dt of a device-resident f (ghosts valid) without any allocation; device
        pointers as integers; asynchronous on the library's stream (sync() waits).  No reference counterpart."""
        self._ck(self.lib.pcb200_cfl_dt_noalloc(C.c_void_p(int(f_dev_ptr)), C.c_void_p(int(dt_out_dev_ptr))))

    def snapshot_begin(self, path, t, grid):
        """wsnap from the device (snapshot.f90:286-302 + io_dist.f90:110-167): returns at once, the file is written by
        a worker thread while the time loop goes on; grid = dict(x, y, z incl. ghost points, dx, dy, dz)"""
        x, y, z = (np.ascontiguousarray(grid[k], dtype=np.float64) for k in ("x", "y", "z"))
        self._ck(self.lib.pcb200_snapshot_begin(str(path).encode(), float(t), x.ctypes.data, y.ctypes.data, z.ctypes.data,
                                                float(grid["dx"]), float(grid["dy"]), float(grid["dz"])))

    def snapshot_wait(self):
        self._ck(self.lib.pcb200_snapshot_wait())

    def enable_kernel_timing(self, on=True):
        self._ck(self.lib.pcb200_enable_kernel_timing(int(on)))

    def rhs_kernel_time(self):
        ms = C.c_float()
        n = C.c_int()
        self._ck(self.lib.pcb200_rhs_kernel_time(C.byref(ms), C.byref(n)))
        return ms.value, n.value

    # -- dust particles (called by the wrapper when lparticles) --------
    def particles_initialize(self, pp):
        self._ck(self.lib.pcb200_particles_initialize(C.byref(pp)))

    def particles_load(self, st):
        """st: setups.ParticleState (fp[6][mpar], ipar[mpar])."""
        self._ck(self.lib.pcb200_particles_load(st.fp.ctypes.data, st.ipar.ctypes.data))

    def particles_store(self, st):
        self._ck(self.lib.pcb200_particles_store(st.fp.ctypes.data, st.dfp.ctypes.data, st.ipar.ctypes.data,
                                                 st.ineargrid.ctypes.data, st.k1_imn.ctypes.data, st.k2_imn.ctypes.data))
        self._ck(self.lib.pcb200_particles_get_aux(st.np_.ctypes.data, st.rhop.ctypes.data))
        n = C.c_int32()
        self._ck(self.lib.pcb200_particles_count(C.byref(n)))
        st.npar = n.value                     # changes when particles migrate between ranks
        return n.value
