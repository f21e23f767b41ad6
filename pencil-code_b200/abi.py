"""ctypes mirror of include/pcb200.h (the C-ABI boundary).

Plain data only: these structures are what the Fortran wrapper
(fortran/gpu_b200.f90) fills from cdata/cparam and the module run_pars, the
replacement for the generated cparam_c.h/cdata_c.h headers of the reference
(src/gpu_astaroth_ansi.c:31-35, src/astaroth/gpu_astaroth.cc:242-317).
"""
import ctypes as C

ABI_VERSION = 1
MAXVAR = 8
NGHOST = 3
NDIAG = 16
NPROF = 3

BC = {"p": 0, "s": 1, "a": 2, "a2": 3, "c1": 4, "cT": 5}
HCOND = {"nothing": 0, "K-profile": 1, "chi-const": 2}
PROF_HCOND, PROF_DLNHCOND, PROF_COOL = 0, 1, 2

DIAG_NAMES = ["u2sum", "u2max", "rhosum", "sssum", "b2sum", "b2max", "j2sum", "j2max",
              "dtc", "dtu", "dtnu", "dtchi", "dt1max", "absum", "jbsum", "ousum"]

_i = C.c_int32
_d = C.c_double


class Params(C.Structure):
    _fields_ = [
        ("abi_version", _i),
        ("nx", _i), ("ny", _i), ("nz", _i),
        ("nprocx", _i), ("nprocy", _i), ("nprocz", _i),
        ("ipx", _i), ("ipy", _i), ("ipz", _i),
        ("lperi", _i * 3),
        ("mvar", _i), ("mfarray", _i),
        ("iux", _i), ("iuy", _i), ("iuz", _i), ("ilnrho", _i), ("iss", _i),
        ("iax", _i), ("iay", _i), ("iaz", _i),
        ("lhydro", _i), ("ldensity", _i), ("lentropy", _i), ("lmagnetic", _i),
        ("lgravz", _i), ("lforcing", _i),
        ("lupw_uu", _i), ("lupw_lnrho", _i), ("lupw_ss", _i),
        ("lbidiagonal_derij", _i),
        ("iheatcond", _i), ("lcooling", _i), ("lviscosity_heat", _i), ("lohmic_heat", _i),
        ("ldt", _i),
        ("reserved_i", _i * 7),
        ("bcx", (_i * MAXVAR) * 2), ("bcy", (_i * MAXVAR) * 2), ("bcz", (_i * MAXVAR) * 2),
        ("dx", _d), ("dy", _d), ("dz", _d),
        ("dx_1", _d), ("dy_1", _d), ("dz_1", _d),
        ("dz2_bound", _d * 7),
        ("xyz0", _d * 3), ("Lxyz", _d * 3),
        ("x0_loc", _d), ("y0_loc", _d), ("z0_loc", _d),
        ("cdt", _d), ("cdtv", _d), ("cdts", _d), ("dtmax", _d), ("dt_fixed", _d),
        ("alpha_ts", _d * 3), ("beta_ts", _d * 3),
        ("nu", _d), ("eta", _d), ("mu0", _d),
        ("cs20", _d), ("gamma", _d), ("cp", _d), ("lnrho0", _d),
        ("Omega", _d), ("gravz", _d), ("chi", _d),
        ("hcond0", _d), ("hcond1", _d), ("hcond2", _d),
        ("z1", _d), ("z2", _d), ("widthss", _d),
        ("cool", _d), ("wcool", _d), ("cs2cool", _d),
        ("ztop", _d), ("Fbot", _d), ("cs2top", _d),
        ("reserved_d", _d * 8),
    ]


class Forcing(C.Structure):
    _fields_ = [
        ("coef1", _d * 3), ("coef2", _d * 3), ("fda", _d * 3), ("kk", _d * 3),
        ("phase", _d), ("fact", _d),
    ]


class ParticleParams(C.Structure):
    """pcb200_particle_params (particles_dust/particles_map run parameters)."""
    _fields_ = [
        ("abi_version", _i),
        ("npar_loc", _i), ("mpar_loc", _i),
        ("ldragforce_dust_par", _i), ("ldragforce_gas_par", _i), ("lcoriolis_force_par", _i),
        ("ldt_adv_par", _i), ("lcheck_exact_frontier", _i),
        ("reserved_i", _i * 8),
        ("tausp", _d), ("rhop_swarm", _d), ("cdtp", _d), ("cdtp_drag", _d),
        ("reserved_d", _d * 8),
    ]


# every symbol include/pcb200.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "pcb200_init": (C.c_int, [C.c_int]),
    "pcb200_initialize": (C.c_int, [C.POINTER(Params)]),
    "pcb200_register": (C.c_int, [C.c_void_p]),
    "pcb200_substep": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(_d)]),
    "pcb200_timestep": (C.c_int, [C.c_int, C.c_int, C.POINTER(_d)]),
    "pcb200_copy_farray": (C.c_int, [C.c_void_p]),
    "pcb200_load_farray": (C.c_int, [C.c_void_p]),
    "pcb200_set_forcing": (C.c_int, [C.POINTER(Forcing)]),
    "pcb200_set_zprofile": (C.c_int, [C.c_int, C.c_void_p, C.c_int]),
    "pcb200_get_diagnostics": (C.c_int, [C.POINTER(_d), C.c_int]),
    "pcb200_nccl_unique_id": (C.c_int, [C.c_char_p]),
    "pcb200_comm_init": (C.c_int, [C.c_char_p, C.c_int, C.c_int]),
    "pcb200_finalize": (C.c_int, []),
    "pcb200_timer_start": (C.c_int, []),
    "pcb200_timer_stop": (C.c_int, [C.POINTER(C.c_float)]),
    "pcb200_sync": (C.c_int, []),
    "pcb200_launch_count": (C.c_int64, []),
    "pcb200_rhs_kernel_time": (C.c_int, [C.POINTER(C.c_float), C.POINTER(C.c_int)]),
    "pcb200_enable_kernel_timing": (C.c_int, [C.c_int]),
    "pcb200_particles_initialize": (C.c_int, [C.POINTER(ParticleParams)]),
    "pcb200_particles_load": (C.c_int, [C.c_void_p, C.c_void_p]),
    "pcb200_particles_store": (C.c_int, [C.c_void_p] * 6),
    "pcb200_particles_get_aux": (C.c_int, [C.c_void_p, C.c_void_p]),
    "pcb200_particles_count": (C.c_int, [C.POINTER(C.c_int32)]),
    "pcb200_kernel_name": (C.c_char_p, []),
    "pcb200_last_error": (C.c_char_p, []),
    "pcb200_abi_version": (C.c_int, []),
    "pcb200_cfl_dt_noalloc": (C.c_int, [C.c_void_p, C.c_void_p]),
    "pcb200_snapshot_begin": (C.c_int, [C.c_char_p, _d, C.c_void_p, C.c_void_p, C.c_void_p, _d, _d, _d]),
    "pcb200_snapshot_wait": (C.c_int, []),
}


def bind(lib):
    """Attach restype/argtypes for every declared symbol; raises if one is missing."""
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)      # AttributeError -> symbol not exported
        fn.restype = res
        fn.argtypes = args
    return lib
