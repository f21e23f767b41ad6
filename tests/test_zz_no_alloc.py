"""pcb200_cfl_dt_noalloc (static device scratch, no allocation) against the existing entry point.

Collected last on purpose.  The existing path: substep 1 of an isothermal
run launches cfl_prepass with cudaMalloc'ed partials and returns dt through pcb200_substep; the new entry
point runs the same kernel on a caller-owned device copy of f with the partials in a file-scope
__device__ array.  Two calls in a row on different states (the array keeps the last call's contents).
Tolerance: 1e-13 relative, the dt bar of tests/test_gpu_parity.py (same kernel, so the bits agree)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def wrap_ghosts(f):
    """periodic ghost zones from the interior, all three directions (what bc_periodic does on the device)"""
    inner = f[:, 3:-3, 3:-3, 3:-3]
    return np.ascontiguousarray(np.pad(inner, ((0, 0), (3, 3), (3, 3), (3, 3)), mode="wrap"))


@pytest.mark.parametrize("case", ["mhd", "hydro"])
def test_noalloc_cfl_dt_matches_substep_dt_twice(pk, dev, case):
    import torch
    P, zprof, grid = (pk.setups.config_mhd if case == "mhd" else pk.setups.config_hydro)(32)
    f0 = pk.setups.synthetic_fields(P, grid)
    # existing entry point: dt of f0, a full step, dt of the stepped state
    dev.initialize_GPU(P, zprof)
    f = f0.copy()
    dev.register_GPU(f)
    dev.set_forcing(None)
    dtA = dev.rhs_GPU(f, 1)
    dev.rhs_GPU(f, 2)
    dev.rhs_GPU(f, 3)
    f1 = dev.copy_farray_from_GPU(f).copy()
    dtB = dev.rhs_GPU(f, 1)
    assert dtA > 0 and dtB > 0 and dtA != dtB
    # new entry point: caller-owned device buffers (torch only allocates them), nothing allocated inside
    out = torch.zeros(2, dtype=torch.float64, device="cuda:0")
    got = []
    for i, state in enumerate((f0, f1)):
        fd = torch.from_numpy(wrap_ghosts(state)).to("cuda:0")
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info(0)[0]
        dev.cfl_dt_noalloc(fd.data_ptr(), out.data_ptr())
        dev.sync()
        if i == 1:      # the first call may make the driver load the two kernels (lazy module loading)
            assert torch.cuda.mem_get_info(0)[0] >= free0      # no device memory taken by the call
        got.append(out.cpu().numpy().copy())
    for (dt, dt1), ref in zip(got, (dtA, dtB)):
        assert abs(dt - ref) <= 1e-13 * ref, (dt, ref)
        assert abs(dt * dt1 - 1.0) < 1e-15


def test_noalloc_rejects_grids_beyond_the_static_scratch(pk, dev):
    import torch
    # 65 x 128 x 1 blocks of 32x8x64 cells = 8320 > PCB200_NOALLOC_MAX_BLOCKS (8192)
    P, zprof, grid = pk.setups.make_params(2080, 1024, 6, magnetic=False)
    dev.initialize_GPU(P, zprof)
    dummy = torch.zeros(16, dtype=torch.float64, device="cuda:0")
    with pytest.raises(pk.gpu.PCB200Error, match="static scratch"):
        dev.cfl_dt_noalloc(dummy.data_ptr(), dummy.data_ptr())
    with pytest.raises(pk.gpu.PCB200Error, match="null"):
        dev.cfl_dt_noalloc(0, dummy.data_ptr())
    # leave the session device on a small grid again
    P, zprof, grid = pk.setups.config_hydro(32)
    dev.initialize_GPU(P, zprof)
