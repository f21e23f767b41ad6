"""Device-side snapshot writer: pcb200_snapshot_begin/wait write var.dat in the reference's layout
(io_dist.f90:110-167) from a staging copy of f while the time loop goes on.  Checked against copy_farray_from_GPU taken
at the same moment, read back with pencil-code_b200/snapshot.py (whose reader/writer are pinned byte for byte to the
reference's own fixture tests/input/serial-1/proc0/var.dat by tests/test_snapshot_layout.py)."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_snapshot_written_while_the_run_goes_on(pk, dev, tmp_path):
    from importlib import import_module
    snap = import_module("pencil_code_b200.snapshot")
    P, zprof, grid = pk.setups.config_mhd(32)
    f = pk.setups.synthetic_fields(P, grid)
    dev.initialize_GPU(P, zprof)
    dev.register_GPU(f)
    dev.set_forcing(None)
    t = 0.0
    for _ in range(2):
        t += dev.time_step()
    path = str(tmp_path / "var.dat")
    dev.snapshot_begin(path, t, grid)               # returns at once
    held = dev.copy_farray_from_GPU(f).copy()       # the same state through the reference's synchronous route
    for _ in range(3):                              # the run goes on while the worker writes
        dev.time_step()
    dev.snapshot_wait()
    m = P.nx + 6
    v = snap.read_var(path, m, m, m, P.mvar)
    assert np.array_equal(v["f"], held)             # ghost zones included, bit for bit
    assert v["t"] == t and np.array_equal(v["x"], grid["x"]) and np.array_equal(v["z"], grid["z"])
    assert (v["dx"], v["dy"], v["dz"]) == (grid["dx"], grid["dy"], grid["dz"])
    # the host-side writer produces the same bytes
    ref = str(tmp_path / "ref.dat")
    snap.write_var(ref, held, t, grid["x"], grid["y"], grid["z"], grid["dx"], grid["dy"], grid["dz"])
    assert open(path, "rb").read() == open(ref, "rb").read()


def test_snapshot_subrecords_and_small_chunks(tmp_path):
    """records above the marker limit are split gfortran-style; exercised in a fresh process with a small limit that is
    not a multiple of 8 (like the real 2147483639) and chunk buffers smaller than a subrecord"""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out, ref = str(tmp_path / "v.dat"), str(tmp_path / "r.dat")
    code = f"""
import sys; sys.path.insert(0, {root!r})
import numpy as np, __graft_entry__ as g
from importlib import import_module
pk = g.package(); snap = import_module("pencil_code_b200.snapshot")
P, zp, grid = pk.setups.config_hydro(32)
f = pk.setups.synthetic_fields(P, grid)
d = pk.GPU(); d.gpu_init(0); d.initialize_GPU(P, zp); d.register_GPU(f); d.load_farray()
d.snapshot_begin({out!r}, 0.25, grid); d.snapshot_wait()
d.copy_farray_from_GPU(f)
snap.write_var({ref!r}, f, 0.25, grid["x"], grid["y"], grid["z"], grid["dx"], grid["dy"], grid["dz"], rec_max=1048573)
d.finalize_GPU()
"""
    env = dict(os.environ, PCB200_SNAP_RECMAX="1048573", PCB200_SNAP_CHUNK_MB="1")
    subprocess.check_call([sys.executable, "-c", code], env=env)
    a, b = open(out, "rb").read(), open(ref, "rb").read()
    assert len(a) == len(b) and a == b
    assert np.frombuffer(a[:4], dtype="<i4")[0] == -1048573
