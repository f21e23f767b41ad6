"""Snapshot record layout: the reader/writer of pencil-code_b200/snapshot.py against the reference's own
fixture tests/input/serial-1/proc0/{var.dat,dim.dat} (copied byte-identical to tests/golden/; single precision,
mx,my,mz,mvar = 10,12,11,5), i.e. io_dist.f90:110-167: record 1 = f(mx,my,mz,mvar), record 2 = t,x,y,z,dx,dy,dz."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def _dims():
    rows = open(os.path.join(HERE, "golden", "serial-1.proc0.dim.dat")).read().split("\n")
    mx, my, mz, mvar = (int(v) for v in rows[0].split()[:4])
    prec = rows[1].strip()
    return mx, my, mz, mvar, (np.float32 if prec == "S" else np.float64)


def test_reader_parses_reference_var_dat(pk):
    from importlib import import_module
    snap = import_module("pencil_code_b200.snapshot")
    mx, my, mz, mvar, dt = _dims()
    path = os.path.join(HERE, "golden", "serial-1.proc0.var.dat")
    recs = snap.read_records(path)
    item = np.dtype(dt).itemsize
    assert [len(r) for r in recs] == [mx * my * mz * mvar * item, (1 + mx + my + mz + 3) * item]
    v = snap.read_var(path, mx, my, mz, mvar, dt)
    assert v["f"].shape == (mvar, mz, my, mx)
    # the mesh of record 2: x and y equidistant with the stated spacings (z is a non-equidistant grid in this fixture)
    assert np.allclose(np.diff(v["x"]), v["dx"], rtol=1e-5) and np.allclose(np.diff(v["y"]), v["dy"], rtol=1e-5)
    assert (np.diff(v["z"]) > 0).all() and v["dz"] > 0
    assert np.isfinite(v["f"]).all() and v["t"] >= 0.0


def test_writer_reproduces_reference_bytes(pk, tmp_path):
    """read the reference file, write it again with our writer: byte-identical"""
    from importlib import import_module
    snap = import_module("pencil_code_b200.snapshot")
    mx, my, mz, mvar, dt = _dims()
    src = os.path.join(HERE, "golden", "serial-1.proc0.var.dat")
    v = snap.read_var(src, mx, my, mz, mvar, dt)
    out = str(tmp_path / "var.dat")
    snap.write_var(out, v["f"], v["t"], v["x"], v["y"], v["z"], v["dx"], v["dy"], v["dz"], dtype=dt)
    assert open(out, "rb").read() == open(src, "rb").read()


def test_subrecords_round_trip(pk, tmp_path):
    """records above the marker limit are split gfortran-style (negative leading marker = continued, negative trailing
    marker = has a predecessor); exercised with a small artificial limit"""
    from importlib import import_module
    snap = import_module("pencil_code_b200.snapshot")
    rng = np.random.default_rng(3)
    f = rng.standard_normal((2, 5, 4, 6))
    x, y, z = np.arange(6.0), np.arange(4.0), np.arange(5.0)
    out = str(tmp_path / "v.dat")
    snap.write_var(out, f, 1.5, x, y, z, 1.0, 1.0, 1.0, rec_max=500)
    raw = open(out, "rb").read()
    first = np.frombuffer(raw[:4], dtype="<i4")[0]
    assert first == -500                       # 1920-byte record -> 500+500+500+420, first three marked "continued"
    v = snap.read_var(out, 6, 4, 5, 2)
    assert np.array_equal(v["f"], f) and v["t"] == 1.5 and np.array_equal(v["z"], z)


REF_DOUBLE = "/root/reference/samples/2d-tests/2d_methane_flame/turbulent_field/data/proc0"


def test_double_precision_reference_snapshot_with_persistent_block(pk, tmp_path):
    """the reference's one double-precision var.dat (38x38x7x4, written by a run with forcing: a persistent block follows
    the two records).  Read straight from the reference tree (324 KB, not copied); skipped where the tree is absent.
    The two snapshot records re-written by our writer are byte-identical to the file's first two records."""
    import pytest
    if not os.path.isfile(os.path.join(REF_DOUBLE, "var.dat")):
        pytest.skip("reference tree not here")
    snap = pk.snapshot
    mx, my, mz, mvar = (int(v) for v in open(os.path.join(REF_DOUBLE, "dim.dat")).read().split()[:4])
    assert open(os.path.join(REF_DOUBLE, "dim.dat")).read().split()[6] == "D"
    path = os.path.join(REF_DOUBLE, "var.dat")
    v = snap.read_var(path, mx, my, mz, mvar, np.float64)
    assert v["f"].shape == (mvar, mz, my, mx) and np.isfinite(v["f"]).all()
    assert np.allclose(np.diff(v["x"]), v["dx"], rtol=1e-12) and np.allclose(np.diff(v["y"]), v["dy"], rtol=1e-12)
    per = snap.read_persist(path)
    assert snap.ID_RECORD_RANDOM_SEEDS in per and len(per[snap.ID_RECORD_RANDOM_SEEDS]) % 4 == 0
    out = str(tmp_path / "var.dat")
    snap.write_var(out, v["f"], v["t"], v["x"], v["y"], v["z"], v["dx"], v["dy"], v["dz"], dtype=np.float64)
    ours, ref = open(out, "rb").read(), open(path, "rb").read()
    assert ref[:len(ours)] == ours and len(ref) - len(ours) == sum(len(r) + 8 for r in snap.read_records(path)[2:])


def test_snapshot_with_persistent_block_is_rewritten_byte_for_byte(pk, tmp_path):
    """samples/backwards-compatible/data/proc0/var.dat (single precision, persistent block with the random-number state
    and the forcing records): records 1-2 by write_var, the block by append_persist -> the same bytes"""
    snap = pk.snapshot
    src = os.path.join(HERE, "golden", "backwards-compatible.var.dat")
    v = snap.read_var(src, 10, 10, 10, 7, np.float32)
    per = snap.read_persist(src)
    assert list(per)[0] == snap.ID_RECORD_RANDOM_SEEDS and len(per) >= 3
    out = str(tmp_path / "var.dat")
    snap.write_var(out, v["f"], v["t"], v["x"], v["y"], v["z"], v["dx"], v["dy"], v["dz"], dtype=np.float32)
    snap.append_persist(out, per)
    assert open(out, "rb").read() == open(src, "rb").read()
