"""var.dat / VAR<n> snapshot files in the reference's layout (io_dist.f90:110-167 output_snap, :535-568 input_snap):
Fortran sequential unformatted, 4-byte record markers (gfortran, incl. its subrecord scheme for records > 2 GiB),
    record 1:  f(mx,my,mz,mvar)            ghost zones included, x fastest
    record 2:  t, x(mx), y(my), z(mz), dx, dy, dz   [, deltay with shear]
Host-side reader/writer used by the tests to check what pcb200_write_snapshot (csrc/snapshot.cuh) puts on disk, and
by callers that want to hand a reference snapshot to the device (pc_start output -> register_GPU)."""
import struct

import numpy as np

FREC_MAX = 2147483639          # libgfortran's maximum subrecord length


def read_records(path):
    """list of logical records (bytes); subrecords (negative leading marker = continued) are joined"""
    out, cur = [], []
    with open(path, "rb") as fh:
        while True:
            head = fh.read(4)
            if not head:
                break
            n, = struct.unpack("<i", head)
            body = fh.read(abs(n))
            tail, = struct.unpack("<i", fh.read(4))
            if abs(tail) != abs(n) or len(body) != abs(n):
                raise ValueError(f"{path}: broken record markers ({n}, {tail})")
            cur.append(body)
            if n >= 0:                       # last (or only) subrecord of this record
                out.append(b"".join(cur))
                cur = []
    if cur:
        raise ValueError(f"{path}: file ends inside a continued record")
    return out


def read_var(path, mx, my, mz, mvar, dtype=np.float64):
    """-> dict(f=(mvar,mz,my,mx) array [== Fortran f(mx,my,mz,mvar)], t, x, y, z, dx, dy, dz[, deltay])"""
    recs = read_records(path)
    if len(recs) < 2:
        raise ValueError(f"{path}: expected 2 records, found {len(recs)}")
    dt = np.dtype(dtype)
    f = np.frombuffer(recs[0], dtype=dt)
    if f.size != mx * my * mz * mvar:
        raise ValueError(f"{path}: record 1 has {f.size} values, expected {mx}*{my}*{mz}*{mvar}")
    tail = np.frombuffer(recs[1], dtype=dt)
    n = 1 + mx + my + mz + 3
    if tail.size not in (n, n + 1):
        raise ValueError(f"{path}: record 2 has {tail.size} values, expected {n} (+1 with shear)")
    out = dict(f=f.reshape(mvar, mz, my, mx), t=float(tail[0]), x=tail[1:1 + mx], y=tail[1 + mx:1 + mx + my],
               z=tail[1 + mx + my:1 + mx + my + mz], dx=float(tail[n - 3]), dy=float(tail[n - 2]), dz=float(tail[n - 1]))
    if tail.size == n + 1:
        out["deltay"] = float(tail[n])
    return out


ID_BLOCK_PERSISTENT, ID_RECORD_RANDOM_SEEDS = 2000, 1          # record_types.h:13,16


def read_persist(path):
    """the persistent block that follows the two snapshot records (io_dist.f90:351-405 init_write_persist /
    write_persist_id, read side :772-846): int32 2000, then pairs of records (id, payload), closed by 2000 again.
    -> {id: payload bytes}; id 1 = the random-number state run.x restores (persist.f90:156-160)."""
    recs = read_records(path)[2:]
    ids = [struct.unpack("<i", r)[0] if len(r) == 4 else None for r in recs]
    if not recs:
        return {}
    if ids[0] != ID_BLOCK_PERSISTENT or ids[-1] != ID_BLOCK_PERSISTENT or len(recs) % 2:
        raise ValueError(f"{path}: persistent block not framed by {ID_BLOCK_PERSISTENT}")
    return {ids[k]: recs[k + 1] for k in range(1, len(recs) - 1, 2)}


def _write_record(fh, payload, rec_max=FREC_MAX):
    total, off, first = len(payload), 0, True
    while True:
        k = min(total - off, rec_max)
        more = off + k < total
        fh.write(struct.pack("<i", -k if more else k))
        fh.write(payload[off:off + k])
        fh.write(struct.pack("<i", k if first else -k))
        off += k
        first = False
        if not more:
            break


def write_var(path, f, t, x, y, z, dx, dy, dz, dtype=np.float64, rec_max=FREC_MAX):
    """host-side writer of the same layout (tests compare its bytes with the device writer's)"""
    dt = np.dtype(dtype)
    with open(path, "wb") as fh:
        _write_record(fh, np.ascontiguousarray(f, dtype=dt).tobytes(), rec_max)
        tail = np.concatenate([[t], x, y, z, [dx, dy, dz]]).astype(dt)
        _write_record(fh, tail.tobytes(), rec_max)


def append_persist(path, records):
    """append the persistent block to a snapshot file that ends after record 2 (what the host does after
    pcb200_snapshot_wait): {id: payload bytes} in the given order, framed by the block id (io_dist.f90:351-405)"""
    with open(path, "ab") as fh:
        _write_record(fh, struct.pack("<i", ID_BLOCK_PERSISTENT))
        for rid, payload in records.items():
            _write_record(fh, struct.pack("<i", rid))
            _write_record(fh, bytes(payload))
        _write_record(fh, struct.pack("<i", ID_BLOCK_PERSISTENT))

