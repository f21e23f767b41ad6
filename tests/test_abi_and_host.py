"""CPU-only checks of the boundary: the shared library loads and exports every
symbol include/pcb200.h declares; the ctypes mirror matches the C struct; the
host logic (set-ups, slab decomposition) is consistent.  No compute calls."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(pk):
    lib = C.CDLL(pk.gpu.LIB_PATH)
    hdr = open(os.path.join(ROOT, "include", "pcb200.h")).read()
    declared = set(re.findall(r"\b(pcb200_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations found"
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/pcb200.h but not exported"
    assert set(pk.abi.SYMBOLS) == declared
    pk.abi.bind(lib)
    assert lib.pcb200_abi_version() == pk.abi.ABI_VERSION
    assert lib.pcb200_kernel_name() == b""      # no substep launched yet (no compute call here: no GPU needed)


def test_struct_layout_matches_c(pk, tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "pcb200.h"\nint main(){printf("%zu %zu %zu %zu %zu\\n",'
                   'sizeof(pcb200_params),offsetof(pcb200_params,bcx),offsetof(pcb200_params,dx),'
                   'offsetof(pcb200_params,nu),sizeof(pcb200_forcing));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["/usr/bin/gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(v) for v in subprocess.check_output([str(exe)]).split()]
    P = pk.abi.Params
    assert got == [C.sizeof(P), P.bcx.offset, P.dx.offset, P.nu.offset, C.sizeof(pk.abi.Forcing)]


def test_no_device_fails_loudly(pk):
    """without a CUDA device the product refuses to run (no CPU fallback)"""
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    d = pk.GPU()
    with pytest.raises(pk.PCB200Error):
        d.gpu_init(0)


def test_product_does_not_import_oracle():
    """the product path never imports, links or calls anything under oracle/"""
    bad = re.compile(r"(import\s+oracle|from\s+oracle|libpcref|pcref_[a-z]|#include\s+\"[^\"]*pcref)")
    for dirpath, _, files in os.walk(os.path.join(ROOT, "pencil-code_b200")):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".f90", "Makefile")):
                txt = open(os.path.join(dirpath, fn), errors="ignore").read()
                assert not bad.search(txt), fn


def test_slab_decomposition_reproduces_global_fields(pk):
    Pg, _, gg = pk.setups.config_mhd(16)
    fg = pk.setups.synthetic_fields(Pg, gg)
    for ipz in range(2):
        P, _, gl = pk.setups.make_params(16, 16, 16, magnetic=True, nprocz=2, ipz=ipz, nu=5e-3, eta=5e-3, cdt=0.4, cdtv=0.3)
        fl = pk.setups.synthetic_fields(P, gl)
        assert P.nz == 8 and P.ipz == ipz
        assert np.allclose(fl, fg[:, ipz * 8: ipz * 8 + 14], atol=1e-14)


def test_two_rank_gloo_slab_exchange_matches_oracle_bc(pk, orc, tmp_path):
    """world_size-2 gloo run of the host-side z-slab logic: exchanging 3 ghost planes
    between slabs reproduces the single-domain periodic boundary condition."""
    script = tmp_path / "w.py"
    script.write_text(f"""
import os, sys
sys.path.insert(0, {ROOT!r})
import numpy as np, torch, torch.distributed as dist
import __graft_entry__ as g
pk = g.package()
rank = int(os.environ['RANK']); world = int(os.environ['WORLD_SIZE'])
dist.init_process_group('gloo', rank=rank, world_size=world)
P, _, grid = pk.setups.make_params(12, 12, 12 * world, magnetic=True, nprocz=world, ipz=rank, nu=5e-3, eta=5e-3)
f = pk.setups.synthetic_fields(P, grid)
ref = f.copy()
f[:, :3] = 0; f[:, -3:] = 0                      # forget the z ghosts
lo, hi = (rank - 1) % world, (rank + 1) % world  # same neighbour rule as exchange_z() in csrc/pcb200.cu
send_lo = torch.from_numpy(f[:, 3:6].copy()); send_hi = torch.from_numpy(f[:, -6:-3].copy())
recv_hi = torch.zeros_like(send_lo); recv_lo = torch.zeros_like(send_hi)
reqs = [dist.isend(send_lo, lo, tag=1), dist.isend(send_hi, hi, tag=2), dist.irecv(recv_hi, hi, tag=1), dist.irecv(recv_lo, lo, tag=2)]
for r in reqs: r.wait()
f[:, -3:] = recv_hi.numpy(); f[:, :3] = recv_lo.numpy()
ok = np.allclose(f, ref, atol=1e-14)
t = torch.tensor([1 if ok else 0]); dist.all_reduce(t, op=dist.ReduceOp.MIN)
dist.destroy_process_group()
sys.exit(0 if int(t[0]) else 1)
""")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533", str(script)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


def test_particle_struct_layout_matches_c(pk, tmp_path):
    src = tmp_path / "szp.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "pcb200.h"\nint main(){printf("%zu %zu %zu\\n",'
                   'sizeof(pcb200_particle_params),offsetof(pcb200_particle_params,tausp),'
                   'offsetof(pcb200_particle_params,cdtp_drag));return 0;}\n')
    exe = tmp_path / "szp"
    subprocess.check_call(["/usr/bin/gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(v) for v in subprocess.check_output([str(exe)]).split()]
    PP = pk.abi.ParticleParams
    assert got == [C.sizeof(PP), PP.tausp.offset, PP.cdtp_drag.offset]


def test_block_decomposition_tiles_the_global_grid(pk):
    """nprocx x nprocy x nprocz blocks with the reference's rank layout
    iproc = ipx + nprocx*(ipy + nprocy*ipz) (mpicomm.f90:367-369) reproduce the global fields."""
    Pg, _, gg = pk.setups.make_params(16, 24, 32, magnetic=True)
    fg = pk.setups.synthetic_fields(Pg, gg)
    nproc = (2, 3, 2)
    seen = np.zeros((32, 24, 16), dtype=int)
    for iproc in range(12):
        P, _, gl = pk.setups.make_params(16, 24, 32, magnetic=True, nproc=nproc, iproc=iproc)
        assert (P.nx, P.ny, P.nz) == (8, 8, 16)
        assert P.ipx + P.nprocx * (P.ipy + P.nprocy * P.ipz) == iproc
        fl = pk.setups.synthetic_fields(P, gl)
        sl = (slice(None), slice(P.ipz * 16, P.ipz * 16 + 22), slice(P.ipy * 8, P.ipy * 8 + 14), slice(P.ipx * 8, P.ipx * 8 + 14))
        assert np.allclose(fl, fg[sl], atol=1e-14)
        seen[P.ipz * 16:(P.ipz + 1) * 16, P.ipy * 8:(P.ipy + 1) * 8, P.ipx * 8:(P.ipx + 1) * 8] += 1
    assert (seen == 1).all()


def test_four_rank_gloo_xyz_exchange_fills_edges_and_corners(pk, tmp_path):
    """world_size-4 gloo run (2x2x1 blocks) of the ghost-zone rule of do_substep()/exchange_packed():
    axis by axis x -> y -> z, each face over the FULL ghost-inclusive extent of the other axes,
    with the periodic neighbour rule of neighbours() -- reproduces every ghost cell, edges and
    corners included, without corner messages."""
    script = tmp_path / "w4.py"
    script.write_text(f"""
import os, sys
sys.path.insert(0, {ROOT!r})
import numpy as np, torch, torch.distributed as dist
import __graft_entry__ as g
pk = g.package()
rank = int(os.environ['RANK']); world = int(os.environ['WORLD_SIZE'])
dist.init_process_group('gloo', rank=rank, world_size=world)
nproc = (2, 2, 1)
P, _, grid = pk.setups.make_params(16, 16, 8, magnetic=True, nproc=nproc, iproc=rank)
f = pk.setups.synthetic_fields(P, grid)          # analytic field: ghosts already hold the periodic images
ref = f.copy()
f[:, :3] = 0; f[:, -3:] = 0; f[:, :, :3] = 0; f[:, :, -3:] = 0; f[:, :, :, :3] = 0; f[:, :, :, -3:] = 0
ip = [P.ipx, P.ipy, P.ipz]
def rank_of(axis, i):
    q = list(ip); q[axis] = i % nproc[axis]
    return q[0] + nproc[0] * (q[1] + nproc[1] * q[2])
for axis in (0, 1, 2):                           # x -> y -> z
    ax = 3 - axis                                # numpy axis of f(mvar, mz, my, mx)
    n = f.shape[ax] - 6
    take = lambda a, b: np.ascontiguousarray(np.take(f, range(a, b), axis=ax))
    def put(a, vals):
        idx = [slice(None)] * 4; idx[ax] = slice(a, a + 3); f[tuple(idx)] = vals
    if nproc[axis] == 1:                         # wraps inside the rank
        put(0, take(n, n + 3)); put(n + 3, take(3, 6))
        continue
    lo, hi = rank_of(axis, ip[axis] - 1), rank_of(axis, ip[axis] + 1)
    s_lo, s_hi = torch.from_numpy(take(3, 6)), torch.from_numpy(take(n, n + 3))
    r_hi, r_lo = torch.zeros_like(s_lo), torch.zeros_like(s_hi)
    reqs = [dist.isend(s_lo, lo, tag=10 + axis), dist.irecv(r_hi, hi, tag=10 + axis),
            dist.isend(s_hi, hi, tag=20 + axis), dist.irecv(r_lo, lo, tag=20 + axis)]
    for r in reqs: r.wait()
    put(n + 3, r_hi.numpy()); put(0, r_lo.numpy())
ok = np.allclose(f, ref, atol=1e-14)
t = torch.tensor([1 if ok else 0]); dist.all_reduce(t, op=dist.ReduceOp.MIN)
dist.destroy_process_group()
sys.exit(0 if int(t[0]) else 1)
""")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "4",
                        "--master-addr", "127.0.0.1", "--master-port", "29534", str(script)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
