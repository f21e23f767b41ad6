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
