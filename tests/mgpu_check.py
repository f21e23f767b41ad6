"""Multi-GPU parity check (run under torchrun, one rank per GPU): every block
decomposition nprocx x nprocy x nprocz of the world size (z-slabs, x/y faces,
2x2x2 ...) with NCCL halo exchange vs the single-domain oracle.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
      --master-port 29511 tests/mgpu_check.py
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pk = entry.package()
    dev = pk.GPU()
    dev.gpu_init(local)
    idt = torch.zeros(128, dtype=torch.uint8)
    if rank == 0:
        idt = torch.frombuffer(bytearray(dev.nccl_unique_id()), dtype=torch.uint8).clone()
    dist.broadcast(idt, 0)
    dev.comm_init(idt.numpy().tobytes(), rank, world)
    ok = True
    nsteps = 5
    # global x,y size: 32 by default; PCB200_MGPU_NXY=208,48 gives >= 3 tile columns and rows per rank, which
    # switches on the overlap of the x/y face exchange with the interior tiles (and has ragged edge tiles)
    gx, gy = (int(v) for v in os.environ.get("PCB200_MGPU_NXY", "32,32").split(","))
    decomps = [(px, py, world // (px * py)) for px in (1, 2, 4, 8) for py in (1, 2, 4, 8)
               if world % (px * py) == 0 and gx // px >= 8 and gy // py >= 8]
    if os.environ.get("PCB200_NPROC"):
        decomps = [tuple(int(v) for v in os.environ["PCB200_NPROC"].split(","))]
    for nproc in decomps:
        gz = 16 * nproc[2]
        cases = {
            "mhd_periodic": lambda npr, ipr: pk.setups.make_params(gx, gy, gz, magnetic=True, nproc=npr, iproc=ipr,
                                                                   nu=5e-3, eta=5e-3, cdt=0.4, cdtv=0.3),
            "convslab_walls": lambda npr, ipr: pk.setups.config_convslab(gx, gy, gz, nproc=npr, iproc=ipr),
        }
        for name, mk in cases.items():
            P, zprof, grid = mk(nproc, rank)
            f = pk.setups.synthetic_fields(P, grid)
            if name.startswith("convslab"):
                f[P.ilnrho - 1] *= 0.1
            dev.initialize_GPU(P, zprof)
            dev.register_GPU(f)
            dts = []
            for _ in range(nsteps):
                for isub in (1, 2, 3):
                    dt = dev.rhs_GPU(f, isub)
                dts.append(dt)
            dev.copy_farray_from_GPU(f)
            inner = torch.from_numpy(np.ascontiguousarray(f[:, 3:-3, 3:-3, 3:-3]))
            parts = [torch.zeros_like(inner) for _ in range(world)] if rank == 0 else None
            dist.gather(inner, parts, 0)
            if rank == 0:
                orc = entry.oracle()
                Pg, zpg, gg = mk((1, 1, 1), 0)
                fo = pk.setups.synthetic_fields(Pg, gg)
                if name.startswith("convslab"):
                    fo[Pg.ilnrho - 1] *= 0.1
                orc.set_zprofiles(zpg)
                dfo = np.zeros_like(fo)
                dto, d = [], 0.0
                for _ in range(nsteps):
                    d = orc.timestep(fo, dfo, Pg, d)
                    dto.append(d)
                ref = fo[:, 3:-3, 3:-3, 3:-3]
                got = np.zeros_like(ref)
                nz, ny, nx = inner.shape[1:]
                for r, part in enumerate(parts):      # iproc = ipx + nprocx*(ipy + nprocy*ipz)
                    ix, iy, iz = r % nproc[0], (r // nproc[0]) % nproc[1], r // (nproc[0] * nproc[1])
                    got[:, iz * nz:(iz + 1) * nz, iy * ny:(iy + 1) * ny, ix * nx:(ix + 1) * nx] = part.numpy()
                err = max(np.abs(got[k] - ref[k]).max() / max(np.abs(ref[k]).max(), 1e-300) for k in range(ref.shape[0]))
                dterr = np.abs(np.array(dts) / np.array(dto) - 1).max()
                good = err < 1e-12 and dterr < 1e-13
                ok &= good
                print(f"[mgpu] {name}: nproc={nproc} rel Linf={err:.3e} dt err={dterr:.3e} {'OK' if good else 'FAIL'}", flush=True)
    flag = torch.tensor([1 if ok else 0])
    dist.broadcast(flag, 0)
    dev.finalize_GPU()
    dist.destroy_process_group()
    sys.exit(0 if int(flag[0]) else 1)


if __name__ == "__main__":
    main()
