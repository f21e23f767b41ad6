"""Multi-GPU particle parity check (run under torchrun, one rank per GPU).

Dust particles on a block-decomposed grid: the TSC deposit goes through the ghost-layer
fold between ranks (fold_df/fold_f) and particles that leave a rank's domain migrate
(migrate_particles).  Compared with the single-domain oracle run on rank 0; particles are
matched by their identity ipar because the storage order is per rank.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
      --master-port 29512 tests/mgpu_particles_check.py
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
    nsteps = int(os.environ.get("PCB200_PSTEPS", "6"))
    npar_glob = 24000
    decomps = [(px, py, world // (px * py)) for px in (1, 2) for py in (1, 2) if world % (px * py) == 0]
    if os.environ.get("PCB200_NPROC"):
        decomps = [tuple(int(v) for v in os.environ["PCB200_NPROC"].split(","))]
    for nproc in decomps:
        n = (32, 32, 16 * nproc[2])
        mk = lambda npr, ipr: pk.setups.make_params(n[0], n[1], n[2], nproc=npr, iproc=ipr, nu=5e-3, cdt=0.4, cdtv=0.3)
        Pg, zpg, gg = mk((1, 1, 1), 0)
        ppg = pk.setups.make_particle_params(Pg, npar_glob, tausp=0.05, eps_dtog=1.0)
        stg = pk.setups.synthetic_particles(Pg, ppg, vampl=0.3)       # fast particles: many cross rank boundaries
        P, zprof, grid = mk(nproc, rank)
        f = pk.setups.synthetic_fields(P, grid)
        # this rank's particles: those inside its block (migrate_particles semantics: [x0_loc, x1_loc) )
        lo = np.array([P.xyz0[j] + [P.ipx, P.ipy, P.ipz][j] * P.Lxyz[j] / nproc[j] for j in range(3)])
        hi = lo + np.array([P.Lxyz[j] / nproc[j] for j in range(3)])
        mine = np.all((stg.fp[:3] >= lo[:, None]) & (stg.fp[:3] < hi[:, None]), axis=0)
        nloc = int(mine.sum())
        mpar = int(npar_glob / world * 1.5) + 64
        pp = pk.setups.make_particle_params(P, nloc, mpar, tausp=0.05, rhop_swarm=ppg.rhop_swarm)
        st = pk.setups.ParticleState(P, nloc, mpar)
        st.fp[:, :nloc] = stg.fp[:, mine]
        st.ipar[:nloc] = stg.ipar[mine]
        dev.initialize_GPU(P, zprof)
        dev.particles_initialize(pp)
        dev.register_GPU(f)
        dev.particles_load(st)
        dts = []
        for _ in range(nsteps):
            for isub in (1, 2, 3):
                dt = dev.rhs_GPU(f, isub)
            dts.append(dt)
        dev.copy_farray_from_GPU(f)
        nnow = dev.particles_store(st)
        # gather the fields and the particles (padded to mpar) on rank 0
        inner = torch.from_numpy(np.ascontiguousarray(f[:, 3:-3, 3:-3, 3:-3]))
        parts = [torch.zeros_like(inner) for _ in range(world)] if rank == 0 else None
        dist.gather(inner, parts, 0)
        pack = np.zeros((8, mpar))
        pack[:6, :nnow] = st.fp[:, :nnow]
        pack[6, :nnow] = st.ipar[:nnow]
        pack[7, 0] = nnow
        tp = torch.from_numpy(pack)
        allp = [torch.zeros_like(tp) for _ in range(world)] if rank == 0 else None
        dist.gather(tp, allp, 0)
        rh = torch.from_numpy(np.ascontiguousarray(st.rhop[3:-3, 3:-3, 3:-3]))
        allr = [torch.zeros_like(rh) for _ in range(world)] if rank == 0 else None
        dist.gather(rh, allr, 0)
        if rank == 0:
            orc = entry.oracle()
            fo = pk.setups.synthetic_fields(Pg, gg)
            dfo = np.zeros_like(fo)
            sto = stg.copy()
            dto, d = [], 0.0
            for _ in range(nsteps):
                d = orc.timestep_particles(fo, dfo, Pg, ppg, sto, d)
                dto.append(d)
            ref = fo[:, 3:-3, 3:-3, 3:-3]
            got = np.zeros_like(ref)
            gotr = np.zeros_like(ref[0])
            nz, ny, nx = inner.shape[1:]
            for r, part in enumerate(parts):
                ix, iy, iz = r % nproc[0], (r // nproc[0]) % nproc[1], r // (nproc[0] * nproc[1])
                got[:, iz * nz:(iz + 1) * nz, iy * ny:(iy + 1) * ny, ix * nx:(ix + 1) * nx] = part.numpy()
                gotr[iz * nz:(iz + 1) * nz, iy * ny:(iy + 1) * ny, ix * nx:(ix + 1) * nx] = allr[r].numpy()
            err = max(np.abs(got[k] - ref[k]).max() / max(np.abs(ref[k]).max(), 1e-300) for k in range(ref.shape[0]))
            dterr = np.abs(np.array(dts) / np.array(dto) - 1).max()
            fpg = np.concatenate([a.numpy()[:6, :int(a.numpy()[7, 0])] for a in allp], axis=1)
            ipg = np.concatenate([a.numpy()[6, :int(a.numpy()[7, 0])] for a in allp]).astype(np.int64)
            counts = [int(a.numpy()[7, 0]) for a in allp]
            complete = fpg.shape[1] == npar_glob and (np.sort(ipg) == np.arange(1, npar_glob + 1)).all()
            perr = np.inf
            if complete:
                a = fpg[:, np.argsort(ipg)]
                b = sto.fp[:, np.argsort(sto.ipar[:npar_glob])]
                L = np.array([Pg.Lxyz[j] for j in range(3)])[:, None]
                dx = np.abs(a[:3] - b[:3]); dx = np.minimum(dx, L - dx)
                perr = max(dx.max(), np.abs(a[3:] - b[3:]).max())
            rerr = np.abs(gotr - sto.rhop[3:-3, 3:-3, 3:-3]).max() / sto.rhop.max()
            good = err < 1e-10 and dterr < 1e-11 and complete and perr < 1e-10 and rerr < 1e-10
            ok &= good
            print(f"[mgpu-part] nproc={nproc} particles/rank={counts} f rel Linf={err:.2e} dt err={dterr:.1e} "
                  f"all particles present={complete} fp err={perr:.2e} rhop err={rerr:.2e} {'OK' if good else 'FAIL'}", flush=True)
    flag = torch.tensor([1 if ok else 0])
    dist.broadcast(flag, 0)
    dev.finalize_GPU()
    dist.destroy_process_group()
    sys.exit(0 if int(flag[0]) else 1)


if __name__ == "__main__":
    main()
