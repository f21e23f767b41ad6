#!/usr/bin/env python
"""bench.py -- Mcell-updates/s per RK3 substep of the fused RHS/RK3 hot path.

  python bench.py --gpus N --steps K --warmup W            (this framework, B200)
  python bench.py --impl reference --steps K --warmup W     (CPU arm: the oracle
        port of the reference's rhs_cpu/time_step on all host cores)

A "step" is one full time step = 3 RK substeps (BCs, RHS, dt, 2N update) of
isothermal MHD (BASELINE config 3) over a 512^3 slab per GPU; at N>1 the
slabs are stacked along z (weak scaling, halo exchange over NCCL).  Prints
ONE JSON line on rank 0.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

B_ALG_PER_VAR = 8.0 * 10.0 / 3.0       # bytes per cell-substep per variable


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for nm, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


PUBLISHED_US_PER_PT_STEP_CORE = 0.70      # reference's own figure, doc/manual.tex:8942 (its hardware); = 4.3 Mcell-substeps/s/core


def timing_oracle(pk):
    """CPU arm: the oracle sources in their TIMING build (-O3 -march=native -fopenmp, compiled on this machine) with
    every host core this process may use -- torchrun's OMP_NUM_THREADS=1 is overridden explicitly."""
    from oracle import pcref
    fast = pcref.build_fast()
    orc = pcref.Oracle(pk.abi, fast)
    orc.set_num_threads(len(os.sched_getaffinity(0)))
    build = "-O3 -march=native -fopenmp" if fast else "-O2 -ffp-contract=off -fopenmp (checker build; timing build failed)"
    return orc, build


def cpu_arm_setup(pk, args):
    """the GPU arm's physics on a bounded n^3 sample of its grid"""
    n = args.cpu_n
    if args.convslab:
        P, zprof, grid = pk.setups.config_convslab(n, n, n)
        name = "conv-slab (hydro + entropy + gravity, walls in z)"
    elif args.hydro:
        P, zprof, grid = pk.setups.config_hydro(n)
        name = "isothermal hydro"
    elif args.entropy:
        P, zprof, grid = pk.setups.config_mhd_entropy(n)
        name = "MHD + entropy"
    else:
        P, zprof, grid = pk.setups.config_mhd(n)
        name = "isothermal MHD"
    return P, zprof, grid, name


def bench_fields(pk, args, P, grid):
    """deterministic synthetic fields; conv-slab: small density contrast so that the stratified walls stay tame"""
    f = pk.setups.synthetic_fields(P, grid)
    if args.convslab:
        f[P.ilnrho - 1] *= 0.1
    return f


def run_reference(args):
    """CPU arm: the reference's own algorithm (oracle port, OpenMP over pencils) on the box's host cores; each step is
    one RK3 step of a bounded sample (cpu_n^3 box, default 256^3) of the workload.  Rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    pk = entry.package()
    orc, build = timing_oracle(pk)
    n = args.cpu_n
    P, zprof, grid, name = cpu_arm_setup(pk, args)
    orc.set_zprofiles(zprof)
    f = bench_fields(pk, args, P, grid)
    df = np.zeros_like(f)
    dt = 0.0
    # pencil-wise evaluation like the reference's mn-loop where it applies (BASELINE configs 1-4), else the point-wise port
    for _ in range(args.warmup):
        dt = orc.timestep_pencil(f, df, P, dt)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        dt = orc.timestep_pencil(f, df, P, dt)
    el = time.perf_counter() - t0
    cells = n ** 3
    val = cells * 3 * args.steps / el / 1e6
    cores = orc.num_threads()
    how = "pencil-wise (oracle/pcref_pencil.c)" if orc.pencil_applies(P) else "point-wise (oracle/pcref.c)"
    sample = f"{n}^3 periodic {name} box (same synthetic fields), {args.steps} full RK3 steps, {how}, built {build}"
    out = {
        "impl": "reference", "metric": "Mcell-updates/s per RK3 substep", "value": val, "unit": "Mcell-updates/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": el / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args, args.n), "sample": sample},
        "cpu_baseline": {"value": val, "unit": "Mcell-updates/s", "cores": cores, "kind": "port", "sample": sample,
                         "per_core": val / cores,
                         "published_per_core": 3.0 / PUBLISHED_US_PER_PT_STEP_CORE,
                         "published_note": "reference manual: 0.70 us/pt/step/core on its own hardware (doc/manual.tex:8942); "
                                           "a step is 3 substeps, so 4.3 Mcell-substeps/s/core"},
        "e2e": {"value": val, "unit": "Mcell-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(out), flush=True)


def cpu_baseline(pk, args, budget_s=15.0):
    orc, build = timing_oracle(pk)
    n = args.cpu_n
    P, zprof, grid, name = cpu_arm_setup(pk, args)
    orc.set_zprofiles(zprof)
    f = bench_fields(pk, args, P, grid)
    df = np.zeros_like(f)
    dt = orc.timestep_pencil(f, df, P, 0.0)
    t0 = time.perf_counter()
    steps = 0
    while steps < 2 or (time.perf_counter() - t0 < budget_s and steps < 200):
        dt = orc.timestep_pencil(f, df, P, dt)
        steps += 1
    el = time.perf_counter() - t0
    val = n ** 3 * 3 * steps / el / 1e6
    return {"value": val, "unit": "Mcell-updates/s", "cores": orc.num_threads(), "kind": "port", "per_core": val / orc.num_threads(),
            "published_per_core": 3.0 / PUBLISHED_US_PER_PT_STEP_CORE,
            "sample": f"{n}^3 periodic {name} box, {steps} full RK3 steps, "
                      + ("pencil-wise evaluation like the reference's mn-loop (oracle/pcref_pencil.c)" if orc.pencil_applies(P)
                         else "point-wise port (oracle/pcref.c)") + f", built {build}, OpenMP over pencils"}


def parity_check(pk, dev, dist, rank, world, nproc, args, nsteps=3):
    """Before anything is timed: the SAME library, physics and decomposition on a small global grid
    (64 x 64 x 16 cells per rank block in x,y; 16 planes per z rank), nsteps RK3 steps, against the single-domain
    oracle run on rank 0 (checker only).  Returns the JSON block; bench refuses to print a value when ok is false."""
    bx, by, bz = 64 // nproc[0] if nproc[0] > 1 else 64, 64 // nproc[1] if nproc[1] > 1 else 64, 16
    bx, by = max(bx, 16), max(by, 16)
    gx, gy, gz = bx * nproc[0], by * nproc[1], bz * nproc[2]
    kw = dict(nu=5e-3, eta=5e-3, cdt=0.4, cdtv=0.3) if not args.hydro else dict(nu=5e-3, cdt=0.4, cdtv=0.3)
    if args.entropy:
        kw.update(entropy=True, chi=1e-4, iheatcond=pk.abi.HCOND["chi-const"])

    def mk(npr, ipr):
        if args.convslab:
            return pk.setups.config_convslab(gx, gy, gz, nproc=npr, iproc=ipr)
        return pk.setups.make_params(gx, gy, gz, magnetic=not args.hydro, nproc=npr, iproc=ipr, **kw)
    P, zprof, grid = mk(nproc, rank)
    f = bench_fields(pk, args, P, grid)
    dev.initialize_GPU(P, zprof)
    dev.register_GPU(f)
    dts = []
    for _ in range(nsteps):
        for isub in (1, 2, 3):
            dt = dev.rhs_GPU(f, isub)
        dts.append(dt)
    dev.copy_farray_from_GPU(f)
    inner = np.ascontiguousarray(f[:, 3:-3, 3:-3, 3:-3])
    parts = None
    if dist is not None:
        import torch
        t = torch.from_numpy(inner)
        parts = [torch.zeros_like(t) for _ in range(world)] if rank == 0 else None
        dist.gather(t, parts, 0)
        parts = [q.numpy() for q in parts] if rank == 0 else None
    else:
        parts = [inner]
    out = None
    if rank == 0:
        orc = entry.oracle()
        Pg, zpg, gg = mk((1, 1, 1), 0)
        fo = bench_fields(pk, args, Pg, gg)
        orc.set_zprofiles(zpg)
        dfo = np.zeros_like(fo)
        dto, d = [], 0.0
        for _ in range(nsteps):
            d = orc.timestep(fo, dfo, Pg, d)
            dto.append(d)
        ref = fo[:, 3:-3, 3:-3, 3:-3]
        got = np.zeros_like(ref)
        nz, ny, nx = inner.shape[1:]
        for r, part in enumerate(parts):          # iproc = ipx + nprocx*(ipy + nprocy*ipz), mpicomm.f90:367-369
            ix, iy, iz = r % nproc[0], (r // nproc[0]) % nproc[1], r // (nproc[0] * nproc[1])
            got[:, iz * nz:(iz + 1) * nz, iy * ny:(iy + 1) * ny, ix * nx:(ix + 1) * nx] = part
        err = max(float(np.abs(got[k] - ref[k]).max() / max(np.abs(ref[k]).max(), 1e-300)) for k in range(ref.shape[0]))
        dterr = float(np.abs(np.array(dts) / np.array(dto) - 1).max())
        out = {"grid": [gx, gy, gz], "decomposition": list(nproc), "steps": nsteps, "rel_linf": err, "dt_err": dterr,
               "tol": {"rel_linf": 1e-12, "dt_err": 1e-13}, "against": "oracle/pcref.c, single domain, rank 0",
               "kernel": dev.kernel_name(), "ok": bool(err < 1e-12 and dterr < 1e-13)}
    if dist is not None:
        import torch
        flag = torch.tensor([1 if (out is None or out["ok"]) else 0])
        dist.broadcast(flag, 0)
    return out


def workload_name(args, n):
    if args.convslab:
        return f"samples/conv-slab physics (BASELINE config 1: hydro + entropy + gravity, K-profile, cooling, walls in z), {n}^3 cells per GPU, mvar=5"
    if args.hydro:
        return f"periodic isothermal hydro box (BASELINE config 2), {n}^3 cells per GPU, mvar=4"
    if args.entropy:
        return f"MHD + entropy box (BASELINE config 4 physics), {n}^3 cells per GPU, mvar=8"
    return f"isothermal helical-MHD box (BASELINE config 3), {n}^3 cells per GPU, mvar=7"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--n", "--ncell", dest="n", type=int, default=512, help="grid points per direction per GPU (use --ncell under torchrun, whose parser trips over --n)")
    ap.add_argument("--cpu-n", type=int, default=256, help="CPU arm: edge of the bounded sample box")
    ap.add_argument("--hydro", action="store_true", help="BASELINE config 2 physics: isothermal hydro, mvar=4 (use with --ncell 256)")
    ap.add_argument("--forcing", action="store_true", help="helical forcing kick after substep 3 of every timed step (config 2/3 are forced runs)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity-check", action="store_true", help="skip the decomposed small-grid check against the oracle before timing")
    ap.add_argument("--particles", type=int, default=0,
                    help="dust particles per GPU (BASELINE config 5 is 256^3 cells + 1.25e7 particles per GPU)")
    ap.add_argument("--entropy", action="store_true", help="BASELINE config 4 physics: MHD + entropy (chi-const), mvar=8")
    ap.add_argument("--convslab", action="store_true", help="BASELINE config 1 physics: samples/conv-slab (use with --ncell 32 or 128; shipped resolution 32^3)")
    ap.add_argument("--decomp", default="", help="nprocx,nprocy,nprocz (default: z-slabs 1,1,N)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: n^3 cells per GPU (default, the driver's scaling run); strong: n^3 cells in total (SURVEY 8e table)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group(backend="cpu:gloo,cuda:nccl", rank=rank, world_size=world)

    pk = entry.package()
    dev = pk.GPU()
    dev.gpu_init(local_rank)
    if world > 1:
        import torch
        idt = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            idt = torch.frombuffer(bytearray(dev.nccl_unique_id()), dtype=torch.uint8).clone()
        dist.broadcast(idt, 0)
        dev.comm_init(bytes(idt.numpy().tobytes()), rank, world)

    nproc = tuple(int(v) for v in args.decomp.split(",")) if args.decomp else (1, 1, world)
    assert nproc[0] * nproc[1] * nproc[2] == world, "--decomp does not multiply to the world size"
    parity = None if args.no_parity_check else parity_check(pk, dev, dist, rank, world, nproc, args)
    n = args.n
    # weak scaling: every GPU owns an n^3 block, global box n*nprocx x n*nprocy x n*nprocz, grid spacing kept equal;
    # strong scaling: the global box is n^3 and every GPU owns n/nprocx x n/nprocy x n/nprocz of it
    L = 2 * np.pi
    strong = args.scaling == "strong"
    gdim = (n, n, n) if strong else (n * nproc[0], n * nproc[1], n * nproc[2])
    Lx = (L, L, L) if strong else (L * nproc[0], L * nproc[1], L * nproc[2])
    extra = dict(entropy=True, chi=1e-4, iheatcond=pk.abi.HCOND["chi-const"]) if args.entropy else {}
    if args.convslab:
        P, zprof, grid = pk.setups.config_convslab(gdim[0], gdim[1], gdim[2], nproc=nproc, iproc=rank)
    else:
        P, zprof, grid = pk.setups.make_params(gdim[0], gdim[1], gdim[2], magnetic=not args.hydro, nproc=nproc, iproc=rank,
                                               Lxyz=Lx, nu=5e-3, eta=5e-3, cdt=0.4, cdtv=0.3,
                                               **extra)
    f = bench_fields(pk, args, P, grid)
    dev.initialize_GPU(P, zprof)
    dev.register_GPU(f)
    dev.load_farray()
    if args.particles:
        pp = pk.setups.make_particle_params(P, args.particles, int(args.particles * 1.25), tausp=0.1, eps_dtog=0.5)
        pst = pk.setups.synthetic_particles(P, pp, block_only=True)
        dev.particles_initialize(pp)
        dev.particles_load(pst)

    def barrier():
        dev.sync()
        if dist is not None:
            dist.barrier()

    cells = gdim[0] * gdim[1] * gdim[2]
    fseq = pk.setups.forcing_sequence(args.steps + args.warmup) if args.forcing else None
    state = {"dt": 1e-3, "i": 0}

    def one_step():
        """one time step; with --forcing the host draws this step's kick first (48 B/cell more traffic per step)"""
        if fseq is not None:
            k, ph, phi = fseq[state["i"] % len(fseq)]
            state["i"] += 1
            dev.set_forcing(pk.setups.helical_forcing_coefs(k, ph, phi, state["dt"]))
        state["dt"] = dev.time_step()
        return state["dt"]

    for _ in range(args.warmup):
        one_step()
    # ---- device-resident timed region ---------------------------------------
    barrier()
    dev.launch_count()
    dev.enable_kernel_timing(True)
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    dev.timer_start()
    for _ in range(args.steps):
        dt = one_step()
    ms = dev.timer_stop()
    barrier()
    clk = clocks.stop() if rank == 0 else None
    launches = dev.launch_count()
    kms, kcount = dev.rhs_kernel_time()
    dev.enable_kernel_timing(False)
    if dist is not None:
        import torch
        t = torch.tensor([ms], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
    value = cells * 3 * args.steps / (ms * 1e-3) / 1e6

    # ---- end to end through the host-facing call: host f -> device, K steps, f back -------
    e2e = None
    if not args.no_e2e:
        barrier()
        t0 = time.perf_counter()
        dev._lvery_first = True                      # `full` upload of the registered host array (pinned once)
        for _ in range(args.steps):
            one_step()
        dev.copy_farray_from_GPU()                   # snapshot: device -> host
        dev.sync()
        el = time.perf_counter() - t0
        if dist is not None:
            import torch
            t = torch.tensor([el], dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            el = float(t[0])
        e2e = {"value": cells * 3 * args.steps / el / 1e6, "unit": "Mcell-updates/s",
               "h2d_bytes_per_step": f.nbytes / args.steps, "d2h_bytes_per_step": f.nbytes / args.steps + 16,
               "note": "registered host f uploaded once (rhs_GPU full=1), K steps, copy_farray_from_GPU once; "
                       "dt returned to the host every step"}

    if os.environ.get("PCB200_TIME_GHOST"):
        dev.finalize_GPU()        # prints the phase timers of this rank to stderr
    if rank != 0:
        return
    peak, peak_src = peaks()
    b_alg = B_ALG_PER_VAR * P.mvar
    cells_local = cells // world
    # kernel time per substep: at N>1 a substep is 3 launches (interior + 2 boundary slabs)
    ksub = kms * kcount / (3 * args.steps) if kcount else 0.0
    ach = cells_local * b_alg / (ksub * 1e-3) / 1e9 if ksub > 0 else None
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get("dram_bytes_per_launch")
        except Exception:
            pass
    out = {
        "metric": "Mcell-updates/s per RK3 substep", "value": value, "unit": "Mcell-updates/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args, n) if not strong else workload_name(args, n).replace("cells per GPU", "cells in total"),
                   "global_grid": list(gdim), "decomposition": list(nproc), "step": "one RK3 time step = 3 substeps incl. BCs, CFL dt, 2N update",
                   "particles_per_gpu": args.particles, "forcing": bool(args.forcing), "l2": f"inputs larger than L2 ({f.nbytes / 1e9:.2f} GB per buffer)", "dt": dt},
        "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak if ach else None,
                     "traffic": traffic, "peak_source": peak_src, "kernel": "%s<MAG=%d,ENT=%d>" % (dev.kernel_name(), int(P.lmagnetic), int(P.lentropy)), "launches_timed": kcount,
                     "avg_launch_ms": kms, "kernel_ms_per_substep": ksub, "algorithmic_bytes_per_cell": b_alg,
                     "step_frac": cells_local * b_alg * 3 / (ms / args.steps * 1e-3) / 1e9 / peak,
                     # where the CFL pre-pass does not apply (entropy, PCB200_PREPASS=0) the f write of substep 1 is
                     # done by axpy_update, not by this kernel: its own share of the SURVEY 8d bytes is then 9/10
                     "frac_kernel_own_bytes": ((0.9 if (P.lentropy or os.environ.get("PCB200_PREPASS", "1") == "0") else 1.0)
                                               * ach / peak) if ach else None},
        "clocks": clk, "gpu_launches": launches, "parity_check": parity,
    }
    if parity is not None and not parity["ok"]:
        # a fast wrong answer is not a result: no value, no roofline
        print(json.dumps({"metric": out["metric"], "value": None, "unit": out["unit"], "n_gpus": world, "parity_check": parity,
                          "error": "parity check against the oracle failed; timing withheld"}), flush=True)
        if dist is not None:
            dist.destroy_process_group()
        sys.exit(1)
    if e2e:
        out["e2e"] = e2e
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(pk, args)
    print(json.dumps(out), flush=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
