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


def run_reference(args):
    """CPU arm: the reference's own algorithm (oracle port, OpenMP over pencils)
    on the box's host cores; each step is a bounded sample (ncpu^3 box) of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    pk = entry.package()
    orc = entry.oracle()
    n = args.cpu_n
    P, zprof, grid = pk.setups.config_mhd(n)
    f = pk.setups.synthetic_fields(P, grid)
    df = np.zeros_like(f)
    dt = 0.0
    for _ in range(args.warmup):
        dt = orc.timestep(f, df, P, dt)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        dt = orc.timestep(f, df, P, dt)
    el = time.perf_counter() - t0
    cells = n ** 3
    val = cells * 3 * args.steps / el / 1e6
    cores = orc.num_threads()
    sample = f"{n}^3 periodic isothermal MHD box (same synthetic fields), {args.steps} full RK3 steps"
    out = {
        "impl": "reference", "metric": "Mcell-updates/s per RK3 substep", "value": val, "unit": "Mcell-updates/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": el / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "isothermal MHD 512^3 per GPU (BASELINE config 3), CPU arm on a bounded sample", "sample": sample},
        "cpu_baseline": {"value": val, "unit": "Mcell-updates/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "Mcell-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(out), flush=True)


def cpu_baseline(pk, n, budget_s=15.0):
    orc = entry.oracle()
    P, zprof, grid = pk.setups.config_mhd(n)
    f = pk.setups.synthetic_fields(P, grid)
    df = np.zeros_like(f)
    dt = orc.timestep(f, df, P, 0.0)
    t0 = time.perf_counter()
    steps = 0
    while steps < 3 or (time.perf_counter() - t0 < budget_s and steps < 50):
        dt = orc.timestep(f, df, P, dt)
        steps += 1
    el = time.perf_counter() - t0
    return {"value": n ** 3 * 3 * steps / el / 1e6, "unit": "Mcell-updates/s", "cores": orc.num_threads(), "kind": "port",
            "sample": f"{n}^3 periodic isothermal MHD box, {steps} full RK3 steps, oracle/pcref.c with OpenMP over pencils"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--n", "--ncell", dest="n", type=int, default=512, help="grid points per direction per GPU (use --ncell under torchrun, whose parser trips over --n)")
    ap.add_argument("--cpu-n", type=int, default=128)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--particles", type=int, default=0,
                    help="dust particles per GPU (BASELINE config 5 is 256^3 cells + 1.25e7 particles per GPU)")
    ap.add_argument("--entropy", action="store_true", help="BASELINE config 4 physics: MHD + entropy (chi-const), mvar=8")
    ap.add_argument("--decomp", default="", help="nprocx,nprocy,nprocz (default: z-slabs 1,1,N)")
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

    n = args.n
    # weak scaling: every GPU owns an n^3 slab, global box n x n x (n*world), dz kept equal
    L = 2 * np.pi
    nproc = tuple(int(v) for v in args.decomp.split(",")) if args.decomp else (1, 1, world)
    assert nproc[0] * nproc[1] * nproc[2] == world, "--decomp does not multiply to the world size"
    extra = dict(entropy=True, chi=1e-4, iheatcond=pk.abi.HCOND["chi-const"]) if args.entropy else {}
    P, zprof, grid = pk.setups.make_params(n * nproc[0], n * nproc[1], n * nproc[2], magnetic=True, nproc=nproc, iproc=rank,
                                           Lxyz=(L * nproc[0], L * nproc[1], L * nproc[2]), nu=5e-3, eta=5e-3, cdt=0.4, cdtv=0.3,
                                           **extra)
    f = pk.setups.synthetic_fields(P, grid)
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

    cells = n ** 3 * world
    for _ in range(args.warmup):
        dev.time_step()
    # ---- device-resident timed region ---------------------------------------
    barrier()
    dev.launch_count()
    dev.enable_kernel_timing(True)
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    dev.timer_start()
    for _ in range(args.steps):
        dt = dev.time_step()
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
            dev.time_step()
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

    if rank != 0:
        return
    peak, peak_src = peaks()
    b_alg = B_ALG_PER_VAR * P.mvar
    cells_local = n ** 3
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
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": (f"MHD + entropy box (BASELINE config 4 physics), {n}^3 cells per GPU, mvar=8" if args.entropy else
                                f"isothermal helical-MHD box (BASELINE config 3), {n}^3 cells per GPU, mvar=7"),
                   "global_grid": [n * nproc[0], n * nproc[1], n * nproc[2]], "decomposition": list(nproc), "step": "one RK3 time step = 3 substeps incl. BCs, CFL dt, 2N update",
                   "particles_per_gpu": args.particles, "l2": f"inputs larger than L2 ({f.nbytes / 1e9:.2f} GB per buffer)", "dt": dt},
        "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak if ach else None,
                     "traffic": traffic, "peak_source": peak_src, "kernel": "%s<MAG=1,ENT=%d>" % (dev.kernel_name(), int(args.entropy)), "launches_timed": kcount,
                     "avg_launch_ms": kms, "kernel_ms_per_substep": ksub, "algorithmic_bytes_per_cell": b_alg,
                     "step_frac": cells_local * b_alg * 3 / (ms / args.steps * 1e-3) / 1e9 / peak,
                     # where the CFL pre-pass does not apply (entropy, PCB200_PREPASS=0) the f write of substep 1 is
                     # done by axpy_update, not by this kernel: its own share of the SURVEY 8d bytes is then 9/10
                     "frac_kernel_own_bytes": ((0.9 if (args.entropy or os.environ.get("PCB200_PREPASS", "1") == "0") else 1.0)
                                               * ach / peak) if ach else None},
        "clocks": clk, "gpu_launches": launches,
    }
    if e2e:
        out["e2e"] = e2e
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(pk, args.cpu_n)
    print(json.dumps(out), flush=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
