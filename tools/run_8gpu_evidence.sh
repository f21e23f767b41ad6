#!/bin/bash
# 8-GPU evidence run (one gpurun --gpus 8 call): every decomposition against the single-domain oracle,
# particle migration/fold, and the bench records of BASELINE configs 3, 4, 5 with the ghost-zone phase timers.
# usage: tools/run_8gpu_evidence.sh <tag>     (writes gpurun_out/<tag>_*)
tag=${1:-r02j}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/${tag}_devices.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533"
timeout 900 $TR tests/mgpu_check.py > gpurun_out/${tag}_mgpu_check_8gpu.log 2>&1; echo "rc=$?" >> gpurun_out/${tag}_mgpu_check_8gpu.log
timeout 900 $TR tests/mgpu_particles_check.py > gpurun_out/${tag}_mgpu_particles_8gpu.log 2>&1; echo "rc=$?" >> gpurun_out/${tag}_mgpu_particles_8gpu.log
export PCB200_TIME_GHOST=1
B="timeout 400 $TR bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu-baseline"
$B --no-e2e > gpurun_out/${tag}_bench_8gpu_zslabs.json 2> gpurun_out/${tag}_bench_8gpu_zslabs.err
$B --no-e2e --decomp 2,2,2 > gpurun_out/${tag}_bench_8gpu_2x2x2.json 2> gpurun_out/${tag}_bench_8gpu_2x2x2.err
$B --no-e2e --decomp 2,2,2 --entropy > gpurun_out/${tag}_bench_8gpu_2x2x2_entropy.json 2> gpurun_out/${tag}_bench_8gpu_2x2x2_entropy.err
$B --no-e2e --decomp 2,2,2 --ncell 256 --particles 12500000 > gpurun_out/${tag}_bench_8gpu_config5.json 2> gpurun_out/${tag}_bench_8gpu_config5.err
$B --no-e2e --scaling strong > gpurun_out/${tag}_bench_8gpu_strong_zslabs.json 2> gpurun_out/${tag}_bench_8gpu_strong_zslabs.err
$B --no-e2e --scaling strong --decomp 2,2,2 > gpurun_out/${tag}_bench_8gpu_strong_2x2x2.json 2> gpurun_out/${tag}_bench_8gpu_strong_2x2x2.err
grep -h "ghost-zone" gpurun_out/${tag}_bench_8gpu_*.err | sort | uniq -c | head -40 > gpurun_out/${tag}_ghost_timers.txt
tail -3 gpurun_out/${tag}_mgpu_check_8gpu.log gpurun_out/${tag}_mgpu_particles_8gpu.log
for f in gpurun_out/${tag}_bench_8gpu_*.json; do echo $f; cut -c1-220 $f; done
