#!/bin/bash
# 4 GPUs: the per-axis communicators (ncclCommSplit) against the single-domain oracle, all decompositions of 4 ranks,
# and the 2x2x1 bench with and without them (PCB200_TIME_GHOST=1)
tag=${1:-r02o}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29566"
timeout 600 $TR tests/mgpu_check.py > gpurun_out/${tag}_mgpu_check_4gpu.log 2>&1; echo "rc=$?" >> gpurun_out/${tag}_mgpu_check_4gpu.log
timeout 600 $TR tests/mgpu_particles_check.py > gpurun_out/${tag}_mgpu_particles_4gpu.log 2>&1; echo "rc=$?" >> gpurun_out/${tag}_mgpu_particles_4gpu.log
tail -n 3 gpurun_out/${tag}_mgpu_check_4gpu.log; tail -n 3 gpurun_out/${tag}_mgpu_particles_4gpu.log
export PCB200_TIME_GHOST=1
B="timeout 300 $TR bench.py --gpus 4 --steps 6 --warmup 2 --no-cpu-baseline --no-e2e --decomp 2,2,1"
run() { name=$1; shift; "$@" > gpurun_out/${tag}_$name.json 2> gpurun_out/${tag}_$name.err; echo "== $name: $(grep -h 'rank 0\] ghost-zone' gpurun_out/${tag}_$name.err | head -1 | cut -c1-300)"; python -c "
import json,sys
for l in open('gpurun_out/${tag}_$name.json'):
    if l.startswith('{'):
        d=json.loads(l); print('   ms/step %.2f kernel %.3f parity %s' % (d['ms_per_step'], d['roofline']['kernel_ms_per_substep'], d['parity_check']))
"; }
run split $B
PCB200_COMM_SPLIT=0 run nosplit $B
