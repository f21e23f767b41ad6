#!/bin/bash
# 8-GPU 2x2x2: ghost-zone phase timers with NCCL point-to-point channel settings; particle phases of config 5
tag=${1:-r02l}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29555"
export PCB200_TIME_GHOST=1
B="timeout 300 $TR bench.py --gpus 8 --steps 6 --warmup 2 --no-cpu-baseline --no-e2e --no-parity-check --decomp 2,2,2"
run() { name=$1; shift; "$@" > gpurun_out/${tag}_$name.json 2> gpurun_out/${tag}_$name.err; echo "== $name: $(grep -h 'rank 0\] ghost-zone' gpurun_out/${tag}_$name.err | head -1 | cut -c1-600)"; python -c "
import json,sys
for l in open('gpurun_out/${tag}_$name.json'):
    if l.startswith('{'):
        d=json.loads(l); print('   ms/step %.2f kernel %.3f' % (d['ms_per_step'], d['roofline']['kernel_ms_per_substep']))
"; }
run default $B
NCCL_NCHANNELS_PER_PEER=8 run cpp8 $B
NCCL_NCHANNELS_PER_PEER=16 run cpp16 $B
run config5 $B --ncell 256 --particles 12500000
