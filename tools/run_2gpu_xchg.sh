#!/bin/bash
# 2-GPU timing of the ghost-zone phases per axis and NCCL point-to-point settings (PCB200_TIME_GHOST=1)
tag=${1:-r02k}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544"
export PCB200_TIME_GHOST=1
B="timeout 300 $TR bench.py --gpus 2 --steps 5 --warmup 2 --no-cpu-baseline --no-e2e --no-parity-check"
run() { name=$1; shift; "$@" > gpurun_out/${tag}_$name.json 2> gpurun_out/${tag}_$name.err; echo "== $name: $(grep -h 'ghost-zone' gpurun_out/${tag}_$name.err | head -1 | cut -c1-300)"; python -c "
import json,sys
for l in open('gpurun_out/${tag}_$name.json'):
    if l.startswith('{'):
        d=json.loads(l); print('   ms/step %.2f kernel %.3f' % (d['ms_per_step'], d['roofline']['kernel_ms_per_substep']))
"; }
run x_default $B --decomp 2,1,1
run y_default $B --decomp 1,2,1
run z_default $B --decomp 1,1,2
NCCL_NCHANNELS_PER_PEER=8 run x_cpp8 $B --decomp 2,1,1
NCCL_NCHANNELS_PER_PEER=32 NCCL_MIN_P2P_NCHANNELS=32 NCCL_MAX_P2P_NCHANNELS=32 run x_cpp32 $B --decomp 2,1,1
NCCL_P2P_USE_CUDA_MEMCPY=1 run x_ce $B --decomp 2,1,1
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,P2P $B --decomp 2,1,1 --steps 1 --warmup 1 2>&1 | grep -E "Channel|P2P|NVLS|channels|nChannels" | head -30 > gpurun_out/${tag}_nccl_info.txt
