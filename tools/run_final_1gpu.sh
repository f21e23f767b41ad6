#!/bin/bash
# final 1-GPU records of the round, most important first (the GPU budget may cut the tail)
tag=${1:-r02r}
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log; tail -n 3 gpurun_out/${tag}_pytest.log
timeout 200 python bench.py > gpurun_out/${tag}_bench_1gpu_default.json 2> gpurun_out/${tag}_bench.err
Q="--no-cpu-baseline --no-e2e --no-parity-check --steps 10"
PCB200_LIB=$PWD/pencil-code_b200/libpcb200_wp0.so timeout 120 python bench.py $Q > gpurun_out/${tag}_bench_wp0.json 2>> gpurun_out/${tag}_bench.err
timeout 120 python bench.py $Q > gpurun_out/${tag}_bench_wp32.json 2>> gpurun_out/${tag}_bench.err
timeout 120 python bench.py --no-cpu-baseline --hydro --ncell 256 --forcing > gpurun_out/${tag}_bench_config2_hydro256_forced.json 2>> gpurun_out/${tag}_bench.err
timeout 120 python bench.py --no-cpu-baseline --convslab --ncell 32 --steps 200 > gpurun_out/${tag}_bench_config1_convslab32.json 2>> gpurun_out/${tag}_bench.err
timeout 120 python bench.py --no-cpu-baseline --convslab --ncell 128 --steps 50 > gpurun_out/${tag}_bench_config1_convslab128.json 2>> gpurun_out/${tag}_bench.err
timeout 120 python bench.py --no-cpu-baseline --forcing --steps 10 > gpurun_out/${tag}_bench_config3_forced.json 2>> gpurun_out/${tag}_bench.err
timeout 150 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-parity-check > gpurun_out/${tag}_ncu_list.log 2>&1
timeout 200 ncu --set full --clock-control none --import-source on -k regex:rhs_dly -s 9 -c 3 -o gpurun_out/${tag}_rhs_dly python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-parity-check > gpurun_out/${tag}_ncu_full.log 2>&1
tail -n 2 gpurun_out/${tag}_bench.err
