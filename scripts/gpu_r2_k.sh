#!/bin/bash
# Round-2: per-launch durations and grids of the tcgen05 product kernel and its operand converter inside a config-3
# training step.
set -u
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum,launch__grid_size --clock-control none -k regex:"gemm_tc_kernel|ts32_from_strided" --launch-skip 600 -c 500 --csv --log-file gpurun_out/k_tc_launches_c3.csv python scripts/bench_train.py --cases c3 --iters 1 > gpurun_out/k_ncu.log 2>&1; echo "rc=$?"
tail -n 3 gpurun_out/k_ncu.log
