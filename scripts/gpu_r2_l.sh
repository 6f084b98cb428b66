#!/bin/bash
# Round-2: training tests and config-3 step timing (gate bits fetched under the MMAs).
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_training.py tests/test_gpu_config4_shapes.py tests/test_gpu_train_rows.py -x -q > gpurun_out/l_tests.log 2>&1; echo "tests rc=$?"; tail -n 5 gpurun_out/l_tests.log
for i in 1 2; do timeout 300 python scripts/bench_train.py --cases c3 --iters 20 2>&1 | tail -n 1; done | tee gpurun_out/l_train.log
