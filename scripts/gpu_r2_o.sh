#!/bin/bash
# Round-2: 16-byte operand copies + wider K split for the small weight-gradient products -- training tests and step times.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_training.py tests/test_gpu_train_rows.py tests/test_gpu_config4_shapes.py -x -q > gpurun_out/o_tests.log 2>&1; echo "tests rc=$?"; tail -n 4 gpurun_out/o_tests.log
for i in 1 2; do timeout 300 python scripts/bench_train.py --cases c2,c3 --iters 30 2>&1 | tail -n 2 | grep -o "train_step_ms\": [0-9.]*"; done
