#!/bin/bash
# Round-2: chunked / clustered dense kernel -- full GPU suite, the small-op microbench, config-2/3 training and MC steps.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/f_tests.log 2>&1; echo "tests rc=$?"; tail -n 15 gpurun_out/f_tests.log
timeout 300 python scripts/bench_small_ops.py > gpurun_out/f_small_ops.log 2>&1; grep "ksplit=0\|net call\|floor" gpurun_out/f_small_ops.log
timeout 300 python scripts/bench_train.py --cases c2,c3 --iters 20 2>&1 | tail -n 2 | tee gpurun_out/f_train.log
