#!/bin/bash
# Round-2: packed two-plaquette cos / projection in the action and observables kernels -- parity and HBM fraction.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_hmc.py tests/test_gpu_sampler.py tests/test_gpu_l2hmc.py -x -q > gpurun_out/o_tests.log 2>&1; echo "tests rc=$?"; tail -n 4 gpurun_out/o_tests.log
timeout 300 python scripts/profile_k1k2.py 2>&1 | tee gpurun_out/o_k1k2.jsonl | grep -i "action\|observ\|hamil\|force" | cut -c1-300
