#!/bin/bash
# Round-2: tcgen05 heads epilogue with hoisted per-output constants -- parity tests and MC-step timings.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_l2hmc.py tests/test_gpu_config4_shapes.py tests/test_gpu_sampler.py tests/test_gpu_kernels.py -x -q > gpurun_out/o_tests.log 2>&1; echo "tests rc=$?"; tail -n 4 gpurun_out/o_tests.log
for i in 1 2; do
  python scripts/profile_c3.py --iters 30 --warm 5 2>&1 | tail -n 1
done | tee gpurun_out/o_c3.log
python scripts/profile_c3.py --arch generic --size 32 --chains 2048 --steps 10 --eps 0.1 --iters 5 --warm 2 2>&1 | tail -n 1
