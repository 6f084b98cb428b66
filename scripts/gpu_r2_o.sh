#!/bin/bash
# Round-2 experiment: K-chunk rule of the dense kernel (slabs per chunk minimum, maximum chunks per tile).
set -u
mkdir -p gpurun_out
for v in "L2HMC_DENSE_SP_MIN=2 L2HMC_DENSE_MAX_CHUNKS=8" "L2HMC_DENSE_SP_MIN=4 L2HMC_DENSE_MAX_CHUNKS=8" "L2HMC_DENSE_SP_MIN=1 L2HMC_DENSE_MAX_CHUNKS=8" "L2HMC_DENSE_SP_MIN=2 L2HMC_DENSE_MAX_CHUNKS=4" "L2HMC_DENSE_SP_MIN=3 L2HMC_DENSE_MAX_CHUNKS=6"; do
  echo "== $v"
  env $v timeout 300 python scripts/bench_small_ops.py 2>&1 | grep "ksplit=0\|path ffma" | grep -v "dW" | cut -c1-140
  env $v timeout 300 python scripts/bench_train.py --cases c2 --iters 30 2>&1 | tail -n 1 | grep -o "train_step_ms\": [0-9.]*"
  env $v python scripts/profile_c3.py --arch generic --size 8 --chains 128 --iters 100 --warm 10 2>&1 | tail -n 1 | grep -o "mc_step_ms\": [0-9.]*"
done | tee gpurun_out/o_chunks.log
