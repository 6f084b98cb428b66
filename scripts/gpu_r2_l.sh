#!/bin/bash
# Round-2: chained tcgen05 backward products off / on, same box.
set -u
mkdir -p gpurun_out
for v in "L2HMC_TC_CHAIN=0" "L2HMC_TC_CHAIN=1" "L2HMC_TC_CHAIN=0" "L2HMC_TC_CHAIN=1"; do
  echo "== $v"; env $v timeout 300 python scripts/bench_train.py --cases c3 --iters 20 2>&1 | tail -n 1
done | tee gpurun_out/l_chain.log
