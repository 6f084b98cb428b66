#!/bin/bash
# Round-2: programmatic dependent launch on the dense kernel -- with / without (L2HMC_PDL=0).
set -u
mkdir -p gpurun_out
for v in "L2HMC_PDL=0" "L2HMC_PDL=1"; do
  echo "== $v"
  env $v timeout 300 python scripts/bench_small_ops.py 2>&1 | grep "ksplit=0\|net call\|floor\|rows=256\|rows=1024 "
  env $v timeout 300 python scripts/bench_train.py --cases c2,c3 --iters 20 2>&1 | tail -n 2
done 2>&1 | tee gpurun_out/g_pdl.log
