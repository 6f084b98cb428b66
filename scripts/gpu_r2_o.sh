#!/bin/bash
# Round-2: deferred log-det reduce off / on, same box (eager MC steps of configs 3 and 2).
set -u
mkdir -p gpurun_out
for v in "L2HMC_DEFER_LOGDET=0" "L2HMC_DEFER_LOGDET=1" "L2HMC_DEFER_LOGDET=0" "L2HMC_DEFER_LOGDET=1"; do
  echo "== $v"
  env $v python scripts/profile_c3.py --iters 30 --warm 5 2>&1 | tail -n 1
  env $v python scripts/profile_c3.py --arch generic --size 8 --chains 128 --iters 100 --warm 10 2>&1 | tail -n 1
done | tee gpurun_out/o_defer.log
