#!/bin/bash
# Round-2: ncu --set full of the fused HMC kernels on the final tree (profiles/r2_traffic.json, r2_hmc_ncu_full.txt).
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-l2hmc"
timeout 300 $CMD > gpurun_out/plain_hmc.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:hmc_kernel -s 3 -c 4 -o gpurun_out/prof_hmc_r2 -f $CMD > gpurun_out/ncu_full_hmc.log 2>&1
echo "ncu hmc rc=$?"
ncu -i gpurun_out/prof_hmc_r2.ncu-rep --page raw --csv > gpurun_out/raw_hmc_r2.csv 2>/dev/null; ls -la gpurun_out/raw_hmc_r2.csv
