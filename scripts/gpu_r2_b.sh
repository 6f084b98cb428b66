#!/bin/bash
# Round-2 GPU check B: GPU tests, bench (both arms), config-3 conv path timing, ncu captures of the HMC kernels (for
# profiles/r2_traffic.json) and of the tcgen05 conv kernel.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_all.log 2>&1; echo rc=$? >> gpurun_out/gpu_all.log; tail -n 6 gpurun_out/gpu_all.log
for P in tc ffma; do
  L2HMC_CONV_PATH=$P timeout 200 python scripts/profile_c3.py --iters 5 > gpurun_out/c3_$P.json 2>&1; cat gpurun_out/c3_$P.json
done
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-l2hmc"
timeout 300 $CMD > gpurun_out/plain_hmc.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:hmc_kernel -s 3 -c 4 -o gpurun_out/prof_hmc_r2 -f $CMD > gpurun_out/ncu_full_hmc.log 2>&1
echo "ncu hmc rc=$?"
ncu -i gpurun_out/prof_hmc_r2.ncu-rep --page raw --csv > gpurun_out/raw_hmc_r2.csv 2>/dev/null
CMD4="python scripts/profile_c3.py --iters 2 --warm 1"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv3d_tc -s 20 -c 1 -o gpurun_out/prof_conv_r2 -f $CMD4 > gpurun_out/ncu_full_conv.log 2>&1; echo ncu conv rc=$?
ncu -i gpurun_out/prof_conv_r2.ncu-rep --page raw --csv > gpurun_out/raw_conv_r2.csv 2>/dev/null
rm -f gpurun_out/prof_hmc_r2.ncu-rep
( time timeout 900 python bench.py > gpurun_out/bench_r2.json 2> gpurun_out/bench_r2.err ) 2> gpurun_out/bench_time.txt; echo bench rc=$?
tail -n 3 gpurun_out/bench_r2.err; cat gpurun_out/bench_time.txt
