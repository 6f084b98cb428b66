#!/bin/bash
# Round-2 GPU check A (run through gpurun): GPU tests, bench, config-3 MC step with conv-2 on tcgen05 vs the FFMA conv
# kernels, launch list of the config-3 step and an ncu --set full capture of the tcgen05 conv kernel.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_all.log 2>&1; echo rc=$? >> gpurun_out/gpu_all.log; tail -n 6 gpurun_out/gpu_all.log
( time timeout 900 python bench.py > gpurun_out/bench_r2.json 2> gpurun_out/bench_r2.err ) 2> gpurun_out/bench_time.txt; echo bench rc=$?
tail -n 3 gpurun_out/bench_r2.err; cat gpurun_out/bench_time.txt
for P in tc ffma; do
  L2HMC_CONV_PATH=$P timeout 200 python scripts/profile_c3.py --iters 5 > gpurun_out/c3_$P.json 2>&1; cat gpurun_out/c3_$P.json
done
CMD4="python scripts/profile_c3.py --iters 2 --warm 1"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches_c3_r2.csv $CMD4 > gpurun_out/ncu_list_c3.log 2>&1; echo ncu rc=$?
python scripts/launch_summary.py gpurun_out/launches_c3_r2.csv > gpurun_out/r2_launch_list_c3_summary.csv; head -12 gpurun_out/r2_launch_list_c3_summary.csv
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv3d_tc -s 20 -c 1 -o gpurun_out/prof_conv_r2 -f $CMD4 > gpurun_out/ncu_full_conv.log 2>&1; echo ncu conv rc=$?
ncu -i gpurun_out/prof_conv_r2.ncu-rep --page raw --csv > gpurun_out/raw_conv_r2.csv 2>/dev/null
rm -f gpurun_out/launches_c3_r2.csv
