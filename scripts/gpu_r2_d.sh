#!/bin/bash
# Round-2 GPU check D: training tests first (conv adjoint from the routing tape), whole GPU suite, training-step timings
# and the launch list of the config-3 training step.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_training.py -m gpu -x -q > gpurun_out/gpu_train.log 2>&1; echo rc=$? >> gpurun_out/gpu_train.log; tail -n 15 gpurun_out/gpu_train.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_all.log 2>&1; echo rc=$? >> gpurun_out/gpu_all.log; tail -n 6 gpurun_out/gpu_all.log
timeout 300 python scripts/bench_train.py --cases c2,c3 --iters 10 > gpurun_out/train_step_r2.jsonl 2>&1; cat gpurun_out/train_step_r2.jsonl
CMD="python scripts/bench_train.py --cases c3 --iters 2"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/launches_train_c3.csv $CMD > gpurun_out/ncu_list_train_c3.log 2>&1; echo ncu rc=$?
python scripts/launch_summary.py gpurun_out/launches_train_c3.csv > gpurun_out/r2_launch_list_train_c3_summary.csv; head -14 gpurun_out/r2_launch_list_train_c3_summary.csv
rm -f gpurun_out/launches_train_c3.csv
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv3d_bwd_tape -s 10 -c 1 -o gpurun_out/prof_convbwd_r2 -f $CMD > gpurun_out/ncu_full_convbwd.log 2>&1; echo ncu convbwd rc=$?
ncu -i gpurun_out/prof_convbwd_r2.ncu-rep --page raw --csv > gpurun_out/raw_convbwd_r2.csv 2>/dev/null
