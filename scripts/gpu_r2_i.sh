#!/bin/bash
# Round-2: GPU suite on the final config-2 training path, its timing, launch list and an ncu --set full capture of the
# chunked dense kernel (clustered) inside a config-2 training step.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/i_tests.log 2>&1; echo "tests rc=$?"; tail -n 6 gpurun_out/i_tests.log
timeout 300 python scripts/bench_train.py --cases c2,c3 --iters 20 2>&1 | tail -n 2 | tee gpurun_out/i_train.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/i_launches_c2.csv python scripts/bench_train.py --cases c2 --iters 2 > gpurun_out/i_ncu.log 2>&1; echo "ncu list rc=$?"
python scripts/launch_summary.py gpurun_out/i_launches_c2.csv > gpurun_out/i_launches_c2_summary.csv 2>&1; head -n 12 gpurun_out/i_launches_c2_summary.csv
timeout 900 ncu --set full --clock-control none --import-source on -k regex:dense_kernel --launch-skip 400 -c 8 -o gpurun_out/i_dense_c2 -f python scripts/bench_train.py --cases c2 --iters 2 > gpurun_out/i_ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i gpurun_out/i_dense_c2.ncu-rep --page raw --csv > gpurun_out/i_dense_c2_raw.csv 2>/dev/null; ls -la gpurun_out/i_dense_c2*
