#!/bin/bash
# Run on the GPU box (via gpurun): tests, bench, ncu launch list and full capture of the two HMC kernels.
set -u
mkdir -p gpurun_out
TAG=${1:-r1}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_$TAG.log
timeout 400 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; cat gpurun_out/bench_$TAG.json
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $CMD > gpurun_out/plain_$TAG.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1
echo "ncu list rc=$?"
timeout 300 $CMD > gpurun_out/plain2_$TAG.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:hmc_kernel -s 3 -c 4 -o gpurun_out/prof_hmc_$TAG -f $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"; tail -3 gpurun_out/ncu_full_$TAG.log
ls -la gpurun_out
