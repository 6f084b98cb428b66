#!/bin/bash
# Run on the GPU box (via gpurun): tests, bench (both arms), K1/K2 stand-alone numbers, ncu launch list of the
# bench command, and --set full captures of the HMC kernels, the tcgen05 GEMM and the K1/K2 kernels.
set -u
mkdir -p gpurun_out
TAG=${1:-r1}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_$TAG.log
timeout 400 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; cat gpurun_out/bench_$TAG.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_$TAG.json 2>&1; echo "ref rc=$?"
timeout 300 python scripts/profile_k1k2.py > gpurun_out/k1k2_$TAG.jsonl 2>&1; echo "k1k2 rc=$?"; cat gpurun_out/k1k2_$TAG.jsonl
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $CMD > gpurun_out/plain_$TAG.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1
echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:hmc_kernel -s 3 -c 4 -o gpurun_out/prof_hmc_$TAG -f $CMD > gpurun_out/ncu_full_hmc_$TAG.log 2>&1
echo "ncu hmc rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 6 -c 3 -o gpurun_out/prof_gemm_c3_$TAG -f $CMD > gpurun_out/ncu_full_gemm_c3_$TAG.log 2>&1
echo "ncu gemm c3 rc=$?"
CMD2="python scripts/bench_nets.py --cases c4s --iters 2"
timeout 300 $CMD2 > gpurun_out/plain2_$TAG.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 9 -c 3 -o gpurun_out/prof_gemm_c4_$TAG -f $CMD2 > gpurun_out/ncu_full_gemm_c4_$TAG.log 2>&1
echo "ncu gemm c4 rc=$?"
CMD3="python scripts/profile_k1k2.py --iters 2"
timeout 300 $CMD3 > gpurun_out/plain3_$TAG.log 2>&1 && \
timeout 900 ncu --set full --clock-control none -k regex:"stencil_kernel|update_|mh_select" -s 21 -c 14 -o gpurun_out/prof_k1k2_$TAG -f $CMD3 > gpurun_out/ncu_full_k1k2_$TAG.log 2>&1
echo "ncu k1k2 rc=$?"
ls -la gpurun_out | tail -30
