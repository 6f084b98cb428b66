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
CMD3="python scripts/profile_k1k2.py --iters 1 --warm 1"
timeout 300 $CMD3 > gpurun_out/plain3_$TAG.log 2>&1 && \
timeout 900 ncu --set full --clock-control none -k regex:"stencil_|update_|mh_select" -c 14 -o gpurun_out/prof_k1k2_$TAG -f $CMD3 > gpurun_out/ncu_full_k1k2_$TAG.log 2>&1
echo "ncu k1k2 rc=$?"
CMD4="python scripts/profile_c3.py --iters 2 --warm 1"
timeout 300 $CMD4 > gpurun_out/c3_$TAG.json 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches_c3_$TAG.csv $CMD4 > gpurun_out/ncu_list_c3_$TAG.log 2>&1
echo "ncu c3 list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv3d_tiled -s 20 -c 1 -o gpurun_out/prof_conv_$TAG -f $CMD4 > gpurun_out/ncu_full_conv_$TAG.log 2>&1
echo "ncu conv rc=$?"
CMD5="python scripts/profile_c3.py --arch generic --size 8 --chains 8192 --steps 10 --eps 0.2 --iters 2 --warm 1"
timeout 300 $CMD5 > gpurun_out/c2_$TAG.json 2>&1; cat gpurun_out/c3_$TAG.json gpurun_out/c2_$TAG.json
# text summaries on the box; the big reports stay behind (gpurun_out/ must stay under 64 MiB)
for r in hmc gemm_c3 gemm_c4 k1k2 conv; do
  f=gpurun_out/prof_${r}_$TAG.ncu-rep
  [ -f $f ] || continue
  ncu -i $f --page raw --csv > gpurun_out/raw_${r}_$TAG.csv 2>/dev/null
  ncu -i $f --page details > gpurun_out/details_${r}_$TAG.txt 2>/dev/null
done
rm -f gpurun_out/prof_k1k2_$TAG.ncu-rep gpurun_out/prof_gemm_c4_$TAG.ncu-rep gpurun_out/prof_hmc_$TAG.ncu-rep
du -sh gpurun_out; ls -la gpurun_out | tail -40
