#!/bin/bash
# Round-2: ncu --set full of the tcgen05 layer / heads kernels inside a config-3 MC step (final tree).
set -u
mkdir -p gpurun_out
python scripts/profile_c3.py --iters 2 > gpurun_out/tc_pre.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel --launch-skip 120 -c 8 -o gpurun_out/tc_c3 -f python scripts/profile_c3.py --iters 2 > gpurun_out/tc_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/tc_c3.ncu-rep --page raw --csv > gpurun_out/tc_c3_raw.csv 2>/dev/null; ls -la gpurun_out/tc_c3*
