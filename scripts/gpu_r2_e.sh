#!/bin/bash
# Round-2 experiment: config-2 training step with the tcgen05 thresholds lowered (diagnostic env knobs), its launch list,
# and the new conv3D gradient parity cases.
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_training.py -x -q -k "oracle_config2_shapes" > gpurun_out/e_tests.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/e_tests.log
for v in "" "L2HMC_TC_MIN_ROWS=256" "L2HMC_TC_MIN_MNK=6e7" "L2HMC_TC_MIN_ROWS=256 L2HMC_TC_MIN_MNK=6e7" "L2HMC_TC_MIN_MNK=3e7"; do
  echo "== $v"; env $v timeout 300 python scripts/bench_train.py --cases c2 --iters 20 2>&1 | tail -n 1
done | tee gpurun_out/e_c2_variants.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/e_launches_c2.csv python scripts/bench_train.py --cases c2 --iters 2 > gpurun_out/e_ncu.log 2>&1; echo "ncu rc=$?"
python scripts/launch_summary.py gpurun_out/e_launches_c2.csv > gpurun_out/e_launches_c2_summary.csv 2>&1; head -n 25 gpurun_out/e_launches_c2_summary.csv
