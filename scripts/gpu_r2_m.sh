#!/bin/bash
# Round-2 evidence pass on the final tree: parity table, smoke(), launch lists of the config-3 MC step and training step.
set -u
mkdir -p gpurun_out
timeout 900 python scripts/parity_errors.py gpurun_out/m_parity_errors.txt > gpurun_out/m_parity.log 2>&1; echo "parity rc=$?"; tail -n 3 gpurun_out/m_parity.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/m_smoke.log 2>&1; tail -n 2 gpurun_out/m_smoke.log
python scripts/profile_c3.py > gpurun_out/m_pc3.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/m_launches_c3.csv python scripts/profile_c3.py > gpurun_out/m_ncu_c3.log 2>&1; echo "ncu c3 rc=$?"
python scripts/launch_summary.py gpurun_out/m_launches_c3.csv > gpurun_out/m_launches_c3_summary.csv 2>&1; head -n 8 gpurun_out/m_launches_c3_summary.csv
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/m_launches_train_c3.csv python scripts/bench_train.py --cases c3 --iters 2 > gpurun_out/m_ncu_t3.log 2>&1; echo "ncu t3 rc=$?"
python scripts/launch_summary.py gpurun_out/m_launches_train_c3.csv > gpurun_out/m_launches_train_c3_summary.csv 2>&1; head -n 10 gpurun_out/m_launches_train_c3_summary.csv
