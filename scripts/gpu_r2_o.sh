#!/bin/bash
# Round-2: resident run reuses the previous step's charges as the charge of its input -- sampler tests and the bench line.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_sampler.py tests/test_gpu_multirank.py -x -q > gpurun_out/o_tests.log 2>&1; echo "tests rc=$?"; tail -n 4 gpurun_out/o_tests.log
timeout 900 python bench.py --steps 20 --warmup 3 --no-l2hmc > gpurun_out/o_bench.json 2> gpurun_out/o_bench.err; echo "rc=$?"
python - <<'P'
import json
for l in open("gpurun_out/o_bench.json"):
    if l.startswith("{"):
        d = json.loads(l)
        print("value", d["value"], "ms", d["ms_per_step"], "e2e", d["e2e"]["value"], "resident", d["e2e"]["resident_run"]["value"], d["e2e"]["resident_run"]["ms_per_step"], d["e2e"]["resident_run"]["frac_of_device_resident_value"])
P
