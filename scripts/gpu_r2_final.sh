#!/bin/bash
# Round-2 final pass on one GPU: full GPU suite, smoke(), both bench arms, parity table.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/final_tests.log 2>&1; echo "tests rc=$?"; tail -n 5 gpurun_out/final_tests.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/final_smoke.log 2>&1; tail -n 2 gpurun_out/final_smoke.log
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_bench_reference.json 2> gpurun_out/final_bench_reference.err; echo "ref rc=$?"
timeout 900 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench rc=$?"
timeout 900 python scripts/parity_errors.py gpurun_out/final_parity_errors.txt > gpurun_out/final_parity.log 2>&1; echo "parity rc=$?"
python - <<'P'
import json
for l in open("gpurun_out/final_bench.json"):
    if l.startswith("{"):
        d = json.loads(l); L = d["l2hmc"]
        c3 = L["config3_16x16_conv3D"]; c4 = L["config4_32x32_generic_nets"]
        print("value", d["value"], "e2e", d["e2e"]["value"], "frac", d["roofline"]["frac"], "launches", d["gpu_launches"], "c2 mc", L["config2_8x8_generic"]["mc_step_ms"], L["config2_8x8_generic"]["graph_replay"]["mc_step_ms"],
              "c3 mc", c3["mc_step_ms"], c3.get("graph_replay", {}).get("mc_step_ms"), "c3 call", c3["net_call"]["ms"], c3["roofline"]["frac"],
              "c4 mc", c4["mc_step_ms"], c4["roofline"]["frac"], "train", L["training"]["config2_8x8_generic"]["train_step_ms"], L["training"]["config3_16x16_conv3D"]["train_step_ms"])
for l in open("gpurun_out/final_bench_reference.json"):
    if l.startswith("{"):
        d = json.loads(l); print("reference arm", d["value"], d["ms_per_step"], d["cpu_baseline"]["cores"])
P
