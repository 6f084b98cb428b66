#!/bin/bash
# Round-2: full GPU suite + bench line (log-det reduces deferred to the end of the trajectory).
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/n_tests.log 2>&1; echo "tests rc=$?"; tail -n 6 gpurun_out/n_tests.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/n_bench.json 2> gpurun_out/n_bench.err; echo "rc=$?"
python - <<'P'
import json
for f in ("gpurun_out/n_bench.json",):
    for l in open(f):
        if l.startswith("{"):
            d = json.loads(l); L = d["l2hmc"]
            c3 = L["config3_16x16_conv3D"]; c4 = L["config4_32x32_generic_nets"]
            print(f, "value", d["value"], "e2e", d["e2e"]["value"], "c2 mc", L["config2_8x8_generic"]["mc_step_ms"], L["config2_8x8_generic"]["graph_replay"]["mc_step_ms"],
                  "c3 mc", c3["mc_step_ms"], c3.get("graph_replay", {}).get("mc_step_ms"), "c3 call", c3["net_call"]["ms"], c3["roofline"]["frac"],
                  "c4 mc", c4["mc_step_ms"], "train", L["training"]["config2_8x8_generic"]["train_step_ms"], L["training"]["config3_16x16_conv3D"]["train_step_ms"])
P
