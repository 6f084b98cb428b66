#!/bin/bash
# Round-2: the bench line with programmatic dependent launch off / on.
set -u
mkdir -p gpurun_out
L2HMC_PDL=0 timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/h_bench_pdl0.json 2> gpurun_out/h_bench_pdl0.err; echo "rc=$?"
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/h_bench_pdl1.json 2> gpurun_out/h_bench_pdl1.err; echo "rc=$?"
python - <<'P'
import json
for f in ("gpurun_out/h_bench_pdl0.json", "gpurun_out/h_bench_pdl1.json"):
    for l in open(f):
        if l.startswith("{"):
            d = json.loads(l); L = d["l2hmc"]
            print(f, "value", d["value"], "c2 mc", L["config2_8x8_generic"]["mc_step_ms"], L["config2_8x8_generic"]["graph_replay"]["mc_step_ms"],
                  "c3 mc", L["config3_16x16_conv3D"]["mc_step_ms"], L["config3_16x16_conv3D"].get("graph_replay", {}).get("mc_step_ms"),
                  "train", L["training"]["config2_8x8_generic"]["train_step_ms"], L["training"]["config3_16x16_conv3D"]["train_step_ms"])
P
