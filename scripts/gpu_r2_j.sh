#!/bin/bash
# Round-2: programmatic dependent launch across the tcgen05 / conv / stencil / training kernels -- GPU suite and the
# bench line with it off / on.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/j_tests.log 2>&1; echo "tests rc=$?"; tail -n 6 gpurun_out/j_tests.log
L2HMC_PDL=0 timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/j_bench_pdl0.json 2> gpurun_out/j_bench_pdl0.err; echo "rc=$?"
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/j_bench_pdl1.json 2> gpurun_out/j_bench_pdl1.err; echo "rc=$?"
python - <<'P'
import json
for f in ("gpurun_out/j_bench_pdl0.json", "gpurun_out/j_bench_pdl1.json"):
    for l in open(f):
        if l.startswith("{"):
            d = json.loads(l); L = d["l2hmc"]
            c3 = L["config3_16x16_conv3D"]; c4 = L["config4_32x32_generic_nets"]
            print(f, "value", d["value"], "c2 mc", L["config2_8x8_generic"]["mc_step_ms"], L["config2_8x8_generic"]["graph_replay"]["mc_step_ms"],
                  "c3 mc", c3["mc_step_ms"], c3.get("graph_replay", {}).get("mc_step_ms"), "c3 call", c3["net_call"]["ms"], c3["roofline"]["frac"],
                  "c4 call", c4["net_call"]["ms"], c4["roofline"]["frac"], "c4 mc", c4["mc_step_ms"],
                  "train", L["training"]["config2_8x8_generic"]["train_step_ms"], L["training"]["config3_16x16_conv3D"]["train_step_ms"])
P
