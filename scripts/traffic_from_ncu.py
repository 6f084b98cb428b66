#!/usr/bin/env python
"""profiles/r2_traffic.json from `ncu --set full` raw CSV pages: per-launch DRAM bytes (read + write), executed warp
instructions and duration of the kernels bench.py quotes static counters for.
usage: traffic_from_ncu.py raw_hmc.csv [raw_gemm.csv ...] > profiles/r2_traffic.json"""
import csv
import json
import re
import sys

NAMES = {r"hmc_kernel<32, 32, (1|true)>": "hmc_kernel<32,32,true>", r"hmc_kernel<32, 32, (0|false)>": "hmc_kernel<32,32,false>",
         r"gemm_tc_kernel<256, 0>": "gemm_tc_kernel<256,0>", r"gemm_tc_kernel<128, 0>": "gemm_tc_kernel<128,0>",
         r"conv3d_tc_kernel<16": "conv3d_tc_kernel<16>"}


def num(s):
    return float(s.replace(",", "")) if s else 0.0


out = {"_source": "ncu --set full --clock-control none of this round's tree (scripts/gpu_r2_*.sh); per launch, last "
                  "captured launch of each kernel; files: " + ", ".join(sys.argv[1:])}
for path in sys.argv[1:]:
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}

    def val(r, key, want_unit=None):
        if key not in idx or r[idx[key]] == "":
            return None
        v, u = num(r[idx[key]]), units[idx[key]]
        scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "ns": 1e-3, "us": 1.0, "ms": 1e3}.get(u, 1.0)
        return v * scale
    for r in rows[2:]:
        name = r[idx["Kernel Name"]]
        for pat, key in NAMES.items():
            if re.search(pat, name):
                out[key] = {"dram_bytes": (val(r, "dram__bytes_read.sum") or 0) + (val(r, "dram__bytes_write.sum") or 0),
                            "inst_executed": val(r, "smsp__inst_executed.sum"),
                            "duration_us": val(r, "gpu__time_duration.sum"),
                            "grid": val(r, "launch__grid_size")}
print(json.dumps(out, indent=1))
