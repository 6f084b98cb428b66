#!/usr/bin/env python
"""Time the S/T/Q network call (K3) on one GPU for the BASELINE shapes, per path, and the L2HMC MC
step built on it.  Prints one JSON object per case.  Developer tool (bench.py carries the official
line); usage: python scripts/bench_nets.py [--cases c2,c3,c4s] [--iters 20]"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from l2hmc_b200 import GaugeDynamics, GaugeLattice, _lib  # noqa: E402

CASES = {
    # name: (arch, T, X, chains, M)
    "c2": ("generic", 8, 8, 128, 5),
    "c3": ("conv3D", 16, 16, 2048, 5),
    "c3g": ("generic", 16, 16, 2048, 5),
    "c4s": ("generic", 32, 32, 2048, 2),     # config-4 net shapes (H = 8192) on a slice of the chains
    "c4c": ("conv3D", 32, 32, 2048, 2),
}


def time_ms(fn, iters, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(iters):
        fn()
    ev[1].record()
    torch.cuda.synchronize()
    return ev[0].elapsed_time(ev[1]) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", default="c2,c3,c3g,c4s")
    ap.add_argument("--iters", type=int, default=10)
    args = ap.parse_args()
    for name in args.cases.split(","):
        arch, T, X, N, M = CASES[name]
        D = 2 * T * X
        lat = GaugeLattice(T, X, 2, "U1", num_samples=1)
        np.random.seed(42)
        dyn = GaugeDynamics(lat, hmc=False, network_arch=arch, num_steps=M, eps=0.1, seed=1)
        net = dyn.position_fn
        K, H = net.num_features, net.num_hidden
        flop = 2.0 * N * (2 * K * H + 2 * H + H * H + 3 * H * D)
        gen = torch.Generator(device="cuda").manual_seed(0)
        x = torch.rand(N, D, device="cuda", generator=gen) * 6.28
        v = torch.randn(N, D, device="cuda", generator=gen)
        out = {"case": name, "arch": arch, "lattice": f"{T}x{X}", "chains": N, "K": K, "H": H, "D": D,
               "gflop_per_call": flop / 1e9}
        ld = torch.zeros(N, device="cuda")
        for pname, path in (("tcgen05", _lib.NET_PATH_TCGEN05), ("ffma", _lib.NET_PATH_FFMA)):
            dyn.position_fn.path = dyn.momentum_fn.path = path
            try:
                ms = time_ms(lambda: net.position_update(x, v, (1.0, 0.0), dyn.mask[0], False, 0.1, 1, ld, dyn._flag),
                             args.iters)
            except _lib.L2hmcError as e:
                out[pname] = str(e)
                continue
            out[pname] = {"ms_per_call": ms, "tflops_fp32_equiv": flop / ms / 1e9}
            ms_mc = time_ms(lambda: dyn.apply_transition(x, 2.0), max(2, args.iters // 4), warm=1)
            out[pname]["mc_step_ms"] = ms_mc
            out[pname]["link_updates_per_s"] = N * D * M / (ms_mc * 1e-3)
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
