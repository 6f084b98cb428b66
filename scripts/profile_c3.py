#!/usr/bin/env python
"""The L2HMC MC step alone (default: BASELINE config 3: 16x16, conv3D nets, 2048 chains, M = 5) for
ncu launch lists:  python scripts/profile_c3.py [--arch conv3D|generic] [--size 16] [--chains 2048]
[--steps 5] [--iters 3].  Prints the CUDA-event time per MC step."""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from l2hmc_b200 import GaugeDynamics, GaugeLattice  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--arch", default="conv3D")
    ap.add_argument("--size", type=int, default=16)
    ap.add_argument("--chains", type=int, default=2048)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--eps", type=float, default=0.25)
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--warm", type=int, default=2)
    a = ap.parse_args()
    T = a.size
    lat = GaugeLattice(T, T, 2, "U1", num_samples=1)
    np.random.seed(42)
    dyn = GaugeDynamics(lat, hmc=False, network_arch=a.arch, num_steps=a.steps, eps=a.eps, seed=7)
    gen = torch.Generator(device="cuda").manual_seed(5)
    x = torch.rand(a.chains, 2 * T * T, device="cuda", generator=gen) * (2 * np.pi)
    for _ in range(a.warm):
        x = dyn.apply_transition(x, 3.0, wrap=True)[3]
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(a.iters):
        x = dyn.apply_transition(x, 3.0, wrap=True)[3]
    ev[1].record()
    torch.cuda.synchronize()
    dyn.raise_if_nonfinite()
    ms = ev[0].elapsed_time(ev[1]) / a.iters
    print(json.dumps({"arch": a.arch, "lattice": f"{T}x{T}", "chains": a.chains, "lf_steps": a.steps,
                      "mc_step_ms": ms, "link_updates_per_s": a.chains * 2 * T * T * a.steps / (ms * 1e-3)}))


if __name__ == "__main__":
    main()
