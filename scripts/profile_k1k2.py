#!/usr/bin/env python
"""Stand-alone K1 / K2 / K4 kernels at the config-4 size (32x32, 65 536 chains): CUDA-event time and
achieved HBM GB/s against their algorithmic bytes.  Also the command ncu wraps for the
per-kernel dram counters in profiles/.  usage: python scripts/profile_k1k2.py [--chains N] [--iters K]"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from l2hmc_b200 import GaugeLattice, _lib  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--chains", type=int, default=65536)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--warm", type=int, default=3)
    args = ap.parse_args()
    T = X = 32
    N, D = args.chains, 2 * T * X
    lib = _lib.get_lib()
    lat = GaugeLattice(T, X, 2, "U1", num_samples=1)
    gen = torch.Generator(device="cuda").manual_seed(0)
    x = torch.rand(N, D, device="cuda", generator=gen) * (2 * np.pi)
    v = torch.randn(N, D, device="cuda", generator=gen)
    S, Tr, Q = (torch.randn(N, D, device="cuda", generator=gen) * 0.1 for _ in range(3))
    g = torch.empty_like(x)
    out = torch.empty_like(x)
    act = torch.empty(N, device="cuda")
    plaq, qf = torch.empty_like(act), torch.empty_like(act)
    qi = torch.empty(N, dtype=torch.int32, device="cuda")
    ld = torch.zeros(N, device="cuda")
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    mask = (torch.rand(D, device="cuda", generator=gen) > 0.5).float()
    p = torch.rand(N, device="cuda", generator=gen)
    u = torch.rand(N, device="cuda", generator=gen)
    acc = torch.empty(N, dtype=torch.uint8, device="cuda")
    P = _lib.dptr
    nd = N * D
    cases = [
        ("stencil force (K1)", 8 * nd,
         lambda: lib.l2hmc_u1_action_force(P(x), 4.0, N, T, X, None, P(g), None)),
        ("stencil action (K1)", 4 * nd,
         lambda: lib.l2hmc_u1_action_force(P(x), 4.0, N, T, X, P(act), None, None)),
        ("observables (K4)", 4 * nd,
         lambda: lib.l2hmc_u1_observables(P(x), N, T, X, P(act), P(plaq), P(qf), P(qi, torch.int32), None)),
        ("hamiltonian", 8 * nd,
         lambda: lib.l2hmc_u1_hamiltonian(P(x), P(v), 4.0, N, T, X, P(act), None)),
        ("update_momentum with S,T,Q (K2)", 24 * nd,
         lambda: lib.l2hmc_update_momentum(P(v), P(g), P(S), P(Tr), P(Q), 0.1, 1, N, D, P(out), P(ld),
                                           P(flag, torch.int32), None)),
        ("update_position with S,T,Q (K2)", 24 * nd,
         lambda: lib.l2hmc_update_position(P(x), P(v), P(S), P(Tr), P(Q), P(mask), 0, 0.1, 1, N, D, P(out), P(ld),
                                           P(flag, torch.int32), None)),
        ("mh_select + wrap (K4)", 12 * nd,
         lambda: lib.l2hmc_mh_select(P(x), P(g), None, None, None, P(p), None, None, P(u), N, D, 1, None, None, None,
                                     P(out), P(acc, torch.uint8), None)),
    ]
    peak = 6543.1
    pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peak = float(json.load(open(pk))["hbm_gbs"])
    for name, nbytes, fn in cases:
        for _ in range(args.warm):
            _lib.check(fn())
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        ev[0].record()
        for _ in range(args.iters):
            _lib.check(fn())
        ev[1].record()
        torch.cuda.synchronize()
        ms = ev[0].elapsed_time(ev[1]) / args.iters
        gbs = nbytes / ms / 1e6
        print(json.dumps({"kernel": name, "chains": N, "ms": ms, "algorithmic_bytes": nbytes, "GBps": gbs,
                          "frac_of_measured_hbm": gbs / peak}), flush=True)


if __name__ == "__main__":
    main()
