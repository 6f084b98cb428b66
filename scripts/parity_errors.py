#!/usr/bin/env python
"""Measured parity table (profiles/r2_parity_errors.txt): for every BASELINE config shape, the error of the CUDA path
against the float64 oracle next to the fp32 oracle's own error against float64, for each quantity the north star
names (links, log-dets, actions, plaquettes; accept decisions and integer charge as mismatch counts).

Relative errors are max|a - ref| / max|ref| (tests/parity_util.rel_err); accept_prob is an absolute error.
Run on the GPU box:  python scripts/parity_errors.py [out.txt]     (test infrastructure: imports oracle/)"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]

from oracle import u1_oracle as orc                       # noqa: E402
from parity_util import format_table, measure_case, oracle_pair, trained_like_weights   # noqa: E402

# (label, arch or None for plain HMC, T, X, N, M, eps, beta, path, head gain, hot start)
CASES = [
    ("C1 8x8 HMC N=64 M=10", None, 8, 8, 64, 10, 0.1, 2.0, None, None, True),
    ("C2 8x8 generic H=512 N=128 M=5 (FFMA)", "generic", 8, 8, 128, 5, 0.25, 2.0, "ffma", 1.0, False),
    ("C2 shape, N=512 (tcgen05)", "generic", 8, 8, 512, 5, 0.25, 2.0, "tc", 1.0, False),
    ("C3 16x16 conv3D F=16 H=1024 N=64 M=5 (FFMA)", "conv3D", 16, 16, 64, 5, 0.25, 3.0, "ffma", 1.0, False),
    ("C3 16x16 conv3D N=512 M=5 (tcgen05)", "conv3D", 16, 16, 512, 5, 0.25, 3.0, "tc", 1.0, False),
    ("C4 32x32 HMC N=64 M=10", None, 32, 32, 64, 10, 0.1, 4.0, None, None, True),
    ("C4 32x32 generic H=8192 N=8 M=10 (FFMA)", "generic", 32, 32, 8, 10, 0.1, 4.0, "ffma", 0.1, False),
    ("C4 32x32 generic H=8192 N=256 M=5 (tcgen05)", "generic", 32, 32, 256, 5, 0.1, 4.0, "tc", 0.3, False),
    ("C4 32x32 conv3D F=32 H=4096 N=8 M=10 (FFMA)", "conv3D", 32, 32, 8, 10, 0.1, 4.0, "ffma", 0.6, False),
    ("C4 32x32 conv3D F=32 N=256 M=5 (tcgen05)", "conv3D", 32, 32, 256, 5, 0.1, 4.0, "tc", 1.2, False),
]


def main(out_path):
    import l2hmc_b200 as pkg
    from l2hmc_b200 import _lib
    rows, cache = [], {}
    for label, arch, T, X, N, M, eps, beta, path, gain, hot in CASES:
        t0 = time.time()
        D = 2 * T * X
        rng = np.random.default_rng(1000 + T + N + M)
        x = (rng.uniform(0, 2 * np.pi, (N, D)) if hot else np.mod(rng.normal(0, 0.45, (N, D)), 2 * np.pi)).astype(np.float32)
        vf, vb = (rng.standard_normal((N, D)).astype(np.float32) for _ in range(2))
        ud, ua = (rng.uniform(size=N).astype(np.float32) for _ in range(2))
        mask = orc.construct_masks(D, M, seed=42)
        ws = None
        if arch is not None:
            key = (arch, T, X, gain)
            if key not in cache:
                cache.clear()                              # one set of config-4 weights at a time (GBs in float64)
                cache[key] = trained_like_weights(arch, T, X, np.random.default_rng(77 + T), gain=gain)
            ws = cache[key]
        od32, od64 = oracle_pair(arch or "generic", T, X, M, eps, mask, ws)
        lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=N)
        np.random.seed(42)
        dyn = pkg.GaugeDynamics(lat, hmc=arch is None, network_arch=arch or "generic", num_steps=M, eps=eps, weights=ws)
        dyn.set_masks(mask)
        if path:
            dyn.position_fn.path = dyn.momentum_fn.path = {"ffma": _lib.NET_PATH_FFMA, "tc": _lib.NET_PATH_TCGEN05}[path]
        errs = measure_case(dyn, od32, od64, x, beta, vf, vb, ud, ua, T, X)
        dyn.raise_if_nonfinite()
        rows.append((label, errs))
        print(f"{label}: {time.time() - t0:.1f} s", flush=True)
        del dyn, od32, od64
    text = ("# CUDA path vs float64 oracle, next to fp32 oracle vs float64 (same seeded inputs).\n"
            "# relative error = max|a - ref| / max|ref|; accept_prob absolute; *_mismatch / *_unsafe are chain counts\n"
            "# (accept_unsafe: chains with |p - u| < 1e-3, excluded from the bit-exact accept comparison).\n"
            "# north star: <= 1e-5 relative in fp32; rule of the tests: <= max(1e-5, 2 x the fp32 oracle's error).\n\n"
            + format_table(rows))
    print(text)
    if out_path:
        with open(out_path, "w") as f:
            f.write(text + "\n")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else None)
