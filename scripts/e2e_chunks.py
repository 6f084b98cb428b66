#!/usr/bin/env python
"""e2e (host buffers in and out every step) of the config-4 HMC MC step against the sampler's chunk size."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from l2hmc_b200 import GaugeDynamics, GaugeLattice  # noqa: E402
from l2hmc_b200.sampler import GaugeSampler  # noqa: E402

T = X = 32
N, M, D = 65536, 10, 2048
lat = GaugeLattice(T, X, 2, "U1", num_samples=1)
dyn = GaugeDynamics(lat, hmc=True, num_steps=M, eps=0.1, seed=1)
x0 = (torch.rand(N, D) * (2 * np.pi)).float()
for chunk in [int(a) for a in (sys.argv[1:] or ["4096", "8192", "16384", "32768", "65536"])]:
    s = GaugeSampler(dyn, N, chunk_chains=chunk)
    host = torch.empty(N, D, dtype=torch.float32).pin_memory()
    host.copy_(x0)
    for _ in range(2):
        res = s.step(host, 4.0)
        host.copy_(res.samples)
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    K = 10
    for _ in range(K):
        res = s.step(host, 4.0)
        host, s.h_out = res.samples, host
    ev[1].record()
    torch.cuda.synchronize()
    ms = ev[0].elapsed_time(ev[1]) / K
    print(json.dumps({"chunk_chains": chunk, "ms_per_step": ms, "link_updates_per_s": N * D * M / (ms * 1e-3),
                      "GBps_each_way": N * D * 4 / ms / 1e6}))
    del s, host
