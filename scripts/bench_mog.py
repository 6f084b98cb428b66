#!/usr/bin/env python
"""Config-5 MoG MC step alone (2^20 chains, k = 2, L = 2, H = 50): CUDA-event time per launch."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from l2hmc_b200 import mog  # noqa: E402

N, k, L, H = 1 << 20, 2, 2, 50
gmm = mog.GMM([[1., 0.], [-1., 0.]], [0.05 * np.eye(2)] * 2, [0.5, 0.5])
np.random.seed(42)
dyn = mog.Dynamics(k, gmm, trajectory_length=L, eps=0.1, seed=11,
                   net_factory=lambda d, scope, factor: mog.network(d, scope, factor, num_nodes=H))
x = torch.randn(N, k, device="cuda") * 0.2 + torch.tensor([1.0, 0.0], device="cuda")
for _ in range(3):
    x = dyn.mc_step(x)["x_out"]
torch.cuda.synchronize()
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 10
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
ev[0].record()
for _ in range(iters):
    x = dyn.mc_step(x)["x_out"]
ev[1].record()
torch.cuda.synchronize()
ms = ev[0].elapsed_time(ev[1]) / iters
flop = N * L * 5 * 2.0 * (2 * k * H + 2 * H + H * H + 3 * H * k)
print(json.dumps({"mc_step_ms": ms, "chain_updates_per_s": N / (ms * 1e-3), "fp32_tflops": flop / ms / 1e9}))
