#!/usr/bin/env python
"""One training step (forward with tape + backward + Adam) for ncu launch lists: --case c2 | c3."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

ap = argparse.ArgumentParser()
ap.add_argument("--case", default="c3")
ap.add_argument("--steps", type=int, default=1)
ap.add_argument("--warm", type=int, default=1)
args = ap.parse_args()
from l2hmc_b200.gauge_model import GaugeModel
prm = {"c2": dict(time_size=8, space_size=8, num_samples=128, num_steps=5, eps=0.2, network_arch="generic"),
       "c3": dict(time_size=16, space_size=16, num_samples=2048, num_steps=5, eps=0.2, network_arch="conv3D")}[args.case]
np.random.seed(42)
model = GaugeModel(dict(prm, rand=True))
x = model._x(model.lattice.samples).contiguous()
for _ in range(args.warm):
    x = model.train_step(x, 2.0)[1]
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
for _ in range(args.steps):
    out = model.train_step(x, 2.0)
    x = out[1]
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("loss", float(out[0]), "eps", float(out[3]))
