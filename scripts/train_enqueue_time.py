#!/usr/bin/env python
"""Is the training step bound by the host's launch rate or by the device?  One native loss + gradient evaluation
(l2hmc_u1_train_grads, one C call) enqueued on an idle device -- the launch queue cannot push back on the host -- gives
the host's enqueue time; K back-to-back calls give the device time per call."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch


def main():
    from l2hmc_b200.gauge_model import GaugeModel
    cases = {"c2": dict(time_size=8, space_size=8, num_samples=128, num_steps=5, eps=0.2, network_arch="generic"),
             "c3": dict(time_size=16, space_size=16, num_samples=2048, num_steps=5, eps=0.2, network_arch="conv3D")}
    for name in (sys.argv[1] if len(sys.argv) > 1 else "c2").split(","):
        np.random.seed(42)
        model = GaugeModel(dict(cases[name], rand=True, train_steps=10000))
        model.dynamics.seed = 3
        x = model._x(model.lattice.samples).contiguous()
        draws = model._draws(x, None, None, None)
        kw = dict(draws_x=draws[0], z=draws[1], draws_z=draws[2])

        def call():
            return model._calc_loss_and_grads(x, 2.0, **kw)
        for _ in range(3):
            call()
        torch.cuda.synchronize()
        K = 20
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(K):
            call()
        e1.record()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        single = []
        for _ in range(5):
            torch.cuda.synchronize()
            a = time.perf_counter()
            call()
            single.append((time.perf_counter() - a) * 1e3)
        torch.cuda.synchronize()
        print(json.dumps({"case": name, "single_call_host_enqueue_ms": sorted(single)[len(single) // 2], "calls": K,
                          "device_ms_per_call": e0.elapsed_time(e1) / K, "wall_ms_per_call": (t2 - t0) / K * 1e3}))


if __name__ == "__main__":
    main()
