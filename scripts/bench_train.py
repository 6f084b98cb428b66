#!/usr/bin/env python
"""Training-step timing: BASELINE config 2 (8x8, generic H = 512, batch 128, M = 5) and config 3
(16x16, conv3D F = 16, H = 1024, 2048 chains per GPU, M = 5).  CUDA-event times of forward-with-tape, backward and the
whole train_step; optionally the CPU oracle (torch autograd through the oracle's transition) beside it."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch


def events_ms(fn, iters, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", default="c2,c3")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--cpu", action="store_true")
    args = ap.parse_args()
    from l2hmc_b200.gauge_model import GaugeModel
    cases = {"c2": dict(time_size=8, space_size=8, num_samples=128, num_steps=5, eps=0.2, network_arch="generic"),
             "c3": dict(time_size=16, space_size=16, num_samples=2048, num_steps=5, eps=0.2, network_arch="conv3D"),
             "c2hmc": dict(time_size=8, space_size=8, num_samples=128, num_steps=5, eps=0.2, hmc=True)}
    for name in args.cases.split(","):
        prm = dict(cases[name], rand=True, train_steps=10000)
        np.random.seed(42)
        model = GaugeModel(prm)
        model.dynamics.seed = 3
        x = [model._x(model.lattice.samples).contiguous()]
        eng = model.engine
        N, D = x[0].shape
        w = dict(model.loss_weights)
        draws = model._draws(x[0], None, None, None)

        def fwd():
            return eng.forward(x[0], 2.0, draws[0], draws[1], draws[2], w)
        fw = [fwd()]

        def bwd():
            eng.backward(fw[0])

        def step():
            x[0] = model.train_step(x[0], 2.0)[1]
        out = {"case": name, "chains": N, "D": D, "M": model.num_steps, "arch": prm.get("network_arch", "hmc"),
               "params": int(eng.params.size)}
        out["forward_tape_ms"] = events_ms(fwd, args.iters)
        l0 = eng.launches + eng.gemm.launches
        bwd()
        out["backward_launches_estimate"] = eng.launches + eng.gemm.launches - l0   # host-side tally (a GEMM with workspace counts 2); the ncu launch list is exact
        out["backward_ms"] = events_ms(bwd, args.iters)
        out["train_step_ms"] = events_ms(step, args.iters)
        t0 = time.perf_counter()
        for _ in range(args.iters):
            step()
        torch.cuda.synchronize()
        out["train_step_wall_ms"] = (time.perf_counter() - t0) / args.iters * 1e3
        out["link_updates_per_s"] = N * D * model.num_steps / (out["train_step_ms"] * 1e-3)
        if args.cpu and name == "c2":
            from oracle import train_oracle as tro
            from oracle import u1_oracle as orc
            torch.set_num_threads(os.cpu_count())
            dyn = model.dynamics
            wx = {k: v.cpu().numpy() for k, v in dyn.position_fn.weights.items()}
            wv = {k: v.cpu().numpy() for k, v in dyn.momentum_fn.weights.items()}
            od = orc.OracleDynamics(8, 8, 5, 0.2, hmc=False, network_arch="generic", mask=dyn.mask.cpu().numpy(), xnet=wx,
                                    vnet=wv)
            cpu = lambda t: tuple(a.cpu().numpy() for a in t)
            xs = x[0].cpu().numpy()
            args_ = (od, xs, 2.0, cpu(draws[0]), draws[1].cpu().numpy(), cpu(draws[2]))
            tro.loss_and_grads(*args_, **w)
            t0 = time.perf_counter()
            for _ in range(3):
                tro.loss_and_grads(*args_, **w)
            out["cpu_oracle_loss_and_grads_ms"] = (time.perf_counter() - t0) / 3 * 1e3
            out["cpu_cores"] = os.cpu_count()
        print(json.dumps(out), flush=True)
        del model, eng, fw
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
