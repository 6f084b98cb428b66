#!/usr/bin/env python
"""Per-launch cost of the config-2-sized operations (256 stacked rows, K = D = 128, H = 512) as a dependent chain sees
them: `reps` calls captured in one CUDA graph (no host enqueue cost), replayed, CUDA-event time / reps.  Shows which of
the small kernels sit above the launch-to-launch floor (the `counter_add` line)."""
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch


def graph_time(fn, reps=50, replays=20):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(reps):
                fn()
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(replays):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (reps * replays)      # us per call


def main():
    import l2hmc_b200 as pkg
    from l2hmc_b200 import _lib
    from l2hmc_b200.training import _Gemm, _ptr
    dev = torch.device("cuda", 0)
    R, K, H, D = 256, 128, 512, 128
    rnd = lambda *s: torch.randn(*s, device=dev)
    gemm = _Gemm(dev)
    out = {}
    lib = _lib.get_lib()
    ctr = torch.zeros(1, dtype=torch.int64, device=dev)
    out["counter_add (launch floor)"] = graph_time(
        lambda: _lib.check(lib.l2hmc_counter_add(ctypes.c_void_p(ctr.data_ptr()), 1, _lib.stream_ptr())))
    dz2, dz1, h1, wh = rnd(R, H), rnd(R, H), rnd(R, H), rnd(H, H)
    dA, ws, wt, wq = rnd(R, 3 * D), rnd(H, D), rnd(H, D), rnd(H, D)
    gx, win = rnd(R, D), rnd(K, H)
    for ks in (0, 1, 2, 4, 8):
        out[f"dz1 = relu'(h1) . (dz2 . Wh^T)  256x512x512 ksplit={ks}"] = graph_time(
            lambda: gemm(R, H, [(_ptr(dz2), H, 1, _ptr(wh), 1, H, H)], _ptr(dz1), H, relu_gate=_ptr(h1), ldg=H, ksplit=ks))
        out[f"dz2 = relu'(h2) . (dA . [Ws Wt Wq]^T)  256x512x(3x128) ksplit={ks}"] = graph_time(
            lambda: gemm(R, H, [(_ptr(dA), 3 * D, 1, _ptr(ws), 1, D, D), (_ptr(dA, D), 3 * D, 1, _ptr(wt), 1, D, D),
                                (_ptr(dA, 2 * D), 3 * D, 1, _ptr(wq), 1, D, D)], _ptr(dz2), H, relu_gate=_ptr(h1), ldg=H,
                         ksplit=ks))
        out[f"gx += dz1 . Win^T  256x128x512 ksplit={ks}"] = graph_time(
            lambda: gemm(R, K, [(_ptr(dz1), H, 1, _ptr(win), 1, H, H)], _ptr(gx), D, accumulate=True, ksplit=ks))
    # small weight gradients over the whole tape: h2^T [512 x 2560] . dA [2560 x 128] (row-major operands, TN form)
    Rt = 2560
    h2t, dAt, gw = rnd(Rt, H), rnd(Rt, 3 * D), rnd(H, D)
    for ks in (0, 8, 19, 26, 32, 40, 64):
        out[f"dW = h2^T . dA  512x128x2560 ksplit={ks}"] = graph_time(
            lambda: gemm(H, D, [(_ptr(h2t), 1, H, _ptr(dAt), 3 * D, 1, Rt)], _ptr(gw), D, ksplit=ks))
    win_t, dz1t, gwi = rnd(Rt, K), rnd(Rt, H), rnd(K, H)
    for ks in (19, 32, 40, 64):
        out[f"dW_in = in^T . dz1  128x512x2560 ksplit={ks}"] = graph_time(
            lambda: gemm(K, H, [(_ptr(win_t), 1, K, _ptr(dz1t), H, 1, Rt)], _ptr(gwi), H, ksplit=ks))
    # forward net call on the FFMA path (3 dense kernels) and the fused position update
    lat = pkg.GaugeLattice(8, 8, 2, "U1", num_samples=R)
    np.random.seed(0)
    dyn = pkg.GaugeDynamics(lat, hmc=False, network_arch="generic", num_steps=5, eps=0.1)
    net = dyn.position_fn
    x, v = rnd(R, D), rnd(R, D)
    out["net call S,T,Q (3 dense kernels), path auto"] = graph_time(lambda: net([v, x, (1.0, 0.0)]))
    for path, nm in ((_lib.NET_PATH_FFMA, "ffma"), (_lib.NET_PATH_TCGEN05, "tcgen05")):
        try:
            net.path = path
            out[f"net call S,T,Q, path {nm}"] = graph_time(lambda: net([v, x, (1.0, 0.0)]))
        except Exception as e:      # noqa
            out[f"net call S,T,Q, path {nm}"] = "n/a: " + str(e)[:80]
    # FFMA layers against the tcgen05 layers by batch size: where l2hmc_net_run's AUTO rule should switch
    for (T, arch, tag) in ((8, "generic", "8x8 generic K=128 H=512 D=128"), (16, "conv3D", "16x16 conv3D K=512 H=1024 D=512")):
        for rows in (256, 512, 1024, 2048, 4096):
            lat = pkg.GaugeLattice(T, T, 2, "U1", num_samples=rows)
            np.random.seed(0)
            dyn = pkg.GaugeDynamics(lat, hmc=False, network_arch=arch, num_steps=5, eps=0.1)
            net = dyn.position_fn
            Dd = 2 * T * T
            x, v = rnd(rows, Dd), rnd(rows, Dd)
            for path, nm in ((_lib.NET_PATH_FFMA, "ffma"), (_lib.NET_PATH_TCGEN05, "tcgen05")):
                net.path = path
                out[f"sweep {tag} rows={rows} path={nm}"] = graph_time(lambda: net([v, x, (1.0, 0.0)]), reps=20, replays=10)
    for k, val in out.items():
        print(json.dumps({"op": k, "us_per_call": val}))


if __name__ == "__main__":
    main()
