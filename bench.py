#!/usr/bin/env python
"""bench.py -- link-updates/s of the U(1) leapfrog path on N B200s.

Workload (BASELINE.json metric / config 4): 32x32 lattice, 65 536 chains PER GPU (weak
scaling; the 1 GB of links+momenta per GPU is ~8x the 126 MB L2), plain HMC, M = 10 leapfrog
steps, eps = 0.1, beta = 4, hot start, free-running Philox momenta.  A "step" is one MC step
(GaugeDynamics.apply_transition + observables) = N*D*M link-updates.

  value      device-resident throughput of the product path (one-launch MC step)
  e2e        same through GaugeSampler.step with pinned HOST buffers (H2D + D2H inside)
  roofline   the fused leapfrog-STEP kernel (x,v streamed once per LF step, 16 B/link-update),
             the kernel the north-star's >= 70 % HBM target names, timed live in its own loop
  roofline_mc  the one-launch MC-step kernel against its own algorithmic HBM bytes
  cpu_baseline the oracle restatement of the reference's TF graph on the host cores

`--impl reference` times that CPU restatement alone (TensorFlow 1.x cannot be installed here,
DESIGN.md section 7), rank 0 only.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

# Only the JSON line may reach stdout: libraries (NCCL prints its version banner there) write to fd 1
# directly, so fd 1 is pointed at stderr for the whole run and the line goes to a saved copy.
_REAL_STDOUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)


def emit(obj):
    _REAL_STDOUT.write(json.dumps(obj) + "\n")
    _REAL_STDOUT.flush()

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

T = X = 32
D = 2 * T * X
M = 10
EPS = 0.1
BETA = 4.0
METRIC = "link-updates/s"


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def ncu_traffic(kernel):
    """dram bytes per launch from the committed ncu capture (profiles/r1_traffic.json), or None."""
    path = os.path.join(ROOT, "profiles", "r1_traffic.json")
    try:
        with open(path) as f:
            return float(json.load(f)[kernel])
    except (OSError, KeyError, ValueError):
        return None


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons while the timed region runs (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.proc = index, [], None

    def run(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
        self.join(timeout=2)
        sm, reasons, mx = [], set(), None
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                for nm, val in zip(names, r[3:7]):
                    if val.lower().startswith("active"):
                        reasons.add(nm)
            except (ValueError, IndexError):
                continue
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_reference_run(steps, warmup, chains):
    """The oracle restatement of the reference's TF graph (apply_transition: BOTH directions
    integrated, one discarded, exactly as gauge_dynamics.py:211-241) on the host cores.
    Returns (link-updates/s, s/step, threads).  Only this function touches oracle/."""
    import numpy as np
    import torch
    from oracle import u1_oracle as orc
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    rng = np.random.default_rng(0)
    dyn = orc.OracleDynamics(T, X, M, EPS, hmc=True, mask=np.zeros((M, D), np.float32))
    x = rng.uniform(0, 2 * np.pi, (chains, D)).astype(np.float32)
    times = []
    with torch.no_grad():
        for i in range(warmup + steps):
            vf = rng.standard_normal((chains, D)).astype(np.float32)
            vb = rng.standard_normal((chains, D)).astype(np.float32)
            ud, ua = rng.uniform(size=chains).astype(np.float32), rng.uniform(size=chains).astype(np.float32)
            t0 = time.perf_counter()
            _, _, _, xo = dyn.apply_transition(x, BETA, vf, vb, ud, ua)
            orc.plaq_observables(xo, T, X)
            x = orc.wrap_2pi(xo.numpy())
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt)
    per_step = sum(times) / len(times)
    return chains * D * M / per_step, per_step, threads


def run_reference(args, rank):
    if rank != 0:
        return
    chains = args.cpu_chains
    value, per_step, threads = cpu_reference_run(args.steps, args.warmup, chains)
    sample = (f"{chains} chains of the 32x32 workload per step (the GPU arm runs {args.chains_per_gpu} per GPU); "
              f"torch-CPU fp32 restatement of the TF graph, both directions integrated like the reference")
    emit(({
        "impl": "reference", "metric": METRIC, "value": value, "unit": "link-updates/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, args.chains_per_gpu),
        "cpu_baseline": {"value": value, "unit": "link-updates/s", "cores": threads, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": "link-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "TensorFlow 1.x is not installable in this image: the reference arm is the oracle port",
    }))


def workload_config(args, chains_per_gpu):
    return {"workload": "U(1) 32x32 plain-HMC MC step (BASELINE config 4), hot start",
            "lattice": "32x32", "chains_per_gpu": chains_per_gpu, "total_chains": chains_per_gpu * args.gpus,
            "lf_steps": M, "eps": EPS, "beta": BETA, "momenta": "Philox4x32-10 in-kernel",
            "observables": "action, plaquette, integer charge every step",
            "cache": "inputs larger than L2 (x and v are 512 MiB each per GPU vs 126 MB L2)",
            "parallelism": f"chains sharded over {args.gpus} GPU(s), no data-path collective"}


def run_gpu(args, rank, local_rank, world):
    import numpy as np
    import torch
    import torch.distributed as dist
    from l2hmc_b200 import GaugeDynamics, GaugeLattice, _lib
    from l2hmc_b200.sampler import GaugeSampler

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    from l2hmc_b200.sharding import bind_to_gpu_numa
    numa_cpus = bind_to_gpu_numa(local_rank) if world > 1 else None     # pinned e2e buffers on the GPU's own node
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.get_lib()                                   # fail loudly if the CUDA library is missing
    N = args.chains_per_gpu
    K, W = args.steps, max(args.warmup, 3)
    lat = GaugeLattice(T, X, 2, "U1", num_samples=1)
    from l2hmc_b200.sharding import shard_range
    chain_lo, chain_hi = shard_range(world * N, rank, world)          # contiguous block of the global batch
    assert chain_hi - chain_lo == N
    dyn = GaugeDynamics(lat, hmc=True, num_steps=M, eps=EPS, seed=2024, chain_offset=chain_lo, device=dev)
    gen = torch.Generator(device=dev).manual_seed(1000 + rank)
    x = torch.rand(N, D, device=dev, generator=gen) * (2 * np.pi)      # synthetic hot start

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def max_over_ranks(ms):
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return ms

    # ------------------------------------------------ device-resident MC steps (the headline `value`)
    xa, xb = x.clone(), torch.empty_like(x)
    p = torch.empty(N, device=dev)
    act, plaq = torch.empty(N, device=dev), torch.empty(N, device=dev)
    q = torch.empty(N, dtype=torch.int32, device=dev)
    sums = torch.zeros(4, dtype=torch.float64, device=dev)
    lib = _lib.get_lib()

    def mc_step(src, dst):
        a = _lib.HmcMcArgs()
        a.x, a.x_out, a.accept_prob = _lib.dptr(src), _lib.dptr(dst), _lib.dptr(p)
        a.beta, a.eps, a.num_steps, a.default_dir, a.wrap_2pi = BETA, EPS, M, 1, 1
        a.seed, a.mc_step, a.chain0 = dyn.seed, dyn._mc_step, dyn.chain_offset
        a.N, a.T, a.X = N, T, X
        a.nonfinite_flag = _lib.dptr(dyn._flag, torch.int32)
        a.obs_action, a.obs_plaq, a.obs_charge = _lib.dptr(act), _lib.dptr(plaq), _lib.dptr(q, torch.int32)
        _lib.check(lib.l2hmc_u1_hmc_mc_step(a, _lib.stream_ptr()))
        dyn._mc_step += 1

    def accumulate():
        return torch.stack([plaq.sum(dtype=torch.float64), act.sum(dtype=torch.float64), p.sum(dtype=torch.float64),
                            n_chains])
    n_chains = torch.tensor(float(N), dtype=torch.float64, device=dev)
    for _ in range(W):
        mc_step(xa, xb)
        xa, xb = xb, xa
    accumulate()                                   # (loads torch's reduction kernels outside the timed region)
    if world > 1:
        dist.all_reduce(torch.zeros(4, dtype=torch.float64, device=dev))
    clocks = ClockSampler(local_rank)
    clocks.start()
    time.sleep(0.3)
    barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev_mc = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    ev[0].record()
    for i in range(K):
        ev_mc[i][0].record()
        mc_step(xa, xb)
        ev_mc[i][1].record()
        xa, xb = xb, xa
    sums += accumulate()
    if world > 1:
        dist.all_reduce(sums)                      # the only collective: O(1) observable accumulators
    ev[1].record()
    barrier()
    ms_total = max_over_ranks(ev[0].elapsed_time(ev[1]))
    step_ms = ms_total / K
    value = world * N * D * M / (step_ms * 1e-3)
    dyn.raise_if_nonfinite()
    obs = {"avg_plaq": float(sums[0] / sums[3]), "accept_prob": float(sums[2] / sums[3])}

    # ------------------------------------------------ the fused LF-step kernel, its own timed loop
    v = torch.randn(N, D, device=dev, generator=gen)
    x1, v1 = torch.empty_like(x), torch.empty_like(v)
    bufs = [(xa, v), (x1, v1)]

    def lf_launch(i):
        (sx, sv), (dx, dv) = bufs[i % 2], bufs[(i + 1) % 2]
        _lib.check(lib.l2hmc_u1_hmc_step(_lib.dptr(sx), _lib.dptr(sv), BETA, EPS, None, N, T, X, _lib.dptr(dx),
                                         _lib.dptr(dv), _lib.stream_ptr()))
    for i in range(4):
        lf_launch(i)
    barrier()
    nl = K * M
    ev[0].record()
    for i in range(nl):
        lf_launch(i)
    ev[1].record()
    barrier()
    lf_ms = max_over_ranks(ev[0].elapsed_time(ev[1])) / nl
    clk = clocks.stop()

    peak, peak_kind = measured_peaks()
    lf_bytes = 16.0 * N * D                           # read x, v; write x, v
    lf_gbs = lf_bytes / (lf_ms * 1e-3) / 1e9
    mc_kernel_ms = sum(a.elapsed_time(b) for a, b in ev_mc) / K
    mc_bytes = 8.0 * N * D                            # MC kernel: read x, write x_out (observables fused)
    mc_gbs = mc_bytes / (mc_kernel_ms * 1e-3) / 1e9
    return dict(value=value, step_ms=step_ms, clk=clk, obs=obs, peak=peak, peak_kind=peak_kind,
                lf=dict(ms=lf_ms, gbs=lf_gbs, bytes=lf_bytes), mc=dict(ms=mc_kernel_ms, gbs=mc_gbs, bytes=mc_bytes),
                ctx=dict(numa_cpus=numa_cpus, dyn=dyn, dev=dev, N=N, K=K, W=W, barrier=barrier, max_over_ranks=max_over_ranks, x=xa,
                         GaugeSampler=GaugeSampler, torch=torch, world=world))


def run_e2e(r):
    """Same MC step through GaugeSampler.step: pinned host links in, host results out, every step."""
    c = r["ctx"]
    torch, N, K = c["torch"], c["N"], c["K"]
    sampler = c["GaugeSampler"](c["dyn"], N, chunk_chains=8192)
    host = torch.empty(N, D, dtype=torch.float32).pin_memory()
    host.copy_(c["x"])
    for _ in range(2):
        res = sampler.step(host, BETA)
        host.copy_(res.samples)
    c["barrier"]()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(K):
        res = sampler.step(host, BETA)          # H2D of the links, MC step, observables, D2H of everything
        host, sampler.h_out = res.samples, host  # next step feeds what this one returned (no host copy)
    ev[1].record()
    c["barrier"]()
    ms = c["max_over_ranks"](ev[0].elapsed_time(ev[1]))   # step() ends with a host sync: covers copies + kernels
    step_ms = ms / K
    return {"value": c["world"] * N * D * M / (step_ms * 1e-3), "unit": "link-updates/s",
            "h2d_bytes_per_step": sampler.h2d_bytes, "d2h_bytes_per_step": sampler.d2h_bytes,
            "ms_per_step": step_ms, "launches_per_step": sampler.launches_per_step,
            "host_cpus_bound": None if not c.get("numa_cpus") else len(c["numa_cpus"]),
            "plaq": float(res.plaqs.mean()), "charge_diff": res.charge_diff}


def run_l2hmc(r, steps):
    """Second leg: the L2HMC path with S/T/Q networks (BASELINE config 3: 16x16, conv3D front-end,
    2048 chains per GPU, M = 5) -- MC-step throughput, and the tensor-pipe roofline of the network
    GEMMs (tcgen05, 3xTF32) at the config-3 and config-4 (32x32 generic, H = 8192) shapes."""
    import numpy as np
    c = r["ctx"]
    torch, dev, world = c["torch"], c["dev"], c["world"]
    from l2hmc_b200 import GaugeDynamics, GaugeLattice, _lib
    peaks = {}
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            peaks = json.load(f)
    bf16_peak, peak_kind = float(peaks.get("bf16_tflops", 1590.0)), ("measured" if peaks else "fallback")
    tf32_peak = bf16_peak / 2.0           # kind::tf32 runs at half the bf16 rate (1.1 vs 2.25 PF nominal)
    out = {}

    def events_ms(fn, iters, warm=2):
        for _ in range(warm):
            fn()
        c["barrier"]()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        ev[0].record()
        for _ in range(iters):
            fn()
        ev[1].record()
        c["barrier"]()
        return c["max_over_ranks"](ev[0].elapsed_time(ev[1])) / iters

    for tag, arch, T_, N, M_, eps, do_mc in (("config3_16x16_conv3D", "conv3D", 16, 2048, 5, 0.25, True),
                                              ("config4_32x32_generic_nets", "generic", 32, 2048, 10, 0.1, True)):
        D_ = 2 * T_ * T_
        lat = GaugeLattice(T_, T_, 2, "U1", num_samples=1)
        np.random.seed(42)
        dyn = GaugeDynamics(lat, hmc=False, network_arch=arch, num_steps=M_, eps=eps, seed=7,
                            chain_offset=c["dyn"].chain_offset // 32, device=dev)
        net = dyn.position_fn
        K_, H_ = net.num_features, net.num_hidden
        gen = torch.Generator(device=dev).manual_seed(5)
        x = torch.rand(N, D_, device=dev, generator=gen) * (2 * np.pi)
        v = torch.randn(N, D_, device=dev, generator=gen)
        ld = torch.zeros(N, device=dev)
        flop = 2.0 * N * (2 * K_ * H_ + 2 * H_ + H_ * H_ + 3 * H_ * D_)
        feats = net._features(v, x, dyn.mask[0], False)       # conv features (or the inputs) outside the GEMM timing
        call = _lib.NetCall()
        call.mode, call.eps, call.dir = _lib.NET_POSITION, eps, 1
        call.in1_mask, call.in1_mask_invert = _lib.dptr(feats[2]), 0
        x_new, mask0 = torch.empty_like(x), dyn.mask[0].contiguous()
        call.x, call.v, call.mask, call.out = _lib.dptr(x), _lib.dptr(v), _lib.dptr(mask0), _lib.dptr(x_new)
        call.logdet, call.nonfinite_flag = _lib.dptr(ld), _lib.dptr(dyn._flag, torch.int32)
        ms = events_ms(lambda: net._run(call, feats[0], feats[1], N, (1.0, 0.0)), max(5, steps // 2))
        tf = flop / ms / 1e9
        entry = {"chains_per_gpu": N, "K": K_, "H": H_, "D": D_, "lf_steps": M_,
                 "net_call": {"what": "3 tcgen05 GEMMs (two-input layer, hidden layer, 3 heads with the position "
                                      "update + log-det in the epilogue) + operand converters, one position_fn call",
                              "ms": ms, "gflop_algorithmic": flop / 1e9},
                 "roofline": {"bound": "tensor", "achieved": tf, "unit": "TFLOP/s",
                              "tensor_pipe_tflops": 3 * tf, "peak": tf32_peak, "peak_kind": peak_kind + " bf16 / 2",
                              "frac": 3 * tf / tf32_peak,
                              "note": "achieved = algorithmic fp32 flops / time; every product is 3 TF32 MMAs "
                                      "(hi.lo + lo.hi + hi.hi), so the tensor pipe runs at 3x that rate"}}
        if do_mc:
            x_mc = [x]

            def mc():
                x_mc[0] = dyn.apply_transition(x_mc[0], 3.0, wrap=True)[3]
            step_ms = events_ms(mc, max(3, steps // 4), warm=1 if T_ >= 32 else 2)
            entry["mc_step_ms"] = step_ms
            entry["link_updates_per_s"] = world * N * D_ * M_ / (step_ms * 1e-3)
            dyn.raise_if_nonfinite()
        out[tag] = entry
        del dyn, net, x, v
        torch.cuda.empty_cache()
    out["config5_mog"] = run_mog(c, events_ms, steps)
    out["training"] = run_training(c, events_ms, steps)
    return out


def run_training(c, events_ms, steps):
    """Training step at BASELINE config 2 (8x8, generic S/T/Q MLPs H = 4D = 512, batch 128, M = 5)
    and config 3 (16x16, conv3D F = 16, H = 1024, 2048 chains per GPU, M = 5): loss over the x- and z-chain with tape,
    explicit backward pass (one C call, csrc/train_step.cu), [gradient all-reduce at N > 1], Adam, observables.
    Rank 0 at N = 1 also times the CPU oracle (torch autograd through the oracle's transition) on the config-2 step."""
    import numpy as np
    from l2hmc_b200.gauge_model import GaugeModel
    torch, dev, world = c["torch"], c["dev"], c["world"]
    out = {}
    for tag, prm in (("config2_8x8_generic", dict(time_size=8, space_size=8, num_samples=128, network_arch="generic")),
                     ("config3_16x16_conv3D", dict(time_size=16, space_size=16, num_samples=2048, network_arch="conv3D"))):
        np.random.seed(42)
        model = GaugeModel(dict(prm, num_steps=5, eps=0.2, rand=True, train_steps=10000), device=dev)
        N, D, M = prm["num_samples"], 2 * prm["time_size"] ** 2, 5
        model.dynamics.seed, model.dynamics.chain_offset = 13, (c["dyn"].chain_offset // 65536) * N
        model.using_hvd = world > 1
        eng = model.engine
        if world > 1:
            from l2hmc_b200.sharding import broadcast_weights
            broadcast_weights([eng.params.flat])
            eng.params.sync_to_dynamics()
        state = [model._x(model.lattice.samples).contiguous()]

        def step():
            state[0] = model.train_step(state[0], 2.0)[1]
        ms = events_ms(step, max(5, steps // 2), warm=2)
        entry = {"chains_per_gpu": N, "H": int(model.dynamics.position_fn.num_hidden), "D": D, "lf_steps": M,
                 "params": int(eng.params.size), "train_step_ms": ms,
                 "link_updates_per_s": world * N * D * M / (ms * 1e-3),
                 "note": "x- and z-chain forward with tape + backward + Adam + observables; the x-chain's advance is "
                         "counted once; fp32: tcgen05 3xTF32 layers / backward products from 1024 stacked rows, FFMA "
                         "(cp.async multi-stage, split-K) below"}
        if tag.startswith("config2") and world == 1 and not c.get("no_cpu"):
            import time
            from oracle import train_oracle as tro
            from oracle import u1_oracle as orc
            torch.set_num_threads(os.cpu_count())
            dyn = model.dynamics
            od = orc.OracleDynamics(8, 8, M, dyn._eps(), hmc=False, network_arch="generic", mask=dyn.mask.cpu().numpy(),
                                    xnet={k: v.cpu().numpy() for k, v in dyn.position_fn.weights.items()},
                                    vnet={k: v.cpu().numpy() for k, v in dyn.momentum_fn.weights.items()})
            rng = np.random.default_rng(3)
            dr = lambda: (rng.standard_normal((N, D)).astype(np.float32), rng.standard_normal((N, D)).astype(np.float32),
                          rng.uniform(size=N).astype(np.float32), rng.uniform(size=N).astype(np.float32))
            args = (od, state[0].cpu().numpy(), 2.0, dr(), rng.standard_normal((N, D)).astype(np.float32), dr())
            tro.loss_and_grads(*args, **model.loss_weights)
            t0 = time.perf_counter()
            for _ in range(3):
                tro.loss_and_grads(*args, **model.loss_weights)
            cpu_ms = (time.perf_counter() - t0) / 3 * 1e3
            entry["cpu_baseline"] = {"value": N * D * M / (cpu_ms * 1e-3), "unit": "link-updates/s", "ms_per_step": cpu_ms,
                                     "cores": os.cpu_count(), "kind": "port",
                                     "sample": "3 loss-and-gradient evaluations of the same config-2 batch (torch "
                                               "autograd through the oracle's transition; no optimiser step)"}
        out[tag] = entry
        del model, eng
        torch.cuda.empty_cache()
    return out


def run_mog(c, events_ms, steps):
    """BASELINE config 5: two-mode Gaussian mixture, k = 2, M = 2 leapfrog steps, H = 50, 2^20 chains
    per GPU; one launch of mog_mc_kernel = one MC step (both nets evaluated in shared memory)."""
    import numpy as np
    from l2hmc_b200 import mog
    torch, dev, world = c["torch"], c["dev"], c["world"]
    N, k, L, H = 1 << 20, 2, 2, 50
    gmm = mog.GMM([[1., 0.], [-1., 0.]], [0.05 * np.eye(2)] * 2, [0.5, 0.5], device=dev)
    np.random.seed(42)
    dyn = mog.Dynamics(k, gmm, trajectory_length=L, eps=0.1, seed=11, device=dev,
                       chain_offset=c["dyn"].chain_offset // 32 * 512,
                       net_factory=lambda d, scope, factor: mog.network(d, scope, factor, num_nodes=H, device=dev))
    gen = torch.Generator(device=dev).manual_seed(3)
    x = [torch.randn(N, k, device=dev, generator=gen) * 0.2 + torch.tensor([1.0, 0.0], device=dev)]

    def mc():
        x[0] = dyn.mc_step(x[0])["x_out"]
    ms = events_ms(mc, max(5, steps))
    # per chain per step: 3 V-net + 2 X-net calls per LF step, each 2(kH + kH + 2H + HH + 3Hk) flops
    flop = N * L * 5 * 2.0 * (2 * k * H + 2 * H + H * H + 3 * H * k)
    return {"chains_per_gpu": N, "k": k, "lf_steps": L, "H": H, "mc_step_ms": ms,
            "chain_updates_per_s": world * N / (ms * 1e-3), "lf_steps_per_s": world * N * L / (ms * 1e-3),
            "ffma_tflops": flop / ms / 1e9,
            "algorithmic_bytes": N * (4 * k * 4 + 4 + 1 + 4), "hbm_gbs": N * (4 * k * 4 + 9) / ms / 1e6,
            "note": "compute-bound: 50x50 hidden layer per net call per chain on the FFMA pipe, nets in shared memory"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--chains-per-gpu", type=int, default=65536)
    ap.add_argument("--cpu-chains", type=int, default=1024)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-l2hmc", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    r = run_gpu(args, rank, local_rank, world)
    e2e = None if args.no_e2e else run_e2e(r)
    r["ctx"]["no_cpu"] = bool(args.no_cpu_baseline) or rank != 0
    l2hmc = None if args.no_l2hmc else run_l2hmc(r, args.steps)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, per_step, threads = cpu_reference_run(steps=10, warmup=1, chains=args.cpu_chains)
        cpu = {"value": v, "unit": "link-updates/s", "cores": threads, "kind": "port",
               "sample": f"{args.cpu_chains} chains of the same 32x32 M=10 workload, 10 MC steps after 1 warm-up, "
                         f"{per_step:.3f} s/step; oracle restatement of the TF graph (TF 1.x not installable)"}
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    K = r["ctx"]["K"]
    launches = (K + r["ctx"]["W"]) + K * M + 4 + (0 if e2e is None else (K + 2) * e2e["launches_per_step"])
    line = {
        "metric": METRIC, "value": r["value"], "unit": "link-updates/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": r["ctx"]["W"], "ms_per_step": r["step_ms"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, args.chains_per_gpu),
        "clocks": r["clk"], "gpu_launches": launches, "e2e": e2e,
        "roofline": {"bound": "hbm", "kernel": "hmc_kernel<32,32,false> (fused leapfrog step, one launch per LF step)",
                     "achieved": r["lf"]["gbs"], "peak": r["peak"], "peak_kind": r["peak_kind"], "unit": "GB/s",
                     "frac": r["lf"]["gbs"] / r["peak"], "traffic": ncu_traffic("hmc_kernel<32,32,false>"), "algorithmic_bytes": r["lf"]["bytes"],
                     "ms_per_launch": r["lf"]["ms"],
                     "link_updates_per_s": args.chains_per_gpu * D / (r["lf"]["ms"] * 1e-3)},
        "roofline_mc": {"bound": "hbm", "kernel": "hmc_kernel<32,32,true> (one launch = one MC step incl. observables; M LF steps "
                        "on-chip: issue-bound, not HBM-bound)",
                        "achieved": r["mc"]["gbs"], "peak": r["peak"], "unit": "GB/s",
                        "frac": r["mc"]["gbs"] / r["peak"], "algorithmic_bytes": r["mc"]["bytes"],
                        "ms_per_step": r["mc"]["ms"]},
        "l2hmc": l2hmc, "cpu_baseline": cpu, "observables": r["obs"],
    }
    emit(line)


if __name__ == "__main__":
    main()
