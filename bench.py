#!/usr/bin/env python
"""bench.py -- link-updates/s of the U(1) leapfrog path on N B200s.

Workload (BASELINE.json metric / config 4): 32x32 lattice, 65 536 chains PER GPU (weak
scaling; the 1 GB of links+momenta per GPU is ~8x the 126 MB L2), plain HMC, M = 10 leapfrog
steps, eps = 0.1, beta = 4, hot start, free-running Philox momenta.  A "step" is one MC step
(GaugeDynamics.apply_transition + observables) = N*D*M link-updates.

  value        device-resident throughput of the product path (one-launch MC step)
  e2e          same through GaugeSampler.step with pinned HOST buffers (H2D + D2H of the links inside, every step),
               next to the measured copy-only ceiling of the same chunked transfer pattern, and -- clearly labelled,
               never instead -- the resident run loop (GaugeSampler.run: links stay on the GPU, observables come back)
  roofline     the fused leapfrog-STEP kernel (x,v streamed once per LF step, 16 B/link-update),
               the kernel the north-star's >= 70 % HBM target names, timed live in its own loop
  roofline_mc  the one-launch MC-step kernel against its real bound: instruction issue
  l2hmc        the legs with S/T/Q networks (configs 2-4), config 1, config 5 (MoG), training steps; each with the CPU
               oracle timed beside it at N = 1
  cpu_baseline the oracle restatement of the reference's TF graph on the host cores (config-4 HMC sample)

`--impl reference` times that CPU restatement alone (TensorFlow 1.x cannot be installed here,
DESIGN.md section 6), rank 0 only; its `config` says how many chains it actually integrates.
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


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f), "measured (MEASURED_PEAKS.json)"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1366.0}, "fallback (B200_PROFILING.md)"


def ncu_static(kernel, key):
    """Per-launch counter of `kernel` from the committed ncu capture of this round's tree
    (profiles/r2_traffic.json, else round 1's): STATIC evidence, not measured in this run."""
    for name in ("r2_traffic.json", "r1_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                d = json.load(f)
            v = d[kernel]
            if isinstance(v, dict):
                return float(v[key]), "static, from profiles/" + name
            if key == "dram_bytes":
                return float(v), "static, from profiles/" + name
        except (OSError, KeyError, ValueError, TypeError):
            continue
    return None, None


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


# =============================================================================================== CPU oracle legs
# Only the functions of this section touch oracle/ (the checker / timed CPU baseline; never the product path).
def _best_of(fn, warm=1, reps=3):
    for _ in range(warm):
        fn()
    best = float("inf")
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        best = min(best, time.perf_counter() - t0)
    return best


def cpu_hmc_run(steps, warmup, chains, T_=T, M_=M, eps=EPS, beta=BETA):
    """The oracle restatement of the reference's TF graph (apply_transition: BOTH directions
    integrated, one discarded, exactly as gauge_dynamics.py:211-241) on the host cores.
    Returns (link-updates/s, s/step, threads)."""
    import numpy as np
    import torch
    from oracle import u1_oracle as orc
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    D_ = 2 * T_ * T_
    rng = np.random.default_rng(0)
    dyn = orc.OracleDynamics(T_, T_, M_, eps, hmc=True, mask=np.zeros((M_, D_), np.float32))
    x = rng.uniform(0, 2 * np.pi, (chains, D_)).astype(np.float32)
    times = []
    with torch.no_grad():
        for i in range(warmup + steps):
            vf = rng.standard_normal((chains, D_)).astype(np.float32)
            vb = rng.standard_normal((chains, D_)).astype(np.float32)
            ud, ua = rng.uniform(size=chains).astype(np.float32), rng.uniform(size=chains).astype(np.float32)
            t0 = time.perf_counter()
            _, _, _, xo = dyn.apply_transition(x, beta, vf, vb, ud, ua)
            orc.plaq_observables(xo, T_, T_)
            x = orc.wrap_2pi(xo.numpy())
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt)
    per_step = sum(times) / len(times)
    return chains * D_ * M_ / per_step, per_step, threads


def cpu_l2hmc_leg(arch, T_, chains, M_, eps, beta, weights, mask):
    """One L2HMC MC step of the oracle (both directions, as the reference) with the GPU leg's weights and masks."""
    import numpy as np
    import torch
    from oracle import u1_oracle as orc
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    D_ = 2 * T_ * T_
    rng = np.random.default_rng(1)
    od = orc.OracleDynamics(T_, T_, M_, eps, hmc=False, network_arch=arch, mask=mask, xnet=weights["XNet"],
                            vnet=weights["VNet"])
    x = rng.uniform(0, 2 * np.pi, (chains, D_)).astype(np.float32)
    vf, vb = (rng.standard_normal((chains, D_)).astype(np.float32) for _ in range(2))
    ud, ua = (rng.uniform(size=chains).astype(np.float32) for _ in range(2))

    def step():
        with torch.no_grad():
            xo = od.apply_transition(x, beta, vf, vb, ud, ua)[3]
            orc.plaq_observables(xo, T_, T_)
    s = _best_of(step)
    return {"value": chains * D_ * M_ / s, "unit": "link-updates/s", "ms_per_step": s * 1e3, "cores": threads,
            "kind": "port", "sample": f"{chains} chains, best of 3 MC steps after 1 warm-up; torch-CPU fp32 oracle of the "
                                      "TF graph (both directions integrated, like the reference), same weights and masks"}


def cpu_mog_leg(chains, L, H):
    import numpy as np
    import torch
    from oracle import mog_oracle as mogo
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    rng = np.random.default_rng(2)
    k = 2
    target = mogo.GMMTarget.from_params(np.array([[1., 0.], [-1., 0.]]), np.stack([0.05 * np.eye(2)] * 2), [0.5, 0.5])
    od = mogo.OracleMogDynamics(target, k, L, 0.1, hmc=False, mask=mogo.init_masks(k, L),
                                xnet=mogo.init_mog_net(k, 2., rng, H), vnet=mogo.init_mog_net(k, 1., rng, H))
    x = (rng.standard_normal((chains, k)) * 0.2 + np.array([1.0, 0.0])).astype(np.float32)
    vf, vb = (rng.standard_normal((chains, k)).astype(np.float32) for _ in range(2))
    dm, ua = (rng.uniform(size=chains) > 0.5).astype(np.float32), rng.uniform(size=chains).astype(np.float32)

    def step():                   # the oracle's GMM force is torch autograd, as the reference's tf.gradients
        od.propose(x, vf, vb, dm, ua)
    s = _best_of(step)
    return {"value": chains / s, "unit": "chain-updates/s", "ms_per_step": s * 1e3, "cores": threads, "kind": "port",
            "sample": f"{chains} chains, best of 3 MC steps; torch-CPU oracle of utils/dynamics.py + sampler.py "
                      "(row-wise GMM energy: the reference's N x N form cannot run at this N)"}


def run_reference(args, rank):
    if rank != 0:
        return
    chains = args.cpu_chains
    value, per_step, threads = cpu_hmc_run(args.steps, args.warmup, chains)
    sample = (f"{chains} chains of the 32x32 M=10 plain-HMC workload per step (a bounded sample: the GPU arm integrates "
              f"{args.chains_per_gpu} chains per GPU per step); torch-CPU fp32 restatement of the TF graph, both "
              "directions integrated like the reference")
    cfg = workload_config(args, chains, gpus=0)
    cfg["momenta"] = "NumPy default_rng normals / uniforms injected every step (the reference draws them in-graph)"
    cfg["parallelism"] = f"{threads} host threads (torch intra-op), no GPU"
    cfg["sample_of"] = {"chains_per_gpu_of_the_gpu_arm": args.chains_per_gpu}
    cfg["cache"] = f"host memory: x, v are {chains * D * 4 >> 20} MiB each per direction"
    emit(({
        "impl": "reference", "metric": METRIC, "value": value, "unit": "link-updates/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": cfg,
        "cpu_baseline": {"value": value, "unit": "link-updates/s", "cores": threads, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": "link-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "TensorFlow 1.x is not installable in this image: the reference arm is the oracle port",
    }))


def workload_config(args, chains_per_gpu, gpus=None):
    gpus = args.gpus if gpus is None else gpus
    return {"workload": "U(1) 32x32 plain-HMC MC step (BASELINE config 4), hot start",
            "lattice": "32x32", "chains_per_gpu": chains_per_gpu, "total_chains": chains_per_gpu * max(gpus, 1),
            "lf_steps": M, "eps": EPS, "beta": BETA, "momenta": "Philox4x32-10 in-kernel",
            "observables": "action, plaquette, integer charge every step",
            "cache": "inputs larger than L2 (x and v are 512 MiB each per GPU vs 126 MB L2)",
            "parallelism": f"chains sharded over {gpus} GPU(s), no data-path collective"}


# =============================================================================================== GPU arm
class Ctx(object):
    """Device, ranks, timing helpers and the launch tally of the GPU arm."""

    def __init__(self, args, rank, local_rank, world):
        import torch
        import torch.distributed as dist
        from l2hmc_b200 import _lib
        self.torch, self.dist, self._lib = torch, dist, _lib
        self.args, self.rank, self.world = args, rank, world
        torch.cuda.set_device(local_rank)
        self.dev = torch.device("cuda", local_rank)
        self.local_rank = local_rank
        from l2hmc_b200.sharding import bind_to_gpu_numa
        self.numa = bind_to_gpu_numa(local_rank) if world > 1 else {"bound": False, "why": "single rank: not bound"}
        if world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        _lib.get_lib()                                   # fail loudly if the CUDA library is missing
        self.launches = {}
        self.no_cpu = bool(args.no_cpu_baseline) or rank != 0 or world != 1

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize(self.dev)

    def max_over_ranks(self, ms):
        if self.world > 1:
            t = self.torch.tensor([ms], dtype=self.torch.float64, device=self.dev)
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
            return float(t.item())
        return ms

    def timed(self, name, fn, iters, warm=2):
        """ms per call of fn(): `warm` untimed calls, barrier + synchronize, CUDA events around `iters` calls on the
        current stream, barrier + synchronize, max over ranks.  The library's launch counter is differenced around the
        timed calls and tallied under `name`."""
        torch = self.torch
        for _ in range(warm):
            fn()
        self.barrier()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        c0 = self._lib.launch_count()
        ev[0].record()
        for _ in range(iters):
            fn()
        ev[1].record()
        self.barrier()
        self.launches[name] = self.launches.get(name, 0) + self._lib.launch_count() - c0
        return self.max_over_ranks(ev[0].elapsed_time(ev[1])) / iters


def measure_tf32_peak(c):
    """kind::tf32 tensor-pipe peak of this GPU, measured the way MEASURED_PEAKS.json measures bf16: a cuBLAS
    8192^3 fp32 matmul with TF32 allowed (2 N^3 flop), best of 10 (burst) and back to back for ~1.5 s (sustained)."""
    torch = c.torch
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        n = 8192
        a = torch.randn(n, n, device=c.dev)
        b = torch.randn(n, n, device=c.dev)
        out = torch.empty(n, n, device=c.dev)
        for _ in range(3):
            torch.matmul(a, b, out=out)
        torch.cuda.synchronize(c.dev)
        flop = 2.0 * n ** 3
        best = float("inf")
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            torch.matmul(a, b, out=out)
            e1.record()
            torch.cuda.synchronize(c.dev)
            best = min(best, e0.elapsed_time(e1))
        reps = max(10, int(1500.0 / best))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            torch.matmul(a, b, out=out)
        e1.record()
        torch.cuda.synchronize(c.dev)
        sustained = e0.elapsed_time(e1) / reps
        del a, b, out
        torch.cuda.empty_cache()
        return {"tf32_tflops": flop / best / 1e9, "tf32_tflops_sustained": flop / sustained / 1e9,
                "how": "torch.matmul fp32 8192^3 with allow_tf32 (cuBLAS), best of 10 / back to back for ~1.5 s, CUDA events"}
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


def run_hmc(c):
    """Headline: device-resident one-launch MC steps; then the fused LF-step kernel in its own timed loop."""
    import numpy as np
    torch, _lib, dev, world, rank = c.torch, c._lib, c.dev, c.world, c.rank
    from l2hmc_b200 import GaugeDynamics, GaugeLattice
    from l2hmc_b200.sharding import shard_range
    N = c.args.chains_per_gpu
    K, W = c.args.steps, max(c.args.warmup, 3)
    lat = GaugeLattice(T, X, 2, "U1", num_samples=1)
    chain_lo, chain_hi = shard_range(world * N, rank, world)          # contiguous block of the global batch
    assert chain_hi - chain_lo == N
    dyn = GaugeDynamics(lat, hmc=True, num_steps=M, eps=EPS, seed=2024, chain_offset=chain_lo, device=dev)
    gen = torch.Generator(device=dev).manual_seed(1000 + rank)
    x = torch.rand(N, D, device=dev, generator=gen) * (2 * np.pi)      # synthetic hot start
    xa, xb = x.clone(), torch.empty_like(x)
    p = torch.empty(N, device=dev)
    act, plaq = torch.empty(N, device=dev), torch.empty(N, device=dev)
    q = torch.empty(N, dtype=torch.int32, device=dev)
    sums = torch.zeros(4, dtype=torch.float64, device=dev)
    lib = _lib.get_lib()
    state = [xa, xb]

    def mc_step():
        src, dst = state
        a = _lib.HmcMcArgs()
        a.x, a.x_out, a.accept_prob = _lib.dptr(src), _lib.dptr(dst), _lib.dptr(p)
        a.beta, a.eps, a.num_steps, a.default_dir, a.wrap_2pi = BETA, EPS, M, 1, 1
        a.seed, a.mc_step, a.chain0 = dyn.seed, dyn._mc_step, dyn.chain_offset
        a.N, a.T, a.X = N, T, X
        a.nonfinite_flag = _lib.dptr(dyn._flag, torch.int32)
        a.obs_action, a.obs_plaq, a.obs_charge = _lib.dptr(act), _lib.dptr(plaq), _lib.dptr(q, torch.int32)
        _lib.check(lib.l2hmc_u1_hmc_mc_step(a, _lib.stream_ptr()))
        dyn._mc_step += 1
        state[0], state[1] = dst, src

    n_chains = torch.tensor(float(N), dtype=torch.float64, device=dev)

    def accumulate():
        return torch.stack([plaq.sum(dtype=torch.float64), act.sum(dtype=torch.float64), p.sum(dtype=torch.float64),
                            n_chains])
    for _ in range(W):
        mc_step()
    accumulate()                                   # (loads torch's reduction kernels outside the timed region)
    if world > 1:
        c.dist.all_reduce(torch.zeros(4, dtype=torch.float64, device=dev))
    clocks = ClockSampler(c.local_rank)
    clocks.start()
    time.sleep(0.3)
    c.barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev_mc = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    c0 = _lib.launch_count()
    ev[0].record()
    for i in range(K):
        ev_mc[i][0].record()
        mc_step()
        ev_mc[i][1].record()
    sums += accumulate()
    if world > 1:
        c.dist.all_reduce(sums)                      # the only collective: O(1) observable accumulators
    ev[1].record()
    c.barrier()
    c.launches["value: K one-launch MC steps"] = _lib.launch_count() - c0
    ms_total = c.max_over_ranks(ev[0].elapsed_time(ev[1]))
    step_ms = ms_total / K
    value = world * N * D * M / (step_ms * 1e-3)
    dyn.raise_if_nonfinite()
    obs = {"avg_plaq": float(sums[0] / sums[3]), "accept_prob": float(sums[2] / sums[3])}

    # ------------------------------------------------ the fused LF-step kernel, its own timed loop
    v = torch.randn(N, D, device=dev, generator=gen)
    x1, v1 = torch.empty_like(x), torch.empty_like(v)
    bufs = [(state[0], v), (x1, v1)]
    it = [0]

    def lf_launch():
        (sx, sv), (dx, dv) = bufs[it[0] % 2], bufs[(it[0] + 1) % 2]
        _lib.check(lib.l2hmc_u1_hmc_step(_lib.dptr(sx), _lib.dptr(sv), BETA, EPS, None, N, T, X, _lib.dptr(dx),
                                         _lib.dptr(dv), _lib.stream_ptr()))
        it[0] += 1
    lf_ms = c.timed("roofline: K*M fused leapfrog-step launches", lf_launch, K * M, warm=4)
    clk = clocks.stop()
    peaks, peak_kind = load_peaks()
    lf_bytes = 16.0 * N * D                           # read x, v; write x, v
    lf_gbs = lf_bytes / (lf_ms * 1e-3) / 1e9
    mc_kernel_ms = sum(a.elapsed_time(b) for a, b in ev_mc) / K
    c.dyn, c.x, c.N, c.K, c.W = dyn, state[0], N, K, W
    return dict(value=value, step_ms=step_ms, clk=clk, obs=obs, peak=float(peaks["hbm_gbs"]), peak_kind=peak_kind,
                lf=dict(ms=lf_ms, gbs=lf_gbs, bytes=lf_bytes), mc=dict(ms=mc_kernel_ms, bytes=8.0 * N * D))


def copy_ceiling(c, N, chunk, iters=5):
    """The host-link ceiling of GaugeSampler.step's transfer pattern: the same pinned buffers, the same chunking over
    three streams, H2D of every chunk and D2H of every chunk -- and NO kernel.  ms per 'step' of N chains."""
    torch, dev = c.torch, c.dev
    h_in = torch.empty(N, D, dtype=torch.float32).pin_memory()
    h_out = torch.empty(N, D, dtype=torch.float32).pin_memory()
    h_in.zero_()
    nbuf = 3
    d = [torch.empty(chunk, D, device=dev) for _ in range(nbuf)]
    streams = [torch.cuda.Stream(device=dev) for _ in range(nbuf)]
    main = torch.cuda.current_stream(dev)

    def step():
        for i, lo in enumerate(range(0, N, chunk)):
            hi = min(lo + chunk, N)
            b = i % nbuf
            streams[b].wait_stream(main)
            with torch.cuda.stream(streams[b]):
                d[b][:hi - lo].copy_(h_in[lo:hi], non_blocking=True)
                h_out[lo:hi].copy_(d[b][:hi - lo], non_blocking=True)
        for st in streams:
            main.wait_stream(st)
        main.synchronize()
    ms = c.timed("copy ceiling (no kernels)", step, iters, warm=2)
    return {"ms_per_step": ms, "GBps_each_way": N * D * 4 / ms / 1e6,
            "what": f"pinned H2D || D2H of {N * D * 4 >> 20} MiB each way in {chunk}-chain chunks over 3 streams, no kernel"}


def run_e2e(c, hmc):
    """Same MC step through GaugeSampler.step: pinned host links in, host results out, every step."""
    torch, N, K = c.torch, c.N, c.K
    from l2hmc_b200.sampler import GaugeSampler
    chunk = c.args.e2e_chunk
    sampler = GaugeSampler(c.dyn, N, chunk_chains=chunk)
    host = [torch.empty(N, D, dtype=torch.float32).pin_memory()]
    host[0].copy_(c.x)
    res = [None]

    def step():
        res[0] = sampler.step(host[0], BETA)       # H2D of the links, MC step, observables, D2H of everything
        host[0], sampler.h_out = res[0].samples, host[0]   # next step feeds what this one returned (no host copy)
    step_ms = c.timed("e2e: K host-buffer steps", step, K, warm=2)   # step() ends with a host sync: covers copies + kernels
    out = {"value": c.world * N * D * M / (step_ms * 1e-3), "unit": "link-updates/s",
           "h2d_bytes_per_step": sampler.h2d_bytes, "d2h_bytes_per_step": sampler.d2h_bytes,
           "ms_per_step": step_ms, "chunk_chains": chunk,
           "host_cpu_binding": c.numa,
           "plaq": float(res[0].plaqs.mean()), "charge_diff": res[0].charge_diff}
    ceil = copy_ceiling(c, N, chunk)
    out["copy_ceiling"] = ceil
    out["frac_of_copy_ceiling"] = ceil["ms_per_step"] / step_ms
    # the resident loop: links stay on the GPU, per-step observables (px, action, plaquette, charge in / out per chain) come back
    r = [None]

    def resident():
        r[0] = sampler.run(K, BETA, samples=None if r[0] is not None else host[0], return_samples=None)
    ms_run = c.timed("e2e_resident: GaugeSampler.run(K)", resident, 1, warm=1)
    rs_ms = ms_run / K
    out["resident_run"] = {"what": "GaugeSampler.run(K): NOT the per-step round trip above -- links resident on the GPU, "
                                   "only the per-chain observables return to pinned host memory every step",
                           "value": c.world * N * D * M / (rs_ms * 1e-3), "unit": "link-updates/s", "ms_per_step": rs_ms,
                           "d2h_bytes_per_step": 5 * N * 4, "h2d_bytes_per_step": 0,
                           "tunneling_rate_last": float(r[0]["tunneling_rate"][-1]),
                           "frac_of_device_resident_value": hmc["step_ms"] / rs_ms}
    return out


def run_l2hmc(c):
    """The legs with S/T/Q networks, config 1, config 5 and the training steps."""
    import numpy as np
    torch, dev, world, _lib, steps = c.torch, c.dev, c.world, c._lib, c.args.steps
    from l2hmc_b200 import GaugeDynamics, GaugeLattice
    from l2hmc_b200.gauge_model import RunStats
    from l2hmc_b200.sampler import GaugeSampler
    peaks, peak_kind = load_peaks()
    tf32 = measure_tf32_peak(c)
    out = {"tf32_peak": tf32}

    # ---------------------------------------------------------------- config 1: 8x8 plain HMC, batch 64 (CPU-assigned config)
    lat = GaugeLattice(8, 8, 2, "U1", num_samples=1)
    dyn = GaugeDynamics(lat, hmc=True, num_steps=10, eps=0.1, seed=3, device=dev)
    xs = [torch.rand(64, 128, device=dev) * (2 * np.pi)]

    def c1():
        xs[0] = dyn.apply_transition(xs[0], 2.0, wrap=True, observables=True)[3]
    ms = c.timed("config1: 8x8 HMC N=64 MC steps", c1, max(20, steps), warm=3)
    out["config1_8x8_hmc"] = {"chains": 64, "lf_steps": 10, "eps": 0.1, "beta": 2.0, "mc_step_ms": ms,
                              "link_updates_per_s": 64 * 128 * 10 / (ms * 1e-3),
                              "note": "one launch per MC step; latency-bound at this size (64 chains = 64 KB of state)"}
    if not c.no_cpu:
        v, per_step, threads = cpu_hmc_run(steps=20, warmup=2, chains=64, T_=8, M_=10, eps=0.1, beta=2.0)
        out["config1_8x8_hmc"]["cpu_baseline"] = {
            "value": v, "unit": "link-updates/s", "ms_per_step": per_step * 1e3, "cores": threads, "kind": "port",
            "sample": "config 1 in full: 64 chains, 20 MC steps after 2 warm-up; oracle restatement of the TF graph"}

    # ---------------------------------------------------------------- nets: (tag, arch, T, N, M, eps, cpu chains)
    for tag, arch, T_, N, M_, eps, cpu_n in (("config2_8x8_generic", "generic", 8, 128, 5, 0.25, 128),
                                             ("config3_16x16_conv3D", "conv3D", 16, 2048, 5, 0.25, 256),
                                             ("config4_32x32_generic_nets", "generic", 32, 2048, 10, 0.1, 0)):
        D_ = 2 * T_ * T_
        lat = GaugeLattice(T_, T_, 2, "U1", num_samples=1)
        np.random.seed(42)
        dyn = GaugeDynamics(lat, hmc=False, network_arch=arch, num_steps=M_, eps=eps, seed=7,
                            chain_offset=c.dyn.chain_offset // 32, device=dev)
        net = dyn.position_fn
        K_, H_ = net.num_features, net.num_hidden
        gen = torch.Generator(device=dev).manual_seed(5)
        x = torch.rand(N, D_, device=dev, generator=gen) * (2 * np.pi)
        v = torch.randn(N, D_, device=dev, generator=gen)
        entry = {"chains_per_gpu": N, "K": K_, "H": H_, "D": D_, "lf_steps": M_}
        if N >= 256:           # the tcgen05 path: tensor-pipe roofline of one position_fn call
            ld = torch.zeros(N, device=dev)
            flop = 2.0 * N * (2 * K_ * H_ + 2 * H_ + H_ * H_ + 3 * H_ * D_)
            feats = net._features(v, x, dyn.mask[0], False)       # conv features (or the inputs) outside the GEMM timing
            call = _lib.NetCall()
            call.mode, call.eps, call.dir = _lib.NET_POSITION, eps, 1
            call.in1_mask, call.in1_mask_invert = _lib.dptr(feats[2]), 0
            x_new, mask0 = torch.empty_like(x), dyn.mask[0].contiguous()
            call.x, call.v, call.mask, call.out = _lib.dptr(x), _lib.dptr(v), _lib.dptr(mask0), _lib.dptr(x_new)
            call.logdet, call.nonfinite_flag = _lib.dptr(ld), _lib.dptr(dyn._flag, torch.int32)
            ms = c.timed(f"{tag}: position_fn calls", lambda: net._run(call, feats[0], feats[1], N, (1.0, 0.0)),
                         max(5, steps // 2))
            tf = flop / ms / 1e9
            entry["net_call"] = {"what": "3 tcgen05 GEMMs (two-input layer, hidden layer, 3 heads with the position "
                                         "update + log-det in the epilogue) + operand converters, one position_fn call",
                                 "ms": ms, "gflop_algorithmic": flop / 1e9}
            entry["roofline"] = {"bound": "tensor", "achieved": tf, "unit": "TFLOP/s", "tensor_pipe_tflops": 3 * tf,
                                 "peak": tf32["tf32_tflops_sustained"], "peak_burst": tf32["tf32_tflops"],
                                 "peak_kind": "measured in this run: cuBLAS TF32 8192^3, sustained (kernel timed inside a step)",
                                 "frac": 3 * tf / tf32["tf32_tflops_sustained"], "frac_of_burst": 3 * tf / tf32["tf32_tflops"],
                                 "frac_of_half_bf16_sustained": 3 * tf / (float(peaks["bf16_tflops_sustained"]) / 2),
                                 "note": "achieved = algorithmic fp32 flops / time; every product is 3 TF32 MMAs "
                                         "(hi.lo + lo.hi + hi.hi), so the tensor pipe runs at 3x that rate"}
        x_mc = [x]

        def mc():
            x_mc[0] = dyn.apply_transition(x_mc[0], 3.0, wrap=True, observables=True)[3]
        step_ms = c.timed(f"{tag}: MC steps", mc, max(3, steps // 4) if T_ >= 32 else max(10, steps),
                          warm=1 if T_ >= 32 else 2)
        entry["mc_step_ms"] = step_ms
        entry["link_updates_per_s"] = world * N * D_ * M_ / (step_ms * 1e-3)
        dyn.raise_if_nonfinite()
        if T_ < 32:            # the same steps as replays of ONE captured CUDA graph (GaugeSampler.run(graph=True))
            smp = GaugeSampler(dyn, N)
            smp.run(3, 3.0, samples=x_mc[0], graph=True, return_samples=None)       # eager step + capture of both directions
            n_rep = max(20, steps)
            g_ms = c.timed(f"{tag}: MC steps, graph replay", lambda: smp.run(n_rep, 3.0, graph=True, return_samples=None),
                           1, warm=1) / n_rep
            entry["graph_replay"] = {"mc_step_ms": g_ms, "link_updates_per_s": world * N * D_ * M_ / (g_ms * 1e-3),
                                     "speedup_vs_eager": step_ms / g_ms,
                                     "what": "GaugeSampler.run: resident links, per-chain observables to the host every "
                                             "step, the step's launches replayed from a CUDA graph (bit-identical)"}
            dyn.raise_if_nonfinite()
            del smp
        if cpu_n and not c.no_cpu:
            ws = {"XNet": dyn.position_fn.state_dict(), "VNet": dyn.momentum_fn.state_dict()}
            entry["cpu_baseline"] = cpu_l2hmc_leg(arch, T_, cpu_n, M_, eps, 3.0, ws, dyn.mask.cpu().numpy())
        if tag.startswith("config4"):
            # an untrained H = 8192 net with the reference's initial head scale moves every link enough that nothing is
            # accepted (measured: <p> = 6e-20); the run uses heads 100x smaller -- synthetic weights either way
            for net_ in (dyn.position_fn, dyn.momentum_fn):
                for nm in ("scale_layer", "translation_layer", "transformation_layer"):
                    net_.weights[nm + ".kernel"].mul_(0.01)
                net_._struct = None                      # re-pack the tensor-core copy of the weights
            entry["run_8192_chains"] = run_config4_nets(c, dyn, GaugeSampler, RunStats)
        out[tag] = entry
        del dyn, net, x, v, x_mc
        torch.cuda.empty_cache()
    out["config5_mog"] = run_mog(c)
    out["training"] = run_training(c)
    return out


def run_config4_nets(c, dyn, GaugeSampler, RunStats):
    """BASELINE config 4 as written: 32x32 L2HMC WITH the (generic) S/T/Q nets, 8192 chains per GPU (65 536 over 8),
    M = 10, >= 20 MC steps with the tunneling-rate and topological-charge statistics of the run
    (gauge_model.py:1379-1388, 1473-1533), device-resident; then the same step through the host-buffer sampler."""
    import numpy as np
    torch, dev, world = c.torch, c.dev, c.world
    N, D_, M_ = 8192, 2048, 10
    n_steps = max(20, c.args.steps)
    dyn.chain_offset = c.rank * N
    sampler = GaugeSampler(dyn, N, chunk_chains=2048)
    gen = torch.Generator(device=dev).manual_seed(77 + c.rank)
    x0 = torch.rand(N, D_, device=dev, generator=gen) * (2 * np.pi)
    sampler.run(2, BETA, samples=x0, return_samples=None)                       # warm-up (2 MC steps)
    stats = RunStats(N, n_steps, therm_frac=10, device=dev)
    res = [None]

    def run():
        res[0] = sampler.run(n_steps, BETA, stats=stats, return_samples="host")
    ms = c.timed("config4 nets: GaugeSampler.run(n_steps), 8192 chains", run, 1, warm=0) / n_steps
    r = res[0]
    (a_m, a_e), (p_m, p_e), (q_m, q_e), (s_m, s_e), probs = stats.finalize()
    dyn.raise_if_nonfinite()
    # all ranks' O(1) accumulators: [sum |dQ|, sum px, sum plaq, chains x steps]
    acc = torch.tensor([float(r["charge_diff"].sum()), float(r["px"].sum()), float(r["plaqs"].sum()),
                        float(n_steps * N)], dtype=torch.float64, device=dev)
    if world > 1:
        c.dist.all_reduce(acc)
    entry = {"chains_per_gpu": N, "total_chains": N * world, "lf_steps": M_, "mc_steps": n_steps, "beta": BETA,
             "mc_step_ms": ms, "link_updates_per_s": world * N * D_ * M_ / (ms * 1e-3),
             "tunneling_rate": float(acc[0] / acc[3]), "accept_prob": float(acc[1] / acc[3]),
             "avg_plaq": float(acc[2] / acc[3]),
             "rank0_run_stats": {"therm_steps": stats.therm_steps, "plaq_mean": float(np.mean(p_m)),
                                 "plaq_err_mean": float(np.mean(p_e)), "charge_mean": float(np.mean(q_m)),
                                 "suscept_mean": float(np.mean(s_m)),
                                 "charge_probabilities": {str(k): float(v) for k, v in probs.items()}},
             "note": "random-init nets (the reference's initialisers, head kernels x 0.01), hot start, free-running Philox draws; "
                     "statistics streamed on the device (csrc/stats.cu), per-chain observables fetched every step"}
    # the same step through the host-buffer sampler: H2D links, one C call per 2048-chain chunk, D2H links + observables
    host = [r["samples"].clone().pin_memory()]

    def step():
        out = sampler.step(host[0], BETA)
        host[0], sampler.h_out = out.samples, host[0]
    e_ms = c.timed("config4 nets: host-buffer steps", step, 3, warm=1)
    entry["e2e"] = {"value": world * N * D_ * M_ / (e_ms * 1e-3), "unit": "link-updates/s", "ms_per_step": e_ms,
                    "h2d_bytes_per_step": sampler.h2d_bytes, "d2h_bytes_per_step": sampler.d2h_bytes,
                    "chunk_chains": sampler.chunk, "frac_of_device_resident": ms / e_ms}
    del sampler
    torch.cuda.empty_cache()
    return entry


def run_training(c):
    """Training step at BASELINE config 2 (8x8, generic S/T/Q MLPs H = 4D = 512, batch 128, M = 5)
    and config 3 (16x16, conv3D F = 16, H = 1024, 2048 chains per GPU, M = 5): loss over the x- and z-chain with tape,
    explicit backward pass (one C call, csrc/train_step.cu), [gradient all-reduce at N > 1], Adam, observables.
    Rank 0 at N = 1 also times the CPU oracle (torch autograd through the oracle's transition) on the config-2 step."""
    import numpy as np
    from l2hmc_b200.gauge_model import GaugeModel
    torch, dev, world, steps = c.torch, c.dev, c.world, c.args.steps
    out = {}
    for tag, prm in (("config2_8x8_generic", dict(time_size=8, space_size=8, num_samples=128, network_arch="generic")),
                     ("config3_16x16_conv3D", dict(time_size=16, space_size=16, num_samples=2048, network_arch="conv3D"))):
        np.random.seed(42)
        model = GaugeModel(dict(prm, num_steps=5, eps=0.2, rand=True, train_steps=10000), device=dev)
        N, D_, M_ = prm["num_samples"], 2 * prm["time_size"] ** 2, 5
        model.dynamics.seed, model.dynamics.chain_offset = 13, c.rank * N
        model.using_hvd = world > 1
        eng = model.engine
        if world > 1:
            from l2hmc_b200.sharding import broadcast_weights
            broadcast_weights([eng.params.flat])
            eng.params.sync_to_dynamics()
        state = [model._x(model.lattice.samples).contiguous()]

        def step():
            state[0] = model.train_step(state[0], 2.0)[1]
        ms = c.timed(f"training {tag}: train steps", step, max(5, steps // 2), warm=2)
        entry = {"chains_per_gpu": N, "H": int(model.dynamics.position_fn.num_hidden), "D": D_, "lf_steps": M_,
                 "params": int(eng.params.size), "train_step_ms": ms,
                 "link_updates_per_s": world * N * D_ * M_ / (ms * 1e-3),
                 "note": "x- and z-chain forward with tape + backward + Adam + observables; the x-chain's advance is "
                         "counted once; fp32: tcgen05 3xTF32 layers / backward products from 1024 stacked rows, chunked "
                         "FFMA kernel (K chunks over a thread-block cluster) below; dependent kernels launched as "
                         "programmatic dependents"}
        if tag.startswith("config2") and not c.no_cpu:
            from oracle import train_oracle as tro
            from oracle import u1_oracle as orc
            torch.set_num_threads(os.cpu_count())
            dyn = model.dynamics
            od = orc.OracleDynamics(8, 8, M_, dyn._eps(), hmc=False, network_arch="generic", mask=dyn.mask.cpu().numpy(),
                                    xnet={k: v.cpu().numpy() for k, v in dyn.position_fn.weights.items()},
                                    vnet={k: v.cpu().numpy() for k, v in dyn.momentum_fn.weights.items()})
            rng = np.random.default_rng(3)
            dr = lambda: (rng.standard_normal((N, D_)).astype(np.float32), rng.standard_normal((N, D_)).astype(np.float32),
                          rng.uniform(size=N).astype(np.float32), rng.uniform(size=N).astype(np.float32))
            a = (od, state[0].cpu().numpy(), 2.0, dr(), rng.standard_normal((N, D_)).astype(np.float32), dr())
            cpu_s = _best_of(lambda: tro.loss_and_grads(*a, **model.loss_weights))
            entry["cpu_baseline"] = {"value": N * D_ * M_ / cpu_s, "unit": "link-updates/s", "ms_per_step": cpu_s * 1e3,
                                     "cores": os.cpu_count(), "kind": "port",
                                     "sample": "best of 3 loss-and-gradient evaluations of the same config-2 batch (torch "
                                               "autograd through the oracle's transition; no optimiser step)"}
        out[tag] = entry
        del model, eng
        torch.cuda.empty_cache()
    return out


def run_mog(c):
    """BASELINE config 5: two-mode Gaussian mixture, k = 2, M = 2 leapfrog steps, H = 50, 2^20 chains
    per GPU; one launch of mog_mc_kernel = one MC step (both nets evaluated in shared memory)."""
    import numpy as np
    from l2hmc_b200 import mog
    torch, dev, world = c.torch, c.dev, c.world
    N, k, L, H = 1 << 20, 2, 2, 50
    gmm = mog.GMM([[1., 0.], [-1., 0.]], [0.05 * np.eye(2)] * 2, [0.5, 0.5], device=dev)
    np.random.seed(42)
    dyn = mog.Dynamics(k, gmm, trajectory_length=L, eps=0.1, seed=11, device=dev, chain_offset=c.rank * N,
                       net_factory=lambda d, scope, factor: mog.network(d, scope, factor, num_nodes=H, device=dev))
    gen = torch.Generator(device=dev).manual_seed(3)
    x = [torch.randn(N, k, device=dev, generator=gen) * 0.2 + torch.tensor([1.0, 0.0], device=dev)]

    def mc():
        x[0] = dyn.mc_step(x[0])["x_out"]
    ms = c.timed("config5 MoG: MC steps", mc, max(5, c.args.steps))
    # per chain per step: 3 V-net + 2 X-net calls per LF step, each 2(kH + kH + 2H + HH + 3Hk) flops
    flop = N * L * 5 * 2.0 * (2 * k * H + 2 * H + H * H + 3 * H * k)
    entry = {"chains_per_gpu": N, "k": k, "lf_steps": L, "H": H, "mc_step_ms": ms,
             "chain_updates_per_s": world * N / (ms * 1e-3), "lf_steps_per_s": world * N * L / (ms * 1e-3),
             "ffma_tflops": flop / ms / 1e9,
             "algorithmic_bytes": N * (4 * k * 4 + 4 + 1 + 4), "hbm_gbs": N * (4 * k * 4 + 9) / ms / 1e6,
             "note": "compute-bound: 50x50 hidden layer per net call per chain on the FFMA pipe, nets in shared memory"}
    if not c.no_cpu:
        entry["cpu_baseline"] = cpu_mog_leg(10000, L, H)
    return entry


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--chains-per-gpu", type=int, default=65536)
    ap.add_argument("--cpu-chains", type=int, default=8192)
    ap.add_argument("--e2e-chunk", type=int, default=8192)
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
    c = Ctx(args, rank, local_rank, world)
    r = run_hmc(c)
    e2e = None if args.no_e2e else run_e2e(c, r)
    l2hmc = None if args.no_l2hmc else run_l2hmc(c)
    cpu = None
    if not c.no_cpu:
        v, per_step, threads = cpu_hmc_run(steps=4, warmup=1, chains=args.cpu_chains)
        cpu = {"value": v, "unit": "link-updates/s", "cores": threads, "kind": "port",
               "sample": f"{args.cpu_chains} chains of the same 32x32 M=10 workload, 4 MC steps after 1 warm-up, "
                         f"{per_step:.3f} s/step; oracle restatement of the TF graph (TF 1.x not installable)"}
    if world > 1:
        c.dist.barrier()
        c.dist.destroy_process_group()
    if rank != 0:
        return
    traffic, traffic_src = ncu_static("hmc_kernel<32,32,false>", "dram_bytes")
    inst, inst_src = ncu_static("hmc_kernel<32,32,true>", "inst_executed")
    sms, _ = c._lib.device_info()
    clk_mhz = r["clk"]["sm_mhz"] or r["clk"]["sm_max_mhz"] or 1965.0
    issue_peak = 4.0 * sms * clk_mhz * 1e6                    # warp instructions / s: 4 schedulers per SM, 1 per clock
    roofline_mc = {"bound": "issue", "kernel": "hmc_kernel<32,32,true> (one launch = one MC step incl. observables; the M "
                   "leapfrog steps run on-chip, so it is bound by instruction issue, not by HBM)",
                   "ms_per_step": r["mc"]["ms"], "algorithmic_bytes": r["mc"]["bytes"],
                   "hbm_gbs": r["mc"]["bytes"] / (r["mc"]["ms"] * 1e-3) / 1e9,
                   "hbm_frac": r["mc"]["bytes"] / (r["mc"]["ms"] * 1e-3) / 1e9 / r["peak"]}
    if inst is not None:
        inst_launch = inst * c.N / 65536.0                    # the capture ran 65 536 chains
        ach = inst_launch / (r["mc"]["ms"] * 1e-3)
        roofline_mc.update({"achieved": ach, "peak": issue_peak, "unit": "warp-instructions/s", "frac": ach / issue_peak,
                            "warp_instructions_per_launch": inst_launch, "instructions_source": inst_src,
                            "thread_instructions_per_link_update": 32.0 * inst_launch / (c.N * D * M),
                            "peak_kind": f"4 schedulers x {sms} SMs x {clk_mhz:.0f} MHz (median SM clock sampled in this run)"})
    line = {
        "metric": METRIC, "value": r["value"], "unit": "link-updates/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": c.W, "ms_per_step": r["step_ms"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, args.chains_per_gpu),
        "clocks": r["clk"],
        "gpu_launches": c.launches["value: K one-launch MC steps"],
        "gpu_launches_source": "l2hmc_launch_count() of the C library differenced around the timed region of `value` "
                               "(rank 0); the other timed regions are in gpu_launches_by_region",
        "gpu_launches_by_region": c.launches,
        "e2e": e2e,
        "roofline": {"bound": "hbm", "kernel": "hmc_kernel<32,32,false> (fused leapfrog step, one launch per LF step)",
                     "achieved": r["lf"]["gbs"], "peak": r["peak"], "peak_kind": r["peak_kind"], "unit": "GB/s",
                     "frac": r["lf"]["gbs"] / r["peak"], "traffic": traffic, "traffic_source": traffic_src,
                     "algorithmic_bytes": r["lf"]["bytes"], "ms_per_launch": r["lf"]["ms"],
                     "link_updates_per_s": args.chains_per_gpu * D / (r["lf"]["ms"] * 1e-3)},
        "roofline_mc": roofline_mc,
        "l2hmc": l2hmc, "cpu_baseline": cpu, "observables": r["obs"],
    }
    emit(line)


if __name__ == "__main__":
    main()
