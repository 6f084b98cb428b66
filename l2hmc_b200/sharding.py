"""Chain sharding across the GPUs of one box.

Chains are independent (no term of the action, the leapfrog or the accept step couples two
chains), so the batch axis is partitioned contiguously, one process per GPU, with NO data-path
collective.  What keeps an N-rank run identical to the 1-rank run on the concatenated batch is the
random stream: Philox is keyed by (seed, GLOBAL chain index, MC step), and `chain_offset` of
GaugeDynamics is the global index of the rank's first chain.

The only collectives are O(1) observable accumulators (all_reduce SUM) -- the reference does not
even do that: under Horovod rank 0 logs its own batch only (gauge_model.py:383-385, 1244-1246) --
and, for training, the gradient all-reduce / initial weight broadcast of
hvd.DistributedOptimizer / BroadcastGlobalVariablesHook (gauge_model.py:941-943, 1008).
Works on any torch.distributed backend (nccl on the GPUs, gloo in the CPU tests).
"""
import torch
import torch.distributed as dist


def world():
    """(rank, world_size) of the default process group, (0, 1) when not initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(total_chains, rank=None, world_size=None):
    """Contiguous [lo, hi) block of chains owned by `rank`; the first `total % world` ranks take one
    extra chain.  lo is the rank's `chain_offset`."""
    if rank is None or world_size is None:
        rank, world_size = world()
    if not 0 <= rank < world_size:
        raise ValueError(f"rank {rank} outside world of {world_size}")
    base, extra = divmod(int(total_chains), int(world_size))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard(batch, rank=None, world_size=None):
    """The rank's rows of a (N, ...) array / tensor and its chain offset."""
    lo, hi = shard_range(len(batch), rank, world_size)
    return batch[lo:hi], lo


class ObservableAccumulator(object):
    """Running sums of per-chain observables over MC steps and ranks.

    add() accumulates on the local device in fp64 with no communication; reduce() does ONE
    all_reduce(SUM) of the (K+1)-vector [sum_0 .. sum_{K-1}, count] and returns the global means.
    Mirrors the columns the reference prints per step (gauge_model.py:1104-1110: accept prob,
    action, plaquette, |dQ|) but over ALL ranks' chains."""

    def __init__(self, names, device="cpu"):
        self.names = tuple(names)
        self.sums = torch.zeros(len(self.names) + 1, dtype=torch.float64, device=device)

    def add(self, **per_chain):
        n = None
        for i, name in enumerate(self.names):
            t = per_chain[name]
            n = t.numel() if n is None else n
            if t.numel() != n:
                raise ValueError("per-chain observables must have the same length")
            self.sums[i] += t.to(device=self.sums.device).sum(dtype=torch.float64)
        self.sums[-1] += float(n)

    def reduce(self, group=None):
        total = self.sums.clone()
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(total, op=dist.ReduceOp.SUM, group=group)
        count = float(total[-1])
        means = {name: float(total[i]) / max(count, 1.0) for i, name in enumerate(self.names)}
        means["chains_x_steps"] = count
        return means


def gather_chains(local, total_chains, group=None):
    """All ranks' per-chain rows concatenated in global chain order (final gather of an inference
    run).  Handles uneven shards by padding to the largest one."""
    rank, ws = world()
    if ws == 1:
        return local
    sizes = [shard_range(total_chains, r, ws)[1] - shard_range(total_chains, r, ws)[0] for r in range(ws)]
    pad = max(sizes)
    buf = torch.zeros((pad,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    buf[:local.shape[0]] = local
    out = [torch.empty_like(buf) for _ in range(ws)]
    dist.all_gather(out, buf, group=group)
    return torch.cat([o[:n] for o, n in zip(out, sizes)], dim=0)


def broadcast_weights(tensors, src=0, group=None):
    """Rank `src`'s weights to every rank, in place (hvd.BroadcastGlobalVariablesHook(0),
    gauge_model.py:1008)."""
    if world()[1] > 1:
        for t in tensors:
            dist.broadcast(t, src=src, group=group)


def allreduce_mean_(flat_grad, group=None, divide=True):
    """In-place average of one flat gradient bucket over ranks (hvd.DistributedOptimizer,
    gauge_model.py:941-943), on the current stream.  divide=False leaves the SUM: the caller folds 1 / world size into
    the optimiser step (l2hmc_adam_step's grad_scale) instead of spending a kernel on it."""
    ws = world()[1]
    if ws > 1:
        dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM, group=group)
        if divide:
            flat_grad /= ws
    return flat_grad


def bind_to_gpu_numa(device_index):
    """Pin this process to the CPUs NVML reports as local to its GPU, BEFORE pinned host buffers are allocated: the
    pages then land on the GPU's NUMA node (first touch) and the per-step host<->device copies of every rank stop
    sharing one socket's memory controller.

    Returns a dict that says what happened: {"bound": bool, "cpus": number of CPUs now allowed, "why": text}.  "bound"
    is False -- with the reason -- when NVML / the affinity call is unavailable, when NVML's CPU set does not intersect
    the CPUs this process may use, or when the GPU's set is the whole machine (a single NUMA node: nothing to gain)."""
    import os
    try:
        allowed_now = sorted(os.sched_getaffinity(0))
    except (AttributeError, OSError) as e:
        return {"bound": False, "cpus": None, "why": f"sched_getaffinity unavailable: {e}"}
    try:
        import pynvml
    except ImportError as e:
        return {"bound": False, "cpus": len(allowed_now), "why": f"pynvml not importable: {e}"}
    try:
        import torch
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(device_index).uuid)
        handle = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(handle, words)
    except Exception as e:                       # NVML errors are their own exception hierarchy
        return {"bound": False, "cpus": len(allowed_now), "why": f"NVML affinity query failed: {type(e).__name__}: {e}"}
    cpus = [w * 64 + b for w, bits in enumerate(mask) for b in range(64) if (int(bits) >> b) & 1]
    target = sorted(set(cpus) & set(allowed_now))
    if not target:
        return {"bound": False, "cpus": len(allowed_now),
                "why": f"NVML's {len(cpus)} GPU-local CPUs do not intersect the {len(allowed_now)} CPUs this process may use"}
    if target == allowed_now:
        return {"bound": False, "cpus": len(allowed_now),
                "why": f"the GPU's local CPU set is all {len(allowed_now)} allowed CPUs (one NUMA node): nothing to bind"}
    try:
        os.sched_setaffinity(0, target)
    except OSError as e:
        return {"bound": False, "cpus": len(allowed_now), "why": f"sched_setaffinity failed: {e}"}
    return {"bound": True, "cpus": len(target), "why": f"bound to the {len(target)} CPUs NVML reports local to GPU {device_index}"}
