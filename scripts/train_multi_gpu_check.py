#!/usr/bin/env python
"""SURVEY 8(e), training: G ranks (one per GPU, NCCL) each take a contiguous shard of the batch; after the gradient
all-reduce (average) the flat gradient must equal the 1-GPU gradient of the concatenated batch, and the weights after
Adam must be identical on every rank.  Launch: python -m torch.distributed.run --nproc-per-node G --master-addr 127.0.0.1
scripts/train_multi_gpu_check.py"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist


def main():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from l2hmc_b200.gauge_model import GaugeModel
    from l2hmc_b200.sharding import allreduce_mean_, broadcast_weights
    T = X = 8
    n_local, M = 32, 3
    N = n_local * world
    D = 2 * T * X
    rng = np.random.default_rng(123)                     # the same global batch and draws on every rank
    x = rng.uniform(0, 2 * np.pi, (N, D)).astype(np.float32)
    z = rng.standard_normal((N, D)).astype(np.float32)
    dr = lambda: (rng.standard_normal((N, D)).astype(np.float32), rng.standard_normal((N, D)).astype(np.float32),
                  rng.uniform(size=N).astype(np.float32), rng.uniform(size=N).astype(np.float32))
    draws_x, draws_z = dr(), dr()

    def build(n):
        np.random.seed(42)
        m = GaugeModel(dict(time_size=T, space_size=X, num_samples=n, num_steps=M, eps=0.2, network_arch="generic"))
        return m
    sl = slice(rank * n_local, (rank + 1) * n_local)
    cut = lambda d: tuple(a[sl] for a in d)
    model = build(n_local)
    broadcast_weights([model.engine.params.flat])
    model.engine.params.sync_to_dynamics()
    loss, grads, _, _, _ = model._calc_loss_and_grads(x[sl], 2.5, draws_x=cut(draws_x), z=z[sl], draws_z=cut(draws_z))
    g = model.engine.params.grad
    allreduce_mean_(g)
    out = {"world": world, "rank": rank}
    if rank == 0:
        full = build(N)
        full.engine.params.flat.copy_(model.engine.params.flat)
        full.engine.params.sync_to_dynamics()
        loss_full, _, _, _, _ = full._calc_loss_and_grads(x, 2.5, draws_x=draws_x, z=z, draws_z=draws_z)
        gf = full.engine.params.grad
        out["grad_rel_err_vs_1gpu"] = float((g - gf).abs().max() / gf.abs().max())
        out["loss_local"], out["loss_full"] = float(loss), float(loss_full)
        out["grad_max"] = float(gf.abs().max())
        out["px_mean"] = float(full.last_forward["px"].mean())
    model._optimizer.step(1e-3)
    flat = model.engine.params.flat
    gathered = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    out["weights_identical_across_ranks"] = all(torch.equal(gathered[0], t) for t in gathered)
    # a few free-running distributed steps through the public loop
    model.dynamics.chain_offset = rank * n_local
    samples, hist = model.train(4 * world, samples_init=x[sl], distributed=True)
    dist.all_gather(gathered, model.engine.params.flat)
    out["weights_identical_after_train_loop"] = all(torch.equal(gathered[0], t) for t in gathered)
    out["losses"] = [float(l) for l in hist]
    if rank == 0:
        print(json.dumps(out), flush=True)
        assert out["grad_rel_err_vs_1gpu"] < 1e-5, out
    assert out["weights_identical_across_ranks"] and out["weights_identical_after_train_loop"], out
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
