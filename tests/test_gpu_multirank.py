"""The CUDA path on 2 ranks (one process per GPU, NCCL) against the 1-rank run on the concatenated batch
(reference semantics: hvd.DistributedOptimizer / BroadcastGlobalVariablesHook, gauge_model.py:941-943,
1008).  Skipped when the box has fewer than 2 GPUs.

* inference: HMC and L2HMC MC steps on the free-running Philox stream are BIT-IDENTICAL to the 1-rank run (the stream
  is keyed by the global chain index; chains never communicate);
* training through the public GaugeModel.train(distributed=True), WITHOUT touching chain_offset by hand: the ranks
  draw different momenta (their losses differ), share rank 0's masks and weights, and stay identical after the steps;
* the averaged gradient equals the 1-rank gradient of the concatenated batch to 1e-6 of its scale."""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

WORLD = 2


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _weights(arch, T, X):
    from parity_util import trained_like_weights
    return trained_like_weights(arch, T, X, np.random.default_rng(5), gain=1.0)


def _mc_run(pkg, arch, T, X, x, chain_offset, steps, beta, ws):
    """`steps` free-running MC steps (wrap + observables) of the chains in x, whose first global index is chain_offset."""
    lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=x.shape[0])
    np.random.seed(42)
    dyn = pkg.GaugeDynamics(lat, hmc=arch is None, network_arch=arch or "generic", num_steps=3, eps=0.15, weights=ws,
                            seed=77, chain_offset=chain_offset)
    xs = torch.as_tensor(x).cuda()
    obs = []
    for _ in range(steps):
        _, _, p, xs = dyn.apply_transition(xs, beta, wrap=True, observables=True)
        obs.append(torch.stack([p, dyn.last_observables[0], dyn.last_observables[1],
                                dyn.last_observables[2].to(torch.float32)], dim=1))
    dyn.raise_if_nonfinite()
    return xs, torch.stack(obs)          # (n, D), (steps, n, 4)


def _worker(rank, port, tmp):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    import torch.distributed as dist
    dist.init_process_group("nccl", rank=rank, world_size=WORLD, device_id=torch.device("cuda", rank))
    import l2hmc_b200 as pkg
    from l2hmc_b200 import sharding
    from l2hmc_b200.gauge_model import GaugeModel
    report = {}

    # ------------------------------------------------------------------ inference: bit-identical to 1 rank
    for tag, arch, T, X, total in (("hmc", None, 16, 16, 301), ("generic", "generic", 8, 8, 600), ("conv3D", "conv3D", 8, 8, 141)):
        D = 2 * T * X
        x_all = np.random.default_rng(3).uniform(0, 2 * np.pi, (total, D)).astype(np.float32)
        ws = None if arch is None else _weights(arch, T, X)
        x_loc, lo = sharding.shard(x_all)
        xs, obs = _mc_run(pkg, arch, T, X, x_loc, lo, 3, 2.5, ws)
        full_x = sharding.gather_chains(xs, total)
        full_obs = sharding.gather_chains(obs.permute(1, 0, 2).contiguous(), total)
        if rank == 0:
            ref_x, ref_obs = _mc_run(pkg, arch, T, X, x_all, 0, 3, 2.5, ws)
            report[tag + ".x_equal"] = bool(torch.equal(full_x, ref_x))
            report[tag + ".obs_equal"] = bool(torch.equal(full_obs, ref_obs.permute(1, 0, 2)))
            report[tag + ".accept_rate"] = float(ref_obs[..., 0].mean())

    # ------------------------------------------------------------------ training through the public train()
    T = X = 8
    n, D = 64, 128
    params = dict(time_size=T, space_size=X, num_samples=n, num_steps=2, eps=0.2, network_arch="generic", hmc=False,
                  annealing=False, beta_init=2.5, beta_final=2.5, train_steps=100, clip_grads=True, clip_value=1e6)
    np.random.seed(1000 + rank)                      # DIFFERENT NumPy streams: the masks must come from rank 0
    model = GaugeModel(params)
    own_mask = model.dynamics.mask.clone()
    if rank == 1:                                    # ... and so must the weights
        for t in model.dynamics.position_fn.variables:
            t.mul_(1.5)
    x_all = np.random.default_rng(9).uniform(0, 2 * np.pi, (WORLD * n, D)).astype(np.float32)
    x_loc, lo = sharding.shard(x_all)
    samples, losses = model.train(2 * WORLD, samples_init=x_loc, distributed=True)     # 2 steps per rank
    report_local = dict(chain_offset=int(model.dynamics.chain_offset), n_steps=len(losses))
    flat = model.engine.params.flat
    both = [torch.empty_like(flat) for _ in range(WORLD)]
    dist.all_gather(both, flat)
    masks = [torch.empty_like(own_mask) for _ in range(WORLD)]
    dist.all_gather(masks, model.dynamics.mask)
    own_masks = [torch.empty_like(own_mask) for _ in range(WORLD)]
    dist.all_gather(own_masks, own_mask)
    loss_t = torch.stack([l.reshape(()) for l in losses])
    all_losses = [torch.empty_like(loss_t) for _ in range(WORLD)]
    dist.all_gather(all_losses, loss_t)
    offs = torch.tensor([report_local["chain_offset"]], device="cuda")
    all_offs = [torch.empty_like(offs) for _ in range(WORLD)]
    dist.all_gather(all_offs, offs)

    # ------------------------------------------------------------------ averaged gradient == 1-rank gradient
    mc = int(model.dynamics._mc_step)
    loss, grads, _, _, _ = model._calc_loss_and_grads(samples, 2.5)
    g_avg = model.engine.params.grad.clone()
    g_local_norm = float(g_avg.norm())
    sharding.allreduce_mean_(g_avg)
    full_samples = sharding.gather_chains(samples, WORLD * n)
    # ------------------------------------------------------------------ overlapped first-bucket all-reduce == one all-reduce
    flats = {}
    for overlap in (True, False):
        np.random.seed(2000)
        m = GaugeModel(dict(params, clip_grads=False))
        m.overlap_allreduce = overlap
        m.train(3 * WORLD, samples_init=x_loc, distributed=True)
        flats[overlap] = m.engine.params.flat.clone()
        if overlap:
            used_overlap = getattr(m, "_bucket_event", None) is not None
    both_o = [torch.empty_like(flats[True]) for _ in range(WORLD)]
    dist.all_gather(both_o, flats[True])
    if rank == 0:
        report["overlap.used"] = bool(used_overlap)
        report["overlap.equals_single_allreduce"] = bool(torch.equal(flats[True], flats[False]))
        report["overlap.ranks_equal"] = bool(torch.equal(both_o[0], both_o[1]))
        report["overlap.moved"] = float((flats[True] - flat).abs().max())
        report["train.offsets"] = [int(o) for o in all_offs]
        report["train.steps_per_rank"] = report_local["n_steps"]
        report["train.weights_equal"] = bool(torch.equal(both[0], both[1]))
        report["train.masks_equal"] = bool(torch.equal(masks[0], masks[1])) and bool(torch.equal(masks[0], own_masks[0]))
        report["train.own_masks_differed"] = not bool(torch.equal(own_masks[0], own_masks[1]))
        report["train.losses_differ"] = bool((all_losses[0] != all_losses[1]).all())
        np.random.seed(1000)
        p1 = dict(params, num_samples=WORLD * n)
        one = GaugeModel(p1)
        one.dynamics.set_masks(model.dynamics.mask.cpu().numpy())
        one.engine.params.flat.copy_(flat)
        one.engine.params.sync_to_dynamics()
        one.dynamics._mc_step = mc
        one.dynamics.chain_offset = 0
        _, _, _, _, _ = one._calc_loss_and_grads(full_samples, 2.5)
        g_one = one.engine.params.grad
        scale = float(g_one.abs().max())
        report["grad.max_abs_diff_over_scale"] = float((g_avg - g_one).abs().max()) / scale
        report["grad.scale"] = scale
        report["grad.local_norm"] = g_local_norm
        torch.save(report, os.path.join(tmp, "report.pt"))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < WORLD, reason="needs 2 GPUs")
def test_two_nccl_ranks_equal_one_rank(tmp_path):
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(_free_port(), str(tmp_path)), nprocs=WORLD, join=True)
    r = torch.load(os.path.join(str(tmp_path), "report.pt"), weights_only=False)
    print(r)
    for tag in ("hmc", "generic", "conv3D"):
        assert r[tag + ".x_equal"], f"{tag}: 2-rank links differ from the 1-rank run"
        assert r[tag + ".obs_equal"], f"{tag}: 2-rank accept probabilities / observables differ from the 1-rank run"
        assert 0.0 < r[tag + ".accept_rate"] <= 1.0
    assert r["train.offsets"] == [0, 64]                      # set by train(), not by the test
    assert r["train.steps_per_rank"] == 2                     # train_steps // world size (gauge_model.py:1094-1095)
    assert r["train.own_masks_differed"] and r["train.masks_equal"]
    assert r["train.weights_equal"]
    assert r["train.losses_differ"]                           # the ranks drew different momenta
    # the first gradient bucket averaged on a side stream under momentum_fn's weight-gradient products: same weights
    assert r["overlap.used"] and r["overlap.equals_single_allreduce"] and r["overlap.ranks_equal"]
    # clip_value = 1e6 >> |g|: the per-rank clip (gauge_model.py:826-827) runs but is the identity, so the average of the
    # ranks' gradients must equal the 1-rank gradient of the concatenated batch
    assert r["grad.local_norm"] < 1e5
    assert r["grad.max_abs_diff_over_scale"] <= 1e-6, r
