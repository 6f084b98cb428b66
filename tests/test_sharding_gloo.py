"""N > 1 host logic on CPU: world_size-2 gloo processes.  The compute stand-in is the
oracle; what is tested is the sharding plumbing of l2hmc_b200/sharding.py: contiguous shards with
global chain offsets, the Philox stream keyed by global chain index, the one all_reduce of
observable accumulators, the final gather, weight broadcast and gradient averaging."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import u1_oracle as orc
from philox_ref import philox_normals, philox_uniforms


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


T = X = 4
D, M, EPS, BETA, SEED = 32, 3, 0.1, 2.0, 1234


def _mc_step(x, chain0, mc_step):
    """One plain-HMC MC step of the oracle with the draws the CUDA path would generate in-kernel."""
    n = x.shape[0]
    z = philox_normals(SEED, mc_step, chain0, n, D)
    ud, ua = philox_uniforms(SEED, mc_step, chain0, n)
    dyn = orc.OracleDynamics(T, X, M, EPS, hmc=True)
    _, _, p, xo = dyn.apply_transition(x, BETA, z, z, ud, ua)
    a, plaq, q = orc.plaq_observables(xo, T, X)
    return xo, p, a, plaq, q


def _worker(rank, world_size, port, total, tmp):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    from l2hmc_b200 import sharding
    torch.set_num_threads(1)
    x_all = torch.as_tensor(np.random.default_rng(0).uniform(0, 2 * np.pi, (total, D)).astype(np.float32))
    x, lo = sharding.shard(x_all)
    assert (lo, lo + len(x)) == sharding.shard_range(total)
    acc = sharding.ObservableAccumulator(["px", "action", "plaq", "charge"])
    for step in range(2):
        x, p, a, plaq, q = _mc_step(x, lo, step)
        acc.add(px=p, action=a, plaq=plaq, charge=q)
        x = torch.as_tensor(orc.wrap_2pi(x.numpy()))
    means = acc.reduce()
    full = sharding.gather_chains(x, total)
    # training collectives: weights from rank 0, gradients averaged
    w = [torch.full((5,), float(rank + 1)), torch.full((2, 2), float(10 * (rank + 1)))]
    sharding.broadcast_weights(w)
    g = sharding.allreduce_mean_(torch.arange(6, dtype=torch.float32) * (rank + 1))
    if rank == 0:
        torch.save({"means": means, "full": full, "w": w, "g": g}, os.path.join(tmp, "out.pt"))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("total", [10, 7])
def test_two_rank_run_equals_single_rank(tmp_path, total):
    from l2hmc_b200 import sharding
    port = _free_port()
    mp.spawn(_worker, args=(2, port, total, str(tmp_path)), nprocs=2, join=True)
    got = torch.load(os.path.join(str(tmp_path), "out.pt"), weights_only=False)
    # single-rank reference on the concatenated batch
    x = torch.as_tensor(np.random.default_rng(0).uniform(0, 2 * np.pi, (total, D)).astype(np.float32))
    acc = sharding.ObservableAccumulator(["px", "action", "plaq", "charge"])
    for step in range(2):
        x, p, a, plaq, q = _mc_step(x, 0, step)
        acc.add(px=p, action=a, plaq=plaq, charge=q)
        x = torch.as_tensor(orc.wrap_2pi(x.numpy()))
    ref = acc.reduce()
    assert torch.equal(got["full"], x)                       # bit-identical chains, in global order
    for k, v in ref.items():
        assert abs(got["means"][k] - v) <= 1e-12 * max(1.0, abs(v)), k
    assert got["means"]["chains_x_steps"] == 2 * total
    assert torch.equal(got["w"][0], torch.full((5,), 1.0)) and torch.equal(got["w"][1], torch.full((2, 2), 10.0))
    assert torch.allclose(got["g"], torch.arange(6, dtype=torch.float32) * 1.5)


def test_shard_ranges_partition_the_batch():
    from l2hmc_b200 import sharding
    for total in (0, 1, 7, 64, 65536, 65537):
        for ws in (1, 2, 3, 4, 8):
            edges = [sharding.shard_range(total, r, ws) for r in range(ws)]
            assert edges[0][0] == 0 and edges[-1][1] == total
            assert all(edges[i][1] == edges[i + 1][0] for i in range(ws - 1))
            sizes = [hi - lo for lo, hi in edges]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_range(10, 3, 2)
    assert sharding.world() == (0, 1)


def test_run_logger_rows_have_the_reference_columns(tmp_path):
    """Rank-0 JSONL log (gauge_model.py:1104-1110, 1190-1221; rank gate utils/file_io.py:27-46): one parsed line."""
    import json
    from l2hmc_b200.lattice import u1_plaq_exact
    from l2hmc_b200.run_logger import COLUMNS, RunLogger
    path = str(tmp_path / "run_info.jsonl")
    N = 6
    out = (torch.tensor(-12.5), None, torch.full((N,), 0.75), torch.tensor(0.21), torch.full((N,), 40.0),
           torch.full((N,), 0.6), torch.zeros(N), 9.6e-4, torch.tensor(0.5))
    log = RunLogger(path, print_steps=2, train_steps=10)
    assert log(0, out, 2.0) is not None and log(1, out, 2.0) is None and log(2, out, 2.5) is not None
    rows = [json.loads(line) for line in open(path)]
    assert [r["STEP"] for r in rows] == [0, 2]
    assert tuple(rows[0].keys()) == COLUMNS
    r = rows[1]
    assert r["LOSS"] == -12.5 and r["% ACC"] == 0.75 and abs(r["EPS"] - 0.21) < 1e-7 and r["BETA"] == 2.5
    assert r["ACTION"] == 40.0 and abs(r["PLAQ"] - 0.6) < 1e-7 and r["dQ"] == 0.5 and r["LR"] == 9.6e-4
    assert abs(r["(EXACT)"] - u1_plaq_exact(2.5)) < 1e-12 and r["t/STEP"] >= 0
    text = log.text(r)
    assert text.split()[0] == "2/10" and len(log.header.splitlines()) == 3 and "(EXACT)" in log.header
    # any other rank writes nothing
    quiet = RunLogger(str(tmp_path / "other.jsonl"), rank=1)
    assert quiet(0, out, 2.0) is None and not (tmp_path / "other.jsonl").exists()
