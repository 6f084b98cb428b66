"""GaugeSampler (l2hmc_b200/sampler.py): the host-buffer step of the reference's run loop (gauge_model.py:1379-1388) and
the device-resident run loop, for plain HMC and for L2HMC with the S/T/Q networks.  Chunked / pipelined / resident
execution must not change a bit: everything is compared with GaugeDynamics.apply_transition on the whole batch."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _make(arch, T, N, M=3, seed=9):
    import l2hmc_b200 as pkg
    from parity_util import trained_like_weights
    lat = pkg.GaugeLattice(T, T, 2, "U1", num_samples=N)
    ws = None if arch is None else trained_like_weights(arch, T, T, np.random.default_rng(3), gain=1.0)
    np.random.seed(42)
    return pkg.GaugeDynamics(lat, hmc=arch is None, network_arch=arch or "generic", num_steps=M, eps=0.15, weights=ws,
                             seed=seed, chain_offset=17)


def _reference_steps(dyn, x, beta, steps):
    """apply_transition on the whole batch, `steps` times: per-step (px, action, plaq, q_out, q_in) and final links."""
    xs, rows = torch.as_tensor(x).cuda(), []
    for _ in range(steps):
        q_in = dyn.lattice.observables(xs)[3]
        _, _, p, xo = dyn.apply_transition(xs, beta, wrap=True, observables=True)
        a, pl, q = dyn.last_observables
        rows.append((p.cpu(), a.cpu(), pl.cpu(), q.cpu(), q_in.cpu()))
        xs = xo
    return rows, xs.cpu()


@pytest.mark.parametrize("arch,T,N,chunk", [(None, 16, 700, 256), (None, 8, 64, 64), ("generic", 8, 200, 64),
                                            ("conv3D", 8, 150, 64), ("generic", 8, 900, 256), ("conv3D", 16, 600, 128),
                                            ("generic", 8, 2100, 256), ("conv3D", 16, 1100, 128)])
def test_host_buffer_step_equals_apply_transition(arch, T, N, chunk):
    from l2hmc_b200.sampler import GaugeSampler
    D, beta, steps = 2 * T * T, 2.5, 3
    x = np.random.default_rng(1).uniform(0, 2 * np.pi, (N, D)).astype(np.float32)
    ref_rows, ref_x = _reference_steps(_make(arch, T, N), x, beta, steps)
    dyn = _make(arch, T, N)
    s = GaugeSampler(dyn, N, chunk_chains=chunk)
    if arch is not None and N >= s.tc_min_rows:      # every chunk stays on the tensor-core path the whole batch takes
        assert s.chunk >= s.tc_min_rows
        assert s.nchunks >= 2 or N < 2 * s.tc_min_rows
    else:                                            # HMC, or FFMA layers (batch-size independent): chunks as asked for
        assert s.nchunks >= 2 or N <= chunk
    host = s.pinned_like(x)
    for k in range(steps):
        r = s.step(host, beta)
        p, a, pl, q, q_in = ref_rows[k]
        assert torch.equal(r.px, p) and torch.equal(r.actions, a) and torch.equal(r.plaqs, pl)
        assert torch.equal(r.charges, q)
        # charge_diff_op (gauge_model.py:849-850): |Q(samples fed) - Q(x_out)| summed over chains / N -- from step 1 on
        assert r.charge_diff == float((q - q_in).abs().sum()) / N
        host = r.samples.clone().pin_memory()
    assert torch.equal(host, ref_x)
    assert dyn._mc_step == steps
    dyn.raise_if_nonfinite()


@pytest.mark.parametrize("arch,T,N", [(None, 16, 500), ("generic", 8, 300), ("conv3D", 8, 140)])
def test_resident_run_equals_stepwise_and_feeds_run_stats(arch, T, N):
    from l2hmc_b200.gauge_model import RunStats
    from l2hmc_b200.sampler import GaugeSampler
    D, beta, steps = 2 * T * T, 2.5, 6
    x = np.random.default_rng(2).uniform(0, 2 * np.pi, (N, D)).astype(np.float32)
    ref_rows, ref_x = _reference_steps(_make(arch, T, N), x, beta, steps)
    dyn = _make(arch, T, N)
    s = GaugeSampler(dyn, N)
    stats = RunStats(N, steps, therm_frac=3)
    out = s.run(4, beta, samples=x, stats=stats)
    out = {k: (v.clone() if isinstance(v, torch.Tensor) else v) for k, v in out.items()}   # views of re-used buffers
    out2 = s.run(2, beta, stats=stats, return_samples="host")          # continues from the resident state
    for k in range(steps):
        o, j = (out, k) if k < 4 else (out2, k - 4)
        p, a, pl, q, q_in = ref_rows[k]
        assert torch.equal(o["px"][j], p) and torch.equal(o["actions"][j], a) and torch.equal(o["plaqs"][j], pl)
        assert torch.equal(o["charges"][j], q) and torch.equal(o["charges_in"][j], q_in)
        assert torch.equal(o["charge_diff"][j], (q - q_in).abs())
        assert float(o["tunneling_rate"][j]) == float((q - q_in).abs().sum()) / N
    assert torch.equal(out["samples"].cpu() if out["samples"].is_cuda else out["samples"], _reference_steps(
        _make(arch, T, N), x, beta, 4)[1])
    assert torch.equal(out2["samples"], ref_x)
    # the statistics accumulated on the device equal calc_observables_stats over the fetched history
    import l2hmc_b200 as pkg
    from l2hmc_b200.gauge_model import GaugeModel
    got = stats.finalize()
    hist = lambda key: {k: (out if k < 4 else out2)[key][k if k < 4 else k - 4].numpy() for k in range(steps)}
    model = GaugeModel(dict(time_size=T, space_size=T, num_samples=N, hmc=True), lattice=dyn.lattice, dynamics=dyn)
    want = model.calc_observables_stats((hist("actions"), hist("plaqs"), hist("charges")), therm_frac=3)
    for g_, w_ in zip(got[:4], want[:4]):
        np.testing.assert_array_equal(g_[0], w_[0])
        np.testing.assert_array_equal(g_[1], w_[1])
    assert got[4] == want[4]


def test_sampler_refuses_what_it_cannot_run():
    import l2hmc_b200 as pkg
    from l2hmc_b200.sampler import GaugeSampler
    lat = pkg.GaugeLattice(6, 10, 2, "U1", num_samples=4)
    dyn = pkg.GaugeDynamics(lat, hmc=True, num_steps=2, eps=0.1)       # no fused kernel for 6x10
    with pytest.raises(NotImplementedError):
        GaugeSampler(dyn, 4)
    s = GaugeSampler(_make(None, 8, 8), 8)
    with pytest.raises(ValueError):
        s.run(1, 2.0)                                                   # nothing to continue from


@pytest.mark.parametrize("arch,T,N", [("generic", 8, 128), ("conv3D", 8, 96), ("generic", 8, 512), (None, 8, 64)])
def test_graph_replay_equals_eager_loop(arch, T, N):
    """run(graph=True): the MC step captured once as a CUDA graph (Philox step counter on the device, advanced inside the
    graph) gives the same bits as the eager loop, also across run() calls and after an eager step in between."""
    from l2hmc_b200.sampler import GaugeSampler
    D, beta = 2 * T * T, 2.5
    x = np.random.default_rng(4).uniform(0, 2 * np.pi, (N, D)).astype(np.float32)
    keys = ("px", "actions", "plaqs", "charges", "charges_in")
    outs = {}
    for mode in (False, True):
        dyn = _make(arch, T, N)
        s = GaugeSampler(dyn, N)
        a = s.run(5, beta, samples=x, graph=mode, return_samples="host")
        a = {k: a[k].clone() for k in keys + ("samples",)}
        b = s.run(1, beta, graph=False, return_samples="host")                 # an eager step: the counters must re-sync
        b = {k: b[k].clone() for k in keys + ("samples",)}
        c = s.run(4, beta, graph=mode, return_samples="host")
        outs[mode] = (a, b, {k: c[k].clone() for k in keys + ("samples",)})
        assert dyn._mc_step == 10
        if mode:
            assert len(s._run_state["graphs"]) == 2                            # one per ping-pong direction, captured once
        dyn.raise_if_nonfinite()
    for part_e, part_g in zip(outs[False], outs[True]):
        for k in keys + ("samples",):
            assert torch.equal(part_e[k], part_g[k]), k
    assert float(outs[True][2]["px"].min()) < 1.0                             # the steps did something
