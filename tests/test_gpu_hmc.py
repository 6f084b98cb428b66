"""Plain-HMC path (config 1 and the 32x32 roofline config) on the GPU: fused one-launch MC step,
fused leapfrog-step kernel and the generic K1+K2 composition, against the reference's golden
vectors, the oracle, and size-independent properties at full size."""
import numpy as np
import pytest
import torch

from oracle import u1_oracle as orc

pytestmark = pytest.mark.gpu


def close(a, b, rtol=1e-5, atol=2e-5):
    a = a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
    b = b.detach().cpu().numpy() if isinstance(b, torch.Tensor) else np.asarray(b)
    np.testing.assert_allclose(a.astype(np.float64), b.astype(np.float64), rtol=rtol, atol=atol)


@pytest.fixture(scope="module")
def pkg():
    import l2hmc_b200
    return l2hmc_b200


def make_dyn(pkg, g, tag, fused):
    T, X, N, M = (int(v) for v in g[tag + ".meta"])
    lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=N)
    np.random.seed(42)
    dyn = pkg.GaugeDynamics(lat, lat.get_energy_function(), hmc=True, num_steps=M, eps=float(g[tag + ".eps"]),
                            use_fused=fused)
    assert np.array_equal(dyn.mask.cpu().numpy(), g[tag + ".mask"])      # same legacy-RNG masks as the reference
    return dyn


@pytest.mark.parametrize("tag,fused", [("8x8", True), ("8x8", False), ("16x16", True), ("16x16", False),
                                       ("6x10", False), ("32x32", True), ("32x32", False)])
def test_hmc_transition_against_reference_goldens(pkg, golden, tag, fused):
    g = golden("hmc_cases")
    dyn = make_dyn(pkg, g, tag, fused)
    assert dyn._fused_ok() == fused
    beta = float(g[tag + ".beta"])
    for fwd, vk, sfx in ((True, ".v_f", "f"), (False, ".v_b", "b")):
        xs, vs, ps = dyn.transition_kernel(g[tag + ".x"], beta, forward=fwd, momentum=g[tag + vk])
        close(xs, g[f"{tag}.x_{sfx}"]); close(vs, g[f"{tag}.v_{sfx}_out"]); close(ps, g[f"{tag}.p_{sfx}"], rtol=2e-4, atol=1e-6)
    xp, vp, p, xo = dyn.apply_transition(g[tag + ".x"], beta, momentum_f=g[tag + ".v_f"], momentum_b=g[tag + ".v_b"],
                                         direction_uniform=g[tag + ".u_dir"], accept_uniform=g[tag + ".u_acc"])
    close(xp, g[tag + ".x_post"]); close(vp, g[tag + ".v_post"])
    close(p, g[tag + ".accept_prob"], rtol=2e-4, atol=1e-6)
    # accept decisions bit-exact (and the margin by which they were taken)
    acc_ref = g[tag + ".accept_prob"] > g[tag + ".u_acc"]
    assert np.array_equal(dyn.last_accepted.cpu().numpy().astype(bool), acc_ref)
    margin = np.abs(g[tag + ".accept_prob"] - g[tag + ".u_acc"]).min()
    assert margin > 1e-4, f"golden case has a near-tie (margin {margin}); regenerate with another seed"
    close(xo, g[tag + ".x_out"])
    dyn.raise_if_nonfinite()


@pytest.mark.parametrize("T,X", [(4, 4), (8, 8), (16, 16), (32, 32), (8, 16), (16, 32)])
@pytest.mark.parametrize("N", [1, 37, 300])
def test_fused_equals_generic_and_oracle(pkg, T, X, N):
    """Ragged chain counts (tail tiles), every fused shape: fused == generic ~= oracle."""
    rng = np.random.default_rng(T * X + N)
    D, M, eps, beta = 2 * T * X, 3, 0.12, 2.5
    x = rng.uniform(0, 2 * np.pi, (N, D)).astype(np.float32)
    vf, vb = (rng.standard_normal((N, D)).astype(np.float32) for _ in range(2))
    ud, ua = (rng.uniform(size=N).astype(np.float32) for _ in range(2))
    lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=N)
    outs = []
    for fused in (True, False):
        dyn = pkg.GaugeDynamics(lat, hmc=True, num_steps=M, eps=eps, use_fused=fused)
        outs.append(dyn.apply_transition(x, beta, momentum_f=vf, momentum_b=vb, direction_uniform=ud,
                                         accept_uniform=ua, wrap=True, observables=True) + (dyn.last_accepted,))
        # observables of x_out (fused into the MC-step launch on the fused path) vs the stencil kernel
        a_, p_, _, q_ = lat.observables(outs[-1][3])
        close(dyn.last_observables[0], a_, rtol=1e-5, atol=1e-4)
        close(dyn.last_observables[1], p_, rtol=1e-5, atol=1e-6)
        assert torch.equal(dyn.last_observables[2], q_)
    for a, b in zip(outs[0][:2], outs[1][:2]):
        close(a, b, rtol=0, atol=2e-6)                 # same op order; sinf vs sincosf may differ by an ulp
    close(outs[0][2], outs[1][2], rtol=1e-4, atol=1e-6)
    od = orc.OracleDynamics(T, X, M, eps, hmc=True)
    xp, vp, p, xo = od.apply_transition(x, beta, vf, vb, ud, ua)
    close(outs[0][0], xp); close(outs[0][1], vp); close(outs[0][2], p, rtol=2e-4, atol=1e-6)
    safe = np.abs(p.numpy() - ua) > 1e-4
    acc = outs[0][4].cpu().numpy().astype(bool)
    assert np.array_equal(acc[safe], (p.numpy() > ua)[safe])
    xo_ref = orc.wrap_2pi(np.where(acc[:, None], xp.numpy(), x))
    d = np.abs(outs[0][3].cpu().numpy() - xo_ref)
    assert np.minimum(d, 2 * np.pi - d).max() < 5e-5      # compared on the circle (wrap boundary)


@pytest.mark.parametrize("T,X", [(8, 8), (32, 32), (16, 32)])
def test_hmc_step_kernel_against_oracle_and_reversible(pkg, T, X):
    rng = np.random.default_rng(9)
    N, D, eps, beta = 45, 2 * T * X, 0.1, 4.0
    x = rng.uniform(0, 2 * np.pi, (N, D)).astype(np.float32)
    v = rng.standard_normal((N, D)).astype(np.float32)
    lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=N)
    dyn = pkg.GaugeDynamics(lat, hmc=True, num_steps=1, eps=eps)
    od = orc.OracleDynamics(T, X, 1, eps, hmc=True)
    b = torch.tensor(beta)
    x1, v1 = dyn.hmc_step(x, v, beta)
    xr, vr, _ = od.forward_lf(torch.as_tensor(x), torch.as_tensor(v), b, 0)
    close(x1, xr); close(v1, vr)
    direction = torch.as_tensor(np.where(rng.uniform(size=N) > 0.5, 1, -1).astype(np.int8)).cuda()
    x2, v2 = dyn.hmc_step(x, v, beta, direction)
    xb, vb, _ = od.backward_lf(torch.as_tensor(x), torch.as_tensor(v), b, 0)
    sel = (direction.cpu().numpy() > 0)[:, None]
    close(x2, np.where(sel, xr.numpy(), xb.numpy())); close(v2, np.where(sel, vr.numpy(), vb.numpy()))
    # backward step undoes the forward step
    neg = torch.full((N,), -1, dtype=torch.int8, device="cuda")
    x3, v3 = dyn.hmc_step(x1, v1, beta, neg)
    close(x3, x, atol=3e-6); close(v3, v, atol=3e-6)


def test_free_running_fused_equals_generic_with_philox_draws(pkg):
    T = X = 8
    N, M = 70, 4
    lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=N)
    x = np.random.default_rng(4).uniform(0, 2 * np.pi, (N, 128)).astype(np.float32)
    res = []
    for fused in (True, False):
        dyn = pkg.GaugeDynamics(lat, hmc=True, num_steps=M, eps=0.1, use_fused=fused, seed=99, chain_offset=500)
        dyn._mc_step = 7
        res.append(dyn.apply_transition(x, 2.0) + (dyn.last_accepted,))
        assert dyn._mc_step == 8
    close(res[0][0], res[1][0], atol=2e-5); close(res[0][1], res[1][1], atol=2e-5)
    close(res[0][2], res[1][2], rtol=2e-4, atol=1e-6)
    assert torch.equal(res[0][4], res[1][4])
    # sharding independence: chains [30, 70) run alone with chain_offset 530 reproduce the slice
    dyn = pkg.GaugeDynamics(lat, hmc=True, num_steps=M, eps=0.1, seed=99, chain_offset=530)
    dyn._mc_step = 7
    part = dyn.apply_transition(x[30:], 2.0)
    for a, b in zip(part, res[0][:4]):
        assert torch.equal(a, b[30:])


def test_full_size_properties_32x32_65536_chains(pkg):
    """BASELINE config 4 size: reversibility, energy conservation O(eps^2), charge integrality."""
    T = X = 32
    N, D, beta = 65536, 2048, 4.0
    lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=1)
    gen = torch.Generator(device="cuda").manual_seed(0)
    x = torch.rand(N, D, device="cuda", generator=gen) * (2 * np.pi)
    v = torch.randn(N, D, device="cuda", generator=gen)
    dyn = pkg.GaugeDynamics(lat, hmc=True, num_steps=10, eps=0.1)
    xs, vs = x, v
    for _ in range(3):
        xs, vs = dyn.hmc_step(xs, vs, beta)
    neg = torch.full((N,), -1, dtype=torch.int8, device="cuda")
    for _ in range(3):
        xs, vs = dyn.hmc_step(xs, vs, beta, neg)
    assert float((xs - x).abs().max()) < 2e-5 and float((vs - v).abs().max()) < 2e-5
    # the step kernel composed M times == the one-launch trajectory
    xs, vs = x, v
    for _ in range(10):
        xs, vs = dyn.hmc_step(xs, vs, beta)
    xp, vp, p = dyn.transition_kernel(x, beta, forward=True, momentum=v)
    assert float((xp - xs).abs().max()) < 5e-6 and float((vp - vs).abs().max()) < 5e-6
    # |dH| shrinks 4x when eps halves (same trajectory length)
    dh = []
    for eps, M in ((0.1, 10), (0.05, 20)):
        d2 = pkg.GaugeDynamics(lat, hmc=True, num_steps=M, eps=eps)
        xq, vq, _ = d2.transition_kernel(x[:4096], beta, forward=True, momentum=v[:4096])
        dh.append(float((d2.hamiltonian(xq, vq, beta) - d2.hamiltonian(x[:4096], v[:4096], beta)).abs().mean()))
    assert dh[1] < dh[0] / 3
    # topological charge: real-valued sum is an integer to rounding, and equals the integer observable
    _, _, qf, qi = lat.observables(xp)
    assert float((qf - torch.round(qf)).abs().max()) < 2e-3
    assert torch.equal(torch.round(qf).to(torch.int32), qi)
    dyn.raise_if_nonfinite()


def test_plaquette_converges_to_exact_value(pkg):
    """<plaq> -> I1(beta)/I0(beta) (the reference prints this check every step, gauge_model.py:1149)."""
    T = X = 8
    N, beta = 4096, 2.0
    lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=1)
    dyn = pkg.GaugeDynamics(lat, hmc=True, num_steps=10, eps=0.1, seed=3)
    x = torch.zeros(N, 128, device="cuda")
    plaqs, accs = [], []
    for step in range(300):
        _, _, p, x = dyn.apply_transition(x, beta, wrap=True)
        if step >= 100:
            plaqs.append(float(lat.observables(x)[1].mean()))
            accs.append(float(p.mean()))
    assert abs(np.mean(plaqs) - pkg.u1_plaq_exact(beta)) < 4e-3, np.mean(plaqs)
    assert 0.5 < np.mean(accs) <= 1.0
    assert float(x.min()) >= 0 and float(x.max()) < 2 * np.pi + 1e-6
