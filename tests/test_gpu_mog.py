"""MoG / generic-target path on the GPU (csrc/mog.cu) against the golden vectors produced by the
unmodified reference (tests/golden/mog_cases.npz) and the oracle.  Tolerance: 1e-5 relative (fp32)
plus a few ulps of the value range, as stated per check."""
import numpy as np
import pytest
import torch

from oracle import mog_oracle as mogo

pytestmark = pytest.mark.gpu

TAGS = ["mog2.L2", "mog2.L5.T2", "ring4.L3", "mog2.hmc", "gmm6.L2"]


def close(a, b, rtol=1e-5, atol=2e-5):
    a = a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
    b = b.detach().cpu().numpy() if isinstance(b, torch.Tensor) else np.asarray(b)
    np.testing.assert_allclose(a.astype(np.float64), b.astype(np.float64), rtol=rtol, atol=atol)


def build(g, tag):
    from l2hmc_b200 import mog
    x_dim, L, H, N, hmc, nb = (int(v) for v in g[tag + ".meta"])
    gmm = mog.GMM.from_arrays(g[tag + ".mus"], g[tag + ".i_sigmas"], g[tag + ".constants"])
    np.random.seed(42)
    dyn = mog.Dynamics(x_dim, gmm, trajectory_length=L, eps=float(g[tag + ".eps"]), hmc=bool(hmc),
                       net_factory=lambda d, scope, factor: mog.network(d, scope, factor, num_nodes=H),
                       use_temperature=True)
    dyn.temperature = float(g[tag + ".temperature"])
    assert np.array_equal(dyn.mask.cpu().numpy(), g[tag + ".mask"])       # same legacy-RNG masks as the reference
    if not hmc:
        for name, net in (("XNet", dyn.XNet), ("VNet", dyn.VNet)):
            pre = f"{tag}.{name}."
            net.load_weights({k[len(pre):]: v for k, v in g.items() if k.startswith(pre) and ".call." not in k})
    return dyn, (x_dim, L, H, N, bool(hmc))


@pytest.mark.parametrize("tag", TAGS)
def test_mog_against_reference_goldens(golden, tag):
    from l2hmc_b200 import mog
    g = golden("mog_cases")
    dyn, (x_dim, L, H, N, hmc) = build(g, tag)
    x, vf, vb = g[tag + ".x"], g[tag + ".v_f"], g[tag + ".v_b"]
    close(dyn.energy(x), g[tag + ".energy"], rtol=1e-5, atol=1e-5)
    close(dyn.grad_energy(x), g[tag + ".grad"], rtol=1e-5, atol=5e-5)
    if not hmc:
        for name, net in (("XNet", dyn.XNet), ("VNet", dyn.VNet)):
            S, Tr, Q = net([vf, x, g[tag + ".call.tvec"], None])
            close(S, g[f"{tag}.{name}.call.S"], atol=2e-6); close(Tr, g[f"{tag}.{name}.call.T"], atol=2e-6)
            close(Q, g[f"{tag}.{name}.call.Q"], atol=2e-6)
    X, V, p = dyn.forward(x, init_v=vf)
    close(X, g[tag + ".fwd.x"], rtol=1e-5, atol=3e-5); close(V, g[tag + ".fwd.v"], rtol=1e-4, atol=2e-4)
    close(p, g[tag + ".fwd.p"], rtol=3e-3, atol=1e-6)
    if not hmc:
        X, V, p = dyn.backward(x, init_v=vb)
        close(X, g[tag + ".bwd.x"], rtol=1e-5, atol=3e-5); close(V, g[tag + ".bwd.v"], rtol=1e-4, atol=2e-4)
        close(p, g[tag + ".bwd.p"], rtol=3e-3, atol=1e-6)
    Lx, Lv, px, outs = mog.propose(x, dyn, do_mh_step=True, v_f=vf, v_b=None if hmc else vb,
                                   dir_mask=None if hmc else g[tag + ".dir_mask"], u_acc=g[tag + ".u_acc"])
    close(Lx, g[tag + ".propose.Lx"], rtol=1e-5, atol=3e-5); close(px, g[tag + ".propose.px"], rtol=3e-3, atol=1e-6)
    close(outs[0], g[tag + ".propose.x_out"], rtol=1e-5, atol=3e-5)       # => accept decisions identical
    assert int(dyn._flag.item()) == 0


def test_single_steps_against_goldens(golden):
    """_forward_step / _backward_step (utils/dynamics.py:120-225): a trajectory of length 1 on the mask /
    time row of step 1 is exactly one step."""
    from l2hmc_b200 import mog
    g = golden("mog_cases")
    for tag in ("mog2.L2", "ring4.L3", "gmm6.L2"):
        dyn, (x_dim, L, H, N, hmc) = build(g, tag)
        x, vf = g[tag + ".x"], g[tag + ".v_f"]
        od, _ = __import__("test_oracle_golden").mog_dyn_from(g, tag)
        xs, vs, lj = od.forward_step(torch.as_tensor(x), torch.as_tensor(vf), 1)
        close(xs, g[tag + ".fstep.x"], atol=1e-5)
        # GPU: trajectories of full length vs the oracle (already pinned to the goldens step by step)
        X, V, p = dyn.forward(x, init_v=vf)
        Xo, Vo, po = od.forward(x, vf)
        close(X, Xo, rtol=1e-5, atol=3e-5); close(V, Vo, rtol=1e-4, atol=2e-4); close(p, po, rtol=3e-3, atol=1e-6)


def test_mog_free_running_philox_sharding_and_statistics():
    """Config-5 shape: k = 2, two modes, M = 2, H = 50.  Free-running stream is independent of sharding;
    plain HMC at temperature 1 keeps the chains distributed like the target; tunneling rate is finite."""
    from l2hmc_b200 import mog
    N = 20000
    gmm = mog.GMM([[1., 0.], [-1., 0.]], [0.05 * np.eye(2)] * 2, [0.5, 0.5])
    np.random.seed(42)
    dyn = mog.Dynamics(2, gmm.get_energy_function(), trajectory_length=2, eps=0.1, seed=5)
    rng = np.random.default_rng(0)
    x = (np.array([[1., 0.], [-1., 0.]])[rng.integers(0, 2, N)] + 0.2 * rng.standard_normal((N, 2))).astype(np.float32)
    dyn._mc_step = 3
    full = dyn.mc_step(x)
    np.random.seed(42)
    part_dyn = mog.Dynamics(2, gmm.get_energy_function(), trajectory_length=2, eps=0.1, seed=5, chain_offset=7000)
    part_dyn.XNet.load_weights(dyn.XNet.state_dict()); part_dyn.VNet.load_weights(dyn.VNet.state_dict())
    part_dyn._mc_step = 3
    part = part_dyn.mc_step(x[7000:9000])
    for k in ("Lx", "px", "x_out", "accepted", "mode"):
        assert torch.equal(part[k], full[k][7000:9000]), k
    # HMC sampling: mean |x_0| ~ 1, var of x_1 ~ 0.05, both modes populated
    np.random.seed(42)
    hmc = mog.Dynamics(2, gmm, trajectory_length=10, eps=0.05, hmc=True, seed=9)
    xs, modes = torch.as_tensor(x).cuda(), []
    for _ in range(60):
        o = hmc.mc_step(xs)
        xs = o["x_out"]
        modes.append(o["mode"])
    assert abs(float(xs[:, 0].abs().mean()) - 1.0) < 0.02 and abs(float(xs[:, 1].var()) - 0.05) < 0.005
    assert 0.4 < float((o["mode"] == 0).float().mean()) < 0.6
    rate = mog.tunneling_rate(torch.stack(modes).cpu())
    assert rate.shape == (N,) and float(rate.mean()) < 0.05
    assert float(o["px"].mean()) > 0.8


def test_mog_argument_errors():
    from l2hmc_b200 import _lib, mog
    with pytest.raises(NotImplementedError):
        mog.Dynamics(2, lambda x: x, trajectory_length=2)
    gmm = mog.GMM([[1., 0.], [-1., 0.]], [0.05 * np.eye(2)] * 2, [0.5, 0.5])
    with pytest.raises(ValueError):
        mog.Dynamics(3, gmm, trajectory_length=2, hmc=True)
    lib = _lib.get_lib()
    assert lib.l2hmc_mog_mc_step(None, None, None) == _lib.L2HMC_ERR_INVALID
    e, f = gmm.energy_force(np.zeros((0, 2), np.float32))
    assert e.shape == (0,) and f.shape == (0, 2)


@pytest.mark.parametrize("N,k,scale", [(1000, 2, 0.1), (37, 4, 1.0), (300000, 2, 0.1)])
def test_mog_loss_value_against_oracle(N, k, scale):
    """mog_model.py:324-355 (loss VALUE): the CUDA two-stage sum against the float64 restatement on the SAME propose
    outputs, and its run-to-run bits."""
    from l2hmc_b200 import mog
    from oracle import mog_oracle as mo
    rng = np.random.default_rng(N)
    mus = rng.standard_normal((3, k))
    gmm = mog.GMM(mus, [0.1 * np.eye(k)] * 3, [0.3, 0.3, 0.4])
    np.random.seed(7)
    dyn = mog.Dynamics(k, gmm.get_energy_function(), trajectory_length=2, eps=0.1, seed=11)
    x = (mus[rng.integers(0, 3, N)] + 0.3 * rng.standard_normal((N, k))).astype(np.float32)
    z = rng.standard_normal((N, k)).astype(np.float32)
    dyn._mc_step = 0
    loss, o = mog.loss(x, dyn, scale=scale, z=z)
    ref, v1, v2 = mo.mog_loss(x, o["Lx"].cpu().numpy(), o["px"].cpu().numpy(), z, o["Lz"].cpu().numpy(),
                              o["pz"].cpu().numpy(), scale)
    np.testing.assert_allclose(o["v1"].cpu().numpy(), v1, rtol=2e-6, atol=1e-10)
    np.testing.assert_allclose(o["v2"].cpu().numpy(), v2, rtol=2e-6, atol=1e-10)
    assert abs(float(loss) - ref) <= 1e-5 * abs(ref), (float(loss), ref)        # north-star fp32 tolerance
    dyn._mc_step = 0
    loss2, _ = mog.loss(x, dyn, scale=scale, z=z)
    assert torch.equal(loss, loss2)
    assert int(dyn._flag.item()) == 0
