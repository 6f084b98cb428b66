"""The CPU oracle against the golden vectors the unmodified reference produced
(tests/golden/make_golden.py).  This is what pins the oracle; CPU only."""
import numpy as np
import pytest
import torch

from oracle import u1_oracle as orc

RTOL = 2e-6   # same ops, same order, same CPU library: expected bit-exact or last-ulp


def t32(a):
    return torch.as_tensor(np.asarray(a, dtype=np.float32))


def close(a, b, rtol=RTOL, atol=2e-6):
    np.testing.assert_allclose(np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64), rtol=rtol, atol=atol)


@pytest.mark.parametrize("shape", ["8x8", "16x16", "6x10", "4x4", "32x32"])
def test_lattice_observables_and_force(golden, shape):
    g = golden("lattice_cases")
    T, X = map(int, shape.split("x"))
    x = t32(g[shape + ".x"])
    beta = torch.tensor(float(g[shape + ".beta"]), dtype=torch.float32)
    close(orc.total_action(x, T, X), g[shape + ".total_action"])
    a, p, q = orc.plaq_observables(x, T, X)
    close(a, g[shape + ".obs_action"])
    close(p, g[shape + ".obs_avg_plaq"])
    assert np.array_equal(q.numpy(), g[shape + ".obs_charge"])          # integer charge: exact
    close(orc.potential_energy(x, beta, T, X), g[shape + ".potential"])
    close(orc.kinetic_energy(t32(g[shape + ".v"])), g[shape + ".kinetic"])
    f = orc.grad_potential(x, beta, T, X)
    close(f, g[shape + ".force"], atol=3e-6)
    close(orc.grad_potential_autograd(x, beta, T, X), g[shape + ".force"], atol=3e-6)
    close(orc.grad_potential_loops(x.numpy(), float(beta), T, X), g[shape + ".force"], rtol=1e-5, atol=2e-5)
    # cold-start row (notebook record: ACTION 0, PLAQ 1)
    assert a[0] == 0 and p[0] == 1 and q[0] == 0


def test_masks_match_reference_rng(golden):
    g = golden("hmc_cases")
    for tag in ("8x8", "16x16", "6x10", "32x32"):
        T, X, N, M = g[tag + ".meta"]
        m = orc.construct_masks(2 * T * X, int(M), seed=42)
        assert np.array_equal(m, g[tag + ".mask"])
        assert (m.sum(axis=1) == T * X).all()


def regenerated_nets(g, tag, arch):
    """Goldens whose weights are a function of a seed (tests/parity_util.trained_like_weights) keep the seed and the
    sha256 of the weights instead of the tensors."""
    from parity_util import trained_like_weights, weights_digest
    T, X = (int(v) for v in g[tag + ".meta"][:2])
    ws = trained_like_weights(arch, T, X, np.random.default_rng(int(g[tag + ".weight_seed"])), gain=1.0)
    assert weights_digest(ws) == bytes(g[tag + ".weight_sha256"]).hex(), "regenerated weights differ from the golden's"
    return ws


def _dyn_from(g, tag, hmc, arch="generic"):
    T, X, N, M = (int(v) for v in g[tag + ".meta"])
    nets = {}
    if not hmc and tag + ".weight_seed" in g:
        nets = regenerated_nets(g, tag, arch)
    elif not hmc:
        for net in ("XNet", "VNet"):
            pre = f"{tag}.{net}."
            nets[net] = {k[len(pre):]: v for k, v in g.items() if k.startswith(pre) and ".call." not in k}
    dyn = orc.OracleDynamics(T, X, M, float(g[tag + ".eps"]), hmc=hmc, network_arch=arch,
                             mask=g[tag + ".mask"], xnet=nets.get("XNet"), vnet=nets.get("VNet"))
    return dyn, (T, X, N, M)


@pytest.mark.parametrize("tag", ["8x8", "16x16", "6x10", "32x32"])
def test_hmc_transition(golden, tag):
    g = golden("hmc_cases")
    dyn, _ = _dyn_from(g, tag, hmc=True)
    beta = float(g[tag + ".beta"])
    xf, vf, pf = dyn.transition_kernel(g[tag + ".x"], beta, g[tag + ".v_f"], forward=True)
    xb, vb, pb = dyn.transition_kernel(g[tag + ".x"], beta, g[tag + ".v_b"], forward=False)
    close(xf, g[tag + ".x_f"], atol=2e-5); close(vf, g[tag + ".v_f_out"], atol=2e-5); close(pf, g[tag + ".p_f"], rtol=1e-4)
    close(xb, g[tag + ".x_b"], atol=2e-5); close(vb, g[tag + ".v_b_out"], atol=2e-5); close(pb, g[tag + ".p_b"], rtol=1e-4)
    xp, vp, p, xo = dyn.apply_transition(g[tag + ".x"], beta, g[tag + ".v_f"], g[tag + ".v_b"],
                                         g[tag + ".u_dir"], g[tag + ".u_acc"])
    close(xp, g[tag + ".x_post"], atol=2e-5); close(vp, g[tag + ".v_post"], atol=2e-5)
    close(p, g[tag + ".accept_prob"], rtol=1e-4)
    # accept decisions bit-exact: x_out rows are either x_post or x
    acc_ref = (g[tag + ".accept_prob"] > g[tag + ".u_acc"])
    acc = (p.numpy() > g[tag + ".u_acc"])
    assert np.array_equal(acc, acc_ref)
    close(xo, g[tag + ".x_out"], atol=2e-5)


@pytest.mark.parametrize("tag,arch", [("generic.4x4", "generic"), ("conv3D.4x4", "conv3D"), ("conv3D.8x8", "conv3D"),
                                      ("conv3D.16x16", "conv3D")])
def test_l2hmc_pieces_and_transition(golden, tag, arch):
    g = golden("l2hmc_config3_case" if tag == "conv3D.16x16" else "l2hmc_cases")   # BASELINE config 3's shape
    dyn, (T, X, N, M) = _dyn_from(g, tag, hmc=False, arch=arch)
    beta = torch.tensor(float(g[tag + ".beta"]))
    x, v = t32(g[tag + ".x"]), t32(g[tag + ".v_f"])
    # bare net calls
    for net, fn in (("XNet", dyn.position_fn), ("VNet", dyn.momentum_fn)):
        S, Tr, Q = fn(v, x, t32(g[f"{tag}.{net}.call.tvec"]))
        close(S, g[f"{tag}.{net}.call.S"], atol=1e-5); close(Tr, g[f"{tag}.{net}.call.T"], atol=1e-5)
        close(Q, g[f"{tag}.{net}.call.Q"], atol=1e-5)
    tv = dyn._tvec(1, N)
    m, mi = dyn._masks(1)
    for nm, fn in (("vf", dyn.update_momentum_forward), ("vb", dyn.update_momentum_backward)):
        vv, ld = fn(x, v, beta, tv)
        close(vv, g[f"{tag}.sub.{nm}.v"], atol=2e-5); close(ld, g[f"{tag}.sub.{nm}.logdet"], atol=2e-5)
    for nm, fn in (("xf", dyn.update_position_forward), ("xb", dyn.update_position_backward)):
        xx, ld = fn(x, v, tv, m, mi)
        close(xx, g[f"{tag}.sub.{nm}.x"], atol=2e-5); close(ld, g[f"{tag}.sub.{nm}.logdet"], atol=2e-5)
    xx, vv, ld = dyn.forward_lf(x, v, beta, 1)
    close(xx, g[tag + ".lf_f.x"], atol=5e-5); close(vv, g[tag + ".lf_f.v"], atol=5e-5); close(ld, g[tag + ".lf_f.logdet"], atol=5e-5)
    xx, vv, ld = dyn.backward_lf(x, v, beta, 1)
    close(xx, g[tag + ".lf_b.x"], atol=5e-5); close(vv, g[tag + ".lf_b.v"], atol=5e-5); close(ld, g[tag + ".lf_b.logdet"], atol=5e-5)
    xp, vp, p, xo = dyn.apply_transition(g[tag + ".x"], float(beta), g[tag + ".v_f"], g[tag + ".v_b"],
                                         g[tag + ".u_dir"], g[tag + ".u_acc"])
    close(xp, g[tag + ".x_post"], rtol=1e-4, atol=2e-4); close(vp, g[tag + ".v_post"], rtol=1e-4, atol=2e-4)
    close(p, g[tag + ".accept_prob"], rtol=1e-3, atol=1e-6)
    assert np.array_equal(p.numpy() > g[tag + ".u_acc"], g[tag + ".accept_prob"] > g[tag + ".u_acc"])
    close(xo, g[tag + ".x_out"], rtol=1e-4, atol=2e-4)


def test_lf_reversibility_and_logdet():
    """_backward_lf(step = M-1-k) inverts _forward_lf(step = k) with opposite log-det (SURVEY 8c iii)."""
    rng = np.random.default_rng(3)
    T = X = 4
    D, N, M = 32, 5, 3
    xw = orc.init_net_weights("generic", T, X, 2., rng)
    vw = orc.init_net_weights("generic", T, X, 1., rng)
    for w in (xw, vw):
        for k in ("scale_layer", "translation_layer", "transformation_layer"):
            w[k + ".kernel"] *= 30
    dyn = orc.OracleDynamics(T, X, M, 0.2, hmc=False, xnet=xw, vnet=vw, dtype=torch.float64)
    x = torch.as_tensor(rng.uniform(0, 2 * np.pi, (N, D)))
    v = torch.as_tensor(rng.standard_normal((N, D)))
    beta = torch.tensor(2.0, dtype=torch.float64)
    for k in range(M):
        x1, v1, ld = dyn.forward_lf(x, v, beta, k)
        x2, v2, ldb = dyn.backward_lf(x1, v1, beta, M - 1 - k)
        close(x2, x, rtol=1e-10, atol=1e-10); close(v2, v, rtol=1e-10, atol=1e-10)
        close(ldb, -ld, rtol=1e-10, atol=1e-10)
        assert float(ld.abs().max()) > 1e-4


def test_hmc_energy_conservation_and_gauge_invariance():
    rng = np.random.default_rng(5)
    T, X, N = 8, 8, 4
    x = rng.uniform(0, 2 * np.pi, (N, 2 * T * X))
    v = rng.standard_normal((N, 2 * T * X))
    errs = []
    for eps in (0.1, 0.05, 0.025):
        dyn = orc.OracleDynamics(T, X, int(round(1.0 / eps)), eps, hmc=True, dtype=torch.float64)
        xt, vt = torch.as_tensor(x), torch.as_tensor(v)
        b = torch.tensor(2.0, dtype=torch.float64)
        h0 = orc.hamiltonian(xt, vt, b, T, X)
        x1, v1, p = dyn.transition_kernel(x, 2.0, v, forward=True)
        errs.append(float((orc.hamiltonian(x1, v1, b, T, X) - h0).abs().max()))
    assert errs[1] < errs[0] / 3 and errs[2] < errs[1] / 3        # |dH| = O(eps^2)
    # gauge transformation U_mu(n) -> U_mu(n) + a(n) - a(n + mu)
    a = rng.uniform(0, 2 * np.pi, (N, T, X))
    s = x.reshape(N, T, X, 2).copy()
    s[..., 0] += a - np.roll(a, -1, axis=1)
    s[..., 1] += a - np.roll(a, -1, axis=2)
    xg = torch.as_tensor(s.reshape(N, -1))
    close(orc.total_action(xg, T, X), orc.total_action(torch.as_tensor(x), T, X), rtol=1e-10, atol=1e-10)
    q = orc.top_charges(torch.as_tensor(x), T, X)
    close(q, torch.round(q), atol=1e-9)                              # sum P = 0 => Q integer


def test_misc_host_helpers():
    assert abs(orc.u1_plaq_exact(2.0) - 0.69777) < 1e-5 and abs(orc.u1_plaq_exact(4.0) - 0.86352) < 1e-5
    assert abs(orc.update_beta(0, 2., 4., 100) - 2.) < 1e-12 and abs(orc.update_beta(100, 2., 4., 100) - 4.) < 1e-12
    w = orc.wrap_2pi(np.array([-0.5, 7.0, 0.0, 6.2831855], np.float32))
    assert (w >= 0).all() and (w < np.float32(2 * np.pi) + 1e-6).all()


# --------------------------------------------------------------------------- MoG path (config 5)
MOG_TAGS = ["mog2.L2", "mog2.L5.T2", "ring4.L3", "mog2.hmc", "gmm6.L2"]


def mog_dyn_from(g, tag, dtype=torch.float32):
    from oracle import mog_oracle as mog
    x_dim, L, H, N, hmc, nb = (int(v) for v in g[tag + ".meta"])
    nets = {}
    if not hmc:
        for net in ("XNet", "VNet"):
            pre = f"{tag}.{net}."
            nets[net] = {k[len(pre):]: v for k, v in g.items() if k.startswith(pre) and ".call." not in k}
    target = mog.GMMTarget(g[tag + ".mus"], g[tag + ".i_sigmas"], g[tag + ".constants"], dtype)
    dyn = mog.OracleMogDynamics(target, x_dim, L, float(g[tag + ".eps"]), hmc=bool(hmc), mask=g[tag + ".mask"],
                                xnet=nets.get("XNet"), vnet=nets.get("VNet"),
                                temperature=float(g[tag + ".temperature"]), dtype=dtype)
    return dyn, (x_dim, L, H, N, bool(hmc))


@pytest.mark.parametrize("tag", MOG_TAGS)
def test_mog_oracle_against_reference_goldens(golden, tag):
    from oracle import mog_oracle as mog
    g = golden("mog_cases")
    dyn, (x_dim, L, H, N, hmc) = mog_dyn_from(g, tag)
    assert np.array_equal(mog.init_masks(x_dim, L, seed=42), g[tag + ".mask"])
    x, vf, vb = t32(g[tag + ".x"]), t32(g[tag + ".v_f"]), t32(g[tag + ".v_b"])
    close(dyn.energy(x), g[tag + ".energy"], rtol=1e-5, atol=1e-5)
    close(dyn.grad_energy(x), g[tag + ".grad"], rtol=1e-5, atol=2e-5)
    if not hmc:
        tv = t32(g[tag + ".call.tvec"])
        for name, fn in (("XNet", dyn.XNet), ("VNet", dyn.VNet)):
            S, Tr, Q = fn(vf, x, tv)
            close(S, g[f"{tag}.{name}.call.S"], atol=2e-6); close(Tr, g[f"{tag}.{name}.call.T"], atol=2e-6)
            close(Q, g[f"{tag}.{name}.call.Q"], atol=2e-6)
    step = 1 if L > 1 else 0
    xo, vo, lj = dyn.forward_step(x, vf, step)
    close(xo, g[tag + ".fstep.x"], atol=1e-5); close(vo, g[tag + ".fstep.v"], rtol=1e-5, atol=2e-5)
    close(lj, g[tag + ".fstep.logjac"], atol=1e-5)
    xo, vo, lj = dyn.backward_step(x, vf, step)
    close(xo, g[tag + ".bstep.x"], atol=1e-5); close(vo, g[tag + ".bstep.v"], rtol=1e-5, atol=2e-5)
    close(lj, g[tag + ".bstep.logjac"], atol=1e-5)
    X, V, p = dyn.forward(x, vf)
    close(X, g[tag + ".fwd.x"], rtol=1e-5, atol=2e-5); close(V, g[tag + ".fwd.v"], rtol=1e-4, atol=1e-4)
    close(p, g[tag + ".fwd.p"], rtol=2e-3, atol=1e-6)
    X, V, p = dyn.backward(x, vb)
    close(X, g[tag + ".bwd.x"], rtol=1e-5, atol=2e-5); close(V, g[tag + ".bwd.v"], rtol=1e-4, atol=1e-4)
    close(p, g[tag + ".bwd.p"], rtol=2e-3, atol=1e-6)
    Lx, Lv, px, xout, acc = dyn.propose(x, vf, vb, g[tag + ".dir_mask"], g[tag + ".u_acc"])
    close(Lx, g[tag + ".propose.Lx"], rtol=1e-5, atol=2e-5); close(px, g[tag + ".propose.px"], rtol=2e-3, atol=1e-6)
    assert np.array_equal(acc.numpy(), (g[tag + ".propose.px"] - g[tag + ".u_acc"]) >= 0)
    close(xout, g[tag + ".propose.x_out"], rtol=1e-5, atol=2e-5)


def test_mog_step_reversibility_and_tunneling_rate():
    from oracle import mog_oracle as mog
    rng = np.random.default_rng(2)
    target = mog.GMMTarget.from_params([[1., 0.], [-1., 0.]], [0.05 * np.eye(2)] * 2, [0.5, 0.5], torch.float64)
    xw, vw = mog.init_mog_net(2, 2., rng), mog.init_mog_net(2, 1., rng)
    for w in (xw, vw):
        for k in ("scale_layer", "translation_layer", "transformation_layer"):
            w[k + ".kernel"] *= 100
    dyn = mog.OracleMogDynamics(target, 2, 3, 0.1, xnet=xw, vnet=vw, dtype=torch.float64)
    x = torch.as_tensor(rng.standard_normal((6, 2)) * 0.3 + np.array([1., 0.]))
    v = torch.as_tensor(rng.standard_normal((6, 2)))
    for k in range(3):
        x1, v1, lj = dyn.forward_step(x, v, k)
        x2, v2, ljb = dyn.backward_step(x1, v1, k)
        close(x2, x, rtol=1e-9, atol=1e-9); close(v2, v, rtol=1e-9, atol=1e-9); close(ljb, -lj, rtol=1e-9, atol=1e-9)
    traj = np.zeros((5, 2, 2))
    traj[:, 0, 0] = [1, -1, -1, 1, 1]        # three switches in four steps
    traj[:, 1, 0] = [1, 1, 1, 1, 1]
    assert np.allclose(mog.calc_tunneling_rate(traj, [[1., 0.], [-1., 0.]], 2), [0.5, 0.0])


# ----------------------------------------------------------------------------- section 8(f) rows: loss value, run statistics
def _train_tags(g, prefix):
    return sorted({k.rsplit(".meta", 1)[0] for k in g if k.startswith(prefix) and k.endswith(".meta")})


def train_dyn_from(g, tag):
    T, X, N, M, hmc = (int(v) for v in g[tag + ".meta"])
    arch = tag.split(".")[1]
    nets = {}
    if not hmc:
        for net in ("XNet", "VNet"):
            pre = f"{tag}.{net}."
            nets[net] = {k[len(pre):]: v for k, v in g.items() if k.startswith(pre)}
    dyn = orc.OracleDynamics(T, X, M, float(g[tag + ".eps"]), hmc=bool(hmc), network_arch="generic" if hmc else arch,
                             mask=g[tag + ".mask"], xnet=nets.get("XNet"), vnet=nets.get("VNet"))
    return dyn, (T, X, N, M, bool(hmc))


def test_loss_value_against_reference_goldens(golden):
    """gauge_model.py:728-803 (_calc_loss) restated in oracle/train_oracle.py vs the unmodified reference."""
    from oracle import train_oracle as tro
    g = golden("train_cases")
    tags = _train_tags(g, "loss")
    assert len(tags) == 6
    for tag in tags:
        dyn, (T, X, N, M, hmc) = train_dyn_from(g, tag)
        metric = tag.split(".")[-1]
        aux, std, chg = (float(v) for v in g[tag + ".weights"])
        beta, ls = float(g[tag + ".beta"]), float(g[tag + ".loss_scale"])
        for nm in ("x", "x_prop", "z"):
            close(tro.top_charges(g[f"{tag}.{nm}"], T, X, fft=True), g[f"{tag}.qfft.{nm}"], rtol=1e-5, atol=2e-5)
            close(tro.top_charges(g[f"{tag}.{nm}"], T, X, fft=False), g[f"{tag}.q.{nm}"], rtol=1e-5, atol=2e-5)
        # the loss terms on the reference's own proposals: isolates the loss arithmetic
        loss, per_chain, x_dq = tro.loss_terms(g[tag + ".x"], g[tag + ".x_prop"], g[tag + ".z"], g[tag + ".px"],
                                               g[tag + ".pz"], g[tag + ".x_out"], T, X, metric, ls, aux, std, chg)
        scale = float(per_chain.abs().max())
        close(loss, g[tag + ".loss"], rtol=2e-5, atol=2e-5 * scale)
        assert np.array_equal(x_dq.numpy(), g[tag + ".x_dq"])
        # end to end through the oracle dynamics (both chains)
        loss2, x_out, px, x_dq2 = tro.calc_loss(
            dyn, g[tag + ".x"], beta, (g[tag + ".v_f"], g[tag + ".v_b"], g[tag + ".u_dir"], g[tag + ".u_acc"]),
            g[tag + ".z"], (g[tag + ".vz_f"], g[tag + ".vz_b"], g[tag + ".uz_dir"], g[tag + ".uz_acc"]),
            metric=metric, loss_scale=ls, aux_weight=aux, std_weight=std, charge_weight=chg)
        close(px, g[tag + ".px"], rtol=1e-3, atol=1e-6)
        close(x_out, g[tag + ".x_out"], rtol=1e-4, atol=2e-4)
        close(loss2, g[tag + ".loss"], rtol=5e-3, atol=5e-3 * scale)     # 1/x_std amplifies the proposals' rounding
        assert np.array_equal(x_dq2.numpy(), g[tag + ".x_dq"])


def test_observables_stats_against_reference_goldens(golden):
    """gauge_model.py:1473-1533 (calc_observables_stats) restated vs the unmodified reference."""
    from oracle import train_oracle as tro
    g = golden("train_cases")
    tags = _train_tags(g, "stats")
    assert len(tags) == 4
    for tag in tags:
        steps, N, tf_ = (int(v) for v in g[tag + ".meta"])
        st = tro.calc_observables_stats(g[tag + ".actions"], g[tag + ".plaqs"], g[tag + ".charges"], therm_frac=tf_)
        for nm in ("actions", "plaqs", "charges", "suscept"):
            np.testing.assert_allclose(st[nm][0], g[f"{tag}.{nm}.mean"], rtol=1e-6, atol=1e-7)
            np.testing.assert_allclose(st[nm][1], g[f"{tag}.{nm}.err"], rtol=1e-5, atol=1e-7)
        assert list(st["charge_probabilities"].keys()) == list(g[tag + ".prob.keys"])
        np.testing.assert_allclose(list(st["charge_probabilities"].values()), g[tag + ".prob.vals"], rtol=1e-12)


def test_loss_gradients_against_reference_goldens(golden):
    """d loss / d eps and d loss / d weights (gauge_model.py:805-832) from autodiff through the oracle vs the
    reference's tf.gradients (run on the stand-in): pins the backward reference for the next round's kernels."""
    from oracle import train_oracle as tro
    g = golden("train_cases")
    for tag in _train_tags(g, "loss"):
        dyn, (T, X, N, M, hmc) = train_dyn_from(g, tag)
        aux, std, chg = (float(v) for v in g[tag + ".weights"])
        loss, grads = tro.loss_and_grads(
            dyn, g[tag + ".x"], float(g[tag + ".beta"]),
            (g[tag + ".v_f"], g[tag + ".v_b"], g[tag + ".u_dir"], g[tag + ".u_acc"]), g[tag + ".z"],
            (g[tag + ".vz_f"], g[tag + ".vz_b"], g[tag + ".uz_dir"], g[tag + ".uz_acc"]),
            metric=tag.split(".")[-1], loss_scale=float(g[tag + ".loss_scale"]), aux_weight=aux, std_weight=std,
            charge_weight=chg)
        names = [k[len(tag) + 6:] for k in g if k.startswith(tag + ".grad.")]
        assert set(names) == set(grads), (sorted(names)[:3], sorted(grads)[:3])
        for nm in names:
            ref = g[f"{tag}.grad.{nm}"]
            scale = max(float(np.abs(ref).max()), 1e-12)
            np.testing.assert_allclose(grads[nm].numpy(), ref, rtol=2e-2, atol=2e-3 * scale, err_msg=f"{tag} {nm}")


def test_mog_loss_restatement():
    """mog_model.py:344-355 on hand-computable numbers."""
    from oracle import mog_oracle as mo
    x = np.array([[0., 0.], [1., 1.]]); Lx = np.array([[1., 0.], [1., 3.]]); px = np.array([0.5, 1.0])
    z = np.array([[0., 0.], [0., 0.]]); Lz = np.array([[0., 2.], [0., 0.]]); pz = np.array([1.0, 0.3])
    loss, v1, v2 = mo.mog_loss(x, Lx, px, z, Lz, pz, 0.1)
    np.testing.assert_allclose(v1, [0.5 + 1e-4, 4.0 + 1e-4])
    np.testing.assert_allclose(v2, [4.0 + 1e-4, 1e-4])
    want = 0.1 * (np.mean(1 / v1) + np.mean(1 / v2)) - (np.mean(v1) + np.mean(v2)) / 0.1
    assert abs(loss - want) < 1e-12
