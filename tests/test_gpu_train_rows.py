"""SURVEY section 8(f) rows on the GPU: the training loss VALUE (csrc/loss.cu) and the streaming run statistics
(csrc/stats.cu), against the goldens produced by the unmodified reference (tests/golden/train_cases.npz) and the
oracle (oracle/train_oracle.py).  Tolerances are stated per check (fp32, 1e-5 relative unless amplified)."""
import numpy as np
import pytest
import torch

from oracle import train_oracle as tro
from oracle import u1_oracle as orc

pytestmark = pytest.mark.gpu


def close(a, b, rtol=1e-5, atol=2e-5):
    a = a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
    b = b.detach().cpu().numpy() if isinstance(b, torch.Tensor) else np.asarray(b)
    np.testing.assert_allclose(a.astype(np.float64), b.astype(np.float64), rtol=rtol, atol=atol)


def tags(g, prefix):
    return sorted({k.rsplit(".meta", 1)[0] for k in g if k.startswith(prefix) and k.endswith(".meta")})


def build_model(g, tag):
    import l2hmc_b200 as pkg
    from l2hmc_b200.gauge_model import GaugeModel
    T, X, N, M, hmc = (int(v) for v in g[tag + ".meta"])
    arch, metric = tag.split(".")[1], tag.split(".")[-1]
    weights = None
    if not hmc:
        weights = {}
        for net in ("XNet", "VNet"):
            pre = f"{tag}.{net}."
            weights[net] = {k[len(pre):]: v for k, v in g.items() if k.startswith(pre)}
    lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=N)
    np.random.seed(42)
    dyn = pkg.GaugeDynamics(lat, hmc=bool(hmc), network_arch="generic" if hmc else arch, num_steps=M,
                            eps=float(g[tag + ".eps"]), weights=weights)
    dyn.set_masks(g[tag + ".mask"])
    aux, std, chg = (float(v) for v in g[tag + ".weights"])
    model = GaugeModel(dict(time_size=T, space_size=X, num_samples=N, metric=metric, loss_scale=float(g[tag + ".loss_scale"]),
                            aux_weight=aux, std_weight=std, charge_weight=chg), lattice=lat, dynamics=dyn)
    return model, (T, X, N, M, bool(hmc), metric)


def test_loss_value_against_reference_goldens(golden):
    from l2hmc_b200 import _lib
    import ctypes
    g = golden("train_cases")
    for tag in tags(g, "loss"):
        model, (T, X, N, M, hmc, metric) = build_model(g, tag)
        for nm in ("x", "x_prop", "z"):
            close(model._calc_top_charges(g[f"{tag}.{nm}"], fft=True), g[f"{tag}.qfft.{nm}"], rtol=1e-5, atol=2e-5)
            close(model._calc_top_charges(g[f"{tag}.{nm}"], fft=False), g[f"{tag}.q.{nm}"], rtol=1e-5, atol=2e-5)
        # the loss kernel on the reference's own proposals (isolates the loss arithmetic)
        dev = {k: torch.as_tensor(g[f"{tag}.{k}"]).cuda().contiguous() for k in ("x", "x_prop", "z", "x_out", "px", "pz")}
        prm = _lib.LossParams()
        prm.metric, prm.loss_scale = _lib.METRICS[metric], float(g[tag + ".loss_scale"])
        prm.aux_weight, prm.std_weight, prm.charge_weight = (float(v) for v in g[tag + ".weights"])
        per_chain, x_dq = torch.empty(N, device="cuda"), torch.empty(N, dtype=torch.int32, device="cuda")
        loss = torch.empty(1, device="cuda")
        _lib.check(_lib.get_lib().l2hmc_u1_loss(ctypes.byref(prm), *(_lib.dptr(dev[k]) for k in
                                                                   ("x", "x_prop", "z", "x_out", "px", "pz")),
                                                N, T, X, _lib.dptr(per_chain), _lib.dptr(x_dq, torch.int32),
                                                _lib.dptr(loss), None))
        _, ref_chain, ref_dq = tro.loss_terms(g[tag + ".x"], g[tag + ".x_prop"], g[tag + ".z"], g[tag + ".px"],
                                              g[tag + ".pz"], g[tag + ".x_out"], T, X, metric, prm.loss_scale,
                                              prm.aux_weight, prm.std_weight, prm.charge_weight)
        scale = float(ref_chain.abs().max())
        close(per_chain, ref_chain, rtol=3e-5, atol=3e-5 * scale)
        close(loss, g[tag + ".loss"], rtol=3e-5, atol=3e-5 * scale)
        # x_dq = int32(|Q(x) - Q(x_out)|): the float is an integer up to summation-order rounding and tf.cast
        # TRUNCATES it (0.9999998 -> 0, 1.0000002 -> 1), so the reference's own value flips with the reduction order
        # (its CPU and GPU kernels would disagree the same way).  Meaningful and checked: the rounded difference
        # is identical, and the truncated one is the rounded one or one less.
        dq_ref = tro.top_charges_diff(g[tag + ".x"], g[tag + ".x_out"], T, X).numpy()
        got = x_dq.cpu().numpy()
        assert np.all((got == np.rint(dq_ref)) | (got == np.rint(dq_ref) - 1))
        assert np.abs(got - g[tag + ".x_dq"]).max() <= 1
        # end to end: both chains through the CUDA dynamics, then the loss
        loss2, x_out, px, x_dq2 = model._calc_loss(
            g[tag + ".x"], float(g[tag + ".beta"]),
            draws_x=(g[tag + ".v_f"], g[tag + ".v_b"], g[tag + ".u_dir"], g[tag + ".u_acc"]), z=g[tag + ".z"],
            draws_z=(g[tag + ".vz_f"], g[tag + ".vz_b"], g[tag + ".uz_dir"], g[tag + ".uz_acc"]))
        close(px, g[tag + ".px"], rtol=1e-3, atol=1e-6)
        close(x_out, g[tag + ".x_out"], rtol=1e-4, atol=2e-4)
        close(loss2, g[tag + ".loss"], rtol=5e-3, atol=5e-3 * scale)    # 1/x_std amplifies the proposals' rounding
        # |Q(x) - Q(x_out)| is an integer up to rounding, and tf.cast truncates it: 0.9999998 -> 0, 1.0000001 -> 1.
        # On identical inputs (above) the truncation is reproduced bit for bit; through two different leapfrog
        # implementations only the rounded value is meaningful.
        dq_ref = tro.top_charges_diff(g[tag + ".x"], g[tag + ".x_out"], T, X).numpy()
        dq_gpu = (model._calc_top_charges(g[tag + ".x"]) - model._calc_top_charges(x_out)).abs().cpu().numpy()
        assert np.array_equal(np.rint(dq_gpu), np.rint(dq_ref))
        assert np.abs(x_dq2.cpu().numpy() - g[tag + ".x_dq"]).max() <= 1
        model.dynamics.raise_if_nonfinite()


@pytest.mark.parametrize("T,X,N,metric", [(32, 32, 300, "cos_diff"), (16, 16, 77, "l2"), (5, 7, 33, "cos")])
def test_loss_kernel_against_oracle_larger(T, X, N, metric):
    from l2hmc_b200 import _lib
    import ctypes
    rng = np.random.default_rng(N)
    D = 2 * T * X
    x = rng.uniform(0, 2 * np.pi, (N, D)).astype(np.float32)
    xp = (x + 0.3 * rng.standard_normal((N, D))).astype(np.float32)
    z = rng.standard_normal((N, D)).astype(np.float32)
    xo = np.where(rng.uniform(size=(N, 1)) > 0.4, xp, x).astype(np.float32)
    px, pz = (rng.uniform(0.05, 1, N).astype(np.float32) for _ in range(2))
    prm = _lib.LossParams()
    prm.metric, prm.loss_scale, prm.aux_weight, prm.std_weight, prm.charge_weight = _lib.METRICS[metric], 0.7, 0.8, 1.1, 0.9
    dev = [torch.as_tensor(a).cuda() for a in (x, xp, z, xo, px, pz)]
    per_chain, x_dq = torch.empty(N, device="cuda"), torch.empty(N, dtype=torch.int32, device="cuda")
    loss = torch.empty(1, device="cuda")
    _lib.check(_lib.get_lib().l2hmc_u1_loss(ctypes.byref(prm), *(_lib.dptr(a) for a in dev), N, T, X,
                                            _lib.dptr(per_chain), _lib.dptr(x_dq, torch.int32), _lib.dptr(loss), None))
    ref_loss, ref_chain, ref_dq = tro.loss_terms(x, xp, z, px, pz, xo, T, X, metric, 0.7, 0.8, 1.1, 0.9)
    ref64 = tro.loss_terms(*(torch.as_tensor(a, dtype=torch.float64) for a in (x, xp, z, px, pz, xo)), T, X, metric,
                           0.7, 0.8, 1.1, 0.9)[1] if False else None
    scale = float(ref_chain.abs().max())
    close(per_chain, ref_chain, rtol=1e-4, atol=1e-5 * scale)
    close(loss, ref_loss, rtol=1e-4, atol=1e-5 * scale)
    dq = tro.top_charges_diff(x, xo, T, X).numpy()          # truncated |dQ|: the rounded value or one less (see above)
    got = x_dq.cpu().numpy()
    assert np.all((got == np.rint(dq)) | (got == np.rint(dq) - 1))


def test_run_stats_against_reference_goldens(golden):
    from l2hmc_b200.gauge_model import GaugeModel
    g = golden("train_cases")
    model = GaugeModel(dict(time_size=4, space_size=4, num_samples=2, hmc=True))
    for tag in tags(g, "stats"):
        steps, N, tf_ = (int(v) for v in g[tag + ".meta"])
        obs = tuple({s: arr[s] for s in range(steps)} for arr in (g[tag + ".actions"], g[tag + ".plaqs"],
                                                                 g[tag + ".charges"]))
        a_st, p_st, q_st, s_st, probs = model.calc_observables_stats(obs, therm_frac=tf_)
        for nm, st in (("actions", a_st), ("plaqs", p_st), ("charges", q_st), ("suscept", s_st)):
            np.testing.assert_allclose(st[0], g[f"{tag}.{nm}.mean"], rtol=2e-6, atol=1e-7)    # reference sums in fp32
            np.testing.assert_allclose(st[1], g[f"{tag}.{nm}.err"], rtol=2e-5, atol=1e-7)
        assert list(probs.keys()) == list(g[tag + ".prob.keys"])
        np.testing.assert_allclose(list(probs.values()), g[tag + ".prob.vals"], rtol=1e-12)


def test_run_stats_streaming_on_real_observables_and_truncation():
    """RunStats fed by the MC loop's device tensors == the oracle on the stacked history; non-integer charges are
    truncated toward zero like np.array(..., dtype=int) (gauge_model.py:1496)."""
    import l2hmc_b200 as pkg
    from l2hmc_b200.gauge_model import RunStats
    T = X = 8
    N, steps = 96, 30
    lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=N)
    dyn = pkg.GaugeDynamics(lat, hmc=True, num_steps=5, eps=0.2, seed=3)
    x = torch.rand(N, 2 * T * X, device="cuda") * (2 * np.pi)
    stats = RunStats(N, steps, therm_frac=10)
    hist = []
    for _ in range(steps):
        x = dyn.apply_transition(x, 3.0, wrap=True)[3]
        a, p, qf, qi = lat.observables(x)
        stats.update(a, p, qf)                       # the run loop stores the float charge (gauge_model.py:1400-1410)
        hist.append((a.cpu().numpy(), p.cpu().numpy(), qf.cpu().numpy()))
    res = stats.finalize()
    ref = tro.calc_observables_stats(*(np.stack([h[i] for h in hist]) for i in range(3)), therm_frac=10)
    for st, nm in zip(res[:4], ("actions", "plaqs", "charges", "suscept")):
        np.testing.assert_allclose(st[0], ref[nm][0], rtol=2e-6, atol=1e-6)
        np.testing.assert_allclose(st[1], ref[nm][1], rtol=2e-5, atol=1e-6)
    assert list(res[4].keys()) == [int(k) for k in ref["charge_probabilities"].keys()]
    np.testing.assert_allclose(list(res[4].values()), list(ref["charge_probabilities"].values()), rtol=1e-12)


def test_mode_switch_counter_matches_tunneling_rate():
    from l2hmc_b200 import _lib, mog
    from oracle import mog_oracle as mogo
    N, steps = 4000, 25
    gmm = mog.GMM([[1., 0.], [-1., 0.]], [0.3 * np.eye(2)] * 2, [0.5, 0.5])
    np.random.seed(42)
    dyn = mog.Dynamics(2, gmm, trajectory_length=8, eps=0.15, hmc=True, seed=1)
    x = torch.randn(N, 2, device="cuda") * 0.3 + torch.tensor([1.0, 0.0], device="cuda")
    prev = torch.empty(N, dtype=torch.int32, device="cuda")
    sw = torch.zeros(N, dtype=torch.int32, device="cuda")
    traj = []
    for s in range(steps):
        o = dyn.mc_step(x)
        x = o["x_out"]
        traj.append(x.cpu().numpy())
        _lib.check(_lib.get_lib().l2hmc_mode_switches(_lib.dptr(o["mode"], torch.int32), _lib.dptr(prev, torch.int32),
                                                      _lib.dptr(sw, torch.int32), int(s > 0), N, None))
    rate = sw.cpu().numpy() / (steps - 1)
    ref = mogo.calc_tunneling_rate(np.stack(traj), np.stack(gmm.mus), 2)
    # nearest-mean ties aside (x_0 == 0 exactly never happens), identical
    assert np.array_equal(rate, ref) and rate.mean() > 0
