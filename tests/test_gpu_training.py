"""SURVEY section 8(f) ranks 1-2 and 4 on the GPU: the backward pass (csrc/train.cu + l2hmc_b200/training.py), Adam and
the saved state.

Oracles: (i) the gradients the UNMODIFIED reference produced with tf.gradients on the TF stand-in
(tests/golden/train_cases.npz, keys `<tag>.grad.*`); (ii) reverse-mode autodiff through the CPU oracle
(oracle/train_oracle.py::loss_and_grads, itself pinned to (i)); (iii) torch-CPU autograd of the oracle's formulas for
each adjoint kernel on its own.  Tolerances: gradients are sums of O(N*D*M) fp32 products whose order differs between
the graph and the kernels, so whole-pass gradients are compared at 5e-4 of each tensor's largest entry (measured: <= 2.2e-5, profiles/r1_train_grad_errors.txt) and single kernels at 1e-5 .. 1e-4 (stated per check)."""
import ctypes
import math
import os

import numpy as np
import pytest
import torch

from oracle import train_oracle as tro
from oracle import u1_oracle as orc

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def cu(a, dtype=torch.float32):
    return torch.as_tensor(np.ascontiguousarray(a)).to(device="cuda", dtype=dtype).contiguous()


def close_scaled(got, want, tol=2e-4, what=""):
    got = got.detach().cpu().numpy().astype(np.float64) if isinstance(got, torch.Tensor) else np.asarray(got, np.float64)
    want = want.detach().cpu().numpy().astype(np.float64) if isinstance(want, torch.Tensor) else np.asarray(want, np.float64)
    scale = max(np.abs(want).max(), 1e-30)
    err = np.abs(got.reshape(want.shape) - want).max()
    if os.path.isdir(os.path.join(ROOT, "gpurun_out")):
        with open(os.path.join(ROOT, "gpurun_out", "train_grad_errors.txt"), "a") as fh:
            fh.write(f"{what}: err/scale = {err / scale:.3e} (scale {scale:.3e}, tol {tol})\n")
    assert err <= tol * scale, f"{what}: max error {err:.3e} vs scale {scale:.3e} (tol {tol})"


# --------------------------------------------------------------------------------------------- SGEMM
@pytest.mark.parametrize("M,N,K,ta,tb", [(70, 130, 50, False, False), (64, 64, 16, False, True), (33, 257, 100, True, False),
                                         (1, 200, 777, True, False), (3, 64, 1000, True, True), (256, 512, 384, False, True),
                                         # M <= 4 over a long K with row-major B: the weighted-column-sum kernel
                                         (2, 1000, 5003, True, False), (4, 130, 2560, False, False),
                                         (1, 64, 40960, True, False),
                                         # both operands K-contiguous: the chunked dense kernel (clustered K chunks for
                                         # few tiles, one CTA walking the chunks for many), ragged in every dimension
                                         (130, 70, 100, False, True), (70, 33, 1030, False, True),
                                         (9700, 200, 300, False, True), (5, 7, 3, False, True),
                                         # A^T . B with both operands row-major (small weight gradients over a long K)
                                         (512, 128, 2560, True, False), (130, 70, 1000, True, False),
                                         (9, 33, 5000, True, False),
                                         # ... with M, N multiples of 4 and aligned rows: 16-byte operand copies
                                         (64, 64, 37, True, False), (132, 68, 1000, True, False)])
def test_sgemm_forms(M, N, K, ta, tb):
    from l2hmc_b200 import _lib
    from l2hmc_b200.training import _Gemm, _ptr
    rng = np.random.default_rng(M * 7 + N)
    A = rng.standard_normal((K, M) if ta else (M, K)).astype(np.float32)
    B = rng.standard_normal((N, K) if tb else (K, N)).astype(np.float32)
    ref = (A.T if ta else A).astype(np.float64) @ (B.T if tb else B).astype(np.float64)
    dA, dB = cu(A), cu(B)
    a_rs, a_cs = (1, M) if ta else (K, 1)
    b_rs, b_cs = (1, K) if tb else (N, 1)
    gemm = _Gemm(torch.device("cuda"))
    for ksplit in (1, 4, 0):
        C = torch.full((M, N), 7.0, device="cuda")
        gemm(M, N, [(_ptr(dA), a_rs, a_cs, _ptr(dB), b_rs, b_cs, K)], _ptr(C), N, ksplit=ksplit)
        close_scaled(C, ref, 1e-5, f"sgemm ksplit={ksplit}")
    # accumulate + ReLU gate + per-row column mask sets
    C0 = rng.standard_normal((M, N)).astype(np.float32)
    gate = rng.standard_normal((M, N)).astype(np.float32)
    m0, m1 = (rng.uniform(size=N) > 0.5).astype(np.float32), (rng.uniform(size=N) > 0.5).astype(np.float32)
    rows = (rng.uniform(size=M) > 0.5).astype(np.uint8)
    scale = np.where(rows[:, None] != 0, 1 - m1[None, :], m0[None, :])
    want = C0 + ref * (gate > 0) * scale
    tg, t0, t1, tr = cu(gate), cu(m0), cu(m1), cu(rows, torch.uint8)
    for ksplit in (1, 3, 0):            # split-K applies the same epilogue in its second stage
        C = cu(C0)
        gemm(M, N, [(_ptr(dA), a_rs, a_cs, _ptr(dB), b_rs, b_cs, K)], _ptr(C), N, accumulate=True, relu_gate=_ptr(tg),
             ldg=N, col_scale=_ptr(t0), invert=False, row_set=_lib.dptr(tr, torch.uint8), col_scale_alt=_ptr(t1),
             invert_alt=True, ksplit=ksplit)
        torch.cuda.synchronize()
        close_scaled(C, want, 1e-5, f"sgemm epilogue ksplit={ksplit}")


def test_sgemm_three_segments():
    from l2hmc_b200.training import _Gemm, _ptr
    rng = np.random.default_rng(5)
    R, D, H = 90, 48, 130
    dA = rng.standard_normal((R, 3 * D)).astype(np.float32)
    Ws = [rng.standard_normal((H, D)).astype(np.float32) for _ in range(3)]
    ref = sum(dA[:, s * D:(s + 1) * D].astype(np.float64) @ Ws[s].T.astype(np.float64) for s in range(3))
    tA, tW = cu(dA), [cu(w) for w in Ws]
    for ksplit in (1, 2, 0):
        C = torch.empty(R, H, device="cuda")
        _Gemm(torch.device("cuda"))(R, H, [(_ptr(tA, s * D), 3 * D, 1, _ptr(tW[s]), 1, D, D) for s in range(3)], _ptr(C), H,
                                    ksplit=ksplit)
        close_scaled(C, ref, 1e-5, f"3-segment NT gemm ksplit={ksplit}")


# --------------------------------------------------------------------------------------------- update adjoints
def _oracle_update(kind, dirn, x, v, g, a_s, a_t, a_q, cs, cq, m, eps):
    """The reference's sub-update (oracle formulas, gauge_dynamics.py:486-590) as a torch function of leaves."""
    S = torch.tanh(a_s) * torch.exp(cs)
    T = a_t
    Q = a_q * torch.exp(cq)
    mi = 1. - m
    if kind == "m":
        if dirn > 0:
            sc = S * (0.5 * eps)
            out = v * torch.exp(sc) - 0.5 * eps * (torch.exp(Q * eps) * g - T)
        else:
            sc = S * (-0.5 * eps)
            out = torch.exp(sc) * (v + 0.5 * eps * (torch.exp(Q * eps) * g - T))
        return out, sc.sum(dim=1), (S, T, Q)
    if dirn > 0:
        sc = S * eps
        tmp = x * torch.exp(sc) + eps * (torch.exp(Q * eps) * v + T)
    else:
        sc = S * (-eps)
        tmp = torch.exp(sc) * (x - eps * (torch.exp(Q * eps) * v + T))
    return m * x + mi * tmp, (mi * sc).sum(dim=1), (S, T, Q)


@pytest.mark.parametrize("kind", ["m", "p"])
@pytest.mark.parametrize("with_nets", [True, False])
def test_update_bwd_against_autograd(kind, with_nets):
    from l2hmc_b200 import _lib
    rng = np.random.default_rng(11 + (kind == "p") + 2 * with_nets)
    N, D, eps = 37, 72, 0.23
    f = lambda *s: torch.tensor(rng.standard_normal(s), dtype=torch.float64)
    x, v, g = f(N, D), f(N, D), f(N, D)
    a_s, a_t, a_q = (0.5 * f(N, D) for _ in range(3))
    if not with_nets:
        a_s, a_t, a_q = (torch.zeros(N, D, dtype=torch.float64) for _ in range(3))
    cs, cq = 0.1 * f(D), 0.1 * f(D)
    if not with_nets:
        cs, cq = torch.zeros(D, dtype=torch.float64), torch.zeros(D, dtype=torch.float64)
    m0 = torch.tensor((rng.uniform(size=D) > 0.5).astype(np.float64))
    m1 = torch.tensor((rng.uniform(size=D) > 0.5).astype(np.float64))
    rows = (rng.uniform(size=N) > 0.5)
    up_out, up_ld = f(N, D), f(N)
    up_other = f(N, D)        # gradient already sitting on the variable the update passes through
    leaves = [t.clone().requires_grad_(True) for t in (x, v, g, a_s, a_t, a_q, cs, cq)]
    epst = torch.tensor(eps, dtype=torch.float64, requires_grad=True)
    lx, lv, lg, las, lat, laq, lcs, lcq = leaves
    total = 0.
    S_all, T_all, Q_all = torch.zeros(N, D, dtype=torch.float64), torch.zeros(N, D, dtype=torch.float64), \
        torch.zeros(N, D, dtype=torch.float64)
    for sel, dirn, mask in ((~rows, +1, m0), (rows, -1, 1. - m1)):
        idx = torch.tensor(np.nonzero(sel)[0])
        out, ld, (S, T, Q) = _oracle_update(kind, dirn, lx[idx], lv[idx], lg[idx], las[idx], lat[idx], laq[idx], lcs, lcq,
                                            mask, epst)
        total = total + (out * up_out[idx]).sum() + (ld * up_ld[idx]).sum()
        S_all[idx], T_all[idx], Q_all[idx] = S.detach(), T.detach(), Q.detach()
    grads = torch.autograd.grad(total, leaves + [epst], allow_unused=True)
    gx_ref, gv_ref, gg_ref, gas, gat, gaq, gcs, gcq, geps = [None if t is None else t for t in grads]

    u = _lib.UpdateBwdArgs()
    u.position, u.N, u.D, u.eps, u.dir, u.dir_alt = int(kind == "p"), N, D, eps, 1, -1
    rs = cu(rows.astype(np.uint8), torch.uint8)
    u.row_set = _lib.dptr(rs, torch.uint8)
    tx, tv, tg = cu(x.numpy()), cu(v.numpy()), cu(g.numpy())
    u.x, u.v, u.g = _lib.dptr(tx), _lib.dptr(tv), _lib.dptr(tg)
    keep = [cu(S_all.numpy()), cu(T_all.numpy()), cu(Q_all.numpy()), cu(cs.numpy()), cu(cq.numpy())]
    dA, ps, pq = torch.zeros(N, 3 * D, device="cuda"), torch.zeros(N, D, device="cuda"), torch.zeros(N, D, device="cuda")
    if with_nets:
        u.S, u.T, u.Q, u.coeff_s, u.coeff_q = (_lib.dptr(t) for t in keep)
        u.dA, u.ps, u.pq = _lib.dptr(dA), _lib.dptr(ps), _lib.dptr(pq)
    tm0, tm1 = cu(m0.numpy()), cu(m1.numpy())
    u.mask, u.invert, u.mask_alt, u.invert_alt = _lib.dptr(tm0), 0, _lib.dptr(tm1), 1
    gld = cu(up_ld.numpy())
    u.gld = _lib.dptr(gld)
    deps = torch.zeros(N, device="cuda")
    u.deps = _lib.dptr(deps)
    dg = torch.zeros(N, D, device="cuda")
    if kind == "m":
        gv = cu(up_out.numpy())
        gx = cu(up_other.numpy())
        u.gv, u.dg = _lib.dptr(gv), _lib.dptr(dg)
    else:
        gx = cu(up_out.numpy())
        gv = cu(up_other.numpy())
        u.gx, u.gv = _lib.dptr(gx), _lib.dptr(gv)
    _lib.check(_lib.get_lib().l2hmc_update_bwd(ctypes.byref(u), None))
    torch.cuda.synchronize()
    if kind == "m":
        close_scaled(gv, gv_ref, 2e-5, "d/dv")
        close_scaled(dg, gg_ref, 2e-5, "d/dg")
    else:
        close_scaled(gx, gx_ref, 2e-5, "d/dx")
        close_scaled(gv, up_other + gv_ref, 2e-5, "d/dv (accumulated)")
    close_scaled(deps.sum(), geps, 5e-5, "d/deps")
    if with_nets:
        close_scaled(dA[:, :D], gas, 2e-5, "d/da_s")
        close_scaled(dA[:, D:2 * D], gat, 2e-5, "d/da_t")
        close_scaled(dA[:, 2 * D:], gaq, 2e-5, "d/da_q")
        close_scaled(ps.sum(dim=0), gcs, 5e-5, "d/dcoeff_scale")
        close_scaled(pq.sum(dim=0), gcq, 5e-5, "d/dcoeff_transformation")


@pytest.mark.parametrize("T,X,N", [(8, 8, 50), (6, 10, 7), (16, 16, 33), (32, 32, 9)])
def test_force_hvp_against_autograd(T, X, N):
    from l2hmc_b200 import _lib
    rng = np.random.default_rng(T * X + N)
    D, beta = 2 * T * X, 2.7
    x = torch.tensor(rng.uniform(0, 2 * np.pi, (N, D)), dtype=torch.float64, requires_grad=True)
    w = torch.tensor(rng.standard_normal((N, D)), dtype=torch.float64)
    out0 = rng.standard_normal((N, D)).astype(np.float32)
    (ref,) = torch.autograd.grad((orc.grad_potential(x, beta, T, X) * w).sum(), x)
    out, tx, tw = cu(out0), cu(x.detach().numpy()), cu(w.numpy())
    _lib.check(_lib.get_lib().l2hmc_u1_force_hvp(_lib.dptr(tx), _lib.dptr(tw), beta, N, T, X, _lib.dptr(out), None))
    torch.cuda.synchronize()
    close_scaled(out, torch.tensor(out0, dtype=torch.float64) + ref, 1e-5, "force HVP")


@pytest.mark.parametrize("metric", ["l1", "l2", "cos", "cos2", "cos_diff"])
def test_loss_bwd_against_autograd(metric):
    from l2hmc_b200 import _lib
    T, X, N = 8, 6, 41
    D = 2 * T * X
    rng = np.random.default_rng(len(metric) * 13 + ord(metric[-1]))
    x = torch.tensor(rng.uniform(0, 2 * np.pi, (N, D)), dtype=torch.float64)
    xp = (x + 0.4 * torch.tensor(rng.standard_normal((N, D)))).requires_grad_(True)
    z = torch.tensor(rng.standard_normal((N, D)), dtype=torch.float64)
    px = torch.tensor(rng.uniform(0.1, 1, N), dtype=torch.float64, requires_grad=True)
    pz = torch.tensor(rng.uniform(0.1, 1, N), dtype=torch.float64, requires_grad=True)
    ls, aw, sw, cw = 0.7, 0.8, 1.1, 0.9

    def qfft(a):        # gauge_model.py:699-713 with fft=True, in double
        return tro.project_angle_approx(orc.plaq_sums(a, T, X)).sum(dim=(1, 2)) / (2 * np.pi)

    def loss64():       # gauge_model.py:762-801
        m = tro.metric_fn(metric)
        e = 1e-3
        qd = lambda a, b: (qfft(a) - qfft(b)).abs()
        x_std = m(x, xp).sum(dim=1) * px + e
        z_std = aw * (m(z, xp).sum(dim=1) * pz + e)
        std = (ls * (1. / x_std + 1. / z_std) - (x_std + z_std) / ls) * sw
        return (std + cw * ((px * qd(x, xp) + e) + aw * (pz * qd(z, xp) + e))).mean()
    gxp, gpx, gpz = torch.autograd.grad(loss64(), [xp, px, pz])
    prm = _lib.LossParams()
    prm.metric, prm.loss_scale, prm.aux_weight, prm.std_weight, prm.charge_weight = _lib.METRICS[metric], ls, aw, sw, cw
    o_xp, o_px, o_pz = torch.empty(N, D, device="cuda"), torch.empty(N, device="cuda"), torch.empty(N, device="cuda")
    dev = [cu(t.detach().numpy()) for t in (x, xp, z, px, pz)]
    _lib.check(_lib.get_lib().l2hmc_u1_loss_bwd(ctypes.byref(prm), *(_lib.dptr(t) for t in dev), N, T, X, _lib.dptr(o_xp),
                                               _lib.dptr(o_px), _lib.dptr(o_pz), None))
    torch.cuda.synchronize()
    close_scaled(o_px, gpx, 1e-4, "d/dpx")
    close_scaled(o_pz, gpz, 1e-4, "d/dpz")
    close_scaled(o_xp, gxp, 1e-4, "d/dx_prop")


def test_accept_bwd_and_traj_init():
    from l2hmc_b200 import _lib
    rng = np.random.default_rng(3)
    N, D = 300, 40
    h0 = torch.tensor(rng.normal(50, 2, N), dtype=torch.float64)
    h1 = (h0 + torch.tensor(rng.normal(0, 1.5, N))).requires_grad_(True)
    sld = torch.tensor(rng.normal(0, 0.3, N), dtype=torch.float64, requires_grad=True)
    gp = torch.tensor(rng.standard_normal(N), dtype=torch.float64)
    p = torch.exp(torch.minimum(h0 - h1 + sld, torch.zeros_like(h0)))
    (g_sld,) = torch.autograd.grad((p * gp).sum(), sld)
    a = torch.empty(N, device="cuda")
    # fp32 inputs: recompute the reference on the rounded values so the sign of (h0 - h1 + sld) agrees
    h0f, h1f, sldf = (t.detach().numpy().astype(np.float32) for t in (h0, h1, sld))
    dev = [cu(t) for t in (h0f, h1f, sldf, gp.numpy())]
    _lib.check(_lib.get_lib().l2hmc_accept_prob_bwd(*(_lib.dptr(t) for t in dev), N, _lib.dptr(a), None))
    torch.cuda.synchronize()
    arg = (h0f - h1f) + sldf
    want = np.where(arg <= 0, gp.numpy() * np.exp(np.minimum(arg.astype(np.float64), 0)), 0.)
    close_scaled(a, want, 1e-5, "accept-prob adjoint")
    force, v, gxp = (rng.standard_normal((N, D)).astype(np.float32) for _ in range(3))
    gx, gv = torch.empty(N, D, device="cuda"), torch.empty(N, D, device="cuda")
    dev = [cu(t) for t in (force, v, gxp)]
    _lib.check(_lib.get_lib().l2hmc_traj_bwd_init(_lib.dptr(a), *(_lib.dptr(t) for t in dev), N, D, _lib.dptr(gx),
                                                 _lib.dptr(gv), None))
    torch.cuda.synchronize()
    an = a.cpu().numpy().astype(np.float64)[:, None]
    close_scaled(gx, gxp - an * force, 1e-6, "gx init")
    close_scaled(gv, -an * v, 1e-6, "gv init")


# --------------------------------------------------------------------------------------------- whole backward pass
def _tags(g, prefix):
    return sorted({k.rsplit(".meta", 1)[0] for k in g if k.startswith(prefix) and k.endswith(".meta")})


def test_gradients_against_reference_goldens(golden):
    """tf.gradients(loss, dynamics.variables) of the unmodified reference (gauge_model.py:823)."""
    from test_gpu_train_rows import build_model
    g = golden("train_cases")
    checked = 0
    for tag in _tags(g, "loss"):
        model, (T, X, N, M, hmc, metric) = build_model(g, tag)
        loss, grads, x_out, px, x_dq = model._calc_loss_and_grads(
            g[tag + ".x"], float(g[tag + ".beta"]),
            draws_x=(g[tag + ".v_f"], g[tag + ".v_b"], g[tag + ".u_dir"], g[tag + ".u_acc"]), z=g[tag + ".z"],
            draws_z=(g[tag + ".vz_f"], g[tag + ".vz_b"], g[tag + ".uz_dir"], g[tag + ".uz_acc"]))
        torch.cuda.synchronize()
        model.dynamics.raise_if_nonfinite()
        scale = abs(float(g[tag + ".loss"]))
        assert abs(float(loss) - float(g[tag + ".loss"])) <= 5e-3 * max(scale, 1.), tag
        np.testing.assert_allclose(px.cpu().numpy(), g[tag + ".px"], rtol=1e-3, atol=1e-6)
        for name, gr in grads.items():
            want = g[f"{tag}.grad.{name}"]
            # the loss divides by x_std ~ O(1e-3 .. 1): its gradient inherits the 5e-3 the loss value is held to
            close_scaled(gr, want, 5e-4, f"{tag} d loss / d {name}")
            checked += 1
    assert checked >= 3 + 2 * 33 + 49


@pytest.mark.parametrize("arch,T,X,N,M,metric", [("generic", 8, 8, 16, 5, "cos_diff"), ("generic", 4, 6, 9, 3, "l2"),
                                                  ("conv3D", 8, 8, 12, 2, "cos_diff"), ("conv3D", 32, 32, 5, 1, "cos_diff")])
def test_gradients_against_oracle_config2_shapes(arch, T, X, N, M, metric):
    """Config-2 network shapes (8x8, H = 4D = 512, M = 5) against autodiff through the CPU oracle (fp32); the conv3D
    front end at 8x8 and at BASELINE config 4's lattice (32x32, F = 32: the adjoint runs from the forward pass's
    routing tape, the only path that fits that shape)."""
    import l2hmc_b200 as pkg
    from l2hmc_b200.gauge_model import GaugeModel
    rng = np.random.default_rng(N + M)
    D, eps, beta = 2 * T * X, 0.2, 2.5
    weights = {}
    for net, factor in (("XNet", 2.), ("VNet", 1.)):
        w = orc.init_net_weights(arch, T, X, factor, rng)
        w["coeff_scale"] = (0.1 * rng.standard_normal((1, D))).astype(np.float32)
        w["coeff_transformation"] = (0.1 * rng.standard_normal((1, D))).astype(np.float32)
        for k in list(w):
            if k.endswith(".bias"):
                w[k] = (0.01 * rng.standard_normal(w[k].shape)).astype(np.float32)
        weights[net] = w
    mask = orc.construct_masks(D, M)
    x = rng.uniform(0, 2 * np.pi, (N, D)).astype(np.float32)
    z = rng.standard_normal((N, D)).astype(np.float32)
    dr = lambda: (rng.standard_normal((N, D)).astype(np.float32), rng.standard_normal((N, D)).astype(np.float32),
                  rng.uniform(size=N).astype(np.float32), rng.uniform(size=N).astype(np.float32))
    draws_x, draws_z = dr(), dr()
    lw = dict(aux_weight=1., std_weight=1., charge_weight=1.)
    od = orc.OracleDynamics(T, X, M, eps, hmc=False, network_arch=arch, mask=mask, xnet=weights["XNet"],
                            vnet=weights["VNet"])
    ref_loss, ref_grads = tro.loss_and_grads(od, x, beta, draws_x, z, draws_z, metric=metric, loss_scale=1., **lw)
    lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=N)
    dyn = pkg.GaugeDynamics(lat, hmc=False, network_arch=arch, num_steps=M, eps=eps, weights=weights)
    dyn.set_masks(mask)
    model = GaugeModel(dict(time_size=T, space_size=X, num_samples=N, metric=metric, loss_scale=1., **lw), lattice=lat,
                       dynamics=dyn)
    loss, grads, _, _, _ = model._calc_loss_and_grads(x, beta, draws_x=draws_x, z=z, draws_z=draws_z)
    torch.cuda.synchronize()
    assert abs(float(loss) - float(ref_loss)) <= 2e-3 * max(abs(float(ref_loss)), 1.)
    assert set(grads) == set(ref_grads)
    for name, gr in grads.items():
        close_scaled(gr, ref_grads[name], 5e-4, f"d loss / d {name}")
    # run to run: bit-identical (no atomics, fixed split-K order)
    loss2, grads2, _, _, _ = model._calc_loss_and_grads(x, beta, draws_x=draws_x, z=z, draws_z=draws_z)
    flat1 = model.engine.params.grad.clone()
    model._calc_loss_and_grads(x, beta, draws_x=draws_x, z=z, draws_z=draws_z)
    assert torch.equal(flat1, model.engine.params.grad)


def test_hmc_eps_gradient_finite_difference():
    """d loss / d eps of plain HMC against a central difference of the CUDA loss itself (fp32 loss, so a loose 2 %)."""
    import l2hmc_b200 as pkg
    from l2hmc_b200.gauge_model import GaugeModel
    T, X, N, M, beta = 8, 8, 64, 4, 3.0
    D = 2 * T * X
    rng = np.random.default_rng(9)
    x = rng.uniform(0, 2 * np.pi, (N, D)).astype(np.float32)
    z = rng.standard_normal((N, D)).astype(np.float32)
    dr = lambda: (rng.standard_normal((N, D)).astype(np.float32), rng.standard_normal((N, D)).astype(np.float32),
                  rng.uniform(size=N).astype(np.float32), rng.uniform(size=N).astype(np.float32))
    draws_x, draws_z = dr(), dr()

    def run(eps):
        lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=N)
        dyn = pkg.GaugeDynamics(lat, hmc=True, num_steps=M, eps=eps)
        model = GaugeModel(dict(time_size=T, space_size=X, num_samples=N, hmc=True), lattice=lat, dynamics=dyn)
        loss, grads, _, _, _ = model._calc_loss_and_grads(x, beta, draws_x=draws_x, z=z, draws_z=draws_z)
        return float(loss), float(grads["eps"][0])
    l0, g0 = run(0.2)
    h = 2e-3
    fd = (run(0.2 + h)[0] - run(0.2 - h)[0]) / (2 * h)
    od = orc.OracleDynamics(T, X, M, 0.2, hmc=True, dtype=torch.float64)
    _, ref = tro.loss_and_grads(od, torch.tensor(x, dtype=torch.float64), beta,
                                tuple(torch.tensor(a, dtype=torch.float64) for a in draws_x), torch.tensor(z, dtype=torch.float64),
                                tuple(torch.tensor(a, dtype=torch.float64) for a in draws_z))
    assert abs(g0 - float(ref["eps"])) <= 2e-3 * abs(float(ref["eps"])), (g0, float(ref["eps"]))
    assert abs(g0 - fd) <= 0.05 * abs(fd) + 1e-3, (g0, fd)


# --------------------------------------------------------------------------------------------- optimiser / state
def test_adam_against_restatement():
    from l2hmc_b200 import _lib
    rng = np.random.default_rng(1)
    n = 100003
    p = rng.standard_normal(n).astype(np.float32)
    m, v = np.zeros(n, np.float64), np.zeros(n, np.float64)
    ref = p.astype(np.float64)
    dp, dm, dv = cu(p), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    sumsq, scratch = torch.zeros(1, device="cuda"), torch.empty(1024, device="cuda")
    lib = _lib.get_lib()
    for t in range(1, 6):
        g = (rng.standard_normal(n) * 3).astype(np.float32)
        lr, b1, b2, e, clip = 1e-3 * 0.96 ** (t // 2), 0.9, 0.999, 1e-8, 50.0
        gg = g.astype(np.float64)
        scale = clip / max(np.sqrt((gg ** 2).sum()), clip)              # tf.clip_by_global_norm
        m, v, ref = tro.adam_update(ref, gg * scale, m, v, t, lr, b1, b2, e)
        dg = cu(g)
        _lib.check(lib.l2hmc_reduce_sum(_lib.dptr(dg), n, 1, _lib.dptr(sumsq), _lib.dptr(scratch), None))
        assert abs(float(sumsq) - (gg ** 2).sum()) <= 1e-5 * (gg ** 2).sum()
        lr_t = lr * math.sqrt(1 - b2 ** t) / (1 - b1 ** t)
        _lib.check(lib.l2hmc_adam_step(_lib.dptr(dp), _lib.dptr(dg), _lib.dptr(dm), _lib.dptr(dv), n, lr_t, b1, b2, e, 1.0,
                                       _lib.dptr(sumsq), clip, None))
    np.testing.assert_allclose(dp.cpu().numpy(), ref, rtol=2e-6, atol=2e-7)
    np.testing.assert_allclose(dm.cpu().numpy(), m, rtol=1e-5, atol=1e-7)


def test_train_loop_and_saved_state(tmp_path):
    """A few training steps at config-2 shapes: parameters move, the loss is finite, the schedules follow the
    reference's formulas, and restore_state() reproduces the next step bit for bit."""
    from l2hmc_b200.gauge_model import GaugeModel
    params = dict(time_size=8, space_size=8, num_samples=32, num_steps=3, eps=0.2, network_arch="generic", rand=True,
                  train_steps=100, lr_init=1e-3, lr_decay_steps=2, lr_decay_rate=0.5, beta_init=2., beta_final=4.)
    np.random.seed(42)
    model = GaugeModel(params)
    model.dynamics.seed = 7
    w0 = model.engine.params.flat.clone()
    samples, hist = model.train(4)
    assert len(hist) == 4 and all(math.isfinite(float(l)) for l in hist)
    assert not torch.equal(w0, model.engine.params.flat)
    assert model.global_step == 4
    assert abs(model.learning_rate() - 1e-3 * 0.5 ** 2) < 1e-12                # staircase decay (:934-939)
    assert abs(model.beta - float(orc.update_beta(3, 2., 4., 100))) < 1e-12     # 1/beta-linear annealing (:1039-1046)
    assert float(samples.min()) >= 0. and float(samples.max()) < 2 * math.pi + 1e-6
    assert abs(model.dynamics._eps() - 0.2) > 0                                 # eps is trained
    path = str(tmp_path / "state.pt")
    model.save_state(path, samples)
    out_a = model.train_step(samples, model.update_beta(4))
    np.random.seed(42)
    model2 = GaugeModel(params)
    model2.dynamics.seed = 7
    restored = model2.restore_state(path)
    assert torch.equal(restored, samples)
    out_b = model2.train_step(restored, model2.update_beta(4))
    assert torch.equal(out_a[1], out_b[1]) and float(out_a[0]) == float(out_b[0])
    assert torch.equal(model.engine.params.flat, model2.engine.params.flat)
    # inference after training uses the updated weights (packed tensor-core copies are rebuilt)
    xp, _, p, _ = model.dynamics.apply_transition(samples, 3.0)
    assert torch.isfinite(p).all()


# --------------------------------------------------------------------------------------------- conv front-end adjoint
# register-tiled kernels: F in {8, 16, 32}; X == F folds the index arithmetic; even X / even pooled X take the
# paired-position phases, odd ones the single-position phases; other F: the generic kernel
@pytest.mark.parametrize("T,X,F,N", [(8, 8, 8, 21), (6, 10, 4, 9), (16, 16, 16, 300), (4, 4, 3, 400), (8, 12, 8, 50),
                                     (6, 10, 16, 40), (8, 7, 8, 30), (8, 8, 32, 20), (32, 32, 32, 5), (5, 7, 6, 12)])
def test_conv3d_backward_against_autograd(T, X, F, N):
    """l2hmc_conv3d_backward vs torch autograd through the oracle's Conv3D/MaxPool3D stack (conv_net.py:247-260),
    with per-chain mask sets and more chains than partial rows (accumulation across CTA passes)."""
    from l2hmc_b200 import _lib
    rng = np.random.default_rng(T * 100 + X + F)
    D = 2 * T * X
    w = {"k1": rng.uniform(-0.4, 0.4, (3, 3, 2, 1, F)), "b1": rng.uniform(-0.1, 0.1, F),
         "k2": rng.uniform(-0.3, 0.3, (2, 2, 2, F, 2 * F)), "b2": rng.uniform(-0.1, 0.1, 2 * F)}
    leaves = {k: torch.tensor(v, dtype=torch.float64, requires_grad=True) for k, v in w.items()}
    x = torch.tensor(rng.standard_normal((N, D)), dtype=torch.float64, requires_grad=True)
    m0 = (rng.uniform(size=D) > 0.5).astype(np.float64)
    m1 = (rng.uniform(size=D) > 0.5).astype(np.float64)
    rows = rng.uniform(size=N) > 0.5
    mrow = torch.tensor(np.where(rows[:, None], 1 - m1[None, :], m0[None, :]))
    h = (mrow * x).reshape(N, T, X, 2, 1)
    h = orc._maxpool3d_same(orc._conv3d_same_relu(h, leaves["k1"], leaves["b1"]))
    h = orc._maxpool3d_same(orc._conv3d_same_relu(h, leaves["k2"], leaves["b2"]))
    feat = h.reshape(N, -1)
    K = feat.shape[1]
    gf = torch.tensor(rng.standard_normal((N, K)), dtype=torch.float64)
    gx_ref, g1, gb1, g2, gb2 = torch.autograd.grad((feat * gf).sum(), [x, leaves["k1"], leaves["b1"], leaves["k2"], leaves["b2"]])
    lib = _lib.get_lib()
    assert K == lib.l2hmc_conv3d_num_features(T, X, F)
    dev = {k: cu(v) for k, v in w.items()}
    cw = _lib.Conv3dWeights()
    cw.w1, cw.b1, cw.w2, cw.b2, cw.F = _lib.dptr(dev["k1"]), _lib.dptr(dev["b1"]), _lib.dptr(dev["k2"]), _lib.dptr(dev["b2"]), F
    prows, pcols = lib.l2hmc_conv3d_bwd_rows(N), lib.l2hmc_conv3d_bwd_cols(F)
    assert pcols == 18 * F + F + 16 * F * F + 2 * F
    part = torch.zeros(prows, pcols, device="cuda")
    g_in0 = rng.standard_normal((N, D)).astype(np.float32)
    tx, tg, tgin, tm0, tm1, trs = cu(x.detach().numpy()), cu(gf.numpy()), cu(g_in0), cu(m0), cu(m1), cu(rows.astype(np.uint8), torch.uint8)
    a = _lib.Conv3dBwdArgs()
    a.w, a.in_, a.N, a.T, a.X = ctypes.pointer(cw), _lib.dptr(tx), N, T, X
    a.mask, a.mask_invert, a.mask_alt, a.mask_invert_alt = _lib.dptr(tm0), 0, _lib.dptr(tm1), 1
    a.row_set = _lib.dptr(trs, torch.uint8)
    a.g_feat, a.g_in, a.partial = _lib.dptr(tg), _lib.dptr(tgin), _lib.dptr(part)
    # recompute path (no tape); it refuses the shapes whose activations do not fit shared memory (32x32 with F = 32)
    rc = lib.l2hmc_conv3d_backward(ctypes.byref(a), None)
    recompute_ok = rc == 0
    assert recompute_ok or (rc == _lib.L2HMC_ERR_UNSUPPORTED and T * X * F >= 32 * 32 * 32)
    torch.cuda.synchronize()

    def check(tgin_, part_, what):
        tot = part_.double().sum(dim=0).cpu()
        close_scaled(tgin_, torch.tensor(g_in0, dtype=torch.float64) + gx_ref, 2e-5, f"conv d/d input ({what})")
        off = 0
        for nm, ref in (("w1", g1), ("b1", gb1), ("w2", g2), ("b2", gb2)):
            n = ref.numel()
            close_scaled(tot[off:off + n], ref.reshape(-1), 5e-5, f"conv d/d {nm} ({what})")
            off += n
    if recompute_ok:
        check(tgin, part, "recompute")
    # tape path: the forward pass records P1 and the two pools' argmax codes, the adjoint runs over the non-zeros only
    n_el = lib.l2hmc_conv3d_tape_elems(T, X, F)
    assert n_el == ((T + 1) // 2) * ((X + 1) // 2) * F
    p1 = torch.full((N, n_el), float("nan"), device="cuda")
    a1 = torch.full((N, n_el), 77, dtype=torch.uint8, device="cuda")
    a2 = torch.full((N, K), 77, dtype=torch.uint8, device="cuda")
    tape = _lib.Conv3dTape()
    tape.p1, tape.arg1, tape.arg2 = _lib.dptr(p1), _lib.dptr(a1, torch.uint8), _lib.dptr(a2, torch.uint8)
    feat_dev = torch.empty(N, K, device="cuda")
    _lib.check(lib.l2hmc_conv3d_features_tape(ctypes.byref(cw), _lib.dptr(tx), _lib.dptr(tm0), 0, _lib.dptr(tm1), 1,
                                              _lib.dptr(trs, torch.uint8), N, T, X, _lib.dptr(feat_dev), ctypes.byref(tape),
                                              None))
    close_scaled(feat_dev, feat.detach(), 2e-5, "features recorded with the tape")
    # the tape itself: P1 is the pooled conv-1 plane; codes are in range, dead exactly where the pooled value is zero
    h1 = orc._maxpool3d_same(orc._conv3d_same_relu((mrow * x).reshape(N, T, X, 2, 1).detach(), leaves["k1"].detach(),
                                                   leaves["b1"].detach())).reshape(N, -1)
    close_scaled(p1, h1, 2e-5, "tape p1")
    c1, c2 = a1.cpu().numpy(), a2.cpu().numpy()
    assert ((c1 < 8) | (c1 == 255)).all() and ((c2 < 4) | (c2 == 255)).all()
    assert ((c1 == 255) == (p1.cpu().numpy() == 0)).all() and ((c2 == 255) == (feat_dev.cpu().numpy() == 0)).all()
    part2 = torch.zeros(prows, pcols, device="cuda")
    tgin2 = cu(g_in0)
    a.g_in, a.partial, a.tape = _lib.dptr(tgin2), _lib.dptr(part2), ctypes.pointer(tape)
    _lib.check(lib.l2hmc_conv3d_backward(ctypes.byref(a), None))
    torch.cuda.synchronize()
    check(tgin2, part2, "tape")


def test_native_step_equals_python_walk(golden):
    """l2hmc_u1_train_grads (one C call, csrc/train_step.cu) issues the same launches as the Python-driven walk of
    training.py: loss, proposals and every gradient are bit-identical (plain HMC, generic and conv3D nets)."""
    from test_gpu_train_rows import build_model
    g = golden("train_cases")
    seen = set()
    for tag in _tags(g, "loss"):
        arch = tag.split(".")[1]
        if arch in seen:
            continue
        seen.add(arch)
        model, _ = build_model(g, tag)
        args = dict(draws_x=(g[tag + ".v_f"], g[tag + ".v_b"], g[tag + ".u_dir"], g[tag + ".u_acc"]), z=g[tag + ".z"],
                    draws_z=(g[tag + ".vz_f"], g[tag + ".vz_b"], g[tag + ".uz_dir"], g[tag + ".uz_acc"]))
        model.native_train = False
        loss_a, _, xo_a, px_a, dq_a = model._calc_loss_and_grads(g[tag + ".x"], float(g[tag + ".beta"]), **args)
        flat_a = model.engine.params.grad.clone()
        model.engine.params.grad.fill_(float("nan"))
        model.native_train = True
        loss_b, _, xo_b, px_b, dq_b = model._calc_loss_and_grads(g[tag + ".x"], float(g[tag + ".beta"]), **args)
        torch.cuda.synchronize()
        flat_b = model.engine.params.grad
        assert float(loss_a) == float(loss_b), tag
        assert torch.equal(xo_a, xo_b) and torch.equal(px_a, px_b) and torch.equal(dq_a, dq_b), tag
        # the arena's 16-byte padding between tensors is never written by either path
        for name in model.engine.params.names:
            va, vb = model.engine.params.view(flat_a, name), model.engine.params.view(flat_b, name)
            assert torch.equal(va, vb), f"{tag} {name}"
    assert seen == {"hmc", "generic", "conv3D"}


# --------------------------------------------------------------------------------------------- tcgen05 training GEMM
def _tc_gemm(M, N, segs, C, ldc, **kw):
    """l2hmc_tc_gemm through ctypes (same argument struct as l2hmc_sgemm)."""
    from l2hmc_b200 import _lib
    a = _lib.GemmArgs()
    a.M, a.N, a.nseg = M, N, len(segs)
    for i, (A, a_rs, a_cs, B, b_rs, b_cs, K) in enumerate(segs):
        s = a.seg[i]
        s.A, s.a_rs, s.a_cs, s.B, s.b_rs, s.b_cs, s.K = A, a_rs, a_cs, B, b_rs, b_cs, K
    a.C, a.ldc, a.accumulate = C, ldc, int(kw.get("accumulate", False))
    a.relu_gate, a.ldg = kw.get("relu_gate"), kw.get("ldg", 0)
    a.col_scale, a.col_scale_invert = kw.get("col_scale"), int(kw.get("invert", False))
    a.row_set, a.col_scale_alt, a.col_scale_invert_alt = kw.get("row_set"), kw.get("col_scale_alt"), int(kw.get("invert_alt", False))
    lib = _lib.get_lib()
    need = int(lib.l2hmc_tc_gemm_workspace_bytes(ctypes.byref(a)))
    ws = torch.empty(need // 4 + 64, device="cuda")
    _lib.check(lib.l2hmc_tc_gemm(ctypes.byref(a), ctypes.c_void_p(ws.data_ptr()), need, None))
    torch.cuda.synchronize()


@pytest.mark.parametrize("M,N,K,ta,tb", [(256, 512, 384, False, True), (130, 70, 100, False, False), (512, 1024, 4096, True, False),
                                         (1024, 512, 333, True, True), (4096, 1024, 512, False, True), (128, 64, 128, False, True),
                                         # long K, few tiles: split-K over the idle SMs (ragged last range; the deferred
                                         # weight-gradient shape of config 3)
                                         (256, 192, 8200, True, False), (1024, 512, 40960, True, False)])
def test_tc_gemm_forms(M, N, K, ta, tb):
    """The tcgen05 3xTF32 product in NN / NT / TN forms against a float64 product: fp32-level accuracy (2e-6 of scale)."""
    from l2hmc_b200 import _lib
    from l2hmc_b200.training import _ptr
    rng = np.random.default_rng(M + 3 * N + K)
    A = rng.standard_normal((K, M) if ta else (M, K)).astype(np.float32)
    B = rng.standard_normal((N, K) if tb else (K, N)).astype(np.float32)
    ref = (A.T if ta else A).astype(np.float64) @ (B.T if tb else B).astype(np.float64)
    dA, dB = cu(A), cu(B)
    a_rs, a_cs = (1, M) if ta else (K, 1)
    b_rs, b_cs = (1, K) if tb else (N, 1)
    C = torch.full((M, N), 7.0, device="cuda")
    _tc_gemm(M, N, [(_ptr(dA), a_rs, a_cs, _ptr(dB), b_rs, b_cs, K)], _ptr(C), N)
    close_scaled(C, ref, 2e-6, "tc gemm")
    C0 = rng.standard_normal((M, N)).astype(np.float32)
    gate = rng.standard_normal((M, N)).astype(np.float32)
    m0, m1 = (rng.uniform(size=N) > 0.5).astype(np.float32), (rng.uniform(size=N) > 0.5).astype(np.float32)
    rows = (rng.uniform(size=M) > 0.5).astype(np.uint8)
    want = C0 + ref * (gate > 0) * np.where(rows[:, None] != 0, 1 - m1[None, :], m0[None, :])
    C, tg, t0, t1, tr = cu(C0), cu(gate), cu(m0), cu(m1), cu(rows, torch.uint8)
    _tc_gemm(M, N, [(_ptr(dA), a_rs, a_cs, _ptr(dB), b_rs, b_cs, K)], _ptr(C), N, accumulate=True, relu_gate=_ptr(tg), ldg=N,
             col_scale=_ptr(t0), invert=False, row_set=_lib.dptr(tr, torch.uint8), col_scale_alt=_ptr(t1), invert_alt=True)
    close_scaled(C, want, 2e-6, "tc gemm epilogue")


def test_tc_gemm_three_segments():
    from l2hmc_b200.training import _ptr
    rng = np.random.default_rng(8)
    R, D, H = 300, 80, 200
    dA = rng.standard_normal((R, 3 * D)).astype(np.float32)
    Ws = [rng.standard_normal((H, D)).astype(np.float32) for _ in range(3)]
    ref = sum(dA[:, s * D:(s + 1) * D].astype(np.float64) @ Ws[s].T.astype(np.float64) for s in range(3))
    tA, tW = cu(dA), [cu(w) for w in Ws]
    C = torch.empty(R, H, device="cuda")
    _tc_gemm(R, H, [(_ptr(tA, s * D), 3 * D, 1, _ptr(tW[s]), 1, D, D) for s in range(3)], _ptr(C), H)
    close_scaled(C, ref, 2e-6, "tc 3-segment gemm")


def test_training_step_with_tcgen05_layers():
    """From 1024 stacked rows the forward layers and the dense backward products run on tcgen05 (3xTF32): native and
    Python-driven walks stay bit-identical, and the gradients match autodiff through the CPU oracle."""
    import l2hmc_b200 as pkg
    from l2hmc_b200.gauge_model import GaugeModel
    T, X, N, M, eps, beta = 8, 8, 512, 2, 0.15, 2.5
    D = 2 * T * X
    rng = np.random.default_rng(77)
    weights = {net: orc.init_net_weights("generic", T, X, f, rng) for net, f in (("XNet", 2.), ("VNet", 1.))}
    mask = orc.construct_masks(D, M)
    x = rng.uniform(0, 2 * np.pi, (N, D)).astype(np.float32)
    z = rng.standard_normal((N, D)).astype(np.float32)
    dr = lambda: (rng.standard_normal((N, D)).astype(np.float32), rng.standard_normal((N, D)).astype(np.float32),
                  rng.uniform(size=N).astype(np.float32), rng.uniform(size=N).astype(np.float32))
    draws_x, draws_z = dr(), dr()
    lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=N)
    dyn = pkg.GaugeDynamics(lat, hmc=False, network_arch="generic", num_steps=M, eps=eps, weights=weights)
    dyn.set_masks(mask)
    model = GaugeModel(dict(time_size=T, space_size=X, num_samples=N), lattice=lat, dynamics=dyn)
    model.native_train = False
    loss_a, _, _, px_a, _ = model._calc_loss_and_grads(x, beta, draws_x=draws_x, z=z, draws_z=draws_z)
    assert model.engine._tc_forward
    flat_a = model.engine.params.grad.clone()
    model.native_train = True
    loss_b, grads, _, px_b, _ = model._calc_loss_and_grads(x, beta, draws_x=draws_x, z=z, draws_z=draws_z)
    torch.cuda.synchronize()
    assert float(loss_a) == float(loss_b) and torch.equal(px_a, px_b)
    for name in model.engine.params.names:
        assert torch.equal(model.engine.params.view(flat_a, name), model.engine.params.view(model.engine.params.grad, name)), name
    od = orc.OracleDynamics(T, X, M, eps, hmc=False, network_arch="generic", mask=mask, xnet=weights["XNet"],
                            vnet=weights["VNet"])
    ref_loss, ref = tro.loss_and_grads(od, x, beta, draws_x, z, draws_z, aux_weight=1., std_weight=1., charge_weight=1.)
    assert abs(float(loss_b) - float(ref_loss)) <= 2e-3 * max(1., abs(float(ref_loss)))
    for name, gr in grads.items():
        close_scaled(gr, ref[name], 5e-4, f"tcgen05 step d loss / d {name}")


@pytest.mark.parametrize("rows,cols,ld", [(1, 5, 5), (31, 130, 130), (2560, 384, 1536), (40960, 1024, 1024), (1000, 77, 90)])
def test_col_sum(rows, cols, ld):
    """l2hmc_col_sum (bias / coeff gradients: batch sums of an activation gradient) against a float64 sum, on a column
    window of a wider matrix; two evaluations are bit-identical (fixed two-stage order)."""
    from l2hmc_b200.training import _Gemm, _ptr
    rng = np.random.default_rng(rows + cols)
    B = rng.standard_normal((rows, ld)).astype(np.float32)
    ref = B[:, :cols].astype(np.float64).sum(axis=0)
    tB = cu(B)
    g = _Gemm(torch.device("cuda"))
    out, out2 = torch.full((cols,), 3.0, device="cuda"), torch.empty(cols, device="cuda")
    g.col_sum(_ptr(tB), ld, rows, cols, _ptr(out))
    g.col_sum(_ptr(tB), ld, rows, cols, _ptr(out2))
    torch.cuda.synchronize()
    assert torch.equal(out, out2)
    err = np.abs(out.cpu().numpy().astype(np.float64) - ref).max()
    assert err <= 2e-6 * np.abs(B[:, :cols]).astype(np.float64).sum(axis=0).max(), err
