"""Shared pieces of the parity tests (test infrastructure; imports the oracle).

* trained_like_weights: the reference's initialisers (oracle.init_net_weights) with O(0.1) heads and non-zero
  biases / coeffs, a pure function of a NumPy Generator -- so a golden file can hold the OUTPUTS the unmodified
  reference produced with these weights without holding the (tens of MB of) weights themselves.
* error measures of the north star: "links, actions, log-dets and plaquettes must agree within 1e-5 relative in
  fp32".  rel_err(a, ref) = max|a - ref| / max|ref| (relative to the array's scale: links are angles in [0, 2 pi),
  an element-wise relative error is meaningless at a zero crossing).
* the tolerance rule of the whole-transition tests: an fp32 result is accepted when its distance from the float64
  oracle is at most max(1e-5 relative, K_NOISE x the fp32 ORACLE's own distance from float64 on the same inputs):
  M leapfrog steps through exp() and the networks amplify fp32 rounding, and no fp32 implementation -- the
  reference's included -- is closer to the exact trajectory than that.
"""
import hashlib

import numpy as np
import torch

from oracle import u1_oracle as orc

K_NOISE = 2.0          # CUDA-vs-fp64 may be at most this many times oracle-fp32-vs-fp64
REL_TOL = 1e-5         # north star


def trained_like_weights(arch, T, X, rng, gain=None, num_hidden=None):
    gain = gain if gain is not None else (1.0 if arch == "generic" else 5.0)   # S,T,Q = O(0.3): p_acc stays > 0
    ws = {}
    for net, factor in (("XNet", 2.), ("VNet", 1.)):
        w = orc.init_net_weights(arch, T, X, factor, rng, num_hidden=num_hidden)
        for k in ("scale_layer", "translation_layer", "transformation_layer"):
            w[k + ".kernel"] *= gain                                       # trained-like heads: S,T,Q = O(0.1)
            w[k + ".bias"] = rng.normal(0, 0.02, w[k + ".bias"].shape).astype(np.float32)
        for k in ("x_layer", "v_layer", "t_layer", "h_layer"):
            w[k + ".bias"] = rng.normal(0, 0.05, w[k + ".bias"].shape).astype(np.float32)
        w["coeff_scale"] = rng.normal(0, 0.1, w["coeff_scale"].shape).astype(np.float32)
        w["coeff_transformation"] = rng.normal(0, 0.1, w["coeff_transformation"].shape).astype(np.float32)
        ws[net] = w
    return ws


def weights_digest(ws):
    """sha256 over every tensor's bytes in name order: a golden file records it, the test re-derives it."""
    h = hashlib.sha256()
    for net in sorted(ws):
        for k in sorted(ws[net]):
            h.update(k.encode())
            h.update(np.ascontiguousarray(ws[net][k], dtype=np.float32).tobytes())
    return h.hexdigest()


def _np64(a):
    if isinstance(a, torch.Tensor):
        a = a.detach().cpu().numpy()
    return np.asarray(a, dtype=np.float64)


def abs_err(a, ref):
    d = np.abs(_np64(a) - _np64(ref))
    return float(d.max()) if d.size else 0.0


def rel_err(a, ref):
    ref = _np64(ref)
    return abs_err(a, ref) / max(float(np.abs(ref).max()) if ref.size else 0.0, 1e-30)


def assert_within_noise(name, got, ref64, oracle32, k=K_NOISE, rel=REL_TOL):
    """got (fp32 result) vs ref64 (float64 oracle), with oracle32 = the fp32 oracle on the same inputs."""
    e, noise = rel_err(got, ref64), rel_err(oracle32, ref64)
    assert e <= max(rel, k * noise), (
        f"{name}: relative error {e:.3e} vs float64 exceeds max({rel:g}, {k:g} x fp32-oracle error {noise:.3e})")
    return e, noise


def oracle_pair(arch, T, X, M, eps, mask, ws):
    mk = lambda dt: orc.OracleDynamics(T, X, M, eps, hmc=(ws is None), network_arch=arch, mask=mask,
                                       xnet=None if ws is None else ws["XNet"], vnet=None if ws is None else ws["VNet"],
                                       dtype=dt)
    return mk(torch.float32), mk(torch.float64)


def _oracle_trajectory(od, x, v, beta, forward):
    """M leapfrog steps one way with the summed log-det (gauge_dynamics.py:284-301), oracle side."""
    x, v = orc._as(x, od.dtype), orc._as(v, od.dtype)
    b = torch.as_tensor(np.float32(beta)).to(od.dtype)
    ld = torch.zeros(x.shape[0], dtype=od.dtype)
    lf = od.forward_lf if forward else od.backward_lf
    for t in range(od.num_steps):
        x, v, j = lf(x, v, b, t)
        ld = ld + j
    return x, v, ld


def _cuda_trajectory(dyn, x, v, beta, forward):
    xs, vs, ld = dyn._x(x), dyn._x(v), None
    for t in range(dyn.num_steps):
        xs, vs, ld = dyn._lf(xs, vs, beta, t, +1 if forward else -1, ld)
    return xs, vs, ld


def measure_case(dyn, od32, od64, x, beta, v_f, v_b, u_dir, u_acc, T, X):
    """Every quantity the north star names, CUDA vs the float64 oracle next to fp32-oracle vs float64.

    Returns {name: (cuda_err, oracle32_err)} with relative errors (rel_err above), plus integer entries
    'accept_mismatch' / 'accept_unsafe' / 'charge_mismatch' / 'charge_same_input_mismatch'."""
    out = {}
    lat = dyn.lattice
    # --- trajectories each way: links, momenta, summed log-det (dyn._lf accumulates in place)
    for name, fwd, v0 in (("fwd", True, v_f), ("bwd", False, v_b)):
        cx, cv, cld = _cuda_trajectory(dyn, x, v0, beta, fwd)
        ax, av, ald = _oracle_trajectory(od32, x, v0, beta, fwd)
        bx, bv, bld = _oracle_trajectory(od64, x, v0, beta, fwd)
        out[f"links[{name}]"] = (rel_err(cx, bx), rel_err(ax, bx))
        out[f"momenta[{name}]"] = (rel_err(cv, bv), rel_err(av, bv))
        if not dyn.hmc:
            out[f"logdet[{name}]"] = (rel_err(cld, bld), rel_err(ald, bld))
    # --- the MC step through the public call (one-launch / one-C-call path)
    cxp, cvp, cp, cxo = dyn.apply_transition(x, beta, momentum_f=v_f, momentum_b=v_b, direction_uniform=u_dir,
                                             accept_uniform=u_acc)
    axp, avp, ap, axo = od32.apply_transition(x, beta, v_f, v_b, u_dir, u_acc)
    bxp, bvp, bp, bxo = od64.apply_transition(x, beta, v_f, v_b, u_dir, u_acc)
    out["links[x_post]"] = (rel_err(cxp, bxp), rel_err(axp, bxp))
    out["accept_prob"] = (abs_err(cp, bp), abs_err(ap, bp))
    ua = np.asarray(u_acc, dtype=np.float32)
    safe = np.abs(bp.numpy() - ua) > 1e-3
    acc = dyn.last_accepted.cpu().numpy().astype(bool)
    out["accept_unsafe"] = int((~safe).sum())
    out["accept_mismatch"] = int((acc[safe] != (bp.numpy() > ua)[safe]).sum())
    # --- observables of the state the step ends in: on the trajectory's own x_out ...
    c_act, c_plaq, c_q = lat.calc_plaq_observables(cxo)
    a_act, a_plaq, a_q = orc.plaq_observables(axo, T, X)
    b_act, b_plaq, b_q = orc.plaq_observables(bxo, T, X)
    ok = acc == (bp.numpy() > ua)                      # chains where both took the same branch
    out["action"] = (rel_err(c_act.cpu().numpy()[ok], b_act.numpy()[ok]), rel_err(a_act.numpy()[ok], b_act.numpy()[ok]))
    out["plaquette"] = (rel_err(c_plaq.cpu().numpy()[ok], b_plaq.numpy()[ok]),
                        rel_err(a_plaq.numpy()[ok], b_plaq.numpy()[ok]))
    out["charge_mismatch"] = int((c_q.cpu().numpy()[ok] != b_q.numpy()[ok]).sum())
    # ... and on identical inputs (the CUDA x_out fed to the fp32 oracle): integer charge must be bit-exact
    s_act, s_plaq, s_q = orc.plaq_observables(cxo.cpu(), T, X)
    out["charge_same_input_mismatch"] = int((c_q.cpu().numpy() != s_q.numpy()).sum())
    out["action[same input]"] = (rel_err(c_act, s_act), 0.0)
    out["plaquette[same input]"] = (rel_err(c_plaq, s_plaq), 0.0)
    return out


def check_case(tag, errs, k=K_NOISE, rel=REL_TOL):
    """The acceptance rule over a measure_case() result."""
    for name, val in errs.items():
        if isinstance(val, tuple):
            e, noise = val
            tol = max(rel, k * noise)
            if name == "accept_prob":            # absolute error of a probability; fp32 H sums of O(1e3) carry ~1e-4
                tol = max(2e-4, k * noise)
            if name.endswith("[same input]"):
                tol = rel
            assert e <= tol, f"{tag}: {name}: error {e:.3e} > {tol:.3e} (fp32 oracle vs float64: {noise:.3e})"
    assert errs["accept_mismatch"] == 0, f"{tag}: accept decisions differ on chains with |p - u| > 1e-3"
    assert errs["charge_same_input_mismatch"] == 0, f"{tag}: integer charge differs from the oracle on the same links"


def format_table(rows):
    """rows: [(config label, measure_case dict)] -> text table (profiles/r2_parity_errors.txt)."""
    lines = [f"{'config':<44} {'quantity':<22} {'CUDA vs fp64':>13} {'oracle32 vs fp64':>17} {'ratio':>7}"]
    for label, errs in rows:
        for name, val in errs.items():
            if isinstance(val, tuple):
                e, n = val
                ratio = f"{e / n:7.2f}" if n > 0 else "      -"
                lines.append(f"{label:<44} {name:<22} {e:13.3e} {n:17.3e} {ratio}")
            else:
                lines.append(f"{label:<44} {name:<22} {val:13d}")
        lines.append("")
    return "\n".join(lines)
