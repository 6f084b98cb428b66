"""CPU oracle for the 'next' rows of SURVEY section 8(f): the training loss VALUE and the run-loop statistics.

TEST INFRASTRUCTURE ONLY: imported by tests/ (and nothing in the l2hmc_b200 package).  Restates, op for op in torch-CPU
fp32 / NumPy, the reference's

  gauge_model.py:78-80      project_angle
  gauge_model.py:94-108     project_angle_approx (Fourier series, N = 5 -> harmonics 1..4)
  gauge_model.py:631-657    metric functions l1 / l2 / cos / cos2 / cos_diff
  gauge_model.py:659-725    _calc_plaq_sums, _calc_top_charges(fft), _calc_top_charges_diff
  gauge_model.py:728-803    _calc_loss
  gauge_model.py:1473-1533  calc_observables_stats

Pinned by tests/test_oracle_golden.py against tests/golden/train_cases.npz, which the unmodified reference produced
(tests/golden/make_golden_train.py).
"""
import collections

import numpy as np
import torch

from . import u1_oracle as orc

TWO_PI = 2 * np.pi


def _as(x, dtype=torch.float32):
    return x.to(dtype) if isinstance(x, torch.Tensor) else torch.as_tensor(np.asarray(x), dtype=dtype)


def project_angle_approx(p, n_terms=5):
    """gauge_model.py:94-108: y = sum_{n=1}^{N-1} (-2/n) (-1)^n sin(n p), accumulated in that order."""
    y = torch.zeros_like(p)
    for n in range(1, n_terms):
        y = y + (-2 / n) * ((-1) ** n) * torch.sin(n * p)
    return y


def top_charges(x, T, X, fft=False):
    """gauge_model.py:699-713: sum of projected plaquette angles / 2 pi (float)."""
    p = orc.plaq_sums(_as(x), T, X)
    proj = project_angle_approx(p) if fft else orc.project_angle(p)
    return proj.sum(dim=(1, 2)) / TWO_PI


def top_charges_diff(x1, x2, T, X, fft=False):
    """gauge_model.py:717-725."""
    return (top_charges(x1, T, X, fft) - top_charges(x2, T, X, fft)).abs()


def metric_fn(name):
    """gauge_model.py:631-657."""
    fns = {"l1": lambda a, b: (a - b).abs(), "l2": lambda a, b: (a - b) ** 2,
           "cos": lambda a, b: (torch.cos(a) - torch.cos(b)).abs(),
           "cos2": lambda a, b: (torch.cos(a) - torch.cos(b)) ** 2,
           "cos_diff": lambda a, b: 1. - torch.cos(a - b)}
    if name not in fns:
        raise AttributeError(f"metric={name}. Expected one of: 'l1', 'l2', 'cos', 'cos2', or 'cos_diff'.")
    return fns[name]


def loss_terms(x, x_prop, z, px, pz, x_out, T, X, metric="cos_diff", loss_scale=1., aux_weight=1., std_weight=1.,
               charge_weight=1.):
    """gauge_model.py:762-801 given the two chains' proposals: returns (loss, per-chain loss, x_dq int32).
    NB the z-terms compare z with x_prop (the x-chain's proposal), as the reference does (:774, :789)."""
    x, x_prop, z, x_out = (_as(a) for a in (x, x_prop, z, x_out))
    px, pz = _as(px), _as(pz)
    eps = 1e-3
    m = metric_fn(metric)
    x_dq = top_charges_diff(x, x_out, T, X, fft=False).to(torch.int32)
    x_std = m(x, x_prop).sum(dim=1) * px + eps
    z_std = aux_weight * (m(z, x_prop).sum(dim=1) * pz + eps)
    std_loss = (loss_scale * (1. / x_std + 1. / z_std) - (x_std + z_std) / loss_scale) * std_weight
    xq_loss = px * top_charges_diff(x, x_prop, T, X, fft=True) + eps
    zq_loss = aux_weight * (pz * top_charges_diff(z, x_prop, T, X, fft=True) + eps)
    per_chain = std_loss + charge_weight * (xq_loss + zq_loss)
    return per_chain.mean(dim=0), per_chain, x_dq


def calc_loss(dyn, x, beta, draws_x, z, draws_z, metric="cos_diff", loss_scale=1., **weights):
    """gauge_model.py:728-803 on an oracle dynamics object (oracle.u1_oracle.OracleDynamics);
    draws_* = (v_f, v_b, u_dir, u_acc) of each chain's apply_transition.  Returns (loss, x_out, px, x_dq)."""
    x_prop, _, px, x_out = dyn.apply_transition(x, beta, *draws_x)
    _, _, pz, _ = dyn.apply_transition(z, beta, *draws_z)
    loss, _, x_dq = loss_terms(x, x_prop, z, px, pz, x_out, dyn.T, dyn.X, metric, loss_scale, **weights)
    return loss, x_out, px, x_dq


def _sem(a):
    """scipy.stats.sem (axis 0, ddof 1) in the array's own dtype, like the call at gauge_model.py:1512-1515."""
    a = np.asarray(a)
    n = a.shape[0]
    return np.std(a, axis=0, ddof=1) / np.sqrt(n)


def calc_observables_stats(actions, plaqs, charges, therm_frac=10):
    """gauge_model.py:1473-1533 on (steps, N) arrays.  Quirk kept: the susceptibility array is built BEFORE the
    thermalisation cut (:1498 vs :1503-1505), so its mean / error run over all steps."""
    actions_arr, plaqs_arr = np.asarray(actions), np.asarray(plaqs)
    charges_arr = np.array(np.asarray(charges), dtype=int)
    suscept_arr = np.array(charges_arr ** 2)
    therm_steps = actions_arr.shape[0] // therm_frac
    actions_arr, plaqs_arr, charges_arr = actions_arr[therm_steps:], plaqs_arr[therm_steps:], charges_arr[therm_steps:]
    stats = {}
    for name, arr in (("actions", actions_arr), ("plaqs", plaqs_arr), ("charges", charges_arr),
                      ("suscept", suscept_arr)):
        stats[name] = (np.mean(arr, axis=0), _sem(arr))
    counts = collections.Counter(list(charges_arr.flatten()))
    total = np.sum(list(counts.values()))
    stats["charge_probabilities"] = collections.OrderedDict(sorted(((k, v / total) for k, v in counts.items()),
                                                                key=lambda kv: kv[0]))
    return stats


def loss_and_grads(dyn, x, beta, draws_x, z, draws_z, metric="cos_diff", loss_scale=1., **weights):
    """gauge_model.py:805-832 (tf.gradients(loss, dynamics.variables)) by reverse-mode autodiff THROUGH the oracle's
    own transition (both chains, both directions, the analytic force included).  Returns (loss, {name: grad}) with
    names 'eps', 'XNet.<layer>.kernel', ... as in the golden fixtures.  This is the reference the backward kernels
    of the next round are checked against."""
    named = {"eps": dyn.eps}
    for net_name, w in (("XNet", dyn.xnet), ("VNet", dyn.vnet)):
        if w is not None:
            for k, v in w.items():
                named[f"{net_name}.{k}"] = v
    for v in named.values():
        v.requires_grad_(True)
    try:
        loss, _, _, _ = calc_loss(dyn, x, beta, draws_x, z, draws_z, metric=metric, loss_scale=loss_scale, **weights)
        grads = torch.autograd.grad(loss, list(named.values()), allow_unused=True)
    finally:
        for v in named.values():
            v.requires_grad_(False)
    return loss.detach(), {k: (torch.zeros_like(v) if g is None else g.detach()) for (k, v), g in zip(named.items(), grads)}


def exponential_decay(lr_init, global_step, decay_steps, decay_rate, staircase=True):
    """tf.train.exponential_decay as called at gauge_model.py:934-939 (TensorFlow 1.x, third party, not vendored:
    decayed = lr * rate ** (step / decay_steps), the exponent floored when staircase)."""
    e = global_step / float(decay_steps)
    return lr_init * decay_rate ** (np.floor(e) if staircase else e)


def adam_update(param, grad, m, v, t, lr, beta1=0.9, beta2=0.999, epsilon=1e-8):
    """tf.train.AdamOptimizer.apply_gradients (gauge_model.py:940-945).  TensorFlow is a third-party dependency that is
    absent from /root/reference (unpinned: no requirements file; API use implies >= 1.12, < 2.0), so this restates the
    update published in the class docstring of tensorflow/python/training/adam.py (TF 1.x), which nothing here can pin:
        t <- t + 1
        lr_t <- learning_rate * sqrt(1 - beta2^t) / (1 - beta1^t)
        m_t <- beta1 * m_{t-1} + (1 - beta1) * g
        v_t <- beta2 * v_{t-1} + (1 - beta2) * g * g
        variable <- variable - lr_t * m_t / (sqrt(v_t) + epsilon)
    i.e. the bias corrections m_hat, v_hat of Kingma & Ba's Algorithm 1 are FOLDED into lr_t and epsilon sits OUTSIDE
    the square root, un-corrected ("epsilon hat" of the paper's section 2) -- not Algorithm 1's epsilon.  Defaults
    beta1 = 0.9, beta2 = 0.999, epsilon = 1e-8 (the reference passes only the learning rate).  float64 NumPy.
    Returns (m, v, param)."""
    lr_t = lr * np.sqrt(1. - beta2 ** t) / (1. - beta1 ** t)
    m = beta1 * m + (1. - beta1) * grad
    v = beta2 * v + (1. - beta2) * grad * grad
    return m, v, param - lr_t * m / (np.sqrt(v) + epsilon)


def clip_by_global_norm(grads, clip_norm):
    """tf.clip_by_global_norm (gauge_model.py:826-827): every tensor scaled by clip_norm / max(global_norm, clip_norm)."""
    norm = np.sqrt(sum(float((np.asarray(g, np.float64) ** 2).sum()) for g in grads))
    scale = clip_norm / max(norm, clip_norm)
    return [np.asarray(g) * scale for g in grads], norm
