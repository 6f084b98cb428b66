"""CPU oracle for the MoG / generic-target L2HMC path (BASELINE config 5).

TEST INFRASTRUCTURE ONLY (see oracle/u1_oracle.py): imported by tests/, smoke() and the
cpu_baseline leg of bench.py, never by the product package.

Parity pin: tests/golden/mog_cases.npz, produced by the UNMODIFIED reference modules
utils/{dynamics,sampler,network,distributions}.py on the torch-backed TensorFlow stand-in
(tests/golden/make_golden_mog.py); tests/test_oracle_golden.py checks every function below
against it.  Citations are file:line under l2hmc/ of the reference.

One deliberate difference: quadratic_gaussian forms an N x N matmul and takes its diagonal
(utils/distributions.py:32-39), which is O(N^2) and impossible at 10^6 chains; the row-wise
form below computes the same diagonal.
"""
import math

import numpy as np
import torch


def _as(x, dtype):
    return torch.as_tensor(np.asarray(x) if not isinstance(x, torch.Tensor) else x).to(dtype)


class GMMTarget:
    """utils/distributions.py:124-158: E(x) = -logsumexp_i(-0.5 (x-mu_i)^T S_i (x-mu_i) + log c_i),
    S_i = inv(sigma_i) (fp32), c_i = pi_i / sqrt((2 pi)^k det sigma_i) (fp32)."""

    def __init__(self, mus, i_sigmas, constants, dtype=torch.float32):
        self.mus, self.i_sigmas = _as(mus, dtype), _as(i_sigmas, dtype)
        self.log_c = torch.log(_as(constants, torch.float32)).to(dtype)      # tf.log on the fp32 constant
        self.k = self.mus.shape[1]

    @classmethod
    def from_params(cls, mus, sigmas, pis, dtype=torch.float32):
        mus = np.asarray(mus, np.float32)
        pis = np.asarray(pis, np.float64)
        pis = pis / pis.sum()
        i_s, cs = [], []
        for i, s in enumerate(sigmas):
            s = np.asarray(s)
            i_s.append(np.linalg.inv(s).astype("float32"))
            det = np.sqrt((2 * np.pi) ** mus.shape[1] * np.linalg.det(s)).astype("float32")
            cs.append((pis[i] / det).astype("float32"))
        return cls(mus, np.stack(i_s), np.asarray(cs, np.float32), dtype)

    def energy(self, x):
        d = x[:, None, :] - self.mus[None, :, :]                             # (N, m, k)
        quad = 0.5 * torch.einsum("nmi,mij,nmj->nm", d, self.i_sigmas, d)
        return -torch.logsumexp(-quad + self.log_c[None, :], dim=1)

    def grad(self, x):
        xr = x.detach().clone().requires_grad_(True)
        (g,) = torch.autograd.grad(self.energy(xr).sum(), xr)
        return g


def mog_net(w, in0, in1, tvec):
    """utils/network.py:89-114: embed_1(in0) + embed_2(in1) + embed_3(t) -> relu -> linear_1 -> relu ->
    [scale_s * tanh(linear_s), linear_t, scale_f * tanh(linear_f)]  (tanh on BOTH S and F)."""
    def dense(name, x):
        return x @ w[name + ".kernel"] + w[name + ".bias"]
    h = torch.relu(dense("v_layer", in0) + dense("x_layer", in1) + dense("t_layer", tvec))
    h = torch.relu(dense("h_layer", h))
    S = torch.exp(w["coeff_scale"]) * torch.tanh(dense("scale_layer", h))
    T = dense("translation_layer", h)
    F = torch.exp(w["coeff_transformation"]) * torch.tanh(dense("transformation_layer", h))
    return S, T, F


def init_mog_net(x_dim, factor, rng, num_nodes=50):
    """Linear initialisers of utils/network.py:359-385 (variance scaling, FAN_IN, truncated normal)."""
    w = {}

    def dense(name, fan_in, units, f):
        std = math.sqrt(1.3 * 2. * f / fan_in)
        a = rng.standard_normal(size=(fan_in, units))
        bad = np.abs(a) > 2
        while bad.any():
            a[bad] = rng.standard_normal(size=int(bad.sum()))
            bad = np.abs(a) > 2
        w[name + ".kernel"] = (a * std).astype(np.float32)
        w[name + ".bias"] = np.zeros((units,), np.float32)
    dense("v_layer", x_dim, num_nodes, 1. / 3)
    dense("x_layer", x_dim, num_nodes, factor / 3.)
    dense("t_layer", 2, num_nodes, 1. / 3)
    dense("h_layer", num_nodes, num_nodes, 1.)
    for nm in ("scale_layer", "translation_layer", "transformation_layer"):
        dense(nm, num_nodes, x_dim, 0.001)
    w["coeff_scale"] = np.zeros((1, x_dim), np.float32)
    w["coeff_transformation"] = np.zeros((1, x_dim), np.float32)
    return w


def init_masks(x_dim, trajectory_length, seed=42):
    """utils/dynamics.py:85-100 (NumPy legacy global RNG)."""
    if seed is not None:
        np.random.seed(seed)
    rows = []
    for _ in range(trajectory_length):
        ind = np.random.permutation(np.arange(x_dim))[:int(x_dim / 2)]
        m = np.zeros((x_dim,))
        m[ind] = 1
        rows.append(m)
    return np.stack(rows).astype(np.float32)


class OracleMogDynamics:
    """utils/dynamics.py:34-319 with every random draw injected."""

    def __init__(self, target, x_dim, trajectory_length, eps, hmc=False, mask=None, xnet=None, vnet=None,
                 temperature=1.0, dtype=torch.float32):
        self.target, self.x_dim, self.L, self.hmc, self.dtype = target, x_dim, int(trajectory_length), hmc, dtype
        # eps = exp(alpha), alpha = log(eps) in fp32 (:51-60)
        self.eps = torch.exp(torch.log(torch.tensor(np.float32(eps)))).to(dtype)
        self.mask = _as(mask if mask is not None else init_masks(x_dim, self.L), dtype)
        self.xnet = {k: _as(a, dtype) for k, a in xnet.items()} if xnet is not None else None
        self.vnet = {k: _as(a, dtype) for k, a in vnet.items()} if vnet is not None else None
        self.temperature = torch.tensor(np.float32(temperature)).to(dtype)

    def _net(self, w, in0, in1, t):
        if self.hmc:
            z = torch.zeros_like(in0)
            return z, z, z
        return mog_net(w, in0, in1, t)

    def XNet(self, in0, in1, t):
        return self._net(self.xnet, in0, in1, t)

    def VNet(self, in0, in1, t):
        return self._net(self.vnet, in0, in1, t)

    def energy(self, x):
        """:227-236"""
        return self.target.energy(x) / self.temperature

    def grad_energy(self, x):
        """:241-242"""
        xr = x.detach().clone().requires_grad_(True)
        (g,) = torch.autograd.grad(self.energy(xr).sum(), xr)
        return g

    def hamiltonian(self, x, v):
        return self.energy(x) + 0.5 * (v ** 2).sum(dim=1)

    def _format_time(self, step, n):
        """:106-111"""
        t = torch.as_tensor(float(step), dtype=self.dtype)
        arg = 2 * math.pi * t / self.L
        return torch.stack([torch.cos(arg), torch.sin(arg)]).unsqueeze(0).repeat(n, 1)

    def forward_step(self, x, v, step):
        """:120-168"""
        eps, t = self.eps, self._format_time(step, x.shape[0])
        grad1 = self.grad_energy(x)
        S1 = self.VNet(x, grad1, t)
        sv1, tv1, fv1 = 0.5 * eps * S1[0], S1[1], eps * S1[2]
        v_h = v * torch.exp(sv1) + 0.5 * eps * (-(torch.exp(fv1) * grad1) + tv1)
        m = self.mask[int(step)]
        mb = 1. - m
        X1 = self.XNet(v_h, m * x, t)
        sx1, tx1, fx1 = eps * X1[0], X1[1], eps * X1[2]
        y = m * x + mb * (x * torch.exp(sx1) + eps * (torch.exp(fx1) * v_h + tx1))
        X2 = self.XNet(v_h, mb * y, t)
        sx2, tx2, fx2 = eps * X2[0], X2[1], eps * X2[2]
        x_o = mb * y + m * (y * torch.exp(sx2) + eps * (torch.exp(fx2) * v_h + tx2))
        grad2 = self.grad_energy(x_o)
        S2 = self.VNet(x_o, grad2, t)
        sv2, tv2, fv2 = 0.5 * eps * S2[0], S2[1], eps * S2[2]
        v_o = v_h * torch.exp(sv2) + 0.5 * eps * (-(torch.exp(fv2) * grad2) + tv2)
        return x_o, v_o, (sv1 + sv2 + mb * sx1 + m * sx2).sum(dim=1)

    def backward_step(self, x_o, v_o, step):
        """:170-225"""
        eps, t = self.eps, self._format_time(step, x_o.shape[0])
        grad1 = self.grad_energy(x_o)
        S1 = self.VNet(x_o, grad1, t)
        sv2, tv2, fv2 = -0.5 * eps * S1[0], S1[1], eps * S1[2]
        v_h = (v_o - 0.5 * eps * (-(torch.exp(fv2) * grad1) + tv2)) * torch.exp(sv2)
        m = self.mask[int(step)]
        mb = 1. - m
        X1 = self.XNet(v_h, mb * x_o, t)
        sx2, tx2, fx2 = -eps * X1[0], X1[1], eps * X1[2]
        y = mb * x_o + m * (torch.exp(sx2) * (x_o - eps * (torch.exp(fx2) * v_h + tx2)))
        X2 = self.XNet(v_h, m * y, t)
        sx1, tx1, fx1 = -eps * X2[0], X2[1], eps * X2[2]
        x = m * y + mb * (torch.exp(sx1) * (y - eps * (torch.exp(fx1) * v_h + tx1)))
        grad2 = self.grad_energy(x)
        S2 = self.VNet(x, grad2, t)
        sv1, tv1, fv1 = -0.5 * eps * S2[0], S2[1], eps * S2[2]
        v = torch.exp(sv1) * (v_h - 0.5 * eps * (-(torch.exp(fv1) * grad2) + tv1))
        return x, v, (sv1 + sv2 + mb * sx1 + m * sx2).sum(dim=1)

    def p_accept(self, x0, v0, x1, v1, log_jac):
        """:312-319"""
        d = self.hamiltonian(x0, v0) - self.hamiltonian(x1, v1) + log_jac
        p = torch.exp(torch.minimum(d, torch.zeros_like(d)))
        return torch.where(torch.isfinite(p), p, torch.zeros_like(p))

    def forward(self, x, init_v):
        """:255-283"""
        x0, v0 = _as(x, self.dtype), _as(init_v, self.dtype)
        X, V, j = x0, v0, torch.zeros(x0.shape[0], dtype=self.dtype)
        for t in range(self.L):
            X, V, lj = self.forward_step(X, V, t)
            j = j + lj
        return X, V, self.p_accept(x0, v0, X, V, j)

    def backward(self, x, init_v):
        """:285-310 (step index L - t - 1)"""
        x0, v0 = _as(x, self.dtype), _as(init_v, self.dtype)
        X, V, j = x0, v0, torch.zeros(x0.shape[0], dtype=self.dtype)
        for t in range(self.L):
            X, V, lj = self.backward_step(X, V, self.L - t - 1)
            j = j + lj
        return X, V, self.p_accept(x0, v0, X, V, j)

    def propose(self, x, v_f, v_b, dir_mask, u_acc):
        """utils/sampler.py:28-59 with do_mh_step=True.  hmc: forward only.  Otherwise both
        directions are integrated and blended with the {0,1} mask; accept is p - u >= 0."""
        x = _as(x, self.dtype)
        if self.hmc:
            Lx, Lv, px = self.forward(x, v_f)
        else:
            Lx1, Lv1, px1 = self.forward(x, v_f)
            Lx2, Lv2, px2 = self.backward(x, v_b)
            m = _as(dir_mask, self.dtype).reshape(-1, 1)
            Lx, Lv = m * Lx1 + (1 - m) * Lx2, m * Lv1 + (1 - m) * Lv2
            px = m[:, 0] * px1 + (1 - m[:, 0]) * px2
        acc = (px.to(torch.float32) - _as(u_acc, torch.float32)) >= 0.
        return Lx, Lv, px, torch.where(acc[:, None], Lx, x), acc


def calc_tunneling_rate(trajectory, means, num_distributions):
    """utils/trajectories.py:63-95: mode switches / (steps - 1) by nearest mean, per step along a
    trajectory (steps, dim) -- restated on arrays (steps, N, dim) -> (N,)."""
    tr = np.asarray(trajectory, np.float64)
    means = np.asarray(means, np.float64)[:num_distributions]
    d = np.linalg.norm(tr[:, :, None, :] - means[None, None, :, :], axis=-1)
    idx = d.argmin(axis=2)                               # (steps, N)
    return (idx[1:] != idx[:-1]).sum(axis=0) / (tr.shape[0] - 1)


def mog_loss(x, Lx, px, z, Lz, pz, scale):
    """mog_model.py:344-355: v = |jump|^2 * p + 1e-4 per chain; loss = scale (mean 1/v1 + mean 1/v2) - (mean v1 +
    mean v2) / scale.  float64 numpy; returns (loss, v1, v2)."""
    f = lambda a: np.asarray(a, np.float64)
    v1 = ((f(x) - f(Lx)) ** 2).sum(axis=1) * f(px) + 1e-4
    v2 = ((f(z) - f(Lz)) ** 2).sum(axis=1) * f(pz) + 1e-4
    loss = scale * ((1.0 / v1).mean() + (1.0 / v2).mean()) + (-v1.mean() - v2.mean()) / scale
    return loss, v1, v2
