"""CPU oracle for the L2HMC / HMC leapfrog path of the 2D U(1) gauge model.

TEST INFRASTRUCTURE ONLY.  This is a torch-CPU restatement of the reference's
algorithm (saforem2/l2hmc, TensorFlow-1 graph) for the hot path named in
SURVEY.md section 8.  Only tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py may import it, and only as
the checker or the timed CPU baseline -- never from the product package.

Parity pin: the reference has no tests and no golden vectors of its own
(SURVEY.md 8c).  The pin is tests/golden/*.npz, produced by running the
UNMODIFIED reference modules on a torch-backed TensorFlow stand-in
(tests/golden/make_golden.py); tests/test_oracle_golden.py checks every
function below against them.

Each function cites the reference file:line (under l2hmc/) it restates, and
keeps the reference's fp32 association order: python scalars fold into fp32
exactly where TF would fold them, sums are per-op, nothing is fused.
`dtype=torch.float64` runs the same sequence in double to flag near-ties.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

PI = math.pi
TWO_PI = 2 * math.pi


def _as(x, dtype):
    return torch.as_tensor(np.asarray(x) if not isinstance(x, torch.Tensor) else x).to(dtype)


# --------------------------------------------------------------------------- lattice
def plaq_sums(x, T, X):
    """P(t,x) = U0(t,x) - U1(t,x) - U0(t,x+1) + U1(t+1,x), left-associated.
    lattice/lattice.py:300-303, 354-357; gauge_model.py:675-681."""
    s = x.reshape(-1, T, X, 2)
    return (s[:, :, :, 0] - s[:, :, :, 1]
            - torch.roll(s[:, :, :, 0], shifts=-1, dims=2)
            + torch.roll(s[:, :, :, 1], shifts=-1, dims=1))


def total_action(x, T, X):
    """S = sum_{t,x} (1 - cos P).  lattice/lattice.py:337-362."""
    return (1. - torch.cos(plaq_sums(x, T, X))).sum(dim=(1, 2))


def project_angle(p):
    """p - 2pi*floor((p + pi)/2pi).  lattice/lattice.py:47-49, gauge_model.py:78-80."""
    return p - TWO_PI * torch.floor((p + PI) / TWO_PI)


def plaq_observables(x, T, X):
    """(total_action, avg_plaq, integer topological charge as float).
    lattice/lattice.py:285-313."""
    ps = plaq_sums(x, T, X)
    local = torch.cos(ps)
    action = (1. - local).sum(dim=(1, 2))
    avg_plaq = local.sum(dim=(1, 2)) / (T * X)
    charge = torch.floor(0.1 + project_angle(ps).sum(dim=(1, 2)) / TWO_PI)
    return action, avg_plaq, charge


def avg_plaqs(x, T, X):
    """gauge_model.py:692-699."""
    return torch.cos(plaq_sums(x, T, X)).sum(dim=(1, 2)) / (T * X)


def top_charges(x, T, X):
    """Real-valued charge sum(project_angle(P)) / 2pi.  gauge_model.py:701-716 (fft=False)."""
    return project_angle(plaq_sums(x, T, X)).sum(dim=(1, 2)) / TWO_PI


def top_charges_diff(x1, x2, T, X):
    """|Q(x1) - Q(x2)|.  gauge_model.py:718-725."""
    return torch.abs(top_charges(x1, T, X) - top_charges(x2, T, X))


def potential_energy(x, beta, T, X):
    """beta * S(x).  dynamics/gauge_dynamics.py:675-681."""
    return beta * total_action(x, T, X)


def kinetic_energy(v):
    """0.5 * sum_d v^2.  dynamics/gauge_dynamics.py:683-689."""
    return 0.5 * (v ** 2).sum(dim=1)


def hamiltonian(x, v, beta, T, X):
    """dynamics/gauge_dynamics.py:691-696."""
    return potential_energy(x, beta, T, X) + kinetic_energy(v)


def grad_potential(x, beta, T, X):
    """d(beta*S)/dx.  The reference differentiates the graph (dynamics/gauge_dynamics.py:698-709);
    the adjoint of lattice.py:354-360 is, per plaquette term b = beta*sin P:
        g0(t,x) = b(t,x) - b(t,x-1),   g1(t,x) = -b(t,x) + b(t-1,x)
    (beta multiplies before the difference, as in the back-propagated graph; the same
    staple difference as lattice/gauge_lattice.py:444-454)."""
    b = beta * torch.sin(plaq_sums(x, T, X))
    g0 = b - torch.roll(b, shifts=1, dims=2)
    g1 = -b + torch.roll(b, shifts=1, dims=1)
    return torch.stack([g0, g1], dim=-1).reshape(x.shape[0], -1)


def grad_potential_autograd(x, beta, T, X):
    """Independent cross-check of grad_potential by automatic differentiation."""
    xr = x.detach().clone().requires_grad_(True)
    (g,) = torch.autograd.grad(potential_energy(xr, beta, T, X).sum(), xr)
    return g


def grad_potential_loops(x, beta, T, X):
    """Second independent cross-check: per-site staple loop of the older lattice class
    (lattice/gauge_lattice.py:427-459), pure numpy float64."""
    s = np.asarray(x, dtype=np.float64).reshape(-1, T, X, 2)
    g = np.zeros_like(s)
    for t in range(T):
        for xx in range(X):
            def P(tt, xs):
                tt, xs = tt % T, xs % X
                return (s[:, tt, xs, 0] - s[:, tt, xs, 1] - s[:, tt, (xs + 1) % X, 0]
                        + s[:, (tt + 1) % T, xs, 1])
            g[:, t, xx, 0] = beta * (np.sin(P(t, xx)) - np.sin(P(t, xx - 1)))
            g[:, t, xx, 1] = beta * (-np.sin(P(t, xx)) + np.sin(P(t - 1, xx)))
    return g.reshape(s.shape[0], -1)


# --------------------------------------------------------------------------- masks / time
def construct_masks(x_dim, num_steps, seed=42):
    """Random half masks from the NumPy legacy RNG.  dynamics/gauge_dynamics.py:651-661
    (seeded by gauge_model.py:62 / globals.py:12).  Pass seed=None to continue the stream."""
    if seed is not None:
        np.random.seed(seed)
    rows = []
    for _ in range(num_steps):
        idx = np.random.permutation(np.arange(x_dim))[:x_dim // 2]
        m = np.zeros((x_dim,))
        m[idx] = 1
        rows.append(m)
    return np.stack(rows).astype(np.float32)


def format_time(step, num_steps, dtype=torch.float32):
    """[cos(2 pi t / M), sin(2 pi t / M)] with t carried in the working precision.
    dynamics/gauge_dynamics.py:625-633 (step is an fp32 loop counter there)."""
    t = torch.as_tensor(float(step), dtype=dtype)
    arg = TWO_PI * t / num_steps
    return torch.stack([torch.cos(arg), torch.sin(arg)])


# --------------------------------------------------------------------------- networks
GENERIC_LAYERS = ("x_layer", "v_layer", "t_layer", "h_layer", "scale_layer", "translation_layer",
                  "transformation_layer")
CONV_LAYERS = ("conv_x1", "conv_v1", "conv_x2", "conv_v2")


def _dense(w, name, x):
    return x @ w[name + ".kernel"] + w[name + ".bias"]


def _heads(w, h):
    """generic_net.py:139-144 / conv_net.py:270-278: tanh only on the scale head."""
    scale = torch.tanh(_dense(w, "scale_layer", h)) * torch.exp(w["coeff_scale"])
    translation = _dense(w, "translation_layer", h)
    transformation = _dense(w, "transformation_layer", h) * torch.exp(w["coeff_transformation"])
    return scale, translation, transformation


def generic_net(w, in0, in1, tvec):
    """GenericNet.call: the FIRST input goes through v_layer, the second through x_layer.
    network/generic_net.py:129-146."""
    h = _dense(w, "v_layer", in0) + _dense(w, "x_layer", in1) + _dense(w, "t_layer", tvec)
    h = torch.relu(h)
    h = torch.relu(_dense(w, "h_layer", h))
    return _heads(w, h)


def _same_pad(size, k, stride):
    out = -(-size // stride)
    total = max((out - 1) * stride + k - size, 0)
    return total // 2, total - total // 2          # TF puts the odd cell at the end


def _conv3d_same_relu(x, kernel, bias):
    """Keras Conv3D(padding='same', activation=relu), channels_last.  x: (N,d0,d1,d2,C);
    kernel (k0,k1,k2,Cin,Cout).  network/conv_net.py:86-96, 112-121."""
    xc = x.permute(0, 4, 1, 2, 3)
    pads = []
    for dim in (3, 2, 1):
        pads += list(_same_pad(int(x.shape[dim]), int(kernel.shape[dim - 1]), 1))
    y = F.conv3d(F.pad(xc, pads), kernel.permute(4, 3, 0, 1, 2), bias)
    return torch.relu(y.permute(0, 2, 3, 4, 1))


def _maxpool3d_same(x):
    """MaxPooling3D((2,2,2), strides=2, 'same').  network/conv_net.py:98-104."""
    xc = x.permute(0, 4, 1, 2, 3)
    pads = []
    for dim in (3, 2, 1):
        pads += list(_same_pad(int(x.shape[dim]), 2, 2))
    y = F.max_pool3d(F.pad(xc, pads, value=float("-inf")), 2, 2)
    return y.permute(0, 2, 3, 4, 1)


def conv3d_net(w, in0, in1, tvec, T, X):
    """ConvNet3D.call, channels_last.  network/conv_net.py:247-280, reshape_5D :300-306."""
    N = in0.shape[0]
    v = in0.reshape(N, T, X, 2, 1)
    x = in1.reshape(N, T, X, 2, 1)
    x = _maxpool3d_same(_conv3d_same_relu(x, w["conv_x1.kernel"], w["conv_x1.bias"]))
    x = _maxpool3d_same(_conv3d_same_relu(x, w["conv_x2.kernel"], w["conv_x2.bias"]))
    x = x.reshape(N, -1)
    v = _maxpool3d_same(_conv3d_same_relu(v, w["conv_v1.kernel"], w["conv_v1.bias"]))
    v = _maxpool3d_same(_conv3d_same_relu(v, w["conv_v2.kernel"], w["conv_v2.bias"]))
    v = v.reshape(N, -1)
    h = torch.relu(_dense(w, "v_layer", v) + _dense(w, "x_layer", x) + _dense(w, "t_layer", tvec))
    h = torch.relu(_dense(w, "h_layer", h))
    return _heads(w, h)


def _truncated_normal(rng, shape, std):
    out = rng.standard_normal(size=shape)
    bad = np.abs(out) > 2.0
    while bad.any():
        out[bad] = rng.standard_normal(size=int(bad.sum()))
        bad = np.abs(out) > 2.0
    return (out * std).astype(np.float32)


def init_net_weights(arch, T, X, factor, rng, num_hidden=None):
    """Weights with the reference's initialisers (Keras layouts).
    Dense: truncated normal, std sqrt(1.3 * 2*f / fan_in) with f = factor/3, 1/3, 1/3, 1, 0.001 x3
    (network/generic_net.py:39-83,149-161; tf.contrib variance_scaling_initializer).
    Conv3D: Keras default glorot-uniform (network/conv_net.py:86-141).
    Sizes: generic H = 4D (dynamics/gauge_dynamics.py:169-187); conv3D F = X, H = 2D (:121-143)."""
    D = 2 * T * X
    w = {}

    def dense(name, fan_in, units, f):
        w[name + ".kernel"] = _truncated_normal(rng, (fan_in, units), math.sqrt(1.3 * 2. * f / fan_in))
        w[name + ".bias"] = np.zeros((units,), np.float32)

    if arch == "generic":
        H = num_hidden or 4 * D
        feat = D
    elif arch == "conv3D":
        H = num_hidden or 2 * D
        nf = X
        for nm, shape in (("conv_x1", (3, 3, 2, 1, nf)), ("conv_v1", (3, 3, 2, 1, nf)),
                          ("conv_x2", (2, 2, 2, nf, 2 * nf)), ("conv_v2", (2, 2, 2, nf, 2 * nf))):
            rec = int(np.prod(shape[:3]))
            lim = math.sqrt(6.0 / (shape[3] * rec + shape[4] * rec))
            w[nm + ".kernel"] = rng.uniform(-lim, lim, size=shape).astype(np.float32)
            w[nm + ".bias"] = np.zeros((shape[4],), np.float32)
        half = lambda n: -(-n // 2)                       # 'same' pooling halves, rounding up
        feat = half(half(T)) * half(half(X)) * 1 * 2 * nf
    else:
        raise AttributeError("network_arch must be one of 'conv3D', 'generic'")
    dense("x_layer", feat, H, factor / 3.)
    dense("v_layer", feat, H, 1. / 3.)
    dense("t_layer", 2, H, 1. / 3.)
    dense("h_layer", H, H, 1.)
    dense("scale_layer", H, D, 0.001)
    dense("translation_layer", H, D, 0.001)
    dense("transformation_layer", H, D, 0.001)
    w["coeff_scale"] = np.zeros((1, D), np.float32)
    w["coeff_transformation"] = np.zeros((1, D), np.float32)
    return w


# --------------------------------------------------------------------------- dynamics
class OracleDynamics:
    """GaugeDynamics restated (dynamics/gauge_dynamics.py:42-709) with all random draws injected."""

    def __init__(self, T, X, num_steps, eps, hmc=True, network_arch="generic", mask=None,
                 xnet=None, vnet=None, dtype=torch.float32):
        self.T, self.X, self.x_dim = T, X, 2 * T * X
        self.num_steps, self.hmc, self.arch, self.dtype = num_steps, hmc, network_arch, dtype
        self.eps = torch.as_tensor(np.float32(eps)).to(dtype)      # fp32 variable, :89-95
        if mask is None:
            mask = construct_masks(self.x_dim, num_steps)
        self.mask = _as(mask, dtype)
        self.xnet = {k: _as(a, dtype) for k, a in xnet.items()} if xnet is not None else None
        self.vnet = {k: _as(a, dtype) for k, a in vnet.items()} if vnet is not None else None
        if not hmc and network_arch not in ("generic", "conv3D"):
            raise AttributeError("network_arch must be one of 'conv3D', 'generic'")

    # -- networks (hmc stubs: gauge_dynamics.py:102-108)
    def _net(self, w, in0, in1, tvec):
        if self.hmc:
            z = torch.zeros_like(in0)
            return z, z, z
        if self.arch == "generic":
            return generic_net(w, in0, in1, tvec)
        return conv3d_net(w, in0, in1, tvec, self.T, self.X)

    def position_fn(self, in0, in1, tvec):
        return self._net(self.xnet, in0, in1, tvec)

    def momentum_fn(self, in0, in1, tvec):
        return self._net(self.vnet, in0, in1, tvec)

    def grad_potential(self, x, beta):
        return grad_potential(x, beta, self.T, self.X)

    def _tvec(self, step, n):
        return format_time(step, self.num_steps, self.dtype).unsqueeze(0).repeat(n, 1)

    def _masks(self, step):
        m = self.mask[int(step)]
        return m, 1. - m

    # -- sub-updates
    def update_momentum_forward(self, x, v, beta, tvec):
        """:486-508"""
        grad = self.grad_potential(x, beta)
        scale, translation, transformed = self.momentum_fn(x, grad, tvec)
        scale = scale * (0.5 * self.eps)
        transformed = transformed * self.eps
        v = v * torch.exp(scale) - 0.5 * self.eps * (torch.exp(transformed) * grad - translation)
        return v, scale.sum(dim=1)

    def update_position_forward(self, x, v, tvec, mask, mask_inv):
        """:511-534"""
        scale, translation, transformed = self.position_fn(v, mask * x, tvec)
        scale = scale * self.eps
        transformed = transformed * self.eps
        tmp = x * torch.exp(scale) + self.eps * (torch.exp(transformed) * v + translation)
        x = mask * x + mask_inv * tmp
        return x, (mask_inv * scale).sum(dim=1)

    def update_momentum_backward(self, x, v, beta, tvec):
        """:537-561"""
        grad = self.grad_potential(x, beta)
        scale, translation, transformed = self.momentum_fn(x, grad, tvec)
        scale = scale * (-0.5 * self.eps)
        transformed = transformed * self.eps
        v = torch.exp(scale) * (v + 0.5 * self.eps * (torch.exp(transformed) * grad - translation))
        return v, scale.sum(dim=1)

    def update_position_backward(self, x, v, tvec, mask, mask_inv):
        """:565-590"""
        scale, translation, transformed = self.position_fn(v, mask * x, tvec)
        scale = scale * (-self.eps)
        transformed = transformed * self.eps
        tmp = torch.exp(scale) * (x - self.eps * (torch.exp(transformed) * v + translation))
        x = x * mask + mask_inv * tmp
        return x, (mask_inv * scale).sum(dim=1)

    # -- leapfrog steps
    def forward_lf(self, x, v, beta, step):
        """:412-445"""
        tvec = self._tvec(step, x.shape[0])
        m, mi = self._masks(step)
        v, ld = self.update_momentum_forward(x, v, beta, tvec)
        sumlogdet = 0. + ld
        x, ld = self.update_position_forward(x, v, tvec, m, mi)
        sumlogdet = sumlogdet + ld
        x, ld = self.update_position_forward(x, v, tvec, mi, m)
        sumlogdet = sumlogdet + ld
        v, ld = self.update_momentum_forward(x, v, beta, tvec)
        sumlogdet = sumlogdet + ld
        return x, v, sumlogdet

    def backward_lf(self, x, v, beta, step):
        """:447-483 (reversed index M - step - 1; masks applied inverse-first)."""
        rstep = self.num_steps - step - 1
        tvec = self._tvec(rstep, x.shape[0])
        m, mi = self._masks(rstep)
        v, ld = self.update_momentum_backward(x, v, beta, tvec)
        sumlogdet = 0. + ld
        x, ld = self.update_position_backward(x, v, tvec, mi, m)
        sumlogdet = sumlogdet + ld
        x, ld = self.update_position_backward(x, v, tvec, m, mi)
        sumlogdet = sumlogdet + ld
        v, ld = self.update_momentum_backward(x, v, beta, tvec)
        sumlogdet = sumlogdet + ld
        return x, v, sumlogdet

    def compute_accept_prob(self, x0, v0, x1, v1, sumlogdet, beta):
        """:592-609"""
        old_h = hamiltonian(x0, v0, beta, self.T, self.X)
        new_h = hamiltonian(x1, v1, beta, self.T, self.X)
        prob = torch.exp(torch.minimum(old_h - new_h + sumlogdet, torch.zeros_like(old_h)))
        return torch.where(torch.isfinite(prob), prob, torch.zeros_like(prob))

    def transition_kernel(self, x, beta, momentum, forward=True):
        """:261-313 with the momentum draw injected."""
        x = _as(x, self.dtype)
        v0 = _as(momentum, self.dtype)
        beta = torch.as_tensor(np.float32(beta)).to(self.dtype)
        lf = self.forward_lf if forward else self.backward_lf
        xp, vp = x, v0
        sumlogdet = torch.zeros(x.shape[0], dtype=self.dtype)
        for t in range(self.num_steps):
            xp, vp, j = lf(xp, vp, beta, t)
            sumlogdet = sumlogdet + j
        return xp, vp, self.compute_accept_prob(x, v0, xp, vp, sumlogdet, beta)

    def apply_transition(self, x, beta, momentum_f, momentum_b, direction_uniform, accept_uniform):
        """:195-259: both directions are always integrated and one is discarded."""
        x = _as(x, self.dtype)
        x_f, v_f, p_f = self.transition_kernel(x, beta, momentum_f, forward=True)
        x_b, v_b, p_b = self.transition_kernel(x, beta, momentum_b, forward=False)
        fm = (_as(direction_uniform, torch.float32) > 0.5).to(self.dtype)
        bm = 1. - fm
        x_post = fm[:, None] * x_f + bm[:, None] * x_b
        v_post = fm[:, None] * v_f + bm[:, None] * v_b
        p = fm * p_f + bm * p_b
        am = (p.to(torch.float32) > _as(accept_uniform, torch.float32)).to(self.dtype)   # strict >
        x_out = am[:, None] * x_post + (1. - am)[:, None] * x
        return x_post, v_post, p, x_out


def wrap_2pi(x):
    """Host-side np.mod(x_out, 2*pi) between MC steps.  gauge_model.py:1180, 1388."""
    return np.mod(np.asarray(x, dtype=np.float32), np.float32(TWO_PI)).astype(np.float32)


def update_beta(step, beta_init, beta_final, train_steps):
    """1/beta linear annealing.  gauge_model.py:1039-1046."""
    temp = ((1. / beta_init - 1. / beta_final) * (1. - step / float(train_steps)) + 1. / beta_final)
    return 1. / temp


def u1_plaq_exact(beta):
    """I1(beta)/I0(beta).  lattice/lattice.py:31-33."""
    from scipy.special import i0, i1
    return i1(beta) / i0(beta)
