"""MoG / generic-target L2HMC sampler: host-side mirror of the reference's utils/ path (config 5).

Mirrors, with the same names and argument meaning:
  utils/distributions.py:124-181  GMM(mus, sigmas, pis).get_energy_function(), gen_ring
  utils/network.py:89-114         network(x_dim, scope, factor, num_nodes) -> net([in0, in1, t, aux])
  utils/dynamics.py:34-319        Dynamics(x_dim, energy_function, trajectory_length, eps, hmc,
                                           net_factory, eps_trainable, use_temperature)
                                  .energy / .grad_energy / .hamiltonian / .forward / .backward / .p_accept
  utils/sampler.py:28-59          propose(x, dynamics, do_mh_step=True)
All arithmetic runs in csrc/mog.cu (one launch per MC step) behind l2hmc_mog_mc_step /
l2hmc_gmm_energy_force; there is no CPU fallback.  Random draws can be injected
(init_v / v_f, v_b, dir_mask, u_acc) or come from the Philox stream keyed by the global chain index.
"""
import ctypes

import numpy as np
import torch

from . import _lib
from .lattice import to_device
from .network import _StqNet


class GMM(object):
    """Gaussian mixture target (utils/distributions.py:124-149)."""

    def __init__(self, mus, sigmas, pis, device="cuda"):
        assert len(mus) == len(sigmas)
        pis = np.array(pis, dtype=np.float64)
        if np.sum(pis) != 1.0:
            pis /= pis.sum()
        self.mus = [np.asarray(m, np.float32) for m in mus]
        self.sigmas, self.pis = sigmas, pis
        self.nb_mixtures = len(pis)
        self.k = self.mus[0].shape[0]
        self.i_sigmas, self.constants = [], []
        for i, sigma in enumerate(sigmas):
            self.i_sigmas.append(np.linalg.inv(sigma).astype('float32'))
            det = np.sqrt((2 * np.pi) ** self.k * np.linalg.det(sigma)).astype('float32')
            self.constants.append((pis[i] / det).astype('float32'))
        self._dev = None
        self.device = device

    @classmethod
    def from_arrays(cls, mus, i_sigmas, constants, device="cuda"):
        """Build from the precomputed fp32 arrays (parity tests)."""
        g = cls.__new__(cls)
        g.mus = [np.asarray(m, np.float32) for m in mus]
        g.i_sigmas = [np.asarray(s, np.float32) for s in i_sigmas]
        g.constants = [np.float32(c) for c in constants]
        g.nb_mixtures, g.k = len(g.mus), g.mus[0].shape[0]
        g._dev, g.device = None, device
        return g

    def _struct(self):
        if self._dev is None:
            _lib.require_cuda()
            dev = torch.device(self.device)
            mus = torch.as_tensor(np.stack(self.mus).astype(np.float32)).to(dev).contiguous()
            isg = torch.as_tensor(np.stack(self.i_sigmas).astype(np.float32)).to(dev).contiguous()
            logc = torch.log(torch.as_tensor(np.asarray(self.constants, np.float32))).to(dev).contiguous()  # tf.log (fp32)
            s = _lib.Gmm()
            s.k, s.m = int(self.k), int(self.nb_mixtures)
            s.mus, s.i_sigmas, s.log_consts = _lib.dptr(mus), _lib.dptr(isg), _lib.dptr(logc)
            self._dev = (s, mus, isg, logc)
        return self._dev[0]

    def energy_force(self, x, temperature=1.0, want_energy=True, want_force=True):
        x = to_device(x, self.device)
        N = x.shape[0]
        e = torch.empty(N, dtype=torch.float32, device=x.device) if want_energy else None
        f = torch.empty_like(x) if want_force else None
        with torch.cuda.device(x.device):
            _lib.check(_lib.get_lib().l2hmc_gmm_energy_force(self._struct(), _lib.dptr(x), float(temperature), N,
                                                             _lib.dptr(e), _lib.dptr(f), _lib.stream_ptr()))
        return e, f

    def get_energy_function(self):
        def fn(x):
            return self.energy_force(x, want_force=False)[0]
        fn.gmm = self
        return fn


def gen_ring(r=1.0, var=1.0, nb_mixtures=2):
    """Mixture of `nb_mixtures` Gaussians on a ring (utils/distributions.py:231-243)."""
    base_points = []
    for t in range(nb_mixtures):
        c, s = np.cos(2 * np.pi * t / nb_mixtures), np.sin(2 * np.pi * t / nb_mixtures)
        base_points.append(np.array([r * c, r * s]))
    sigmas = [var * np.eye(2) for _ in range(nb_mixtures)]
    pis = [1. / nb_mixtures] * nb_mixtures
    pis[0] += 1 - sum(pis)
    return sigmas, GMM(base_points, sigmas, pis)


class MogNet(_StqNet):
    """utils/network.py:89-114: embed_1(in0) + embed_2(in1) + embed_3(t) -> relu -> linear_1 -> relu ->
    [scale_s * tanh(linear_s), linear_t, scale_f * tanh(linear_f)].  Weight names follow the gauge
    nets (v_layer = embed_1, x_layer = embed_2, t_layer = embed_3, h_layer = linear_1, ...)."""

    tanh_on_q = 1
    use_tensor_cores = False

    def __init__(self, x_dim, scope, factor, num_nodes=50, device="cuda", seed=0):
        _lib.require_cuda()
        self.name = self.name_scope = scope
        self.x_dim = self.num_features = int(x_dim)
        self.num_hidden, self.factor = int(num_nodes), float(factor)
        self.device = torch.device(device)
        self.weights, self._struct, self._ws = {}, None, None
        rng = np.random.default_rng([int(seed), sum(map(ord, scope))])
        self.load_weights(self._init_dense(self.num_features, rng))

    def _features(self, in0, in1, mask, invert, alt=None):
        m = None if mask is None else mask.reshape(-1).contiguous()
        return in0, in1, m, None

    def __call__(self, inputs, **kw):
        return self.call(list(inputs[:3]), **kw)        # the 4th entry is `aux` (always None for the MoG model)


def network(x_dim, scope, factor, num_nodes=50, **kw):
    return MogNet(x_dim, scope, factor, num_nodes, **kw)


class Dynamics(object):
    """utils/dynamics.py:34-319."""

    def __init__(self, x_dim, energy_function, trajectory_length=10, eps=0.1, hmc=False, net_factory=None,
                 eps_trainable=True, use_temperature=False, device="cuda", seed=0, chain_offset=0):
        _lib.require_cuda()
        gmm = energy_function if isinstance(energy_function, GMM) else getattr(energy_function, "gmm", None)
        if gmm is None:
            raise NotImplementedError("l2hmc_b200 Dynamics accelerates the Gaussian-mixture target only: pass a GMM "
                                      "or GMM.get_energy_function()")
        if gmm.k != x_dim:
            raise ValueError(f"x_dim = {x_dim} but the target has k = {gmm.k}")
        self.gmm, self.x_dim, self.device = gmm, int(x_dim), torch.device(device)
        gmm.device = self.device
        self.use_temperature = use_temperature
        self.temperature = 1.0                       # the reference feeds a placeholder per sess.run
        self.eps = float(np.exp(np.log(np.float32(eps))))      # eps = exp(alpha), alpha = log(eps) (:51-60)
        self.trajectory_length, self.hmc = int(trajectory_length), bool(hmc)
        self.seed, self.chain_offset, self._mc_step = int(seed), int(chain_offset), 0
        self._init_mask()
        self._flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        if hmc:
            self.XNet = self.VNet = None
        else:
            net_factory = net_factory or network
            self.XNet = net_factory(x_dim, scope='XNet', factor=2.0)
            self.VNet = net_factory(x_dim, scope='VNet', factor=1.0)

    def _init_mask(self):
        """Random half masks from the NumPy legacy global stream (:85-100)."""
        rows = []
        for _ in range(self.trajectory_length):
            ind = np.random.permutation(np.arange(self.x_dim))[:int(self.x_dim / 2)]
            m = np.zeros((self.x_dim,))
            m[ind] = 1
            rows.append(m)
        self.set_masks(np.stack(rows))

    def set_masks(self, mask):
        mask = np.ascontiguousarray(np.asarray(mask, np.float32))
        assert mask.shape == (self.trajectory_length, self.x_dim)
        self.mask = torch.from_numpy(mask).to(self.device)

    def _temp(self):
        return float(self.temperature) if self.use_temperature else 1.0

    def energy(self, x, aux=None):
        """(:227-236)"""
        return self.gmm.energy_force(x, self._temp(), want_force=False)[0]

    def grad_energy(self, x, aux=None):
        """(:241-242)"""
        return self.gmm.energy_force(x, self._temp(), want_energy=False)[1]

    def kinetic(self, v):
        v = to_device(v, self.device).reshape(-1, self.x_dim)
        out = torch.empty(v.shape[0], dtype=torch.float32, device=v.device)
        with torch.cuda.device(v.device):
            _lib.check(_lib.get_lib().l2hmc_kinetic_energy(_lib.dptr(v), v.shape[0], v.shape[1], _lib.dptr(out),
                                                           _lib.stream_ptr()))
        return out

    def mc_step(self, x, v_f=None, v_b=None, dir_mask=None, u_acc=None, want=("Lx", "Lv", "px", "x_out")):
        """One launch: trajectory in the direction each chain selects + accept (csrc/mog.cu).
        Returns a dict of CUDA tensors: Lx, Lv, px, x_out, accepted, mode."""
        x = to_device(x, self.device).reshape(-1, self.x_dim)
        N, k = x.shape
        a = _lib.MogArgs()
        vf = None if v_f is None else to_device(v_f, self.device).reshape(N, k).contiguous()
        vb = None if v_b is None else to_device(v_b, self.device).reshape(N, k).contiguous()
        dm = None if dir_mask is None else to_device(dir_mask, self.device).reshape(N).contiguous()
        ua = None if u_acc is None else to_device(u_acc, self.device).reshape(N).contiguous()
        out = {"Lx": torch.empty_like(x), "Lv": torch.empty_like(x), "x_out": torch.empty_like(x),
               "px": torch.empty(N, dtype=torch.float32, device=x.device),
               "accepted": torch.empty(N, dtype=torch.uint8, device=x.device),
               "mode": torch.empty(N, dtype=torch.int32, device=x.device)}
        a.x, a.v_f, a.v_b, a.dir_mask, a.u_acc = _lib.dptr(x), _lib.dptr(vf), _lib.dptr(vb), _lib.dptr(dm), _lib.dptr(ua)
        a.mask, a.eps, a.temperature = _lib.dptr(self.mask), self.eps, self._temp()
        a.L, a.hmc, a.N = self.trajectory_length, int(self.hmc), N
        a.seed, a.mc_step, a.chain0 = self.seed, self._mc_step, self.chain_offset
        if not self.hmc:
            a.xnet = ctypes.pointer(self.XNet._weights_struct())
            a.vnet = ctypes.pointer(self.VNet._weights_struct())
        a.Lx, a.Lv, a.px, a.x_out = _lib.dptr(out["Lx"]), _lib.dptr(out["Lv"]), _lib.dptr(out["px"]), _lib.dptr(out["x_out"])
        a.accepted, a.mode = _lib.dptr(out["accepted"], torch.uint8), _lib.dptr(out["mode"], torch.int32)
        a.nonfinite_flag = _lib.dptr(self._flag, torch.int32)
        with torch.cuda.device(self.device):
            _lib.check(_lib.get_lib().l2hmc_mog_mc_step(self.gmm._struct(), a, _lib.stream_ptr()))
        if v_f is None:
            self._mc_step += 1
        return out

    def _one_way(self, x, init_v, forward):
        N = len(x)
        if init_v is None:
            raise ValueError("forward/backward need init_v here (the reference draws it with tf.random_normal); "
                             "use mc_step() for the free-running Philox stream")
        dm = np.full((N,), 1.0 if forward else 0.0, np.float32)
        if self.hmc and not forward:
            raise NotImplementedError("hmc=True integrates forward only (utils/sampler.py:30-32)")
        o = self.mc_step(x, v_f=init_v, v_b=init_v, dir_mask=dm)
        return o["Lx"], o["Lv"], o["px"]

    def forward(self, x, init_v=None, aux=None):
        """L forward steps + accept probability (:255-283)."""
        return self._one_way(x, init_v, True)

    def backward(self, x, init_v=None, aux=None):
        """L backward steps with the reversed step index (:285-310)."""
        return self._one_way(x, init_v, False)


def propose(x, dynamics, init_v=None, aux=None, do_mh_step=False, v_f=None, v_b=None, dir_mask=None, u_acc=None):
    """utils/sampler.py:28-55.  Returns (Lx, Lv, px, outputs) with outputs = [x_out] when do_mh_step.
    v_f / v_b / dir_mask / u_acc inject the reference's in-graph draws; all-None runs on Philox."""
    if dynamics.hmc and v_f is None and init_v is not None:
        v_f = init_v
    o = dynamics.mc_step(x, v_f=v_f, v_b=v_b, dir_mask=dir_mask, u_acc=u_acc)
    return o["Lx"], o["Lv"], o["px"], ([o["x_out"]] if do_mh_step else [])


def tunneling_rate(modes):
    """Mean over chains of (mode switches) / (steps - 1) for a (steps, N) int tensor of nearest-mean
    indices (utils/trajectories.py:63-95 applied per chain)."""
    modes = torch.as_tensor(modes)
    return (modes[1:] != modes[:-1]).float().sum(dim=0) / (modes.shape[0] - 1)


def loss(x, dynamics, scale=0.1, z=None, draws_x=None, draws_z=None):
    """Value of GaussianMixtureModel's training loss (mog_model.py:324-355): Lx, px = propose(x, do_mh_step=True) and
    Lz, pz = propose(z) with z ~ N(0, 1) (drawn here with torch when not given; the reference uses tf.random_normal),
    then  loss = scale (mean 1/v1 + mean 1/v2) - (mean v1 + mean v2) / scale  with v = |jump|^2 p + 1e-4.
    draws_x / draws_z: optional dicts of injected draws for the two propose calls (v_f, v_b, dir_mask, u_acc).
    Returns (loss [1] CUDA tensor, dict(Lx, px, Lz, pz, x_out, v1, v2, z)).  The value only -- no gradient."""
    x = to_device(x, dynamics.device).reshape(-1, dynamics.x_dim).contiguous()
    N, k = x.shape
    if z is None:
        z = torch.randn(N, k, dtype=torch.float32, device=x.device)
    z = to_device(z, dynamics.device).reshape(N, k).contiguous()
    ox = dynamics.mc_step(x, **(draws_x or {}))
    oz = dynamics.mc_step(z, **(draws_z or {}))
    lib = _lib.get_lib()
    need = int(lib.l2hmc_mog_loss_workspace_bytes())
    ws = torch.empty((need + 3) // 4, dtype=torch.float32, device=x.device)
    v1, v2 = torch.empty(N, dtype=torch.float32, device=x.device), torch.empty(N, dtype=torch.float32, device=x.device)
    out = torch.empty(1, dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(lib.l2hmc_mog_loss(_lib.dptr(x), _lib.dptr(ox["Lx"]), _lib.dptr(ox["px"]), _lib.dptr(z), _lib.dptr(oz["Lx"]),
                                      _lib.dptr(oz["px"]), N, k, float(scale), _lib.dptr(v1), _lib.dptr(v2), _lib.dptr(out),
                                      _lib.dptr(ws), need, _lib.stream_ptr()))
    return out, dict(Lx=ox["Lx"], px=ox["px"], Lz=oz["Lx"], pz=oz["px"], x_out=ox["x_out"], v1=v1, v2=v2, z=z)
