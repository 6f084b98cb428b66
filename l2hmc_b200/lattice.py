"""GaugeLattice: host-side mirror of the reference's lattice class for the 2D U(1) model.

Same constructor, attribute names and method names as the reference
(l2hmc/lattice/lattice.py:61-362); the arithmetic runs in the sm_100a kernels of
csrc/stencil.cu through the C ABI (include/l2hmc_b200.h).  Inputs may be numpy arrays or
torch tensors, flat (N, D) or (N, T, X, 2) (the reference reshapes, :297-298, 351-352);
results are torch CUDA tensors.

Out of scope: SU(2)/SU(3) link types and the per-site NumPy
Metropolis helpers (:404-562) -- unreachable from the U(1) hot path.
"""
import numpy as np
import torch

from . import _lib

TWO_PI = 2 * np.pi


def u1_plaq_exact(beta):
    """Expected average plaquette I1(beta)/I0(beta) (lattice.py:31-33)."""
    from scipy.special import i0, i1
    return i1(beta) / i0(beta)


def pbc(tup, shape):
    """tup % shape, periodic boundary conditions (lattice.py:35-37)."""
    return list(np.mod(tup, shape))


def project_angle(x):
    """x - 2pi*floor((x + pi)/2pi) for a torch tensor (lattice.py:47-49).  Host/torch helper
    only; the device version is fused into the observables kernel."""
    return x - TWO_PI * torch.floor((x + np.pi) / TWO_PI)


def to_device(a, device=None):
    """numpy / torch (any device, any float dtype) -> contiguous fp32 CUDA tensor."""
    _lib.require_cuda()
    if isinstance(a, torch.Tensor):
        t = a
    else:
        t = torch.from_numpy(np.ascontiguousarray(np.asarray(a, dtype=np.float32)))
    dev = torch.device(device) if device is not None else (t.device if t.is_cuda else torch.device("cuda"))
    return t.to(device=dev, dtype=torch.float32).contiguous()


class GaugeLattice(object):
    """Lattice with U(1) gauge field on the links (lattice.py:61-157)."""

    def __init__(self, time_size, space_size, dim=2, link_type='U1', num_samples=None, rand=False):
        assert link_type.upper() in ['U1', 'SU2', 'SU3'], (
            "Invalid link_type. Possible values: U1', 'SU2', 'SU3'")
        if link_type.upper() != 'U1' or dim != 2:
            raise NotImplementedError("l2hmc_b200 accelerates the 2D U(1) path only "
                                      "(SU(2)/SU(3) are unreachable from gauge_model.py)")
        self.time_size = time_size
        self.space_size = space_size
        self.dim = dim
        self.link_type = link_type
        self.link_shape = ()
        self._init_lattice(num_samples, rand)

    def _init_lattice(self, num_samples, rand):
        sites_shape = (self.time_size, self.space_size)
        links_shape = (self.time_size, self.space_size, self.dim)
        self.sites = np.zeros(sites_shape, dtype=np.float32)
        self.links = np.zeros(links_shape, dtype=np.float32)
        if rand:
            self.links = np.array(np.random.uniform(0, TWO_PI, links_shape), dtype=np.float32)
        self.site_idxs = self.sites.shape
        self.link_idxs = self.links.shape
        self.num_sites = int(np.cumprod(self.sites.shape)[-1])
        self.num_links = self.num_sites * self.dim
        self.num_plaquettes = self.time_size * self.space_size
        self.bases = np.eye(self.dim, dtype=int)
        if num_samples:
            self.num_samples = num_samples
            self.samples = self.get_links_samples(num_samples, rand=rand, link_type=self.link_type)
            self.samples[0] = self.links

    # ------------------------------------------------------------------ sample generation
    def _generate_links(self, rand=False, link_type=None):
        """Cold start zeros / hot start 2*pi*U[0,1) (lattice.py:198-234)."""
        if rand:
            return TWO_PI * np.random.rand(*self.links.shape)
        return np.zeros(self.links.shape)

    def get_links_samples(self, num_samples, rand=False, link_type=None):
        """`num_samples` link arrays, shape (N, T, X, 2) (lattice.py:236-241)."""
        return np.array([self._generate_links(rand, link_type) for _ in range(num_samples)])

    def iter_sites(self):
        for i in range(self.num_sites):
            indices = []
            for dim in self.site_idxs:
                indices.append(i % dim)
                i = i // dim
            yield tuple(indices)

    def iter_links(self):
        for site in self.iter_sites():
            for u in range(self.dim):
                yield tuple(list(site) + [u])

    # ------------------------------------------------------------------ device arithmetic
    def _flat(self, samples):
        x = to_device(samples)
        D = self.num_links
        if x.dim() == 1:
            x = x.reshape(1, -1)
        per_chain = int(np.prod(x.shape[1:]))
        if per_chain != D:
            raise ValueError(f"samples have {per_chain} links per chain, lattice has {D}")
        return x.reshape(x.shape[0], D)

    def get_energy_function(self, samples=None):
        """fn(samples) -> total action per chain (lattice.py:274-283)."""
        def fn(samples):
            return self.total_action(samples)
        return fn

    def total_action(self, samples):
        """S = sum (1 - cos P) per chain, (N,) (lattice.py:337-362)."""
        x = self._flat(samples)
        out = torch.empty(x.shape[0], device=x.device, dtype=torch.float32)
        with torch.cuda.device(x.device):
            _lib.check(_lib.get_lib().l2hmc_u1_action_force(
                _lib.dptr(x), 0.0, x.shape[0], self.time_size, self.space_size, _lib.dptr(out), None,
                _lib.stream_ptr()))
        return out

    def grad_action(self, samples, beta=1.0):
        """d(beta*S)/dx, (N, D): what the reference gets from tf.gradients
        (dynamics/gauge_dynamics.py:698-709)."""
        x = self._flat(samples)
        out = torch.empty_like(x)
        with torch.cuda.device(x.device):
            _lib.check(_lib.get_lib().l2hmc_u1_action_force(
                _lib.dptr(x), float(beta), x.shape[0], self.time_size, self.space_size, None, _lib.dptr(out),
                _lib.stream_ptr()))
        return out

    def action_and_grad(self, samples, beta):
        x = self._flat(samples)
        act = torch.empty(x.shape[0], device=x.device, dtype=torch.float32)
        out = torch.empty_like(x)
        with torch.cuda.device(x.device):
            _lib.check(_lib.get_lib().l2hmc_u1_action_force(
                _lib.dptr(x), float(beta), x.shape[0], self.time_size, self.space_size, _lib.dptr(act),
                _lib.dptr(out), _lib.stream_ptr()))
        return act, out

    def calc_plaq_observables(self, samples):
        """(total_action, avg_plaq, topological_charge) per chain (lattice.py:285-313);
        the charge is floor(0.1 + sum proj(P)/2pi) returned as float like the reference."""
        a, p, _, qi = self.observables(samples)
        return a, p, qi.to(torch.float32)

    def observables(self, samples):
        """(action, avg_plaq, real-valued charge, integer charge int32): the reference's four
        observable ops (gauge_model.py:683-716, lattice.py:300-311) in one pass."""
        x = self._flat(samples)
        N = x.shape[0]
        a = torch.empty(N, device=x.device, dtype=torch.float32)
        p = torch.empty_like(a)
        qf = torch.empty_like(a)
        qi = torch.empty(N, device=x.device, dtype=torch.int32)
        with torch.cuda.device(x.device):
            _lib.check(_lib.get_lib().l2hmc_u1_observables(
                _lib.dptr(x), N, self.time_size, self.space_size, _lib.dptr(a), _lib.dptr(p), _lib.dptr(qf),
                _lib.dptr(qi, torch.int32), _lib.stream_ptr()))
        return a, p, qf, qi
