"""GaugeDynamics: host-side mirror of the reference's dynamics engine.

Same constructor kwargs, attributes, method names and return tuples as
l2hmc/dynamics/gauge_dynamics.py:42-709 of the reference; every arithmetic step runs in the
sm_100a kernels behind include/l2hmc_b200.h.  There is no eager / CPU fallback.

Differences from the reference, all deliberate and documented in DESIGN.md:
  * random draws can be INJECTED (momentum_f, momentum_b, direction_uniform, accept_uniform)
    so results can be compared with the reference on identical inputs; when they are not
    given they come from a counter-based Philox stream keyed by (seed, global chain index,
    MC step) -- TF's own Philox stream cannot be reproduced;
  * apply_transition integrates only the direction each chain selects.  The reference
    integrates both and blends with exact 0/1 masks (:211-241), so the selected result is
    bit-identical unless the discarded direction is non-finite;
  * hmc=True on a lattice shape with a fused kernel runs the whole MC step in ONE launch.
"""
import ctypes

import numpy as np
import numpy.random as npr
import torch

from . import _lib
from .lattice import to_device


def _f32(x):
    return float(np.float32(x))


def time_table(num_steps):
    """(M, 2) fp32 rows [cos, sin](2 pi i / M), each step rounded to fp32 as in the graph
    (gauge_dynamics.py:625-633: i is an fp32 loop counter, 2*pi folds to fp32)."""
    rows = []
    for i in range(num_steps):
        arg = np.float32(np.float32(2 * np.pi) * np.float32(i)) / np.float32(num_steps)
        rows.append([np.cos(np.float32(arg)), np.sin(np.float32(arg))])
    return np.asarray(rows, dtype=np.float32).reshape(num_steps, 2)


def construct_masks(x_dim, num_steps):
    """(M, D) random half masks drawn from the NumPy legacy GLOBAL stream with the reference's
    exact call sequence (gauge_dynamics.py:651-661; seeded by gauge_model.py:62)."""
    mask_per_step = []
    for _ in range(num_steps):
        idx = npr.permutation(np.arange(x_dim))[:x_dim // 2]
        mask = np.zeros((x_dim,))
        mask[idx] = 1
        mask_per_step.append(mask)
    return np.stack(mask_per_step).astype(np.float32)


class GaugeDynamics(object):
    """Dynamics engine of the L2HMC sampler (gauge_dynamics.py:42-119)."""

    def __init__(self, lattice, potential_fn=None, **kwargs):
        self.lattice = lattice
        self.potential = potential_fn if potential_fn is not None else lattice.get_energy_function()
        self.batch_size = lattice.samples.shape[0] if hasattr(lattice, "samples") else None
        self.x_dim = lattice.num_links
        self.T, self.X = lattice.time_size, lattice.space_size

        # defaults as GaugeModel._create_dynamics passes them (gauge_model.py:596-603)
        self.num_steps = 5
        self.network_arch = 'generic'
        self.hmc = False
        self.eps_trainable = True
        self.data_format = 'channels_last'
        self.seed = 0            # Philox key of the free-running mode
        self.chain_offset = 0    # global index of local chain 0 (chain sharding)
        self.native_step = True  # nets: one C call enqueues the whole MC step (csrc/l2hmc_step.cu)
        self._step_ws = None
        self.use_fused = True    # one-launch HMC path when the shape has a kernel
        self.fuse_updates = True # L2HMC: sub-updates run in the head GEMM's epilogue (no S,T,Q in HBM)
        self.weights = None      # optional {'XNet': {...}, 'VNet': {...}} in Keras layouts
        for key, val in kwargs.items():
            if key != 'eps':
                setattr(self, key, val)
        _lib.require_cuda()
        self.device = torch.device(kwargs.get('device', 'cuda'))
        self.eps = kwargs.get('eps', 0.4)   # property: fp32 device scalar + host copy (no sync per launch)
        self._mc_step = 0
        self._flag = torch.zeros(1, dtype=torch.int32, device=self.device)

        self._construct_time()
        self._construct_masks_while()

        if self.hmc:
            self.position_fn = lambda inp: [torch.zeros_like(inp[0]) for _ in range(3)]
            self.momentum_fn = lambda inp: [torch.zeros_like(inp[0]) for _ in range(3)]
        else:
            if self.network_arch == 'conv3D':
                self._build_conv_nets_3D()
            elif self.network_arch == 'generic':
                self._build_generic_nets()
            elif self.network_arch == 'conv2D':
                raise NotImplementedError(
                    "conv2D cannot run with GaugeDynamics in the reference either: ConvNet2D wants "
                    "4-D input and returns (N,T,X,2) while dynamics passes flat (N,D) "
                    "(network/conv_net.py:484-497); out of scope")
            else:
                raise AttributeError("`self._network_arch` must be one of "
                                     "`'conv3D', 'conv2D', 'generic'.`")

    # ------------------------------------------------------------------ construction
    def _construct_time(self):
        """[cos, sin](2 pi i / M) per step, in fp32 like the graph (gauge_dynamics.py:625-633)."""
        self._ts_host = time_table(self.num_steps)
        self.ts = torch.from_numpy(self._ts_host).to(self.device)

    def _construct_masks_while(self):
        """Random half masks from the NumPy legacy global RNG, exactly the reference's draw
        sequence (gauge_dynamics.py:651-661; the stream is seeded by gauge_model.py:62)."""
        self.set_masks(construct_masks(self.x_dim, self.num_steps))

    def set_masks(self, mask):
        """Install given (M, D) masks (parity tests; restored checkpoints)."""
        mask = np.asarray(mask, dtype=np.float32)
        assert mask.shape == (self.num_steps, self.x_dim), mask.shape
        self.mask = torch.from_numpy(np.ascontiguousarray(mask)).to(self.device)

    def _build_generic_nets(self):
        """XNet (factor 2) and VNet (factor 1), H = 4*D (gauge_dynamics.py:169-187)."""
        from .network import GenericNet
        kw = dict(x_dim=self.x_dim, links_shape=self.lattice.links.shape, num_hidden=int(4 * self.x_dim),
                  name_scope='position', factor=2., device=self.device)
        self.position_fn = GenericNet(model_name='XNet', **kw)
        kw.update(name_scope='momentum', factor=1.)
        self.momentum_fn = GenericNet(model_name='VNet', **kw)
        self._load_given_weights()

    def _build_conv_nets_3D(self):
        """ConvNet3D pair: F = space_size filters, H = 2*D (gauge_dynamics.py:121-143)."""
        from .network import ConvNet3D
        bs = self.batch_size if self.batch_size is not None else 1
        kw = dict(_input_shape=(bs, *self.lattice.links.shape), links_shape=self.lattice.links.shape,
                  x_dim=self.lattice.num_links, factor=2., spatial_size=self.lattice.space_size,
                  num_hidden=2 * self.lattice.num_links, num_filters=int(self.lattice.space_size),
                  filter_sizes=[(3, 3, 2), (2, 2, 2)], name_scope='position',
                  data_format=self.data_format, device=self.device)
        self.position_fn = ConvNet3D(model_name='XNet', **kw)
        kw.update(name_scope='momentum', factor=1.)
        self.momentum_fn = ConvNet3D(model_name='VNet', **kw)
        self._load_given_weights()

    def _load_given_weights(self):
        if self.weights:
            self.position_fn.load_weights(self.weights['XNet'])
            self.momentum_fn.load_weights(self.weights['VNet'])

    @property
    def variables(self):
        out = [self.eps]
        if not self.hmc:
            out += self.position_fn.variables + self.momentum_fn.variables
        return out

    trainable_variables = variables

    # ------------------------------------------------------------------ small helpers
    def _lib_call(self, fn, *args):
        _lib.check(fn(*args, _lib.stream_ptr()))

    def _x(self, a):
        t = to_device(a, self.device)
        return t.reshape(t.shape[0], self.x_dim) if t.dim() != 2 else t

    def _beta(self, beta):
        return _f32(beta.item() if isinstance(beta, torch.Tensor) else beta)

    @property
    def eps(self):
        """Step size: trainable fp32 scalar (gauge_dynamics.py:89-95), kept on the device."""
        return self._eps_dev

    @eps.setter
    def eps(self, value):
        self._eps_host = _f32(value.item() if isinstance(value, torch.Tensor) else value)
        self._eps_dev = torch.tensor(self._eps_host, dtype=torch.float32, device=self.device)

    def _eps(self):
        if self._eps_host is None:          # stale after an optimiser step (training.FlatParams.sync_to_dynamics)
            self._eps_host = float(self._eps_dev.item())
        return self._eps_host

    def raise_if_nonfinite(self):
        """Device-side mirror of tf.check_numerics (gauge_dynamics.py:26-28): one sync."""
        if int(self._flag.item()) != 0:
            self._flag.zero_()
            raise FloatingPointError("exp() produced a non-finite value in the leapfrog update "
                                     "(the reference raises InvalidArgumentError from check_numerics)")

    def _format_time(self, i, tile=1):
        """(tile, 2) rows [cos, sin](2 pi i / M) (gauge_dynamics.py:625-633)."""
        i = float(i.item() if isinstance(i, torch.Tensor) else i)
        arg = np.float32(2 * np.pi) * np.float32(i) / np.float32(self.num_steps)
        row = torch.tensor([[np.cos(np.float32(arg)), np.sin(np.float32(arg))]], dtype=torch.float32,
                           device=self.device)
        return row.repeat(int(tile), 1)

    def _get_time(self, i):
        return self.ts[int(i)][None, :]

    def _get_mask_while(self, step):
        """(m, 1 - m) rows of the step (gauge_dynamics.py:671-673)."""
        m = self.mask[int(step.item() if isinstance(step, torch.Tensor) else step)]
        return m, 1. - m

    # ------------------------------------------------------------------ energies / force
    def potential_energy(self, position, beta):
        """beta * S(x) (gauge_dynamics.py:675-681)."""
        return self._beta(beta) * self.potential(self._x(position))

    def kinetic_energy(self, v):
        """0.5 * sum v^2 (gauge_dynamics.py:683-689)."""
        v = self._x(v)
        out = torch.empty(v.shape[0], dtype=torch.float32, device=v.device)
        self._lib_call(_lib.get_lib().l2hmc_kinetic_energy, _lib.dptr(v), v.shape[0], v.shape[1], _lib.dptr(out))
        return out

    def hamiltonian(self, position, momentum, beta):
        """beta*S(x) + 0.5*sum v^2 in one pass (gauge_dynamics.py:691-696)."""
        x, v = self._x(position), self._x(momentum)
        out = torch.empty(x.shape[0], dtype=torch.float32, device=x.device)
        self._lib_call(_lib.get_lib().l2hmc_u1_hamiltonian, _lib.dptr(x), _lib.dptr(v), self._beta(beta),
                       x.shape[0], self.T, self.X, _lib.dptr(out))
        return out

    def grad_potential(self, position, beta, check_numerics=True):
        """d(beta*S)/dx: analytic force stencil in place of tf.gradients (gauge_dynamics.py:698-709)."""
        return self.lattice.grad_action(self._x(position), self._beta(beta))

    # ------------------------------------------------------------------ the four sub-updates
    def _momentum_update(self, position, momentum, beta, t, direction, logdet=None, alt=None):
        x, v = self._x(position), self._x(momentum)
        N, D = x.shape
        grad = self.grad_potential(x, beta)
        ld = torch.zeros(N, dtype=torch.float32, device=x.device) if logdet is None else logdet
        if not self.hmc and self.fuse_updates and hasattr(self.momentum_fn, 'momentum_update'):
            return self.momentum_fn.momentum_update(x, v, grad, t, self._eps(), direction, ld, self._flag,
                                                    alt=alt), ld
        assert alt is None, "mixed-direction batches need the fused network updates"
        if self.hmc:
            S = Tr = Q = None
        else:
            S, Tr, Q = self.momentum_fn([x, grad, t])
        out = torch.empty_like(v)
        self._lib_call(_lib.get_lib().l2hmc_update_momentum, _lib.dptr(v), _lib.dptr(grad), _lib.dptr(S),
                       _lib.dptr(Tr), _lib.dptr(Q), self._eps(), direction, N, D, _lib.dptr(out), _lib.dptr(ld),
                       _lib.dptr(self._flag, torch.int32))
        return out, ld

    def _position_update(self, position, momentum, t, mask_row, invert, direction, logdet=None, alt=None):
        x, v = self._x(position), self._x(momentum)
        N, D = x.shape
        ld = torch.zeros(N, dtype=torch.float32, device=x.device) if logdet is None else logdet
        if not self.hmc and self.fuse_updates and hasattr(self.position_fn, 'position_update'):
            return self.position_fn.position_update(x, v, t, mask_row, invert, self._eps(), direction, ld,
                                                    self._flag, alt=alt), ld
        assert alt is None, "mixed-direction batches need the fused network updates"
        if self.hmc:
            S = Tr = Q = None
        else:
            mx = torch.empty_like(x)
            self._lib_call(_lib.get_lib().l2hmc_mask_mul, _lib.dptr(x), _lib.dptr(mask_row), int(invert), N, D,
                           _lib.dptr(mx))
            S, Tr, Q = self.position_fn([v, mx, t])
        out = torch.empty_like(x)
        self._lib_call(_lib.get_lib().l2hmc_update_position, _lib.dptr(x), _lib.dptr(v), _lib.dptr(S),
                       _lib.dptr(Tr), _lib.dptr(Q), _lib.dptr(mask_row), int(invert), self._eps(), direction, N, D,
                       _lib.dptr(out), _lib.dptr(ld), _lib.dptr(self._flag, torch.int32))
        return out, ld

    def _mask_args(self, mask, mask_inv):
        """The reference passes (mask, mask_inv) tensors; map them back to (row, invert)."""
        m = to_device(mask, self.device).reshape(-1).contiguous()
        return m, 0

    def _update_momentum_forward(self, position, momentum, beta, t):
        """(gauge_dynamics.py:486-508) -> (momentum, logdet)."""
        return self._momentum_update(position, momentum, beta, t, +1)

    def _update_momentum_backward(self, position, momentum, beta, t):
        """(gauge_dynamics.py:537-561) -> (momentum, logdet)."""
        return self._momentum_update(position, momentum, beta, t, -1)

    def _update_position_forward(self, position, momentum, t, mask, mask_inv):
        """(gauge_dynamics.py:511-534) -> (position, logdet)."""
        m, inv = self._mask_args(mask, mask_inv)
        return self._position_update(position, momentum, t, m, inv, +1)

    def _update_position_backward(self, position, momentum, t, mask, mask_inv):
        """(gauge_dynamics.py:565-590) -> (position, logdet)."""
        m, inv = self._mask_args(mask, mask_inv)
        return self._position_update(position, momentum, t, m, inv, -1)

    # ------------------------------------------------------------------ leapfrog steps
    def _lf(self, position, momentum, beta, step, direction, ld=None):
        """One augmented leapfrog step (gauge_dynamics.py:412-483): v-half, x(m), x(mbar), v-half;
        the backward step uses the reversed index and applies the masks inverse-first."""
        x, v = self._x(position), self._x(momentum)
        step = int(step.item() if isinstance(step, torch.Tensor) else step)
        idx = step if direction > 0 else self.num_steps - step - 1
        t = (float(self._ts_host[idx, 0]), float(self._ts_host[idx, 1]))   # host floats: no sync per launch
        row = self.mask[idx]
        if ld is None:
            ld = torch.zeros(x.shape[0], dtype=torch.float32, device=x.device)
        first_inv, second_inv = (0, 1) if direction > 0 else (1, 0)
        v, ld = self._momentum_update(x, v, beta, t, direction, ld)
        x, ld = self._position_update(x, v, t, row, first_inv, direction, ld)
        x, ld = self._position_update(x, v, t, row, second_inv, direction, ld)
        v, ld = self._momentum_update(x, v, beta, t, direction, ld)
        return x, v, ld

    def _lf_mixed(self, x, v, beta, step, backward_rows, ld):
        """One leapfrog step of a batch whose chains run in different directions: chains with
        backward_rows[n] == 0 take the forward step `step`, the others the backward step (reversed
        index, inverse mask first) -- both kinds in the same launches (l2hmc_net_call.row_set)."""
        from .network import AltSet
        ib = self.num_steps - step - 1
        tf_ = (float(self._ts_host[step, 0]), float(self._ts_host[step, 1]))
        tb_ = (float(self._ts_host[ib, 0]), float(self._ts_host[ib, 1]))
        row_f, row_b = self.mask[step], self.mask[ib]
        v, ld = self._momentum_update(x, v, beta, tf_, +1, ld, alt=AltSet(backward_rows, tb_, -1))
        x, ld = self._position_update(x, v, tf_, row_f, 0, +1, ld, alt=AltSet(backward_rows, tb_, -1, row_b, 1))
        x, ld = self._position_update(x, v, tf_, row_f, 1, +1, ld, alt=AltSet(backward_rows, tb_, -1, row_b, 0))
        v, ld = self._momentum_update(x, v, beta, tf_, +1, ld, alt=AltSet(backward_rows, tb_, -1))
        return x, v, ld

    def _forward_lf(self, position, momentum, beta, step):
        """(gauge_dynamics.py:412-445) -> (position, momentum, sumlogdet)."""
        return self._lf(position, momentum, beta, step, +1)

    def _backward_lf(self, position, momentum, beta, step):
        """(gauge_dynamics.py:447-483) -> (position, momentum, sumlogdet)."""
        return self._lf(position, momentum, beta, step, -1)

    def _compute_accept_prob(self, position, momentum, position_post, momentum_post, sumlogdet, beta):
        """exp(min(H_old - H_new + sumlogdet, 0)), non-finite -> 0 (gauge_dynamics.py:592-609)."""
        old_h = self.hamiltonian(position, momentum, beta)
        new_h = self.hamiltonian(position_post, momentum_post, beta)
        p = torch.empty_like(old_h)
        sld = None if sumlogdet is None else to_device(sumlogdet, self.device)
        self._lib_call(_lib.get_lib().l2hmc_accept_prob, _lib.dptr(old_h), _lib.dptr(new_h), _lib.dptr(sld),
                       old_h.shape[0], _lib.dptr(p), _lib.dptr(self._flag, torch.int32))
        return p

    # ------------------------------------------------------------------ random draws
    def _draws(self, N, want_normals=True, want_uniforms=True):
        """Momenta / uniforms of the current MC step from the Philox stream (same numbers the
        fused kernel generates in registers)."""
        D = self.x_dim
        z = torch.empty(N, D, dtype=torch.float32, device=self.device) if want_normals else None
        ud = torch.empty(N, dtype=torch.float32, device=self.device) if want_uniforms else None
        ua = torch.empty(N, dtype=torch.float32, device=self.device) if want_uniforms else None
        self._lib_call(_lib.get_lib().l2hmc_philox_draws, int(self.seed), int(self._mc_step), int(self.chain_offset),
                       N, D, _lib.dptr(z), _lib.dptr(ud), _lib.dptr(ua))
        return z, ud, ua

    # ------------------------------------------------------------------ transition kernel
    def _fused_ok(self):
        return self.hmc and self.use_fused and (self.T, self.X) in _lib.FUSED_HMC_SHAPES

    def _fused_mc(self, x, beta, v_f, v_b, u_dir, u_acc, default_dir, want_post=True, want_out=True,
                  wrap=False, want_h=False, want_obs=False):
        N, D = x.shape
        dev = x.device
        new = lambda *shape: torch.empty(*shape, dtype=torch.float32, device=dev)
        x_post = new(N, D) if want_post else None
        v_post = new(N, D) if want_post else None
        x_out = new(N, D) if want_out else None
        p = new(N)
        acc = torch.empty(N, dtype=torch.uint8, device=dev)
        h0 = new(N) if want_h else None
        h1 = new(N) if want_h else None
        a = _lib.HmcMcArgs()
        a.x, a.v_f, a.v_b = _lib.dptr(x), _lib.dptr(v_f), _lib.dptr(v_b)
        a.u_dir, a.u_acc = _lib.dptr(u_dir), _lib.dptr(u_acc)
        a.beta, a.eps, a.num_steps = self._beta(beta), self._eps(), int(self.num_steps)
        a.default_dir, a.wrap_2pi = int(default_dir), int(bool(wrap))
        a.seed, a.mc_step, a.chain0 = int(self.seed), int(self._mc_step), int(self.chain_offset)
        a.N, a.T, a.X = N, self.T, self.X
        a.x_post, a.v_post, a.accept_prob, a.x_out = _lib.dptr(x_post), _lib.dptr(v_post), _lib.dptr(p), _lib.dptr(x_out)
        a.accepted, a.h_old, a.h_new = _lib.dptr(acc, torch.uint8), _lib.dptr(h0), _lib.dptr(h1)
        a.nonfinite_flag = _lib.dptr(self._flag, torch.int32)
        if want_obs:   # observables of the state each chain ends in, fused into the same launch
            self.last_observables = (new(N), new(N), torch.empty(N, dtype=torch.int32, device=dev))
            a.obs_action, a.obs_plaq = _lib.dptr(self.last_observables[0]), _lib.dptr(self.last_observables[1])
            a.obs_charge = _lib.dptr(self.last_observables[2], torch.int32)
        with torch.cuda.device(dev):
            self._lib_call(_lib.get_lib().l2hmc_u1_hmc_mc_step, ctypes.byref(a))
        return x_post, v_post, p, x_out, acc, h0, h1

    def transition_kernel(self, position, beta, forward=True, momentum=None):
        """M leapfrog steps one way + accept probability (gauge_dynamics.py:261-313).
        `momentum` injects the tf.random_normal draw of :269."""
        x = self._x(position)
        N = x.shape[0]
        direction = +1 if forward else -1
        if momentum is None:
            v0, _, _ = self._draws(N, want_uniforms=False)
            self._mc_step += 1
        else:
            v0 = self._x(momentum)
        if self._fused_ok():
            xp, vp, p, _, _, _, _ = self._fused_mc(x, beta, v0, None, None, None, direction, want_out=False)
            return xp, vp, p
        xp, vp = x, v0
        sumlogdet = torch.zeros(N, dtype=torch.float32, device=x.device)
        for step in range(self.num_steps):   # the kernels accumulate into sumlogdet in place
            xp, vp, sumlogdet = self._lf(xp, vp, beta, step, direction, sumlogdet)
        p = self._compute_accept_prob(x, v0, xp, vp, sumlogdet, beta)
        return xp, vp, p

    # ------------------------------------------------------------------ MC step
    def apply_transition(self, position, beta, momentum_f=None, momentum_b=None, direction_uniform=None,
                         accept_uniform=None, wrap=False, observables=False):
        """Propose and accept/reject (gauge_dynamics.py:195-259).

        Returns (position_post, momentum_post, accept_prob, position_out).  The four keyword
        arguments inject the reference's in-graph draws (:269 twice, :223, :246); all-None runs the
        free-running Philox stream.  wrap=True also applies the host loop's np.mod(x_out, 2 pi)
        (gauge_model.py:1180,1388) to position_out.  observables=True leaves (action, avg_plaq,
        integer charge) of position_out in self.last_observables (lattice.py:285-313).
        """
        x = self._x(position)
        N, D = x.shape
        inject = momentum_f is not None
        if inject:
            if direction_uniform is None or accept_uniform is None:
                raise ValueError("injecting momentum_f also needs direction_uniform and accept_uniform")
            v_f = self._x(momentum_f)
            v_b = self._x(momentum_b) if momentum_b is not None else v_f
            u_dir = to_device(direction_uniform, self.device).reshape(-1)
            u_acc = to_device(accept_uniform, self.device).reshape(-1)
        else:
            v_f = v_b = u_dir = u_acc = None

        if self._fused_ok():
            x_post, v_post, p, x_out, self.last_accepted, _, _ = self._fused_mc(
                x, beta, v_f, v_b if inject and momentum_b is not None else None, u_dir, u_acc, +1, wrap=wrap,
                want_obs=observables)
            if not inject:
                self._mc_step += 1
            return x_post, v_post, p, x_out

        if self.fuse_updates and self.native_step and not self.hmc and hasattr(self.position_fn, '_weights_struct'):
            return self._native_mc_step(x, beta, v_f if inject else None, v_b if inject else None,
                                        u_dir if inject else None, u_acc if inject else None, wrap, observables)
        if not inject:
            z, u_dir, u_acc = self._draws(N)
            self._mc_step += 1
            v_f = v_b = z
        if self.fuse_updates and not self.hmc and hasattr(self.position_fn, 'position_update'):
            # every chain runs only the direction it selects, all of them in the same launches
            lib = _lib.get_lib()
            if v_b is v_f:
                v0 = v_f
            else:                      # injected, distinct draws: v0 = v_f where u_dir > 0.5 else v_b (exact 0/1 blend)
                v0, ones = torch.empty_like(x), torch.ones(N, dtype=torch.float32, device=x.device)
                self._lib_call(lib.l2hmc_mh_select, _lib.dptr(v_f), _lib.dptr(v_f), _lib.dptr(v_b), None, None,
                               _lib.dptr(ones), _lib.dptr(ones), _lib.dptr(u_dir), None, N, D, 0, _lib.dptr(v0),
                               None, None, None, None)
            backward_rows = (u_dir <= 0.5).to(torch.uint8)
            x_post, v_post = x, v0
            sumlogdet = torch.zeros(N, dtype=torch.float32, device=x.device)
            for step in range(self.num_steps):
                x_post, v_post, sumlogdet = self._lf_mixed(x_post, v_post, beta, step, backward_rows, sumlogdet)
            p_post = self._compute_accept_prob(x, v0, x_post, v_post, sumlogdet, beta)
            return self._finish_transition(x, x_post, v_post, p_post, u_acc, wrap, observables)
        # unfused K2 composition: gather the chains of each direction, integrate, scatter
        fwd = u_dir > 0.5
        idx_f = torch.nonzero(fwd, as_tuple=False).reshape(-1)
        idx_b = torch.nonzero(~fwd, as_tuple=False).reshape(-1)
        x_post, v_post = torch.empty_like(x), torch.empty_like(x)
        p_post = torch.empty(N, dtype=torch.float32, device=x.device)
        for idx, vsrc, forward in ((idx_f, v_f, True), (idx_b, v_b, False)):
            if idx.numel() == 0:
                continue
            xs, vs, ps = self.transition_kernel(x.index_select(0, idx), beta, forward=forward,
                                                momentum=vsrc.index_select(0, idx))
            x_post.index_copy_(0, idx, xs)
            v_post.index_copy_(0, idx, vs)
            p_post.index_copy_(0, idx, ps)
        return self._finish_transition(x, x_post, v_post, p_post, u_acc, wrap, observables)

    def _native_mc_step(self, x, beta, v_f, v_b, u_dir, u_acc, wrap, observables):
        """The whole MC step enqueued by one C call (csrc/l2hmc_step.cu: l2hmc_u1_l2hmc_mc_step)."""
        N, D = x.shape
        dev = x.device
        new = lambda *shape: torch.empty(*shape, dtype=torch.float32, device=dev)
        x_post, v_post, x_out, p = new(N, D), new(N, D), new(N, D), new(N)
        acc = torch.empty(N, dtype=torch.uint8, device=dev)
        a = _lib.L2hmcMcArgs()
        a.N, a.T, a.X, a.M = N, self.T, self.X, int(self.num_steps)
        a.eps, a.beta = self._eps(), self._beta(beta)
        mask = self.mask if self.mask.is_contiguous() else self.mask.contiguous()
        a.masks = _lib.dptr(mask)
        ts = np.ascontiguousarray(self._ts_host, dtype=np.float32)
        a.ts = ts.ctypes.data_as(ctypes.c_void_p)
        xs, vs = self.position_fn._weights_struct(), self.momentum_fn._weights_struct()
        a.xnet, a.vnet = ctypes.pointer(xs), ctypes.pointer(vs)
        if hasattr(self.position_fn, '_conv'):
            convs = [self.position_fn._conv("v"), self.position_fn._conv("x"), self.momentum_fn._conv("v"),
                     self.momentum_fn._conv("x")]
            a.xconv_in0, a.xconv_in1, a.vconv_in0, a.vconv_in1 = (ctypes.pointer(c) for c in convs)
        a.net_path = int(self.position_fn.path)
        a.x, a.v_f, a.v_b = _lib.dptr(x), _lib.dptr(v_f), _lib.dptr(None if v_b is v_f else v_b)
        a.u_dir, a.u_acc = _lib.dptr(u_dir), _lib.dptr(u_acc)
        a.seed, a.mc_step, a.chain0 = int(self.seed), int(self._mc_step), int(self.chain_offset)
        a.wrap_2pi = int(bool(wrap))
        a.x_post, a.v_post, a.accept_prob, a.x_out = _lib.dptr(x_post), _lib.dptr(v_post), _lib.dptr(p), _lib.dptr(x_out)
        a.accepted = _lib.dptr(acc, torch.uint8)
        if observables:
            self.last_observables = (new(N), new(N), torch.empty(N, dtype=torch.int32, device=dev))
            a.obs_action, a.obs_plaq = _lib.dptr(self.last_observables[0]), _lib.dptr(self.last_observables[1])
            a.obs_charge = _lib.dptr(self.last_observables[2], torch.int32)
        a.nonfinite_flag = _lib.dptr(self._flag, torch.int32)
        lib = _lib.get_lib()
        need = int(lib.l2hmc_u1_l2hmc_mc_workspace_bytes(ctypes.byref(a)))
        if self._step_ws is None or self._step_ws.numel() < need:
            self._step_ws = torch.empty(need, dtype=torch.uint8, device=dev)
        a.workspace, a.workspace_bytes = ctypes.c_void_p(self._step_ws.data_ptr()), need
        with torch.cuda.device(dev):
            self._lib_call(lib.l2hmc_u1_l2hmc_mc_step, ctypes.byref(a))
        if v_f is None:
            self._mc_step += 1
        self.last_accepted = acc
        return x_post, v_post, p, x_out

    def _finish_transition(self, x, x_post, v_post, p_post, u_acc, wrap, observables):
        """Metropolis-Hastings select (gauge_dynamics.py:243-259) [+ np.mod 2 pi] [+ observables]."""
        N, D = x.shape
        x_out = torch.empty_like(x)
        acc = torch.empty(N, dtype=torch.uint8, device=x.device)
        self._lib_call(_lib.get_lib().l2hmc_mh_select, _lib.dptr(x), _lib.dptr(x_post), None, None, None,
                       _lib.dptr(p_post), None, None, _lib.dptr(u_acc), N, D, int(bool(wrap)), None, None, None,
                       _lib.dptr(x_out), _lib.dptr(acc, torch.uint8))
        self.last_accepted = acc
        if observables:
            a_, p_, _, q_ = self.lattice.observables(x_out)
            self.last_observables = (a_, p_, q_)
        return x_post, v_post, p_post, x_out

    def call(self, position, beta, **inject):
        return self.apply_transition(position, beta, **inject)

    __call__ = call

    def hmc_step(self, position, momentum, beta, direction=None):
        """One plain leapfrog step for all chains in one launch (the 16 B/link-update kernel).
        `direction` is an optional int8 (N,) tensor of +1/-1."""
        x, v = self._x(position), self._x(momentum)
        xo, vo = torch.empty_like(x), torch.empty_like(v)
        self._lib_call(_lib.get_lib().l2hmc_u1_hmc_step, _lib.dptr(x), _lib.dptr(v), self._beta(beta), self._eps(),
                       _lib.dptr(direction, torch.int8), x.shape[0], self.T, self.X, _lib.dptr(xo), _lib.dptr(vo))
        return xo, vo
