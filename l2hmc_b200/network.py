"""GenericNet / ConvNet3D: host-side mirrors of the reference's S/T/Q networks.

Same constructor kwargs, layer / weight names and call convention as
l2hmc/network/generic_net.py:20-161 and l2hmc/network/conv_net.py:57-311 of the reference:
`net([in0, in1, t]) -> (scale, translation, transformation)`, each (N, D), where the FIRST
input goes through `v_layer` and the second through `x_layer` (generic_net.py:129-134).
Weights are kept as CUDA fp32 tensors in Keras layouts -- Dense kernels (in, out), Conv3D
kernels (kd, kh, kw, cin, cout) -- so arrays dumped from a reference run load unchanged
(`load_weights`).  All arithmetic runs in the sm_100a kernels behind l2hmc_net_run /
l2hmc_conv3d_features (include/l2hmc_b200.h); there is no torch or CPU fallback.

Besides the reference's `call`, the nets expose `momentum_update` / `position_update`:
the network evaluation with the leapfrog sub-update fused into the head GEMM's epilogue
(gauge_dynamics.py:486-590), which is what GaugeDynamics uses.
"""
import math

import numpy as np
import torch

from . import _lib

DENSE_LAYERS = ("x_layer", "v_layer", "t_layer", "h_layer", "scale_layer", "translation_layer",
                "transformation_layer")
CONV_LAYERS = ("conv_x1", "conv_v1", "conv_x2", "conv_v2")


def _truncated_normal(rng, shape, std):
    """tf.truncated_normal: redraw beyond two standard deviations."""
    out = rng.standard_normal(size=shape)
    bad = np.abs(out) > 2.0
    while bad.any():
        out[bad] = rng.standard_normal(size=int(bad.sum()))
        bad = np.abs(out) > 2.0
    return (out * std).astype(np.float32)


def _same_out(n, stride):
    return -(-n // stride)


class _StqNet(object):
    """Dense trunk + heads shared by both architectures."""

    tanh_on_q = 0
    use_tensor_cores = True      # pack weights for the tcgen05 path (taken when N >= 256 and the trunk is wide)
    path = _lib.NET_PATH_AUTO    # NET_PATH_FFMA / NET_PATH_TCGEN05 force one path (tests, benchmarks)

    def _init_dense(self, feat, rng):
        """`_custom_dense` initialisers (generic_net.py:149-161): truncated normal with
        std = sqrt(1.3 * 2*f / fan_in), f = factor/3, 1/3, 1/3, 1, 0.001 x 3; zero biases
        and zero coeff_* (:39-91)."""
        H, D = int(self.num_hidden), int(self.x_dim)
        w = {}

        def dense(name, fan_in, units, f):
            w[name + ".kernel"] = _truncated_normal(rng, (fan_in, units), math.sqrt(1.3 * 2. * f / fan_in))
            w[name + ".bias"] = np.zeros((units,), np.float32)
        dense("x_layer", feat, H, self.factor / 3.)
        dense("v_layer", feat, H, 1. / 3.)
        dense("t_layer", 2, H, 1. / 3.)
        dense("h_layer", H, H, 1.)
        dense("scale_layer", H, D, 0.001)
        dense("translation_layer", H, D, 0.001)
        dense("transformation_layer", H, D, 0.001)
        w["coeff_scale"] = np.zeros((1, D), np.float32)
        w["coeff_transformation"] = np.zeros((1, D), np.float32)
        return w

    # ------------------------------------------------------------------ weights
    def load_weights(self, weights):
        """Install weights from a dict of arrays keyed '<layer>.kernel' / '<layer>.bias' /
        'coeff_scale' / 'coeff_transformation' (Keras layouts)."""
        for key in self._weight_keys():
            if key not in weights:
                raise KeyError(f"{self.name}: missing weight '{key}'")
            a = weights[key]
            t = a.detach() if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(
                np.asarray(a, dtype=np.float32)))
            t = t.to(device=self.device, dtype=torch.float32).contiguous()
            want = self._weight_shape(key)
            if tuple(t.shape) != want:
                raise ValueError(f"{self.name}: weight '{key}' has shape {tuple(t.shape)}, expected {want}")
            self.weights[key] = t
        self._struct = None

    def _weight_keys(self):
        keys = []
        for nm in self._layers():
            keys += [nm + ".kernel", nm + ".bias"]
        return keys + ["coeff_scale", "coeff_transformation"]

    def _layers(self):
        return DENSE_LAYERS

    def _weight_shape(self, key):
        H, D, K = int(self.num_hidden), int(self.x_dim), int(self.num_features)
        kern = {"x_layer": (K, H), "v_layer": (K, H), "t_layer": (2, H), "h_layer": (H, H),
                "scale_layer": (H, D), "translation_layer": (H, D), "transformation_layer": (H, D)}
        if key in ("coeff_scale", "coeff_transformation"):
            return (1, D)
        name, kind = key.rsplit(".", 1)
        if name in kern:
            return kern[name] if kind == "kernel" else (kern[name][1],)
        return self._conv_shape(name, kind)

    @property
    def variables(self):
        return [self.weights[k] for k in self._weight_keys()]

    trainable_variables = variables

    def state_dict(self):
        return {k: v.detach().cpu().numpy() for k, v in self.weights.items()}

    # ------------------------------------------------------------------ C-ABI plumbing
    def _weights_struct(self):
        if self._struct is None:
            w, s = self.weights, _lib.NetWeights()
            s.K, s.H, s.D = int(self.num_features), int(self.num_hidden), int(self.x_dim)
            p = lambda k: _lib.dptr(w[k])
            s.w_in0, s.b_in0 = p("v_layer.kernel"), p("v_layer.bias")      # first input -> v_layer
            s.w_in1, s.b_in1 = p("x_layer.kernel"), p("x_layer.bias")
            s.w_t, s.b_t = p("t_layer.kernel"), p("t_layer.bias")
            s.w_h, s.b_h = p("h_layer.kernel"), p("h_layer.bias")
            s.w_s, s.b_s = p("scale_layer.kernel"), p("scale_layer.bias")
            s.w_tr, s.b_tr = p("translation_layer.kernel"), p("translation_layer.bias")
            s.w_q, s.b_q = p("transformation_layer.kernel"), p("transformation_layer.bias")
            s.coeff_s, s.coeff_q = p("coeff_scale"), p("coeff_transformation")
            s.tanh_on_q = int(self.tanh_on_q)
            s.packed = self._packed_ptr()
            self._struct = s
        return self._struct

    def _packed_ptr(self):
        """Pre-tiled TF32-split weight copy for the tcgen05 path (packed lazily, re-packed by
        load_weights); only built when the trunk is wide enough to ever take that path."""
        K, H, D = int(self.num_features), int(self.num_hidden), int(self.x_dim)
        if not self.use_tensor_cores or K < 64 or H < 128 or D < 64 or D % 16:
            self._packed = None
            return None
        lib = _lib.get_lib()
        nbytes = int(lib.l2hmc_net_pack_bytes(K, H, D))
        self._packed = torch.empty((nbytes + 3) // 4, dtype=torch.float32, device=self.device)
        s = _lib.NetWeights()
        s.K, s.H, s.D = K, H, D
        p = lambda k: _lib.dptr(self.weights[k])
        s.w_in0, s.w_in1, s.w_h = p("v_layer.kernel"), p("x_layer.kernel"), p("h_layer.kernel")
        s.w_s, s.w_tr, s.w_q = p("scale_layer.kernel"), p("translation_layer.kernel"), p("transformation_layer.kernel")
        with torch.cuda.device(self.device):
            _lib.check(lib.l2hmc_net_pack(s, _lib.dptr(self._packed), nbytes, _lib.stream_ptr()))
        return _lib.dptr(self._packed)

    def _workspace(self, N):
        need = int(_lib.get_lib().l2hmc_net_workspace_bytes(int(self.num_features), int(self.num_hidden),
                                                            int(self.x_dim), int(N)))
        if self._ws is None or self._ws.numel() * 4 < need:
            self._ws = torch.empty((need + 3) // 4, dtype=torch.float32, device=self.device)
        return self._ws, need

    @staticmethod
    def _tau(t):
        """(cos, sin) of the time embedding: the reference tiles ONE row over the batch
        (gauge_dynamics.py:625-633), so a (N, 2) tensor is read at row 0."""
        if isinstance(t, (tuple, list)) and len(t) == 2 and not isinstance(t[0], (torch.Tensor, np.ndarray)):
            return float(t[0]), float(t[1])
        a = t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)
        a = a.reshape(-1, 2)
        return float(a[0, 0]), float(a[0, 1])

    def _dev(self, a):
        from .lattice import to_device
        t = to_device(a, self.device)
        return t.reshape(t.shape[0], -1)

    def _run(self, call, feats0, feats1, N, t, alt=None):
        ws, need = self._workspace(N)
        call.in0, call.in1 = _lib.dptr(feats0), _lib.dptr(feats1)
        call.tau_cos, call.tau_sin = self._tau(t)
        if alt is not None:
            call.row_set = _lib.dptr(alt.row_set, torch.uint8)
            call.tau_cos_alt, call.tau_sin_alt = self._tau(alt.t)
            call.dir_alt = int(alt.direction)
        call.N = int(N)
        call.workspace, call.workspace_bytes = _lib.dptr(ws), need
        call.path = int(self.path)
        with torch.cuda.device(self.device):
            _lib.check(_lib.get_lib().l2hmc_net_run(self._weights_struct(), call, _lib.stream_ptr()))

    # ------------------------------------------------------------------ the reference's call
    def call(self, inputs, mask=None, mask_invert=False):
        """(scale, translation, transformation) = net([in0, in1, t]).  `mask` (a (D,) 0/1 row)
        multiplies the SECOND input on load -- the `mask * position` argument of position_fn
        (gauge_dynamics.py:515-517) without materialising it."""
        in0, in1, t = inputs
        in0, in1 = self._dev(in0), self._dev(in1)
        N, D = in0.shape[0], int(self.x_dim)
        if mask is not None:
            mask = self._dev(mask).reshape(-1)
        f0, f1, fmask, _ = self._features(in0, in1, mask, mask_invert)
        S, Tr, Q = (torch.empty(N, D, dtype=torch.float32, device=self.device) for _ in range(3))
        c = _lib.NetCall()
        c.mode = _lib.NET_STQ
        c.in1_mask, c.in1_mask_invert = _lib.dptr(fmask), int(bool(mask_invert))
        c.S, c.T, c.Q = _lib.dptr(S), _lib.dptr(Tr), _lib.dptr(Q)
        self._run(c, f0, f1, N, t)
        return S, Tr, Q

    __call__ = call

    # ------------------------------------------------------------------ fused sub-updates
    def momentum_update(self, x, v, g, t, eps, direction, logdet, flag, out=None, alt=None):
        """v' and log-det of _update_momentum_{forward,backward} (gauge_dynamics.py:486-508,
        537-561): S,T,Q = momentum_fn([x, g, t]) consumed in the head GEMM's epilogue.
        `alt` (an AltSet) gives the chains flagged in alt.row_set their own time row / direction."""
        N = x.shape[0]
        f0, f1, _, _ = self._features(x, g, None, False)
        out = torch.empty_like(v) if out is None else out
        c = _lib.NetCall()
        c.mode, c.eps, c.dir = _lib.NET_MOMENTUM, float(eps), int(direction)
        c.v, c.g, c.out = _lib.dptr(v), _lib.dptr(g), _lib.dptr(out)
        c.logdet, c.nonfinite_flag = _lib.dptr(logdet), _lib.dptr(flag, torch.int32)
        self._run(c, f0, f1, N, t, alt)
        return out

    def position_update(self, x, v, t, mask_row, invert, eps, direction, logdet, flag, out=None, alt=None):
        """x' and log-det of _update_position_{forward,backward} (gauge_dynamics.py:511-534,
        565-590): S,T,Q = position_fn([v, m*x, t]); the update keeps m*x and moves mbar*x."""
        N = x.shape[0]
        f0, f1, fmask, fmask_alt = self._features(v, x, mask_row, invert, alt)
        out = torch.empty_like(x) if out is None else out
        c = _lib.NetCall()
        c.mode, c.eps, c.dir = _lib.NET_POSITION, float(eps), int(direction)
        c.in1_mask, c.in1_mask_invert = _lib.dptr(fmask), int(bool(invert))
        c.x, c.v, c.mask, c.mask_invert, c.out = _lib.dptr(x), _lib.dptr(v), _lib.dptr(mask_row), int(bool(invert)), \
            _lib.dptr(out)
        if alt is not None:
            c.in1_mask_alt, c.in1_mask_invert_alt = _lib.dptr(fmask_alt), int(bool(alt.invert))
            c.mask_alt, c.mask_invert_alt = _lib.dptr(alt.mask_row), int(bool(alt.invert))
        c.logdet, c.nonfinite_flag = _lib.dptr(logdet), _lib.dptr(flag, torch.int32)
        self._run(c, f0, f1, N, t, alt)
        return out


class AltSet(object):
    """Second parameter set of a mixed-direction batch (l2hmc_net_call.row_set and *_alt): chains with
    row_set[n] != 0 use time row `t`, `direction` and (for position updates) `mask_row` / `invert`."""

    def __init__(self, row_set, t, direction, mask_row=None, invert=False):
        self.row_set, self.t, self.direction = row_set, t, int(direction)
        self.mask_row, self.invert = mask_row, bool(invert)


class GenericNet(_StqNet):
    """Fully-connected S/T/Q net (generic_net.py:20-146).

    kwargs as the reference passes them (gauge_dynamics.py:169-187): x_dim, links_shape,
    num_hidden, name_scope, factor.  Extra: device, seed (initialiser stream)."""

    def __init__(self, model_name='GenericNet', **kwargs):
        self.name = model_name
        self.name_scope = None
        self.factor = 1.
        self.seed = 0
        self.device = 'cuda'
        for key, val in kwargs.items():
            setattr(self, key, val)
        if self.name_scope is None:
            self.name_scope = model_name
        _lib.require_cuda()
        self.device = torch.device(self.device)
        self.num_features = int(self.x_dim)
        self.weights, self._struct, self._ws = {}, None, None
        rng = np.random.default_rng([int(self.seed), sum(map(ord, model_name))])
        self.load_weights(self._init_dense(self.num_features, rng))

    def _features(self, in0, in1, mask, invert, alt=None):
        m = None if mask is None else mask.reshape(-1).contiguous()
        ma = None if alt is None or alt.mask_row is None else alt.mask_row.reshape(-1).contiguous()
        return in0, in1, m, ma


class ConvNet3D(_StqNet):
    """Conv3D front-end + the same dense trunk and heads (conv_net.py:57-311).

    kwargs as the reference passes them (gauge_dynamics.py:121-143): _input_shape, links_shape,
    x_dim, factor, spatial_size, num_hidden, num_filters, filter_sizes, name_scope, data_format.
    Only data_format='channels_last' with filter_sizes [(3,3,2), (2,2,2)] -- the one
    configuration GaugeDynamics builds -- has kernels."""

    def __init__(self, model_name='ConvNet3D', **kwargs):
        self.name = model_name
        self.name_scope = None
        self.factor = 1.
        self.seed = 0
        self.device = 'cuda'
        self.data_format = 'channels_last'
        self.filter_sizes = [(3, 3, 2), (2, 2, 2)]
        for key, val in kwargs.items():
            setattr(self, key, val)
        if self.name_scope is None:
            self.name_scope = model_name
        if self.data_format != 'channels_last':
            raise NotImplementedError("ConvNet3D: only data_format='channels_last' (the reference default, "
                                      "gauge_model.py:603) is implemented")
        if [tuple(f) for f in self.filter_sizes] != [(3, 3, 2), (2, 2, 2)]:
            raise NotImplementedError("ConvNet3D: filter_sizes must be [(3,3,2), (2,2,2)] (gauge_dynamics.py:131)")
        _lib.require_cuda()
        self.device = torch.device(self.device)
        self.T, self.X = int(self.links_shape[0]), int(self.links_shape[1])
        if int(self.links_shape[2]) != 2 or 2 * self.T * self.X != int(self.x_dim):
            raise ValueError("ConvNet3D: links_shape must be (T, X, 2) with x_dim = 2*T*X")
        self.num_filters = int(self.num_filters)
        half = lambda n: _same_out(n, 2)
        self.num_features = half(half(self.T)) * half(half(self.X)) * 2 * self.num_filters
        self.weights, self._struct, self._ws, self._conv_structs = {}, None, None, None
        self._feat = [None, None]
        rng = np.random.default_rng([int(self.seed), sum(map(ord, model_name))])
        w = self._init_dense(self.num_features, rng)
        nf = self.num_filters
        for nm, shape in (("conv_x1", (3, 3, 2, 1, nf)), ("conv_v1", (3, 3, 2, 1, nf)),
                          ("conv_x2", (2, 2, 2, nf, 2 * nf)), ("conv_v2", (2, 2, 2, nf, 2 * nf))):
            rec = int(np.prod(shape[:3]))                       # Keras default glorot_uniform
            lim = math.sqrt(6.0 / (shape[3] * rec + shape[4] * rec))
            w[nm + ".kernel"] = rng.uniform(-lim, lim, size=shape).astype(np.float32)
            w[nm + ".bias"] = np.zeros((shape[4],), np.float32)
        self.load_weights(w)

    def _layers(self):
        return CONV_LAYERS + DENSE_LAYERS

    def _conv_shape(self, name, kind):
        nf = self.num_filters
        shapes = {"conv_x1": (3, 3, 2, 1, nf), "conv_v1": (3, 3, 2, 1, nf),
                  "conv_x2": (2, 2, 2, nf, 2 * nf), "conv_v2": (2, 2, 2, nf, 2 * nf)}
        return shapes[name] if kind == "kernel" else (shapes[name][4],)

    def load_weights(self, weights):
        super().load_weights(weights)
        self._conv_structs = None

    def _conv(self, which):
        if self._conv_structs is None:
            self._conv_structs = {}
            for s in ("x", "v"):
                c = _lib.Conv3dWeights()
                c.w1, c.b1 = _lib.dptr(self.weights[f"conv_{s}1.kernel"]), _lib.dptr(self.weights[f"conv_{s}1.bias"])
                c.w2, c.b2 = _lib.dptr(self.weights[f"conv_{s}2.kernel"]), _lib.dptr(self.weights[f"conv_{s}2.bias"])
                c.F = self.num_filters
                self._conv_structs[s] = c
        return self._conv_structs[which]

    def _features(self, in0, in1, mask, invert, alt=None):
        """Conv features of both streams: the first input through conv_v*, the second (masked on
        load) through conv_x* (conv_net.py:252-260)."""
        N = in0.shape[0]
        lib = _lib.get_lib()
        feats = []
        for i, (a, which, m) in enumerate(((in0, "v", None), (in1, "x", mask))):
            if self._feat[i] is None or self._feat[i].shape[0] < N:
                self._feat[i] = torch.empty(N, self.num_features, dtype=torch.float32, device=self.device)
            f = self._feat[i][:N]
            mm = None if m is None else m.reshape(-1).contiguous()
            with torch.cuda.device(self.device):
                if m is not None and alt is not None:
                    ma = alt.mask_row.reshape(-1).contiguous()
                    _lib.check(lib.l2hmc_conv3d_features_sets(
                        self._conv(which), _lib.dptr(a), _lib.dptr(mm), int(bool(invert)), _lib.dptr(ma),
                        int(bool(alt.invert)), _lib.dptr(alt.row_set, torch.uint8), N, self.T, self.X, _lib.dptr(f),
                        _lib.stream_ptr()))
                else:
                    _lib.check(lib.l2hmc_conv3d_features(self._conv(which), _lib.dptr(a), _lib.dptr(mm),
                                                         int(bool(invert)), N, self.T, self.X, _lib.dptr(f),
                                                         _lib.stream_ptr()))
            feats.append(f)
        return feats[0], feats[1], None, None
