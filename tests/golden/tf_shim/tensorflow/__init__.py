"""Minimal torch-CPU stand-in for the TensorFlow-1.x symbols the reference's
hot path touches.

TEST INFRASTRUCTURE ONLY.  TensorFlow 1.x cannot be installed in this image
(no py3.12 wheel, no network), so `tests/golden/make_golden.py` puts this
package on `sys.path` under the name `tensorflow` and then imports the
UNMODIFIED reference modules from /root/reference:

    l2hmc/lattice/lattice.py, l2hmc/dynamics/gauge_dynamics.py,
    l2hmc/network/generic_net.py, l2hmc/network/conv_net.py

so that the golden vectors under tests/golden/ are produced by the
reference's own Python control flow and op sequence; only the leaf tensor ops
(`tf.roll`, `tf.cos`, `tf.reduce_sum`, `Dense`, `Conv3D`, ...) are supplied
here, each as the single torch-CPU op with the documented TF semantics.

Differences from real TF that matter for parity, all deliberate:
  * random draws (`tf.random_normal`, `tf.random_uniform`) are popped from an
    injection queue (`inject_randoms`) instead of TF's Philox stream, so the
    same numbers can be fed to the CUDA path;
  * `tf.gradients` is `torch.autograd.grad` (the reference differentiates the
    Wilson action to get the force, gauge_dynamics.py:698-709);
  * `tf.while_loop` is a Python loop;
  * 'same' padding follows TF: total pad k-1 (stride 1), the odd cell goes at
    the END; max-pool 'same' pads with -inf.
Nothing here is imported by the product package.
"""
import contextlib
import math as _pymath
import types

import numpy as np
import torch
import torch.nn.functional as F

__version__ = "1.15.shim"

float32 = torch.float32
float64 = torch.float64
int32 = torch.int32
int64 = torch.int64

_STATE = {"normals": [], "uniforms": [], "init_rng": np.random.RandomState(1234)}


def inject_randoms(normals=(), uniforms=()):
    """Queue the arrays successive tf.random_normal / tf.random_uniform calls return."""
    _STATE["normals"] = [torch.as_tensor(np.asarray(a), dtype=torch.float32) for a in normals]
    _STATE["uniforms"] = [torch.as_tensor(np.asarray(a), dtype=torch.float32) for a in uniforms]


def set_init_seed(seed):
    _STATE["init_rng"] = np.random.RandomState(seed)


def _t(x, dtype=None):
    if isinstance(x, torch.Tensor):
        return x if dtype is None else x.to(dtype)
    if isinstance(x, (list, tuple)) and len(x) and isinstance(x[0], torch.Tensor):
        return torch.stack([_t(v, dtype) for v in x])
    return torch.as_tensor(np.asarray(x), dtype=dtype)


@contextlib.contextmanager
def name_scope(*_a, **_k):
    yield


variable_scope = name_scope


def executing_eagerly():
    return False


def set_random_seed(_seed):
    pass


def constant(value, dtype=None, name=None, shape=None):
    if dtype is None and isinstance(value, float):
        dtype = torch.float32
    return _t(value, dtype)


def convert_to_tensor(value, dtype=None, name=None):
    return _t(value, dtype)


def Variable(initial_value, name=None, dtype=None, trainable=True):
    v = _t(initial_value, dtype).clone().detach()
    v._shim_trainable = trainable
    v._shim_name = name
    return v


def zeros(shape, dtype=torch.float32, name=None):
    shape = [int(s) for s in (shape if isinstance(shape, (list, tuple, torch.Size)) else _t(shape).tolist())]
    return torch.zeros(shape, dtype=dtype)


def zeros_like(x, name=None):
    return torch.zeros_like(x)


def shape(x, name=None):
    return torch.tensor(list(x.shape), dtype=torch.int64)


def _shape_list(s):
    if isinstance(s, torch.Tensor):
        return [int(v) for v in s.tolist()]
    return [int(v) for v in s]


def reshape(x, shape, name=None):
    return torch.reshape(x, _shape_list(shape))


def cast(x, dtype, name=None):
    x = _t(x)
    if dtype in (torch.int32, torch.int64):
        return torch.trunc(x).to(dtype) if x.is_floating_point() else x.to(dtype)
    return x.to(dtype)


def roll(x, shift, axis, name=None):
    return torch.roll(x, shifts=shift, dims=axis)


def cos(x, name=None):
    return torch.cos(_t(x))


def sin(x, name=None):
    return torch.sin(_t(x))


def exp(x, name=None):
    return torch.exp(_t(x))


def floor(x, name=None):
    return torch.floor(x)


def abs(x, name=None):  # noqa: A001 - mirrors tf.abs
    return torch.abs(x)


def square(x, name=None):
    return x * x


def minimum(x, y, name=None):
    return torch.minimum(x, _t(y, x.dtype))


def multiply(x, y, name=None):
    return x * y


def less(x, y, name=None):
    return _t(x) < y


def is_finite(x, name=None):
    return torch.isfinite(x)


def where(cond, x, y, name=None):
    if cond.dim() == 1 and _t(x).dim() > 1:        # TF-1: a vector condition selects whole rows
        cond = cond.reshape((-1,) + (1,) * (_t(x).dim() - 1))
    return torch.where(cond, x, y)


def check_numerics(x, message, name=None):
    if not bool(torch.isfinite(x).all()):
        raise FloatingPointError(message)
    return x


def _axes(axis):
    if axis is None:
        return None
    return tuple(axis) if isinstance(axis, (tuple, list)) else (axis,)


def reduce_sum(x, axis=None, name=None, keepdims=False):
    ax = _axes(axis)
    return x.sum() if ax is None else x.sum(dim=ax, keepdim=keepdims)


def reduce_mean(x, axis=None, name=None, keepdims=False):
    ax = _axes(axis)
    return x.mean() if ax is None else x.mean(dim=ax, keepdim=keepdims)


def gather(params, indices, name=None):
    return params[indices.long()] if isinstance(indices, torch.Tensor) else params[indices]


def expand_dims(x, axis, name=None):
    return torch.unsqueeze(_t(x), axis)


def squeeze(x, axis=None, name=None):
    x = _t(x)
    return torch.squeeze(x) if axis is None else torch.squeeze(x, axis)


def tile(x, multiples, name=None):
    return x.repeat(*_shape_list(multiples))


def transpose(x, perm=None):
    return x.t() if perm is None else x.permute(*perm)


def conj(x):
    return torch.conj(x)


def mod(x, y):
    return torch.remainder(_t(x), _t(y))


def random_normal(shape, mean=0.0, stddev=1.0, dtype=torch.float32, name=None):
    out = _STATE["normals"].pop(0)
    assert list(out.shape) == _shape_list(shape), (out.shape, shape)
    # a graph node other ops may differentiate with respect to (the auxiliary z-chain of
    # gauge_model.py:757-758 feeds a drawn tensor into grad_potential)
    return torch.as_tensor(out).detach().clone().requires_grad_(True)


def random_uniform(shape, minval=0, maxval=None, dtype=torch.float32, name=None):
    out = _STATE["uniforms"].pop(0)
    assert list(out.shape) == _shape_list(shape), (out.shape, shape)
    return out


def gradients(ys, xs, name=None):
    single = not isinstance(xs, (list, tuple))
    xs_l = [xs] if single else list(xs)
    # create_graph: the loss gradient (gauge_model.py:823) differentiates THROUGH the force, itself a
    # tf.gradients node (gauge_dynamics.py:707); TF keeps every gradient op differentiable
    g = torch.autograd.grad(ys.sum(), xs_l, retain_graph=True, create_graph=True, allow_unused=True)
    return list(g)


def while_loop(cond, body, loop_vars, **_k):
    state = list(loop_vars)
    while bool(cond(*state)):
        state = list(body(*state))
    return state


def constant_initializer(value=0.0, dtype=torch.float32):
    def init(shape):
        return torch.full(tuple(shape), float(value), dtype=torch.float32)
    return init


def _truncated_normal(shape, std, rng):
    # tf.truncated_normal: resample anything beyond two standard deviations
    out = rng.standard_normal(size=shape)
    bad = np.abs(out) > 2.0
    while bad.any():
        out[bad] = rng.standard_normal(size=int(bad.sum()))
        bad = np.abs(out) > 2.0
    return torch.as_tensor(out * std, dtype=torch.float32)


def _variance_scaling_initializer(factor=2.0, mode="FAN_IN", uniform=False, seed=None, dtype=torch.float32):
    # tf.contrib.layers.variance_scaling_initializer, non-uniform branch:
    # truncated normal with stddev sqrt(1.3 * factor / n), n = fan_in.
    assert mode == "FAN_IN" and not uniform

    def init(shape):
        fan_in = int(np.prod(shape[:-1]))
        return _truncated_normal(tuple(shape), _pymath.sqrt(1.3 * factor / fan_in), _STATE["init_rng"])
    return init


def _glorot_uniform(shape):
    receptive = int(np.prod(shape[:-2])) if len(shape) > 2 else 1
    fan_in, fan_out = shape[-2] * receptive, shape[-1] * receptive
    lim = _pymath.sqrt(6.0 / (fan_in + fan_out))
    return torch.as_tensor(_STATE["init_rng"].uniform(-lim, lim, size=tuple(shape)), dtype=torch.float32)


class _Layer:
    def __init__(self, name=None):
        self.name = name
        self.weights = []


class _Dense(_Layer):
    def __init__(self, units, use_bias=True, kernel_initializer=None, bias_initializer=None,
                 activation=None, name=None):
        super().__init__(name)
        self.units, self.use_bias = units, use_bias
        self.kernel_initializer = kernel_initializer or _glorot_uniform
        self.bias_initializer = bias_initializer or constant_initializer(0.0)
        self.activation = activation
        self.kernel = self.bias = None

    def __call__(self, x):
        if self.kernel is None:
            self.kernel = self.kernel_initializer((int(x.shape[-1]), self.units))
            self.bias = self.bias_initializer((self.units,))
            self.weights = [self.kernel, self.bias]
        y = x @ self.kernel + self.bias
        return self.activation(y) if self.activation else y


class _Flatten(_Layer):
    def __call__(self, x):
        return x.reshape(x.shape[0], -1)


def _same_pad(size, k, stride):
    out = -(-size // stride)
    total = max((out - 1) * stride + k - size, 0)
    return total // 2, total - total // 2


class _Conv3D(_Layer):
    def __init__(self, filters, kernel_size, activation=None, padding="valid", name=None,
                 dtype=None, data_format="channels_last", **_k):
        super().__init__(name)
        assert data_format == "channels_last" and padding == "same"
        self.filters, self.kernel_size, self.activation = filters, tuple(kernel_size), activation
        self.kernel = self.bias = None

    def __call__(self, x):  # x: (N, d0, d1, d2, C)
        if self.kernel is None:
            self.kernel = _glorot_uniform(self.kernel_size + (int(x.shape[-1]), self.filters))
            self.bias = torch.zeros(self.filters)
            self.weights = [self.kernel, self.bias]
        xc = x.permute(0, 4, 1, 2, 3)
        pads = []
        for dim in (3, 2, 1):  # F.pad wants the last dim first
            lo, hi = _same_pad(int(x.shape[dim]), self.kernel_size[dim - 1], 1)
            pads += [lo, hi]
        xc = F.pad(xc, pads)
        y = F.conv3d(xc, self.kernel.permute(4, 3, 0, 1, 2), self.bias)
        y = y.permute(0, 2, 3, 4, 1)
        return self.activation(y) if self.activation else y


class _MaxPooling3D(_Layer):
    def __init__(self, pool_size=(2, 2, 2), strides=None, padding="valid", name=None, **_k):
        super().__init__(name)
        assert padding == "same"
        self.pool = tuple(pool_size)
        s = strides if strides is not None else pool_size
        self.strides = (s, s, s) if isinstance(s, int) else tuple(s)

    def __call__(self, x):
        xc = x.permute(0, 4, 1, 2, 3)
        pads = []
        for dim in (3, 2, 1):
            lo, hi = _same_pad(int(x.shape[dim]), self.pool[dim - 1], self.strides[dim - 1])
            pads += [lo, hi]
        xc = F.pad(xc, pads, value=float("-inf"))
        y = F.max_pool3d(xc, self.pool, self.strides)
        return y.permute(0, 2, 3, 4, 1)


class _Model:
    def __init__(self, name=None):
        self.name = name

    def __call__(self, *args, **kwargs):
        return self.call(*args, **kwargs)

    def _collect(self, seen):
        out = []
        for val in self.__dict__.values():
            if isinstance(val, torch.Tensor) and hasattr(val, "_shim_trainable"):
                if id(val) not in seen:
                    seen.add(id(val))
                    out.append(val)
            elif isinstance(val, _Layer):
                for w in val.weights:
                    if id(w) not in seen:
                        seen.add(id(w))
                        out.append(w)
            elif isinstance(val, _Model) and val is not self:
                out += val._collect(seen)
        return out

    @property
    def variables(self):
        return self._collect(set())

    trainable_variables = variables


keras = types.SimpleNamespace(
    Model=_Model,
    layers=types.SimpleNamespace(Dense=_Dense, Flatten=_Flatten, Conv3D=_Conv3D,
                                 MaxPooling3D=_MaxPooling3D),
)
nn = types.SimpleNamespace(relu=torch.relu, tanh=torch.tanh)
math = types.SimpleNamespace(floor=floor, cos=cos, sin=sin, exp=exp)  # noqa: A001 (tf.math)
contrib = types.SimpleNamespace(
    layers=types.SimpleNamespace(variance_scaling_initializer=_variance_scaling_initializer),
    eager=types.SimpleNamespace(),
)
logging = types.SimpleNamespace(set_verbosity=lambda *_a: None, INFO=0)


# --------------------------------------------------------------------------- MoG path additions
# Symbols used by the reference's utils/{dynamics,sampler,network,distributions}.py (config 5).
import sys as _sys


def placeholder(dtype=torch.float32, shape=None, name=None):
    """A feedable scalar: the golden script sets it in place (`.fill_`).  Starts at 1."""
    return torch.ones(tuple(shape or ()), dtype=dtype)


@contextlib.contextmanager
def variable_scope(*_a, **_k):
    yield types.SimpleNamespace(reuse_variables=lambda: None)


def get_variable(name, shape=None, initializer=None, trainable=True, dtype=None):
    if isinstance(initializer, torch.Tensor):
        return initializer.clone()
    if callable(initializer):
        return initializer(list(shape)).clone()
    return torch.zeros(tuple(shape))


def log(x, name=None):
    return torch.log(_t(x))


def sqrt(x, name=None):
    return torch.sqrt(_t(x))


def matmul(a, b, name=None):
    return torch.matmul(_t(a), _t(b))


def add(a, b, name=None):
    return _t(a) + _t(b)


def diag_part(x, name=None):
    return torch.diagonal(x)


def concat(values, axis=0, name=None):
    return torch.cat(list(values), dim=axis)


def reduce_logsumexp(x, axis=None, name=None, keepdims=False):
    return torch.logsumexp(x, dim=axis, keepdim=keepdims)


def reduce_max(x, axis=None, name=None):
    return x.max() if axis is None else x.max(dim=axis).values


def reduce_min(x, axis=None, name=None):
    return x.min() if axis is None else x.min(dim=axis).values


def stop_gradient(x, name=None):
    return x.detach()


summary = types.SimpleNamespace(scalar=lambda *_a, **_k: None, histogram=lambda *_a, **_k: None)
pi = _pymath.pi

for _name in ("tensorflow.python", "tensorflow.python.ops", "tensorflow.python.ops.variable_scope",
              "tensorflow.python.ops.array_ops", "tensorflow.python.ops.math_ops", "tensorflow.python.ops.nn_ops"):
    _sys.modules.setdefault(_name, types.ModuleType(_name))
_sys.modules["tensorflow.python"].ops = _sys.modules["tensorflow.python.ops"]
for _sub in ("variable_scope", "array_ops", "math_ops", "nn_ops"):
    setattr(_sys.modules["tensorflow.python.ops"], _sub, _sys.modules["tensorflow.python.ops." + _sub])


# ---- what `import gauge_model` touches at module level (gauge_model.py:52-66): never executed as TF
__version__ = "1.13.1-shim"
logging = types.SimpleNamespace(INFO=20, set_verbosity=lambda *_a, **_k: None, info=lambda *_a, **_k: None)


def set_random_seed(seed):
    set_init_seed(int(seed))


for _name in ("tensorflow.python.debug", "tensorflow.python.client", "tensorflow.python.client.timeline",
              "tensorflow.core", "tensorflow.core.protobuf", "tensorflow.core.protobuf.rewriter_config_pb2"):
    _sys.modules.setdefault(_name, types.ModuleType(_name))
_sys.modules["tensorflow.python"].debug = _sys.modules["tensorflow.python.debug"]
_sys.modules["tensorflow.python"].client = _sys.modules["tensorflow.python.client"]
_sys.modules["tensorflow.python.client"].timeline = _sys.modules["tensorflow.python.client.timeline"]
_sys.modules["tensorflow.core"].protobuf = _sys.modules["tensorflow.core.protobuf"]
_sys.modules["tensorflow.core.protobuf"].rewriter_config_pb2 = \
    _sys.modules["tensorflow.core.protobuf.rewriter_config_pb2"]
