"""TensorFlow 'SAME' padding semantics of Conv3D / MaxPooling3D, stated a second, independent way.

The oracle (oracle/u1_oracle.py:_conv3d_same_relu/_maxpool3d_same) and the TensorFlow stand-in that generated the
goldens (tests/golden/tf_shim) both pad explicitly and call torch's conv3d / max_pool3d, with the same helper text.
Here the same semantics are written WITHOUT padding and WITHOUT torch: explicit NumPy index loops over output
positions and kernel taps, straight from TF's documented SAME rule (tensorflow/core/framework/common_shape_fns.cc,
GetWindowedOutputSizeVerbose; the `tf.nn.convolution` docs):

    out        = ceil(in / stride)
    pad_total  = max((out - 1) * stride + k - in, 0)
    pad_before = pad_total // 2          (the odd cell goes at the END)
    input index of tap j at output o:  i = o * stride + j - pad_before;  taps outside [0, in) contribute 0 to a
    convolution and are IGNORED by a max-pool (they are not zeros: an all-negative window keeps its negative maximum).

Checked: hand-computed known answers for the (2,2,2) even kernel and the odd-extent pool, then the oracle and the
stand-in against the loops on the reference's own layer shapes (network/conv_net.py:86-141: Conv3D(F,(3,3,2)) ->
MaxPool(2,2) -> Conv3D(2F,(2,2,2)) -> MaxPool(2,2), channels_last)."""
import importlib.util
import os

import numpy as np
import pytest
import torch

from oracle import u1_oracle as orc

HERE = os.path.dirname(os.path.abspath(__file__))


def _pad_before(size, k, stride):
    out = -(-size // stride)
    total = max((out - 1) * stride + k - size, 0)
    return out, total // 2


def conv3d_same_loops(x, kernel, bias):
    """x (N,d0,d1,d2,Cin) float64, kernel (k0,k1,k2,Cin,Cout): stride-1 SAME convolution (cross-correlation, as TF)."""
    N, d0, d1, d2, cin = x.shape
    k0, k1, k2, _, cout = kernel.shape
    (_, p0), (_, p1), (_, p2) = (_pad_before(d, k, 1) for d, k in ((d0, k0), (d1, k1), (d2, k2)))
    out = np.zeros((N, d0, d1, d2, cout))
    for o0 in range(d0):
        for o1 in range(d1):
            for o2 in range(d2):
                acc = np.tile(bias.astype(np.float64), (N, 1))
                for j0 in range(k0):
                    i0 = o0 + j0 - p0
                    if not 0 <= i0 < d0:
                        continue
                    for j1 in range(k1):
                        i1 = o1 + j1 - p1
                        if not 0 <= i1 < d1:
                            continue
                        for j2 in range(k2):
                            i2 = o2 + j2 - p2
                            if not 0 <= i2 < d2:
                                continue
                            acc += x[:, i0, i1, i2, :] @ kernel[j0, j1, j2]
                out[:, o0, o1, o2, :] = acc
    return out


def maxpool3d_same_loops(x, pool=2, stride=2):
    N, d0, d1, d2, c = x.shape
    (n0, p0), (n1, p1), (n2, p2) = (_pad_before(d, pool, stride) for d in (d0, d1, d2))
    out = np.full((N, n0, n1, n2, c), -np.inf)
    for o0 in range(n0):
        for o1 in range(n1):
            for o2 in range(n2):
                for j0 in range(pool):
                    i0 = o0 * stride + j0 - p0
                    if not 0 <= i0 < d0:
                        continue
                    for j1 in range(pool):
                        i1 = o1 * stride + j1 - p1
                        if not 0 <= i1 < d1:
                            continue
                        for j2 in range(pool):
                            i2 = o2 * stride + j2 - p2
                            if not 0 <= i2 < d2:
                                continue
                            out[:, o0, o1, o2, :] = np.maximum(out[:, o0, o1, o2, :], x[:, i0, i1, i2, :])
    return out


@pytest.fixture(scope="module")
def shim():
    """The TensorFlow stand-in of tests/golden, loaded under a private module name (it must not become `tensorflow`
    for the rest of the test session)."""
    path = os.path.join(HERE, "golden", "tf_shim", "tensorflow", "__init__.py")
    spec = importlib.util.spec_from_file_location("_l2hmc_tf_shim_for_tests", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_hand_computed_known_answers():
    # --- even kernel, extent 2 (the link-direction axis of every lattice: (T, X, 2)): pad_total = 1, all of it at the
    # END.  1-D view along the last axis, kernel taps (a, b) = (10, 1): out[0] = a x0 + b x1, out[1] = a x1 + b * 0.
    x = np.zeros((1, 1, 1, 2, 1)); x[0, 0, 0, :, 0] = (3., 5.)
    k = np.zeros((1, 1, 2, 1, 1)); k[0, 0, :, 0, 0] = (10., 1.)
    got = conv3d_same_loops(x, k, np.zeros(1))[0, 0, 0, :, 0]
    assert got.tolist() == [35., 50.]            # a pad at the FRONT would give [3, 35]
    # --- odd kernel 3 on extent 4: one cell each side.  taps (1, 10, 100) on x = (1,2,3,4)
    x = np.zeros((1, 4, 1, 1, 1)); x[0, :, 0, 0, 0] = (1., 2., 3., 4.)
    k = np.zeros((3, 1, 1, 1, 1)); k[:, 0, 0, 0, 0] = (1., 10., 100.)
    got = conv3d_same_loops(x, k, np.zeros(1))[0, :, 0, 0, 0]
    assert got.tolist() == [210., 321., 432., 43.]
    # --- max-pool 2/2 on an odd extent 3: out = 2, pad_total = 1 at the END; the padded cell is ignored, not zero
    x = np.zeros((1, 3, 1, 1, 1)); x[0, :, 0, 0, 0] = (-1., -7., -5.)
    got = maxpool3d_same_loops(x)[0, :, 0, 0, 0]
    assert got.tolist() == [-1., -5.]            # zero padding would give [-1, 0]
    x[0, :, 0, 0, 0] = (-7., -1., -5.)
    assert maxpool3d_same_loops(x)[0, :, 0, 0, 0].tolist() == [-1., -5.]   # front padding would give [-7, -1]
    # --- max-pool on extent 2 -> 1 (the link axis after conv-1) and 1 -> 1 (after pool-1: pad_total = 1, window = the cell)
    x = np.zeros((1, 1, 1, 2, 1)); x[0, 0, 0, :, 0] = (-3., -4.)
    assert maxpool3d_same_loops(x)[0, 0, 0, :, 0].tolist() == [-3.]
    x = np.full((1, 1, 1, 1, 1), -2.)
    assert maxpool3d_same_loops(x)[0, 0, 0, :, 0].tolist() == [-2.]


def test_hand_computed_answers_hold_for_oracle_and_shim(shim):
    """The same known answers through the two padded implementations."""
    x = torch.zeros(1, 1, 1, 2, 1); x[0, 0, 0, :, 0] = torch.tensor([3., 5.])
    k = torch.zeros(1, 1, 2, 1, 1); k[0, 0, :, 0, 0] = torch.tensor([10., 1.])
    assert orc._conv3d_same_relu(x, k, torch.zeros(1))[0, 0, 0, :, 0].tolist() == [35., 50.]
    conv = shim.keras.layers.Conv3D(1, (1, 1, 2), padding="same")
    conv.kernel, conv.bias = k, torch.zeros(1)
    assert conv(x)[0, 0, 0, :, 0].tolist() == [35., 50.]
    xp = torch.zeros(1, 3, 1, 1, 1); xp[0, :, 0, 0, 0] = torch.tensor([-7., -1., -5.])
    assert orc._maxpool3d_same(xp)[0, :, 0, 0, 0].tolist() == [-1., -5.]
    pool = shim.keras.layers.MaxPooling3D((2, 2, 2), 2, padding="same")
    assert pool(xp)[0, :, 0, 0, 0].tolist() == [-1., -5.]


@pytest.mark.parametrize("T,X,F", [(4, 4, 4), (8, 8, 8), (6, 10, 10), (5, 7, 3), (16, 16, 16)])
def test_oracle_and_shim_conv_stack_against_index_loops(shim, T, X, F):
    """The reference's conv front-end (network/conv_net.py:247-260) on its own layer shapes, even and odd extents."""
    rng = np.random.default_rng(T * 31 + X)
    N = 2
    x = rng.normal(0, 1, (N, T, X, 2, 1))
    w1, b1 = rng.normal(0, 0.4, (3, 3, 2, 1, F)), rng.normal(0, 0.3, F)
    w2, b2 = rng.normal(0, 0.2, (2, 2, 2, F, 2 * F)), rng.normal(0, 0.3, 2 * F)
    # loops, float64
    a1 = np.maximum(conv3d_same_loops(x, w1, b1), 0)
    p1 = maxpool3d_same_loops(a1)
    a2 = np.maximum(conv3d_same_loops(p1, w2, b2), 0)
    p2 = maxpool3d_same_loops(a2)
    half = lambda n: -(-n // 2)
    assert p1.shape == (N, half(T), half(X), 1, F) and p2.shape == (N, half(half(T)), half(half(X)), 1, 2 * F)
    t = lambda a: torch.as_tensor(a, dtype=torch.float64)
    # oracle
    o1 = orc._conv3d_same_relu(t(x), t(w1), t(b1))
    np.testing.assert_allclose(o1.numpy(), a1, rtol=0, atol=1e-12)
    q1 = orc._maxpool3d_same(o1)
    np.testing.assert_allclose(q1.numpy(), p1, rtol=0, atol=1e-12)
    o2 = orc._conv3d_same_relu(q1, t(w2), t(b2))
    np.testing.assert_allclose(o2.numpy(), a2, rtol=0, atol=1e-12)
    np.testing.assert_allclose(orc._maxpool3d_same(o2).numpy(), p2, rtol=0, atol=1e-12)
    # the stand-in's Keras layers with the same weights
    relu = shim.nn.relu
    c1 = shim.keras.layers.Conv3D(F, (3, 3, 2), activation=relu, padding="same")
    c1.kernel, c1.bias = t(w1), t(b1)
    c2 = shim.keras.layers.Conv3D(2 * F, (2, 2, 2), activation=relu, padding="same")
    c2.kernel, c2.bias = t(w2), t(b2)
    pool = shim.keras.layers.MaxPooling3D((2, 2, 2), 2, padding="same")
    s1 = c1(t(x))
    np.testing.assert_allclose(s1.numpy(), a1, rtol=0, atol=1e-12)
    s2 = c2(pool(s1))
    np.testing.assert_allclose(s2.numpy(), a2, rtol=0, atol=1e-12)
    np.testing.assert_allclose(pool(s2).numpy(), p2, rtol=0, atol=1e-12)
