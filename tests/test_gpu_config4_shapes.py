"""Parity at the network shapes of BASELINE config 4 (32x32 lattice, D = 2048 links per chain) -- the shapes bench.py
times: `generic` nets (K = D = 2048, H = 4D = 8192; dynamics/gauge_dynamics.py:169-187) and `conv3D` nets (F = 32
filters, 4096 conv features, H = 2D = 4096; :121-143), each on the FFMA path at a small batch and on the tcgen05
3xTF32 path at N = 256.

Checked against oracle/u1_oracle.py in fp32 AND float64 on the same seeded inputs: bare S/T/Q, the four sub-updates,
one leapfrog step each way (1e-5 relative to the array's scale), then whole trajectories (links, momenta, summed
log-det), the MC step, accept decisions and the observables under the rule of tests/parity_util.py: error vs float64
<= max(1e-5, 2 x the fp32 oracle's own error vs float64).  Accept decisions and integer charge: bit-exact."""
import numpy as np
import pytest
import torch

from oracle import u1_oracle as orc
from parity_util import check_case, measure_case, oracle_pair, rel_err, trained_like_weights

pytestmark = pytest.mark.gpu

T = X = 32
D = 2 * T * X
M, EPS, BETA = 2, 0.1, 4.0          # config 4: eps = 0.1, beta = 4; M = 2 keeps the fp64 oracle in seconds


@pytest.fixture(scope="module", params=["generic", "conv3D"])
def case(request):
    arch = request.param
    # near-thermalised links and heads of O(0.1) so that accept probabilities are neither 0 nor all 1
    ws = trained_like_weights(arch, T, X, np.random.default_rng(4000 + len(arch)), gain={"generic": 0.5, "conv3D": 2.5}[arch])
    mask = orc.construct_masks(D, M, seed=42)
    od32, od64 = oracle_pair(arch, T, X, M, EPS, mask, ws)
    return arch, ws, mask, od32, od64


def _inputs(N, seed):
    rng = np.random.default_rng(seed)
    x = np.mod(rng.normal(0, 0.45, (N, D)), 2 * np.pi).astype(np.float32)
    vf, vb = (rng.standard_normal((N, D)).astype(np.float32) for _ in range(2))
    ud, ua = (rng.uniform(size=N).astype(np.float32) for _ in range(2))
    return x, vf, vb, ud, ua


def _dyn(arch, ws, mask, N, path):
    import l2hmc_b200 as pkg
    from l2hmc_b200 import _lib
    lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=N)
    np.random.seed(42)
    dyn = pkg.GaugeDynamics(lat, hmc=False, network_arch=arch, num_steps=M, eps=EPS, weights=ws)
    dyn.set_masks(mask)
    dyn.position_fn.path = dyn.momentum_fn.path = {"ffma": _lib.NET_PATH_FFMA, "tc": _lib.NET_PATH_TCGEN05}[path]
    if arch == "generic":
        assert (dyn.position_fn.num_features, dyn.position_fn.num_hidden) == (2048, 8192)
    else:
        assert (dyn.position_fn.num_features, dyn.position_fn.num_hidden, dyn.position_fn.num_filters) == (4096, 4096, 32)
    return dyn


@pytest.mark.parametrize("path,N", [("ffma", 8), ("tc", 256)])
def test_config4_network_shapes_against_oracle(case, path, N):
    arch, ws, mask, od32, od64 = case
    dyn = _dyn(arch, ws, mask, N, path)
    x, vf, vb, ud, ua = _inputs(N, 17 + N)
    tx, tv = torch.as_tensor(x), torch.as_tensor(vf)
    tol = lambda ref: 1e-5 * max(float(ref.abs().max()), 1e-3)

    # ---- bare network calls (generic_net.py:129-146 / conv_net.py:247-280) at time row 1
    tvec32, tvec64 = od32._tvec(1, N), od64._tvec(1, N)
    for net, fn32, fn64 in ((dyn.position_fn, od32.position_fn, od64.position_fn),
                            (dyn.momentum_fn, od32.momentum_fn, od64.momentum_fn)):
        got = net([vf, x, dyn._format_time(1, tile=N)])
        r32, r64 = fn32(tv, tx, tvec32), fn64(tv.double(), tx.double(), tvec64)
        for g_, a_, b_ in zip(got, r32, r64):
            assert rel_err(g_, b_) <= max(1e-5, 2 * rel_err(a_, b_)), (arch, path, rel_err(g_, b_), rel_err(a_, b_))
            assert float(b_.abs().max()) > 1e-3                     # heads are alive

    # ---- the four sub-updates at step 1 (gauge_dynamics.py:486-590) and one LF step each way (:412-483)
    t1 = dyn._format_time(1, tile=N)
    m, mi = dyn._get_mask_while(1)
    b32 = torch.tensor(np.float32(BETA))
    m32, mi32 = od32._masks(1)
    for cu, orc_fn in ((dyn._update_momentum_forward, od32.update_momentum_forward),
                       (dyn._update_momentum_backward, od32.update_momentum_backward)):
        vv, ld = cu(x, vf, BETA, t1)
        rv, rld = orc_fn(tx, tv, b32, tvec32)
        assert float((vv.cpu() - rv).abs().max()) <= tol(rv) and float((ld.cpu() - rld).abs().max()) <= tol(rld)
    for cu, orc_fn in ((dyn._update_position_forward, od32.update_position_forward),
                       (dyn._update_position_backward, od32.update_position_backward)):
        xx, ld = cu(x, vf, t1, m, mi)
        rx, rld = orc_fn(tx, tv, tvec32, m32, mi32)
        assert float((xx.cpu() - rx).abs().max()) <= tol(rx) and float((ld.cpu() - rld).abs().max()) <= tol(rld)
    for cu, orc_fn in ((dyn._forward_lf, od32.forward_lf), (dyn._backward_lf, od32.backward_lf)):
        xx, vv, ld = cu(x, vf, BETA, 1)
        rx, rv, rld = orc_fn(tx, tv, b32, 1)
        assert float((xx.cpu() - rx).abs().max()) <= 2 * tol(rx)
        assert float((vv.cpu() - rv).abs().max()) <= 2 * tol(rv)
        assert float((ld.cpu() - rld).abs().max()) <= 2 * tol(rld)

    # ---- whole trajectories, MC step, accept decisions, observables: measured-noise rule
    errs = measure_case(dyn, od32, od64, x, BETA, vf, vb, ud, ua, T, X)
    check_case(f"config4 {arch} {path} N={N}", errs)
    dyn.raise_if_nonfinite()
