"""L2HMC path with S/T/Q networks on the GPU (K3 + fused K2 epilogues) against the reference's
golden vectors (tests/golden/l2hmc_cases.npz, produced by the unmodified reference) and the oracle.

Tolerances: north_star asks 1e-5 relative in fp32 for links / log-dets; the `atol` terms are a few
fp32 ulps of the value range (links ~ 2 pi, momenta ~ 3, log-dets ~ 1)."""
import numpy as np
import pytest
import torch

from oracle import u1_oracle as orc

pytestmark = pytest.mark.gpu


def close(a, b, rtol=1e-5, atol=2e-5):
    a = a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
    b = b.detach().cpu().numpy() if isinstance(b, torch.Tensor) else np.asarray(b)
    np.testing.assert_allclose(a.astype(np.float64), b.astype(np.float64), rtol=rtol, atol=atol)


@pytest.fixture(scope="module")
def pkg():
    import l2hmc_b200
    return l2hmc_b200


def golden_nets(g, tag):
    if tag + ".weight_seed" in g:     # weights regenerated from a seed, sha256-checked (config-3 case)
        from test_oracle_golden import regenerated_nets
        return regenerated_nets(g, tag, tag.split(".")[0])
    nets = {}
    for net in ("XNet", "VNet"):
        pre = f"{tag}.{net}."
        nets[net] = {k[len(pre):]: v for k, v in g.items() if k.startswith(pre) and ".call." not in k}
    return nets


def make_dyn(pkg, g, tag, arch, **kw):
    T, X, N, M = (int(v) for v in g[tag + ".meta"])
    lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=N)
    np.random.seed(42)
    dyn = pkg.GaugeDynamics(lat, lat.get_energy_function(), hmc=False, network_arch=arch, num_steps=M,
                            eps=float(g[tag + ".eps"]), weights=golden_nets(g, tag), **kw)
    assert np.array_equal(dyn.mask.cpu().numpy(), g[tag + ".mask"])
    return dyn, (T, X, N, M)


CASES = [("generic.4x4", "generic"), ("conv3D.4x4", "conv3D"), ("conv3D.8x8", "conv3D"),
         ("conv3D.16x16", "conv3D")]     # the last one is BASELINE config 3's shape (M = 5, eps = 0.25, F = 16, H = 1024)


def _within_noise(name, got, ref32, ref64):
    """Whole-transition rule (tests/parity_util.py): `got` must be within max(1e-5, 2 x noise) of the float64 oracle,
    relative to the array's scale, where noise is the distance of the REFERENCE's own fp32 result (the golden) from
    float64 -- M leapfrog steps amplify fp32 rounding, the reference's included; and within 3 x that of the golden."""
    from parity_util import K_NOISE, REL_TOL, rel_err
    e, noise = rel_err(got, ref64), rel_err(ref32, ref64)
    assert e <= max(REL_TOL, K_NOISE * noise), f"{name}: {e:.3e} vs float64 > max(1e-5, 2 x {noise:.3e})"
    d = rel_err(got, ref32)
    assert d <= max(REL_TOL, (K_NOISE + 1) * noise), f"{name}: {d:.3e} from the reference's fp32 result, noise {noise:.3e}"


@pytest.mark.parametrize("tag,arch", CASES)
@pytest.mark.parametrize("fuse", [True, False])
def test_l2hmc_against_reference_goldens(pkg, golden, tag, arch, fuse):
    g = golden("l2hmc_config3_case" if tag == "conv3D.16x16" else "l2hmc_cases")
    dyn, (T, X, N, M) = make_dyn(pkg, g, tag, arch, fuse_updates=fuse)
    beta = float(g[tag + ".beta"])
    x, v = g[tag + ".x"], g[tag + ".v_f"]
    # bare network calls (generic_net.py:129-146 / conv_net.py:247-280)
    for name, net in (("XNet", dyn.position_fn), ("VNet", dyn.momentum_fn)):
        S, Tr, Q = net([v, x, g[f"{tag}.{name}.call.tvec"]])
        close(S, g[f"{tag}.{name}.call.S"], atol=2e-6); close(Tr, g[f"{tag}.{name}.call.T"], atol=2e-6)
        close(Q, g[f"{tag}.{name}.call.Q"], atol=2e-6)
    # the four sub-updates at step 1 (gauge_dynamics.py:486-590)
    tv = dyn._format_time(1, tile=N)
    m, mi = dyn._get_mask_while(1)
    for nm, fn in (("vf", dyn._update_momentum_forward), ("vb", dyn._update_momentum_backward)):
        vv, ld = fn(x, v, beta, tv)
        close(vv, g[f"{tag}.sub.{nm}.v"]); close(ld, g[f"{tag}.sub.{nm}.logdet"])
    for nm, fn in (("xf", dyn._update_position_forward), ("xb", dyn._update_position_backward)):
        xx, ld = fn(x, v, tv, m, mi)
        close(xx, g[f"{tag}.sub.{nm}.x"]); close(ld, g[f"{tag}.sub.{nm}.logdet"])
    # one leapfrog step each way (gauge_dynamics.py:412-483)
    xx, vv, ld = dyn._forward_lf(x, v, beta, 1)
    close(xx, g[tag + ".lf_f.x"], atol=5e-5); close(vv, g[tag + ".lf_f.v"], atol=5e-5)
    close(ld, g[tag + ".lf_f.logdet"], atol=5e-5)
    xx, vv, ld = dyn._backward_lf(x, v, beta, 1)
    close(xx, g[tag + ".lf_b.x"], atol=5e-5); close(vv, g[tag + ".lf_b.v"], atol=5e-5)
    close(ld, g[tag + ".lf_b.logdet"], atol=5e-5)
    # whole transitions and the MC step: tolerance from the reference's own fp32 distance to the float64 oracle
    od64 = orc.OracleDynamics(T, X, M, float(g[tag + ".eps"]), hmc=False, network_arch=arch, mask=g[tag + ".mask"],
                              xnet=golden_nets(g, tag)["XNet"], vnet=golden_nets(g, tag)["VNet"], dtype=torch.float64)
    for fwd, vk, sfx in ((True, ".v_f", "f"), (False, ".v_b", "b")):
        xs, vs, ps = dyn.transition_kernel(x, beta, forward=fwd, momentum=g[tag + vk])
        x64, v64, p64 = od64.transition_kernel(x, beta, g[tag + vk], forward=fwd)
        _within_noise(f"{tag} x_{sfx}", xs, g[f"{tag}.x_{sfx}"], x64)
        _within_noise(f"{tag} v_{sfx}", vs, g[f"{tag}.v_{sfx}_out"], v64)
        # accept probability p = exp(min(H_old - H_new + sum log-det, 0)): compared in the exponent, where the error
        # is additive.  H = beta S + v.v/2 is an fp32 sum of O(1e2..1e3): the north star's 1e-5 relative on actions
        # would allow 1e-5 |H| here; held to 1e-6 |H| (or twice the reference's own fp32 error, if larger)
        lp, lp32, lp64 = (np.log(np.maximum(np.asarray(a, np.float64), 1e-300)) for a in
                          (ps.cpu().numpy(), g[f"{tag}.p_{sfx}"], p64.numpy()))
        b64 = torch.tensor(beta, dtype=torch.float64)
        hscale = max(float(orc.hamiltonian(torch.as_tensor(x).double(), torch.as_tensor(g[tag + vk]).double(), b64, T, X).abs().max()),
                     float(orc.hamiltonian(x64, v64, b64, T, X).abs().max()))
        live = lp64 > -80                                   # below that p underflows towards fp32 denormals
        noise = np.abs(lp32 - lp64)[live].max()
        assert np.abs(lp - lp64)[live].max() <= max(1e-6 * hscale, 2 * noise), (np.abs(lp - lp64).max(), noise, hscale)
    xp, vp, p, xo = dyn.apply_transition(x, beta, momentum_f=g[tag + ".v_f"], momentum_b=g[tag + ".v_b"],
                                         direction_uniform=g[tag + ".u_dir"], accept_uniform=g[tag + ".u_acc"])
    xp64, vp64, pp64, xo64 = od64.apply_transition(x, beta, g[tag + ".v_f"], g[tag + ".v_b"], g[tag + ".u_dir"],
                                                   g[tag + ".u_acc"])
    _within_noise(f"{tag} x_post", xp, g[tag + ".x_post"], xp64)
    _within_noise(f"{tag} v_post", vp, g[tag + ".v_post"], vp64)
    acc_ref = g[tag + ".accept_prob"] > g[tag + ".u_acc"]
    assert np.array_equal(dyn.last_accepted.cpu().numpy().astype(bool), acc_ref)     # accept decisions bit-exact
    assert float(np.abs(pp64.numpy() - g[tag + ".u_acc"]).min()) > 1e-4              # ... with margin
    _within_noise(f"{tag} x_out", xo, g[tag + ".x_out"], xo64)
    if tag + ".obs_charge" in g:     # observables of the state the step ends in (lattice.py:285-313)
        a_, p_, q_ = dyn.lattice.calc_plaq_observables(xo)
        assert np.array_equal(q_.cpu().numpy(), g[tag + ".obs_charge"])               # integer charge bit-exact
        a64, p64_, _ = orc.plaq_observables(xo64, T, X)
        _within_noise(f"{tag} action", a_, g[tag + ".obs_action"], a64)
        _within_noise(f"{tag} plaquette", p_, g[tag + ".obs_avg_plaq"], p64_)
    dyn.raise_if_nonfinite()


def _oracle_weights(arch, T, X, rng, gain=None):
    from parity_util import trained_like_weights
    return trained_like_weights(arch, T, X, rng, gain)


@pytest.mark.parametrize("arch,T,X,N,M", [("generic", 8, 8, 128, 5), ("generic", 6, 10, 37, 2), ("generic", 3, 5, 70, 2),
                                          ("generic", 16, 16, 200, 2), ("conv3D", 16, 16, 33, 2),
                                          ("conv3D", 6, 10, 9, 2), ("conv3D", 8, 4, 130, 3)])
def test_l2hmc_transition_against_oracle(pkg, arch, T, X, N, M):
    """Config-2 size (8x8 generic, N=128, M=5) and ragged shapes: fused and unfused paths vs the oracle."""
    rng = np.random.default_rng(100 + T * X + N)
    D, eps, beta = 2 * T * X, 0.2, 2.5
    ws = _oracle_weights(arch, T, X, rng)
    x = rng.uniform(0, 2 * np.pi, (N, D)).astype(np.float32)
    vf, vb = (rng.standard_normal((N, D)).astype(np.float32) for _ in range(2))
    ud, ua = (rng.uniform(size=N).astype(np.float32) for _ in range(2))
    lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=N)
    np.random.seed(42)
    mask = orc.construct_masks(D, M, seed=42)
    od = orc.OracleDynamics(T, X, M, eps, hmc=False, network_arch=arch, mask=mask, xnet=ws["XNet"], vnet=ws["VNet"])
    od64 = orc.OracleDynamics(T, X, M, eps, hmc=False, network_arch=arch, mask=mask, xnet=ws["XNet"],
                              vnet=ws["VNet"], dtype=torch.float64)
    rxp, rvp, rp, rxo = od.apply_transition(x, beta, vf, vb, ud, ua)
    dxp, dvp, dp, _ = od64.apply_transition(x, beta, vf, vb, ud, ua)
    # the fp32 oracle's own distance from fp64 sets the scale of a fair tolerance
    noise = max(float((rxp.double() - dxp).abs().max()), 1e-6)
    outs = []
    for fuse in (True, False):
        np.random.seed(42)
        dyn = pkg.GaugeDynamics(lat, hmc=False, network_arch=arch, num_steps=M, eps=eps, weights=ws, fuse_updates=fuse)
        xp, vp, p, xo = dyn.apply_transition(x, beta, momentum_f=vf, momentum_b=vb, direction_uniform=ud,
                                             accept_uniform=ua)
        assert float((xp.cpu().double() - dxp).abs().max()) < 4 * noise + 2e-5
        assert float((vp.cpu().double() - dvp).abs().max()) < 4 * noise + 2e-5
        close(xp, rxp, rtol=1e-5, atol=4 * noise + 2e-5); close(vp, rvp, rtol=1e-5, atol=4 * noise + 2e-5)
        close(p, rp, rtol=2e-3, atol=1e-6)
        safe = np.abs(dp.numpy() - ua) > 1e-3
        assert np.array_equal(dyn.last_accepted.cpu().numpy().astype(bool)[safe], (dp.numpy() > ua)[safe])
        dyn.raise_if_nonfinite()
        outs.append((xp, vp, p))
    close(outs[0][0], outs[1][0], rtol=0, atol=1e-5)        # fused epilogue == separate S,T,Q + K2 kernels
    close(outs[0][1], outs[1][1], rtol=0, atol=1e-5)
    assert float(np.abs(rp.numpy() - 1).max()) > 1e-3        # the nets actually moved the accept probability


def test_net_reversibility_on_device(pkg):
    """_backward_lf(step = M-1-k) undoes _forward_lf(step = k), log-dets opposite (SURVEY 8c iii)."""
    rng = np.random.default_rng(8)
    T = X = 8
    N, M, D = 64, 3, 128
    ws = _oracle_weights("generic", T, X, rng)
    lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=N)
    dyn = pkg.GaugeDynamics(lat, hmc=False, network_arch="generic", num_steps=M, eps=0.2, weights=ws)
    x = rng.uniform(0, 2 * np.pi, (N, D)).astype(np.float32)
    v = rng.standard_normal((N, D)).astype(np.float32)
    for k in range(M):
        x1, v1, ld = dyn._forward_lf(x, v, 2.0, k)
        x2, v2, ldb = dyn._backward_lf(x1, v1, 2.0, M - 1 - k)
        close(x2, x, rtol=0, atol=3e-5); close(v2, v, rtol=0, atol=3e-5)
        close(ldb, -ld, rtol=1e-4, atol=2e-5)
        assert float(ld.abs().max()) > 1e-3


def test_net_errors(pkg):
    from l2hmc_b200 import _lib
    from l2hmc_b200.network import GenericNet
    net = GenericNet(model_name="XNet", x_dim=32, links_shape=(4, 4, 2), num_hidden=64, name_scope="position",
                     factor=2.)
    with pytest.raises(KeyError):
        net.load_weights({"x_layer.kernel": np.zeros((32, 64), np.float32)})
    w = net.state_dict()
    w["h_layer.kernel"] = np.zeros((3, 3), np.float32)
    with pytest.raises(ValueError):
        net.load_weights(w)
    lib = _lib.get_lib()
    assert lib.l2hmc_net_run(None, None, None) == _lib.L2HMC_ERR_INVALID
    # zero-initialised heads => S = T = Q = 0 exactly like the reference's freshly built nets are ~1e-3
    S, Tr, Q = net([np.zeros((5, 32), np.float32), np.zeros((5, 32), np.float32), np.array([[1.0, 0.0]], np.float32)])
    assert S.shape == (5, 32) and float(S.abs().max()) < 0.05
    # check_numerics mirror: huge scale head -> exp overflow -> flag
    w = net.state_dict()
    w["scale_layer.bias"] = np.full((32,), 50.0, np.float32)
    w["coeff_scale"] = np.full((1, 32), 8.0, np.float32)
    net.load_weights(w)
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    ld = torch.zeros(5, device="cuda")
    z = torch.zeros(5, 32, device="cuda")
    net.momentum_update(z, z + 1, z, (1.0, 0.0), 1.0, 1, ld, flag)
    assert int(flag.item()) == 1


# ------------------------------------------------------------------ tcgen05 path
@pytest.mark.parametrize("M,Nc,K,bn", [(128, 256, 32, 256), (128, 128, 64, 128), (256, 512, 512, 256),
                                       (300, 200, 100, 128), (1000, 1024, 1024, 256), (130, 70, 40, 256)])
def test_tc_selftest_gemm(pkg, M, Nc, K, bn):
    """relu(A.W) through the tcgen05 3xTF32 layer kernel + TS32 converters vs float64."""
    from l2hmc_b200 import _lib
    lib = _lib.get_lib()
    rng = np.random.default_rng(M + Nc + K)
    A = rng.standard_normal((M, K)).astype(np.float32)
    W = (rng.standard_normal((K, Nc)) / np.sqrt(K)).astype(np.float32)
    dA, dW = torch.as_tensor(A).cuda(), torch.as_tensor(W).cuda()
    C = torch.empty(M, Nc, device="cuda")
    scratch = torch.empty(int(lib.l2hmc_tc_selftest_bytes(M, Nc, K)) // 4 + 64, device="cuda")
    _lib.check(lib.l2hmc_tc_selftest(_lib.dptr(dA), _lib.dptr(dW), M, Nc, K, bn, _lib.dptr(C), _lib.dptr(scratch), None))
    torch.cuda.synchronize()
    ref = np.maximum(A.astype(np.float64) @ W.astype(np.float64), 0)
    err = np.abs(C.cpu().numpy() - ref).max()
    fp32 = np.abs(np.maximum(A @ W, 0) - ref).max()               # what plain fp32 accumulation achieves
    assert err < max(2e-6, 4 * fp32), f"max abs err {err}, fp32 matmul {fp32} (plain TF32 would be ~1e-3)"


@pytest.mark.parametrize("arch,T,X,N,M", [("generic", 8, 8, 256, 2), ("generic", 16, 16, 300, 2),
                                          ("conv3D", 16, 16, 512, 2), ("generic", 6, 8, 257, 1)])
def test_tc_path_equals_ffma_path_and_oracle(pkg, arch, T, X, N, M):
    from l2hmc_b200 import _lib
    rng = np.random.default_rng(7 + T * X + N)
    D, eps, beta = 2 * T * X, 0.2, 2.5
    ws = _oracle_weights(arch, T, X, rng)
    x = rng.uniform(0, 2 * np.pi, (N, D)).astype(np.float32)
    vf, vb = (rng.standard_normal((N, D)).astype(np.float32) for _ in range(2))
    ud, ua = (rng.uniform(size=N).astype(np.float32) for _ in range(2))
    lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=N)
    mask = orc.construct_masks(D, M, seed=42)
    od = orc.OracleDynamics(T, X, M, eps, hmc=False, network_arch=arch, mask=mask, xnet=ws["XNet"], vnet=ws["VNet"])
    rxp, rvp, rp, _ = od.apply_transition(x, beta, vf, vb, ud, ua)
    res = {}
    for path in (_lib.NET_PATH_TCGEN05, _lib.NET_PATH_FFMA):
        np.random.seed(42)
        dyn = pkg.GaugeDynamics(lat, hmc=False, network_arch=arch, num_steps=M, eps=eps, weights=ws)
        dyn.position_fn.path = dyn.momentum_fn.path = path
        # bare S,T,Q through each path
        S, Tr, Q = dyn.position_fn([vf, x, (1.0, 0.0)], mask=dyn.mask[0], mask_invert=True)
        xp, vp, p, xo = dyn.apply_transition(x, beta, momentum_f=vf, momentum_b=vb, direction_uniform=ud,
                                             accept_uniform=ua)
        res[path] = (S, Tr, Q, xp, vp, p)
        close(xp, rxp, rtol=1e-5, atol=5e-5); close(vp, rvp, rtol=1e-5, atol=5e-5)
        close(p, rp, rtol=2e-3, atol=1e-6)
        dyn.raise_if_nonfinite()
    a, b = res[_lib.NET_PATH_TCGEN05], res[_lib.NET_PATH_FFMA]
    for i in range(3):
        close(a[i], b[i], rtol=1e-5, atol=2e-6)          # 3xTF32 == fp32 FFMA to ~1e-6
    close(a[3], b[3], rtol=0, atol=3e-5); close(a[4], b[4], rtol=0, atol=3e-5)
    tS, tT, tQ = od.position_fn(torch.as_tensor(vf), (1 - torch.as_tensor(mask[0])) * torch.as_tensor(x),
                                torch.tensor([[1.0, 0.0]]).repeat(N, 1))
    close(a[0], tS, rtol=1e-5, atol=2e-6); close(a[1], tT, rtol=1e-5, atol=2e-6); close(a[2], tQ, rtol=1e-5, atol=2e-6)


def test_tc_path_refuses_small_shapes(pkg):
    from l2hmc_b200 import _lib
    from l2hmc_b200.network import GenericNet
    net = GenericNet(model_name="XNet", x_dim=32, links_shape=(4, 4, 2), num_hidden=64, name_scope="position",
                     factor=2.)
    net.path = _lib.NET_PATH_TCGEN05
    with pytest.raises(_lib.L2hmcError, match="no tcgen05 path"):
        net([np.zeros((5, 32), np.float32), np.zeros((5, 32), np.float32), (1.0, 0.0)])


@pytest.mark.parametrize("T,X,F,N", [(16, 16, 16, 67), (8, 8, 8, 3), (7, 9, 8, 5), (32, 32, 32, 6), (5, 6, 16, 4),
                                     (6, 10, 10, 9), (4, 4, 4, 2), (16, 16, 16, 1), (12, 8, 8, 11), (16, 32, 16, 7),
                                     (8, 8, 8, 1000)])
def test_conv3d_features_against_oracle(T, X, F, N):
    """l2hmc_conv3d_features (register-tiled kernel for F in {8,16,32}, generic otherwise) against the oracle's
    Conv3D/MaxPool3D 'same' stack (network/conv_net.py:247-260), with and without the mask on load."""
    from l2hmc_b200 import _lib
    rng = np.random.default_rng(T * 100 + X + F)
    w1 = rng.normal(0, 0.4, (3, 3, 2, 1, F)).astype(np.float32)
    w2 = rng.normal(0, 0.2, (2, 2, 2, F, 2 * F)).astype(np.float32)
    b1, b2 = rng.normal(0, 0.1, F).astype(np.float32), rng.normal(0, 0.1, 2 * F).astype(np.float32)
    x = rng.uniform(0, 2 * np.pi, (N, 2 * T * X)).astype(np.float32)
    mask = (rng.uniform(size=2 * T * X) > 0.5).astype(np.float32)
    dev = [torch.as_tensor(a).cuda().contiguous() for a in (w1, b1, w2, b2, x, mask)]
    c = _lib.Conv3dWeights()
    c.w1, c.b1, c.w2, c.b2, c.F = _lib.dptr(dev[0]), _lib.dptr(dev[1]), _lib.dptr(dev[2]), _lib.dptr(dev[3]), F
    lib = _lib.get_lib()
    nf = lib.l2hmc_conv3d_num_features(T, X, F)
    for m, inv in ((None, 0), (dev[5], 0), (dev[5], 1)):
        out = torch.full((N, nf), float("nan"), device="cuda")
        _lib.check(lib.l2hmc_conv3d_features(c, _lib.dptr(dev[4]), _lib.dptr(m), inv, N, T, X, _lib.dptr(out), None))
        xin = torch.as_tensor(x) if m is None else torch.as_tensor(x * (1 - mask if inv else mask))
        ref = xin.reshape(N, T, X, 2, 1)
        ref = orc._maxpool3d_same(orc._conv3d_same_relu(ref, torch.as_tensor(w1), torch.as_tensor(b1)))
        ref = orc._maxpool3d_same(orc._conv3d_same_relu(ref, torch.as_tensor(w2), torch.as_tensor(b2)))
        ref = ref.reshape(N, -1)
        assert ref.shape[1] == nf
        close(out, ref, rtol=1e-5, atol=3e-5)


@pytest.mark.parametrize("arch,T,N,path", [("generic", 8, 300, "ffma"), ("generic", 8, 300, "tc"), ("conv3D", 8, 70, "ffma"),
                                           ("conv3D", 16, 260, "tc")])
def test_mixed_direction_batch_equals_per_direction_runs(pkg, arch, T, N, path):
    """apply_transition runs forward and backward chains in the same launches (l2hmc_net_call.row_set);
    every chain must come out bit-identical to the one-direction transition_kernel of its own direction
    (the reference computes both for all chains and blends with 0/1 masks, gauge_dynamics.py:211-241)."""
    from l2hmc_b200 import _lib
    M, eps, beta = 3, 0.15, 2.5
    D = 2 * T * T
    rng = np.random.default_rng(T + N)
    ws = _oracle_weights(arch, T, T, rng)
    x = rng.uniform(0, 2 * np.pi, (N, D)).astype(np.float32)
    vf, vb = (rng.standard_normal((N, D)).astype(np.float32) for _ in range(2))
    ud, ua = (rng.uniform(size=N).astype(np.float32) for _ in range(2))
    lat = pkg.GaugeLattice(T, T, 2, "U1", num_samples=N)
    np.random.seed(42)
    dyn = pkg.GaugeDynamics(lat, hmc=False, network_arch=arch, num_steps=M, eps=eps, weights=ws)
    dyn.position_fn.path = dyn.momentum_fn.path = _lib.NET_PATH_TCGEN05 if path == "tc" else _lib.NET_PATH_FFMA
    xp, vp, p, xo = dyn.apply_transition(x, beta, momentum_f=vf, momentum_b=vb, direction_uniform=ud, accept_uniform=ua)
    xf, vfo, pf = dyn.transition_kernel(x, beta, forward=True, momentum=vf)
    xb, vbo, pb = dyn.transition_kernel(x, beta, forward=False, momentum=vb)
    fwd = torch.as_tensor(ud > 0.5).cuda()
    assert 0 < int(fwd.sum()) < N
    assert torch.equal(xp, torch.where(fwd[:, None], xf, xb))
    assert torch.equal(vp, torch.where(fwd[:, None], vfo, vbo))
    assert torch.equal(p, torch.where(fwd, pf, pb))
    dyn.raise_if_nonfinite()


@pytest.mark.parametrize("arch,T,N,path,inject", [("generic", 8, 150, "ffma", True), ("generic", 8, 300, "tc", True),
                                                  ("conv3D", 8, 70, "ffma", True), ("conv3D", 16, 260, "tc", False),
                                                  ("generic", 4, 9, "ffma", False)])
def test_native_mc_step_equals_python_composition(pkg, arch, T, N, path, inject):
    """l2hmc_u1_l2hmc_mc_step (one C call enqueues the whole MC step) against the same kernels driven one by one from
    Python: identical bits, injected or Philox draws, with the mod 2 pi and the observables."""
    from l2hmc_b200 import _lib
    M, eps, beta = 3, 0.15, 2.5
    D = 2 * T * T
    rng = np.random.default_rng(T * N)
    ws = _oracle_weights(arch, T, T, rng)
    x = rng.uniform(0, 2 * np.pi, (N, D)).astype(np.float32)
    kw = {}
    if inject:
        kw = dict(momentum_f=rng.standard_normal((N, D)).astype(np.float32),
                  momentum_b=rng.standard_normal((N, D)).astype(np.float32),
                  direction_uniform=rng.uniform(size=N).astype(np.float32),
                  accept_uniform=rng.uniform(size=N).astype(np.float32))
    lat = pkg.GaugeLattice(T, T, 2, "U1", num_samples=N)
    outs = []
    for native in (True, False):
        np.random.seed(42)
        dyn = pkg.GaugeDynamics(lat, hmc=False, network_arch=arch, num_steps=M, eps=eps, weights=ws, seed=5, chain_offset=11)
        dyn.native_step = native
        dyn.position_fn.path = dyn.momentum_fn.path = _lib.NET_PATH_TCGEN05 if path == "tc" else _lib.NET_PATH_FFMA
        dyn._mc_step = 4
        res = dyn.apply_transition(x, beta, wrap=True, observables=True, **kw)
        outs.append(res + (dyn.last_accepted,) + tuple(dyn.last_observables))
        assert dyn._mc_step == (4 if inject else 5)
        dyn.raise_if_nonfinite()
    for a, b in zip(*outs):
        assert torch.equal(a, b)
    assert 0 < int(outs[0][4].sum()) <= N


@pytest.mark.parametrize("arch,T,X", [("generic", 8, 8), ("generic", 6, 10), ("conv3D", 8, 8)])
def test_ffma_layers_do_not_depend_on_batch_size(arch, T, X):
    """The chunked FFMA layers (nets.cu): a chain's S, T, Q and its fused sub-updates are the same BITS whether it is
    launched alone, in a small batch (K chunks spread over a thread-block cluster) or in a large one (one CTA walks the
    chunks) -- what lets the sampler chunk a batch and the ranks shard it without changing a result."""
    import l2hmc_b200 as pkg
    from l2hmc_b200 import _lib
    from parity_util import trained_like_weights
    D = 2 * T * X
    rng = np.random.default_rng(T * X)
    big = 9700                                   # > 148 output tiles: the one-CTA-per-tile path
    x = rng.uniform(0, 2 * np.pi, (big, D)).astype(np.float32)
    v = rng.standard_normal((big, D)).astype(np.float32)
    ws = trained_like_weights(arch, T, X, np.random.default_rng(3), gain=1.0)
    outs = {}
    for n in (3, 200, big):
        lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=n)
        np.random.seed(1)
        dyn = pkg.GaugeDynamics(lat, hmc=False, network_arch=arch, num_steps=2, eps=0.1, weights=ws)
        net = dyn.position_fn
        net.path = _lib.NET_PATH_FFMA
        xs, vs = torch.as_tensor(x[:n]).cuda(), torch.as_tensor(v[:n]).cuda()
        S, Tr, Q = net([vs, xs, (0.3, 0.7)], mask=dyn.mask[0], mask_invert=True)
        ld = torch.zeros(n, device="cuda")
        xo = net.position_update(xs, vs, (0.3, 0.7), dyn.mask[0], False, 0.1, 1, ld, dyn._flag)
        outs[n] = [t.cpu() for t in (S, Tr, Q, xo, ld)]
    for n in (3, 200):
        for a, b, nm in zip(outs[n], outs[big], ("S", "T", "Q", "x_out", "logdet")):
            assert torch.equal(a, b[:n]), f"{arch} {T}x{X}: {nm} of the first {n} chains differs between a {n}- and a {big}-chain launch"
