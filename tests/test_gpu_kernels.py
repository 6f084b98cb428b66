"""CUDA kernels (through the C ABI / host mirror) against the golden vectors and the oracle:
stencil (K1), observables (K4), leapfrog updates (K2), accept/select, Philox draws."""
import numpy as np
import pytest
import torch

from oracle import u1_oracle as orc

pytestmark = pytest.mark.gpu

RTOL = 1e-5   # north_star: links, actions, log-dets, plaquettes within 1e-5 relative in fp32


def dev(a, dtype=torch.float32):
    return torch.as_tensor(np.asarray(a)).to("cuda", dtype).contiguous()


def close(a, b, rtol=RTOL, atol=1e-5):
    a = a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
    b = b.detach().cpu().numpy() if isinstance(b, torch.Tensor) else np.asarray(b)
    np.testing.assert_allclose(a.astype(np.float64), b.astype(np.float64), rtol=rtol, atol=atol)


@pytest.fixture(scope="module")
def pkg():
    import l2hmc_b200
    from l2hmc_b200 import _lib
    _lib.get_lib()
    return l2hmc_b200


@pytest.mark.parametrize("shape", ["8x8", "16x16", "6x10", "4x4", "32x32"])
def test_stencil_against_reference_goldens(pkg, golden, shape):
    g = golden("lattice_cases")
    T, X = map(int, shape.split("x"))
    x = g[shape + ".x"]
    lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=x.shape[0], rand=False)
    close(lat.total_action(x), g[shape + ".total_action"], atol=2e-5)
    a, p, q = lat.calc_plaq_observables(x.reshape(-1, T, X, 2))        # 4-D input is reshaped like the reference
    close(a, g[shape + ".obs_action"], atol=2e-5)
    close(p, g[shape + ".obs_avg_plaq"], atol=1e-6)
    assert np.array_equal(q.cpu().numpy(), g[shape + ".obs_charge"])    # integer charge bit-exact
    close(lat.grad_action(x, float(g[shape + ".beta"])), g[shape + ".force"], atol=2e-6)
    act, f = lat.action_and_grad(x, float(g[shape + ".beta"]))
    close(act, g[shape + ".total_action"], atol=2e-5)
    close(f, g[shape + ".force"], atol=2e-6)
    a2, p2, qf, qi = lat.observables(x)
    assert qi.dtype == torch.int32 and np.array_equal(qi.cpu().numpy(), g[shape + ".obs_charge"].astype(np.int32))
    close(qf, orc.top_charges(torch.as_tensor(x), T, X), atol=2e-5)
    assert float(a[0]) == 0.0 and float(p[0]) == 1.0 and int(qi[0]) == 0  # cold-start row


@pytest.mark.parametrize("T,X,N", [(1, 1, 3), (2, 2, 5), (3, 5, 7), (5, 3, 33), (64, 64, 2), (8, 8, 1000)])
def test_stencil_odd_shapes_against_oracle(pkg, T, X, N):
    rng = np.random.default_rng(T * 100 + X)
    x = rng.uniform(-2 * np.pi, 4 * np.pi, (N, 2 * T * X)).astype(np.float32)
    v = rng.standard_normal(x.shape).astype(np.float32)
    xt = torch.as_tensor(x)
    lat = pkg.GaugeLattice(T, X, 2, "U1", num_samples=N)
    close(lat.total_action(x), orc.total_action(xt, T, X), atol=1e-4)
    close(lat.grad_action(x, 3.0), orc.grad_potential(xt, torch.tensor(3.0), T, X), atol=5e-6)
    a, p, qf, qi = lat.observables(x)
    ao, po, qo = orc.plaq_observables(xt, T, X)
    close(p, po, atol=1e-6)
    assert np.array_equal(qi.cpu().numpy(), qo.numpy().astype(np.int32))
    dyn = pkg.GaugeDynamics(lat, hmc=True, num_steps=1, eps=0.1)
    close(dyn.hamiltonian(x, v, 3.0), orc.hamiltonian(xt, torch.as_tensor(v), torch.tensor(3.0), T, X), atol=2e-4)
    close(dyn.kinetic_energy(v), orc.kinetic_energy(torch.as_tensor(v)), atol=1e-4)


def test_empty_batch_and_bad_arguments(pkg):
    from l2hmc_b200 import _lib
    lat = pkg.GaugeLattice(8, 8, 2, "U1", num_samples=1)
    assert lat.total_action(np.zeros((0, 128), np.float32)).shape == (0,)
    with pytest.raises(ValueError):
        lat.total_action(np.zeros((2, 100), np.float32))
    lib = _lib.get_lib()
    rc = lib.l2hmc_u1_action_force(None, 1.0, 1, 8, 8, None, None, None)
    assert rc == _lib.L2HMC_ERR_INVALID and b"null" in lib.l2hmc_last_error()
    x = dev(np.zeros((1, 2 * 7 * 9)))
    rc = lib.l2hmc_u1_hmc_step(_lib.dptr(x), _lib.dptr(x), 1.0, 0.1, None, 1, 7, 9, _lib.dptr(x), _lib.dptr(x), None)
    assert rc == _lib.L2HMC_ERR_UNSUPPORTED
    with pytest.raises(TypeError):
        _lib.dptr(torch.zeros(3))                      # CPU tensor: no silent fallback
    with pytest.raises(NotImplementedError):
        pkg.GaugeLattice(4, 4, 2, "SU2")
    with pytest.raises(AttributeError):
        pkg.GaugeDynamics(lat, hmc=False, network_arch="nope")


@pytest.mark.parametrize("D,N", [(128, 64), (60, 9), (2048, 40), (2, 5)])
@pytest.mark.parametrize("direction", [1, -1])
def test_update_kernels_against_oracle(pkg, D, N, direction):
    """K2 with explicit S,T,Q inputs: same formulas as gauge_dynamics.py:486-590."""
    from l2hmc_b200 import _lib
    lib = _lib.get_lib()
    rng = np.random.default_rng(D + N)
    x, v, g = (rng.standard_normal((N, D)).astype(np.float32) * s for s in (3.0, 1.0, 2.0))
    S, Tr, Q = (rng.standard_normal((N, D)).astype(np.float32) * 0.4 for _ in range(3))
    mask = (rng.uniform(size=D) > 0.5).astype(np.float32)
    eps = np.float32(0.3)
    tx, tv, tg, tS, tT, tQ, tm = map(torch.as_tensor, (x, v, g, S, Tr, Q, mask))
    e = torch.tensor(eps)
    # momentum
    if direction > 0:
        sc = tS * (0.5 * e)
        v_ref = tv * torch.exp(sc) - 0.5 * e * (torch.exp(tQ * e) * tg - tT)
    else:
        sc = tS * (-0.5 * e)
        v_ref = torch.exp(sc) * (tv + 0.5 * e * (torch.exp(tQ * e) * tg - tT))
    out = torch.empty(N, D, device="cuda")
    ld = torch.full((N,), 0.5, device="cuda")
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    ins = [dev(a) for a in (v, g, S, Tr, Q)]        # keep the device copies alive across the launch
    _lib.check(lib.l2hmc_update_momentum(*map(_lib.dptr, ins), float(eps), direction, N, D,
                                         _lib.dptr(out), _lib.dptr(ld), _lib.dptr(flag, torch.int32), None))
    close(out, v_ref, atol=1e-6)
    close(ld, 0.5 + sc.sum(1), atol=1e-5)                 # accumulated, not overwritten
    # position, both mask orientations
    for invert in (0, 1):
        m = 1 - tm if invert else tm
        mi = 1 - m
        if direction > 0:
            sc = tS * e
            tmp = tx * torch.exp(sc) + e * (torch.exp(tQ * e) * tv + tT)
        else:
            sc = tS * (-e)
            tmp = torch.exp(sc) * (tx - e * (torch.exp(tQ * e) * tv + tT))
        x_ref = m * tx + mi * tmp
        ld = torch.zeros(N, device="cuda")
        ins = [dev(a) for a in (x, v, S, Tr, Q, mask)]
        _lib.check(lib.l2hmc_update_position(*map(_lib.dptr, ins), invert, float(eps),
                                             direction, N, D, _lib.dptr(out), _lib.dptr(ld),
                                             _lib.dptr(flag, torch.int32), None))
        close(out, x_ref, atol=1e-6)
        close(ld, (mi * sc).sum(1), atol=1e-5)
        mm = torch.empty(N, D, device="cuda")
        _lib.check(lib.l2hmc_mask_mul(_lib.dptr(ins[0]), _lib.dptr(ins[5]), invert, N, D, _lib.dptr(mm), None))
        assert torch.equal(mm.cpu(), m * tx)
    assert int(flag.item()) == 0
    # overflow in exp() raises the check_numerics flag
    big = dev(np.full((N, D), 400.0, np.float32))
    _lib.check(lib.l2hmc_update_momentum(_lib.dptr(ins[1]), _lib.dptr(ins[1]), _lib.dptr(big), _lib.dptr(ins[3]),
                                         _lib.dptr(ins[4]), 1.0, 1, N, D, _lib.dptr(out), None,
                                         _lib.dptr(flag, torch.int32), None))
    assert int(flag.item()) == 1


def test_accept_prob_and_select(pkg):
    from l2hmc_b200 import _lib
    lib = _lib.get_lib()
    rng = np.random.default_rng(2)
    N, D = 257, 36
    h0 = rng.standard_normal(N).astype(np.float32) * 3
    h1 = rng.standard_normal(N).astype(np.float32) * 3
    sld = rng.standard_normal(N).astype(np.float32)
    h0[:3] = [np.nan, np.inf, -np.inf]
    p = torch.empty(N, device="cuda")
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    hs = [dev(a) for a in (h0, h1, sld)]
    _lib.check(lib.l2hmc_accept_prob(_lib.dptr(hs[0]), _lib.dptr(hs[1]), _lib.dptr(hs[2]), N, _lib.dptr(p),
                                     _lib.dptr(flag, torch.int32), None))
    d = torch.as_tensor(h0) - torch.as_tensor(h1) + torch.as_tensor(sld)
    ref = torch.exp(torch.minimum(d, torch.zeros(())))
    ref = torch.where(torch.isfinite(ref), ref, torch.zeros(()))
    close(p, ref, atol=1e-7)
    assert int(flag.item()) == 1            # the NaN row
    # blends: both directions given (reference formulation) == select
    x, xf, xb, vf, vb = (rng.standard_normal((N, D)).astype(np.float32) * 7 for _ in range(5))
    pf, pb, ud, ua = (rng.uniform(size=N).astype(np.float32) for _ in range(4))
    outs = [torch.empty(N, D, device="cuda") for _ in range(3)]
    pp = torch.empty(N, device="cuda")
    acc = torch.empty(N, dtype=torch.uint8, device="cuda")
    ins = [dev(a) for a in (x, xf, xb, vf, vb, pf, pb, ud, ua)]
    _lib.check(lib.l2hmc_mh_select(*map(_lib.dptr, ins), N, D, 1,
                                   _lib.dptr(outs[0]), _lib.dptr(outs[1]), _lib.dptr(pp), _lib.dptr(outs[2]),
                                   _lib.dptr(acc, torch.uint8), None))
    fm = (ud > 0.5)
    xp = np.where(fm[:, None], xf, xb)
    prob = np.where(fm, pf, pb)
    am = prob > ua
    assert np.array_equal(outs[0].cpu().numpy(), xp) and np.array_equal(outs[1].cpu().numpy(), np.where(fm[:, None], vf, vb))
    assert np.array_equal(pp.cpu().numpy(), prob) and np.array_equal(acc.cpu().numpy().astype(bool), am)
    xo = np.where(am[:, None], xp, x)
    assert np.array_equal(outs[2].cpu().numpy(), orc.wrap_2pi(xo))   # np.mod(x, 2 pi) bit-exact


def test_philox_draws_match_host_philox_and_are_standard_normal(pkg):
    from l2hmc_b200 import _lib
    from philox_ref import philox_normals, philox_uniforms
    lib = _lib.get_lib()
    N, D, seed, step, chain0 = 37, 40, 0x1234567890ABCDEF, 5, 1000
    z = torch.empty(N, D, device="cuda")
    ud, ua = torch.empty(N, device="cuda"), torch.empty(N, device="cuda")
    _lib.check(lib.l2hmc_philox_draws(seed, step, chain0, N, D, _lib.dptr(z), _lib.dptr(ud), _lib.dptr(ua), None))
    u_ref = philox_uniforms(seed, step, chain0, N)
    assert np.array_equal(ud.cpu().numpy(), u_ref[0]) and np.array_equal(ua.cpu().numpy(), u_ref[1])
    close(z, philox_normals(seed, step, chain0, N, D), rtol=1e-4, atol=1e-5)
    # sharding independence: chains [10, 20) drawn alone equal the slice of the full draw
    z2 = torch.empty(10, D, device="cuda")
    _lib.check(lib.l2hmc_philox_draws(seed, step, chain0 + 10, 10, D, _lib.dptr(z2), None, None, None))
    assert torch.equal(z2, z[10:20])
    big = torch.empty(4096, 512, device="cuda")
    _lib.check(lib.l2hmc_philox_draws(7, 0, 0, 4096, 512, _lib.dptr(big), None, None, None))
    assert abs(float(big.mean())) < 3e-3 and abs(float(big.var()) - 1) < 5e-3
    assert abs(float((big ** 4).mean()) - 3) < 0.05
