#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ by running the UNMODIFIED
reference (read from /root/reference, never copied) on the torch-CPU
TensorFlow stand-in in tests/golden/tf_shim/.

Run in the build container only (the GPU box has no /root/reference):

    python tests/golden/make_golden.py

Every fixture holds the inputs (links, momenta, masks, uniforms, weights) and
the outputs of the reference's own methods, so the oracle and the CUDA path
can be checked against them without the reference being present.

Reference entry points exercised (file:line under /root/reference/l2hmc):
  lattice/lattice.py:285-313   GaugeLattice.calc_plaq_observables
  lattice/lattice.py:337-362   GaugeLattice.total_action
  dynamics/gauge_dynamics.py:195-259  GaugeDynamics.apply_transition
  dynamics/gauge_dynamics.py:261-313  transition_kernel
  dynamics/gauge_dynamics.py:412-483  _forward_lf / _backward_lf
  dynamics/gauge_dynamics.py:486-590  the four _update_* methods
  dynamics/gauge_dynamics.py:592-609  _compute_accept_prob
  dynamics/gauge_dynamics.py:698-709  grad_potential (tf.gradients)
  network/generic_net.py:95-146       GenericNet.call
  network/conv_net.py:210-280         ConvNet3D.call
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/l2hmc"

sys.path.insert(0, os.path.join(HERE, "tf_shim"))
sys.path.insert(0, REF)
if not hasattr(np, "int"):
    np.int = int  # lattice.py:148 uses the alias numpy >= 1.24 removed

import tensorflow as tf  # noqa: E402  (the shim)
from lattice.lattice import GaugeLattice  # noqa: E402
from dynamics.gauge_dynamics import GaugeDynamics  # noqa: E402

torch.set_grad_enabled(True)
torch.set_num_threads(1)


def npf(t):
    return t.detach().cpu().numpy().astype(np.float32) if isinstance(t, torch.Tensor) else np.asarray(t, np.float32)


def leaf(a):
    return torch.tensor(np.asarray(a, dtype=np.float32), requires_grad=True)


def make_lattice(T, X, N):
    return GaugeLattice(time_size=T, space_size=X, dim=2, link_type="U1", num_samples=N, rand=False)


def make_dynamics(lat, **kw):
    args = dict(num_steps=5, eps=0.25, network_arch="generic", hmc=False, eps_trainable=True,
                data_format="channels_last")
    args.update(kw)
    np.random.seed(42)          # gauge_model.py:62 seeds the legacy RNG the masks come from
    tf.set_init_seed(1234)
    return GaugeDynamics(lattice=lat, potential_fn=lat.get_energy_function(lat.samples), **args)


def net_layers(net):
    names = ["x_layer", "v_layer", "t_layer", "h_layer", "scale_layer", "translation_layer",
             "transformation_layer"]
    if hasattr(net, "conv_x1"):
        names = ["conv_x1", "conv_v1", "conv_x2", "conv_v2"] + names
    return names


def liven_net(net, rng, head_gain=3.0):
    """Make an initialised net 'trained-like': O(0.1) heads, non-zero biases/coeffs."""
    for nm in net_layers(net):
        lyr = getattr(net, nm)
        if nm in ("scale_layer", "translation_layer", "transformation_layer"):
            lyr.kernel.mul_(head_gain)
        lyr.bias.copy_(torch.as_tensor(rng.normal(0, 0.02, size=tuple(lyr.bias.shape)), dtype=torch.float32))
    net.coeff_scale.copy_(torch.as_tensor(rng.normal(0, 0.1, size=tuple(net.coeff_scale.shape)), dtype=torch.float32))
    net.coeff_transformation.copy_(
        torch.as_tensor(rng.normal(0, 0.1, size=tuple(net.coeff_transformation.shape)), dtype=torch.float32))


def dump_net(net, prefix, out):
    for nm in net_layers(net):
        lyr = getattr(net, nm)
        out[f"{prefix}.{nm}.kernel"] = npf(lyr.kernel)
        out[f"{prefix}.{nm}.bias"] = npf(lyr.bias)
    out[f"{prefix}.coeff_scale"] = npf(net.coeff_scale)
    out[f"{prefix}.coeff_transformation"] = npf(net.coeff_transformation)


def gen_lattice_cases():
    """Plaquette sums -> action / avg plaquette / integer Q, and the autodiff force."""
    out = {}
    rng = np.random.default_rng(7)
    for (T, X, N) in [(8, 8, 6), (16, 16, 3), (6, 10, 4), (4, 4, 5), (32, 32, 2)]:
        lat = make_lattice(T, X, N)
        dyn = make_dynamics(lat, hmc=True, num_steps=2, eps=0.1)
        x = rng.uniform(0, 2 * np.pi, size=(N, 2 * T * X)).astype(np.float32)
        x[N // 2:] = rng.uniform(-3 * np.pi, 5 * np.pi, size=x[N // 2:].shape).astype(np.float32)  # unwrapped links
        x[0] = 0.0  # cold start row: S=0, plaq=1, Q=0 (gauge_model_graph_mode.ipynb:257)
        xt = leaf(x)
        beta = tf.constant(2.5, dtype=tf.float32)
        act = lat.total_action(xt)
        act2, plaq, q = lat.calc_plaq_observables(xt)
        force = dyn.grad_potential(xt, beta)
        key = f"{T}x{X}"
        out[key + ".x"] = x
        out[key + ".beta"] = np.float32(2.5)
        out[key + ".total_action"] = npf(act)
        out[key + ".obs_action"] = npf(act2)
        out[key + ".obs_avg_plaq"] = npf(plaq)
        out[key + ".obs_charge"] = npf(q)
        out[key + ".force"] = npf(force)
        out[key + ".potential"] = npf(dyn.potential_energy(xt, beta))
        v = rng.standard_normal(size=x.shape).astype(np.float32)
        out[key + ".v"] = v
        out[key + ".kinetic"] = npf(dyn.kinetic_energy(torch.as_tensor(v)))
    np.savez_compressed(os.path.join(HERE, "lattice_cases.npz"), **out)
    print("lattice_cases.npz", len(out), "arrays")


def run_transition(dyn, x, beta, v_f, v_b, u_dir, u_acc, out, tag):
    """apply_transition plus the two transition_kernel calls it is built from."""
    tf.inject_randoms(normals=[v_f, v_b], uniforms=[u_dir, u_acc])
    xp, vp, p, xo = dyn.apply_transition(leaf(x), tf.constant(beta, dtype=tf.float32))
    out[tag + ".x_post"], out[tag + ".v_post"] = npf(xp), npf(vp)
    out[tag + ".accept_prob"], out[tag + ".x_out"] = npf(p), npf(xo)
    tf.inject_randoms(normals=[v_f])
    xf, vf, pf = dyn.transition_kernel(leaf(x), tf.constant(beta, dtype=tf.float32), forward=True)
    tf.inject_randoms(normals=[v_b])
    xb, vb, pb = dyn.transition_kernel(leaf(x), tf.constant(beta, dtype=tf.float32), forward=False)
    out[tag + ".x_f"], out[tag + ".v_f_out"], out[tag + ".p_f"] = npf(xf), npf(vf), npf(pf)
    out[tag + ".x_b"], out[tag + ".v_b_out"], out[tag + ".p_b"] = npf(xb), npf(vb), npf(pb)


def common_inputs(rng, N, D, out, tag, hot=True):
    x = (rng.uniform(0, 2 * np.pi, size=(N, D)) if hot else np.zeros((N, D))).astype(np.float32)
    v_f = rng.standard_normal(size=(N, D)).astype(np.float32)
    v_b = rng.standard_normal(size=(N, D)).astype(np.float32)
    u_dir = rng.uniform(size=(N,)).astype(np.float32)
    u_acc = rng.uniform(size=(N,)).astype(np.float32)
    for k, a in dict(x=x, v_f=v_f, v_b=v_b, u_dir=u_dir, u_acc=u_acc).items():
        out[f"{tag}.{k}"] = a
    return x, v_f, v_b, u_dir, u_acc


def gen_hmc_cases():
    """Config 1: U(1) 8x8 plain HMC, beta=2, M=10, eps=0.1 (+ a 16x16 and a ragged case)."""
    out = {}
    rng = np.random.default_rng(11)
    for (T, X, N, M, eps, beta) in [(8, 8, 16, 10, 0.1, 2.0), (16, 16, 6, 3, 0.2921, 4.0),
                                    (6, 10, 5, 4, 0.15, 3.0), (32, 32, 3, 10, 0.1, 4.0)]:
        lat = make_lattice(T, X, N)
        dyn = make_dynamics(lat, hmc=True, num_steps=M, eps=eps)
        tag = f"{T}x{X}"
        out[tag + ".meta"] = np.array([T, X, N, M], dtype=np.int64)
        out[tag + ".eps"], out[tag + ".beta"] = np.float32(eps), np.float32(beta)
        out[tag + ".mask"] = npf(dyn.mask)
        x, v_f, v_b, u_dir, u_acc = common_inputs(rng, N, 2 * T * X, out, tag)
        run_transition(dyn, x, beta, v_f, v_b, u_dir, u_acc, out, tag)
    np.savez_compressed(os.path.join(HERE, "hmc_cases.npz"), **out)
    print("hmc_cases.npz", len(out), "arrays")


def gen_l2hmc_cases():
    """Augmented leapfrog with S/T/Q nets: per-sub-update, per-LF-step and whole transition."""
    out = {}
    rng = np.random.default_rng(23)
    for (arch, T, X, N, M, eps, beta) in [("generic", 4, 4, 6, 5, 0.25, 2.0),
                                          ("conv3D", 4, 4, 6, 3, 0.2, 3.0),
                                          ("conv3D", 8, 8, 4, 2, 0.1, 2.0)]:
        lat = make_lattice(T, X, N)
        dyn = make_dynamics(lat, hmc=False, num_steps=M, eps=eps, network_arch=arch)
        D = 2 * T * X
        tag = f"{arch}.{T}x{X}"
        out[tag + ".meta"] = np.array([T, X, N, M], dtype=np.int64)
        out[tag + ".eps"], out[tag + ".beta"] = np.float32(eps), np.float32(beta)
        out[tag + ".mask"] = npf(dyn.mask)
        x, v_f, v_b, u_dir, u_acc = common_inputs(rng, N, D, out, tag)
        # build the lazily-created layers, then make the nets trained-like
        tvec = dyn._format_time(tf.constant(0.0, dtype=tf.float32), tile=N)
        dyn.position_fn([torch.as_tensor(v_f), torch.as_tensor(x), tvec])
        dyn.momentum_fn([torch.as_tensor(x), torch.as_tensor(v_f), tvec])
        liven_net(dyn.position_fn, rng)
        liven_net(dyn.momentum_fn, rng)
        dump_net(dyn.position_fn, tag + ".XNet", out)
        dump_net(dyn.momentum_fn, tag + ".VNet", out)
        # --- a bare network call (generic_net.py:129-146 / conv_net.py:247-280)
        for name, net in (("XNet", dyn.position_fn), ("VNet", dyn.momentum_fn)):
            tv = dyn._format_time(tf.constant(1.0, dtype=tf.float32), tile=N)
            S, Tr, Q = net([torch.as_tensor(v_f), torch.as_tensor(x), tv])
            out[f"{tag}.{name}.call.tvec"] = npf(tv)
            out[f"{tag}.{name}.call.S"], out[f"{tag}.{name}.call.T"], out[f"{tag}.{name}.call.Q"] = \
                npf(S), npf(Tr), npf(Q)
        # --- the four sub-updates at step 1 (gauge_dynamics.py:486-590)
        b = tf.constant(beta, dtype=tf.float32)
        step = tf.constant(1.0, dtype=tf.float32)
        tv = dyn._format_time(step, tile=N)
        m, mi = dyn._get_mask_while(step)
        for nm, fn in (("vf", dyn._update_momentum_forward), ("vb", dyn._update_momentum_backward)):
            vv, ld = fn(leaf(x), torch.as_tensor(v_f), b, tv)
            out[f"{tag}.sub.{nm}.v"], out[f"{tag}.sub.{nm}.logdet"] = npf(vv), npf(ld)
        for nm, fn in (("xf", dyn._update_position_forward), ("xb", dyn._update_position_backward)):
            xx, ld = fn(leaf(x), torch.as_tensor(v_f), tv, m, mi)
            out[f"{tag}.sub.{nm}.x"], out[f"{tag}.sub.{nm}.logdet"] = npf(xx), npf(ld)
        # --- one LF step each way (gauge_dynamics.py:412-483)
        xx, vv, ld = dyn._forward_lf(leaf(x), torch.as_tensor(v_f), b, step)
        out[tag + ".lf_f.x"], out[tag + ".lf_f.v"], out[tag + ".lf_f.logdet"] = npf(xx), npf(vv), npf(ld)
        xx, vv, ld = dyn._backward_lf(leaf(x), torch.as_tensor(v_f), b, step)
        out[tag + ".lf_b.x"], out[tag + ".lf_b.v"], out[tag + ".lf_b.logdet"] = npf(xx), npf(vv), npf(ld)
        # --- whole MC step
        run_transition(dyn, x, beta, v_f, v_b, u_dir, u_acc, out, tag)
    np.savez_compressed(os.path.join(HERE, "l2hmc_cases.npz"), **out)
    print("l2hmc_cases.npz", len(out), "arrays")


def load_net(net, w):
    """Install given weights (Keras layouts, oracle key names) into a built reference net."""
    for nm in net_layers(net):
        lyr = getattr(net, nm)
        assert tuple(lyr.kernel.shape) == w[nm + ".kernel"].shape, (nm, tuple(lyr.kernel.shape), w[nm + ".kernel"].shape)
        lyr.kernel.copy_(torch.as_tensor(w[nm + ".kernel"]))
        lyr.bias.copy_(torch.as_tensor(w[nm + ".bias"]))
    net.coeff_scale.copy_(torch.as_tensor(w["coeff_scale"]).reshape(net.coeff_scale.shape))
    net.coeff_transformation.copy_(torch.as_tensor(w["coeff_transformation"]).reshape(net.coeff_transformation.shape))


def gen_l2hmc_config3_case():
    """BASELINE config 3's shape through the unmodified reference: 16x16, conv3D nets (F = 16, H = 1024), M = 5,
    eps = 0.25 (globals.py:27-30), N = 4.  The 2 x 3.7 M weights are NOT stored: they are a pure function of
    default_rng(WEIGHT_SEED) (tests/parity_util.trained_like_weights); the file keeps their sha256."""
    sys.path.insert(0, os.path.dirname(HERE))
    from parity_util import trained_like_weights, weights_digest
    out = {}
    arch, T, X, N, M, eps, beta, wseed = "conv3D", 16, 16, 4, 5, 0.25, 3.0, 316
    rng = np.random.default_rng(31)
    lat = make_lattice(T, X, N)
    dyn = make_dynamics(lat, hmc=False, num_steps=M, eps=eps, network_arch=arch)
    D = 2 * T * X
    tag = f"{arch}.{T}x{X}"
    out[tag + ".meta"] = np.array([T, X, N, M], dtype=np.int64)
    out[tag + ".eps"], out[tag + ".beta"] = np.float32(eps), np.float32(beta)
    out[tag + ".mask"] = npf(dyn.mask)
    x, v_f, v_b, u_dir, u_acc = common_inputs(rng, N, D, out, tag)
    # links near a thermalised configuration (fluctuations around the cold start) and heads of O(0.1): the accept
    # probabilities then spread over (1e-4, 1) instead of underflowing as they do from a hot start at this eps
    x = np.mod(rng.normal(0, 0.5, size=(N, D)), 2 * np.pi).astype(np.float32)
    out[tag + ".x"] = x
    tvec = dyn._format_time(tf.constant(0.0, dtype=tf.float32), tile=N)
    dyn.position_fn([torch.as_tensor(v_f), torch.as_tensor(x), tvec])     # builds the lazily-created layers
    dyn.momentum_fn([torch.as_tensor(x), torch.as_tensor(v_f), tvec])
    ws = trained_like_weights(arch, T, X, np.random.default_rng(wseed), gain=1.0)
    load_net(dyn.position_fn, ws["XNet"])
    load_net(dyn.momentum_fn, ws["VNet"])
    out[tag + ".weight_seed"] = np.int64(wseed)
    out[tag + ".weight_sha256"] = np.frombuffer(bytes.fromhex(weights_digest(ws)), dtype=np.uint8)
    b = tf.constant(beta, dtype=tf.float32)
    step = tf.constant(1.0, dtype=tf.float32)
    tv = dyn._format_time(step, tile=N)
    for name, net in (("XNet", dyn.position_fn), ("VNet", dyn.momentum_fn)):
        S, Tr, Q = net([torch.as_tensor(v_f), torch.as_tensor(x), tv])
        out[f"{tag}.{name}.call.tvec"] = npf(tv)
        out[f"{tag}.{name}.call.S"], out[f"{tag}.{name}.call.T"], out[f"{tag}.{name}.call.Q"] = npf(S), npf(Tr), npf(Q)
    m, mi = dyn._get_mask_while(step)
    for nm, fn in (("vf", dyn._update_momentum_forward), ("vb", dyn._update_momentum_backward)):
        vv, ld = fn(leaf(x), torch.as_tensor(v_f), b, tv)
        out[f"{tag}.sub.{nm}.v"], out[f"{tag}.sub.{nm}.logdet"] = npf(vv), npf(ld)
    for nm, fn in (("xf", dyn._update_position_forward), ("xb", dyn._update_position_backward)):
        xx, ld = fn(leaf(x), torch.as_tensor(v_f), tv, m, mi)
        out[f"{tag}.sub.{nm}.x"], out[f"{tag}.sub.{nm}.logdet"] = npf(xx), npf(ld)
    xx, vv, ld = dyn._forward_lf(leaf(x), torch.as_tensor(v_f), b, step)
    out[tag + ".lf_f.x"], out[tag + ".lf_f.v"], out[tag + ".lf_f.logdet"] = npf(xx), npf(vv), npf(ld)
    xx, vv, ld = dyn._backward_lf(leaf(x), torch.as_tensor(v_f), b, step)
    out[tag + ".lf_b.x"], out[tag + ".lf_b.v"], out[tag + ".lf_b.logdet"] = npf(xx), npf(vv), npf(ld)
    run_transition(dyn, x, beta, v_f, v_b, u_dir, u_acc, out, tag)
    # observables of the state the step ends in (lattice.py:285-313)
    act, plaq, q = lat.calc_plaq_observables(torch.as_tensor(out[tag + ".x_out"]))
    out[tag + ".obs_action"], out[tag + ".obs_avg_plaq"], out[tag + ".obs_charge"] = npf(act), npf(plaq), npf(q)
    np.savez_compressed(os.path.join(HERE, "l2hmc_config3_case.npz"), **out)
    print("l2hmc_config3_case.npz", len(out), "arrays")


if __name__ == "__main__":
    assert os.path.isdir(REF), "the reference tree is only present in the build container"
    gen_lattice_cases()
    gen_hmc_cases()
    gen_l2hmc_cases()
    gen_l2hmc_config3_case()
