#!/usr/bin/env python
"""Golden vectors for the MoG / generic-target L2HMC path (BASELINE config 5), produced by the
UNMODIFIED reference modules (read from /root/reference, never copied) on the torch-CPU TensorFlow
stand-in in tests/golden/tf_shim/:

  utils/distributions.py:124-158   GMM.get_energy_function (quadratic_gaussian :32-39)
  utils/network.py:89-114,359-454  network() factory, Linear / Zip / Parallel / Sequential / ScaleTanh
  utils/dynamics.py:120-319        Dynamics._forward_step / _backward_step / forward / backward / p_accept
  utils/sampler.py:28-59           propose, tf_accept

Run in the build container only:  python tests/golden/make_golden_mog.py  -> tests/golden/mog_cases.npz
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/l2hmc"
sys.path.insert(0, os.path.join(HERE, "tf_shim"))
sys.path.insert(0, REF)

import tensorflow as tf  # noqa: E402  (the shim)
from utils.distributions import GMM, gen_ring  # noqa: E402
from utils.dynamics import Dynamics  # noqa: E402
from utils.network import network  # noqa: E402
from utils.sampler import propose  # noqa: E402

torch.set_num_threads(1)


def npf(t):
    return t.detach().cpu().numpy().astype(np.float32) if isinstance(t, torch.Tensor) else np.asarray(t, np.float32)


def leaf(a):
    return torch.tensor(np.asarray(a, dtype=np.float32), requires_grad=True)


def net_linears(net):
    """The reference's Sequential -> the seven Linear layers and two ScaleTanh, by role."""
    zip_, lin1, par = net.layers[0], net.layers[3], net.layers[5]
    return {"v_layer": zip_.layers[0], "x_layer": zip_.layers[1], "t_layer": zip_.layers[2], "h_layer": lin1,
            "scale_layer": par.layers[0].layers[0], "translation_layer": par.layers[1],
            "transformation_layer": par.layers[2].layers[0]}, par.layers[0].layers[1], par.layers[2].layers[1]


def liven_and_dump(net, rng, prefix, out, head_gain):
    lin, st_s, st_f = net_linears(net)
    for nm, lyr in lin.items():
        if nm in ("scale_layer", "translation_layer", "transformation_layer"):
            lyr.W.mul_(head_gain)
        lyr.b.copy_(torch.as_tensor(rng.normal(0, 0.05, size=tuple(lyr.b.shape)), dtype=torch.float32))
        out[f"{prefix}.{nm}.kernel"], out[f"{prefix}.{nm}.bias"] = npf(lyr.W), npf(lyr.b)
    for key, st in (("coeff_scale", st_s), ("coeff_transformation", st_f)):
        coeff = torch.as_tensor(rng.normal(0, 0.2, size=tuple(st.scale.shape)), dtype=torch.float32)
        st.scale = torch.exp(coeff)                     # ScaleTanh holds exp(variable) (network.py:436-441)
        out[f"{prefix}.{key}"] = npf(coeff)


def gen_case(tag, gmm, x_dim, L, eps, H, temp, N, hmc, rng, out):
    np.random.seed(42)                                  # Dynamics._init_mask draws from the legacy global RNG
    tf.set_init_seed(77)
    factory = (lambda x_dim, scope, factor: network(x_dim, scope, factor, num_nodes=H))
    with torch.no_grad():
        dyn = Dynamics(x_dim, gmm.get_energy_function(), trajectory_length=L, eps=eps, hmc=hmc,
                       net_factory=factory, use_temperature=True)
        if not hmc:
            liven_and_dump(dyn.XNet, rng, tag + ".XNet", out, 10.0)
            liven_and_dump(dyn.VNet, rng, tag + ".VNet", out, 10.0)
    dyn.temperature.fill_(temp)
    out[tag + ".meta"] = np.array([x_dim, L, H, N, int(hmc), gmm.nb_mixtures], dtype=np.int64)
    out[tag + ".eps"], out[tag + ".temperature"] = np.float32(eps), np.float32(temp)
    out[tag + ".mask"] = npf(dyn.mask)
    out[tag + ".mus"] = np.asarray(gmm.mus, np.float32)
    out[tag + ".i_sigmas"] = np.asarray(gmm.i_sigmas, np.float32)
    out[tag + ".constants"] = np.asarray(gmm.constants, np.float32)
    centres = np.asarray(gmm.mus, np.float32)[np.arange(N) % gmm.nb_mixtures]     # start near the modes
    x = (centres + 0.25 * rng.standard_normal((N, x_dim))).astype(np.float32)
    v_f = rng.standard_normal((N, x_dim)).astype(np.float32)
    v_b = rng.standard_normal((N, x_dim)).astype(np.float32)
    dir_mask = rng.integers(0, 2, size=(N, 1)).astype(np.float32)
    u_acc = rng.uniform(size=(N,)).astype(np.float32)
    for k, a in dict(x=x, v_f=v_f, v_b=v_b, dir_mask=dir_mask, u_acc=u_acc).items():
        out[f"{tag}.{k}"] = a
    out[tag + ".energy"] = npf(dyn.energy(leaf(x)))
    out[tag + ".grad"] = npf(dyn.grad_energy(leaf(x)))
    if not hmc:
        tv = dyn._format_time(tf.constant(1.0, dtype=tf.float32), tile=N)
        out[tag + ".call.tvec"] = npf(tv)
        for name, net in (("XNet", dyn.XNet), ("VNet", dyn.VNet)):
            S, Tr, Q = net([torch.as_tensor(v_f), torch.as_tensor(x), tv, None])
            out[f"{tag}.{name}.call.S"], out[f"{tag}.{name}.call.T"], out[f"{tag}.{name}.call.Q"] = \
                npf(S), npf(Tr), npf(Q)
    step = tf.constant(1.0 if L > 1 else 0.0, dtype=tf.float32)
    xo, vo, lj = dyn._forward_step(leaf(x), torch.as_tensor(v_f), step)
    out[tag + ".fstep.x"], out[tag + ".fstep.v"], out[tag + ".fstep.logjac"] = npf(xo), npf(vo), npf(lj)
    xo, vo, lj = dyn._backward_step(leaf(x), torch.as_tensor(v_f), step)
    out[tag + ".bstep.x"], out[tag + ".bstep.v"], out[tag + ".bstep.logjac"] = npf(xo), npf(vo), npf(lj)
    X, V, p = dyn.forward(leaf(x), init_v=torch.as_tensor(v_f))
    out[tag + ".fwd.x"], out[tag + ".fwd.v"], out[tag + ".fwd.p"] = npf(X), npf(V), npf(p)
    X, V, p = dyn.backward(leaf(x), init_v=torch.as_tensor(v_b))
    out[tag + ".bwd.x"], out[tag + ".bwd.v"], out[tag + ".bwd.p"] = npf(X), npf(V), npf(p)
    if hmc:
        tf.inject_randoms(normals=[torch.as_tensor(v_f)], uniforms=[torch.as_tensor(u_acc)])
    else:
        tf.inject_randoms(normals=[torch.as_tensor(v_f), torch.as_tensor(v_b)],
                          uniforms=[torch.as_tensor(dir_mask), torch.as_tensor(u_acc)])
    Lx, _, px, outs = propose(leaf(x), dyn, do_mh_step=True)
    out[tag + ".propose.Lx"], out[tag + ".propose.px"], out[tag + ".propose.x_out"] = npf(Lx), npf(px), npf(outs[0])


def main():
    assert os.path.isdir(REF), "the reference tree is only present in the build container"
    out = {}
    rng = np.random.default_rng(31)
    means = np.array([[1.0, 0.0], [-1.0, 0.0]], np.float32)          # mog_model.py:1086-1094
    two = GMM(means, [0.05 * np.eye(2).astype(np.float32)] * 2, [0.5, 0.5])
    gen_case("mog2.L2", two, 2, 2, 0.1, 50, 1.0, 12, False, rng, out)   # config 5 shape: k=2, M=2, H=50
    gen_case("mog2.L5.T2", two, 2, 5, 0.2, 50, 2.5, 9, False, rng, out)
    _, ring = gen_ring(r=1.5, var=0.1, nb_mixtures=4)
    gen_case("ring4.L3", ring, 2, 3, 0.15, 10, 1.0, 7, False, rng, out)
    gen_case("mog2.hmc", two, 2, 4, 0.05, 50, 1.0, 8, True, rng, out)
    # a 6-dimensional target: the same code path is not special-cased to k = 2
    mus6 = [rng.standard_normal(6).astype(np.float32) for _ in range(3)]
    g6 = GMM(mus6, [0.5 * np.eye(6).astype(np.float32)] * 3, [0.2, 0.3, 0.5])
    gen_case("gmm6.L2", g6, 6, 2, 0.1, 20, 1.0, 5, False, rng, out)
    np.savez_compressed(os.path.join(HERE, "mog_cases.npz"), **out)
    print("mog_cases.npz", len(out), "arrays")


if __name__ == "__main__":
    main()
