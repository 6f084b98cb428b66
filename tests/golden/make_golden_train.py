#!/usr/bin/env python
"""Golden vectors for the 'next' rows of SURVEY section 8(f), produced by the UNMODIFIED reference
(read from /root/reference, never copied) on the torch-CPU TensorFlow stand-in in tests/golden/tf_shim/:

  gauge_model.py:728-803    GaugeModel._calc_loss  (std loss on the metric, Fourier-smoothed charge loss, the x- and the
                            auxiliary z-chain; :94-108 project_angle_approx, :631-657 metric functions, :692-725 charges)
  gauge_model.py:805-832    tf.gradients(loss, dynamics.variables): d loss / d eps and d loss / d weights, stored for
                            the backward kernels of the next round (keys `<tag>.grad.*`)
  gauge_model.py:1473-1533  GaugeModel.calc_observables_stats (thermalisation cut, mean / scipy sem, charge histogram)

Run in the build container only:  python tests/golden/make_golden_train.py  -> tests/golden/train_cases.npz
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402  (sets the paths: tf shim first, then the reference tree)

import tensorflow as tf  # noqa: E402  (the shim)
import gauge_model as gm  # noqa: E402  (the reference module)

npf, leaf = mg.npf, mg.leaf


class _NumpyForGraphCode(object):
    """gauge_model.py:105-107 accumulates graph tensors into `np.zeros(x.shape)` with `+=`; under TF-1 the
    result is a float32 graph tensor.  NumPy refuses that in-place add for torch tensors, so the module-level
    name `np` of the (unmodified) reference module resolves `zeros(<tensor shape>)` to a float32 tensor of
    zeros and everything else to NumPy."""

    def __getattr__(self, name):
        return getattr(np, name)

    @staticmethod
    def zeros(shape, *a, **k):
        if isinstance(shape, torch.Size):
            return torch.zeros(tuple(shape), dtype=torch.float32)
        return np.zeros(shape, *a, **k)


gm.np = _NumpyForGraphCode()


def bare_model(lat, dyn, N, metric, loss_scale):
    """A GaugeModel carrying only what _calc_loss reads (its __init__ builds sessions, files, plots)."""
    model = object.__new__(gm.GaugeModel)
    model.lattice, model.dynamics = lat, dyn
    model.batch_size = model.num_samples = N
    model.metric_fn = gm.GaugeModel._create_metric_fn(metric)
    model.loss_scale = loss_scale
    return model


def gen_loss_cases(out):
    rng = np.random.default_rng(51)
    cases = [("hmc", 8, 8, 12, 5, 0.3, 4.0, "cos_diff", 1.0, dict(aux_weight=1., std_weight=1., charge_weight=1.)),
             ("generic", 4, 4, 10, 3, 0.2, 2.5, "cos_diff", 1.0, dict(aux_weight=1., std_weight=1., charge_weight=1.)),
             ("generic", 4, 4, 10, 3, 0.2, 2.5, "l2", 0.1, dict(aux_weight=0.5, std_weight=2., charge_weight=0.25)),
             ("hmc", 6, 10, 7, 3, 0.15, 3.0, "cos", 1.0, dict(aux_weight=1., std_weight=1., charge_weight=0.)),
             ("hmc", 8, 8, 9, 4, 0.35, 4.0, "l1", 0.5, dict(aux_weight=0.25, std_weight=1., charge_weight=1.)),
             ("conv3D", 8, 8, 5, 2, 0.1, 2.0, "cos2", 1.0, dict(aux_weight=1., std_weight=0.5, charge_weight=2.))]
    for ci, (arch, T, X, N, M, eps, beta, metric, ls, wts) in enumerate(cases):
        lat = mg.make_lattice(T, X, N)
        hmc = arch == "hmc"
        dyn = mg.make_dynamics(lat, hmc=hmc, num_steps=M, eps=eps, network_arch="generic" if hmc else arch)
        D = 2 * T * X
        tag = f"loss{ci}.{arch}.{T}x{X}.{metric}"
        out[tag + ".meta"] = np.array([T, X, N, M, int(hmc)], dtype=np.int64)
        out[tag + ".eps"], out[tag + ".beta"], out[tag + ".loss_scale"] = np.float32(eps), np.float32(beta), np.float32(ls)
        out[tag + ".weights"] = np.array([wts["aux_weight"], wts["std_weight"], wts["charge_weight"]], np.float32)
        out[tag + ".mask"] = npf(dyn.mask)
        x, v_f, v_b, u_dir, u_acc = mg.common_inputs(rng, N, D, out, tag)
        z = rng.standard_normal((N, D)).astype(np.float32)
        vz_f, vz_b = (rng.standard_normal((N, D)).astype(np.float32) for _ in range(2))
        uz_dir, uz_acc = (rng.uniform(size=N).astype(np.float32) for _ in range(2))
        for k, a in dict(z=z, vz_f=vz_f, vz_b=vz_b, uz_dir=uz_dir, uz_acc=uz_acc).items():
            out[f"{tag}.{k}"] = a
        if not hmc:
            tvec = dyn._format_time(tf.constant(0.0, dtype=tf.float32), tile=N)
            dyn.position_fn([torch.as_tensor(v_f), torch.as_tensor(x), tvec])
            dyn.momentum_fn([torch.as_tensor(x), torch.as_tensor(v_f), tvec])
            mg.liven_net(dyn.position_fn, rng)
            mg.liven_net(dyn.momentum_fn, rng)
            mg.dump_net(dyn.position_fn, tag + ".XNet", out)
            mg.dump_net(dyn.momentum_fn, tag + ".VNet", out)
        model = bare_model(lat, dyn, N, metric, ls)
        named = {"eps": dyn.eps}
        if not hmc:
            for net_name, net in (("XNet", dyn.position_fn), ("VNet", dyn.momentum_fn)):
                for nm in mg.net_layers(net):
                    named[f"{net_name}.{nm}.kernel"], named[f"{net_name}.{nm}.bias"] = getattr(net, nm).kernel, getattr(net, nm).bias
                named[f"{net_name}.coeff_scale"] = net.coeff_scale
                named[f"{net_name}.coeff_transformation"] = net.coeff_transformation
        for v in named.values():
            v.requires_grad_(True)
        # draw order inside _calc_loss (gauge_model.py:754-760 over gauge_dynamics.py:211-248):
        #   x-chain: v_f, v_b | u_dir, u_acc ; z ; z-chain: vz_f, vz_b | uz_dir, uz_acc
        tf.inject_randoms(normals=[v_f, v_b, z, vz_f, vz_b], uniforms=[u_dir, u_acc, uz_dir, uz_acc])
        loss, x_out, px, x_dq = model._calc_loss(leaf(x), tf.constant(beta, dtype=tf.float32), **wts)
        out[tag + ".loss"], out[tag + ".x_out"], out[tag + ".px"] = npf(loss), npf(x_out), npf(px)
        out[tag + ".x_dq"] = x_dq.detach().cpu().numpy().astype(np.int32)
        grads = tf.gradients(loss, list(named.values()))          # gauge_model.py:823
        for (nm, v), gr in zip(named.items(), grads):
            out[f"{tag}.grad.{nm}"] = npf(torch.zeros_like(v) if gr is None else gr)
        # the pieces, for localising a mismatch
        tf.inject_randoms(normals=[v_f, v_b], uniforms=[u_dir, u_acc])
        x_, _, px2, _ = dyn(leaf(x), tf.constant(beta, dtype=tf.float32))
        tf.inject_randoms(normals=[vz_f, vz_b], uniforms=[uz_dir, uz_acc])
        z_, _, pz, _ = dyn(leaf(z), tf.constant(beta, dtype=tf.float32))
        out[tag + ".x_prop"], out[tag + ".pz"] = npf(x_), npf(pz)
        for nm, a in (("x", x), ("x_prop", npf(x_)), ("z", z)):
            out[f"{tag}.qfft.{nm}"] = npf(model._calc_top_charges(torch.as_tensor(a), fft=True))
            out[f"{tag}.q.{nm}"] = npf(model._calc_top_charges(torch.as_tensor(a), fft=False))


def gen_stats_cases(out):
    rng = np.random.default_rng(77)
    for ci, (steps, N, therm_frac) in enumerate([(50, 8, 10), (23, 5, 4), (200, 64, 10), (9, 3, 10)]):
        tag = f"stats{ci}"
        actions = (rng.normal(60, 5, (steps, N))).astype(np.float32)
        plaqs = rng.uniform(0.3, 0.9, (steps, N)).astype(np.float32)
        charges = np.rint(rng.normal(0, 1.5, (steps, N))).astype(np.float32)        # the run loop stores floats
        obs = tuple({s: arr[s] for s in range(steps)} for arr in (actions, plaqs, charges))
        a_st, p_st, q_st, s_st, probs = gm.GaugeModel.calc_observables_stats(None, obs, therm_frac=therm_frac)
        out[tag + ".meta"] = np.array([steps, N, therm_frac], np.int64)
        out[tag + ".actions"], out[tag + ".plaqs"], out[tag + ".charges"] = actions, plaqs, charges
        for nm, st in (("actions", a_st), ("plaqs", p_st), ("charges", q_st), ("suscept", s_st)):
            out[f"{tag}.{nm}.mean"] = np.asarray(st[0], np.float64)
            out[f"{tag}.{nm}.err"] = np.asarray(st[1], np.float64)
        out[tag + ".prob.keys"] = np.array(list(probs.keys()), np.int64)
        out[tag + ".prob.vals"] = np.array(list(probs.values()), np.float64)


def main():
    assert os.path.isdir(mg.REF), "the reference tree is only present in the build container"
    out = {}
    gen_loss_cases(out)
    gen_stats_cases(out)
    np.savez_compressed(os.path.join(HERE, "train_cases.npz"), **out)
    print("train_cases.npz", len(out), "arrays")
    for k in out:
        if k.endswith(".loss"):
            print(k, out[k], "px", out[k[:-5] + ".px"].mean())


if __name__ == "__main__":
    main()
