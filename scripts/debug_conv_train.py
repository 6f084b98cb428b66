import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from test_gpu_train_rows import build_model
g = dict(np.load(os.path.join(ROOT, "tests/golden/train_cases.npz")))
tag = [k[:-5] for k in g if k.startswith("loss5") and k.endswith(".meta")][0]
model, meta = build_model(g, tag)
args = dict(draws_x=(g[tag + ".v_f"], g[tag + ".v_b"], g[tag + ".u_dir"], g[tag + ".u_acc"]), z=g[tag + ".z"],
            draws_z=(g[tag + ".vz_f"], g[tag + ".vz_b"], g[tag + ".uz_dir"], g[tag + ".uz_acc"]))
l_inf, xo, px, dq = model._calc_loss(g[tag + ".x"], float(g[tag + ".beta"]), **args)
print("inference-path loss", float(l_inf), "golden", float(g[tag + ".loss"]))
print("px inf", px.cpu().numpy(), "golden", g[tag + ".px"])
loss, grads, x_out, px2, x_dq = model._calc_loss_and_grads(g[tag + ".x"], float(g[tag + ".beta"]), **args)
fw = model.last_forward
print("train-path loss", float(loss))
print("px train", px2.cpu().numpy(), "pz train", fw["pz"].cpu().numpy(), "golden pz", g[tag + ".pz"])
print("x_prop err", np.abs(fw["x_prop"].cpu().numpy() - g[tag + ".x_prop"]).max())
print("u_dir", g[tag + ".u_dir"], g[tag + ".uz_dir"])
for k, v in grads.items():
    w = g[f"{tag}.grad.{k}"]
    print(k, float(np.abs(v.cpu().numpy().reshape(w.shape) - w).max() / max(np.abs(w).max(), 1e-30)), float(np.abs(w).max()))
