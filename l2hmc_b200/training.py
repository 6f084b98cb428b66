"""Training step of the L2HMC sampler: loss, explicit backward pass, Adam -- SURVEY section 8(f) ranks 1, 2 and 4.

Mirrors, with the same meaning (reference l2hmc/gauge_model.py):
  :728-803    _calc_loss                      (forward: both chains, recorded on a tape)
  :805-832    _calc_loss_and_grads            tf.gradients(loss, dynamics.variables) [+ clip_by_global_norm]
  :925-945    exponential_decay (staircase) learning rate, tf.train.AdamOptimizer, Horovod's lr * size / averaged grads
  :1039-1046  update_beta                     (GaugeModel.update_beta)
  :519-556, 1104-1110  saved state            (weights, eps, Adam moments, samples, step, beta -> one torch.save file)

TensorFlow derives the gradients by graph autodiff; here every adjoint is a kernel of csrc/train.cu and this module
only walks the tape (which call, which pointers).  Layout of one training step on the device:

  * the x-chain and the auxiliary z-chain are stacked into ONE batch of R = 2N rows, and chains going forward and
    backward share every launch (row_set), so a step is 4M network calls regardless of directions;
  * per network the tape holds, for its 2M calls, both inputs, h1, h2, S, T, Q and -- filled by the backward sweep --
    the gradients of the head pre-activations and of both hidden layers.  Every weight gradient is then ONE
    transposed GEMM over the whole tape (K = 2M * R) instead of 2M small ones, split along K in fixed order;
  * parameters, gradients and Adam moments are three flat arenas; the networks' weight tensors are views into the
    first, so the optimiser is one launch and the multi-GPU gradient exchange one all-reduce.
There is no CPU or autograd fallback: every number comes from the sm_100a kernels behind include/l2hmc_b200.h.
"""
import ctypes
import math

import numpy as np
import torch

from . import _lib


def _ptr(t, offset=0):
    """Device address of element `offset` (in floats) of a CUDA fp32 tensor."""
    return ctypes.c_void_p(t.data_ptr() + 4 * int(offset))


class _Gemm(object):
    """Thin caller of l2hmc_gemm (tcgen05 for dense contractions, FFMA otherwise) with a cached workspace."""

    def __init__(self, device):
        self.device = device
        self._ws = None
        self.launches = 0

    def __call__(self, M, N, segs, C, ldc, accumulate=False, relu_gate=None, ldg=0, col_scale=None, invert=False,
                 row_set=None, col_scale_alt=None, invert_alt=False, ksplit=0):
        a = _lib.GemmArgs()
        a.M, a.N, a.nseg = int(M), int(N), len(segs)
        for i, (A, a_rs, a_cs, B, b_rs, b_cs, K) in enumerate(segs):
            s = a.seg[i]
            s.A, s.a_rs, s.a_cs, s.B, s.b_rs, s.b_cs, s.K = A, int(a_rs), int(a_cs), B, int(b_rs), int(b_cs), int(K)
        a.C, a.ldc, a.accumulate = C, int(ldc), int(bool(accumulate))
        a.relu_gate, a.ldg = relu_gate, int(ldg)
        a.col_scale, a.col_scale_invert = col_scale, int(bool(invert))
        a.row_set = row_set
        a.col_scale_alt, a.col_scale_invert_alt = col_scale_alt, int(bool(invert_alt))
        lib = _lib.get_lib()
        ws = None
        need = int(lib.l2hmc_gemm_workspace_bytes(ctypes.byref(a), int(ksplit)))      # ksplit = 0: the library chooses
        if need:
            if self._ws is None or self._ws.numel() * 4 < need:
                self._ws = torch.empty((need + 3) // 4, dtype=torch.float32, device=self.device)
            ws = _ptr(self._ws)
        _lib.check(lib.l2hmc_gemm(ctypes.byref(a), int(ksplit), ws, need, _lib.stream_ptr()))
        self.launches += 2 if need else 1


    def col_sum(self, B, ld, rows, cols, out):
        """out[c] = sum_r B[r*ld + c] (l2hmc_col_sum: fixed two-stage order)."""
        lib = _lib.get_lib()
        need = int(lib.l2hmc_col_sum_workspace_bytes(int(rows), int(cols)))
        if self._ws is None or self._ws.numel() * 4 < need:
            self._ws = torch.empty((need + 3) // 4, dtype=torch.float32, device=self.device)
        _lib.check(lib.l2hmc_col_sum(B, int(ld), int(rows), int(cols), out, _ptr(self._ws), need, _lib.stream_ptr()))
        self.launches += 2


class _NetTape(object):
    """Activations and activation gradients of one network's 2M calls (rows = calls * R)."""

    def __init__(self, calls, R, K, H, D, device):
        f = lambda *s: torch.empty(*s, dtype=torch.float32, device=device)
        self.calls, self.R, self.K, self.H, self.D = calls, R, K, H, D
        self.in0, self.in1 = f(calls, R, K), f(calls, R, K)
        self.h1, self.h2 = f(calls, R, H), f(calls, R, H)
        self.S, self.T, self.Q = f(calls, R, D), f(calls, R, D), f(calls, R, D)
        self.dA = f(calls, R, 3 * D)
        self.dz2, self.dz1 = f(calls, R, H), f(calls, R, H)
        self.ps, self.pq = f(calls, R, D), f(calls, R, D)
        self.tau1 = f(calls, R, 3)          # (cos, sin, 1) per row: t_layer's input and the bias "input"


class FlatParams(object):
    """dynamics.variables (gauge_dynamics.py `variables`: eps, then XNet, then VNet) as ONE fp32 arena; the nets'
    weight tensors are re-pointed to views of it.  `names` follows the golden fixtures ('eps', 'XNet.<key>', ...)."""

    def __init__(self, dynamics):
        self.dynamics = dynamics
        dev = dynamics.device
        entries = [("eps", dynamics.eps.reshape(1))]
        if not dynamics.hmc:
            for net_name, net in (("XNet", dynamics.position_fn), ("VNet", dynamics.momentum_fn)):
                for k in net._weight_keys():
                    entries.append((f"{net_name}.{k}", net.weights[k]))
        self.names = [n for n, _ in entries]
        self.shapes = [tuple(t.shape) for _, t in entries]
        self.offsets, off = [], 0
        for _, t in entries:
            self.offsets.append(off)
            off += (t.numel() + 3) // 4 * 4                      # 16-byte aligned views
        self.size = off
        self.flat = torch.zeros(off, dtype=torch.float32, device=dev)
        for (n, t), o in zip(entries, self.offsets):
            self.flat[o:o + t.numel()].copy_(t.reshape(-1))
        self.grad = torch.zeros_like(self.flat)
        dynamics._eps_dev = self.flat[0]
        self._repoint()

    def view(self, buf, name):
        i = self.names.index(name)
        n = int(np.prod(self.shapes[i])) if self.shapes[i] else 1
        return buf[self.offsets[i]:self.offsets[i] + n].view(self.shapes[i])

    def _repoint(self):
        dyn = self.dynamics
        if dyn.hmc:
            return
        for net_name, net in (("XNet", dyn.position_fn), ("VNet", dyn.momentum_fn)):
            for k in net._weight_keys():
                net.weights[k] = self.view(self.flat, f"{net_name}.{k}")
            net._struct = None
            if hasattr(net, "_conv_structs"):
                net._conv_structs = None

    def sync_to_dynamics(self):
        """After an optimiser step: the host copy of eps is stale (fetched -- one sync -- only when something asks for
        it: the log row, an inference call; the reference fetches dynamics.eps every step, gauge_model.py:1166) and
        the packed tensor-core weight copies are dropped."""
        dyn = self.dynamics
        dyn._eps_dev = self.flat[0]
        dyn._eps_host = None
        if not dyn.hmc:
            for net in (dyn.position_fn, dyn.momentum_fn):
                net._struct = None

    def grads_dict(self):
        return {n: self.view(self.grad, n) for n in self.names}


class TrainEngine(object):
    """Forward with tape + backward sweep for one GaugeDynamics (generic nets or plain HMC)."""

    def __init__(self, dynamics, metric="cos_diff", loss_scale=1.):
        if not dynamics.hmc and dynamics.network_arch not in ("generic", "conv3D"):
            raise NotImplementedError("the backward pass is built for network_arch in ('generic', 'conv3D') and hmc=True")
        self.dyn = dynamics
        self.conv = (not dynamics.hmc) and dynamics.network_arch == "conv3D"
        self._conv_partial = {}
        self._conv_tapes = {}
        self.device = dynamics.device
        self.metric, self.loss_scale = metric, float(loss_scale)
        self.params = FlatParams(dynamics)
        self.gemm = _Gemm(self.device)
        self._tapes, self._structs = {}, {}
        self._tc_forward = False
        self._scratch = torch.empty(1024, dtype=torch.float32, device=self.device)
        self.launches = 0

    # ------------------------------------------------------------------ small launch helpers
    def _call(self, fn, *args):
        _lib.check(fn(*args, _lib.stream_ptr()))
        self.launches += 1

    def _new(self, *shape):
        return torch.empty(*shape, dtype=torch.float32, device=self.device)

    def _tape(self, name, R):
        dyn = self.dyn
        key = (name, R)
        if key not in self._tapes:
            net = dyn.position_fn if name == "XNet" else dyn.momentum_fn
            self._tapes[key] = _NetTape(2 * dyn.num_steps, R, int(net.num_features), int(net.num_hidden),
                                        int(net.x_dim), self.device)
        return self._tapes[key]

    def _net_struct(self, net):
        """l2hmc_net_weights without the packed tensor-core copy (training runs the FFMA path, whose hidden
        activations stay on the tape).  The weights are views of the flat arena, updated in place, so the
        pointers stay valid across optimiser steps."""
        key = id(net)
        if key not in self._structs:
            w, s = net.weights, _lib.NetWeights()
            s.K, s.H, s.D = int(net.num_features), int(net.num_hidden), int(net.x_dim)
            p = lambda k: _lib.dptr(w[k])
            s.w_in0, s.b_in0 = p("v_layer.kernel"), p("v_layer.bias")
            s.w_in1, s.b_in1 = p("x_layer.kernel"), p("x_layer.bias")
            s.w_t, s.b_t = p("t_layer.kernel"), p("t_layer.bias")
            s.w_h, s.b_h = p("h_layer.kernel"), p("h_layer.bias")
            s.w_s, s.b_s = p("scale_layer.kernel"), p("scale_layer.bias")
            s.w_tr, s.b_tr = p("translation_layer.kernel"), p("translation_layer.bias")
            s.w_q, s.b_q = p("transformation_layer.kernel"), p("transformation_layer.bias")
            s.coeff_s, s.coeff_q = p("coeff_scale"), p("coeff_transformation")
            s.tanh_on_q, s.packed = 0, None
            self._structs[key] = s
        return self._structs[key]

    def _pack_for_forward(self, R):
        """Forward layers on tcgen05 when the stacked batch has >= 1024 rows and the shapes tile (the rule of
        csrc/train_step.cu): refresh the TS32 copies of the dense weights, which moved since the last step."""
        dyn, lib = self.dyn, _lib.get_lib()
        nets = (dyn.position_fn, dyn.momentum_fn)
        fits = lambda n: n.num_features >= 64 and n.num_hidden >= 128 and n.x_dim >= 64 and n.x_dim % 16 == 0
        if R < 1024 or not all(fits(n) for n in nets):
            return False
        if not hasattr(self, "_packed_structs"):
            self._packed_structs, self._packed_bufs = {}, {}
        for net in nets:
            base = self._net_struct(net)
            nbytes = int(lib.l2hmc_net_pack_bytes(base.K, base.H, base.D))
            if id(net) not in self._packed_bufs:
                self._packed_bufs[id(net)] = torch.empty((nbytes + 3) // 4, dtype=torch.float32, device=self.device)
                s = _lib.NetWeights()
                ctypes.memmove(ctypes.byref(s), ctypes.byref(base), ctypes.sizeof(s))
                s.packed = _lib.dptr(self._packed_bufs[id(net)])
                self._packed_structs[id(net)] = s
            _lib.check(lib.l2hmc_net_pack(base, _lib.dptr(self._packed_bufs[id(net)]), nbytes, _lib.stream_ptr()))
        return True

    def _net_forward(self, net, tape, j, R, tau, tau_alt, row_set):
        lib = _lib.get_lib()
        c = _lib.NetCall()
        c.mode = _lib.NET_STQ
        c.in0, c.in1 = _ptr(tape.in0[j]), _ptr(tape.in1[j])
        c.tau_cos, c.tau_sin = float(tau[0]), float(tau[1])
        c.tau_cos_alt, c.tau_sin_alt = float(tau_alt[0]), float(tau_alt[1])
        c.row_set = _lib.dptr(row_set, torch.uint8)
        c.dir = c.dir_alt = 1
        c.N = int(R)
        c.S, c.T, c.Q = _ptr(tape.S[j]), _ptr(tape.T[j]), _ptr(tape.Q[j])
        c.h1_out, c.h2_out = _ptr(tape.h1[j]), _ptr(tape.h2[j])
        ws, need = net._workspace(R)
        c.workspace, c.workspace_bytes = _lib.dptr(ws), need
        struct = self._net_struct(net)
        c.path = _lib.NET_PATH_FFMA
        if self._tc_forward:              # tcgen05 layers; h1 / h2 come back row-major for the tape
            struct = self._packed_structs[id(net)]
            c.path = _lib.NET_PATH_TCGEN05
        _lib.check(lib.l2hmc_net_run(struct, c, _lib.stream_ptr()))
        self.launches += 3
        # (cos, sin, 1) rows of this call: the t_layer input of each chain
        t0 = torch.tensor([tau[0], tau[1], 1.], dtype=torch.float32, device=self.device)
        t1 = torch.tensor([tau_alt[0], tau_alt[1], 1.], dtype=torch.float32, device=self.device)
        tape.tau1[j] = torch.where(row_set[:, None] != 0, t1[None, :], t0[None, :])

    def _conv_tape(self, key, net, R):
        """The routing tape (l2hmc_conv3d_tape) of one conv call: pooled conv-1 plane + the two pools' argmax codes."""
        if key not in self._conv_tapes:
            n = int(_lib.get_lib().l2hmc_conv3d_tape_elems(self.dyn.T, self.dyn.X, int(net.num_filters)))
            u8 = dict(dtype=torch.uint8, device=self.device)
            self._conv_tapes[key] = (self._new(R, n), torch.empty(R, n, **u8), torch.empty(R, int(net.num_features), **u8))
        p1, a1, a2 = self._conv_tapes[key]
        t = _lib.Conv3dTape()
        t.p1, t.arg1, t.arg2 = _lib.dptr(p1), _lib.dptr(a1, torch.uint8), _lib.dptr(a2, torch.uint8)
        return t

    def _conv_features(self, name, j, net, which, raw, feat, R, masks=None):
        """Conv front-end of one input stream into the tape slot `feat`, recording its routing tape (csrc/conv3d.cu)."""
        lib, dyn = _lib.get_lib(), self.dyn
        tape = self._conv_tape((name, which, j), net, R)
        m, inv, ma, inva, rs = masks if masks is not None else (None, 0, None, 0, None)
        self._call(lib.l2hmc_conv3d_features_tape, net._conv(which), _lib.dptr(raw), _lib.dptr(m), int(inv),
                   _lib.dptr(ma), int(inva), rs, R, dyn.T, dyn.X, _ptr(feat), ctypes.byref(tape))

    def _conv_backward(self, name, j, net, which, raw, g_feat, g_in, R, masks=None):
        """Adjoint of _conv_features: g_in += mask * d loss / d input; weight gradients into the stream's partial rows."""
        lib, dyn = _lib.get_lib(), self.dyn
        tape = self._conv_tape((name, which, j), net, R)
        key = (name, which)
        if key not in self._conv_partial:
            rows, cols = int(lib.l2hmc_conv3d_bwd_rows(R)), int(lib.l2hmc_conv3d_bwd_cols(int(net.num_filters)))
            self._conv_partial[key] = torch.zeros(rows, cols, dtype=torch.float32, device=self.device)
        a = _lib.Conv3dBwdArgs()
        a.w = ctypes.pointer(net._conv(which))
        a.in_, a.N, a.T, a.X = _lib.dptr(raw), R, dyn.T, dyn.X
        if masks is not None:
            m, inv, ma, inva, rs = masks
            a.mask, a.mask_invert, a.mask_alt, a.mask_invert_alt, a.row_set = _lib.dptr(m), int(inv), _lib.dptr(ma), \
                int(inva), rs
        a.g_feat, a.g_in, a.partial = _lib.dptr(g_feat), _lib.dptr(g_in), _lib.dptr(self._conv_partial[key])
        a.tape = ctypes.pointer(tape)
        self._call(lib.l2hmc_conv3d_backward, ctypes.byref(a))

    def _conv_weight_grads(self, name, net, tp):
        """Sum the per-CTA rows of the conv partials (fixed order) into the flat gradient."""
        P, F = self.params, int(net.num_filters)
        for which in ("x", "v"):
            part = self._conv_partial[(name, which)]
            rows, cols = part.shape
            off = 0
            for key, n in ((f"conv_{which}1.kernel", 18 * F), (f"conv_{which}1.bias", F),
                           (f"conv_{which}2.kernel", 16 * F * F), (f"conv_{which}2.bias", 2 * F)):
                self.gemm.col_sum(_ptr(part, off), cols, rows, n, _ptr(P.view(P.grad, f"{name}.{key}")))
                off += n

    # ------------------------------------------------------------------ native: one C call per step
    def _grads_struct(self, name, net):
        P = self.params
        g = _lib.NetGrads()
        p = lambda k: _ptr(P.view(P.grad, f"{name}.{k}"))
        g.w_in0, g.b_in0 = p("v_layer.kernel"), p("v_layer.bias")           # first input -> v_layer
        g.w_in1, g.b_in1 = p("x_layer.kernel"), p("x_layer.bias")
        g.w_t, g.b_t, g.w_h, g.b_h = p("t_layer.kernel"), p("t_layer.bias"), p("h_layer.kernel"), p("h_layer.bias")
        g.w_s, g.b_s = p("scale_layer.kernel"), p("scale_layer.bias")
        g.w_tr, g.b_tr = p("translation_layer.kernel"), p("translation_layer.bias")
        g.w_q, g.b_q = p("transformation_layer.kernel"), p("transformation_layer.bias")
        g.coeff_s, g.coeff_q = p("coeff_scale"), p("coeff_transformation")
        convs = {}
        if self.conv:
            for which in ("x", "v"):
                c = _lib.Conv3dGrads()
                c.w1, c.b1 = p(f"conv_{which}1.kernel"), p(f"conv_{which}1.bias")
                c.w2, c.b2 = p(f"conv_{which}2.kernel"), p(f"conv_{which}2.bias")
                convs[which] = c
        return g, convs

    def native(self, x, beta, draws_x, z, draws_z, weights, forward_only=False, first_bucket_event=None):
        """Loss and gradients of one step enqueued by ONE C call (csrc/train_step.cu: l2hmc_u1_train_grads) -- the same
        launches as forward() + backward(), issued from C++.  Returns the same dict keys the training loop reads."""
        dyn, lib = self.dyn, _lib.get_lib()
        N, D = x.shape
        dev = self.device
        new = lambda *s: torch.empty(*s, dtype=torch.float32, device=dev)
        a = _lib.TrainArgs()
        a.N, a.T, a.X, a.M = N, dyn.T, dyn.X, int(dyn.num_steps)
        # the step size is read by the kernels where the optimiser keeps it (flat[0]): no host copy, no sync per step
        a.eps, a.beta = 0.0, dyn._beta(beta)
        a.eps_dev = _lib.dptr(dyn._eps_dev)      # flat[0] once the arena exists; the tensor of the eps setter otherwise
        mask = dyn.mask if dyn.mask.is_contiguous() else dyn.mask.contiguous()
        a.masks = _lib.dptr(mask)
        ts = np.ascontiguousarray(dyn._ts_host, dtype=np.float32)
        a.ts = ts.ctypes.data_as(ctypes.c_void_p)
        keep = [mask, ts]
        if not dyn.hmc:
            if getattr(self, "_native_structs", None) is None:
                xs, vs = self._net_struct(dyn.position_fn), self._net_struct(dyn.momentum_fn)
                xg, xc = self._grads_struct("XNet", dyn.position_fn)
                vg, vc = self._grads_struct("VNet", dyn.momentum_fn)
                self._native_structs = (xs, vs, xg, vg, xc, vc)
            xs, vs, xg, vg, xc, vc = self._native_structs
            a.xnet, a.vnet = ctypes.pointer(xs), ctypes.pointer(vs)
            a.xnet_g, a.vnet_g = ctypes.pointer(xg), ctypes.pointer(vg)
            if self.conv:
                a.xconv_in0, a.xconv_in1 = ctypes.pointer(dyn.position_fn._conv("v")), ctypes.pointer(dyn.position_fn._conv("x"))
                a.vconv_in0, a.vconv_in1 = ctypes.pointer(dyn.momentum_fn._conv("v")), ctypes.pointer(dyn.momentum_fn._conv("x"))
                a.xconv_in0_g, a.xconv_in1_g = ctypes.pointer(xc["v"]), ctypes.pointer(xc["x"])
                a.vconv_in0_g, a.vconv_in1_g = ctypes.pointer(vc["v"]), ctypes.pointer(vc["x"])
        v_f = torch.cat([draws_x[0], draws_z[0]], dim=0)
        v_b = v_f if (draws_x[1] is draws_x[0] and draws_z[1] is draws_z[0]) else torch.cat([draws_x[1], draws_z[1]], dim=0)
        u_dir = torch.cat([draws_x[2], draws_z[2]], dim=0)
        u_acc = draws_x[3].contiguous()
        a.x, a.z, a.v_f, a.v_b = _lib.dptr(x), _lib.dptr(z), _lib.dptr(v_f), _lib.dptr(v_b)
        a.u_dir, a.u_acc = _lib.dptr(u_dir), _lib.dptr(u_acc)
        a.loss.metric, a.loss.loss_scale = _lib.METRICS[self.metric], self.loss_scale
        a.loss.aux_weight, a.loss.std_weight, a.loss.charge_weight = (float(weights[k]) for k in
                                                                      ("aux_weight", "std_weight", "charge_weight"))
        a.forward_only = int(bool(forward_only))
        out = dict(N=N, D=D, loss=new(1), per_chain=new(N), x_out=new(N, D), px=new(N), pz=new(N), x_prop=new(N, D),
                   x_dq=torch.empty(N, dtype=torch.int32, device=dev), accepted=torch.empty(N, dtype=torch.uint8, device=dev))
        a.loss_out, a.per_chain, a.x_out, a.px, a.pz, a.x_prop = (_lib.dptr(out[k]) for k in
                                                                  ("loss", "per_chain", "x_out", "px", "pz", "x_prop"))
        a.x_dq, a.accepted = _lib.dptr(out["x_dq"], torch.int32), _lib.dptr(out["accepted"], torch.uint8)
        P = self.params
        a.g_eps = _ptr(P.view(P.grad, "eps"))
        a.nonfinite_flag = _lib.dptr(dyn._flag, torch.int32)
        need = int(lib.l2hmc_u1_train_workspace_bytes(ctypes.byref(a)))
        if getattr(self, "_native_ws", None) is None or self._native_ws.numel() < need:
            self._native_ws = torch.empty(need, dtype=torch.uint8, device=dev)
        a.workspace, a.workspace_bytes = ctypes.c_void_p(self._native_ws.data_ptr()), need
        if first_bucket_event is not None:     # a torch.cuda.Event that has been recorded once (its handle exists)
            a.first_bucket_event = ctypes.c_void_p(int(first_bucket_event.cuda_event))
        _lib.check(lib.l2hmc_u1_train_grads(ctypes.byref(a), _lib.stream_ptr()))
        del keep
        return out

    # ------------------------------------------------------------------ forward with tape
    def forward(self, x, beta, draws_x, z, draws_z, weights):
        """Both chains' transitions + the loss.  draws_* = (v_f, v_b, u_dir, u_acc)."""
        dyn, lib = self.dyn, _lib.get_lib()
        N, D = x.shape
        R, M = 2 * N, int(dyn.num_steps)
        beta, eps = dyn._beta(beta), dyn._eps()
        nets = not dyn.hmc
        X0 = torch.cat([x, z], dim=0)
        v_f = torch.cat([draws_x[0], draws_z[0]], dim=0)
        v_b = torch.cat([draws_x[1], draws_z[1]], dim=0)
        u_dir = torch.cat([draws_x[2], draws_z[2]], dim=0).contiguous()
        row_set = (u_dir <= 0.5).to(torch.uint8)                    # 1: the chain integrates backward
        ones = torch.ones(R, dtype=torch.float32, device=self.device)
        v0 = self._new(R, D)                                        # v_f where u_dir > 0.5 else v_b (exact 0/1 blend)
        self._call(lib.l2hmc_mh_select, _lib.dptr(v_f), _lib.dptr(v_f), _lib.dptr(v_b), None, None, _lib.dptr(ones),
                   _lib.dptr(ones), _lib.dptr(u_dir), None, R, D, 0, _lib.dptr(v0), None, None, None, None)
        tapes = {n: self._tape(n, R) for n in ("XNet", "VNet")} if nets else {}
        self._tc_forward = nets and self._pack_for_forward(R)
        sld = torch.zeros(R, dtype=torch.float32, device=self.device)
        flag = dyn._flag
        calls, jx, jv = [], 0, 0
        xs, vs = X0, v0
        rs = _lib.dptr(row_set, torch.uint8)
        for step in range(M):
            ib = M - step - 1
            tau, tau_alt = dyn._ts_host[step], dyn._ts_host[ib]
            row_f, row_b = dyn.mask[step], dyn.mask[ib]
            for kind, inv, inv_alt in (("m", 0, 0), ("p", 0, 1), ("p", 1, 0), ("m", 0, 0)):
                if kind == "m":
                    rec = dict(kind="m", x=xs, v=vs, j=jv)
                    if nets and not self.conv:
                        tp = tapes["VNet"]
                        tp.in0[jv].copy_(xs)
                        g = tp.in1[jv]
                    else:
                        g = self._new(R, D)
                    self._call(lib.l2hmc_u1_action_force, _lib.dptr(xs), beta, R, dyn.T, dyn.X, None, _ptr(g))
                    rec["g"] = g
                    S = Tr = Q = None
                    if nets:
                        tp = tapes["VNet"]
                        if self.conv:        # momentum_fn([x, grad, t]): x through conv_v*, the force through conv_x*
                            self._conv_features("VNet", jv, dyn.momentum_fn, "v", xs, tp.in0[jv], R)
                            self._conv_features("VNet", jv, dyn.momentum_fn, "x", g, tp.in1[jv], R)
                        self._net_forward(dyn.momentum_fn, tp, jv, R, tau, tau_alt, row_set)
                        S, Tr, Q = _ptr(tp.S[jv]), _ptr(tp.T[jv]), _ptr(tp.Q[jv])
                        jv += 1
                    out = self._new(R, D)
                    self._call(lib.l2hmc_train_update_fwd, 0, None, _lib.dptr(vs), _ptr(g), S, Tr, Q, None, 0, None, 0,
                               rs, eps, 1, -1, R, D, _lib.dptr(out), _lib.dptr(sld), _lib.dptr(flag, torch.int32))
                    vs = out
                else:
                    rec = dict(kind="p", x=xs, v=vs, j=jx, mask=row_f, inv=inv, mask_alt=row_b, inv_alt=inv_alt)
                    S = Tr = Q = None
                    if nets and self.conv:       # position_fn([v, mask * x, t]): v through conv_v*, m*x through conv_x*
                        tp = tapes["XNet"]
                        self._conv_features("XNet", jx, dyn.position_fn, "v", vs, tp.in0[jx], R)
                        self._conv_features("XNet", jx, dyn.position_fn, "x", xs, tp.in1[jx], R, (row_f, inv, row_b, inv_alt, rs))
                    elif nets:
                        tp = tapes["XNet"]
                        tp.in0[jx].copy_(vs)
                        self._call(lib.l2hmc_mask_mul_sets, _lib.dptr(xs), _lib.dptr(row_f), inv, _lib.dptr(row_b),
                                   inv_alt, rs, R, D, _ptr(tp.in1[jx]))
                    if nets:
                        self._net_forward(dyn.position_fn, tp, jx, R, tau, tau_alt, row_set)
                        S, Tr, Q = _ptr(tp.S[jx]), _ptr(tp.T[jx]), _ptr(tp.Q[jx])
                        jx += 1
                    out = self._new(R, D)
                    self._call(lib.l2hmc_train_update_fwd, 1, _lib.dptr(xs), _lib.dptr(vs), None, S, Tr, Q,
                               _lib.dptr(row_f), inv, _lib.dptr(row_b), inv_alt, rs, eps, 1, -1, R, D, _lib.dptr(out),
                               _lib.dptr(sld), _lib.dptr(flag, torch.int32))
                    xs = out
                calls.append(rec)
        h_old, h_new, p = self._new(R), self._new(R), self._new(R)
        self._call(lib.l2hmc_u1_hamiltonian, _lib.dptr(X0), _lib.dptr(v0), beta, R, dyn.T, dyn.X, _lib.dptr(h_old))
        self._call(lib.l2hmc_u1_hamiltonian, _lib.dptr(xs), _lib.dptr(vs), beta, R, dyn.T, dyn.X, _lib.dptr(h_new))
        self._call(lib.l2hmc_accept_prob, _lib.dptr(h_old), _lib.dptr(h_new), _lib.dptr(sld), R, _lib.dptr(p),
                   _lib.dptr(flag, torch.int32))
        x_prop, px, pz = xs[:N], p[:N], p[N:]
        x_out = self._new(N, D)
        acc = torch.empty(N, dtype=torch.uint8, device=self.device)
        u_acc = draws_x[3].contiguous()
        self._call(lib.l2hmc_mh_select, _lib.dptr(x), _lib.dptr(x_prop), None, None, None, _lib.dptr(px), None, None,
                   _lib.dptr(u_acc), N, D, 0, None, None, None, _lib.dptr(x_out), _lib.dptr(acc, torch.uint8))
        prm = _lib.LossParams()
        prm.metric, prm.loss_scale = _lib.METRICS[self.metric], self.loss_scale
        prm.aux_weight, prm.std_weight, prm.charge_weight = (float(weights[k]) for k in
                                                             ("aux_weight", "std_weight", "charge_weight"))
        per_chain = self._new(N)
        x_dq = torch.empty(N, dtype=torch.int32, device=self.device)
        loss = self._new(1)
        self._call(lib.l2hmc_u1_loss, ctypes.byref(prm), _lib.dptr(x), _ptr(x_prop), _lib.dptr(z), _lib.dptr(x_out),
                   _ptr(px), _ptr(pz), N, dyn.T, dyn.X, _lib.dptr(per_chain), _lib.dptr(x_dq, torch.int32),
                   _lib.dptr(loss))
        self.launches += 1
        return dict(N=N, D=D, R=R, beta=beta, eps=eps, x=x, z=z, xM=xs, vM=vs, h_old=h_old, h_new=h_new, sld=sld, p=p,
                    x_prop=x_prop, px=px, pz=pz, x_out=x_out, accepted=acc, x_dq=x_dq, loss=loss, per_chain=per_chain,
                    calls=calls, tapes=tapes, row_set=row_set, prm=prm)

    # ------------------------------------------------------------------ backward sweep
    def _net_backward(self, net, tp, j, R):
        """dA[j] -> dz2[j], dz1[j] (activation gradients of one call; weight gradients are deferred)."""
        w, H, D = net.weights, tp.H, tp.D
        dA = tp.dA[j]
        segs = [(_ptr(dA, s * D), 3 * D, 1, _lib.dptr(w[k]), 1, D, D)
                for s, k in enumerate(("scale_layer.kernel", "translation_layer.kernel", "transformation_layer.kernel"))]
        self.gemm(R, H, segs, _ptr(tp.dz2[j]), H, relu_gate=_ptr(tp.h2[j]), ldg=H)
        self.gemm(R, H, [(_ptr(tp.dz2[j]), H, 1, _lib.dptr(w["h_layer.kernel"]), 1, H, H)], _ptr(tp.dz1[j]), H,
                  relu_gate=_ptr(tp.h1[j]), ldg=H)

    def _feat_grads(self, net, tp, dz1, R):
        """d loss / d conv features of both streams: dz1 . W_in0^T and dz1 . W_in1^T."""
        H, K = tp.H, tp.K
        if getattr(self, "_fg", None) is None or self._fg[0].shape != (R, K):
            self._fg = (self._new(R, K), self._new(R, K))
        for buf, layer in zip(self._fg, ("v_layer.kernel", "x_layer.kernel")):
            self.gemm(R, K, [(dz1, H, 1, _lib.dptr(net.weights[layer]), 1, H, H)], _lib.dptr(buf), K)
        return self._fg

    def backward(self, fw):
        """d loss / d (eps, weights) into self.params.grad; returns its dict of views."""
        dyn, lib = self.dyn, _lib.get_lib()
        for part in self._conv_partial.values():
            part.zero_()
        N, D, R, beta, eps = fw["N"], fw["D"], fw["R"], fw["beta"], fw["eps"]
        nets = not dyn.hmc
        rs = _lib.dptr(fw["row_set"], torch.uint8)
        gx_post = torch.zeros(R, D, dtype=torch.float32, device=self.device)   # z rows: z_prop is not in the loss
        g_p = self._new(R)
        self._call(lib.l2hmc_u1_loss_bwd, ctypes.byref(fw["prm"]), _lib.dptr(fw["x"]), _ptr(fw["x_prop"]),
                   _lib.dptr(fw["z"]), _ptr(fw["px"]), _ptr(fw["pz"]), N, dyn.T, dyn.X, _ptr(gx_post), _ptr(g_p),
                   _ptr(g_p, N))
        a = self._new(R)
        self._call(lib.l2hmc_accept_prob_bwd, _lib.dptr(fw["h_old"]), _lib.dptr(fw["h_new"]), _lib.dptr(fw["sld"]),
                   _lib.dptr(g_p), R, _lib.dptr(a))
        force = self._new(R, D)
        self._call(lib.l2hmc_u1_action_force, _lib.dptr(fw["xM"]), beta, R, dyn.T, dyn.X, None, _lib.dptr(force))
        gx, gv = self._new(R, D), self._new(R, D)
        self._call(lib.l2hmc_traj_bwd_init, _lib.dptr(a), _lib.dptr(force), _lib.dptr(fw["vM"]), _lib.dptr(gx_post), R,
                   D, _lib.dptr(gx), _lib.dptr(gv))
        deps = torch.zeros(R, dtype=torch.float32, device=self.device)
        dg = self._new(R, D)
        for rec in reversed(fw["calls"]):
            j = rec["j"]
            u = _lib.UpdateBwdArgs()
            u.N, u.D, u.eps, u.dir, u.dir_alt, u.row_set = R, D, eps, 1, -1, rs
            u.v, u.gv, u.gld, u.deps = _lib.dptr(rec["v"]), _lib.dptr(gv), _lib.dptr(a), _lib.dptr(deps)
            net = tp = None
            if nets:
                net = dyn.momentum_fn if rec["kind"] == "m" else dyn.position_fn
                tp = fw["tapes"]["VNet" if rec["kind"] == "m" else "XNet"]
                u.S, u.T, u.Q = _ptr(tp.S[j]), _ptr(tp.T[j]), _ptr(tp.Q[j])
                u.coeff_s, u.coeff_q = _lib.dptr(net.weights["coeff_scale"]), _lib.dptr(net.weights["coeff_transformation"])
                u.dA, u.ps, u.pq = _ptr(tp.dA[j]), _ptr(tp.ps[j]), _ptr(tp.pq[j])
            if rec["kind"] == "m":
                u.position, u.g, u.dg = 0, _ptr(rec["g"]), _lib.dptr(dg)
                self._call(lib.l2hmc_update_bwd, ctypes.byref(u))
                if nets:
                    self._net_backward(net, tp, j, R)
                    H, K = tp.H, tp.K
                    dz1 = _ptr(tp.dz1[j])
                    # momentum_fn([x, grad, t]): first input x, second the force
                    if self.conv:
                        f0, f1 = self._feat_grads(net, tp, dz1, R)
                        self._conv_backward("VNet", j, net, "v", rec["x"], f0, gx, R)
                        self._conv_backward("VNet", j, net, "x", rec["g"], f1, dg, R)
                    else:
                        self.gemm(R, K, [(dz1, H, 1, _lib.dptr(net.weights["v_layer.kernel"]), 1, H, H)], _lib.dptr(gx),
                                  D, accumulate=True)
                        self.gemm(R, K, [(dz1, H, 1, _lib.dptr(net.weights["x_layer.kernel"]), 1, H, H)], _lib.dptr(dg),
                                  D, accumulate=True)
                self._call(lib.l2hmc_u1_force_hvp, _lib.dptr(rec["x"]), _lib.dptr(dg), beta, R, dyn.T, dyn.X,
                           _lib.dptr(gx))
            else:
                u.position, u.x, u.gx = 1, _lib.dptr(rec["x"]), _lib.dptr(gx)
                u.mask, u.invert = _lib.dptr(rec["mask"]), int(rec["inv"])
                u.mask_alt, u.invert_alt = _lib.dptr(rec["mask_alt"]), int(rec["inv_alt"])
                self._call(lib.l2hmc_update_bwd, ctypes.byref(u))
                if nets:
                    self._net_backward(net, tp, j, R)
                    H, K = tp.H, tp.K
                    dz1 = _ptr(tp.dz1[j])
                    # position_fn([v, mask * x, t]): first input v, second the masked links
                    if self.conv:
                        f0, f1 = self._feat_grads(net, tp, dz1, R)
                        self._conv_backward("XNet", j, net, "v", rec["v"], f0, gv, R)
                        self._conv_backward("XNet", j, net, "x", rec["x"], f1, gx, R,
                                            (rec["mask"], rec["inv"], rec["mask_alt"], rec["inv_alt"], rs))
                    else:
                        self.gemm(R, K, [(dz1, H, 1, _lib.dptr(net.weights["v_layer.kernel"]), 1, H, H)], _lib.dptr(gv),
                                  D, accumulate=True)
                        self.gemm(R, K, [(dz1, H, 1, _lib.dptr(net.weights["x_layer.kernel"]), 1, H, H)], _lib.dptr(gx),
                                  D, accumulate=True, col_scale=_lib.dptr(rec["mask"]), invert=rec["inv"], row_set=rs,
                                  col_scale_alt=_lib.dptr(rec["mask_alt"]), invert_alt=rec["inv_alt"])
        # ---- parameter gradients
        P = self.params
        P.grad.zero_()
        self._call(lib.l2hmc_reduce_sum, _lib.dptr(deps), R, 0, _ptr(P.view(P.grad, "eps")), _lib.dptr(self._scratch))
        self.launches += 1
        if nets:
            for name, net in (("XNet", dyn.position_fn), ("VNet", dyn.momentum_fn)):
                self._weight_grads(name, fw["tapes"][name])
                if self.conv:
                    self._conv_weight_grads(name, net, fw["tapes"][name])
        return P.grads_dict()

    def _ksplit(self, M, N, K):
        tiles = -(-M // 64) * -(-N // 64)
        return int(max(1, min(64, -(-592 // tiles), K // 64)))      # the same rule as csrc/train_step.cu (bit-identical walks)

    def _weight_grads(self, name, tp):
        """One transposed GEMM over the whole tape per weight tensor (K = calls * R rows); biases and coeff_* are
        column sums of the activation gradients."""
        P, H, D, Kin = self.params, tp.H, tp.D, tp.K
        Rt = tp.calls * tp.R
        g = lambda key: _ptr(P.view(P.grad, f"{name}.{key}"))
        dz2, dz1 = _ptr(tp.dz2), _ptr(tp.dz1)
        for s, layer in enumerate(("scale_layer", "translation_layer", "transformation_layer")):
            B = (_ptr(tp.dA, s * D), 3 * D, 1)
            self.gemm(H, D, [(_ptr(tp.h2), 1, H) + B + (Rt,)], g(layer + ".kernel"), D, ksplit=self._ksplit(H, D, Rt))
            self.gemm.col_sum(_ptr(tp.dA, s * D), 3 * D, Rt, D, g(layer + ".bias"))
        self.gemm(H, H, [(_ptr(tp.h1), 1, H, dz2, H, 1, Rt)], g("h_layer.kernel"), H, ksplit=self._ksplit(H, H, Rt))
        self.gemm.col_sum(dz2, H, Rt, H, g("h_layer.bias"))
        self.gemm(Kin, H, [(_ptr(tp.in0), 1, Kin, dz1, H, 1, Rt)], g("v_layer.kernel"), H,
                  ksplit=self._ksplit(Kin, H, Rt))
        self.gemm(Kin, H, [(_ptr(tp.in1), 1, Kin, dz1, H, 1, Rt)], g("x_layer.kernel"), H,
                  ksplit=self._ksplit(Kin, H, Rt))
        self.gemm(2, H, [(_ptr(tp.tau1), 1, 3, dz1, H, 1, Rt)], g("t_layer.kernel"), H, ksplit=self._ksplit(2, H, Rt))
        self.gemm.col_sum(dz1, H, Rt, H, g("v_layer.bias"))          # the three first-layer biases share one sum
        for layer in ("x_layer", "t_layer"):
            P.view(P.grad, f"{name}.{layer}.bias").copy_(P.view(P.grad, f"{name}.v_layer.bias"))
        self.gemm.col_sum(_ptr(tp.ps), D, Rt, D, g("coeff_scale"))
        self.gemm.col_sum(_ptr(tp.pq), D, Rt, D, g("coeff_transformation"))


def exponential_decay(lr_init, global_step, decay_steps, decay_rate, staircase=True):
    """tf.train.exponential_decay (gauge_model.py:925-936)."""
    e = global_step / float(decay_steps)
    if staircase:
        e = math.floor(e)
    return lr_init * decay_rate ** e


class Adam(object):
    """tf.train.AdamOptimizer over the flat arena (gauge_model.py:937-945): beta1 0.9, beta2 0.999, epsilon 1e-8;
    clip_norm > 0 applies tf.clip_by_global_norm (:826-827) inside the same launch."""

    def __init__(self, params, beta1=0.9, beta2=0.999, epsilon=1e-8):
        self.params = params
        self.beta1, self.beta2, self.epsilon = beta1, beta2, epsilon
        self.m = torch.zeros_like(params.flat)
        self.v = torch.zeros_like(params.flat)
        self.t = 0
        self._sumsq = torch.zeros(1, dtype=torch.float32, device=params.flat.device)
        self._scratch = torch.empty(1024, dtype=torch.float32, device=params.flat.device)

    def clip_by_global_norm(self, clip_norm):
        """tf.clip_by_global_norm on the gradient arena, in place (gauge_model.py:826-827: inside
        _calc_loss_and_grads, so under Horovod every rank clips its own gradient BEFORE the average)."""
        P, lib = self.params, _lib.get_lib()
        _lib.check(lib.l2hmc_reduce_sum(_lib.dptr(P.grad), P.size, 1, _lib.dptr(self._sumsq),
                                        _lib.dptr(self._scratch), _lib.stream_ptr()))
        _lib.check(lib.l2hmc_clip_by_global_norm(_lib.dptr(P.grad), P.size, _lib.dptr(self._sumsq), float(clip_norm),
                                                 _lib.stream_ptr()))

    def step(self, lr, grad_scale=1., clip_norm=0.):
        P, lib = self.params, _lib.get_lib()
        self.t += 1
        lr_t = lr * math.sqrt(1. - self.beta2 ** self.t) / (1. - self.beta1 ** self.t)
        sumsq = None
        if clip_norm and clip_norm > 0:
            _lib.check(lib.l2hmc_reduce_sum(_lib.dptr(P.grad), P.size, 1, _lib.dptr(self._sumsq),
                                            _lib.dptr(self._scratch), _lib.stream_ptr()))
            sumsq = _lib.dptr(self._sumsq)
        _lib.check(lib.l2hmc_adam_step(_lib.dptr(P.flat), _lib.dptr(P.grad), _lib.dptr(self.m), _lib.dptr(self.v), P.size,
                                       float(lr_t), float(self.beta1), float(self.beta2), float(self.epsilon),
                                       float(grad_scale), sumsq, float(clip_norm or 0.), _lib.stream_ptr()))
        P.sync_to_dynamics()
