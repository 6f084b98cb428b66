"""GaugeSampler: the reference's inference step with HOST buffers.

Mirrors what one iteration of GaugeModel.run does around the dynamics call
(reference l2hmc/gauge_model.py:1379-1388): feed `samples` from the host, fetch
[x_out, px, actions, plaqs, charges, charge_diff], then `samples = np.mod(x_out, 2 pi)`.
Here the mod and the observables are fused into the one-launch MC-step kernel, and
the host<->device copies are pipelined against compute in chain chunks over three CUDA
streams (chains are independent, so chunking changes nothing numerically: the Philox stream
is keyed by the global chain index).
"""
import numpy as np
import torch

from . import _lib


class StepResult(object):
    __slots__ = ("samples", "px", "actions", "plaqs", "charges", "charge_diff")


class GaugeSampler(object):
    def __init__(self, dynamics, num_chains, chunk_chains=8192):
        _lib.require_cuda()
        if not dynamics._fused_ok():
            raise NotImplementedError("GaugeSampler drives the one-launch HMC path; use "
                                      "GaugeDynamics.apply_transition for L2HMC nets / other shapes")
        self.dyn = dynamics
        self.lat = dynamics.lattice
        self.N, self.D = int(num_chains), dynamics.x_dim
        self.chunk = int(min(chunk_chains, self.N))
        self.dev = dynamics.device
        nchunks = (self.N + self.chunk - 1) // self.chunk
        self.nbuf = min(3, nchunks)
        f32 = dict(dtype=torch.float32, device=self.dev)
        self.d_in = [torch.empty(self.chunk, self.D, **f32) for _ in range(self.nbuf)]
        self.d_out = [torch.empty(self.chunk, self.D, **f32) for _ in range(self.nbuf)]
        # per-chain scalars for the whole batch stay on the device until the end of the step
        self.d_px = torch.empty(self.N, **f32)
        self.d_act = torch.empty(self.N, **f32)
        self.d_plaq = torch.empty(self.N, **f32)
        self.d_q = torch.empty(self.N, dtype=torch.int32, device=self.dev)
        self.d_q_prev = torch.zeros(self.N, dtype=torch.int32, device=self.dev)
        self.h_out = torch.empty(self.N, self.D, dtype=torch.float32).pin_memory()
        self.h_scal = torch.empty(3, self.N, dtype=torch.float32).pin_memory()
        self.h_q = torch.empty(2, self.N, dtype=torch.int32).pin_memory()
        self.streams = [torch.cuda.Stream(device=self.dev) for _ in range(self.nbuf)]
        self.h2d_bytes = self.N * self.D * 4
        self.d2h_bytes = self.N * self.D * 4 + 3 * self.N * 4 + 2 * self.N * 4
        self.launches_per_step = nchunks

    def pinned_like(self, samples):
        """A pinned host copy of `samples` (N, D) to feed step()."""
        h = torch.empty(self.N, self.D, dtype=torch.float32).pin_memory()
        h.copy_(torch.as_tensor(np.asarray(samples, dtype=np.float32)).reshape(self.N, self.D))
        return h

    def step(self, samples_host, beta):
        """One MC step for all chains.  `samples_host`: pinned (N, D) fp32 host tensor.
        Returns StepResult of HOST tensors (views of this sampler's pinned buffers)."""
        dyn, lib = self.dyn, _lib.get_lib()
        main = torch.cuda.current_stream(self.dev)
        base_offset = dyn.chain_offset
        for i, lo in enumerate(range(0, self.N, self.chunk)):
            hi = min(lo + self.chunk, self.N)
            n = hi - lo
            b = i % self.nbuf
            st = self.streams[b]
            st.wait_stream(main)
            with torch.cuda.stream(st):
                xin, xout = self.d_in[b][:n], self.d_out[b][:n]
                xin.copy_(samples_host[lo:hi], non_blocking=True)
                a = _lib.HmcMcArgs()
                a.x, a.x_out = _lib.dptr(xin), _lib.dptr(xout)
                a.beta, a.eps, a.num_steps = dyn._beta(beta), dyn._eps(), int(dyn.num_steps)
                a.default_dir, a.wrap_2pi = 1, 1
                a.seed, a.mc_step, a.chain0 = int(dyn.seed), int(dyn._mc_step), int(base_offset + lo)
                a.N, a.T, a.X = n, dyn.T, dyn.X
                a.accept_prob = _lib.dptr(self.d_px[lo:hi])
                a.nonfinite_flag = _lib.dptr(dyn._flag, torch.int32)
                a.obs_action, a.obs_plaq = _lib.dptr(self.d_act[lo:hi]), _lib.dptr(self.d_plaq[lo:hi])
                a.obs_charge = _lib.dptr(self.d_q[lo:hi], torch.int32)
                _lib.check(lib.l2hmc_u1_hmc_mc_step(a, _lib.stream_ptr()))
                self.h_out[lo:hi].copy_(xout, non_blocking=True)
        for st in self.streams:
            main.wait_stream(st)
        self.h_scal[0].copy_(self.d_px, non_blocking=True)
        self.h_scal[1].copy_(self.d_act, non_blocking=True)
        self.h_scal[2].copy_(self.d_plaq, non_blocking=True)
        self.h_q[0].copy_(self.d_q, non_blocking=True)
        self.h_q[1].copy_(self.d_q_prev, non_blocking=True)
        self.d_q_prev.copy_(self.d_q)
        main.synchronize()
        dyn._mc_step += 1
        r = StepResult()
        r.samples, r.px, r.actions, r.plaqs = self.h_out, self.h_scal[0], self.h_scal[1], self.h_scal[2]
        r.charges = self.h_q[0]
        r.charge_diff = float((self.h_q[0] - self.h_q[1]).abs().sum()) / self.N
        return r
