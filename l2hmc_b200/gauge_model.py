"""Host-side mirror of the parts of the reference's GaugeModel that sit next to the hot path
(SURVEY section 8(f)): the training loss VALUE, the observable ops and the run-loop statistics.

Mirrors, with the same names and argument meaning (reference l2hmc/gauge_model.py):
  :585-613    _create_lattice / _create_dynamics (from the params dict of globals.py:22-47)
  :631-657    _create_metric_fn  (here: the metric's name selects the kernel's formula)
  :683-725    _calc_total_actions / _calc_avg_plaqs / _calc_top_charges(fft) / _calc_top_charges_diff
  :728-803    _calc_loss(x, beta, **weights) -> (loss, x_out, px, x_dq)
  :1039-1046  update_beta (1/beta-linear annealing)
  :1473-1533  calc_observables_stats(observables, therm_frac)
  :805-832    _calc_loss_and_grads(x, beta, **weights) -> (loss, grads, x_out, px, x_dq)   [training.py, csrc/train.cu]
  :925-945    learning-rate schedule + Adam (train_op)   :1155-1190  one iteration of the training loop (train_step)
  :519-556    saved state (save_state / restore_state)
All arithmetic runs in csrc/loss.cu, csrc/stats.cu, csrc/train.cu and the path's own kernels; there is no CPU fallback.
The backward pass covers network_arch='generic', 'conv3D' (csrc/conv3d_bwd.cu) and hmc=True.
"""
import collections
import ctypes

import numpy as np
import torch

from . import _lib
from .dynamics import GaugeDynamics
from .lattice import GaugeLattice, to_device

PARAMS = {  # globals.py:22-47 (the keys this mirror reads)
    'time_size': 8, 'space_size': 8, 'link_type': 'U1', 'dim': 2, 'num_samples': 128, 'rand': False,
    'num_steps': 5, 'eps': 0.25, 'loss_scale': 1., 'beta_init': 2., 'beta_final': 4., 'train_steps': 10000,
    'hmc': False, 'network_arch': 'generic', 'metric': 'cos_diff', 'std_weight': 1., 'aux_weight': 1.,
    'charge_weight': 1., 'annealing': True, 'data_format': 'channels_last',
    'lr_init': 1e-3, 'lr_decay_steps': 1000, 'lr_decay_rate': 0.96, 'clip_grads': False, 'clip_value': 1.,
}


class RunStats(object):
    """Streaming calc_observables_stats (gauge_model.py:1473-1533): fold each MC step's (N,) observables into
    device accumulators; nothing but the final statistics leaves the GPU.  `num_steps` is the planned run length
    (the thermalisation cut is its first num_steps // therm_frac steps)."""

    def __init__(self, num_chains, num_steps, therm_frac=10, qmax=64, device="cuda"):
        _lib.require_cuda()
        self.N, self.num_steps, self.therm_steps = int(num_chains), int(num_steps), int(num_steps) // int(therm_frac)
        self.device, self.qmax, self.step = torch.device(device), int(qmax), 0
        self.sum = torch.zeros(4, self.N, dtype=torch.float64, device=self.device)
        self.sumsq = torch.zeros(4, self.N, dtype=torch.float64, device=self.device)
        self.hist = torch.zeros(2 * self.qmax + 2, dtype=torch.int64, device=self.device)
        s = _lib.RunStats()
        s.N, s.qmax = self.N, self.qmax
        s.sum, s.sumsq = _lib.dptr(self.sum, torch.float64), _lib.dptr(self.sumsq, torch.float64)
        s.hist = _lib.dptr(self.hist, torch.int64)        # non-negative counts: same bits as unsigned long long
        self._s = s

    def update(self, actions, plaqs, charges):
        """One MC step's observables, (N,) each (device tensors or arrays; charges float or int32)."""
        a, p = to_device(actions, self.device).reshape(-1), to_device(plaqs, self.device).reshape(-1)
        q = charges if isinstance(charges, torch.Tensor) else torch.as_tensor(np.asarray(charges))
        q = q.to(device=self.device, dtype=torch.float32).reshape(-1).contiguous()
        counted = int(self.step >= self.therm_steps)
        with torch.cuda.device(self.device):
            _lib.check(_lib.get_lib().l2hmc_stats_update(ctypes.byref(self._s), _lib.dptr(a), _lib.dptr(p), _lib.dptr(q),
                                                         counted, _lib.stream_ptr()))
        self.step += 1

    def finalize(self):
        """((mean, err) of actions, plaqs, charges, suscept -- (N,) fp64 host arrays each -- , charge_probabilities)."""
        counted = self.step - min(self.step, self.therm_steps)
        mean = torch.empty(4, self.N, dtype=torch.float64, device=self.device)
        err = torch.empty_like(mean)
        with torch.cuda.device(self.device):
            _lib.check(_lib.get_lib().l2hmc_stats_finalize(ctypes.byref(self._s), counted, self.step,
                                                           _lib.dptr(mean, torch.float64), _lib.dptr(err, torch.float64),
                                                           _lib.stream_ptr()))
        mean, err, hist = mean.cpu().numpy(), err.cpu().numpy(), self.hist.cpu().numpy()
        if hist[-1]:
            raise OverflowError(f"{int(hist[-1])} charges outside [-{self.qmax}, {self.qmax}]: raise qmax")
        total = hist[:-1].sum()
        probs = collections.OrderedDict((int(i) - self.qmax, hist[i] / total) for i in np.nonzero(hist[:-1])[0])
        return tuple((mean[i], err[i]) for i in range(4)) + (probs,)


class GaugeModel(object):
    def __init__(self, params=None, lattice=None, dynamics=None, device="cuda"):
        _lib.require_cuda()
        self.params = dict(PARAMS)
        self.params.update(params or {})
        for key, val in self.params.items():
            setattr(self, key, val)
        self.device = torch.device(device)
        self.lattice = lattice if lattice is not None else self._create_lattice()
        self.batch_size = self.num_samples = self.lattice.samples.shape[0] if lattice is not None else self.num_samples
        self.x_dim = self.lattice.num_links
        self.dynamics = dynamics if dynamics is not None else self._create_dynamics(self.lattice)
        self.loss_weights = {k: getattr(self, k) for k in ('std_weight', 'aux_weight', 'charge_weight')}
        self.metric_fn = self._create_metric_fn(self.metric)
        self.beta = self.beta_init
        self._z_step = 0
        self._engine = self._optimizer = None
        self.global_step = 0
        self.native_train = True        # loss + gradients of a step in one C call (csrc/train_step.cu)
        self.using_hvd = False          # set by train(): lr * world size, gradients averaged (gauge_model.py:939-941)
        self.overlap_allreduce = True   # average the first gradient bucket under the tail of the backward pass

    # ------------------------------------------------------------------ construction (:585-613)
    def _create_lattice(self):
        return GaugeLattice(time_size=self.time_size, space_size=self.space_size, dim=self.dim,
                            link_type=self.link_type, num_samples=self.num_samples, rand=self.rand)

    def _create_dynamics(self, lattice, **kwargs):
        kw = dict(num_steps=self.num_steps, eps=self.eps, network_arch=self.network_arch, hmc=self.hmc,
                  eps_trainable=True, data_format=self.data_format, device=self.device)
        kw.update(kwargs)
        return GaugeDynamics(lattice=lattice, potential_fn=lattice.get_energy_function(), **kw)

    @staticmethod
    def _create_metric_fn(metric):
        """(:631-657) The returned object only names the formula; the kernel evaluates it."""
        if metric not in _lib.METRICS:
            raise AttributeError(f"metric={metric}. Expected one of: 'l1', 'l2', 'cos', 'cos2', or 'cos_diff'.")
        return metric

    def update_beta(self, step):
        """(:1039-1046) 1/beta-linear annealing from beta_init to beta_final over train_steps."""
        temp = ((1. / self.beta_init - 1. / self.beta_final) * (1. - step / float(self.train_steps))
                + 1. / self.beta_final)
        return 1. / temp

    # ------------------------------------------------------------------ observable ops (:683-725)
    def _x(self, x):
        return to_device(x, self.device).reshape(-1, self.x_dim)

    def _calc_total_actions(self, x):
        return self.lattice.observables(self._x(x))[0]

    def _calc_avg_plaqs(self, x):
        return self.lattice.observables(self._x(x))[1]

    def _calc_top_charges(self, x, fft=False):
        x = self._x(x)
        if not fft:
            return self.lattice.observables(x)[2]
        q = torch.empty(x.shape[0], dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(_lib.get_lib().l2hmc_u1_top_charges_fft(_lib.dptr(x), x.shape[0], self.lattice.time_size,
                                                               self.lattice.space_size, _lib.dptr(q), _lib.stream_ptr()))
        return q

    # ------------------------------------------------------------------ loss value (:728-803)
    def _calc_loss(self, x, beta, draws_x=None, z=None, draws_z=None, **weights):
        """(loss, x_out, px, x_dq).  draws_x / draws_z = (v_f, v_b, u_dir, u_acc) and `z` inject the reference's
        in-graph draws (:754-758 over gauge_dynamics.py:211-248); all-None runs on the Philox stream."""
        x = self._x(x)
        N, D = x.shape
        w = dict(self.loss_weights)
        w.update(weights)
        dyn = self.dynamics
        # the same draws, from the same counters, as the training path (_calc_loss_and_grads)
        draws_x, z, draws_z = self._draws(x, draws_x, z, draws_z)
        inj = lambda d: dict(momentum_f=d[0], momentum_b=d[1], direction_uniform=d[2], accept_uniform=d[3])
        x_prop, _, px, x_out = dyn(x, beta, **inj(draws_x))
        _, _, pz, _ = dyn(z, beta, **inj(draws_z))
        prm = _lib.LossParams()
        prm.metric, prm.loss_scale = _lib.METRICS[self.metric_fn], float(self.loss_scale)
        prm.aux_weight, prm.std_weight, prm.charge_weight = (float(w[k]) for k in ('aux_weight', 'std_weight',
                                                                                   'charge_weight'))
        per_chain = torch.empty(N, dtype=torch.float32, device=x.device)
        x_dq = torch.empty(N, dtype=torch.int32, device=x.device)
        loss = torch.empty(1, dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(_lib.get_lib().l2hmc_u1_loss(
                ctypes.byref(prm), _lib.dptr(x), _lib.dptr(x_prop), _lib.dptr(z), _lib.dptr(x_out), _lib.dptr(px),
                _lib.dptr(pz), N, self.lattice.time_size, self.lattice.space_size, _lib.dptr(per_chain),
                _lib.dptr(x_dq, torch.int32), _lib.dptr(loss), _lib.stream_ptr()))
        self.last_per_chain_loss = per_chain
        return loss[0], x_out, px, x_dq

    # ------------------------------------------------------------------ backward pass + optimiser (:805-832, :925-945)
    @property
    def engine(self):
        """Tape / backward engine; created on first use (it re-points the nets' weights into one flat arena, so load
        weights into the dynamics BEFORE the first training call)."""
        if self._engine is None:
            from .training import Adam, TrainEngine
            self._engine = TrainEngine(self.dynamics, metric=self.metric_fn, loss_scale=self.loss_scale)
            self._optimizer = Adam(self._engine.params)
        return self._engine

    def _draws(self, x, draws_x, z, draws_z):
        """The in-graph draws of _calc_loss (:754-758): injected, or from the Philox stream."""
        dyn, lib = self.dynamics, _lib.get_lib()
        N, D = x.shape
        dev = x.device

        def philox(seed):
            v = torch.empty(N, D, dtype=torch.float32, device=dev)
            ud, ua = (torch.empty(N, dtype=torch.float32, device=dev) for _ in range(2))
            with torch.cuda.device(dev):
                _lib.check(lib.l2hmc_philox_draws(int(seed), int(dyn._mc_step), int(dyn.chain_offset), N, D,
                                                  _lib.dptr(v), _lib.dptr(ud), _lib.dptr(ua), _lib.stream_ptr()))
            return v, ud, ua
        drew = False
        if draws_x is None:
            v, ud, ua = philox(dyn.seed)
            draws_x, drew = (v, v, ud, ua), True
        if z is None:
            z, drew = philox(int(dyn.seed) ^ 0x5A5A5A5A)[0], True
        if draws_z is None:
            v, ud, ua = philox(int(dyn.seed) ^ 0x3C3C3C3C)
            draws_z, drew = (v, v, ud, ua), True
        if drew:                       # one counter tick per call whenever any of the three sets came from Philox
            dyn._mc_step += 1
        dv = lambda a: to_device(a, self.device).reshape(N, -1).contiguous()
        d1 = lambda a: to_device(a, self.device).reshape(-1).contiguous()
        fix = lambda d: (dv(d[0]), dv(d[1] if d[1] is not None else d[0]), d1(d[2]), d1(d[3]))
        return fix(draws_x), dv(z), fix(draws_z)

    def _calc_loss_and_grads(self, x, beta, draws_x=None, z=None, draws_z=None, _first_bucket_event=None, **weights):
        """(loss, grads, x_out, px, x_dq) -- gauge_model.py:805-832.  `grads` is a dict of views into the flat
        gradient arena keyed 'eps', 'XNet.<layer>.kernel', ... in the order of dynamics.variables; with clip_grads
        they are already clipped by their global norm (:826-827), as the reference returns them."""
        x = self._x(x).contiguous()
        w = dict(self.loss_weights)
        w.update(weights)
        draws_x, z, draws_z = self._draws(x, draws_x, z, draws_z)
        eng = self.engine
        with torch.cuda.device(self.device):
            if self.native_train:          # one C call enqueues the whole step (csrc/train_step.cu)
                fw = eng.native(x, beta, draws_x, z, draws_z, w, first_bucket_event=_first_bucket_event)
                grads = eng.params.grads_dict()
            else:                          # the same launches driven from Python (kept for tests / debugging)
                fw = eng.forward(x, beta, draws_x, z, draws_z, w)
                grads = eng.backward(fw)
            if self.clip_grads and not fw.get("forward_only"):   # tf.clip_by_global_norm, per rank (:826-827)
                self._optimizer.clip_by_global_norm(self.clip_value)
        self.last_per_chain_loss, self.last_forward = fw["per_chain"], fw
        return fw["loss"][0], grads, fw["x_out"], fw["px"], fw["x_dq"]

    def learning_rate(self, global_step=None):
        """tf.train.exponential_decay(..., staircase=True) (:934-939), times the world size under Horovod (:940)."""
        from .training import exponential_decay
        step = self.global_step if global_step is None else global_step
        lr = exponential_decay(self.lr_init, step, self.lr_decay_steps, self.lr_decay_rate, staircase=True)
        if self.using_hvd:
            import torch.distributed as dist
            lr *= dist.get_world_size()
        return lr

    def train_step(self, samples, beta=None, **inject):
        """One iteration of the training loop (:1155-1190): loss + gradients, [all-reduce], Adam, observables of the
        new samples.  Returns (loss, samples_out, px, eps, actions, plaqs, charges, lr, charge_diff) with
        samples_out = mod(x_out, 2 pi) as the loop feeds it back (:1180)."""
        beta = self.beta if beta is None else beta
        eng, opt = self.engine, self._optimizer
        # hvd.DistributedOptimizer (:941-943): average of the per-rank (clipped) gradients.  Two buckets of the flat
        # arena, eps | XNet and VNet: the first is averaged on a side stream while momentum_fn's weight-gradient products
        # still run (the C call records an event in between).  A global-norm clip needs the whole gradient first.
        scale = 1.
        overlap = (self.using_hvd and self.native_train and not self.clip_grads and not self.dynamics.hmc
                   and self.overlap_allreduce)
        ev = None
        if overlap:
            if getattr(self, "_bucket_event", None) is None:
                self._bucket_event = torch.cuda.Event()
                self._bucket_event.record()                           # creates the handle the C call records into
                self._comm_stream = torch.cuda.Stream(device=self.device)
                names = eng.params.names
                self._bucket_split = int(eng.params.offsets[next(i for i, n in enumerate(names) if n.startswith("VNet."))])
            ev = self._bucket_event
        loss, grads, x_out, px, x_dq = self._calc_loss_and_grads(samples, beta, _first_bucket_event=ev, **inject)
        if self.using_hvd:
            from .sharding import allreduce_mean_
            if overlap:
                main = torch.cuda.current_stream(self.device)
                split = self._bucket_split
                self._comm_stream.wait_event(ev)
                with torch.cuda.stream(self._comm_stream):
                    allreduce_mean_(eng.params.grad[:split], divide=False)
                allreduce_mean_(eng.params.grad[split:], divide=False)
                main.wait_stream(self._comm_stream)
            else:
                allreduce_mean_(eng.params.grad, divide=False)
            import torch.distributed as dist
            scale = 1.0 / dist.get_world_size()        # the SUM becomes the average inside the Adam kernel
        lr = self.learning_rate()
        with torch.cuda.device(self.device):
            opt.step(lr, grad_scale=scale)   # the clip (if any) was applied by _calc_loss_and_grads, before the average
        self.global_step += 1
        x_new = self._wrap(x_out)
        actions, plaqs, _, charges = self.lattice.observables(x_new)
        charge_diff = x_dq.sum().to(torch.float32) / self.num_samples
        return loss, x_new, px, self.dynamics.eps, actions, plaqs, charges, lr, charge_diff

    def _wrap(self, x):
        """np.mod(x_out, 2 pi) of the host loop (:1180) on the device (exact, l2hmc_mh_select's wrap)."""
        N, D = x.shape
        out = torch.empty_like(x)
        consts = getattr(self, "_wrap_consts", None)          # accept probability 1 > uniform 0: every chain "accepts"
        if consts is None or consts[0].numel() != N or consts[0].device != x.device:
            consts = self._wrap_consts = (torch.ones(N, dtype=torch.float32, device=x.device),
                                          torch.zeros(N, dtype=torch.float32, device=x.device))
        ones, zeros = consts
        with torch.cuda.device(x.device):
            _lib.check(_lib.get_lib().l2hmc_mh_select(_lib.dptr(x), _lib.dptr(x), None, None, None, _lib.dptr(ones), None,
                                                      None, _lib.dptr(zeros), N, D, 1, None, None, None, _lib.dptr(out),
                                                      None, _lib.stream_ptr()))
        return out

    def train(self, train_steps, samples_init=None, beta_init=None, distributed=False, log=None):
        """The training loop (:1112-1230) without files / TensorBoard: beta annealing (:1158), train_step, feed the
        wrapped samples back.  `log`: a callable (step, train_step's tuple, beta), a run_logger.RunLogger, or a file path
        (rank-0 JSONL rows STEP LOSS t/STEP %ACC EPS BETA ACTION PLAQ (EXACT) dQ LR, :1104-1110).  distributed=True: one process per GPU, weights broadcast from rank 0 at the start
        (BroadcastGlobalVariablesHook), lr * world size, gradients averaged, steps // world size (:2085-2096)."""
        self.using_hvd = bool(distributed)
        samples = self._x(self.lattice.samples if samples_init is None else samples_init).contiguous()
        eng = self.engine
        if self.using_hvd:
            import torch.distributed as dist
            from .sharding import broadcast_weights
            rank, ws = dist.get_rank(), dist.get_world_size()
            # each rank owns num_samples chains of the global batch: key its Philox draws by the GLOBAL chain index
            # (otherwise every rank would draw the same momenta / uniforms and the average would change nothing)
            if self.dynamics.chain_offset == 0:
                self.dynamics.chain_offset = rank * self.num_samples
            elif rank > 0 and self.dynamics.chain_offset < self.num_samples:
                raise ValueError(f"rank {rank}: chain_offset {self.dynamics.chain_offset} overlaps rank 0's chains")
            # the masks come from each process's NumPy global RNG: rank 0's are THE model's (they are state, :651-661)
            mask = self.dynamics.mask.clone()
            broadcast_weights([eng.params.flat, mask])
            self.dynamics.set_masks(mask.cpu().numpy())
            eng.params.sync_to_dynamics()
            train_steps = train_steps // ws
        if isinstance(log, str):       # a path: rank-0 JSONL rows with the reference's columns (:1104-1110, 1190-1221)
            from .run_logger import RunLogger
            log = RunLogger(log, print_steps=getattr(self, "print_steps", 1), train_steps=train_steps)
        history = []
        for step in range(self.global_step, train_steps):
            self.beta = self.update_beta(step) if self.annealing else (self.beta_init if beta_init is None else beta_init)
            out = self.train_step(samples, self.beta)
            samples = out[1]
            if log is not None:
                log(step, out, self.beta)
            history.append(out[0])
        self.samples = samples
        return samples, history

    # ------------------------------------------------------------------ saved state (:519-556, :1104-1110)
    def save_state(self, path, samples=None):
        """Weights, eps, Adam moments, masks, step, beta and the current samples in one torch.save file (the
        reference pickles the same items next to a TF checkpoint; its on-disk format is not kept)."""
        eng, opt = self.engine, self._optimizer
        state = {"names": list(eng.params.names), "shapes": [list(sh) for sh in eng.params.shapes],
                 "offsets": [int(o) for o in eng.params.offsets],
                 "flat": eng.params.flat.cpu(), "adam_m": opt.m.cpu(), "adam_v": opt.v.cpu(), "adam_t": opt.t,
                 "global_step": self.global_step, "beta": float(self.beta), "lr": self.learning_rate(),
                 "mask": self.dynamics.mask.cpu(), "mc_step": self.dynamics._mc_step, "z_step": self._z_step,
                 "samples": None if samples is None else to_device(samples, self.device).cpu(),
                 "params": {k: v for k, v in self.params.items() if isinstance(v, (int, float, str, bool, type(None)))}}
        torch.save(state, path)

    def restore_state(self, path):
        """Inverse of save_state; returns the saved samples (or None)."""
        state = torch.load(path, map_location="cpu", weights_only=True)   # plain tensors / lists / scalars only
        eng, opt = self.engine, self._optimizer
        if list(state["names"]) != list(eng.params.names) or list(state["offsets"]) != list(eng.params.offsets):
            raise ValueError("restore_state: the saved variables do not match this model")
        eng.params.flat.copy_(state["flat"])
        opt.m.copy_(state["adam_m"])
        opt.v.copy_(state["adam_v"])
        opt.t = int(state["adam_t"])
        self.global_step, self.beta = int(state["global_step"]), float(state["beta"])
        self.dynamics.set_masks(state["mask"].numpy())
        self.dynamics._mc_step, self._z_step = int(state["mc_step"]), int(state["z_step"])
        eng.params.sync_to_dynamics()
        return None if state["samples"] is None else state["samples"].to(self.device)

    # ------------------------------------------------------------------ statistics (:1473-1533)
    def calc_observables_stats(self, observables, therm_frac=10):
        """observables = (actions_dict, plaqs_dict, charges_dict) of {step: (N,) array}; returns
        (actions_stats, plaqs_stats, charges_stats, suscept_stats, charge_probabilities)."""
        actions, plaqs, charges = (list(d.values()) for d in observables[:3])
        stats = RunStats(len(np.asarray(actions[0]).reshape(-1)) if not isinstance(actions[0], torch.Tensor)
                         else actions[0].numel(), len(actions), therm_frac, device=self.device)
        for a, p, q in zip(actions, plaqs, charges):
            stats.update(a, p, q)
        return stats.finalize()
