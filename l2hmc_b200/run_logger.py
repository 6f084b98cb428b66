"""Rank-0 training log with the reference's columns.

The reference prints, every `print_steps`, the row  STEP LOSS t/STEP %ACC EPS BETA ACTION PLAQ (EXACT) dQ LR
(gauge_model.py:1104-1110, filled at :1190-1221) and appends it to run_info.txt, both gated on hvd.rank() == 0
(utils/file_io.py:27-46).  RunLogger writes the same row as one JSON object per line (machine-readable) and, on request,
the reference's fixed-width text line; only rank 0 of the default process group writes anything.  Host bookkeeping:
nothing here touches the GPU except reading the step's (already computed) means.
"""
import json
import time

import torch

from .lattice import u1_plaq_exact
from .sharding import world

COLUMNS = ("STEP", "LOSS", "t/STEP", "% ACC", "EPS", "BETA", "ACTION", "PLAQ", "(EXACT)", "dQ", "LR")


def _f(v):
    if isinstance(v, torch.Tensor):
        return float(v.detach().float().mean().item())
    return float(v)


class RunLogger(object):
    def __init__(self, path=None, print_steps=1, train_steps=None, echo=False, rank=None):
        self.rank = world()[0] if rank is None else int(rank)
        self.path, self.print_steps, self.train_steps, self.echo = path, max(1, int(print_steps)), train_steps, bool(echo)
        self._t = time.time()
        self._wrote_header = False
        if self.active and path:
            open(path, "a").close()

    @property
    def active(self):
        return self.rank == 0

    @property
    def header(self):
        """The reference's train_header (gauge_model.py:1104-1110)."""
        h = ("{:^12s}" + "{:^10s}" * 10).format(*COLUMNS)
        dash = (len(h) + 1) * "-"
        return dash + "\n" + h + "\n" + dash

    def row(self, step, out, beta):
        """`out` = GaugeModel.train_step's tuple (loss, samples, px, eps, actions, plaqs, charges, lr, charge_diff)."""
        loss, _, px, eps, actions, plaqs, _, lr, charge_diff = out
        now = time.time()
        rec = {"STEP": int(step), "LOSS": _f(loss), "t/STEP": now - self._t, "% ACC": _f(px), "EPS": _f(eps),
               "BETA": float(beta), "ACTION": _f(actions), "PLAQ": _f(plaqs), "(EXACT)": float(u1_plaq_exact(beta)),
               "dQ": _f(charge_diff), "LR": float(lr)}
        self._t = now
        return rec

    def text(self, rec):
        """The reference's data_str (gauge_model.py:1207-1217)."""
        total = self.train_steps if self.train_steps is not None else 0
        return (f"{rec['STEP']:>5g}/{total:<6g} {rec['LOSS']:^9.4g} {rec['t/STEP']:^9.4g} {rec['% ACC']:^9.4g}"
                f"{rec['EPS']:^9.4g} {rec['BETA']:^9.4g} {rec['ACTION']:^9.4g} {rec['PLAQ']:^9.4g} "
                f"{rec['(EXACT)']:^9.4g} {rec['dQ']:^9.4g} {rec['LR']:^9.4g}")

    def __call__(self, step, out, beta):
        """The `log=` callback of GaugeModel.train.  Reading the means synchronises with the step, so rows are only
        built on rank 0 and only every print_steps (as the reference's `step % self.print_steps == 0`)."""
        if not self.active or step % self.print_steps:
            self._t = time.time() if not self.active else self._t
            return None
        rec = self.row(step, out, beta)
        if self.path:
            with open(self.path, "a") as f:
                f.write(json.dumps(rec) + "\n")
        if self.echo:
            if not self._wrote_header:
                print(self.header)
                self._wrote_header = True
            print(self.text(rec))
        return rec
