"""Multi-tensor Adam on the C ABI (pn_adam_step) with the interface subset of ``torch.optim.Adam`` the reference uses.

The reference builds ``torch.optim.Adam(params, lr=..., betas=(0.9, 0.999))`` twice - the MLP (``optimizer``) and the
neural point tables + ``bg_color`` (``neural_point_optimizer``), models/mvs_points_volumetric_model.py:47-100 - and steps
them in ``backward`` (:103-125).  ``Adam`` here takes the same arguments, keeps the same per-parameter state
(``step``, ``exp_avg``, ``exp_avg_sq``) and updates every tensor of a group in ONE launch; the arithmetic follows
torch's single-tensor Adam op by op, so a reference run can switch optimisers mid-training.  No CPU path.
"""
import torch

from . import _lib
from ._lib import AdamSegment


class Adam:
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False):
        if weight_decay != 0 or amsgrad:
            raise NotImplementedError("the reference uses plain Adam (no weight decay, no amsgrad)")
        params = list(params)
        if params and isinstance(params[0], dict):
            self.param_groups = [dict(lr=lr, betas=betas, eps=eps, **g) for g in params]
        else:
            self.param_groups = [dict(params=params, lr=lr, betas=betas, eps=eps)]
        for g in self.param_groups:
            g["params"] = list(g["params"])
        self.state = {}
        _lib.lib()  # fail now, loudly, if the CUDA library is missing

    def zero_grad(self, set_to_none=True):
        for g in self.param_groups:
            for p in g["params"]:
                if p.grad is not None:
                    if set_to_none:
                        p.grad = None
                    else:
                        p.grad.zero_()

    @torch.no_grad()
    def step(self):
        lib = _lib.lib()
        for g in self.param_groups:
            by_step = {}
            for p in g["params"]:
                if p.grad is None:
                    continue
                st = self.state.get(p)
                if st is None:
                    st = self.state[p] = dict(step=0, exp_avg=torch.zeros_like(p, memory_format=torch.contiguous_format),
                                              exp_avg_sq=torch.zeros_like(p, memory_format=torch.contiguous_format))
                st["step"] += 1
                by_step.setdefault(st["step"], []).append((p, st))
            for t, items in by_step.items():          # (parameters added later have their own step count)
                segs = (AdamSegment * len(items))()
                keep = []
                for s, (p, st) in zip(segs, items):
                    grad = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                    keep.append(grad)
                    s.param, s.grad = _lib.ptr(p.data), _lib.ptr(grad)
                    s.exp_avg, s.exp_avg_sq, s.n = _lib.ptr(st["exp_avg"]), _lib.ptr(st["exp_avg_sq"]), p.numel()
                _lib.check(lib.pn_adam_step(segs, len(items), float(g["lr"]), float(g["betas"][0]), float(g["betas"][1]),
                                            float(g["eps"]), int(t), _lib.stream()), "pn_adam_step")


class RowAdam:
    """Lazy Adam over the per-point tables: only the rows a step touched are updated (pn_adam_rows, the arithmetic of
    ``torch.optim.SparseAdam``).  NOT the reference's optimiser: its dense ``torch.optim.Adam`` keeps moving every point on
    its momentum each step (``Adam`` above reproduces that exactly); at 20 M points the dense update is 21.8 GB of traffic
    per step, more than the forward + backward.  Use it knowingly: untouched rows keep their moments and do not move.

    params: up to four parameters shaped [1, N, w] / [N, w] / [N] that share the row index (points_embeding, points_conf,
    points_dir, points_color).  ``step(ids, stride=1)``: ids int32 device tensor, entries < 0 are skipped (the
    ``touched_ids`` list of the fused backward, or the id column of the gathered rows of a sharded exchange)."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        self.params = [p for p in params if p is not None]
        if not 1 <= len(self.params) <= 4:
            raise ValueError("RowAdam: 1..4 tables")
        self.lr, self.betas, self.eps, self.t = float(lr), (float(betas[0]), float(betas[1])), float(eps), 0
        self.state = [dict(exp_avg=torch.zeros_like(p, memory_format=torch.contiguous_format),
                           exp_avg_sq=torch.zeros_like(p, memory_format=torch.contiguous_format)) for p in self.params]
        n = {int(p.numel() // self._width(p)) for p in self.params}
        if len(n) != 1:
            raise ValueError("RowAdam: the tables must have the same number of rows")
        _lib.lib()

    @staticmethod
    def _width(p):
        return int(p.shape[-1]) if p.dim() >= 2 else 1

    @torch.no_grad()
    def step(self, ids, stride=1, n=None):
        self.t += 1
        tabs = (_lib.AdamRowsTable * len(self.params))()
        keep = []
        for tb, p, st in zip(tabs, self.params, self.state):
            g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
            keep.append(g)
            tb.param, tb.grad, tb.exp_avg, tb.exp_avg_sq, tb.width = _lib.ptr(p.data), _lib.ptr(g), _lib.ptr(st["exp_avg"]), _lib.ptr(st["exp_avg_sq"]), self._width(p)
        n = int(ids.numel() // stride) if n is None else int(n)
        _lib.check(_lib.lib().pn_adam_rows(_lib.ptr(ids), n, int(stride), tabs, len(self.params), self.lr, self.betas[0], self.betas[1],
                                           self.eps, self.t, _lib.stream()), "pn_adam_rows")
