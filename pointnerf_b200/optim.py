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
