"""LeakyReLU-kink probe for gradient parity (TEST INFRASTRUCTURE, see oracle/__init__.py).

The aggregator MLP (point_aggregators.py:276-348) is piecewise linear in its hidden units.  The FORWARD is
continuous, so two fp32 implementations agree to rounding (~1e-7).  The BACKWARD is not: a hidden unit whose
pre-activation is within rounding of zero has derivative 1 in one implementation and 0.01 in the other, which
moves that pair's gradient by ~10 % and every upstream weight gradient by ~1e-3 - the reference's own module run
in fp32 and in fp64 on the same inputs differs by 4e-4 .. 5e-3 on such inputs (measured, DESIGN.md "Parity").
"Gradients within 1e-4" is therefore only a well-posed statement on inputs whose hidden units are not AT a kink.

This probe evaluates the oracle aggregator once in fp64 with forward hooks on its Linear layers and reports, per
sample, whether any hidden unit on that sample's gradient path (block1 / block3 units of its K neighbour pairs,
colour-branch units of the sample itself) has |pre-activation| < tau * rms(layer).  Tests give such samples (or
the rays that contain them) ZERO loss weight: the forward is still computed and compared everywhere, and every
gradient tensor - per-point and per-weight - is then compared at the full 1e-4 tolerance, because no ambiguous
unit carries gradient in either implementation.
"""
import torch
import torch.nn as nn

TAU = 2e-5   # several times the forward rounding error of an fp32-class implementation (3xTF32 tensor cores: ~2e-6 of max) relative to the layer rms


# Frozen-pattern parity (torch_port.LeakyReLUProbe, pipeline_ref.run(frozen_signs=...)): the oracle's backward is evaluated
# under the activation pattern of the implementation being checked.  The pattern may differ from the oracle's own sign test
# only on units with |pre-activation| < PATTERN_TOL * rms(layer): flipping such a unit moves the FORWARD by less than the
# forward tolerance itself (1e-4 relative), i.e. both patterns are correct answers at the stated precision.
PATTERN_TOL = 1e-4


class KinkProbe:
    def __init__(self, agg, tau=TAU):
        self.tau = tau
        self.pair_min, self.sample_min = None, None
        self._handles = []
        pair_layers = [agg.block1[0], agg.block1[2], agg.block3[0], agg.block3[2]]
        sample_layers = [agg.color_branch[0], agg.color_branch[2], agg.color_branch[4]]
        for m in pair_layers:
            self._handles.append(m.register_forward_hook(self._hook("pair_min")))
        for m in sample_layers:
            self._handles.append(m.register_forward_hook(self._hook("sample_min")))

    def _hook(self, which):
        def fn(_mod, _inp, out):
            o = out.detach().double()
            rms = o.pow(2).mean().sqrt().clamp_min(1e-30)
            margin = (o.abs() / rms).min(dim=-1)[0]
            cur = getattr(self, which)
            setattr(self, which, margin if cur is None else torch.minimum(cur, margin))
        return fn

    def close(self):
        for h in self._handles:
            h.remove()

    def ambiguous_samples(self, sample_pnt_mask):
        """sample_pnt_mask bool [..., K] as given to the aggregator -> bool [...] (True = at a kink)."""
        K = sample_pnt_mask.shape[-1]
        pm = sample_pnt_mask.reshape(-1)
        sv = sample_pnt_mask.reshape(-1, K).any(-1)
        amb = torch.zeros(sv.shape[0], dtype=torch.bool)
        if self.pair_min is not None:
            pair_amb = torch.zeros(pm.shape[0], dtype=torch.bool)
            pair_amb[pm] = self.pair_min < self.tau
            amb |= pair_amb.view(-1, K).any(-1)
        if self.sample_min is not None:
            s_amb = torch.zeros(sv.shape[0], dtype=torch.bool)
            s_amb[sv] = self.sample_min < self.tau
            amb |= s_amb
        return amb.view(sample_pnt_mask.shape[:-1])


def probe(agg_port, forward_fn, sample_pnt_mask, tau=TAU):
    """Runs ``forward_fn(agg64)`` (one forward of a float64 copy of ``agg_port``) and returns the ambiguous-sample mask."""
    import copy
    agg64 = copy.deepcopy(agg_port).double()
    p = KinkProbe(agg64, tau)
    try:
        with torch.no_grad():
            forward_fn(agg64)
        return p.ambiguous_samples(sample_pnt_mask)
    finally:
        p.close()
