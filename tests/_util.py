import numpy as np
import torch

RTOL = 1e-4  # BASELINE.json north_star: "within 1e-4 relative in fp32"


def rel_err(got, want):
    """max |got - want| / max |want|  (relative to the tensor's scale, fp32)."""
    got = torch.as_tensor(got).detach().cpu().double()
    want = torch.as_tensor(want).detach().cpu().double()
    assert got.shape == want.shape, (got.shape, want.shape)
    if want.numel() == 0:
        return 0.0
    return float((got - want).abs().max() / want.abs().max().clamp_min(1e-30))


def assert_close(got, want, name, tol=RTOL, want64=None):
    """``want`` = the reference evaluated in fp32 on the CPU.  ``want64`` (optional) = the same reference code on the
    same fp32 inputs evaluated in fp64; it adjudicates LeakyReLU kinks (oracle/pipeline_ref.py): a pre-activation that
    rounding puts on the other side of zero moves a gradient by a finite amount in any fp32 implementation, the
    reference's own included, so the result must be within ``tol`` of the fp32 OR the fp64 evaluation."""
    e = rel_err(got, want)
    if e > tol and want64 is not None:
        e = min(e, rel_err(got, want64))
    assert e <= tol, "%s: relative error %.3g > %.1g" % (name, e, tol)


def canon_neighbours(pidx, xyz, loc_w):
    """Sort each sample's K neighbour indices by (d^2, index) so that sets compare independent of slot order."""
    pidx = np.asarray(pidx)
    xyz = np.asarray(xyz, np.float64)
    loc = np.asarray(loc_w, np.float64)
    d = xyz[np.clip(pidx, 0, None)] - loc[..., None, :]
    d2 = (d * d).sum(-1)
    d2 = np.where(pidx >= 0, d2, np.inf)
    order = np.lexsort((pidx, d2), axis=-1)
    return np.take_along_axis(pidx, order, axis=-1)


def group_linear_grads(grads):
    """{name: grad} -> {name: grad} with every Linear's ``X.weight`` / ``X.bias`` pair merged into one ``X.[weight|bias]``
    tensor [out, in + 1]: the bias is the weight of a constant-1 input column, so its gradient is compared on the scale of
    the layer's gradient (a single-element bias gradient such as ``alpha_branch.0.bias`` is a sum of cancelling per-pair
    terms, and "relative to its own value" would measure the cancellation, not the arithmetic)."""
    out = {}
    for n, g in grads.items():
        if n.endswith(".weight") and n[:-7] + ".bias" in grads:
            out[n[:-7] + ".[weight|bias]"] = torch.cat([torch.as_tensor(g).detach().cpu().double(),
                                                        torch.as_tensor(grads[n[:-7] + ".bias"]).detach().cpu().double()[:, None]], dim=1)
        elif n.endswith(".bias") and n[:-5] + ".weight" in grads:
            continue
        else:
            out[n] = g
    return out
