"""Drop-in ``ray_march`` / ``alpha_ray_march`` on the C ABI (pn_composite_forward / _backward).

Mirrors models/rendering/diff_ray_marching.py:509-573 of the reference with the only render / blend
functions its scripts select - ``radiance_render`` (``feat[..., 1:]``, diff_render_func.py:48-50) and
``alpha_blend`` (``opacity * acc_transmission``, :36-37): same signature, same 7-tuple / 5-tuple, same
dtypes and shapes, differentiable w.r.t. ``ray_features`` and ``bg_color``.  Also ``ray_dist_from_samples``,
the cummax / diff / clamp block of models/neural_points_volumetric_model.py:293-301.

The reference hard-codes 128 colour channels (``.view(B, 1, 128)``, :546); here C = features - 1.
There is no CPU path: CPU tensors raise.
"""
import torch

from . import _lib


def radiance_render(ray_feature):
    """diff_render_func.py:48-50."""
    return ray_feature[..., 1:]


def alpha_blend(opacity, acc_transmission):
    """diff_render_func.py:36-37."""
    return opacity * acc_transmission


def find_render_function(name):
    if name == "radiance":
        return radiance_render
    raise NotImplementedError("render function %r is outside the hot path (all scripts: which_render_func='radiance')" % name)


def find_blend_function(name):
    if name == "alpha":
        return alpha_blend
    raise NotImplementedError("blend function %r is outside the hot path (all scripts: which_blend_func='alpha')" % name)


def _check_funcs(render_func, blend_func):
    if render_func is not None and getattr(render_func, "__name__", "") != "radiance_render":
        raise NotImplementedError("ray_march: only radiance_render is implemented, got %r" % (render_func,))
    if getattr(blend_func, "__name__", "") != "alpha_blend":
        raise NotImplementedError("ray_march: only alpha_blend is implemented, got %r" % (blend_func,))


class _Composite(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ray_features, ray_dist, ray_valid_u8, bg_color, want_color, slot_row=None, n_rows_dev=None):
        lib = _lib.lib()
        R, SR = (int(v) for v in ray_dist.shape)       # ray_features: [R,SR,C+1], or [S,C+1] with slot_row (compact)
        C1 = int(ray_features.shape[-1])
        C = C1 - 1
        dev = ray_features.device
        feats = ray_features.detach().contiguous()
        dist = ray_dist.detach().contiguous().float()
        bg = bg_color.detach().contiguous().float().view(-1) if bg_color is not None else None
        if bg is not None and bg.numel() != C:
            raise ValueError("bg_color has %d entries, features have %d colour channels" % (bg.numel(), C))
        ray_color = torch.empty((R, C), dtype=torch.float32, device=dev) if want_color else None
        opacity = torch.empty((R, SR), dtype=torch.float32, device=dev)
        acc_T = torch.empty((R, SR), dtype=torch.float32, device=dev)
        blend_w = torch.empty((R, SR), dtype=torch.float32, device=dev)
        bg_T = torch.empty((R,), dtype=torch.float32, device=dev)
        feats_k = feats if feats.numel() else feats.new_zeros((1, C1))     # compact layout with no valid sample at all
        if R > 0:
            _lib.check(lib.pn_composite_forward(_lib.ptr(feats_k), _lib.ptr(dist), _lib.ptr(ray_valid_u8), _lib.ptr(slot_row), _lib.ptr(bg), R, SR, C,
                                                _lib.ptr(ray_color), _lib.ptr(opacity), _lib.ptr(acc_T), _lib.ptr(blend_w),
                                                _lib.ptr(bg_T), _lib.stream()), "pn_composite_forward")
        ctx.set_materialize_grads(False)    # unused outputs arrive as None (NULL for the kernel), not as zero tensors to fill and read
        ctx.save_for_backward(feats, dist, ray_valid_u8, bg if bg is not None else torch.empty(0, device=dev), opacity, acc_T, bg_T)
        ctx.has_bg, ctx.want_color, ctx.slot_row, ctx.n_rows_dev = bg is not None, want_color, slot_row, n_rows_dev
        ctx.bg_shape = tuple(bg_color.shape) if bg_color is not None else None
        if ray_color is None:
            ray_color = torch.empty(0, device=dev)
        return ray_color, opacity, acc_T, blend_w, bg_T

    @staticmethod
    def backward(ctx, g_color, g_opacity, g_acc_T, g_blend_w, g_bg_T):
        lib = _lib.lib()
        feats, dist, valid, bg, opacity, acc_T, bg_T = ctx.saved_tensors
        R, SR = (int(v) for v in dist.shape)
        C1 = int(feats.shape[-1])
        C = C1 - 1
        dev = feats.device

        def c(t):
            return t.contiguous() if t is not None else None

        g_color = c(g_color) if ctx.want_color else None
        g_feat = torch.empty_like(feats)
        need_bg = ctx.has_bg and ctx.needs_input_grad[3]
        g_bg = torch.empty((C,), dtype=torch.float32, device=dev) if need_bg else None
        if R == 0:
            return g_feat, None, None, (torch.zeros(ctx.bg_shape, device=dev) if need_bg else None), None, None, None
        n_rows = int(feats.shape[0]) if ctx.slot_row is not None else 0
        ws_bytes = lib.pn_composite_workspace_bytes(R, C, n_rows)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        feats_k = feats if feats.numel() else feats.new_zeros((1, C1))
        g_k = g_feat if g_feat.numel() else g_feat.new_zeros((1, C1))
        _lib.check(lib.pn_composite_backward(_lib.ptr(feats_k), _lib.ptr(dist), _lib.ptr(valid), _lib.ptr(ctx.slot_row),
                                             _lib.ptr(bg) if ctx.has_bg else None,
                                             _lib.ptr(opacity), _lib.ptr(acc_T), _lib.ptr(bg_T), _lib.ptr(g_color),
                                             _lib.ptr(c(g_opacity)), _lib.ptr(c(g_acc_T)), _lib.ptr(c(g_blend_w)),
                                             _lib.ptr(c(g_bg_T)), R, SR, C, n_rows, _lib.ptr(ctx.n_rows_dev), _lib.ptr(g_k), _lib.ptr(g_bg),
                                             _lib.ptr(ws), ws_bytes, _lib.stream()), "pn_composite_backward")
        return g_feat, None, None, (g_bg.view(ctx.bg_shape) if need_bg else None), None, None, None


def _as_u8(mask):
    """bool / integer validity mask -> contiguous uint8 (a bool tensor is reinterpreted, not converted)."""
    mask = mask.contiguous()
    return mask.view(torch.uint8) if mask.dtype in (torch.bool, torch.int8) else mask.to(torch.uint8)


def ray_march_compact(ray_dist, ray_valid, features_compact, slot_row, bg_color=None, n_rows_dev=None):
    """Fused-path variant of ``ray_march``: ``features_compact`` [S,C+1] holds one row per valid sample and
    ``slot_row`` int32 [R,SR] maps sample (r,s) to its row (-1: none).  Same 7-tuple minus ``point_color`` (None).
    ``n_rows_dev`` (device int32[1]): the number of valid rows when the host never learned it (``features_compact`` is then
    allocated at capacity)."""
    B, R, SR = ray_dist.shape
    if B != 1:
        raise NotImplementedError("B = 1 only, like every caller of the reference")
    valid = _as_u8(ray_valid.reshape(R, SR))
    ray_color, opacity, acc_T, blend_w, bg_T = _Composite.apply(features_compact, ray_dist.reshape(R, SR), valid, bg_color, True,
                                                                slot_row.reshape(R, SR).contiguous(), n_rows_dev)
    bg_T = bg_T.view(B, R, 1)
    return (ray_color.view(B, R, features_compact.shape[-1] - 1), None, opacity.view(B, R, SR), acc_T.view(B, R, SR),
            blend_w.view(B, R, SR, 1), bg_T, bg_T)      # background_blend_weight = 1 * T: the same values


def _flatten(ray_dist, ray_valid, ray_features):
    B, R, SR = ray_dist.shape
    if B != 1:
        raise NotImplementedError("B = 1 only, like every caller of the reference")
    return (ray_features.reshape(R, SR, ray_features.shape[-1]), ray_dist.reshape(R, SR),
            _as_u8(ray_valid.reshape(R, SR)), B, R, SR)


def ray_march(ray_dist, ray_valid, ray_features, render_func, blend_func, bg_color=None):
    """diff_ray_marching.py:509-555.  ray_dist [1,R,SR], ray_valid bool [1,R,SR], ray_features [1,R,SR,C+1],
    bg_color [.., C] or None -> (ray_color [1,R,C], point_color [1,R,SR,C], opacity [1,R,SR],
    acc_transmission [1,R,SR], blend_weight [1,R,SR,1], background_transmission [1,R,1],
    background_blend_weight [1,R,1])."""
    _check_funcs(render_func, blend_func)
    feats, dist, valid, B, R, SR = _flatten(ray_dist, ray_valid, ray_features)
    point_color = ray_features[..., 1:]
    ray_color, opacity, acc_T, blend_w, bg_T = _Composite.apply(feats, dist, valid, bg_color, True)
    bg_T = bg_T.view(B, R, 1)
    return (ray_color.view(B, R, feats.shape[-1] - 1), point_color, opacity.view(B, R, SR), acc_T.view(B, R, SR), blend_w.view(B, R, SR, 1),
            bg_T, 1 * bg_T)


def alpha_ray_march(ray_dist, ray_valid, ray_features, blend_func):
    """diff_ray_marching.py:558-573 -> (opacity, acc_transmission, blend_weight, background_transmission,
    background_blend_weight)."""
    _check_funcs(None, blend_func)
    feats, dist, valid, B, R, SR = _flatten(ray_dist, ray_valid, ray_features)
    _, opacity, acc_T, blend_w, bg_T = _Composite.apply(feats, dist, valid, None, False)
    bg_T = bg_T.view(B, R, 1)
    return opacity.view(B, R, SR), acc_T.view(B, R, SR), blend_w.view(B, R, SR, 1), bg_T, 1 * bg_T


def ray_dist_from_samples(sample_loc, ray_valid, vsize, raydist_mode_unit=1):
    """neural_points_volumetric_model.py:293-301.  sample_loc [1,R,SR,3] (perspective coords: [...,2] is the
    camera depth), ray_valid bool [1,R,SR] -> ray_dist [1,R,SR].  No gradient (the reference's flows only to xyz)."""
    lib = _lib.lib()
    B, R, SR = ray_valid.shape
    loc = sample_loc.detach().reshape(B * R, SR, 3).contiguous()
    valid = _as_u8(ray_valid.reshape(B * R, SR))
    out = torch.empty((B * R, SR), dtype=torch.float32, device=loc.device)
    if B * R == 0:
        return out.view(B, R, SR)
    _lib.check(lib.pn_ray_dist(_lib.ptr(loc), _lib.ptr(valid), B * R, SR, float(vsize[2]), int(raydist_mode_unit),
                               _lib.ptr(out), _lib.stream()), "pn_ray_dist")
    return out.view(B, R, SR)


class _ColorMSE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, target, ray_weight, inv_norm, ray_mask=None, inv_norm_dev=None):
        lib = _lib.lib()
        C = int(pred.shape[-1])
        p = pred.detach().reshape(-1, C).contiguous().float()
        t = target.detach().reshape(-1, C).contiguous().float()
        if p.shape != t.shape:
            raise ValueError("color_mse: prediction %s and target %s differ" % (tuple(pred.shape), tuple(target.shape)))
        w = ray_weight.detach().reshape(-1).contiguous().float() if ray_weight is not None else None
        if w is not None and w.numel() != p.shape[0]:
            raise ValueError("color_mse: %d ray weights for %d rays" % (w.numel(), p.shape[0]))
        loss = torch.empty((), dtype=torch.float32, device=p.device)
        grad = torch.empty_like(p) if ctx.needs_input_grad[0] else None
        ws_bytes = lib.pn_color_mse_workspace_bytes()
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=p.device)
        mk = ray_mask.detach().reshape(-1).contiguous() if ray_mask is not None else None
        if mk is not None and (mk.dtype != torch.int32 or mk.numel() != p.shape[0]):
            raise ValueError("color_mse: ray_mask must be int32 with one entry per ray")
        _lib.check(lib.pn_color_mse(_lib.ptr(p), _lib.ptr(t), _lib.ptr(w), _lib.ptr(mk), int(p.shape[0]), C, float(inv_norm),
                                    _lib.ptr(inv_norm_dev), _lib.ptr(loss), _lib.ptr(grad), _lib.ptr(ws), ws_bytes, _lib.stream()),
                   "pn_color_mse")
        ctx.grad, ctx.shape = grad, tuple(pred.shape)
        return loss

    @staticmethod
    def backward(ctx, g):
        grad, ctx.grad = ctx.grad, None
        return (grad * g).view(ctx.shape), None, None, None, None, None


def color_mse(pred, target, ray_weight=None, norm=None, ray_mask=None, inv_norm_dev=None):
    """Colour loss of ``compute_losses`` (base_rendering_model.py:543-550: ``torch.nn.MSELoss`` of ``coarse_raycolor`` against
    ``gt_image``) with its gradient produced in the same pass.  pred, target [..., C]; ray_weight [...] or None (the
    ``ray_masked_*`` items: a {0,1} mask, ``norm`` = number of selected elements); norm: the divisor (default: pred.numel()).
    ``ray_mask`` int32 [...]: rays with mask <= 0 get weight 0 (the query's per-ray neighbour flag, no float conversion);
    ``inv_norm_dev`` (device float[1]): an extra factor of the normaliser that stays on the device (1 / masked count).
    Returns the 0-d loss; deterministic (fixed reduction tree)."""
    if not pred.is_cuda:
        raise _lib.LibraryError("color_mse needs CUDA tensors (there is no CPU path in this package)")
    n = float(norm) if norm is not None else float(max(pred.numel(), 1))
    return _ColorMSE.apply(pred, target, ray_weight, 1.0 / n, ray_mask, inv_norm_dev)
