"""Torch-CPU restatement of the torch side of the reference hot path.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Every function cites the
reference lines it restates; ``tests/test_oracle_vs_reference.py`` checks each
against the unmodified reference modules (imported through
``oracle/ref_import.py``) in the build container, and ``tests/golden/`` holds
reference outputs for the GPU box, where ``/root/reference`` does not exist.

Only the configuration every shipped script uses is restated
(``which_agg_model=viewmlp``, ``agg_intrp_order=2``, ``agg_distance_kernel=
linear``, ``agg_dist_pers=20``, ``apply_pnt_mask=1``).
"""
import math

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F


# --------------------------------------------------------------------------
# helpers
# --------------------------------------------------------------------------
def positional_encoding(x, freqs, ori=False):
    """models/helpers/networks.py:175-189.

    ori=False: per (dim d, freq f) interleaved [sin, cos] at column (d*F+f)*2.
    ori=True : [x | sin(all d*F+f) | cos(all d*F+f)].
    Frequencies are 2**f (no pi).
    """
    bands = (2.0 ** torch.arange(freqs, dtype=x.dtype)).to(x.device)
    ang = (x[..., None] * bands).reshape(x.shape[:-1] + (freqs * x.shape[-1],))
    if ori:
        return torch.cat([x, torch.sin(ang), torch.cos(ang)], dim=-1)
    return torch.stack([torch.sin(ang), torch.cos(ang)], dim=-1).reshape(ang.shape[:-1] + (ang.shape[-1] * 2,))


def w2pers_points(xyz, camrotc2w, campos):
    """models/neural_points/neural_points.py:687-693.  xyz [N,3] -> [B,N,3] (x/z, y/z, z) in camera frame."""
    shift = xyz[None, ...] - campos[:, None, :]
    cam = torch.sum(camrotc2w[:, None, :, :] * shift[:, :, :, None], dim=-2)
    return torch.stack([cam[..., 0] / cam[..., 2], cam[..., 1] / cam[..., 2], cam[..., 2]], dim=-1)


def w2pers_samples(loc_w, camrotc2w, campos):
    """models/neural_points/query_point_indices_worldcoords.py:102-109.  loc_w [B,R,SR,3]."""
    shift = loc_w - campos[:, None, :]
    cam = torch.sum(shift[..., None, :] * torch.transpose(camrotc2w, 1, 2)[:, None, None, ...], dim=-1)
    return torch.stack([cam[..., 0] / cam[..., 2], cam[..., 1] / cam[..., 2], cam[..., 2]], dim=-1)


def sample_ts(raydir, D, near, far, jitter=0.0, uniform=None):
    """Mid-point depths of models/rendering/diff_ray_marching.py:349-393 (``middle_point_ts``).

    ``uniform`` stands for the ``torch.rand((B,R,D))`` draw of :373-375 so that the
    oracle and the CUDA path consume the same numbers.
    """
    B, R = raydir.shape[0], raydir.shape[1]
    t = torch.linspace(0, 1, D + 1, device=raydir.device).view(1, -1)
    t = near * (1 - t) + far * t
    if uniform is None:
        uniform = torch.full((B, R, D), 0.5, device=raydir.device)
    seg = (t[..., 1:] - t[..., :-1]) * (1 + jitter * (uniform - 0.5))
    ends = torch.cumsum(seg, dim=2)
    ends = torch.cat([torch.zeros((B, R, 1), device=raydir.device), ends], dim=2)
    ends = near + ends
    return (ends[:, :, :-1] + ends[:, :, 1:]) / 2


def raypos_from_ts(campos, raydir, ts):
    """diff_ray_marching.py:387: separate multiply then add (no contraction)."""
    return campos[:, None, None, :] + raydir[:, :, None, :] * ts[:, :, :, None]


# --------------------------------------------------------------------------
# gather (NeuralPoints.forward)
# --------------------------------------------------------------------------
def gather_point_tables(sample_pidx, xyz, xyz_pers, embedding, color, pdir, conf):
    """models/neural_points/neural_points.py:788-799.

    sample_pidx int [B,R,SR,K] (-1 = empty); xyz [N,3]; xyz_pers [B,N,3];
    embedding [1,N,F]; color/pdir [1,N,3]; conf [1,N,1].
    Empty slots gather point 0 (clamp) and are masked later.
    """
    mask = sample_pidx >= 0
    B, R, SR, K = sample_pidx.shape
    flat = torch.clamp(sample_pidx, min=0).view(-1).long()
    table = torch.cat([xyz[None, ...], xyz_pers, embedding], dim=-1)
    g = torch.index_select(table, 1, flat).view(B, R, SR, K, table.shape[-1])
    s_color = torch.index_select(color, 1, flat).view(B, R, SR, K, 3)
    s_dir = torch.index_select(pdir, 1, flat).view(B, R, SR, K, 3)
    s_conf = torch.index_select(conf, 1, flat).view(B, R, SR, K, 1)
    return dict(sampled_color=s_color, sampled_dir=s_dir, sampled_conf=s_conf,
                sampled_embedding=g[..., 6:], sampled_xyz_pers=g[..., 3:6], sampled_xyz=g[..., :3],
                sample_pnt_mask=mask)


# --------------------------------------------------------------------------
# aggregator
# --------------------------------------------------------------------------
class LeakyReLUProbe(nn.LeakyReLU):
    """nn.LeakyReLU (the reference's activation) that can (a) record its pre-activations and (b) run with a FROZEN
    activation pattern: ``pattern`` bool [rows, units] - True: derivative 1, False: negative_slope - replaces the sign
    test.  Forward and backward are then linear in the input, the same linear map a kernel whose forward produced that
    pattern differentiates through; where pattern == (x > 0) nothing changes.  Units at a kink (|x| within rounding of 0)
    are where two correct fp32 implementations may disagree on the pattern, and ONLY there may it differ (checked by the
    tests through ``last_pre``)."""
    pattern = None
    record = False
    last_pre = None

    def forward(self, x):
        if self.record:
            self.last_pre = x.detach()
        if self.pattern is None:
            return F.leaky_relu(x, self.negative_slope)
        pat = self.pattern.to(x.device)
        assert pat.shape == x.shape, (pat.shape, x.shape)
        return x * torch.where(pat, torch.ones((), dtype=x.dtype), torch.full((), self.negative_slope, dtype=x.dtype))


class AggregatorPort(nn.Module):
    """Restates PointAggregator (models/aggregators/point_aggregators.py:218-348, 421-429, 488-646,
    724-816) for the in-scope configuration, with the reference's parameter names so that
    ``load_state_dict`` of a reference module works unchanged."""

    def __init__(self, opt):
        super().__init__()
        self.opt = opt
        F_ = opt.point_features_dim
        W = opt.shading_feature_num
        self.dist_dim = 6  # agg_dist_pers == 20 (:243)
        in1 = F_ + 2 * opt.num_feat_freqs * F_ + 2 * opt.dist_xyz_freq * self.dist_dim  # :278-280
        in3 = W + 3 + 4                                                              # :308-309
        vch = 2 * opt.num_viewdir_freqs * 3 + opt.view_ori * 3                        # :252
        inc = W + vch                                                                # :333
        C = opt.shading_color_channel_num
        assert opt.shading_feature_mlp_layer1 == 2 and opt.shading_feature_mlp_layer3 == 2
        assert opt.shading_alpha_mlp_layer == 1 and opt.shading_color_mlp_layer == 4
        assert int(W / 2) == C, "fork: colour branch ends at shading_feature_num/2 channels (:335,343)"

        def act():
            return LeakyReLUProbe(inplace=False)

        self.block1 = nn.Sequential(nn.Linear(in1, W), act(), nn.Linear(W, W), act())
        self.block3 = nn.Sequential(nn.Linear(in3, W), act(), nn.Linear(W, W), act())
        self.alpha_branch = nn.Sequential(nn.Linear(W, 1))
        self.color_branch = nn.Sequential(nn.Linear(inc, C), act(), nn.Linear(C, C), act(),
                                          nn.Linear(C, C), act())

    def activations(self):
        """The seven LeakyReLUs in evaluation order: block1 (2), block3 (2) on pair rows; color_branch (3) on sample rows."""
        return [self.block1[1], self.block1[3], self.block3[1], self.block3[3],
                self.color_branch[1], self.color_branch[3], self.color_branch[5]]

    def freeze_pattern(self, pair_signs, sample_signs, record=True):
        """pair_signs [4,M,W], sample_signs [3,S,C] (bool / uint8; pn_agg_activation_signs layout) or (None, None)."""
        acts = self.activations()
        for i, a in enumerate(acts):
            src = pair_signs if i < 4 else sample_signs
            a.pattern = None if src is None else torch.as_tensor(src[i if i < 4 else i - 4]).bool().cpu()
            a.record = record

    def pattern_disagreement(self):
        """After a frozen forward with record=True: (units whose frozen pattern differs from sign(pre-activation), total
        units, max |pre| / rms(layer) over the differing units)."""
        n_diff, n_all, worst = 0, 0, 0.0
        for a in self.activations():
            if a.pattern is None or a.last_pre is None:
                continue
            x = a.last_pre.double()
            diff = a.pattern != (x > 0)
            n_diff += int(diff.sum())
            n_all += x.numel()
            if diff.any():
                rms = x.pow(2).mean().sqrt().clamp_min(1e-30)
                worst = max(worst, float((x.abs()[diff] / rms).max()))
        return n_diff, n_all, worst

    def forward(self, sampled_color, sampled_Rw2c, sampled_dir, sampled_conf, sampled_embedding,
                sampled_xyz_pers, sampled_xyz, sample_pnt_mask, sample_loc, sample_loc_w,
                sample_ray_dirs, vsize, grid_vox_sz):
        opt = self.opt
        C1 = opt.shading_color_channel_num + 1
        sample_valid = torch.any(sample_pnt_mask, dim=-1).view(-1)                     # :743
        shape3 = sample_loc_w.shape[:-1]
        if sample_valid.numel() == 0 or int(sample_valid.sum()) == 0:                   # :746-748
            return torch.zeros(shape3 + (C1,)), sample_valid.view(shape3), None, None
        B, R, SR, K = sample_pnt_mask.shape

        # 6-D distance vectors, agg_dist_pers == 20 (:775-783)
        pz = sampled_xyz_pers[..., 2]
        sz = sample_loc[:, :, :, None, 2]
        d_pers = torch.stack([sampled_xyz_pers[..., 0] * pz - sample_loc[:, :, :, None, 0] * sz,
                              sampled_xyz_pers[..., 1] * pz - sample_loc[:, :, :, None, 1] * sz,
                              pz - sz], dim=-1)
        dists = torch.cat([sampled_xyz - sample_loc_w[..., None, :], d_pers], dim=-1)

        # inverse-distance weights (:421-429) normalised over K (:803-804)
        w = 1.0 / torch.clamp(torch.norm(dists[..., :3], dim=-1), min=1e-6)
        w = sample_pnt_mask * w
        w = w / torch.clamp(torch.sum(w, dim=-1, keepdim=True), min=1e-8)
        # straight-through clamp of confidence (:724-726, 809-811)
        c0 = sampled_conf[..., 0]
        conf = c0 - (c0 - torch.clamp(c0, min=1e-4, max=1.0)).detach()
        wc = (w * conf).view(B * R * SR, K, 1)

        pmask = sample_pnt_mask.view(-1)
        Rt = sampled_Rw2c.transpose(-1, -2)                                            # :492
        view = sample_ray_dirs.view(-1, 3) @ Rt                                        # :506
        view_pe = positional_encoding(view, opt.num_viewdir_freqs, ori=True)           # :507-509
        ori_view, view_pe = view_pe[..., :3], view_pe[..., 3:]
        view_pe = view_pe[sample_valid, :]                                             # :512

        dflat = dists.view(-1, 6)[pmask, :]                                            # :521-523
        if opt.dist_xyz_deno != 0.0:                                                   # :524-525
            dflat = dflat / float(opt.dist_xyz_deno * np.linalg.norm(vsize))
        dflat = torch.cat([dflat[:, :3] @ Rt, dflat[:, 3:]], dim=-1)                   # :526
        d_pe = positional_encoding(dflat, opt.dist_xyz_freq)                           # :527-529
        emb = sampled_embedding.reshape(-1, sampled_embedding.shape[-1])[pmask, :]      # :530-534
        x = torch.cat([emb, positional_encoding(emb, opt.num_feat_freqs), d_pe], dim=-1)  # :536-538
        x = self.block1(x)                                                             # :547
        col = sampled_color.reshape(-1, 3)[pmask, :]                                   # :557-561
        pd = sampled_dir.reshape(-1, 3)[pmask, :] @ Rt                                 # :563-566
        ov = ori_view[:, None, :].repeat(1, K, 1).view(-1, 3)[pmask, :]                # :567-569
        x = torch.cat([x, col, pd - ov, torch.sum(pd * ov, dim=-1, keepdim=True)], dim=-1)  # :570
        x = self.block3(x)                                                             # :571

        f_full = torch.zeros([B * R * SR * K, x.shape[-1]], dtype=x.dtype)
        f_full[pmask, :] = x                                                           # :622-624 (order 1: :575-579)
        feat = torch.sum(f_full.view(B * R * SR, K, -1) * wc, dim=-2)[sample_valid, :]  # :627-628 (order 1: :580)
        if opt.agg_intrp_order == 1:          # interpolate the features, then ONE density per sample (:573-588)
            alpha = F.softplus(self.alpha_branch(feat) - 1)
        else:                                 # order 2: a density per neighbour, interpolated (:603-614)
            a = F.softplus(self.alpha_branch(x) - 1)                                   # :262-265, 605
            a_full = torch.zeros([B * R * SR * K, 1], dtype=a.dtype)
            a_full[pmask, :] = a                                                       # :608-610
            alpha = torch.sum(a_full.view(B * R * SR, K, 1) * wc, dim=-2)[sample_valid, :]  # :613-614
        colour = self.color_branch(torch.cat([feat, view_pe], dim=-1))                 # :635-636 (order 2 in this fork: no sigmoid)
        if opt.agg_intrp_order == 1:
            colour = torch.sigmoid(colour) * (1 + 2 * 0.001) - 0.001                   # :596 raw2out_color (:269-273, act_super)
        out = torch.zeros([B * R * SR, C1], dtype=x.dtype)
        out[sample_valid] = torch.cat([alpha, colour], dim=-1)                         # :644-646
        ret_w = (opt.sparse_loss_weight > 0) or ("conf_coefficient" in opt.zero_one_loss_items) or opt.prob != 0
        return (out.view(shape3 + (C1,)), sample_valid.view(shape3),
                w if ret_w else None, conf if ret_w else None)                         # :814-816


# --------------------------------------------------------------------------
# ray_dist / compositing / fill_invalid
# --------------------------------------------------------------------------
def ray_dist_from_samples(sample_loc, sample_valid, vsize, raydist_mode_unit=1):
    """models/neural_points_volumetric_model.py:293-301."""
    z = torch.cummax(sample_loc[..., 2], dim=-1)[0]
    d = torch.cat([z[..., 1:] - z[..., :-1],
                   torch.full((z.shape[0], z.shape[1], 1), vsize[2], dtype=z.dtype)], dim=-1)
    bad = d < 1e-8
    if raydist_mode_unit > 0:
        bad = torch.logical_or(bad, d > 2 * vsize[2])
    bad = bad.to(d.dtype)
    d = d * (1.0 - bad) + bad * vsize[2]
    return d * sample_valid.to(d.dtype)


def ray_march(ray_dist, ray_valid, ray_features, bg_color=None):
    """models/rendering/diff_ray_marching.py:509-555 with radiance_render (diff_render_func.py:48-50)
    and alpha_blend (:36-37).  Returns the reference's 7-tuple."""
    point_color = ray_features[..., 1:]
    sigma = ray_features[..., 0] * ray_valid.to(ray_features.dtype)
    opacity = 1 - torch.exp(-sigma * ray_dist)
    T = torch.cumprod(1.0 - opacity + 1e-10, dim=-1)
    bg_T = T[:, :, [-1]]
    T = torch.cat([torch.ones(opacity.shape[0:2] + (1,), dtype=T.dtype), T[:, :, :-1]], dim=-1)
    bw = (opacity * T)[..., None]
    ray_color = torch.sum(point_color * bw, dim=-2)
    if bg_color is not None:
        ray_color = ray_color + bg_color.to(bg_T.dtype).view(bg_T.shape[0], 1, -1) * bg_T
    return ray_color, point_color, opacity, T, bw, bg_T, 1 * bg_T


def fill_invalid(ray_mask, bg_color, coarse_raycolor, coarse_is_background, coarse_point_opacity,
                 queried_shading):
    """models/neural_points_volumetric_model.py:354-389 (tone map 'off')."""
    B, OR = ray_mask.shape
    idx = torch.nonzero(ray_mask)
    dt = coarse_raycolor.dtype
    is_bg = torch.ones([B, OR, 1], dtype=dt)
    is_bg[idx[..., 0], idx[..., 1], :] = coarse_is_background
    colour = torch.ones([B, OR, coarse_raycolor.shape[-1]], dtype=dt) * bg_color[None, ...]
    colour[idx[..., 0], idx[..., 1], :] = coarse_raycolor
    opac = torch.zeros([B, OR, coarse_point_opacity.shape[2]], dtype=dt)
    opac[idx[..., 0], idx[..., 1], :] = coarse_point_opacity
    shading = torch.ones([B, OR, queried_shading.shape[2]], dtype=dt)
    shading[idx[..., 0], idx[..., 1], :] = queried_shading
    return dict(coarse_is_background=is_bg, coarse_mask=1 - is_bg, coarse_raycolor=colour,
                coarse_point_opacity=opac, queried_shading=shading)
