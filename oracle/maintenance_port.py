"""Torch-CPU restatement of the point-cloud maintenance steps between passes.

TEST INFRASTRUCTURE (see oracle/__init__.py).  The reference modules holding these steps import pycuda at module
level (models/neural_points/neural_points.py:3-4, neural_points_volumetric_model_ori.py through its base class), so
they cannot be imported here; the few lines are restated with the same torch ops, each citing its source.  Parity
unpinned by reference-run outputs for this file: the checks are against this restatement only.
"""
import torch


def probe_outputs(coarse_point_opacity, sample_loc_w, weight, conf_coefficient, sampled_xyz, sampled_color, sampled_dir,
                  sampled_conf, sampled_embedding):
    """models/neural_points_volumetric_model_ori.py:351-372 (the ``opt.prob == 1`` block) on compacted rays.
    coarse_point_opacity [1,R,SR]; sample_loc_w [1,R,SR,3]; weight, conf_coefficient [1,R,SR,K]; sampled_* [1,R,SR,K,*]."""
    out = {}
    top, ind = torch.max(coarse_point_opacity, dim=-1, keepdim=True)              # :354  [1,R,1]
    out["ray_max_shading_opacity"] = top
    out["ray_max_slot"] = ind[..., 0]

    def at_max(t):
        """t [1,R,SR,...] -> t at the max-opacity sample of every ray [1,R,...] (the torch.gather calls of :357-367)."""
        idx = ind.view(ind.shape + (1,) * (t.dim() - 3)).expand((-1, -1, -1) + tuple(t.shape[3:]))
        return torch.gather(t, 2, idx).squeeze(2)

    loc = at_max(sample_loc_w)                                                   # :357
    out["ray_max_sample_loc_w"] = loc
    w = at_max(weight * conf_coefficient)[..., None]                             # :358  [1,R,K,1]
    nb_xyz = at_max(sampled_xyz)                                                 # :361  [1,R,K,3]
    out["ray_max_far_dist"] = torch.min(torch.norm(nb_xyz - loc[..., None, :], dim=-1), dim=-1, keepdim=True)[0]   # :362
    out["shading_avg_color"] = torch.sum(at_max(sampled_color) * w, dim=-2)       # :364,369
    out["shading_avg_dir"] = torch.sum(at_max(sampled_dir) * w, dim=-2)           # :365,370
    out["shading_avg_conf"] = torch.sum(at_max(sampled_conf) * w, dim=-2)         # :366,371
    out["shading_avg_embedding"] = torch.sum(at_max(sampled_embedding) * w, dim=-2)   # :367,372
    return out


def unmask(ray_mask, outputs):
    """models/neural_points_volumetric_model_ori.py:131-137: scatter [1,R'',...] results into zero tensors over all R rays."""
    idx = torch.nonzero(ray_mask > 0)
    full = {}
    for k, v in outputs.items():
        t = torch.zeros((ray_mask.shape[0], ray_mask.shape[1]) + tuple(v.shape[2:]), dtype=v.dtype)
        t[idx[:, 0], idx[:, 1]] = v[0]
        full[k] = t
    return full


def prune(tables, thresh):
    """models/neural_points/neural_points.py:350-373: boolean-mask every table with points_conf >= thresh."""
    mask = tables["points_conf"][0, ..., 0] >= thresh
    out = dict(tables)
    out["xyz"] = tables["xyz"][mask, :]
    for k in ("points_embeding", "points_conf", "points_dir", "points_color"):
        out[k] = tables[k][:, mask, :]
    return out


def grow_points(tables, add_xyz, add_embedding, add_color, add_dir, add_conf):
    """models/neural_points/neural_points.py:376-402: concatenate the new rows behind every table."""
    out = dict(tables)
    out["xyz"] = torch.cat([tables["xyz"], add_xyz], dim=0)
    for k, a in (("points_embeding", add_embedding), ("points_conf", add_conf), ("points_dir", add_dir), ("points_color", add_color)):
        out[k] = torch.cat([tables[k], a[None, ...]], dim=1)
    return out
