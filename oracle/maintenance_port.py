"""Torch-CPU restatement of the point-cloud maintenance steps between passes.

TEST INFRASTRUCTURE (see oracle/__init__.py).  The reference modules holding these steps import pycuda at module
level (models/neural_points/neural_points.py:3-4, neural_points_volumetric_model_ori.py through its base class), so
they cannot be imported here; the few lines are restated with the same torch ops, each citing its source.  Pinned in
the build container by executing the reference's own source fragments on CPU tensors (oracle/ref_fragments.py;
tests/test_oracle.py: prune / grow_points / the opt.prob block / unmask bit for bit; the voxel down-sampling through the
reference's call site with torch_scatter 2.0.8's two calls restated - that dependency is absent).
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


def construct_vox_points_closest(xyz_val, vox_res, space_min=None, space_max=None):
    """models/mvs/mvs_utils.py:537-561 with torch_scatter's two calls restated: ``scatter_mean`` = serial index_add_ in
    ascending point index / count, ``scatter_min`` (arg) = first index of the smallest residual per voxel.  With a box the
    filter applies to the coordinates too (the reference only works when nothing is filtered).
    Returns (xyz_centroid [U,3], sparse_grid_idx int32 [U,3], min_idx int64 [U], inv_idx int64 [N], residual [N])."""
    import numpy as np
    if space_min is None:
        xyz_min, xyz_max = torch.min(xyz_val, dim=-2)[0], torch.max(xyz_val, dim=-2)[0]
        space_edge = torch.max(xyz_max - xyz_min) * 1.05
        xyz_mid = (xyz_max + xyz_min) / 2
        space_min = xyz_mid - space_edge / 2
    else:
        space_edge = space_max - space_min
        mask = torch.prod((xyz_val - space_min[None, ...]) * (space_max[None, ...] - xyz_val), dim=-1) > 0
        xyz_val = xyz_val[mask, :]
    construct_vox_sz = space_edge / vox_res
    xyz_shift = xyz_val - space_min[None, ...]
    coords = torch.floor(xyz_shift / construct_vox_sz[None, ...]).to(torch.int32)
    sparse_grid_idx, inv_idx = torch.unique(coords, dim=0, return_inverse=True)
    U = sparse_grid_idx.shape[0]
    acc = np.zeros((U, 3), np.float32)
    np.add.at(acc, inv_idx.numpy(), xyz_val.numpy())          # unbuffered: a serial accumulation in ascending point index
    count = torch.from_numpy(np.bincount(inv_idx.numpy(), minlength=U).astype(np.float32))
    xyz_centroid = torch.from_numpy(acc) / count[:, None]
    xyz_residual = torch.norm(xyz_val - xyz_centroid[inv_idx, :], dim=-1)
    order = np.lexsort((np.arange(len(xyz_val)), xyz_residual.numpy(), inv_idx.numpy()))
    first = np.r_[True, inv_idx.numpy()[order][1:] != inv_idx.numpy()[order][:-1]]
    min_idx = torch.from_numpy(order[first].astype(np.int64))
    return xyz_centroid, sparse_grid_idx, min_idx, inv_idx, xyz_residual
