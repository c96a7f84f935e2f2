"""Host arithmetic of the world-coordinate querier (no torch.cuda, no library): grid hyper-parameters of
query_point_indices_worldcoords.py:48-81.  Imported by pointnerf_b200.query (product) and by the synthetic-scene generator."""
import numpy as np


def voxel_coords(xyz, shift, scaled_vsize):
    """(int) floor((p - shift) / vsize) in IEEE float32, as worldcoords.py:265-267."""
    xyz = np.asarray(xyz, np.float32)
    q = (xyz - np.asarray(shift, np.float32)[None, :]) / np.asarray(scaled_vsize, np.float32)[None, :]
    return np.floor(q).astype(np.int64)


def grid_hyperparameters(opt, min_xyz, max_xyz):
    """Host arithmetic of ``get_hyperparameters`` (worldcoords.py:58-81) given the point bbox.

    min_xyz/max_xyz: float32 arrays [3] (the torch.min/max over the points).  The float32/float64
    mixing of the reference is kept because it decides ``scaled_vdim`` and ``shift``."""
    f32 = np.float32
    vsize = tuple(opt.vsize)
    vscale = np.array(opt.vscale, dtype=np.int32)
    scaled_vsize = (np.asarray(vsize, np.float64) * vscale).astype(f32)
    lo, hi = np.asarray(min_xyz, f32).copy(), np.asarray(max_xyz, f32).copy()
    rg = opt.ranges
    if rg is not None and not (rg[0] >= rg[3]):
        lo = np.maximum(lo, np.asarray(rg[:3], f32))
        hi = np.minimum(hi, np.asarray(rg[3:], f32))
    pad = (scaled_vsize.astype(np.float64) * np.asarray(opt.kernel_size, np.int64) / 2).astype(f32)
    lo = (lo - pad).astype(f32)
    hi = (hi + pad).astype(f32)
    vdim = (hi - lo).astype(f32).astype(np.float64) / np.asarray(vsize, np.float64)
    scaled_vdim = np.ceil(vdim / vscale).astype(np.int32)
    radius = f32(opt.radius_limit_scale * max(vsize[0], vsize[1]))
    return dict(ranges=np.concatenate([lo, hi]).astype(f32), scaled_vsize=scaled_vsize, scaled_vdim=scaled_vdim,
                radius_limit=radius, radius_limit2=f32(radius * radius),
                kernel_size=np.asarray(opt.kernel_size, np.int32), query_size=np.asarray(opt.query_size, np.int32))
