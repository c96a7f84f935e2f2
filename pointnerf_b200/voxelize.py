"""Voxel down-sampling of a point cloud on the C ABI (pn_voxel_downsample).

Mirrors ``construct_vox_points_closest`` of the reference (models/mvs/mvs_utils.py:537-561; the one real user of
``torch_scatter`` next to the hot path: ``scatter_mean`` / ``scatter_min`` after ``torch.unique``), called when a point
cloud is (re)voxelised (run/train_ft.py:141-146,693; train_waymo.py:531).  Same name, arguments and return triple.
There is no CPU path: CPU tensors raise.
"""
import ctypes
import math

import torch

from . import _lib


def construct_vox_points_closest(xyz_val, vox_res, partition_xyz=None, space_min=None, space_max=None, return_inverse=False):
    """xyz_val [N,3] float32 CUDA; vox_res: number (or [3] tensor); optional bounding box ``space_min`` / ``space_max`` [3].
    Returns (xyz_centroid [U,3], sparse_grid_idx int32 [U,3] in lexicographic order, min_idx int64 [U]): the occupied
    voxels, their centroids and the index of the point closest to each centroid - as the reference
    (``return_inverse=True`` appends inv_idx int32 [N], the voxel of every point: torch.unique's inverse).

    With a bounding box the reference filters ``xyz_val`` by the (open) box but keeps computing the voxel coordinates from
    the unfiltered array (:539,547,552), which only works when nothing is filtered; here the filter applies to both and
    ``min_idx`` indexes the FILTERED array like the reference's ``xyz_val``."""
    if partition_xyz is not None:
        raise NotImplementedError("partition_xyz is never passed by the reference's callers")
    lib = _lib.lib()
    xyz_val = xyz_val.detach().float()
    dev = xyz_val.device
    if space_min is None:
        xyz_min, xyz_max = torch.min(xyz_val, dim=-2)[0], torch.max(xyz_val, dim=-2)[0]      # :541-544
        space_edge = torch.max(xyz_max - xyz_min) * 1.05
        xyz_mid = (xyz_max + xyz_min) / 2
        space_min = xyz_mid - space_edge / 2
    else:
        space_min, space_max = space_min.to(dev).float(), space_max.to(dev).float()
        space_edge = space_max - space_min                                                     # :546-550
        mask = torch.prod((xyz_val - space_min[None, ...]) * (space_max[None, ...] - xyz_val), dim=-1) > 0
        xyz_val = xyz_val[mask, :]
    vox_sz = (space_edge / torch.as_tensor(vox_res, device=dev)).float().expand(3)            # :551
    # voxel-coordinate range of the (filtered) points, from their bbox: the map p -> floor((p - min) / size) is monotonic
    if xyz_val.shape[0]:
        ext = torch.stack([torch.min(xyz_val, dim=0)[0], torch.max(xyz_val, dim=0)[0]])
        crange = torch.floor((ext - space_min.expand(3)[None]) / vox_sz[None])
    else:
        crange = torch.zeros(2, 3, device=dev)
    host = torch.stack([space_min.expand(3).float(), vox_sz, crange[0], crange[1]]).cpu()    # one small D2H
    if not bool(torch.isfinite(host).all()):
        raise ValueError("construct_vox_points_closest: degenerate extent / voxel size (the reference yields NaN coordinates here)")
    smin = (ctypes.c_float * 3)(*[float(v) for v in host[0]])
    vsz = (ctypes.c_float * 3)(*[float(v) for v in host[1]])
    cbase = (ctypes.c_int32 * 3)(*[int(v) for v in host[2]])
    bits = max(1, int(math.ceil(math.log2(float((host[3] - host[2]).max()) + 2.0))))
    if bits > 21:
        raise NotImplementedError("voxel grid wider than 2^21 cells per axis")
    xyz_c = xyz_val.contiguous()
    N = int(xyz_c.shape[0])
    centroid = torch.empty((N, 3), dtype=torch.float32, device=dev)
    grid_idx = torch.empty((N, 3), dtype=torch.int32, device=dev)
    min_idx = torch.empty((N,), dtype=torch.int64, device=dev)
    inv_idx = torch.empty((N,), dtype=torch.int32, device=dev)
    counts = torch.zeros(2, dtype=torch.int32, device=dev)
    ws_bytes = lib.pn_voxel_downsample_workspace_bytes(N)
    ws = torch.empty(max(ws_bytes, 1), dtype=torch.uint8, device=dev)
    _lib.check(lib.pn_voxel_downsample(_lib.ptr(xyz_c), N, smin, vsz, cbase, bits, _lib.ptr(centroid), _lib.ptr(grid_idx), _lib.ptr(min_idx),
                                       _lib.ptr(inv_idx), counts.data_ptr(), counts.data_ptr() + 4, _lib.ptr(ws), ws_bytes,
                                       _lib.stream()), "pn_voxel_downsample")
    U, outside = (int(v) for v in counts.cpu())      # host sync: the number of occupied voxels sizes the result
    if outside:
        raise _lib.LibraryError("pn_voxel_downsample: %d points fall outside the %d-bit voxel key range" % (outside, bits))
    if return_inverse:
        return centroid[:U], grid_idx[:U], min_idx[:U], inv_idx
    return centroid[:U], grid_idx[:U], min_idx[:U]
