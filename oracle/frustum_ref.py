"""Host side of the perspective-frustum querier oracle, numpy / torch-CPU.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Restates the Python glue of the reference's second
``lighting_fast_querier`` (models/neural_points/query_point_indices.py): ``get_hyperparameters`` :52-81,
``query_points`` :84-108, ``pers2w`` :110-123, the ``uniform`` jitter :131-135 and ``query_grid_point_index`` :617-786,
for B = 1, around the reference's OWN four kernels compiled for the host (oracle/build_ref.py ->
oracle/_ref/libpn_ref_querier_frustum.so; serial, ascending thread index = the canonical schedule).  There is no
separate C port of these kernels: without the built library this oracle is unavailable, and the committed golden
vectors (tests/golden/frustum_*.npz, written by oracle/make_golden.py from this file) stand in.
"""
import ctypes
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libpn_ref_querier_frustum.so")
_lib = None
I, F_ = ctypes.c_int, ctypes.c_float
THREADS = 1024  # opt.gpu_maxthr default; only shapes the serial loop


def available():
    return os.path.isfile(REF_SO)


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(REF_SO)
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _grid(n):
    return int((n + THREADS - 1) / THREADS)


def get_hyperparameters(opt, h, w, intrinsic, near_depth, far_depth):
    """query_point_indices.py:52-81 (numpy, same expressions and dtypes)."""
    inverse = opt.inverse
    x_rl, x_rh = -intrinsic[0, 2] / intrinsic[0, 0], (w - intrinsic[0, 2]) / intrinsic[0, 0]
    y_rl, y_rh = -intrinsic[1, 2] / intrinsic[1, 1], (h - intrinsic[1, 2]) / intrinsic[1, 1]
    z_r = (far_depth - near_depth) if inverse == 0 else (1.0 / near_depth - 1.0 / far_depth)
    ranges = np.array([x_rl, y_rl, near_depth, x_rh, y_rh, far_depth], dtype=np.float32) if inverse == 0 else \
        np.array([x_rl, y_rl, 1.0 / far_depth, x_rh, y_rh, 1.0 / near_depth], dtype=np.float32)
    vdim = np.array([w, h, opt.z_depth_dim], dtype=np.int32)
    vsize = np.array([(x_rh - x_rl) / vdim[0], (y_rh - y_rl) / vdim[1], z_r / vdim[2]], dtype=np.float32)
    vscale = np.array(opt.vscale, dtype=np.int32)
    scaled_vdim = np.ceil(vdim / vscale).astype(np.int32)
    scaled_vsize = (vsize * vscale).astype(np.float32)
    radius_limit = opt.radius_limit_scale * max(vsize[0], vsize[1])
    depth_limit = opt.depth_limit_scale * vsize[2]
    return dict(radius_limit=np.float32(radius_limit), depth_limit=np.float32(depth_limit), ranges=ranges, vsize=vsize,
                vdim=vdim, scaled_vsize=scaled_vsize, scaled_vdim=scaled_vdim, vscale=vscale,
                kernel_size=np.asarray(opt.kernel_size, dtype=np.int32), query_size=np.asarray(opt.query_size, dtype=np.int32))


def query_grid_point_index(opt, hp, pixel_idx, xyz_pers, stages=None):
    """query_point_indices.py:617-786.  pixel_idx int32 [R,2] (x, y); xyz_pers float32 [N,3].
    -> sample_pidx int32 [R',SR,K], sample_loc float32 [R',SR,3], pixel_idx_cur [R',2], ray_mask int8 [R]."""
    L = lib()
    SR, K, P, NN, inverse = int(opt.SR), int(opt.K), int(opt.P), int(opt.NN), int(opt.inverse)
    assert K <= L.pn_ref_kn() and NN > 0
    xyz = np.ascontiguousarray(xyz_pers, dtype=np.float32)
    N = xyz.shape[0]
    pix = np.ascontiguousarray(pixel_idx, dtype=np.int32).reshape(-1, 2)
    R = pix.shape[0]
    dims = [int(v) for v in hp["scaled_vdim"]]
    pixel_size = dims[0] * dims[1]
    vol = pixel_size * dims[2]
    shift = np.ascontiguousarray(hp["ranges"][:3], dtype=np.float32)
    svs = np.ascontiguousarray(hp["scaled_vsize"], dtype=np.float32)
    sdim = np.ascontiguousarray(hp["scaled_vdim"], dtype=np.int32)
    vscale = np.ascontiguousarray(hp["vscale"], dtype=np.int32)
    ksz, qsz = np.ascontiguousarray(hp["kernel_size"]), np.ascontiguousarray(hp["query_size"])
    n_act = np.array([N], np.int32)

    coor_occ = np.zeros(dims, np.uint8)                                        # :635
    loc_cnt = np.zeros(dims, np.int8)
    near_id = np.full(dims[:2], dims[2], np.int32)
    far_id = np.full(dims[:2], -1, np.int32)
    L.pn_ref_get_occ_vox(I(_grid(N)), I(THREADS), _p(xyz), _p(n_act), I(1), I(N), _p(shift), _p(svs), _p(sdim), _p(qsz),
                         I(pixel_size), I(vol), _p(coor_occ), _p(loc_cnt), _p(near_id), _p(far_id), I(inverse))   # :640-656

    v2z = np.full(dims[:2] + [SR], -1, np.int16)                               # :662
    pixel_map = np.zeros(dims[:2], np.uint8)
    ray_mask = np.zeros(R, np.int8)
    if R > 0:
        L.pn_ref_near_vox_full(I(_grid(R)), I(THREADS), I(1), I(SR), _p(pix), I(R), _p(vscale), _p(sdim), I(pixel_size),
                               I(vol), _p(qsz), _p(pixel_map), _p(ray_mask), _p(coor_occ), _p(loc_cnt), _p(near_id), _p(far_id),
                               _p(v2z))                                         # :665-683
    if stages is not None:
        stages.update(coor_occ=coor_occ.copy(), near_id=near_id.copy(), far_id=far_id.copy(), v2z=v2z.copy())

    pix_cur = np.ascontiguousarray(pix[ray_mask > 0])                           # :691-692
    Rc = int(ray_mask.sum())                                                    # :695 (B = 1)
    loc = torch.from_numpy((loc_cnt > 0).astype(np.int8))                       # :697
    loc = loc * torch.cumsum(loc, dtype=torch.int8, dim=-1) - 1                 # :698 (int8 running sum, as the reference)
    loc_cnt = np.ascontiguousarray(loc.numpy())
    max_o = opt.max_o
    if max_o is None:
        max_o = int(loc_cnt.max()) + 1                                          # :701
    max_o = max(int(max_o), 1)
    if int(loc_cnt.max()) >= max_o:
        raise RuntimeError("frustum oracle: a pixel column needs %d voxels > max_o = %d (the reference indexes out of "
                           "bounds here)" % (int(loc_cnt.max()) + 1, max_o))

    vpc = np.zeros(dims[:2] + [max_o], np.int16)                                # :704
    v2p = np.zeros(dims[:2] + [max_o, P], np.int32)
    ray_vs = (svs / vscale).astype(np.float32)                                  # :707
    L.pn_ref_insert_vox_points(I(_grid(N)), I(THREADS), _p(xyz), _p(n_act), I(1), I(N), I(P), I(max_o), I(pixel_size),
                               I(vol), _p(shift), _p(sdim), _p(svs), _p(loc_cnt), _p(vpc), _p(v2p), I(inverse))   # :710-727
    if stages is not None:
        stages.update(loc=loc_cnt.copy(), vpc=vpc.copy(), max_o=max_o)

    pidx = np.full((Rc, SR, K), -1, np.int32)                                   # :734
    sloc = np.zeros((Rc, SR, 3), np.float32)
    if Rc > 0:
        L.pn_ref_query_along_ray(I(_grid(Rc * SR)), I(THREADS), _p(xyz), I(1), I(SR), I(Rc), I(max_o), I(P), I(K),
                                 I(pixel_size), I(vol), F_(float(np.float32(hp["radius_limit"] ** 2))),
                                 F_(float(np.float32(hp["depth_limit"] ** 2))), _p(shift), _p(sdim), _p(svs), _p(ray_vs),
                                 _p(vscale), _p(ksz), _p(pix_cur), _p(loc_cnt), _p(v2z), _p(vpc), _p(v2p), _p(pidx), _p(sloc),
                                 I(NN), I(inverse))                             # :742-772
    return pidx, sloc, pix_cur, ray_mask


def pers2w(point_xyz_pers, camrotc2w, campos):
    """:110-123 (torch CPU, same ops)."""
    x_pers = point_xyz_pers[..., 0] * point_xyz_pers[..., 2]
    y_pers = point_xyz_pers[..., 1] * point_xyz_pers[..., 2]
    z_pers = point_xyz_pers[..., 2]
    xyz_c = torch.stack([x_pers, y_pers, z_pers], dim=-1)
    xyz_w_shift = torch.sum(xyz_c[..., None, :] * camrotc2w, dim=-1)
    ray_dirs = xyz_w_shift / (torch.linalg.norm(xyz_w_shift, dim=-1, keepdims=True) + 1e-7)
    xyz_w = xyz_w_shift + campos[:, None, :]
    return xyz_w, ray_dirs


def query_points(opt, pixel_idx_tensor, point_xyz_pers_tensor, h, w, intrinsic, near_depth, far_depth, cam_pos_tensor,
                 cam_rot_tensor, uniform=None):
    """:84-108 -> the reference's 7-tuple (torch CPU tensors, B = 1).  ``uniform`` [1,R',SR] in [0,1) stands for the
    torch.rand draw of the ``uniform`` jitter (:131-135) when opt.is_train."""
    hp = get_hyperparameters(opt, h, w, intrinsic, near_depth, far_depth)
    pidx, sloc, pix_cur, ray_mask = query_grid_point_index(opt, hp, pixel_idx_tensor.reshape(-1, 2).numpy(),
                                                           point_xyz_pers_tensor.reshape(-1, 3).numpy())
    sample_loc = torch.from_numpy(sloc)[None]
    if opt.is_train:
        assert opt.shpnt_jitter == "uniform" and uniform is not None
        sample_loc[..., 2] = sample_loc[..., 2] + (uniform - 0.5) * hp["vsize"][2]
    sample_loc_w, sample_ray_dirs = pers2w(sample_loc, cam_rot_tensor, cam_pos_tensor)
    return (torch.from_numpy(pidx)[None], sample_loc, sample_loc_w, sample_ray_dirs, torch.from_numpy(ray_mask)[None],
            hp["vsize"], hp["ranges"])
