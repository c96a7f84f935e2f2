"""Drop-in ``lighting_fast_querier`` for ``wcoord_query = 0``: the perspective-frustum neighbour query on the C ABI.

Mirrors models/neural_points/query_point_indices.py of the reference: class
``lighting_fast_querier`` (:34-786) with ``__init__(device, opt)``, ``clean_up()``, ``get_hyperparameters`` (:52-81),
``query_points`` (:84-108), ``pers2w`` (:110-123), the ``gaussian`` / ``uniform`` / ``passfunc`` sample jitters
(:125-135) and ``query_grid_point_index`` (:617-786).  The voxel grid lives in the camera's perspective space and is
rebuilt for every view (as in the reference); what changes underneath: one CSR over the frustum cells instead of the
uint8 / int8 / int16 dense tables and the per-column slot table, a bitset walk instead of near_vox_full, a
warp-per-sample K-NN, no pycuda.  Results equal the reference's under the canonical schedule when no voxel holds more
than ``P`` points (pointnerf_b200.h section 3b); options whose reference behaviour is not reproducible raise.
"""
import ctypes
import warnings

import numpy as np
import torch

from . import _lib
from ._lib import FrustumQuery, GridParams


def frustum_hyperparameters(opt, h, w, intrinsic, near_depth, far_depth):
    """query_point_indices.py:52-81 for inverse = 0, same numpy expressions and dtypes (host arithmetic)."""
    x_rl, x_rh = -intrinsic[0, 2] / intrinsic[0, 0], (w - intrinsic[0, 2]) / intrinsic[0, 0]
    y_rl, y_rh = -intrinsic[1, 2] / intrinsic[1, 1], (h - intrinsic[1, 2]) / intrinsic[1, 1]
    z_r = far_depth - near_depth
    ranges = np.array([x_rl, y_rl, near_depth, x_rh, y_rh, far_depth], dtype=np.float32)
    vdim = np.array([w, h, opt.z_depth_dim], dtype=np.int32)
    vsize = np.array([(x_rh - x_rl) / vdim[0], (y_rh - y_rl) / vdim[1], z_r / vdim[2]], dtype=np.float32)
    vscale = np.array(opt.vscale, dtype=np.int32)
    scaled_vdim = np.ceil(vdim / vscale).astype(np.int32)
    scaled_vsize = (vsize * vscale).astype(np.float32)
    radius_limit = np.float32(opt.radius_limit_scale * max(vsize[0], vsize[1]))
    depth_limit = np.float32(opt.depth_limit_scale * vsize[2])
    return dict(radius_limit=radius_limit, depth_limit=depth_limit, ranges=ranges, vsize=vsize, vdim=vdim,
                scaled_vsize=scaled_vsize, scaled_vdim=scaled_vdim, vscale=vscale,
                kernel_size=np.asarray(opt.kernel_size, dtype=np.int32), query_size=np.asarray(opt.query_size, dtype=np.int32))


class lighting_fast_querier:
    """Same constructor, methods and return structure as the reference class (query_point_indices.py:34-108)."""

    def __init__(self, device, opt):
        self.device = device
        self.opt = opt
        self.inverse = opt.inverse
        self.last_info = None
        _lib.lib()  # fail now, loudly, if the CUDA library is missing

    def clean_up(self):
        self.last_info = None

    def get_hyperparameters(self, h, w, intrinsic, near_depth, far_depth):
        return frustum_hyperparameters(self.opt, h, w, intrinsic, near_depth, far_depth)

    def grid_report(self):
        """Statistics of the last view's grid (one small device read-back, so not part of ``query_points``): claimed voxels,
        points inside the frustum, the fullest voxel and how many voxels hold more than ``P`` points - there the reference
        keeps a time-seeded random subset (:404-411) and this implementation the first P by index."""
        if self.last_info is None:
            return None
        info = self.last_info.cpu().numpy()
        rep = dict(n_occ=int(info[_lib.GI_N_OCC]), n_inside=int(info[_lib.GI_N_INSIDE]), max_count=int(info[_lib.GI_MAX_COUNT]),
                   over_P=int(info[_lib.GI_OVER_P]))
        if rep["over_P"]:
            warnings.warn("pointnerf_b200: %d frustum voxels hold more than P=%d points (fullest: %d): the reference keeps a "
                          "time-seeded random subset there, this implementation the first P by index"
                          % (rep["over_P"], self.opt.P, rep["max_count"]))
        return rep

    def _check(self, hp):
        o = self.opt
        if o.inverse > 0:
            raise NotImplementedError("disparity-linear frustum depth (opt.inverse > 0) is not built (DESIGN.md 6)")
        if o.NN <= 0:
            raise NotImplementedError("opt.NN <= 0 selects the reference's time-seeded random neighbour pick "
                                      "(query_rand_along_ray, query_point_indices.py:416-492), which has no reproducible result")
        k, q = hp["kernel_size"], hp["query_size"]
        if q[0] < k[0] or q[1] < k[0] or q[2] < k[2]:
            raise NotImplementedError("query_size %s smaller than the cube kernel_size %s visits: the reference then only "
                                      "sees voxels that some other ray of the batch happened to mark" % (tuple(q), tuple(k)))

    # -- the kernels ----------------------------------------------------------------------------------
    def query_raw(self, pixel_idx, xyz_pers, hp, count_candidates=False):
        """pixel_idx int32 [R,2], xyz_pers float32 [N,3] (CUDA) -> uncompacted device results over all R rays, no host sync."""
        lib, opt = _lib.lib(), self.opt
        self._check(hp)
        dev = xyz_pers.device
        R, N, SR, K = int(pixel_idx.shape[0]), int(xyz_pers.shape[0]), int(opt.SR), int(opt.K)
        dims = [int(v) for v in hp["scaled_vdim"]]
        V = dims[0] * dims[1] * dims[2]
        gp = GridParams()
        for i in range(3):
            gp.shift[i] = float(hp["ranges"][i])
            gp.vsize[i] = float(hp["scaled_vsize"][i])
            gp.dims[i] = dims[i]
            gp.dilate[i] = int(hp["query_size"][i])
        gp.P, gp.max_o = int(opt.P), 2 ** 31 - 1
        ws_bytes = lib.pn_frustum_grid_workspace_bytes(N, V)
        # one allocation: cell_start | bucket | occ_bits | info (all rebuilt per view) + the build's workspace
        n_int = (V + 1) + max(N, 1) + (V + 31) // 32 + _lib.GRID_INFO_LEN
        store = torch.empty(n_int, dtype=torch.int32, device=dev)
        cell_start, bucket, occ_bits, info = torch.split(store, [V + 1, max(N, 1), (V + 31) // 32, _lib.GRID_INFO_LEN])
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        st = _lib.stream()
        _lib.check(lib.pn_frustum_grid_build(_lib.ptr(xyz_pers), N, ctypes.byref(gp), _lib.ptr(cell_start), _lib.ptr(bucket),
                                             _lib.ptr(occ_bits), _lib.ptr(info), _lib.ptr(ws), ws_bytes, st), "pn_frustum_grid_build")
        sample_z = torch.empty((R, SR), dtype=torch.int32, device=dev)
        hit = torch.empty((R,), dtype=torch.int32, device=dev)
        mask = torch.empty((R,), dtype=torch.int32, device=dev)
        pidx = torch.empty((R, SR, K), dtype=torch.int32, device=dev)
        loc = torch.empty((R, SR, 3), dtype=torch.float32, device=dev)
        cand = torch.zeros(1, dtype=torch.int64, device=dev) if count_candidates else None
        if R > 0:
            vs = (ctypes.c_int32 * 2)(int(hp["vscale"][0]), int(hp["vscale"][1]))
            _lib.check(lib.pn_frustum_sample(_lib.ptr(pixel_idx), R, SR, ctypes.byref(gp), vs, _lib.ptr(occ_bits),
                                             _lib.ptr(sample_z), _lib.ptr(hit), _lib.ptr(mask), st), "pn_frustum_sample")
            q = FrustumQuery()
            ray_vs = (hp["scaled_vsize"] / hp["vscale"]).astype(np.float32)            # :707 (float64 quotient, then float32)
            q.ray_vsize[0], q.ray_vsize[1] = float(ray_vs[0]), float(ray_vs[1])
            q.vscale[0], q.vscale[1] = int(hp["vscale"][0]), int(hp["vscale"][1])
            q.radius_limit2 = float(np.float32(hp["radius_limit"] ** 2))               # :753-754
            q.depth_limit2 = float(np.float32(hp["depth_limit"] ** 2))
            q.n_layers = (int(hp["kernel_size"][0]) + 1) // 2
            q.z_layers = (int(hp["kernel_size"][2]) + 1) // 2
            q.nn = int(opt.NN)
            _lib.check(lib.pn_frustum_knn_query(_lib.ptr(xyz_pers), ctypes.byref(gp), ctypes.byref(q), _lib.ptr(cell_start),
                                                _lib.ptr(bucket), _lib.ptr(pixel_idx), _lib.ptr(sample_z), _lib.ptr(hit), R, SR, K,
                                                _lib.ptr(pidx), _lib.ptr(loc), _lib.ptr(cand), st), "pn_frustum_knn_query")
        self.last_info = info.clone()          # 8 ints; a view would keep the whole per-view store alive until the next call
        res = dict(sample_pidx=pidx, sample_loc=loc, sample_z=sample_z, hit_count=hit, ray_mask=mask, grid_info=info,
                   occ_bits=occ_bits, cell_start=cell_start, bucket_pidx=bucket, dims=dims)
        if count_candidates:
            res["candidates"] = cand
        return res

    def query_grid_point_index(self, pixel_idx_tensor, point_xyz_pers_tensor, hp):
        """:617-786 -> sample_pidx [1,R',SR,K], sample_loc [1,R',SR,3], pixel_idx_cur [1,R',2], ray_mask int8 [1,R]."""
        pix = pixel_idx_tensor.reshape(-1, 2).to(torch.int32).contiguous()
        raw = self.query_raw(pix, point_xyz_pers_tensor.reshape(-1, 3).contiguous(), hp)
        sel = torch.nonzero(raw["ray_mask"], as_tuple=False).squeeze(1)          # host sync: data-dependent R' (:691-695)
        return (raw["sample_pidx"].index_select(0, sel)[None], raw["sample_loc"].index_select(0, sel)[None],
                pix.index_select(0, sel)[None], raw["ray_mask"].to(torch.int8)[None])

    # -- the reference entry point ------------------------------------------------------------------------
    def query_points(self, pixel_idx_tensor, point_xyz_pers_tensor, point_xyz_w_tensor, actual_numpoints_tensor, h, w,
                     intrinsic, near_depth, far_depth, ray_dirs_tensor, cam_pos_tensor, cam_rot_tensor, uniform=None):
        opt = self.opt
        assert point_xyz_pers_tensor.shape[0] == 1, "B = 1 only, like every caller of the reference"
        # h, w, near, far are used as given (the reference passes 0-d numpy arrays, neural_points.py:786, and numpy's own
        # promotion rules decide the rounding of vsize)
        hp = self.get_hyperparameters(h, w, np.asarray(intrinsic), near_depth, far_depth)
        self.radius_limit, self.depth_limit, self.ranges, self.vsize, self.vdim = \
            hp["radius_limit"], hp["depth_limit"], hp["ranges"], hp["vsize"], hp["vdim"]
        self.scaled_vsize, self.scaled_vdim, self.vscale = hp["scaled_vsize"], hp["scaled_vdim"], hp["vscale"]
        sample_pidx_tensor, sample_loc_tensor, pixel_idx_cur_tensor, ray_mask_tensor = \
            self.query_grid_point_index(pixel_idx_tensor, point_xyz_pers_tensor, hp)
        self.inverse = opt.inverse
        if opt.is_train:
            sample_loc_tensor = getattr(self, opt.shpnt_jitter)(sample_loc_tensor, hp["vsize"], uniform=uniform)
        sample_loc_w_tensor, sample_ray_dirs_tensor = self.pers2w(sample_loc_tensor, cam_rot_tensor, cam_pos_tensor)
        return (sample_pidx_tensor, sample_loc_tensor, sample_loc_w_tensor, sample_ray_dirs_tensor, ray_mask_tensor,
                hp["vsize"], hp["ranges"])

    def pers2w(self, point_xyz_pers, camrotc2w, campos):
        """:110-123, the same torch ops on the device (elementwise + a 3-term sum)."""
        x_pers = point_xyz_pers[..., 0] * point_xyz_pers[..., 2]
        y_pers = point_xyz_pers[..., 1] * point_xyz_pers[..., 2]
        z_pers = point_xyz_pers[..., 2]
        xyz_c = torch.stack([x_pers, y_pers, z_pers], dim=-1)
        xyz_w_shift = torch.sum(xyz_c[..., None, :] * camrotc2w, dim=-1)
        ray_dirs = xyz_w_shift / (torch.linalg.norm(xyz_w_shift, dim=-1, keepdims=True) + 1e-7)
        xyz_w = xyz_w_shift + campos[:, None, :]
        return xyz_w, ray_dirs

    # -- sample jitters (:125-135); ``uniform`` lets a test supply the random draw ------------------------
    def passfunc(self, input, vsize, uniform=None):
        return input

    def gaussian(self, input, vsize, uniform=None):
        B, R, SR, _ = input.shape
        jitters = torch.normal(mean=torch.zeros([B, R, SR], dtype=torch.float32, device=input.device),
                               std=torch.full([B, R, SR], float(vsize[2] / 4), dtype=torch.float32, device=input.device))
        input[..., 2] = input[..., 2] + torch.clamp(jitters, min=float(-vsize[2] / 2), max=float(vsize[2] / 2))
        return input

    def uniform(self, input, vsize, uniform=None):
        B, R, SR, _ = input.shape
        draw = uniform if uniform is not None else torch.rand([B, R, SR], dtype=torch.float32, device=input.device)
        input[..., 2] = input[..., 2] + (draw - 0.5) * float(vsize[2])
        return input
