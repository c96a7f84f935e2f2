"""Drop-in ``lighting_fast_querier`` (world-coordinate voxel neighbour query) on the C ABI.

Mirrors models/neural_points/query_point_indices_worldcoords.py of the reference:
class ``lighting_fast_querier`` (:31-721) with ``__init__(device, opt)``, ``clean_up()``,
``get_hyperparameters`` (:48-81), ``query_points`` (:84-99), ``w2pers`` (:102-109).
What changes underneath: no pycuda context / JIT; the occupancy grid and point buckets are a CSR
built once per point set (cached on the xyz tensor's identity + version, where the reference
rebuilds them every forward, :626), candidate samples are tested against a bitset, and the K-NN is
a warp-per-sample kernel.  Results are bit-identical to the reference under the canonical
schedule for scenes that do not overflow ``P`` / ``max_o``.
"""
import ctypes
import warnings

import numpy as np
import torch

from . import _lib
from ._lib import GridParams
from .query_host import grid_hyperparameters


_EDGE_CACHE = {}


def _depth_edges(dev, D, near, far):
    """[1, D+1] bin edges near*(1-t)+far*t of :369-372; a pure function of (D, near, far), kept per device so that a
    training loop does not re-issue the five tiny torch ops every batch."""
    key = (str(dev), D, near, far)
    e = _EDGE_CACHE.get(key)
    if e is None:
        if len(_EDGE_CACHE) > 64:
            _EDGE_CACHE.clear()
        # evaluated by the host's torch ops (the canonical chain) and uploaded once per (D, near, far)
        t = torch.linspace(0, 1, D + 1).view(1, -1)
        e = near * (1 - t) + far * t
        e = (e.to(dev), (e[..., 1:] - e[..., :-1]).reshape(-1).contiguous().to(dev))
        _EDGE_CACHE[key] = e
    return e


def middle_point_ts(raydir, D, near, far, jitter=0.0, uniform=None):
    """Depths of the candidate samples: the scalar chain of near_far_linear_ray_generation
    (models/rendering/diff_ray_marching.py:369-386) as the reference's torch ops evaluate it on the HOST (the canonical
    chain: fp32 ops rounded one by one, the cumsum accumulated in double and rounded per element - torch's device cumsum
    is a tree scan whose association is not reproducible).  Without jitter every ray has the same depths and one row
    [1,1,D] is returned; with jitter [1,R,D].  On the device both come from ONE kernel (pn_sample_depths), bit-identical
    to the host chain (tests/test_gpu_query.py); ``uniform`` stands for the torch.rand draw of :373."""
    dev = raydir.device
    t, base = _depth_edges(dev, int(D), float(near), float(far))
    R = raydir.shape[1]
    if jitter != 0.0 and uniform is None:
        uniform = torch.rand((raydir.shape[0], R, D), device=dev)
    if raydir.is_cuda and raydir.shape[0] == 1:
        rows = R if jitter != 0.0 else 1
        u = uniform.reshape(R, D).float().contiguous() if jitter != 0.0 else None
        ts = torch.empty((1, rows, D), dtype=torch.float32, device=dev)
        _lib.check(_lib.lib().pn_sample_depths(_lib.ptr(base), _lib.ptr(u), rows, D, float(near), float(jitter), _lib.ptr(ts),
                                               _lib.stream()), "pn_sample_depths")
        return ts
    if jitter == 0.0:
        seg = (t[..., 1:] - t[..., :-1]).view(1, 1, D) * torch.ones((1, 1, D), device=dev)
        rows = 1
    else:
        seg = (t[..., 1:] - t[..., :-1]) * (1 + jitter * (uniform - 0.5))
        rows = R
    ends = torch.cumsum(seg, dim=2)
    ends = torch.cat([torch.zeros((1, rows, 1), device=dev), ends], dim=2)
    ends = near + ends
    return ((ends[:, :, :-1] + ends[:, :, 1:]) / 2).contiguous()


class VoxelGrid:
    """Device-resident grid of one point set: cell_start / bucket_pidx / occ_bits / info (pointnerf_b200.h section 1)."""

    def __init__(self, xyz, hp, P, max_o):
        lib = _lib.lib()
        dev = xyz.device
        self.hp = hp
        dims = [int(v) for v in hp["scaled_vdim"]]
        self.V = dims[0] * dims[1] * dims[2]
        self.N = int(xyz.shape[0])
        gp = GridParams()
        for i in range(3):
            gp.shift[i] = float(hp["ranges"][i])
            gp.vsize[i] = float(hp["scaled_vsize"][i])
            gp.dims[i] = dims[i]
            gp.dilate[i] = int(hp["query_size"][i])
        gp.P, gp.max_o = int(P), int(max_o if max_o is not None else 2 ** 31 - 1)
        self.params = gp
        self.cell_start = torch.empty(self.V + 1, dtype=torch.int32, device=dev)
        self.bucket_pidx = torch.empty(max(self.N, 1), dtype=torch.int32, device=dev)
        self.occ_bits = torch.empty((self.V + 31) // 32, dtype=torch.int32, device=dev)
        self.info = torch.empty(_lib.GRID_INFO_LEN, dtype=torch.int32, device=dev)
        ws_bytes = lib.pn_grid_workspace_bytes(self.N, self.V)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        _lib.check(lib.pn_grid_build(_lib.ptr(xyz), self.N, ctypes.byref(gp), _lib.ptr(self.cell_start),
                                     _lib.ptr(self.bucket_pidx), _lib.ptr(self.occ_bits), _lib.ptr(self.info),
                                     _lib.ptr(ws), ws_bytes, _lib.stream()), "pn_grid_build")
        self._info_host = None

    def info_host(self):
        if self._info_host is None:
            self._info_host = self.info.cpu().numpy()
        return self._info_host


def ray_sample(grid, campos, raydir, ts, SR, want_t=False):
    """campos [3], raydir [R,3], ts [rows,D] -> sample_loc_w [R,SR,3], hit_count [R] (, sample_t [R,SR])."""
    lib = _lib.lib()
    R, D = int(raydir.shape[0]), int(ts.shape[-1])
    dev = raydir.device
    loc = torch.empty((R, SR, 3), dtype=torch.float32, device=dev)
    hit = torch.empty((R,), dtype=torch.int32, device=dev)
    st = torch.empty((R, SR), dtype=torch.float32, device=dev) if want_t else None
    _lib.check(lib.pn_ray_sample(_lib.ptr(campos), _lib.ptr(raydir), _lib.ptr(ts), int(ts.shape[0]), R, D, SR,
                                 ctypes.byref(grid.params), _lib.ptr(grid.occ_bits), _lib.ptr(loc), _lib.ptr(st),
                                 _lib.ptr(hit), _lib.stream()), "pn_ray_sample")
    return (loc, hit, st) if want_t else (loc, hit)


def knn_query(grid, xyz, sample_loc_w, hit_count, K, n_layers, radius_limit2, count_candidates=False):
    """-> sample_pidx int32 [R,SR,K], ray_any int32 [R] (, candidates visited)."""
    lib = _lib.lib()
    R, SR = int(sample_loc_w.shape[0]), int(sample_loc_w.shape[1])
    dev = xyz.device
    pidx = torch.empty((R, SR, K), dtype=torch.int32, device=dev)
    ray_any = torch.empty((R,), dtype=torch.int32, device=dev)
    cand = torch.zeros(1, dtype=torch.int64, device=dev) if count_candidates else None
    _lib.check(lib.pn_knn_query(_lib.ptr(xyz), ctypes.byref(grid.params), _lib.ptr(grid.cell_start),
                                _lib.ptr(grid.bucket_pidx), _lib.ptr(grid.info), _lib.ptr(sample_loc_w),
                                _lib.ptr(hit_count), R, SR, int(K), int(n_layers), float(radius_limit2),
                                _lib.ptr(pidx), _lib.ptr(ray_any), _lib.ptr(cand), _lib.stream()), "pn_knn_query")
    return (pidx, ray_any, cand) if count_candidates else (pidx, ray_any)


def w2pers(xyz_w, camrotc2w, campos):
    """(x/z, y/z, z) camera-frame coordinates of [...,3] world positions; camrotc2w [3,3], campos [3]."""
    lib = _lib.lib()
    flat = xyz_w.reshape(-1, 3).contiguous()
    out = torch.empty_like(flat)
    if flat.shape[0] == 0:
        return out.view(xyz_w.shape)
    _lib.check(lib.pn_w2pers(_lib.ptr(flat), int(flat.shape[0]), _lib.ptr(camrotc2w.contiguous()),
                             _lib.ptr(campos.contiguous()), _lib.ptr(out), _lib.stream()), "pn_w2pers")
    return out.view(xyz_w.shape)


class lighting_fast_querier:
    """Same constructor, methods and return structure as the reference class (worldcoords.py:31-99)."""

    def __init__(self, device, opt):
        self.device = device
        self.opt = opt
        self.inverse = opt.inverse
        self.count = 0
        self._grid = None
        self._grid_key = None
        self._grid_xyz = None       # the tensor the cached grid was built from: held so that its address cannot be recycled
        _lib.lib()  # fail now, loudly, if the CUDA library is missing

    def clean_up(self):
        self._grid = None
        self._grid_key = None
        self._grid_xyz = None

    # -- grid cache -----------------------------------------------------------------------------
    def _key(self, xyz):
        o = self.opt
        return (xyz.data_ptr(), xyz._version, tuple(xyz.shape), tuple(o.vsize), tuple(o.vscale), tuple(o.kernel_size),
                tuple(o.query_size), tuple(o.ranges) if o.ranges is not None else None, o.P, o.max_o)

    def get_hyperparameters(self, vsize_np, point_xyz_w_tensor, ranges=None):
        """worldcoords.py:48-81: bbox of the points (clipped to opt.ranges, padded), voxel dims, radius."""
        flat = point_xyz_w_tensor.reshape(-1, 3)
        mm = torch.stack([flat.min(dim=0)[0], flat.max(dim=0)[0]]).cpu().numpy()  # one small D2H per point set
        return grid_hyperparameters(self.opt, mm[0], mm[1])

    def grid_for(self, xyz):
        """xyz [N,3] float32 CUDA.  Rebuilds only when the point tensor (storage address + autograd version) or the grid
        options changed.  The querier keeps a reference to the keyed tensor, so the caching allocator cannot hand its
        address to a different cloud while the grid is cached (the reference rebuilds on every forward, :626).  In-place
        edits that bypass the version counter (``.data``, raw pointers) need ``clean_up()``."""
        key = self._key(xyz)
        if self._grid is None or key != self._grid_key:
            hp = self.get_hyperparameters(self.opt.vsize, xyz[None], ranges=self.opt.ranges)
            self._grid = VoxelGrid(xyz, hp, self.opt.P, self.opt.max_o)
            self._grid_key = key
            self._grid_xyz = xyz
            info = self._grid.info_host()
            if info[_lib.GI_OVER_MAX_O] or info[_lib.GI_OVER_P]:
                warnings.warn("pointnerf_b200: %d occupied voxels (max_o=%s), %d voxels over P=%d points: the reference "
                              "resolves these overflows with a time-seeded random choice; this implementation keeps "
                              "every voxel and the first P points by index" %
                              (info[_lib.GI_N_OCC], self.opt.max_o, info[_lib.GI_OVER_P], self.opt.P))
        return self._grid

    # -- the reference entry point -------------------------------------------------------------------
    def query_points(self, pixel_idx_tensor, point_xyz_pers_tensor, point_xyz_w_tensor, actual_numpoints_tensor, h, w,
                     intrinsic, near_depth, far_depth, ray_dirs_tensor, cam_pos_tensor, cam_rot_tensor, uniform=None):
        opt = self.opt
        if opt.inverse > 0:
            raise NotImplementedError("disparity-linear sampling (opt.inverse > 0) is outside the hot path (SURVEY.md 8)")
        assert point_xyz_w_tensor.shape[0] == 1, "B = 1 only, like every caller of the reference"
        near_depth, far_depth = np.asarray(near_depth).item(), np.asarray(far_depth).item()
        xyz = point_xyz_w_tensor[0].contiguous()
        raw = self.query_raw(xyz, ray_dirs_tensor, cam_pos_tensor, near_depth, far_depth, uniform=uniform)
        grid, loc_w, pidx, ray_any = raw["grid"], raw["sample_loc_w"], raw["sample_pidx"], raw["ray_any"]
        sel = torch.nonzero(ray_any, as_tuple=False).squeeze(1)         # host sync: data-dependent R'' (:716)
        ray_mask_tensor = (ray_any > 0).to(torch.int8)[None]
        sample_pidx_tensor = pidx.index_select(0, sel)[None]
        sample_loc_w_tensor = loc_w.index_select(0, sel)[None]
        sample_ray_dirs_tensor = ray_dirs_tensor[0].index_select(0, sel)[None, :, None, :] \
            .expand(-1, -1, opt.SR, -1).contiguous()
        sample_loc_tensor = self.w2pers(sample_loc_w_tensor, cam_rot_tensor, cam_pos_tensor)
        return (sample_pidx_tensor, sample_loc_tensor, sample_loc_w_tensor, sample_ray_dirs_tensor, ray_mask_tensor,
                opt.vsize, grid.hp["ranges"])

    def query_raw(self, xyz, ray_dirs_tensor, cam_pos_tensor, near_depth, far_depth, uniform=None, ts=None,
                  count_candidates=False):
        """Uncompacted device results [R,SR,...] with no host sync (used by the fused pipeline)."""
        opt = self.opt
        grid = self.grid_for(xyz)
        if ts is None:
            ts = middle_point_ts(ray_dirs_tensor, opt.z_depth_dim, near_depth, far_depth,
                                 jitter=0.3 if opt.is_train > 0 else 0.0, uniform=uniform)
        ts2 = ts.reshape(-1, ts.shape[-1]).contiguous()
        raydir = ray_dirs_tensor[0].contiguous()
        if raydir.shape[0] == 0:                                          # no rays: nothing to launch
            dev = xyz.device
            res = dict(grid=grid, sample_loc_w=torch.zeros((0, opt.SR, 3), device=dev), ts=ts,
                       hit_count=torch.zeros((0,), dtype=torch.int32, device=dev),
                       sample_pidx=torch.zeros((0, opt.SR, opt.K), dtype=torch.int32, device=dev),
                       ray_any=torch.zeros((0,), dtype=torch.int32, device=dev))
            if count_candidates:
                res["candidates"] = torch.zeros(1, dtype=torch.int64, device=dev)
            return res
        loc_w, hit = ray_sample(grid, cam_pos_tensor[0].contiguous(), raydir, ts2, opt.SR)
        n_layers = (int(opt.kernel_size[0]) + 1) // 2                     # only kernel_size[0] (:482)
        out = knn_query(grid, xyz, loc_w, hit, opt.K, n_layers, float(grid.hp["radius_limit2"]),
                        count_candidates=count_candidates)
        res = dict(grid=grid, sample_loc_w=loc_w, hit_count=hit, sample_pidx=out[0], ray_any=out[1], ts=ts)
        if count_candidates:
            res["candidates"] = out[2]
        return res

    def w2pers(self, point_xyz_w, camrotc2w, campos):
        return w2pers(point_xyz_w, camrotc2w[0], campos[0])
