"""Host side of the neighbour-query oracle (numpy / torch-CPU), around two kernel back ends.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Restates the Python glue of
``lighting_fast_querier`` (models/neural_points/query_point_indices_worldcoords.py):
``get_hyperparameters`` :48-81, ``query_points`` :84-99, ``w2pers`` :102-109, ``build_occ_vox``
:546-611 and ``query_grid_point_index`` :614-721, for B = 1.

Kernel back ends (same six entry points):
  * ``"ref"``  - oracle/_ref/libpn_ref_querier.so: the reference's own kernel string compiled for
                 the host by oracle/build_ref.py (build container; travels as a built file);
  * ``"port"`` - oracle/_build/libpn_port_querier.so from oracle/querier_port.c (always buildable).
"""
import ctypes
import os
import subprocess

import numpy as np
import torch

from . import torch_port

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libpn_ref_querier.so")
PORT_SO = os.path.join(HERE, "_build", "libpn_port_querier.so")
_libs = {}


def build_port(force=False):
    src = os.path.join(HERE, "querier_port.c")
    if force or not os.path.isfile(PORT_SO) or os.path.getmtime(PORT_SO) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(PORT_SO), exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-std=c11", "-shared", "-fPIC", "-mfma", "-ffp-contract=off",
                               src, "-o", PORT_SO, "-lm"])
    return PORT_SO


def have_backend(name):
    return os.path.isfile(REF_SO if name == "ref" else PORT_SO) or name == "port"


def _lib(name):
    if name not in _libs:
        path = REF_SO if name == "ref" else build_port()
        _libs[name] = ctypes.CDLL(path)
    return _libs[name]


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


I, F_, L = ctypes.c_int, ctypes.c_float, ctypes.c_int64
THREADS = 1024  # opt.gpu_maxthr default (neural_points.py:55-58); only shapes the serial loop


class _Kernels:
    """Uniform calling convention over the two libraries."""

    def __init__(self, backend):
        self.backend = backend
        self.lib = _lib(backend)
        self.overflow = np.zeros(1, np.int32)
        self.candidates = np.zeros(1, np.int64)

    @staticmethod
    def _grid(n):
        return int((n + THREADS - 1) / THREADS)

    def claim_occ(self, xyz, n_act, N, shift, vsz, gsz, vol, max_o, occ_idx, c2o, o2c):
        if self.backend == "ref":
            self.lib.pn_ref_claim_occ(I(self._grid(N)), I(THREADS), _p(xyz), _p(n_act), I(1), I(N), _p(shift), _p(vsz),
                                      _p(gsz), I(vol), I(max_o), _p(occ_idx), _p(c2o), _p(o2c))
        else:
            self.lib.pn_port_claim_occ(_p(xyz), I(int(n_act[0])), _p(shift), _p(vsz), _p(gsz), I(max_o), _p(occ_idx),
                                       _p(c2o), _p(o2c), _p(self.overflow))

    def map_coor2occ(self, gsz, dil, vol, max_o, occ_idx, occ, c2o, o2c):
        if self.backend == "ref":
            self.lib.pn_ref_map_coor2occ(I(self._grid(max_o)), I(THREADS), I(1), _p(gsz), _p(dil), I(vol), I(max_o),
                                         _p(occ_idx), _p(occ), _p(c2o), _p(o2c))
        else:
            self.lib.pn_port_map_coor2occ(_p(gsz), _p(dil), I(max_o), _p(occ_idx), _p(occ), _p(c2o), _p(o2c))

    def fill_occ2pnts(self, xyz, n_act, N, P, shift, vsz, gsz, vol, max_o, c2o, o2p, onum):
        if self.backend == "ref":
            self.lib.pn_ref_fill_occ2pnts(I(self._grid(N)), I(THREADS), _p(xyz), _p(n_act), I(1), I(N), I(P), _p(shift),
                                          _p(vsz), _p(gsz), I(vol), I(max_o), _p(c2o), _p(o2p), _p(onum))
        else:
            self.lib.pn_port_fill_occ2pnts(_p(xyz), I(int(n_act[0])), I(P), _p(shift), _p(vsz), _p(gsz), _p(c2o),
                                           _p(o2p), _p(onum), _p(self.overflow))

    def mask_raypos(self, raypos, occ, R, D, vol, shift, gsz, vsz, mask):
        if self.backend == "ref":
            # the reference sizes this grid as ceil(R*D/threads); threads past R*D are cut by the
            # i_batch >= B test (:403-404)
            self.lib.pn_ref_mask_raypos(I(self._grid(R * D)), I(THREADS), _p(raypos), _p(occ), I(1), I(R), I(D), I(vol),
                                        _p(shift), _p(gsz), _p(vsz), _p(mask))
        else:
            self.lib.pn_port_mask_raypos(_p(raypos), _p(occ), L(R * D), _p(shift), _p(gsz), _p(vsz), _p(mask))

    def get_shadingloc(self, raypos, slot, R, D, SR, loc, loc_mask, grid_threads):
        if self.backend == "ref":
            # launched with the grid of the previous (uncompacted) launch (:634,676): extra threads exit
            self.lib.pn_ref_get_shadingloc(I(self._grid(grid_threads)), I(THREADS), _p(raypos), _p(slot), I(1), I(R),
                                           I(D), I(SR), _p(loc), _p(loc_mask))
        else:
            self.lib.pn_port_get_shadingloc(_p(raypos), _p(slot), I(R), I(D), I(SR), _p(loc), _p(loc_mask))

    def query_along_ray(self, xyz, SR, R, max_o, P, K, vol, r2, shift, gsz, vsz, ksz, onum, o2p, c2o, loc, loc_mask,
                        pidx, NN):
        if self.backend == "ref":
            self.lib.pn_ref_query_along_ray(I(self._grid(R * SR)), I(THREADS), _p(xyz), I(1), I(SR), I(R), I(max_o), I(P),
                                            I(K), I(vol), F_(r2), _p(shift), _p(gsz), _p(vsz), _p(ksz), _p(onum),
                                            _p(o2p), _p(c2o), _p(loc), _p(loc_mask), _p(pidx), I(NN))
        else:
            self.lib.pn_port_query_along_ray(_p(xyz), L(R * SR), I(P), I(K), F_(r2), _p(shift), _p(gsz), _p(vsz), _p(ksz),
                                             _p(onum), _p(o2p), _p(c2o), _p(loc), _p(loc_mask), _p(pidx),
                                             _p(self.candidates))


def get_hyperparameters(opt, xyz_w):
    """worldcoords.py:48-81, including its float32/float64 mixing.  xyz_w: torch f32 [1,N,3]."""
    vsize_np = opt.vsize
    min_xyz, max_xyz = torch.min(xyz_w, dim=-2)[0][0], torch.max(xyz_w, dim=-2)[0][0]
    vscale_np = np.array(opt.vscale, dtype=np.int32)
    scaled_vsize_np = (vsize_np * vscale_np).astype(np.float32)
    ranges = opt.ranges
    if ranges is not None and ranges[0] >= ranges[3]:
        ranges = None
    if ranges is not None:
        min_xyz = torch.max(torch.stack([min_xyz, torch.as_tensor(ranges[:3], dtype=torch.float32)], dim=0), dim=0)[0]
        max_xyz = torch.min(torch.stack([max_xyz, torch.as_tensor(ranges[3:], dtype=torch.float32)], dim=0), dim=0)[0]
    pad = torch.as_tensor(scaled_vsize_np * opt.kernel_size / 2, dtype=torch.float32)
    min_xyz = min_xyz - pad
    max_xyz = max_xyz + pad
    ranges_np = torch.cat([min_xyz, max_xyz], dim=-1).numpy().astype(np.float32)
    vdim_np = (max_xyz - min_xyz).numpy() / vsize_np
    scaled_vdim_np = np.ceil(vdim_np / vscale_np).astype(np.int32)
    radius_limit = np.asarray(opt.radius_limit_scale * max(vsize_np[0], vsize_np[1])).astype(np.float32)
    return dict(ranges=ranges_np, scaled_vsize=scaled_vsize_np, scaled_vdim=scaled_vdim_np,
                radius_limit=radius_limit, radius_limit2=np.float32(radius_limit ** 2),
                kernel_size=np.asarray(opt.kernel_size, dtype=np.int32),
                query_size=np.asarray(opt.query_size, dtype=np.int32))


def build_occ_vox(kern, xyz_np, N, P, max_o, hp):
    """worldcoords.py:546-611 (fills, claim_occ, refill, map_coor2occ with query_size, fill_occ2pnts)."""
    gsz = hp["scaled_vdim"]
    vol = int(gsz[0]) * int(gsz[1]) * int(gsz[2])
    shift = np.ascontiguousarray(hp["ranges"][:3])
    n_act = np.array([N], np.int32)
    occ = np.zeros(vol, np.int32)
    o2p = np.full(max_o * P, -1, np.int32)
    o2c = np.full(max_o * 3, -1, np.int32)
    onum = np.zeros(max_o, np.int32)
    c2o = np.full(vol, -1, np.int32)
    occ_idx = np.zeros(1, np.int32)
    kern.claim_occ(xyz_np, n_act, N, shift, hp["scaled_vsize"], gsz, vol, max_o, occ_idx, c2o, o2c)
    c2o = np.full(vol, -1, np.int32)                                        # :575
    kern.map_coor2occ(gsz, hp["query_size"], vol, max_o, occ_idx, occ, c2o, o2c)  # dilation = query_size (:626)
    kern.fill_occ2pnts(xyz_np, n_act, N, P, shift, hp["scaled_vsize"], gsz, vol, max_o, c2o, o2p, onum)
    return dict(coor_occ=occ, occ_2_coor=o2c, coor_2_occ=c2o, occ_idx=occ_idx, occ_numpnts=onum, occ_2_pnts=o2p,
                shift=shift, vol=vol)


def query_grid_point_index(kern, opt, raypos, xyz_np, hp, grid=None):
    """worldcoords.py:614-721.  raypos: np f32 [R,D,3].  Returns compacted arrays + masks + grid."""
    N = xyz_np.shape[0]
    R, D = raypos.shape[0], raypos.shape[1]
    SR, K, P, max_o = opt.SR, opt.K, opt.P, opt.max_o
    g = grid if grid is not None else build_occ_vox(kern, xyz_np, N, P, max_o, hp)
    gsz, vsz, shift, vol = hp["scaled_vdim"], hp["scaled_vsize"], g["shift"], g["vol"]
    mask = np.zeros(R * D, np.int32)
    kern.mask_raypos(np.ascontiguousarray(raypos), g["coor_occ"], R, D, vol, shift, gsz, vsz, mask)
    mask = mask.reshape(R, D)
    ray_mask = mask.max(axis=-1) > 0 if D > 0 else np.zeros(R, bool)         # :655
    R1 = int(ray_mask.sum())
    loc = np.zeros((R1, SR, 3), np.float32)
    pidx = np.full((R1, SR, K), -1, np.int32)
    hit_slots = np.zeros((R1, SR), np.int32)
    if R1 > 0:
        rp = np.ascontiguousarray(raypos[ray_mask])                           # :660
        m = mask[ray_mask]                                                    # :661
        cum = np.cumsum(m, axis=-1).astype(np.int32)                          # :664
        slot = np.ascontiguousarray((m * cum * (cum <= SR)) - 1).astype(np.int32)  # :665
        kern.get_shadingloc(rp, slot, R1, D, SR, loc, hit_slots, R * D)
        kern.query_along_ray(xyz_np, SR, R1, max_o, P, K, vol, float(hp["radius_limit2"]), shift, gsz, vsz,
                             hp["kernel_size"], g["occ_numpnts"], g["occ_2_pnts"], g["coor_2_occ"], loc, hit_slots,
                             pidx, opt.NN)
        valid_ray = (pidx.reshape(R1, -1) >= 0).sum(axis=-1) > 0              # :715
        full = ray_mask.copy()
        full[ray_mask] = valid_ray                                            # :717 masked_scatter_
        ray_mask = full
        pidx_unc, loc_unc = pidx, loc
        pidx = np.ascontiguousarray(pidx[valid_ray])                          # :718
        loc = np.ascontiguousarray(loc[valid_ray])                            # :719
    return dict(sample_pidx=pidx, sample_loc_w=loc, ray_mask=ray_mask.astype(np.int8), grid=g,
                n_hit_rays=R1)


def query_points(opt, xyz_w, raydir, campos, camrotc2w, near, far, backend="port", uniform=None, ts=None,
                 grid=None, hp=None):
    """worldcoords.py:84-99.  Tensors are torch-CPU f32: xyz_w [1,N,3], raydir [1,R,3], campos [1,3],
    camrotc2w [1,3,3].  ``uniform`` replaces the torch.rand draw of the jittered sampler;
    ``ts`` (middle_point_ts [1,R,D]) overrides the sampler entirely.

    Returns the reference's 7-tuple (as torch tensors) plus a dict of extras."""
    kern = _Kernels(backend)
    hp = hp if hp is not None else get_hyperparameters(opt, xyz_w)
    if ts is None:
        ts = torch_port.sample_ts(raydir, opt.z_depth_dim, float(near), float(far),
                                  jitter=0.3 if opt.is_train > 0 else 0.0, uniform=uniform)
    raypos = torch_port.raypos_from_ts(campos, raydir, ts)
    xyz_np = np.ascontiguousarray(xyz_w[0].numpy())
    q = query_grid_point_index(kern, opt, raypos[0].numpy(), xyz_np, hp, grid=grid)
    pidx = torch.from_numpy(q["sample_pidx"])[None]
    loc_w = torch.from_numpy(q["sample_loc_w"])[None]
    ray_mask = torch.from_numpy(q["ray_mask"])[None]
    dirs = torch.masked_select(raydir, ray_mask[..., None] > 0).reshape(1, -1, 3)[..., None, :] \
        .expand(-1, -1, opt.SR, -1).contiguous()                              # :97
    loc_pers = torch_port.w2pers_samples(loc_w, camrotc2w, campos)            # :99
    extras = dict(hp=hp, grid=q["grid"], ts=ts, n_hit_rays=q["n_hit_rays"], overflow=int(kern.overflow[0]),
                  candidates=int(kern.candidates[0]))
    return (pidx, loc_pers, loc_w, dirs, ray_mask, opt.vsize, hp["ranges"]), extras
