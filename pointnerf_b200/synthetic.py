"""Seeded synthetic scenes, cameras and rays of the shapes BASELINE.json names.

There are no datasets or checkpoints on the GPU box, so every workload is generated: points on
analytic surfaces inside ``opt.ranges``, thinned deterministically BY POINT INDEX so that no
scaled voxel holds more than ``P`` points and no more than ``max_o`` voxels are occupied (that keeps
the reference's time-seeded reservoir branches unreachable, worldcoords.py:290-298,378-383), per
point attributes as the reference initialises them, and pinhole cameras looking at the scene.

Everything here is host-side numpy/torch-CPU; it is input generation, not the product path.
"""
import math
from argparse import Namespace

import numpy as np
import torch

from .options import make_opt

# BASELINE.json configs -> generator parameters
CONFIGS = {
    # tiny: unit tests, golden fixtures
    "tiny": dict(n_points=6000, opt=dict(SR=24, K=8, P=9, max_o=20000, z_depth_dim=96,
                                         ranges=(-0.12, -0.12, -0.12, 0.12, 0.12, 0.12)),
                 near=0.3, far=0.9, cam_dist=0.6, h=48, w=48, focal=110.0, surfaces="small"),
    # miniatures of cfg3 / cfg4 for the oracle-parity tests: their voxel sizes, bucket capacities, SR, near / far and
    # camera-inside-the-scene geometry on a few thousand points
    "tiny3": dict(n_points=40000, opt=dict(SR=24, K=8, P=30, max_o=30000, z_depth_dim=128, vsize=(0.008, 0.008, 0.008),
                                          ranges=(-0.4, -0.4, -0.15, 0.4, 0.4, 0.15), radius_limit_scale=4.0),
                  near=0.1, far=0.9, cam_dist=0.0, h=48, w=64, focal=57.0, surfaces="room"),
    "tiny4": dict(n_points=12000, opt=dict(SR=24, K=8, P=26, max_o=30000, z_depth_dim=160, vsize=(0.08, 0.08, 0.08),
                                           ranges=(-10.0, -5.0, -1.0, 10.0, 5.0, 1.0)),
                  near=0.1, far=12.0, cam_dist=0.0, h=64, w=96, focal=100.0, surfaces="street"),
    "cfg1": dict(n_points=200_000, opt=dict(SR=40, K=8, P=9, max_o=400_000), near=2.0, far=6.0, cam_dist=4.0,
                 h=400, w=400, focal=555.0, surfaces="lego"),
    "cfg2": dict(n_points=1_000_000, opt=dict(SR=80, K=8, P=9, max_o=830_000), near=2.0, far=6.0, cam_dist=4.0,
                 h=800, w=800, focal=1111.0, surfaces="lego"),
    "cfg3": dict(n_points=3_000_000,
                 opt=dict(SR=24, K=8, P=30, max_o=2_000_000, vsize=(0.008, 0.008, 0.008),
                          ranges=(-4.0, -4.0, -1.5, 4.0, 4.0, 1.5), radius_limit_scale=4.0),
                 near=0.1, far=8.0, cam_dist=0.0, h=480, w=640, focal=577.0, surfaces="room"),
    "cfg4": dict(n_points=20_000_000,
                 opt=dict(SR=24, K=8, P=26, max_o=6_000_000, vsize=(0.08, 0.08, 0.08),
                          ranges=(-100.0, -50.0, -10.0, 100.0, 50.0, 10.0), z_depth_dim=400),
                 near=0.1, far=100.0, cam_dist=0.0, h=1280, w=1920, focal=2000.0, surfaces="street"),
    "cfg5": dict(n_points=1_000_000, opt=dict(SR=40, K=8, P=9, max_o=830_000), near=2.0, far=6.0, cam_dist=4.0,
                 h=800, w=800, focal=1111.0, surfaces="lego"),
}


def config_opt(name, **overrides) -> Namespace:
    kw = dict(CONFIGS[name]["opt"])
    kw.update(overrides)
    return make_opt(**kw)


from .query_host import grid_hyperparameters, voxel_coords  # noqa: E402,F401  (product host arithmetic; the scene generator thins clouds with it)


def _surface_points(kind, n, rng, ranges):
    lo, hi = np.asarray(ranges[:3]), np.asarray(ranges[3:])
    ctr, ext = (lo + hi) / 2, (hi - lo) / 2
    parts = []

    def shell(c, radii, m):
        v = rng.normal(size=(m, 3))
        v /= np.linalg.norm(v, axis=1, keepdims=True)
        return c + v * radii

    def plane(z, m, over=1.1):
        p = np.empty((m, 3))
        p[:, 0] = ctr[0] + ext[0] * over * rng.uniform(-1, 1, m)
        p[:, 1] = ctr[1] + ext[1] * over * rng.uniform(-1, 1, m)
        p[:, 2] = z
        return p

    if kind in ("lego", "small"):
        parts.append(shell(ctr, ext * 0.80, int(n * 0.40)))
        parts.append(shell(ctr + ext * [0.1, -0.1, 0.05], ext * 0.55, int(n * 0.22)))
        parts.append(shell(ctr - ext * [0.2, 0.1, 0.1], ext * 0.30, int(n * 0.10)))
        parts.append(plane(lo[2] + 0.15 * ext[2], n - sum(len(p) for p in parts)))
    elif kind == "room":
        m = n // 6
        parts.append(plane(lo[2] + 0.02, m, 1.0))
        parts.append(plane(hi[2] - 0.02, m, 1.0))
        for ax, side in ((0, lo), (0, hi), (1, lo)):
            p = plane(0.0, m, 1.0)
            p[:, 2] = ctr[2] + ext[2] * rng.uniform(-1, 1, m)
            p[:, ax] = side[ax] + (0.02 if side is lo else -0.02)
            parts.append(p)
        parts.append(shell(ctr + [0.5, 0.3, -0.8], np.array([0.8, 0.6, 0.5]), n - sum(len(p) for p in parts)))
    elif kind == "street":
        m = n // 2
        parts.append(plane(lo[2] + 1.0, m, 1.0))
        for side in (-1, 1):
            k = (n - m) // 2 if side < 0 else n - m - (n - m) // 2
            p = plane(0.0, k, 1.0)
            p[:, 2] = lo[2] + 1.0 + (ext[2] * 1.6) * rng.uniform(0, 1, k)
            p[:, 1] = ctr[1] + side * ext[1] * 0.6
            parts.append(p)
    else:
        raise KeyError(kind)
    return np.concatenate(parts, axis=0)[:n]


def thin_by_index(xyz, opt):
    """Keeps, per scaled voxel, the first P points by index and at most max_o voxels (by first point)."""
    xyz = np.asarray(xyz, np.float32)
    tx = torch.from_numpy(xyz)
    hp = grid_hyperparameters(opt, tx.min(0)[0].numpy(), tx.max(0)[0].numpy())
    c = voxel_coords(xyz, hp["ranges"][:3], hp["scaled_vsize"])
    dims = hp["scaled_vdim"].astype(np.int64)
    inside = np.all((c >= 0) & (c < dims[None, :]), axis=1)
    key = (c[:, 0] * dims[1] + c[:, 1]) * dims[2] + c[:, 2]
    key = np.where(inside, key, -1)
    # a stable sort has one answer: torch's multi-threaded sort returns numpy's permutation 6x sooner at 20 M points
    order = torch.sort(torch.from_numpy(np.ascontiguousarray(key)), stable=True).indices.numpy()
    ks = key[order]
    first = np.r_[True, ks[1:] != ks[:-1]]
    start = np.maximum.accumulate(np.where(first, np.arange(len(ks)), 0))
    rank = np.arange(len(ks)) - start
    keep_sorted = (rank < opt.P) | (ks < 0)
    # voxel budget: occupied voxels in order of first point index
    vox_first = order[first & (ks >= 0)]
    if len(vox_first) > opt.max_o:
        allowed = set(ks[first & (ks >= 0)][np.argsort(vox_first, kind="stable")[:opt.max_o]].tolist())
        keep_sorted &= np.array([(k < 0) or (k in allowed) for k in ks])
    keep = np.zeros(len(xyz), bool)
    keep[order] = keep_sorted
    return keep


def make_scene(name, seed=None, n_points=None, opt=None):
    """Returns (opt, dict of torch-CPU point tables in the reference's layouts).

    xyz [N,3]; points_embeding [1,N,F] ~ U(-.5,.5) (neural_points.py:292); points_conf [1,N,1];
    points_dir [1,N,3] unit; points_color [1,N,3]; Rw2c [3,3] = I."""
    cfg = CONFIGS[name]
    opt = opt if opt is not None else config_opt(name)
    n = int(n_points if n_points is not None else cfg["n_points"])
    seed = 1234 + sum(map(ord, name)) if seed is None else seed
    rng = np.random.default_rng(seed)
    sv = float(opt.vsize[0]) * float(opt.vscale[0])
    xyz = np.empty((0, 3), np.float32)
    want = n
    for _ in range(6):  # top up until n points survive the thinning
        cand = _surface_points(cfg["surfaces"], int(want * 1.15) + 64, rng, opt.ranges)
        cand = (cand + rng.normal(scale=0.25 * sv, size=cand.shape)).astype(np.float32)
        xyz = np.concatenate([xyz, cand], axis=0)
        xyz = xyz[thin_by_index(xyz, opt)]
        if len(xyz) >= n:
            break
        want = n - len(xyz)
    xyz = xyz[:n]
    while True:  # the bbox (hence the voxel origin) can move when points are dropped: iterate to a fixed point
        keep = thin_by_index(xyz, opt)
        if keep.all():
            break
        xyz = xyz[keep]
    N = len(xyz)
    g = torch.Generator().manual_seed(seed)
    F = opt.point_features_dim
    d = torch.randn(1, N, 3, generator=g)
    tables = dict(
        xyz=torch.from_numpy(np.ascontiguousarray(xyz)),
        points_embeding=torch.rand(1, N, F, generator=g) - 0.5,
        points_conf=torch.rand(1, N, 1, generator=g) * 0.9 + 0.1,
        points_dir=d / d.norm(dim=-1, keepdim=True),
        points_color=torch.randn(1, N, 3, generator=g),
        Rw2c=torch.eye(3),
    )
    return opt, tables


def make_camera(name, seed=0, view=0):
    """Pinhole camera on a sphere around the scene centre (or inside the room/street), c2w rotation."""
    cfg = CONFIGS[name]
    opt_ranges = CONFIGS[name]["opt"].get("ranges", make_opt().ranges)
    lo, hi = np.asarray(opt_ranges[:3]), np.asarray(opt_ranges[3:])
    ctr = (lo + hi) / 2
    rng = np.random.default_rng(1000 * seed + view + 7)
    if cfg["cam_dist"] > 0:
        v = rng.normal(size=3)
        v[2] = abs(v[2]) * 0.7 + 0.2
        v /= np.linalg.norm(v)
        pos = ctr + v * cfg["cam_dist"]
        target = ctr
    else:
        pos = ctr + (hi - lo) * rng.uniform(-0.15, 0.15, 3)
        ang = rng.uniform(0, 2 * math.pi)
        target = pos + np.array([math.cos(ang), math.sin(ang), -0.1])
    fwd = target - pos
    fwd /= np.linalg.norm(fwd)
    up = np.array([0.0, 0.0, 1.0])
    right = np.cross(fwd, up)
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    rot = np.stack([right, down, fwd], axis=1)  # columns = camera axes in world (c2w)
    h, w, f = cfg["h"], cfg["w"], cfg["focal"]
    intrinsic = np.array([[f, 0, w / 2], [0, f, h / 2], [0, 0, 1]], np.float32)
    return dict(campos=torch.tensor(pos, dtype=torch.float32)[None], camrotc2w=torch.tensor(rot, dtype=torch.float32)[None],
                intrinsic=intrinsic, h=h, w=w, near=float(cfg["near"]), far=float(cfg["far"]))


def make_rays(cam, pixel_idx):
    """Ray directions with the semantics of data/data_utils.py:161-175 (dir_norm=0: not normalised).

    pixel_idx: int array [R,2] of (x, y).  Returns raydir f32 [1,R,3], pixel_idx int32 [1,R,2]."""
    K = cam["intrinsic"]
    px = torch.as_tensor(np.asarray(pixel_idx), dtype=torch.float32)
    d = torch.stack([(px[:, 0] + 0.5 - K[0, 2]) / K[0, 0], (px[:, 1] + 0.5 - K[1, 2]) / K[1, 1],
                     torch.ones(px.shape[0])], dim=-1)
    raydir = d @ cam["camrotc2w"][0].T
    return raydir[None].contiguous(), torch.as_tensor(np.asarray(pixel_idx), dtype=torch.int32)[None]


def random_pixels(cam, n, seed=0):
    rng = np.random.default_rng(seed + 31)
    return np.stack([rng.integers(0, cam["w"], n), rng.integers(0, cam["h"], n)], axis=1)


def all_pixels(cam):
    ys, xs = np.meshgrid(np.arange(cam["h"]), np.arange(cam["w"]), indexing="ij")
    return np.stack([xs.reshape(-1), ys.reshape(-1)], axis=1)


# ---- perspective-frustum querier cases: the "tiny" cloud seen through two cameras ------------------
FRUSTUM_CASES = {
    # the 63-channel DTU feed-forward settings (dev_scripts/ete/dtu_dgt_d012_img0123_conf_color_dir_agg2.sh:39-46) in small
    "dtu": dict(opt=dict(vscale=(2, 2, 1), kernel_size=(3, 3, 3), query_size=(3, 3, 3), SR=24, K=8, P=20, NN=2,
                         radius_limit_scale=0.0, depth_limit_scale=0.0, z_depth_dim=96), crop=None, n_pix=500),
    # a cropped view: the left / top / near faces of the frustum cut through the cloud (points up to one voxel below a low
    # face are filed by truncation), odd vscale, radius + depth limits, perspective-space distances, SR < occupied depths
    "crop": dict(opt=dict(vscale=(3, 2, 2), kernel_size=(5, 5, 3), query_size=(5, 5, 3), SR=6, K=4, P=48, NN=1,
                          radius_limit_scale=3.0, depth_limit_scale=2.0, z_depth_dim=64), crop=(17, 11, 40, 30, 0.6, 0.8),
                 n_pix=None),
}


def frustum_case(name):
    """-> opt, xyz_w [N,3], cam dict, pixel_idx int32 [1,R,2] for a perspective-querier parity case."""
    c = FRUSTUM_CASES[name]
    opt = config_opt("tiny", wcoord_query=0, max_o=None, **c["opt"])
    _, tables = make_scene("tiny")
    cam = make_camera("tiny")
    if c["crop"] is not None:
        x0, y0, w, h, near, far = c["crop"]
        K = cam["intrinsic"].copy()
        K[0, 2] -= x0
        K[1, 2] -= y0
        cam = dict(cam, intrinsic=K, w=w, h=h, near=near, far=far)
    pix = all_pixels(cam) if c["n_pix"] is None else random_pixels(cam, c["n_pix"])
    return opt, tables["xyz"], cam, torch.as_tensor(pix, dtype=torch.int32)[None]
