"""Voxel grid, ray sampling and K-NN on the GPU vs the querier oracle (bit-exact: integer / index work)."""
import numpy as np
import pytest
import torch

from oracle import querier_ref as qr
from pointnerf_b200 import _lib, query, synthetic as syn

pytestmark = pytest.mark.gpu


def _oracle(opt, tables, cam, raydir, backend="port", uniform=None):
    return qr.query_points(opt, tables["xyz"][None], raydir, cam["campos"], cam["camrotc2w"], cam["near"], cam["far"],
                           backend=backend, uniform=uniform)


def _buckets_from_oracle(grid, P):
    """cell -> list of point indices from the reference tables (occ_2_pnts rows via coor_2_occ)."""
    c2o, onum, o2p = grid["coor_2_occ"], grid["occ_numpnts"], grid["occ_2_pnts"].reshape(-1, P)
    cells = np.nonzero(c2o >= 0)[0]
    return {int(c): o2p[c2o[c], :min(onum[c2o[c]], P)].tolist() for c in cells}


def test_hyperparameters_match_oracle(tiny_scene, cuda_dev):
    opt, tables, cam = tiny_scene
    q = query.lighting_fast_querier(cuda_dev, opt)
    hp = q.get_hyperparameters(opt.vsize, tables["xyz"][None].to(cuda_dev))
    want = qr.get_hyperparameters(opt, tables["xyz"][None])
    for k in ("ranges", "scaled_vsize", "scaled_vdim", "radius_limit2"):
        assert np.array_equal(np.asarray(hp[k]), np.asarray(want[k])), k


def test_grid_matches_oracle(tiny_scene, cuda_dev):
    opt, tables, cam = tiny_scene
    q = query.lighting_fast_querier(cuda_dev, opt)
    xyz = tables["xyz"].to(cuda_dev)
    g = q.grid_for(xyz)
    raydir, _ = syn.make_rays(cam, syn.random_pixels(cam, 8))
    _, ex = _oracle(opt, tables, cam, raydir)
    og = ex["grid"]
    info = g.info_host()
    assert info[_lib.GI_N_OCC] == int(og["occ_idx"][0])
    assert info[_lib.GI_OVER_P] == 0 and info[_lib.GI_OVER_MAX_O] == 0
    # dilated occupancy
    bits = g.occ_bits.cpu().numpy().view(np.uint32)
    occ = ((bits[:, None] >> np.arange(32, dtype=np.uint32)[None, :]) & 1).reshape(-1)[:g.V]
    assert np.array_equal(occ.astype(np.int32), og["coor_occ"])
    # buckets: same members in the same (ascending index) order, except the slot-0 voxel which the reference leaves empty
    start = g.cell_start.cpu().numpy()
    bucket = g.bucket_pidx.cpu().numpy()
    want = _buckets_from_oracle(og, opt.P)
    slot0 = int(info[_lib.GI_SLOT0_CELL])
    assert og["coor_2_occ"][slot0] == 0
    cells = np.nonzero(start[1:] - start[:-1])[0]
    assert set(cells.tolist()) == set(want.keys())
    for c in cells:
        mine = bucket[start[c]:start[c + 1]].tolist()
        if c == slot0:
            assert want[c] == []  # `if (voxel_idx > 0)` quirk: slot 0 never receives points
        else:
            assert mine == want[c], c


@pytest.mark.parametrize("train", [0, 1])
def test_query_points_matches_oracle(tiny_scene, cuda_dev, train):
    opt, tables, cam = tiny_scene
    import copy
    opt = copy.copy(opt)
    opt.is_train = train
    raydir, pix = syn.make_rays(cam, syn.random_pixels(cam, 700, seed=1))
    uniform = torch.rand(1, raydir.shape[1], opt.z_depth_dim, generator=torch.Generator().manual_seed(9)) if train else None
    (pidx, loc_pers, loc_w, dirs, ray_mask, vsize, ranges), ex = _oracle(opt, tables, cam, raydir, uniform=uniform)
    q = query.lighting_fast_querier(cuda_dev, opt)
    d = cuda_dev
    # the candidate depths come from torch ops on the device; they must be bit-identical to the CPU chain for the
    # sample positions (and therefore every index) to agree
    ts_dev = query.middle_point_ts(raydir.to(d), opt.z_depth_dim, cam["near"], cam["far"], jitter=0.3 if train else 0.0,
                                   uniform=uniform.to(d) if train else None)
    ts_ref = ex["ts"]
    assert torch.equal(ts_dev.cpu().expand_as(ts_ref), ts_ref), "pn_sample_depths differs from the host chain (double-accumulated cumsum)"
    out = q.query_points(pix.to(d), None, tables["xyz"][None].to(d), None, cam["h"], cam["w"], cam["intrinsic"],
                         np.float32(cam["near"]), np.float32(cam["far"]), raydir.to(d), cam["campos"].to(d),
                         cam["camrotc2w"].to(d), uniform=uniform.to(d) if train else None)
    assert torch.equal(out[4].cpu(), ray_mask)
    assert torch.equal(out[2].cpu(), loc_w)
    assert torch.equal(out[0].cpu(), pidx), "neighbour slots differ from the oracle"
    assert out[0].dtype == torch.int32 and out[4].dtype == torch.int8
    assert torch.equal(out[3].cpu(), dirs)
    torch.testing.assert_close(out[1].cpu(), loc_pers, rtol=1e-6, atol=1e-7)
    assert np.array_equal(out[6], ranges)


def test_query_with_given_ts(tiny_scene, cuda_dev):
    """Same depths handed to both sides: positions, masks and neighbour slots must be bit-identical."""
    opt, tables, cam = tiny_scene
    raydir, pix = syn.make_rays(cam, syn.all_pixels(cam)[::3])
    (pidx, loc_pers, loc_w, dirs, ray_mask, vsize, ranges), ex = _oracle(opt, tables, cam, raydir)
    q = query.lighting_fast_querier(cuda_dev, opt)
    d = cuda_dev
    ts = ex["ts"][:, :1].contiguous().to(d)  # no jitter: one row
    raw = q.query_raw(tables["xyz"].to(d), raydir.to(d), cam["campos"].to(d), cam["near"], cam["far"], ts=ts, count_candidates=True)
    sel = raw["ray_any"].cpu() > 0
    assert torch.equal(sel.to(torch.int8)[None], ray_mask)
    assert torch.equal(raw["sample_loc_w"].cpu()[sel][None], loc_w)
    assert torch.equal(raw["sample_pidx"].cpu()[sel][None], pidx)
    # rays that hit occupancy but found no neighbour: all -1
    assert (raw["sample_pidx"].cpu()[~sel] == -1).all()
    assert int(raw["candidates"].item()) > 0


def test_query_with_given_jittered_ts(tiny_scene, cuda_dev):
    """The oracle's PER-RAY jittered depths [1,R,D] handed to the device query: positions, mask and neighbour slots
    bit-identical (independent of how either side produces its depths)."""
    import copy
    opt, tables, cam = tiny_scene
    opt = copy.copy(opt)
    opt.is_train = 1
    raydir, pix = syn.make_rays(cam, syn.random_pixels(cam, 900, seed=4))
    uniform = torch.rand(1, raydir.shape[1], opt.z_depth_dim, generator=torch.Generator().manual_seed(21))
    (pidx, loc_pers, loc_w, dirs, ray_mask, vsize, ranges), ex = _oracle(opt, tables, cam, raydir, uniform=uniform)
    assert tuple(ex["ts"].shape) == (1, raydir.shape[1], opt.z_depth_dim)
    d = cuda_dev
    q = query.lighting_fast_querier(d, opt)
    raw = q.query_raw(tables["xyz"].to(d), raydir.to(d), cam["campos"].to(d), cam["near"], cam["far"], ts=ex["ts"].contiguous().to(d))
    sel = raw["ray_any"].cpu() > 0
    assert torch.equal(sel.to(torch.int8)[None], ray_mask)
    assert torch.equal(raw["sample_loc_w"].cpu()[sel][None], loc_w)
    assert torch.equal(raw["sample_pidx"].cpu()[sel][None], pidx)
    assert (raw["sample_pidx"].cpu()[~sel] == -1).all()


@pytest.mark.parametrize("D,near,far,jitter", [(400, 2.0, 6.0, 0.3), (400, 0.1, 8.0, 0.3), (300, 0.0, 100.0, 0.3), (37, 2.0, 6.0, 0.0),
                                               (400, 2.0, 6.0, 0.0), (400, 0.0, 100.0, 0.0)])
def test_sample_depths_equal_host_chain(cuda_dev, D, near, far, jitter):
    """pn_sample_depths vs the reference's own sampler run on the host (oracle.ref_import is not available on the GPU
    box, so the chain of diff_ray_marching.py:369-386 is spelled out with the same torch ops): bit-identical for the
    near / far / D of every BASELINE config, with and without jitter."""
    R = 513
    u = torch.rand(1, R, D, generator=torch.Generator().manual_seed(D)) if jitter else None
    t = torch.linspace(0, 1, D + 1).view(1, -1)
    t = near * (1 - t) + far * t
    seg = (t[..., 1:] - t[..., :-1]) * (1 + jitter * ((u if jitter else torch.full((1, 1, D), 0.5)) - 0.5))
    ends = torch.cumsum(seg, dim=2)
    ends = near + torch.cat([torch.zeros((1, ends.shape[1], 1)), ends], dim=2)
    want = (ends[:, :, :-1] + ends[:, :, 1:]) / 2
    raydir = torch.zeros(1, R, 3, device=cuda_dev)
    got = query.middle_point_ts(raydir, D, near, far, jitter=jitter, uniform=u.to(cuda_dev) if jitter else None)
    assert torch.equal(got.cpu(), want)


def test_reference_kernels_agree_with_port(tiny_scene):
    """oracle/_ref (the reference's own kernel string, host-compiled) vs the C port, when the built file travelled."""
    opt, tables, cam = tiny_scene
    if not qr.have_backend("ref") or not __import__("os").path.isfile(qr.REF_SO):
        pytest.skip("oracle/_ref not built")
    raydir, _ = syn.make_rays(cam, syn.random_pixels(cam, 300, seed=2))
    a, _ = _oracle(opt, tables, cam, raydir, backend="ref")
    b, _ = _oracle(opt, tables, cam, raydir, backend="port")
    for x, y in zip(a[:5], b[:5]):
        assert torch.equal(x, y)


def test_empty_and_missing_rays(tiny_scene, cuda_dev):
    opt, tables, cam = tiny_scene
    d = cuda_dev
    q = query.lighting_fast_querier(d, opt)
    # rays pointing away from the scene: nothing hit, R'' = 0, shapes as the reference (:657-659)
    raydir = -syn.make_rays(cam, syn.random_pixels(cam, 64))[0]
    out = q.query_points(None, None, tables["xyz"][None].to(d), None, cam["h"], cam["w"], cam["intrinsic"], cam["near"],
                         cam["far"], raydir.to(d), cam["campos"].to(d), cam["camrotc2w"].to(d))
    assert tuple(out[0].shape) == (1, 0, opt.SR, opt.K) and tuple(out[2].shape) == (1, 0, opt.SR, 3)
    assert int(out[4].sum()) == 0 and tuple(out[4].shape) == (1, 64)
    # zero rays
    out = q.query_points(None, None, tables["xyz"][None].to(d), None, cam["h"], cam["w"], cam["intrinsic"], cam["near"],
                         cam["far"], raydir[:, :0].to(d), cam["campos"].to(d), cam["camrotc2w"].to(d))
    assert tuple(out[0].shape) == (1, 0, opt.SR, opt.K)


def test_grid_cache_invalidation(tiny_scene, cuda_dev):
    opt, tables, cam = tiny_scene
    q = query.lighting_fast_querier(cuda_dev, opt)
    xyz = tables["xyz"].to(cuda_dev).clone()
    g1 = q.grid_for(xyz)
    assert q.grid_for(xyz) is g1           # unchanged tensor: reused (the reference rebuilds every forward)
    xyz.add_(0.0)                          # in-place write bumps the version counter
    assert q.grid_for(xyz) is not g1
    q.clean_up()
    assert q._grid is None


@pytest.mark.parametrize("K", [1, 8, 16, 40])
def test_knn_other_K(tiny_scene, cuda_dev, K):
    import copy
    opt, tables, cam = tiny_scene
    opt = copy.copy(opt)
    opt.K = K
    raydir, pix = syn.make_rays(cam, syn.random_pixels(cam, 200, seed=4))
    (pidx, _, loc_w, _, ray_mask, _, _), ex = _oracle(opt, tables, cam, raydir)
    q = query.lighting_fast_querier(cuda_dev, opt)
    d = cuda_dev
    raw = q.query_raw(tables["xyz"].to(d), raydir.to(d), cam["campos"].to(d), cam["near"], cam["far"], ts=ex["ts"][:, :1].contiguous().to(d))
    sel = raw["ray_any"].cpu() > 0
    assert torch.equal(sel.to(torch.int8)[None], ray_mask)
    assert torch.equal(raw["sample_pidx"].cpu()[sel][None], pidx)


@pytest.mark.parametrize("name", ["tiny3", "tiny4"])
def test_query_on_other_scene_shapes(cuda_dev, name):
    """Miniatures of BASELINE configs 3 / 4 (16 mm voxels with P = 30 inside a room; 16 cm voxels with P = 26 on a street
    slab; SR = 24, near = 0.1, camera inside the point cloud): neighbour slots, positions and masks bit-exact against the
    oracle, with the sample depths handed to both sides."""
    opt, tables = syn.make_scene(name)
    cam = syn.make_camera(name)
    raydir, pix = syn.make_rays(cam, syn.random_pixels(cam, 600, seed=3))
    (pidx, loc_pers, loc_w, dirs, ray_mask, vsize, ranges), ex = _oracle(opt, tables, cam, raydir)
    assert int(ray_mask.sum()) > 50 and int((pidx >= 0).sum()) > 2000
    d = cuda_dev
    q = query.lighting_fast_querier(d, opt)
    raw = q.query_raw(tables["xyz"].to(d), raydir.to(d), cam["campos"].to(d), cam["near"], cam["far"], ts=ex["ts"].to(d))
    info = raw["grid"].info_host()
    assert info[_lib.GI_OVER_P] == 0 and info[_lib.GI_OVER_MAX_O] == 0
    if name == "tiny3":
        assert int(ex["grid"]["occ_numpnts"].max()) > 9      # buckets deeper than the lego P = 9
    sel = torch.nonzero(raw["ray_any"] > 0).squeeze(1)
    assert torch.equal((raw["ray_any"] > 0).cpu()[None].to(torch.int8), ray_mask)
    assert torch.equal(raw["sample_loc_w"][sel].cpu()[None], loc_w)
    assert torch.equal(raw["sample_pidx"][sel].cpu()[None], pidx), "neighbour slots differ from the oracle"
    assert np.array_equal(raw["grid"].hp["ranges"], ranges)


def test_sample_depths_sequential_fallback_on_wild_segments(cuda_dev):
    """pn_sample_depths takes a warp-parallel scan when the double running sum is provably exact (segments within 2^20 of each
    other) and the sequential chain otherwise: segment lengths spanning 40 binary orders (and denormals) must still give the
    host chain's bits."""
    D, R = 400, 64
    g = torch.Generator().manual_seed(5)
    base = (torch.rand(D, generator=g) * 2.0 ** torch.randint(-60, -20, (D,), generator=g).float()).float()
    base[::7] = 1e-41                                    # denormals
    u = torch.rand(1, R, D, generator=g)
    near, jitter = 0.25, 0.3
    seg = base.view(1, 1, D) * (1 + jitter * (u - 0.5))
    ends = near + torch.cat([torch.zeros(1, R, 1), torch.cumsum(seg, dim=2)], dim=2)
    want = (ends[:, :, :-1] + ends[:, :, 1:]) / 2
    ts = torch.empty((1, R, D), dtype=torch.float32, device=cuda_dev)
    _lib.check(_lib.lib().pn_sample_depths(_lib.ptr(base.to(cuda_dev)), _lib.ptr(u.reshape(R, D).contiguous().to(cuda_dev)), R, D, near, jitter,
                                           _lib.ptr(ts), _lib.stream()), "pn_sample_depths")
    assert torch.equal(ts.cpu(), want)
