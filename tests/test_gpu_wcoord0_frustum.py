"""Perspective-frustum querier (wcoord_query = 0) on the GPU vs the reference's own kernels (golden vectors written by
oracle/make_golden.py from oracle/frustum_ref.py) - bit-exact for indices, masks and sample centres (integer / index work
and a fixed float chain), 1e-5 relative for the torch-op tail (pers2w), 1e-4 for the rendered step."""
import copy
import os

import numpy as np
import pytest
import torch

from oracle import frustum_model as fm, frustum_ref, kinks, pipeline_ref, torch_port as tp
from pointnerf_b200 import query_frustum, synthetic as syn
from pointnerf_b200.neural_points import NeuralPointsRayMarching
from _util import assert_close, group_linear_grads, rel_err

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["dtu", "crop"]


def _case(name):
    opt, xyz, cam, pix = syn.frustum_case(name)
    xyz_pers = tp.w2pers_points(xyz, cam["camrotc2w"], cam["campos"])          # CPU chain: the golden file's input, bit for bit
    g = np.load(os.path.join(GOLDEN, "frustum_%s_ref.npz" % name))
    assert np.array_equal(xyz_pers.numpy(), g["xyz_pers"])
    return opt, cam, pix, xyz_pers, g


def _query(q, cam, pix, xyz_pers, dev, uniform=None):
    return q.query_points(pix.to(dev), xyz_pers.to(dev), None, None, cam["h"], cam["w"], cam["intrinsic"], cam["near"], cam["far"],
                          None, cam["campos"].to(dev), cam["camrotc2w"].to(dev), uniform=uniform)


@pytest.mark.parametrize("name", CASES)
def test_frustum_stages_match_model(name, cuda_dev):
    """Grid (CSR, dilated occupancy), column walk and ray mask against the numpy statement of the same structures."""
    opt, cam, pix, xyz_pers, g = _case(name)
    q = query_frustum.lighting_fast_querier(cuda_dev, opt)
    hp = q.get_hyperparameters(cam["h"], cam["w"], cam["intrinsic"], cam["near"], cam["far"])
    xp, px = xyz_pers.reshape(-1, 3), pix[0]
    raw = q.query_raw(px.to(cuda_dev).contiguous(), xp.to(cuda_dev).contiguous(), hp)
    cells, occ, dims = fm.build(xp.numpy(), hp)
    V = dims[0] * dims[1] * dims[2]
    bits = np.unpackbits(raw["occ_bits"].cpu().numpy().view(np.uint8), bitorder="little")[:V].reshape(dims).astype(bool)
    assert np.array_equal(bits, occ), "dilated occupancy differs in %d cells" % int((bits != occ).sum())
    cs, bk = raw["cell_start"].cpu().numpy(), raw["bucket_pidx"].cpu().numpy()
    cnt = np.diff(cs)
    assert set(np.nonzero(cnt)[0].tolist()) == set(cells), "claimed cells differ"
    for c, members in cells.items():
        assert bk[cs[c]:cs[c + 1]].tolist() == members, "bucket of cell %d" % c
    sz, mask = fm.sample(px.numpy(), occ, dims, hp["vscale"], opt.SR)
    assert np.array_equal(raw["ray_mask"].cpu().numpy(), mask), "ray mask"
    assert np.array_equal(raw["sample_z"].cpu().numpy(), sz), "depth indices"
    assert np.array_equal(raw["hit_count"].cpu().numpy(), (sz >= 0).sum(1)), "hit count"
    rep = q.grid_report()
    assert rep["n_occ"] == len(cells) and rep["max_count"] == max(len(v) for v in cells.values()) and rep["over_P"] == 0
    assert rep["n_inside"] == sum(len(v) for v in cells.values())


@pytest.mark.parametrize("name", CASES)
def test_frustum_query_points_equals_reference(name, cuda_dev):
    opt, cam, pix, xyz_pers, g = _case(name)
    opt.is_train = 0
    q = query_frustum.lighting_fast_querier(cuda_dev, opt)
    pidx, loc, loc_w, dirs, mask, vsize, ranges = _query(q, cam, pix, xyz_pers, cuda_dev)
    assert np.array_equal(mask.cpu().numpy(), g["ray_mask"]) and mask.dtype == torch.int8
    assert pidx.shape == g["sample_pidx"].shape and pidx.dtype == torch.int32
    bad = int((pidx.cpu().numpy() != g["sample_pidx"]).sum())
    assert bad == 0, "%d of %d neighbour slots differ" % (bad, pidx.numel())
    assert np.array_equal(loc.cpu().numpy(), g["sample_loc"]), "sample centres must be bit-identical (fixed float chain, :532-536)"
    assert np.array_equal(np.asarray(vsize), g["vsize"]) and np.array_equal(np.asarray(ranges), g["ranges"])
    assert rel_err(loc_w, g["sample_loc_w"]) < 1e-5 and rel_err(dirs, g["ray_dirs"]) < 1e-5     # torch ops on the device
    assert int((pidx >= 0).sum()) > 10000
    if frustum_ref.available():   # the reference's kernels, live (oracle/_ref travels with the snapshot)
        live = frustum_ref.query_points(opt, pix, xyz_pers, cam["h"], cam["w"], cam["intrinsic"], cam["near"], cam["far"],
                                        cam["campos"], cam["camrotc2w"])
        assert np.array_equal(live[0].numpy(), pidx.cpu().numpy()) and np.array_equal(live[1].numpy(), loc.cpu().numpy())


def test_frustum_uniform_jitter(cuda_dev):
    opt, cam, pix, xyz_pers, g = _case("dtu")
    opt.is_train, opt.shpnt_jitter = 1, "uniform"
    q = query_frustum.lighting_fast_querier(cuda_dev, opt)
    out = _query(q, cam, pix, xyz_pers, cuda_dev, uniform=torch.from_numpy(g["uniform"]).to(cuda_dev))
    assert np.array_equal(out[0].cpu().numpy(), g["sample_pidx"])
    assert rel_err(out[1], g["jitter_sample_loc"]) < 1e-6 and rel_err(out[2], g["jitter_sample_loc_w"]) < 1e-5
    opt.shpnt_jitter = "gaussian"                       # random draw: only the bound of :128 can be checked
    out2 = _query(q, cam, pix, xyz_pers, cuda_dev)
    dz = (out2[1][..., 2].cpu() - torch.from_numpy(g["sample_loc"])[..., 2]).abs().max()
    assert 0 < float(dz) <= float(g["vsize"][2]) / 2 * (1 + 1e-4)


def test_frustum_edge_cases_and_refusals(cuda_dev):
    opt, cam, pix, xyz_pers, g = _case("crop")
    opt.is_train = 0
    q = query_frustum.lighting_fast_querier(cuda_dev, opt)
    # no rays
    out = _query(q, cam, pix[:, :0], xyz_pers, cuda_dev)
    assert out[0].shape == (1, 0, opt.SR, opt.K) and out[4].shape == (1, 0)
    # no points in the frustum: every ray masked out
    far_away = xyz_pers.clone()
    far_away[..., 2] += 100.0
    out = _query(q, cam, pix, far_away, cuda_dev)
    assert out[0].shape[1] == 0 and int(out[4].sum()) == 0
    # a single ray equals its row of the full query
    first = int(np.nonzero(g["ray_mask"][0])[0][0])
    out = _query(q, cam, pix[:, first:first + 1], xyz_pers, cuda_dev)
    assert np.array_equal(out[0].cpu().numpy()[0, 0], g["sample_pidx"][0, 0])
    for kw in (dict(NN=0), dict(inverse=1), dict(query_size=(3, 3, 3))):
        o2 = syn.config_opt("tiny", **{**syn.FRUSTUM_CASES["crop"]["opt"], **kw})
        with pytest.raises(NotImplementedError):
            _query(query_frustum.lighting_fast_querier(cuda_dev, o2), cam, pix, xyz_pers, cuda_dev)


@pytest.mark.skipif(not frustum_ref.available(), reason="oracle/_ref/libpn_ref_querier_frustum.so not built")
def test_step_with_frustum_querier_matches_oracle(cuda_dev):
    """wcoord_query = 0 through the reference seams (NeuralPoints -> PointAggregator -> ray_march) and through the fused
    driver, forward and backward, against the CPU oracle pipeline on the frustum querier: 1e-4 relative (fp32 or fp64
    evaluation at LeakyReLU kinks)."""
    opt, xyz, cam, pix = syn.frustum_case("dtu")
    opt = copy.copy(opt)
    opt.is_train = 0
    _, tables = syn.make_scene("tiny")
    pix = pix[:, :200]
    R = pix.shape[1]
    g = torch.Generator().manual_seed(2)
    bg = torch.rand(opt.shading_color_channel_num, generator=g)
    wt = torch.randn(1, R, opt.shading_color_channel_num, generator=g)
    torch.manual_seed(0)
    model = NeuralPointsRayMarching(opt, device=cuda_dev)
    with torch.no_grad():
        model.aggregator.alpha_branch[0].bias.fill_(60.0)
        for n, p in model.aggregator.named_parameters():
            if p.dim() == 1 and not n.startswith("alpha"):
                p.normal_(0, 0.05)
    model.neural_points.set_points(tables["xyz"], tables["points_embeding"], tables["points_color"], tables["points_dir"],
                                   tables["points_conf"], Rw2c=tables["Rw2c"])
    sd = model.aggregator.state_dict()
    model.aggregator.export_signs = True      # gradients are compared under the forward's own LeakyReLU pattern, on every ray
    keep = torch.ones(1, R, 1)
    loss_fn = lambda o: ((o["coarse_raycolor"] * wt.to(o["coarse_raycolor"].dtype)).sum()
                         + (o["coarse_point_opacity"] ** 2 * keep.to(o["coarse_raycolor"].dtype)).sum() * 0.1)
    ref = pipeline_ref.run(opt, tables, cam, None, bg, sd, need_grads=False, frustum_pix=pix)
    ref64 = pipeline_ref.run(opt, tables, cam, None, bg, sd, need_grads=False, dtype=torch.float64, frustum_pix=pix)
    assert float(ref["coarse_is_background"].min()) < 0.5
    dev = cuda_dev
    for fused in (False, True):   # the reference seams, then the fused driver (no ray compaction, no [R,SR,K,*] gathers)
        tag = "frustum fused " if fused else "frustum seams "
        for p_ in model.parameters():
            p_.grad = None
        bg_d = bg.to(dev).clone().requires_grad_(True)
        out = model(cam["campos"].to(dev), None, bg_color=bg_d, camrotc2w=cam["camrotc2w"].to(dev), pixel_idx=pix.to(dev),
                    near=cam["near"], far=cam["far"], h=cam["h"], w=cam["w"], intrinsic=cam["intrinsic"], fused=fused)
        assert torch.equal(out["ray_mask"].cpu(), ref["ray_mask"])
        for k in ("coarse_raycolor", "coarse_point_opacity", "coarse_is_background", "coarse_mask", "queried_shading"):
            assert_close(out[k], ref[k], tag + k, want64=ref64[k])
        loss = (out["coarse_raycolor"] * wt.to(dev)).sum() + (out["coarse_point_opacity"] ** 2 * keep.to(dev)).sum() * 0.1
        loss.backward()
        grads = {n: p.grad.detach().cpu().clone() for n, p in model.aggregator.named_parameters()}
        for n in ("points_embeding", "points_conf", "points_dir", "points_color"):
            grads[n] = getattr(model.neural_points, n).grad.detach().cpu().clone()
        grads["bg_color"] = bg_d.grad.detach().cpu().clone()
        signs = tuple(t.cpu() for t in model.aggregator.last_call.signs)
        cnd = pipeline_ref.run(opt, tables, cam, None, bg, sd, loss_fn=loss_fn, frustum_pix=pix, frozen_signs=signs)
        cnd64 = pipeline_ref.run(opt, tables, cam, None, bg, sd, loss_fn=loss_fn, dtype=torch.float64, frustum_pix=pix, frozen_signs=signs)
        n_diff, n_all, worst = cnd64["pattern_disagreement"]
        assert n_all > 0 and n_diff <= 1e-4 * n_all and worst < kinks.PATTERN_TOL, (tag, n_diff, n_all, worst)
        g_ref, g_ref64 = group_linear_grads(cnd["grads"]), group_linear_grads(cnd64["grads"])
        for n, gr in group_linear_grads(grads).items():
            assert_close(gr, g_ref[n], tag + "grad " + n, want64=g_ref64[n])
    with pytest.raises(NotImplementedError):   # the frustum querier is addressed by pixel, not by ray direction
        model(cam["campos"].to(dev), torch.zeros(1, R, 3, device=dev), bg_color=bg_d, camrotc2w=cam["camrotc2w"].to(dev),
              near=cam["near"], far=cam["far"])


def test_frustum_fused_equals_seams_with_jitter(cuda_dev):
    """Training-time sample jitter: the fused driver draws for all R rays, the seams for the R' kept ones; fed the same draw
    they render the same colours."""
    opt, xyz, cam, pix = syn.frustum_case("dtu")
    opt = copy.copy(opt)
    opt.is_train, opt.shpnt_jitter = 1, "uniform"
    _, tables = syn.make_scene("tiny")
    dev = cuda_dev
    pix = pix[:, :300]
    R = pix.shape[1]
    torch.manual_seed(0)
    model = NeuralPointsRayMarching(opt, device=dev)
    with torch.no_grad():
        model.aggregator.alpha_branch[0].bias.fill_(60.0)
    model.neural_points.set_points(tables["xyz"], tables["points_embeding"], tables["points_color"], tables["points_dir"],
                                   tables["points_conf"], Rw2c=tables["Rw2c"])
    bg = torch.rand(opt.shading_color_channel_num, generator=torch.Generator().manual_seed(4)).to(dev)
    u = torch.rand(1, R, opt.SR, generator=torch.Generator().manual_seed(9)).to(dev)
    kw = dict(bg_color=bg, camrotc2w=cam["camrotc2w"].to(dev), pixel_idx=pix.to(dev), near=cam["near"], far=cam["far"], h=cam["h"],
              w=cam["w"], intrinsic=cam["intrinsic"])
    with torch.no_grad():
        of = model(cam["campos"].to(dev), None, fused=True, uniform=u, **kw)
        keep = of["ray_mask"][0] > 0
        assert 0 < int(keep.sum()) < R
        os_ = model(cam["campos"].to(dev), None, fused=False, uniform=u[:, keep], **kw)
    assert torch.equal(of["ray_mask"], os_["ray_mask"])
    for k in ("coarse_raycolor", "coarse_point_opacity", "coarse_is_background", "queried_shading"):
        assert_close(of[k], os_[k], "jitter fused vs seams " + k, tol=1e-5)
    opt.is_train = 0
    with torch.no_grad():
        o0 = model(cam["campos"].to(dev), None, fused=True, **kw)
    assert rel_err(o0["coarse_raycolor"], of["coarse_raycolor"]) > 1e-5      # the jitter did move the samples


def test_render_frame_with_frustum_querier_equals_one_forward(cuda_dev):
    """render_frame on the perspective-frustum querier (rays addressed by pixel index, chunked, query of chunk i + 1 on the
    side stream): same image as one un-chunked forward over all pixels."""
    opt, xyz, cam, pix = syn.frustum_case("dtu")
    opt = copy.copy(opt)
    opt.is_train = 0
    _, tables = syn.make_scene("tiny")
    dev = cuda_dev
    pix = pix[:, :900]
    torch.manual_seed(0)
    model = NeuralPointsRayMarching(opt, device=dev)
    with torch.no_grad():
        model.aggregator.alpha_branch[0].bias.fill_(60.0)
    model.neural_points.set_points(tables["xyz"], tables["points_embeding"], tables["points_color"], tables["points_dir"],
                                   tables["points_conf"], Rw2c=tables["Rw2c"])
    bg = torch.rand(opt.shading_color_channel_num, generator=torch.Generator().manual_seed(4)).to(dev)
    kw = dict(h=cam["h"], w=cam["w"], intrinsic=cam["intrinsic"])
    with torch.no_grad():
        whole = model(cam["campos"].to(dev), None, bg_color=bg, camrotc2w=cam["camrotc2w"].to(dev), pixel_idx=pix.to(dev), near=cam["near"],
                      far=cam["far"], **kw)
        for side in (False, True):
            fr = model.render_frame(cam["campos"].to(dev), None, bg, cam["camrotc2w"].to(dev), cam["near"], cam["far"], chunk=256,
                                    keys=("coarse_raycolor", "ray_mask", "coarse_point_opacity"), side_stream=side, pixel_idx=pix.to(dev), **kw)
            assert torch.equal(fr["ray_mask"], whole["ray_mask"])
            assert torch.equal(fr["coarse_raycolor"], whole["coarse_raycolor"])
            assert torch.equal(fr["coarse_point_opacity"], whole["coarse_point_opacity"])
    assert int((whole["ray_mask"] > 0).sum()) > 100
    with pytest.raises(NotImplementedError):
        model.render_frame(cam["campos"].to(dev), torch.zeros(1, 10, 3, device=dev), bg, cam["camrotc2w"].to(dev), cam["near"], cam["far"])
