"""Whole per-ray step on the GPU (fused path and seam-by-seam path) vs the CPU oracle pipeline."""
import copy

import pytest
import torch

from oracle import kinks, pipeline_ref
from pointnerf_b200 import synthetic as syn
from pointnerf_b200.neural_points import NeuralPointsRayMarching
from _util import assert_close, group_linear_grads

pytestmark = pytest.mark.gpu


def _model(opt, tables, dev, alpha_bias=60.0, seed=0):
    torch.manual_seed(seed)
    m = NeuralPointsRayMarching(opt, device=dev)
    with torch.no_grad():
        # a live density head: sigma*dist ~ 0.5 so that transmittance actually decays along the ray
        m.aggregator.alpha_branch[0].bias.fill_(alpha_bias)
        for n, p in m.aggregator.named_parameters():
            if p.dim() == 1 and not n.startswith("alpha"):
                p.normal_(0, 0.05)
    m.neural_points.set_points(tables["xyz"], tables["points_embeding"], tables["points_color"], tables["points_dir"],
                               tables["points_conf"], Rw2c=tables["Rw2c"])
    return m


def _step(model, cam, raydir, bg, wt, dev, fused, uniform=None, keep=None):
    for p in model.parameters():
        p.grad = None
    bg_d = bg.to(dev).clone().requires_grad_(True)
    out = model(cam["campos"].to(dev), raydir.to(dev), bg_color=bg_d, camrotc2w=cam["camrotc2w"].to(dev), near=cam["near"],
                far=cam["far"], h=cam["h"], w=cam["w"], intrinsic=cam["intrinsic"], fused=fused,
                uniform=uniform.to(dev) if uniform is not None else None)
    keep = torch.ones(1, raydir.shape[1], 1) if keep is None else keep
    loss = (out["coarse_raycolor"] * wt.to(dev)).sum() + (out["coarse_point_opacity"] ** 2 * keep.to(dev)).sum() * 0.1
    loss.backward()
    grads = {n: p.grad.detach().cpu().clone() for n, p in model.aggregator.named_parameters()}
    nps = model.neural_points
    for n in ("points_embeding", "points_conf", "points_dir", "points_color"):
        grads[n] = getattr(nps, n).grad.detach().cpu().clone()
    grads["bg_color"] = bg_d.grad.detach().cpu().clone()
    return out, grads


@pytest.mark.parametrize("train", [0, 1])
def test_step_matches_oracle(tiny_scene, cuda_dev, train):
    opt, tables, cam = tiny_scene
    opt = copy.copy(opt)
    opt.is_train = train
    R = 300
    raydir, _ = syn.make_rays(cam, syn.random_pixels(cam, R, seed=5))
    g = torch.Generator().manual_seed(2)
    bg = torch.rand(opt.shading_color_channel_num, generator=g)
    wt = torch.randn(1, R, opt.shading_color_channel_num, generator=g)
    uniform = torch.rand(1, R, opt.z_depth_dim, generator=g) if train else None
    model = _model(opt, tables, cuda_dev)
    model.aggregator.export_signs = True     # pn_agg_activation_signs: the LeakyReLU pattern of each forward
    keep = torch.ones(1, R, 1)               # EVERY ray carries loss weight (no kink masking)
    loss_fn = lambda o: ((o["coarse_raycolor"] * wt.to(o["coarse_raycolor"].dtype)).sum()
                         + (o["coarse_point_opacity"] ** 2 * keep.to(o["coarse_raycolor"].dtype)).sum() * 0.1)
    sd = model.aggregator.state_dict()
    ref = pipeline_ref.run(opt, tables, cam, raydir, bg, sd, uniform=uniform, need_grads=False)
    ref64 = pipeline_ref.run(opt, tables, cam, raydir, bg, sd, uniform=uniform, need_grads=False, dtype=torch.float64)
    assert float(ref["coarse_is_background"].min()) < 0.5, "test scene must be opaque enough to exercise compositing"
    hit = int((ref["ray_mask"] > 0).sum())
    assert hit > 0.5 * R
    for fused in (True, False):
        out, grads = _step(model, cam, raydir, bg, wt, cuda_dev, fused, uniform, keep)
        tag = "fused" if fused else "seams"
        assert torch.equal(out["ray_mask"].cpu(), ref["ray_mask"]), tag + ": ray mask differs from the oracle"
        # forward: against the UNCONDITIONED reference, on every ray
        for k in ("coarse_raycolor", "coarse_point_opacity", "coarse_is_background", "coarse_mask", "queried_shading"):
            assert_close(out[k], ref[k], "%s %s" % (tag, k), want64=ref64[k])
        # backward: the reference's autograd under the activation pattern this forward produced (the MLP is piecewise
        # linear; see oracle/kinks.py) - all rays weighted, every gradient tensor at 1e-4
        signs = tuple(t.cpu() for t in model.aggregator.last_call.signs)
        cnd = pipeline_ref.run(opt, tables, cam, raydir, bg, sd, uniform=uniform, loss_fn=loss_fn, frozen_signs=signs)
        cnd64 = pipeline_ref.run(opt, tables, cam, raydir, bg, sd, uniform=uniform, loss_fn=loss_fn, frozen_signs=signs,
                                 dtype=torch.float64)
        n_diff, n_all, worst = cnd64["pattern_disagreement"]
        assert n_all > 0 and n_diff <= 1e-4 * n_all and worst < kinks.PATTERN_TOL, (tag, n_diff, n_all, worst)
        for k in ("coarse_raycolor", "coarse_point_opacity"):     # freezing moved the oracle's forward by less than the tolerance
            assert_close(cnd[k], ref[k], "%s conditioned oracle %s" % (tag, k))
        g_ref, g_ref64 = group_linear_grads(cnd["grads"]), group_linear_grads(cnd64["grads"])
        for n, gr in group_linear_grads(grads).items():
            assert_close(gr, g_ref[n], "%s grad %s" % (tag, n), want64=g_ref64[n])


def test_fused_equals_seams_and_is_deterministic(tiny_scene, cuda_dev):
    opt, tables, cam = tiny_scene
    R = 500
    raydir, _ = syn.make_rays(cam, syn.random_pixels(cam, R, seed=8))
    g = torch.Generator().manual_seed(4)
    bg = torch.rand(opt.shading_color_channel_num, generator=g)
    wt = torch.randn(1, R, opt.shading_color_channel_num, generator=g)
    model = _model(opt, tables, cuda_dev)
    o1, g1 = _step(model, cam, raydir, bg, wt, cuda_dev, True)
    o2, g2 = _step(model, cam, raydir, bg, wt, cuda_dev, True)
    o3, g3 = _step(model, cam, raydir, bg, wt, cuda_dev, False)
    assert torch.equal(o1["coarse_raycolor"], o2["coarse_raycolor"])
    for n in g1:
        assert torch.equal(g1[n], g2[n]), "non-deterministic gradient " + n   # no atomics anywhere on the path
    assert torch.equal(o1["ray_mask"], o3["ray_mask"])
    assert_close(o1["coarse_raycolor"], o3["coarse_raycolor"], "fused vs seams colour", tol=1e-5)
    for n in g1:
        assert_close(g1[n], g3[n], "fused vs seams grad " + n, tol=2e-5)


def test_all_rays_miss(tiny_scene, cuda_dev):
    opt, tables, cam = tiny_scene
    model = _model(opt, tables, cuda_dev)
    raydir = -syn.make_rays(cam, syn.random_pixels(cam, 32))[0]
    bg = torch.rand(opt.shading_color_channel_num)
    for fused in (True, False):
        out = model(cam["campos"].to(cuda_dev), raydir.to(cuda_dev), bg_color=bg.to(cuda_dev), camrotc2w=cam["camrotc2w"].to(cuda_dev),
                    near=cam["near"], far=cam["far"], fused=fused)
        assert int(out["ray_mask"].sum()) == 0
        assert torch.equal(out["coarse_raycolor"].cpu(), bg[None, None, :].expand(1, 32, -1))
        assert float(out["coarse_is_background"].min()) == 1.0


def test_smoke_entry(cuda_dev):
    import __graft_entry__ as ge
    ge.smoke()


def test_prefetched_plan_gives_identical_step(tiny_scene, cuda_dev):
    """RayShardedStep(next_batch=...) queues the next batch's sampling / query between forward and backward; the step that
    consumes the plan must produce the same bits as a step that samples inside its own forward."""
    from pointnerf_b200.ray_shard import RayShardedStep
    opt, tables, cam = tiny_scene
    opt = copy.copy(opt)
    opt.is_train = 1
    d = cuda_dev
    R = 200
    g = torch.Generator().manual_seed(4)
    batches = []
    for k in range(2):
        raydir, _ = syn.make_rays(cam, syn.random_pixels(cam, R, seed=10 + k))
        batches.append(dict(cam=dict(campos=cam["campos"].to(d), camrotc2w=cam["camrotc2w"].to(d), near=cam["near"], far=cam["far"]),
                            raydir=raydir.to(d), uniform=torch.rand(1, R, opt.z_depth_dim, generator=g).to(d),
                            gt=torch.rand(1, R, opt.shading_color_channel_num, generator=g).to(d)))
    model = _model(opt, tables, d)
    bg = torch.nn.Parameter(torch.rand(opt.shading_color_channel_num, generator=g).to(d))
    st = RayShardedStep(model, bg)
    b0, b1 = batches
    st(b0["cam"], b0["raydir"], b0["gt"], uniform=b0["uniform"], next_batch=b1)
    assert st._plan is not None
    l1 = st(b1["cam"], b1["raydir"], b1["gt"], uniform=b1["uniform"])           # consumes the plan
    assert st._plan is None
    g1 = st.flat_grads().clone()
    l2 = st(b1["cam"], b1["raydir"], b1["gt"], uniform=b1["uniform"])           # samples inside the forward
    g2 = st.flat_grads().clone()
    assert torch.equal(l1, l2) and torch.equal(g1, g2)
    assert float(g1.abs().max()) > 0


@pytest.mark.parametrize("side", [True, False])
def test_render_frame_equals_one_forward(tiny_scene, cuda_dev, side):
    """Chunked full-frame inference (run/train_ft.py:test loop) gives the bits of ONE forward over all rays: a row's result
    does not depend on which other rows share its GEMM tile, nor on the stream its query ran on."""
    opt, tables, cam = tiny_scene
    opt = copy.copy(opt)
    opt.is_train = 0
    raydir, _ = syn.make_rays(cam, syn.all_pixels(cam))
    R = raydir.shape[1]
    bg = torch.rand(opt.shading_color_channel_num, generator=torch.Generator().manual_seed(4)).to(cuda_dev)
    model = _model(opt, tables, cuda_dev)
    args = (cam["campos"].to(cuda_dev), raydir.to(cuda_dev), bg, cam["camrotc2w"].to(cuda_dev), cam["near"], cam["far"])
    with torch.no_grad():
        whole = model(args[0], args[1], bg_color=bg, camrotc2w=args[3], near=cam["near"], far=cam["far"])
    keys = ("coarse_raycolor", "coarse_point_opacity", "ray_mask", "coarse_is_background")
    for chunk in (R, 1000, 333):
        out = model.render_frame(*args, chunk=chunk, keys=keys, side_stream=side)
        for k in keys:
            assert out[k].shape == whole[k].shape, (k, chunk)
            assert torch.equal(out[k], whole[k]), (k, chunk)
    empty = model.render_frame(args[0], args[1][:, :0], bg, args[3], cam["near"], cam["far"])
    assert empty["coarse_raycolor"].shape[1] == 0


@pytest.mark.parametrize("name", ["tiny3", "tiny4"])
def test_forward_on_other_scene_shapes(cuda_dev, name):
    """Whole forward on the miniatures of BASELINE configs 3 / 4 against the CPU oracle pipeline."""
    opt, tables = syn.make_scene(name)
    cam = syn.make_camera(name)
    raydir, _ = syn.make_rays(cam, syn.random_pixels(cam, 300, seed=8))
    bg = torch.rand(opt.shading_color_channel_num, generator=torch.Generator().manual_seed(2))
    model = _model(opt, tables, cuda_dev, alpha_bias=20.0 if name == "tiny4" else 60.0)
    sd = model.aggregator.state_dict()
    sd = {k: v.detach().cpu() for k, v in sd.items()}
    with torch.no_grad():
        out = model(cam["campos"].to(cuda_dev), raydir.to(cuda_dev), bg_color=bg.to(cuda_dev), camrotc2w=cam["camrotc2w"].to(cuda_dev),
                    near=cam["near"], far=cam["far"])
    ref = pipeline_ref.run(opt, tables, cam, raydir, bg, sd, need_grads=False)
    ref64 = pipeline_ref.run(opt, tables, cam, raydir, bg, sd, need_grads=False, dtype=torch.float64)
    assert torch.equal(out["ray_mask"].cpu(), ref["ray_mask"])
    assert_close(out["coarse_raycolor"], ref["coarse_raycolor"], "coarse_raycolor", want64=ref64["coarse_raycolor"])
    assert_close(out["coarse_point_opacity"], ref["coarse_point_opacity"], "coarse_point_opacity", want64=ref64["coarse_point_opacity"])
    assert_close(out["coarse_is_background"], ref["coarse_is_background"], "coarse_is_background", want64=ref64["coarse_is_background"])
