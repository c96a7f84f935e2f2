"""FusedTrainStep (pointnerf_b200/train_step.py): the training step without autograd, with the sample / pair counts kept on
the device under capacities, captured into CUDA graphs - against the autograd path of NeuralPointsRayMarching (itself held
to the CPU oracle by test_gpu_pipeline.py) and against the reference's colour-loss arithmetic (base_rendering_model.py:543-606)."""
import copy
import warnings

import pytest
import torch

from pointnerf_b200 import synthetic as syn
from pointnerf_b200.neural_points import NeuralPointsRayMarching
from pointnerf_b200.train_step import FusedTrainStep, color_loss_ray_weights
from _util import assert_close

pytestmark = pytest.mark.gpu


def _setup(dev, R, n_batches, train=1, seed=0):
    opt, tables = syn.make_scene("tiny")
    opt = copy.copy(opt)
    opt.is_train = train
    cam0 = syn.make_camera("tiny")
    torch.manual_seed(seed)
    model = NeuralPointsRayMarching(opt, device=dev)
    with torch.no_grad():
        model.aggregator.alpha_branch[0].bias.fill_(60.0)
        for n, p in model.aggregator.named_parameters():
            if p.dim() == 1 and not n.startswith("alpha"):
                p.normal_(0, 0.05)
    model.neural_points.set_points(tables["xyz"], tables["points_embeding"], tables["points_color"], tables["points_dir"],
                                   tables["points_conf"], Rw2c=tables["Rw2c"])
    g = torch.Generator().manual_seed(seed + 1)
    C = opt.shading_color_channel_num
    batches = []
    for b in range(n_batches):
        cam = syn.make_camera("tiny", view=b % 3)
        raydir, _ = syn.make_rays(cam, syn.random_pixels(cam, R, seed=20 + b))
        batches.append(dict(cam=dict(campos=cam["campos"].to(dev), camrotc2w=cam["camrotc2w"].to(dev), near=cam["near"], far=cam["far"]),
                            raydir=raydir.to(dev), gt=torch.rand(1, R, C, generator=g).to(dev),
                            uniform=torch.rand(1, R, opt.z_depth_dim, generator=g).to(dev) if train else None))
    bg = torch.nn.Parameter(torch.rand(C, generator=g).to(dev))
    return opt, model, batches, bg, cam0


def _reference_losses(out, gt, items, weights):
    """compute_losses (base_rendering_model.py:543-606) spelled with the reference's own ops on the model's outputs."""
    l2 = torch.nn.MSELoss()
    C = gt.shape[-1]
    total = 0
    for name, wt in zip(items, weights):
        if name.startswith("ray_masked"):
            sel = (out["ray_mask"] > 0)[..., None].expand(-1, -1, C)
            a, b = torch.masked_select(out["coarse_raycolor"], sel).reshape(1, -1, C), torch.masked_select(gt, sel).reshape(1, -1, C)
            loss = l2(a, b) if a.shape[1] > 0 else torch.tensor(0.0, device=gt.device)
        elif name.startswith("ray_miss"):
            sel = (out["ray_mask"] == 0)[..., None].expand(-1, -1, C)
            a, b = torch.masked_select(out["coarse_raycolor"], sel).reshape(1, -1, C), torch.masked_select(gt, sel).reshape(1, -1, C)
            loss = l2(a, b) * b.shape[1] if a.shape[1] > 0 else torch.tensor(0.0, device=gt.device)
        else:
            loss = l2(out["coarse_raycolor"], gt)
        total = total + (loss * wt + 1e-6)
    return total


def _autograd_step(model, bg, b, items, weights):
    for p in list(model.parameters()) + [bg]:
        p.grad = None
    out = model(b["cam"]["campos"], b["raydir"], bg_color=bg, camrotc2w=b["cam"]["camrotc2w"], near=b["cam"]["near"], far=b["cam"]["far"],
                uniform=b["uniform"])
    loss = _reference_losses(out, b["gt"], items, weights)
    loss.backward()
    grads = {n: p.grad.detach().clone() for n, p in model.aggregator.named_parameters()}
    for n in ("points_embeding", "points_conf", "points_dir", "points_color"):
        grads[n] = getattr(model.neural_points, n).grad.detach().clone()
    grads["bg_color"] = bg.grad.detach().clone()
    return out, loss.detach(), grads


def _fused_grads(model, bg):
    grads = {n: p.grad.detach().clone() for n, p in model.aggregator.named_parameters()}
    for n in ("points_embeding", "points_conf", "points_dir", "points_color"):
        grads[n] = getattr(model.neural_points, n).grad.detach().clone()
    grads["bg_color"] = bg.grad.detach().clone()
    return grads


@pytest.mark.parametrize("items,weights", [(("coarse_raycolor",), (1.0,)),
                                           (("ray_masked_coarse_raycolor", "ray_miss_coarse_raycolor", "coarse_raycolor"), (1.0, 0.0, 0.0)),   # lego.sh:152-153
                                           (("ray_masked_coarse_raycolor", "ray_miss_coarse_raycolor", "coarse_raycolor"), (1.0, 0.5, 0.25))])
def test_fused_step_equals_autograd_step_with_reference_losses(cuda_dev, items, weights):
    R = 300
    opt, model, batches, bg, cam = _setup(cuda_dev, R, 2)
    step = FusedTrainStep(model, bg, R, cam["near"], cam["far"], color_loss_items=items, color_loss_weights=weights, graph=False)
    caps = step.calibrate(batches, headroom=1.5, granule=256)
    assert caps[0] < R * opt.SR and caps[1] < R * opt.SR * opt.K
    n_const = 1e-6 * len(items)
    for b in batches:
        out, loss_ref, g_ref = _autograd_step(model, bg, b, items, weights)
        assert int((out["ray_mask"] > 0).sum()) > 0.5 * R and int((out["ray_mask"] == 0).sum()) > 0
        loss = step(b["cam"], b["raydir"], b["gt"], b["uniform"])
        assert abs(float(loss) + n_const - float(loss_ref)) <= 2e-6 * abs(float(loss_ref)) + 1e-9
        assert torch.equal(step.out["coarse_raycolor"], out["coarse_raycolor"])        # same kernels, same inputs: same bits
        assert torch.equal(step.out["coarse_point_opacity"], out["coarse_point_opacity"])
        got = _fused_grads(model, bg)
        for n in g_ref:
            # the loss gradient reaches the backward through torch's MSELoss chain on one side and pn_color_mse on the other
            # (different rounding of the same weights); the weight-gradient splits are cut for the capacity: 1e-5
            assert_close(got[n].reshape(g_ref[n].shape), g_ref[n], "grad " + n, tol=1e-5)


def test_graph_replay_equals_eager_sequence_bit_for_bit(cuda_dev):
    R = 256
    opt, model, batches, bg, cam = _setup(cuda_dev, R, 4)
    items, weights = ("ray_masked_coarse_raycolor", "coarse_raycolor"), (1.0, 0.3)
    eager = FusedTrainStep(model, bg, R, cam["near"], cam["far"], color_loss_items=items, color_loss_weights=weights, graph=False)
    caps = eager.calibrate(batches, headroom=1.3, granule=256)
    graphed = FusedTrainStep(model, bg, R, cam["near"], cam["far"], color_loss_items=items, color_loss_weights=weights, caps=caps, graph=True)
    for rnd in range(2):                           # second round: replays of graphs captured in the first
        for b in batches:
            l0 = eager(b["cam"], b["raydir"], b["gt"], b["uniform"]).clone()
            g0 = _fused_grads(model, bg)
            l1 = graphed(b["cam"], b["raydir"], b["gt"], b["uniform"]).clone()
            g1 = _fused_grads(model, bg)
            assert torch.equal(l0, l1)
            for n in g0:
                assert torch.equal(g0[n], g1[n]), n
    torch.cuda.synchronize()
    c = graphed.counts_host.tolist()
    assert c[0] == c[2] and c[1] == c[3] and 0 < c[0] <= caps[0] and 0 < c[1] <= caps[1]
    assert graphed.overflows == 0


@pytest.mark.parametrize("graph", [False, True])
def test_pipelined_steps_equal_plain_steps_bit_for_bit(cuda_dev, graph):
    """pipeline=True: the next batch's query runs inside the step, beside the backward; every batch's loss and gradients are
    the ones the plain sequence produces."""
    R = 256
    opt, model, batches, bg, cam = _setup(cuda_dev, R, 4)
    items, weights = ("ray_masked_coarse_raycolor", "ray_miss_coarse_raycolor", "coarse_raycolor"), (1.0, 0.2, 0.1)
    plain = FusedTrainStep(model, bg, R, cam["near"], cam["far"], color_loss_items=items, color_loss_weights=weights, graph=False)
    caps = plain.calibrate(batches, headroom=1.3, granule=256)
    want = []
    for b in batches:
        loss = plain(b["cam"], b["raydir"], b["gt"], b["uniform"]).clone()
        want.append((loss, _fused_grads(model, bg), plain.out["coarse_raycolor"].clone()))
    piped = FusedTrainStep(model, bg, R, cam["near"], cam["far"], color_loss_items=items, color_loss_weights=weights, caps=caps,
                           graph=graph, pipeline=True)
    with pytest.raises(RuntimeError):
        piped(batches[0]["cam"], batches[0]["raydir"], batches[0]["gt"], batches[0]["uniform"])
    b0 = batches[0]
    piped.prime(b0["cam"], b0["raydir"], b0["gt"], b0["uniform"])
    n = len(batches)
    for i in range(2 * n):                        # step(batch i + 1) works on batch i
        nb = batches[(i + 1) % n]
        loss = piped(nb["cam"], nb["raydir"], nb["gt"], nb["uniform"]).clone()
        got = _fused_grads(model, bg)
        wl, wg, wc = want[i % n]
        assert torch.equal(loss, wl), i
        assert torch.equal(piped.out["coarse_raycolor"], wc)
        for k in wg:
            assert torch.equal(got[k], wg[k]), (i, k)


def test_capacity_overflow_truncates_in_slot_order_is_flagged_and_grows(cuda_dev):
    R = 256
    opt, model, batches, bg, cam = _setup(cuda_dev, R, 1, train=0)
    b = batches[0]
    full = FusedTrainStep(model, bg, R, cam["near"], cam["far"], graph=False)
    full.calibrate(batches, headroom=1.0, granule=1)
    full(b["cam"], b["raydir"], b["gt"])
    torch.cuda.synchronize()
    S, M = full.counts_host.tolist()[:2]
    small = FusedTrainStep(model, bg, R, cam["near"], cam["far"], caps=(S // 2, M), graph=True)
    loss = small(b["cam"], b["raydir"], b["gt"])
    torch.cuda.synchronize()
    c = small.counts_host.tolist()
    assert c[2] == S and c[3] == M and c[0] == S // 2 and c[1] < M          # kept a prefix, reported the totals
    assert torch.isfinite(loss)
    # rays are processed in slot order: the rays whose samples all fit are rendered exactly as in the full run
    slot_row = small.out["slot_row"]
    any_full = full.out["slot_row"] >= 0
    intact = ((slot_row >= 0) == any_full).all(-1)
    assert int(intact.sum()) > 10 and int((~intact).sum()) > 10
    assert torch.equal(small.out["coarse_raycolor"][0][intact], full.out["coarse_raycolor"][0][intact])
    with warnings.catch_warnings(record=True) as wlist:
        warnings.simplefilter("always")
        small(b["cam"], b["raydir"], b["gt"])                               # the next call sees the flag, grows and re-captures
    assert small.overflows == 1 and any("capacities" in str(w.message) for w in wlist)
    torch.cuda.synchronize()
    c = small.counts_host.tolist()
    assert c[0] == S and c[1] == M
    assert torch.equal(small.out["coarse_raycolor"], full.out["coarse_raycolor"])


def test_color_loss_ray_weights_formula():
    ray_any = torch.tensor([1, 0, 3, 0, 0, 2], dtype=torch.int32)
    w = color_loss_ray_weights(ray_any, ("ray_masked_x", "ray_miss_x", "x"), (2.0, 3.0, 5.0), channels=4, global_rays=12)
    m = torch.tensor([1., 0, 1, 0, 0, 1])
    want = m * 2.0 / (3 * 4) + (1 - m) * 3.0 / 4 + 5.0 / (12 * 4)
    assert torch.allclose(w, want, rtol=1e-6)
    w2 = color_loss_ray_weights(ray_any, ("ray_masked_x",), (1.0,), channels=4, global_rays=12, masked_count=torch.tensor([7.0]))
    assert torch.allclose(w2, m / (7 * 4), rtol=1e-6)


@pytest.mark.parametrize("graph", [False, True])
@pytest.mark.parametrize("items,weights", [(("coarse_raycolor",), (1.0,)),
                                           (("ray_masked_coarse_raycolor", "ray_miss_coarse_raycolor", "coarse_raycolor"), (1.0, 0.5, 0.25))])
def test_fused_step_on_frustum_querier(cuda_dev, graph, items, weights):
    """wcoord_query = 0 (the DTU feed-forward scripts): rays by pixel index, the per-view grid built inside the query graph,
    per-sample directions from pers2w, the view's own voxel size, uniform depth jitter fed as an input - against the
    autograd path of NeuralPointsRayMarching on the same querier (held to the CPU oracle by test_gpu_wcoord0_frustum.py)."""
    dev = cuda_dev
    opt, xyz, cam, pix = syn.frustum_case("dtu")
    opt = copy.copy(opt)
    opt.is_train, opt.shpnt_jitter = 1, "uniform"
    _, tables = syn.make_scene("tiny")
    torch.manual_seed(0)
    model = NeuralPointsRayMarching(opt, device=dev)
    with torch.no_grad():
        model.aggregator.alpha_branch[0].bias.fill_(60.0)
    model.neural_points.set_points(tables["xyz"], tables["points_embeding"], tables["points_color"], tables["points_dir"],
                                   tables["points_conf"], Rw2c=tables["Rw2c"])
    R, C = 300, opt.shading_color_channel_num
    g = torch.Generator().manual_seed(3)
    bg = torch.nn.Parameter(torch.rand(C, generator=g).to(dev))
    view = dict(h=cam["h"], w=cam["w"], intrinsic=cam["intrinsic"])
    camd = dict(campos=cam["campos"].to(dev), camrotc2w=cam["camrotc2w"].to(dev), near=cam["near"], far=cam["far"])
    step = FusedTrainStep(model, bg, R, cam["near"], cam["far"], color_loss_items=items, color_loss_weights=weights, graph=graph, view=view)
    batches = []
    for b in range(3):
        sel = torch.randperm(pix.shape[1], generator=g)[:R]
        batches.append(dict(cam=camd, pixel_idx=pix[:, sel].to(dev), gt=torch.rand(1, R, C, generator=g).to(dev),
                            uniform=torch.rand(1, R, opt.SR, generator=g).to(dev)))
    step.calibrate(batches, headroom=1.5, granule=256)
    n_const = 1e-6 * len(items)
    for b in batches:
        for p in list(model.parameters()) + [bg]:
            p.grad = None
        out = model(camd["campos"], None, bg_color=bg, camrotc2w=camd["camrotc2w"], pixel_idx=b["pixel_idx"], near=cam["near"], far=cam["far"],
                    h=cam["h"], w=cam["w"], intrinsic=cam["intrinsic"], uniform=b["uniform"].clone())
        assert 0 < int((out["ray_mask"] > 0).sum()) < R
        loss_ref = _reference_losses(out, b["gt"], items, weights)
        loss_ref.backward()
        g_ref = _fused_grads(model, bg)
        loss = step(b["cam"], None, b["gt"], b["uniform"], pixel_idx=b["pixel_idx"])
        loss_ref = float(loss_ref.detach())
        assert abs(float(loss) + n_const - loss_ref) <= 2e-6 * abs(loss_ref) + 1e-9
        assert torch.equal(step.out["coarse_raycolor"], out["coarse_raycolor"].detach())
        got = _fused_grads(model, bg)
        for n in g_ref:
            assert_close(got[n].reshape(g_ref[n].shape), g_ref[n], "frustum step grad " + n, tol=1e-5)
    assert step.overflows == 0
