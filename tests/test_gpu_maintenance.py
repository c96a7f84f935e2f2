"""Point-cloud maintenance between passes: probe outputs, prune, grow - CUDA vs the CPU oracle."""
import copy

import pytest
import torch

from oracle import maintenance_port as mp, pipeline_ref, torch_port as tp
from pointnerf_b200 import synthetic as syn
from pointnerf_b200.neural_points import NeuralPointsRayMarching
from _util import assert_close

pytestmark = pytest.mark.gpu


def _model(opt, tables, dev):
    torch.manual_seed(0)
    m = NeuralPointsRayMarching(opt, device=dev)
    with torch.no_grad():
        m.aggregator.alpha_branch[0].bias.fill_(60.0)
    m.neural_points.set_points(tables["xyz"], tables["points_embeding"], tables["points_color"], tables["points_dir"],
                               tables["points_conf"], Rw2c=tables["Rw2c"])
    return m


def test_probe_outputs_match_oracle(tiny_scene, cuda_dev):
    opt, tables, cam = tiny_scene
    opt = copy.copy(opt)
    opt.is_train, opt.prob = 0, 1
    R = 400
    raydir, _ = syn.make_rays(cam, syn.random_pixels(cam, R, seed=9))
    bg = torch.rand(opt.shading_color_channel_num, generator=torch.Generator().manual_seed(1))
    model = _model(opt, tables, cuda_dev)
    with torch.no_grad():
        out = model(cam["campos"].to(cuda_dev), raydir.to(cuda_dev), bg_color=bg.to(cuda_dev), camrotc2w=cam["camrotc2w"].to(cuda_dev),
                    near=cam["near"], far=cam["far"])
    sd = {k: v.detach().cpu() for k, v in model.aggregator.state_dict().items()}
    ref = pipeline_ref.run(opt, tables, cam, raydir, bg, sd, need_grads=False)
    xyz_pers = tp.w2pers_points(tables["xyz"], cam["camrotc2w"], cam["campos"])
    g = tp.gather_point_tables(ref["sample_pidx"], tables["xyz"], xyz_pers, tables["points_embeding"], tables["points_color"],
                               tables["points_dir"], tables["points_conf"])
    want = mp.probe_outputs(ref["compact"]["opacity"], ref["sample_loc_w"], ref["weight"], ref["conf_coefficient"], g["sampled_xyz"],
                            g["sampled_color"], g["sampled_dir"], g["sampled_conf"], g["sampled_embedding"])
    want = mp.unmask(ref["ray_mask"], want)
    hit = ref["ray_mask"][0] > 0
    assert int(hit.sum()) > 50
    # (a) the probe kernel on the ORACLE's opacities / weights / indices: the arg-max slot is index work - bit-exact on every
    #     hit ray (tie-break: the first maximum, like torch.max)
    R_, SR_, K_ = raydir.shape[1], opt.SR, opt.K
    dev = cuda_dev

    def full(t, fill):      # [1,R'',...] of the kept rays -> [R,...]
        o = torch.full((R_,) + tuple(t.shape[2:]), fill, dtype=t.dtype)
        o[hit] = t[0]
        return o
    with torch.no_grad():
        exact = model.probe_outputs(ref["coarse_point_opacity"][0].to(dev), ref["ray_mask"][0].to(torch.int32).to(dev),
                                    full(ref["sample_loc_w"], 0.0).to(dev), full(ref["sample_pidx"].to(torch.int32), -1).to(dev),
                                    full(ref["weight"].float(), 0.0).to(dev))
    assert torch.equal(exact["ray_max_slot"].cpu()[0][hit], want["ray_max_slot"][0][hit].int())
    assert torch.equal(exact["ray_max_sample_loc_w"].cpu()[0][hit], want["ray_max_sample_loc_w"][0][hit])
    # (b) end to end: OUR opacities differ from the oracle's by fp32 rounding, so the arg-max may move between two samples
    #     whose reference opacities tie to rounding (the reference's own choice there depends on its summation order)
    got_slot = out["ray_max_slot"].cpu()[0]
    same = (got_slot == want["ray_max_slot"][0].int()) & hit
    diff = hit & ~same
    if bool(diff.any()):
        op = ref["coarse_point_opacity"][0]
        a = op[diff].gather(1, got_slot[diff].long()[:, None])
        b = op[diff].gather(1, want["ray_max_slot"][0][diff][:, None])
        assert float((a - b).abs().max()) < 1e-6
    assert bool((got_slot[~hit] == -1).all())
    for k in ("ray_max_shading_opacity", "ray_max_sample_loc_w", "ray_max_far_dist", "shading_avg_color", "shading_avg_dir",
              "shading_avg_conf", "shading_avg_embedding"):
        got = out[k].cpu()
        assert got.shape == want[k].shape, (k, got.shape, want[k].shape)
        assert bool((got[0][~hit] == 0).all()), k
        assert_close(got[0][same], want[k][0][same], k)
    assert torch.equal(out["ray_max_sample_loc_w"].cpu()[0][same], want["ray_max_sample_loc_w"][0][same])


@pytest.mark.parametrize("thresh,n_add", [(0.4, 0), (0.4, 257), (None, 100), (2.0, 0), (2.0, 5), (0.0, 0)])
def test_prune_and_grow_match_oracle(tiny_scene, cuda_dev, thresh, n_add):
    opt, tables, cam = tiny_scene
    model = _model(copy.copy(opt), tables, cuda_dev)
    nps = model.neural_points
    g = torch.Generator().manual_seed(3)
    F_ = tables["points_embeding"].shape[-1]
    add = dict(add_xyz=torch.rand(n_add, 3, generator=g) * 0.1, add_embedding=torch.rand(n_add, F_, generator=g), add_color=torch.rand(n_add, 3, generator=g),
               add_dir=torch.rand(n_add, 3, generator=g), add_conf=torch.rand(n_add, 1, generator=g))
    want = {k: v for k, v in tables.items()}
    if thresh is not None:
        want = mp.prune(want, thresh)
    if n_add:
        want = mp.grow_points(want, **add)
    flags = {n: getattr(nps, n).requires_grad for n in ("points_embeding", "points_conf", "points_dir", "points_color")}
    if thresh is not None and n_add == 0:
        kept = nps.prune(thresh)
        assert kept == want["xyz"].shape[0]
    elif thresh is None:
        nps.grow_points(**{k: v.to(cuda_dev) for k, v in add.items()})
    else:
        nps.prune_and_grow(thresh, **{k: v.to(cuda_dev) for k, v in add.items()})
    assert torch.equal(nps.xyz.detach().cpu(), want["xyz"])
    for n in ("points_embeding", "points_conf", "points_dir", "points_color"):
        assert torch.equal(getattr(nps, n).detach().cpu(), want[n]), n
        assert getattr(nps, n).requires_grad == flags[n]
    assert isinstance(nps.xyz, torch.nn.Parameter) and not nps.xyz.requires_grad
    if want["xyz"].shape[0] > 100:
        # the grid of the old point set is gone: a query on the new set matches the oracle on the new tables
        raydir, _ = syn.make_rays(cam, syn.random_pixels(cam, 128, seed=2))
        bg = torch.zeros(opt.shading_color_channel_num)
        with torch.no_grad():
            out = model(cam["campos"].to(cuda_dev), raydir.to(cuda_dev), bg_color=bg.to(cuda_dev), camrotc2w=cam["camrotc2w"].to(cuda_dev),
                        near=cam["near"], far=cam["far"])
        assert bool(torch.isfinite(out["coarse_raycolor"]).all())
        if n_add == 0:   # (random new points may overfill a voxel beyond P, where the reference picks a random victim)
            sd = {k: v.detach().cpu() for k, v in model.aggregator.state_dict().items()}
            ref = pipeline_ref.run(opt, want, cam, raydir, bg, sd, need_grads=False)
            assert torch.equal(out["ray_mask"].cpu(), ref["ray_mask"])
            assert_close(out["coarse_raycolor"], ref["coarse_raycolor"], "coarse_raycolor after prune")


def test_prune_full_size_property(cuda_dev):
    """1M rows (cfg2 scale): the one-pass compaction equals torch's boolean-mask indexing of every table."""
    from pointnerf_b200.neural_points import NeuralPoints
    from pointnerf_b200.synthetic import config_opt
    opt = config_opt("cfg2")
    N, F_ = 1_000_000, 32
    g = torch.Generator(device="cuda").manual_seed(1)
    xyz = torch.rand(N, 3, device=cuda_dev, generator=g)
    emb = torch.rand(1, N, F_, device=cuda_dev, generator=g)
    conf, pdir, col = torch.rand(1, N, 1, device=cuda_dev, generator=g), torch.rand(1, N, 3, device=cuda_dev, generator=g), torch.rand(1, N, 3, device=cuda_dev, generator=g)
    nps = NeuralPoints(F_, 0, opt, cuda_dev)
    nps.set_points(xyz, emb, col, pdir, conf)
    mask = conf[0, :, 0] >= 0.1
    kept = nps.prune(0.1)
    assert kept == int(mask.sum())
    assert torch.equal(nps.xyz.detach(), xyz[mask]) and torch.equal(nps.points_embeding.detach(), emb[:, mask])
    assert torch.equal(nps.points_conf.detach(), conf[:, mask]) and torch.equal(nps.points_dir.detach(), pdir[:, mask])
    assert torch.equal(nps.points_color.detach(), col[:, mask])
