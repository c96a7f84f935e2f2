"""Writes tests/golden/*.npz from the UNMODIFIED reference (run in the build container: needs /root/reference).

TEST INFRASTRUCTURE (see oracle/__init__.py).  ``python -m oracle.make_golden`` from the repository root.
  aggregator_ref.npz  PointAggregator (models/aggregators/point_aggregators.py) outputs + autograd gradients
  aggregator_order1_ref.npz  the same with agg_intrp_order=1, point_features_dim=63
  ray_march_ref.npz   ray_march (models/rendering/diff_ray_marching.py:509-555) outputs + gradients
  querier_ref.npz     lighting_fast_querier kernels (the reference's CUDA string host-compiled by oracle/build_ref.py,
                      canonical ascending-thread schedule) on the seeded ``tiny`` scene
  driver_{world,frustum}_ref.npz  the reference's WHOLE per-ray driver (NeuralPointsRayMarching.forward + fill_invalid over
                      NeuralPoints.forward over lighting_fast_querier over its kernels, unmodified PointAggregator / ray_march;
                      oracle/ref_fragments.reference_hot_path) on 200 rays of the tiny scene: inputs, aggregator weights, outputs
                      (``python -m oracle.make_golden driver`` writes only these)
  frustum_{dtu,crop}_ref.npz  the perspective-frustum lighting_fast_querier (query_point_indices.py) the same way, through
                      oracle/frustum_ref.py, on synthetic.frustum_case(...)   (``python -m oracle.make_golden frustum`` writes
                      only these)
"""
import copy
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import build_ref, frustum_ref as fr, querier_ref as qr, ref_fragments as rf, ref_import, torch_port as tp  # noqa: E402
from pointnerf_b200 import synthetic as syn  # noqa: E402
from pointnerf_b200.options import make_opt  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def aggregator(name="aggregator_ref.npz", F=32, **opt_kw):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from test_oracle import _agg_inputs
    ns = ref_import.load()
    opt = make_opt(point_features_dim=F, **opt_kw)
    opt.agg_axis_weight = None
    torch.manual_seed(0)
    ref = ns.PointAggregator(opt)
    with torch.no_grad():
        for p in ref.parameters():
            if p.dim() == 1:
                p.normal_(0, 0.1)
    ins = _agg_inputs(seed=11, R=6, SR=4, K=8, F=F)
    gk = ["sampled_color", "sampled_dir", "sampled_conf", "sampled_embedding"]
    for k in gk:
        ins[k].requires_grad_(True)
    o = ref(*[ins[k] for k in ins], (0.004,) * 3, 0)
    wt = torch.randn(o[0].shape, generator=torch.Generator().manual_seed(3))
    loss = (o[0] * wt).sum() + (o[3] ** 2).sum()
    grads = torch.autograd.grad(loss, [ins[k] for k in gk] + list(ref.parameters()))
    d = {"in/" + k: v.detach().numpy() for k, v in ins.items()}
    d.update({"sd/" + k: v.numpy() for k, v in ref.state_dict().items()})
    d.update({"out/decoded": o[0].detach().numpy(), "out/ray_valid": o[1].numpy(), "out/weight": o[2].detach().numpy(),
              "out/conf_coefficient": o[3].detach().numpy(), "loss_weight": wt.numpy()})
    for n, g in zip(gk + [n for n, _ in ref.named_parameters()], grads):
        d["grad/" + n] = g.numpy()
    np.savez_compressed(os.path.join(OUT, name), **d)


def ray_march():
    ns = ref_import.load()
    g = torch.Generator().manual_seed(21)
    R, SR, C = 13, 24, 128
    feats = torch.randn(1, R, SR, C + 1, generator=g)
    feats[..., 0] = torch.rand(1, R, SR, generator=g) * 60
    dist = torch.rand(1, R, SR, generator=g) * 0.02
    valid = torch.rand(1, R, SR, generator=g) < 0.7
    valid[0, 0] = False
    bg = torch.rand(C, generator=g)
    feats.requires_grad_(True)
    bg.requires_grad_(True)
    o = ns.ray_march(dist, valid, feats, ns.radiance_render, ns.alpha_blend, bg)
    w_color, w_op = torch.randn(1, R, C, generator=g), torch.randn(1, R, SR, generator=g)
    gf, gb = torch.autograd.grad((o[0] * w_color).sum() + (o[2] * w_op).sum(), [feats, bg])
    d = {"in/feats": feats.detach().numpy(), "in/dist": dist.numpy(), "in/valid": valid.numpy(), "in/bg": bg.detach().numpy(),
         "w_color": w_color.numpy(), "w_op": w_op.numpy(), "grad/feats": gf.numpy(), "grad/bg": gb.numpy()}
    for i, n in enumerate(("ray_color", "point_color", "opacity", "acc_T", "blend_w", "bg_T", "bg_w")):
        d["out/" + n] = o[i].detach().numpy()
    np.savez_compressed(os.path.join(OUT, "ray_march_ref.npz"), **d)


def querier():
    build_ref.build()
    opt, tables = syn.make_scene("tiny")
    cam = syn.make_camera("tiny")
    raydir, _ = syn.make_rays(cam, syn.random_pixels(cam, 160, seed=12))
    (pidx, loc_pers, loc_w, dirs, ray_mask, _, ranges), ex = qr.query_points(
        opt, tables["xyz"][None], raydir, cam["campos"], cam["camrotc2w"], cam["near"], cam["far"], backend="ref")
    n_occ = int(ex["grid"]["occ_idx"][0])
    np.savez_compressed(os.path.join(OUT, "querier_ref.npz"), xyz=tables["xyz"].numpy(), raydir=raydir.numpy(),
                        sample_pidx=pidx.numpy(), sample_loc_w=loc_w.numpy(), ray_mask=ray_mask.numpy(), ranges=ranges,
                        occ_numpnts=ex["grid"]["occ_numpnts"][:n_occ])


def frustum():
    build_ref.build()
    for name in syn.FRUSTUM_CASES:
        opt, xyz, cam, pix = syn.frustum_case(name)
        xyz_pers = tp.w2pers_points(xyz, cam["camrotc2w"], cam["campos"])
        opt.is_train = 0
        pidx, loc, loc_w, dirs, ray_mask, vsize, ranges = fr.query_points(
            opt, pix, xyz_pers, cam["h"], cam["w"], cam["intrinsic"], cam["near"], cam["far"], cam["campos"], cam["camrotc2w"])
        opt.is_train = 1
        u = torch.rand(loc.shape[:3], generator=torch.Generator().manual_seed(5))
        jit = fr.query_points(opt, pix, xyz_pers, cam["h"], cam["w"], cam["intrinsic"], cam["near"], cam["far"], cam["campos"],
                              cam["camrotc2w"], uniform=u)
        np.savez_compressed(os.path.join(OUT, "frustum_%s_ref.npz" % name), xyz_pers=xyz_pers.numpy(), pixel_idx=pix.numpy(),
                            sample_pidx=pidx.numpy(), sample_loc=loc.numpy(), sample_loc_w=loc_w.numpy(), ray_dirs=dirs.numpy(),
                            ray_mask=ray_mask.numpy(), vsize=vsize, ranges=ranges, uniform=u.numpy(),
                            jitter_sample_loc=jit[1].numpy(), jitter_sample_loc_w=jit[2].numpy())


def driver_case(kind, R=200):
    """Seeded inputs of the driver golden files (also used by the tests that replay them)."""
    import copy
    if kind == "world":
        opt, tables = syn.make_scene("tiny")
        cam = syn.make_camera("tiny")
        raydir, pix = syn.make_rays(cam, syn.random_pixels(cam, R, seed=5))
    else:
        opt, _, cam, pix = syn.frustum_case("dtu")
        _, tables = syn.make_scene("tiny")
        pix, raydir = pix[:, :R].contiguous(), torch.zeros(1, R, 3)
    opt = copy.copy(opt)
    opt.is_train = 0
    g = torch.Generator().manual_seed(1)
    bg = torch.rand(opt.shading_color_channel_num, generator=g)
    return opt, tables, cam, raydir, pix, bg


def driver():
    torch.manual_seed(0)
    agg = tp.AggregatorPort(syn.config_opt("tiny"))
    g = torch.Generator().manual_seed(3)
    with torch.no_grad():
        agg.alpha_branch[0].bias.fill_(60.0)                  # a live density head (transmittance decays along the ray)
        for n, p in agg.named_parameters():
            if p.dim() == 1 and not n.startswith("alpha"):
                p.normal_(0, 0.05, generator=g)
    sd = agg.state_dict()
    np.savez_compressed(os.path.join(OUT, "driver_weights.npz"), **{k: v.numpy() for k, v in sd.items()})
    for kind in ("world", "frustum"):
        opt, tables, cam, raydir, pix, bg = driver_case(kind)
        drv, _ = rf.reference_hot_path(kind, opt, tables, sd)
        f32 = lambda v: torch.tensor([v], dtype=torch.float32)
        with torch.no_grad():
            out = drv.forward(cam["campos"], raydir, bg_color=bg, camrotc2w=cam["camrotc2w"], pixel_idx=pix, near=f32(cam["near"]),
                              far=f32(cam["far"]), h=torch.tensor([cam["h"]]), w=torch.tensor([cam["w"]]),
                              intrinsic=torch.from_numpy(cam["intrinsic"])[None])
        assert float(out["coarse_is_background"].min()) < 0.5 and 0 < int(out["ray_mask"].sum()) < pix.shape[1]
        np.savez_compressed(os.path.join(OUT, "driver_%s_ref.npz" % kind), raydir=raydir.numpy(), pixel_idx=pix.numpy(), bg=bg.numpy(),
                            **{k: out[k].numpy() for k in ("coarse_raycolor", "coarse_point_opacity", "coarse_is_background",
                                                           "coarse_mask", "queried_shading", "ray_mask")})


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    if sys.argv[1:] == ["driver"]:
        driver()
        sys.exit(0)
    if sys.argv[1:] == ["frustum"]:
        frustum()
        sys.exit(0)
    aggregator()
    aggregator("aggregator_order1_ref.npz", F=63, agg_intrp_order=1)      # the DTU feed-forward configuration (63-d features)
    ray_march()
    querier()
    frustum()
    driver()
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))
