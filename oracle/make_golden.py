"""Writes tests/golden/*.npz from the UNMODIFIED reference (run in the build container: needs /root/reference).

TEST INFRASTRUCTURE (see oracle/__init__.py).  ``python -m oracle.make_golden`` from the repository root.
  aggregator_ref.npz  PointAggregator (models/aggregators/point_aggregators.py) outputs + autograd gradients
  aggregator_order1_ref.npz  the same with agg_intrp_order=1, point_features_dim=63
  ray_march_ref.npz   ray_march (models/rendering/diff_ray_marching.py:509-555) outputs + gradients
  querier_ref.npz     lighting_fast_querier kernels (the reference's CUDA string host-compiled by oracle/build_ref.py,
                      canonical ascending-thread schedule) on the seeded ``tiny`` scene
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

from oracle import build_ref, frustum_ref as fr, querier_ref as qr, ref_import, torch_port as tp  # noqa: E402
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


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    if sys.argv[1:] == ["frustum"]:
        frustum()
        sys.exit(0)
    aggregator()
    aggregator("aggregator_order1_ref.npz", F=63, agg_intrp_order=1)      # the DTU feed-forward configuration (63-d features)
    ray_march()
    querier()
    frustum()
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))
