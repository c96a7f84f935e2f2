"""CPU oracle of the whole per-ray step: query -> gather -> aggregate -> ray_dist -> ray_march -> fill_invalid.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Restates NeuralPointsRayMarching.forward
(models/neural_points_volumetric_model.py:272-352) on top of the querier oracle (oracle/querier_ref.py) and
the torch restatements (oracle/torch_port.py), and differentiates it with autograd exactly as the
reference's training step does (loss.backward through ray_march <- aggregator <- index_select).

Pinned in the build container against the reference's own driver assembled from its source and run whole on the CPU
(oracle/ref_fragments.reference_hot_path, tests/test_oracle.py::test_reference_hot_path_whole_equals_pipeline_oracle):
outputs bit for bit, gradients to 1e-6, with the world-coordinate and the perspective-frustum querier.
"""
import torch

from . import querier_ref, torch_port as tp


def run(opt, tables, cam, raydir, bg_color, agg_state, loss_weight=None, backend="port", uniform=None, need_grads=True,
        loss_fn=None, dtype=torch.float32, kink_tau=None, frustum_pix=None, frozen_signs=None):
    """tables: dict from pointnerf_b200.synthetic.make_scene (torch-CPU); cam: make_camera; raydir [1,R,3];
    bg_color [C]; agg_state: state_dict of the aggregator (reference names).  loss = sum(coarse_raycolor * loss_weight)
    unless ``loss_fn(out)`` is given.  Returns the fill_invalid dictionary plus query intermediates and ``grads``.

    ``dtype=torch.float64`` evaluates the SAME fp32 inputs and the same fp32 neighbour query with the aggregator and
    compositing arithmetic in double precision.  It adjudicates LeakyReLU kinks: a pre-activation that fp32 rounding puts on
    the other side of zero changes a gradient by a finite amount in ANY fp32 implementation (the reference's own CPU fp32
    run differs from its fp64 run by up to ~5e-4 on such inputs), so a result is accepted when it is within tolerance of
    the reference evaluated in fp32 OR in fp64 (tests/_util.py).

    ``kink_tau``: additionally report ``out["kink_rays"]`` (bool [1,R_full]) - rays holding a sample with a hidden unit at
    a LeakyReLU kink (oracle/kinks.py); gradient tests give those rays zero loss weight.

    ``frozen_signs`` = (pair_signs [4,M,W], sample_signs [3,S,C]) from pn_agg_activation_signs: the aggregator's LeakyReLUs
    run with that activation pattern frozen (torch_port.LeakyReLUProbe), i.e. the reference backward CONDITIONED on the
    pattern the checked forward produced; ``out["pattern_disagreement"]`` = (units where it differs from the oracle's own
    sign test, all units, the largest |pre-activation| / rms(layer) among those) - the tests hold the last number to
    kinks.TAU, so freezing can only adjudicate units at a kink, never hide an error.

    ``frustum_pix`` (int32 [1,R,2]): query with the perspective-frustum querier (oracle/frustum_ref.py, wcoord_query = 0)
    instead of the world-coordinate one; ``raydir`` is then unused (pers2w gives per-sample directions) and ``uniform`` is
    the [1,R',SR] draw of its sample jitter."""
    dt = dtype
    xyz = tables["xyz"]
    leaves = {k: tables[k].clone().to(dt).requires_grad_(need_grads) for k in ("points_embeding", "points_conf", "points_dir", "points_color")}
    bg = bg_color.clone().to(dt).requires_grad_(need_grads)
    agg = tp.AggregatorPort(opt)
    agg.load_state_dict({k: v.detach().cpu().clone() for k, v in agg_state.items()})
    agg = agg.to(dt)
    if frozen_signs is not None:
        agg.freeze_pattern(frozen_signs[0], frozen_signs[1])

    if frustum_pix is not None:
        from . import frustum_ref
        pidx, loc_pers, loc_w, dirs, ray_mask, vsize, ranges = frustum_ref.query_points(
            opt, frustum_pix, tp.w2pers_points(xyz, cam["camrotc2w"], cam["campos"]), cam["h"], cam["w"], cam["intrinsic"],
            cam["near"], cam["far"], cam["campos"], cam["camrotc2w"], uniform=uniform)
        extras = {}
    else:
        (pidx, loc_pers, loc_w, dirs, ray_mask, vsize, ranges), extras = querier_ref.query_points(
            opt, xyz[None], raydir, cam["campos"], cam["camrotc2w"], cam["near"], cam["far"], backend=backend, uniform=uniform)
    xyz_pers = tp.w2pers_points(xyz, cam["camrotc2w"], cam["campos"]).to(dt)   # fp32 arithmetic, as the inputs of the aggregator
    loc_pers, loc_w, dirs = loc_pers.to(dt), loc_w.to(dt), dirs.to(dt)
    g = tp.gather_point_tables(pidx, xyz.to(dt), xyz_pers, leaves["points_embeding"], leaves["points_color"], leaves["points_dir"],
                               leaves["points_conf"])
    probe = None
    if kink_tau is not None:
        from . import kinks
        probe = kinks.KinkProbe(agg, kink_tau)
    decoded, ray_valid, weight, conf_coef = agg(g["sampled_color"], tables["Rw2c"].to(dt), g["sampled_dir"], g["sampled_conf"],
                                                g["sampled_embedding"], g["sampled_xyz_pers"], g["sampled_xyz"],
                                                g["sample_pnt_mask"], loc_pers, loc_w, dirs, vsize, 0)
    kink_rays = None
    if probe is not None:
        amb = probe.ambiguous_samples(g["sample_pnt_mask"])          # [1,R'',SR]
        probe.close()
        kink_rays = torch.zeros(ray_mask.shape, dtype=torch.bool)
        kink_rays[ray_mask > 0] = amb.any(-1).view(-1)
    ray_dist = tp.ray_dist_from_samples(loc_pers, ray_valid, vsize, opt.raydist_mode_unit)
    ray_color, _, opacity, acc_T, blend_w, bg_T, _ = tp.ray_march(ray_dist, ray_valid, decoded, bg)
    shading = torch.logical_not(torch.any(ray_valid, dim=-1, keepdim=True)).repeat(1, 1, 3).to(dt)
    out = tp.fill_invalid(ray_mask, bg, ray_color, bg_T, opacity, shading)
    if frozen_signs is not None:
        out["pattern_disagreement"] = agg.pattern_disagreement()
    out.update(kink_rays=kink_rays, ray_mask=ray_mask, sample_pidx=pidx, sample_loc_w=loc_w, sample_loc=loc_pers, decoded_features=decoded.detach(),
               ray_valid=ray_valid, ray_dist=ray_dist.detach(), weight=weight, conf_coefficient=conf_coef, extras=extras,
               compact=dict(ray_color=ray_color.detach(), opacity=opacity.detach(), acc_T=acc_T.detach(), bg_T=bg_T.detach()))
    if need_grads:
        if loss_fn is not None:
            loss = loss_fn(out)
        else:
            loss = (out["coarse_raycolor"] * loss_weight).sum()
        params = dict(agg.named_parameters())
        names = list(leaves) + ["bg_color"] + list(params)
        tensors = list(leaves.values()) + [bg] + list(params.values())
        grads = torch.autograd.grad(loss, tensors, allow_unused=True)
        out["grads"] = {n: (gr if gr is not None else torch.zeros_like(t)) for n, gr, t in zip(names, grads, tensors)}
        out["loss"] = loss.detach()
    for k in ("coarse_raycolor", "coarse_point_opacity", "coarse_is_background", "coarse_mask", "queried_shading"):
        out[k] = out[k].detach()
    return out
