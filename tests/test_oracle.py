"""The oracle itself (CPU): restatements vs the unmodified reference (build container) and vs committed golden vectors."""
import copy
import os

import numpy as np
import pytest
import torch

from oracle import querier_ref as qr, ref_import, torch_port as tp
from pointnerf_b200 import synthetic as syn
from pointnerf_b200.options import make_opt
from _util import rel_err

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
needs_ref = pytest.mark.skipif(not ref_import.available(), reason="reference tree not present (GPU box)")


def _agg_inputs(seed=1, R=7, SR=5, K=8, F=32):
    g = torch.Generator().manual_seed(seed)

    def rn(*s):
        return torch.randn(*s, generator=g)
    B = 1
    mask = torch.rand(B, R, SR, K, generator=g) > 0.4
    mask[0, 1] = False
    mask[0, 2, 3] = False
    return dict(sampled_color=rn(B, R, SR, K, 3), sampled_Rw2c=torch.linalg.qr(rn(3, 3))[0], sampled_dir=rn(B, R, SR, K, 3),
                sampled_conf=torch.rand(B, R, SR, K, 1, generator=g) * 1.2 - 0.1, sampled_embedding=rn(B, R, SR, K, F) * 0.3,
                sampled_xyz_pers=rn(B, R, SR, K, 3) + torch.tensor([0, 0, 4.]), sampled_xyz=rn(B, R, SR, K, 3) * 0.01,
                sample_pnt_mask=mask, sample_loc=rn(B, R, SR, 3) + torch.tensor([0, 0, 4.]), sample_loc_w=rn(B, R, SR, 3) * 0.01,
                sample_ray_dirs=rn(B, R, SR, 3))


@needs_ref
@pytest.mark.parametrize("kw", [{}, dict(agg_intrp_order=1), dict(agg_intrp_order=1, point_features_dim=63)])
def test_aggregator_port_equals_reference(kw):
    ns = ref_import.load()
    opt = make_opt(**kw)
    opt_ref = copy.copy(opt)
    opt_ref.agg_axis_weight = None          # the reference builds this tensor on "cuda" (:247); None takes the same branch
    torch.manual_seed(0)
    ref = ns.PointAggregator(opt_ref)
    port = tp.AggregatorPort(opt)
    port.load_state_dict(ref.state_dict())
    ins = _agg_inputs(F=opt.point_features_dim)
    gk = ["sampled_color", "sampled_dir", "sampled_conf", "sampled_embedding"]
    for k in gk:
        ins[k].requires_grad_(True)
    res = []
    wt = None
    for m in (ref, port):
        o = m(*[ins[k] for k in ins], (0.004,) * 3, 0)
        if wt is None:
            wt = torch.randn(o[0].shape, generator=torch.Generator().manual_seed(3))
        loss = (o[0] * wt).sum() + (o[3] ** 2).sum()
        res.append((o, torch.autograd.grad(loss, [ins[k] for k in gk] + list(m.parameters()))))
    (o1, g1), (o2, g2) = res
    assert torch.equal(o1[0], o2[0]) and torch.equal(o1[1], o2[1]) and torch.equal(o1[2], o2[2]) and torch.equal(o1[3], o2[3])
    for a, b in zip(g1, g2):
        assert rel_err(b, a) < 1e-6


@needs_ref
def test_ray_march_port_equals_reference():
    ns = ref_import.load()
    g = torch.Generator().manual_seed(2)
    R, SR, C = 11, 24, 128
    feats = torch.randn(1, R, SR, C + 1, generator=g)
    feats[..., 0] = torch.rand(1, R, SR, generator=g) * 40
    dist = torch.rand(1, R, SR, generator=g) * 0.02
    valid = torch.rand(1, R, SR, generator=g) < 0.7
    bg = torch.rand(C, generator=g)
    a = ns.ray_march(dist, valid, feats, ns.radiance_render, ns.alpha_blend, bg)
    b = tp.ray_march(dist, valid, feats, bg)
    for x, y in zip(a, b):
        assert torch.equal(x, y)
    a5 = ns.alpha_ray_march(dist, valid, feats, ns.alpha_blend)
    for x, y in zip(a5, (b[2], b[3], b[4], b[5], b[6])):
        assert torch.equal(x, y)


@needs_ref
def test_sampler_and_pe_port_equal_reference():
    ns = ref_import.load()
    g = torch.Generator().manual_seed(4)
    raydir = torch.randn(1, 9, 3, generator=g)
    campos = torch.randn(1, 3, generator=g)
    raypos, _, _, ts = ns.near_far_linear_ray_generation(campos, raydir, 64, torch.tensor([[[2.0]]]), torch.tensor([[[6.0]]]), jitter=0.)
    ts_p = tp.sample_ts(raydir, 64, 2.0, 6.0)
    assert torch.equal(ts, ts_p)
    assert torch.equal(raypos, tp.raypos_from_ts(campos, raydir, ts_p))
    x = torch.randn(5, 7, 3, generator=g)
    for freqs, ori in ((4, True), (5, False), (3, False)):
        assert torch.equal(ns.positional_encoding(x, freqs, ori=ori), tp.positional_encoding(x, freqs, ori=ori))


def test_querier_backends_agree(tiny_scene):
    """The reference's own kernel string compiled for the host (oracle/_ref) vs the C port: identical tables and results."""
    opt, tables, cam = tiny_scene
    if not os.path.isfile(qr.REF_SO):
        pytest.skip("oracle/_ref not built (needs the reference tree once)")
    raydir, _ = syn.make_rays(cam, syn.random_pixels(cam, 400, seed=1))
    outs = {}
    for be in ("ref", "port"):
        outs[be] = qr.query_points(opt, tables["xyz"][None], raydir, cam["campos"], cam["camrotc2w"], cam["near"], cam["far"], backend=be)
    (a, ea), (b, eb) = outs["ref"], outs["port"]
    for i in (0, 1, 2, 3, 4):
        assert torch.equal(a[i], b[i]), i
    for k in ("coor_occ", "coor_2_occ", "occ_numpnts", "occ_2_pnts", "occ_2_coor"):
        assert np.array_equal(ea["grid"][k], eb["grid"][k]), k
    assert a[0].shape[1] > 100 and int((a[0] >= 0).sum()) > 10000


def test_hyperparameters_host_arithmetic(tiny_scene):
    """pointnerf_b200.synthetic.grid_hyperparameters (used by the product) == the oracle's get_hyperparameters."""
    opt, tables, cam = tiny_scene
    xyz = tables["xyz"]
    a = qr.get_hyperparameters(opt, xyz[None])
    b = syn.grid_hyperparameters(opt, xyz.min(0)[0].numpy(), xyz.max(0)[0].numpy())
    for k in ("ranges", "scaled_vsize", "scaled_vdim", "radius_limit2"):
        assert np.array_equal(np.asarray(a[k]), np.asarray(b[k])), k
    for name in ("cfg2", "cfg3"):
        o = syn.config_opt(name)
        lo, hi = np.asarray(o.ranges[:3], np.float32) * 0.9, np.asarray(o.ranges[3:], np.float32) * 0.9
        pts = torch.tensor(np.stack([lo, hi]))
        a = qr.get_hyperparameters(o, pts[None])
        b = syn.grid_hyperparameters(o, lo, hi)
        for k in ("ranges", "scaled_vsize", "scaled_vdim", "radius_limit2"):
            assert np.array_equal(np.asarray(a[k]), np.asarray(b[k])), (name, k)


def test_knn_oracle_properties(tiny_scene):
    """Size-independent properties of the reference's layered query (documented quirks)."""
    opt, tables, cam = tiny_scene
    raydir, _ = syn.make_rays(cam, syn.random_pixels(cam, 300, seed=6))
    (pidx, _, loc_w, _, ray_mask, _, _), ex = qr.query_points(opt, tables["xyz"][None], raydir, cam["campos"], cam["camrotc2w"],
                                                              cam["near"], cam["far"])
    xyz = tables["xyz"].numpy()
    p = pidx[0].numpy()
    valid = p >= 0
    d = np.linalg.norm(xyz[np.clip(p, 0, None)] - loc_w[0].numpy()[..., None, :], axis=-1)
    assert (d[valid] <= float(ex["hp"]["radius_limit"]) * (1 + 1e-6)).all()          # radius limit respected
    # filled slots are a prefix (first K accepted fill slots in arrival order)
    assert (np.diff(valid.astype(np.int8), axis=-1) <= 0).all()
    # no duplicate neighbours within a sample
    srt = np.sort(np.where(valid, p, -np.arange(1, p.shape[-1] + 1)), axis=-1)
    assert (np.diff(srt, axis=-1) != 0).all()
    # the slot-0 voxel's points are never returned (fill_occ2pnts `voxel_idx > 0`)
    g = ex["grid"]
    c = syn.voxel_coords(xyz, g["shift"], ex["hp"]["scaled_vsize"])
    dims = ex["hp"]["scaled_vdim"]
    inside = np.all((c >= 0) & (c < np.asarray(dims)[None, :]), axis=1)
    lin = np.where(inside, (c[:, 0] * dims[1] + c[:, 1]) * dims[2] + c[:, 2], 0)
    slot0_pts = np.nonzero(inside & (g["coor_2_occ"][lin] == 0))[0]
    assert len(slot0_pts) >= 1 and 0 in slot0_pts
    assert not np.isin(p[valid], slot0_pts).any()
    assert ray_mask.sum() == pidx.shape[1]


def _load(name):
    path = os.path.join(GOLDEN, name)
    if not os.path.isfile(path):
        pytest.skip("golden fixture %s not generated" % name)
    return np.load(path)


@pytest.mark.parametrize("name,kw", [("aggregator_ref.npz", {}), ("aggregator_order1_ref.npz", dict(agg_intrp_order=1, point_features_dim=63))])
def test_golden_aggregator(name, kw):
    """AggregatorPort reproduces outputs + gradients recorded from the UNMODIFIED reference module."""
    z = _load(name)
    opt = make_opt(**kw)
    port = tp.AggregatorPort(opt)
    port.load_state_dict({k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("sd/")})
    ins = {k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("in/")}
    order = ["sampled_color", "sampled_Rw2c", "sampled_dir", "sampled_conf", "sampled_embedding", "sampled_xyz_pers", "sampled_xyz",
             "sample_pnt_mask", "sample_loc", "sample_loc_w", "sample_ray_dirs"]
    gk = ["sampled_color", "sampled_dir", "sampled_conf", "sampled_embedding"]
    for k in gk:
        ins[k].requires_grad_(True)
    o = port(*[ins[k] for k in order], (0.004,) * 3, 0)
    wt = torch.from_numpy(z["loss_weight"])
    loss = (o[0] * wt).sum() + (o[3] ** 2).sum()
    grads = torch.autograd.grad(loss, [ins[k] for k in gk] + list(port.parameters()))
    assert rel_err(o[0], z["out/decoded"]) < 1e-6
    assert np.array_equal(o[1].numpy(), z["out/ray_valid"])
    assert rel_err(o[2], z["out/weight"]) < 1e-6
    names = gk + [n for n, _ in port.named_parameters()]
    for n, gr in zip(names, grads):
        assert rel_err(gr, z["grad/" + n]) < 1e-5, n


def test_golden_ray_march():
    z = _load("ray_march_ref.npz")
    feats = torch.from_numpy(z["in/feats"]).requires_grad_(True)
    bg = torch.from_numpy(z["in/bg"]).requires_grad_(True)
    o = tp.ray_march(torch.from_numpy(z["in/dist"]), torch.from_numpy(z["in/valid"]), feats, bg)
    for i, n in enumerate(("ray_color", "point_color", "opacity", "acc_T", "blend_w", "bg_T", "bg_w")):
        assert rel_err(o[i], z["out/" + n]) < 1e-6, n
    loss = (o[0] * torch.from_numpy(z["w_color"])).sum() + (o[2] * torch.from_numpy(z["w_op"])).sum()
    gf, gb = torch.autograd.grad(loss, [feats, bg])
    assert rel_err(gf, z["grad/feats"]) < 1e-5 and rel_err(gb, z["grad/bg"]) < 1e-5


def test_golden_querier(tiny_scene):
    """The C port reproduces results recorded from the reference's own kernels (oracle/_ref build) on the tiny scene."""
    z = _load("querier_ref.npz")
    opt, tables, cam = tiny_scene
    assert np.array_equal(tables["xyz"].numpy(), z["xyz"]), "synthetic scene generator changed: regenerate the golden files"
    raydir = torch.from_numpy(z["raydir"])
    (pidx, loc_pers, loc_w, dirs, ray_mask, _, ranges), ex = qr.query_points(
        opt, tables["xyz"][None], raydir, cam["campos"], cam["camrotc2w"], cam["near"], cam["far"], backend="port")
    assert np.array_equal(pidx.numpy(), z["sample_pidx"])
    assert np.array_equal(loc_w.numpy(), z["sample_loc_w"])
    assert np.array_equal(ray_mask.numpy(), z["ray_mask"])
    assert np.array_equal(ranges, z["ranges"])
    assert np.array_equal(ex["grid"]["occ_numpnts"][:int(ex["grid"]["occ_idx"][0])], z["occ_numpnts"])


def test_probe_restatement_against_per_ray_loop():
    """oracle/maintenance_port.probe_outputs (vectorised gathers of neural_points_volumetric_model_ori.py:351-372) against
    a plain per-ray loop of the same definition."""
    from oracle import maintenance_port as mp
    g = torch.Generator().manual_seed(4)
    R, SR, K, F_ = 7, 5, 4, 6
    op = torch.rand(1, R, SR, generator=g)
    op[0, 2] = 0.0                                   # all-equal opacities: first index wins
    loc = torch.randn(1, R, SR, 3, generator=g)
    w, cc = torch.rand(1, R, SR, K, generator=g), torch.rand(1, R, SR, K, generator=g)
    sx, sc, sd = (torch.randn(1, R, SR, K, 3, generator=g) for _ in range(3))
    sf, se = torch.rand(1, R, SR, K, 1, generator=g), torch.randn(1, R, SR, K, F_, generator=g)
    out = mp.probe_outputs(op, loc, w, cc, sx, sc, sd, sf, se)
    for r in range(R):
        s = int(torch.argmax(op[0, r]))
        assert int(out["ray_max_slot"][0, r]) == s and (s == 0 if r == 2 else True)
        assert torch.equal(out["ray_max_sample_loc_w"][0, r], loc[0, r, s])
        ww = (w[0, r, s] * cc[0, r, s])[:, None]
        assert torch.allclose(out["shading_avg_embedding"][0, r], (se[0, r, s] * ww).sum(0))
        assert torch.allclose(out["shading_avg_conf"][0, r], (sf[0, r, s] * ww).sum(0))
        assert torch.allclose(out["ray_max_far_dist"][0, r, 0], (sx[0, r, s] - loc[0, r, s]).norm(dim=-1).min())
    mask = torch.tensor([[1, 0, 1, 1, 0, 1, 1, 1, 0, 1]], dtype=torch.int8)
    full = mp.unmask(mask, {"a": out["shading_avg_color"]})["a"]
    assert full.shape == (1, 10, 3) and bool((full[0, 1] == 0).all()) and torch.equal(full[0, 2], out["shading_avg_color"][0, 1])


# ---- perspective-frustum querier -------------------------------------------------------------------
def _frustum_inputs(name):
    from oracle import frustum_ref as fr
    opt, xyz, cam, pix = syn.frustum_case(name)
    xyz_pers = tp.w2pers_points(xyz, cam["camrotc2w"], cam["campos"])
    hp = fr.get_hyperparameters(opt, cam["h"], cam["w"], cam["intrinsic"], cam["near"], cam["far"])
    return opt, cam, pix, xyz_pers, hp


@pytest.mark.parametrize("name", ["dtu", "crop"])
def test_golden_frustum_querier(name):
    """The reference's own four kernels (host-compiled, canonical schedule) still give the committed vectors."""
    from oracle import frustum_ref as fr
    if not fr.available():
        pytest.skip("oracle/_ref/libpn_ref_querier_frustum.so not built")
    g = np.load(os.path.join(GOLDEN, "frustum_%s_ref.npz" % name))
    opt, cam, pix, xyz_pers, hp = _frustum_inputs(name)
    assert np.array_equal(xyz_pers.numpy(), g["xyz_pers"]) and np.array_equal(pix.numpy(), g["pixel_idx"])
    out = fr.query_points(opt, pix, xyz_pers, cam["h"], cam["w"], cam["intrinsic"], cam["near"], cam["far"], cam["campos"],
                          cam["camrotc2w"])
    assert np.array_equal(out[0].numpy(), g["sample_pidx"]) and np.array_equal(out[1].numpy(), g["sample_loc"])
    assert np.array_equal(out[4].numpy(), g["ray_mask"]) and np.array_equal(out[2].numpy(), g["sample_loc_w"])
    assert np.array_equal(out[6], g["ranges"]) and np.array_equal(out[5], g["vsize"])


@pytest.mark.parametrize("name", ["dtu", "crop"])
def test_frustum_csr_formulation_equals_reference_tables(name):
    """The product's data structure (one CSR over claimed cells, (int)-truncated membership, bitset column walk - restated in
    numpy by oracle/frustum_model.py) reproduces the reference's dense tables / per-column buckets: neighbour indices,
    slot arrangement, sample centres and ray mask bit for bit against the golden vectors."""
    from oracle import frustum_model as fm
    g = np.load(os.path.join(GOLDEN, "frustum_%s_ref.npz" % name))
    opt, cam, pix, xyz_pers, hp = _frustum_inputs(name)
    xp, px = xyz_pers.reshape(-1, 3).numpy(), pix[0].numpy()
    cells, occ, dims = fm.build(xp, hp)
    sz, mask = fm.sample(px, occ, dims, hp["vscale"], opt.SR)
    assert np.array_equal(mask, g["ray_mask"][0])
    sel = mask > 0
    pidx, loc = fm.knn(xp, cells, dims, hp, px[sel], sz[sel], opt)
    assert np.array_equal(pidx, g["sample_pidx"][0]) and np.array_equal(loc, g["sample_loc"][0])
    if name == "crop":
        # the case must expose the truncation rule of insert_vox_points (:394-399): a floor()-only grid gives other neighbours
        t = (xp - hp["ranges"][:3][None]) / hp["scaled_vsize"][None]
        assert int(((t > -1) & (t < 0)).any(1).sum()) > 100
        keep = np.all(np.floor(t) >= 0, axis=1)
        cells_f, _, _ = fm.build(np.where(keep[:, None], xp, np.float32(1e9)), hp)
        assert not np.array_equal(fm.knn(xp, cells_f, dims, hp, px[sel], sz[sel], opt)[0], pidx)
        assert int((sz >= 0).sum(1).max()) == opt.SR


# ---- maintenance steps: the restatements vs the reference's own source, executed as fragments (oracle/ref_fragments.py) ------
def _maint_tables(n=500, F=32, seed=3):
    g = torch.Generator().manual_seed(seed)
    return {"xyz": torch.randn(n, 3, generator=g), "points_embeding": torch.randn(1, n, F, generator=g),
            "points_conf": torch.rand(1, n, 1, generator=g), "points_dir": torch.randn(1, n, 3, generator=g),
            "points_color": torch.randn(1, n, 3, generator=g)}


@needs_ref
def test_prune_and_grow_restatements_equal_reference_methods():
    import types
    from oracle import maintenance_port as mp, ref_fragments as rf
    rel = "models/neural_points/neural_points.py"
    opt = types.SimpleNamespace(xyz_grad=0, feat_grad=1, conf_grad=1, dir_grad=1, color_grad=1)
    t = _maint_tables()
    me = types.SimpleNamespace(opt=opt, eulers=None, Rw2c=torch.nn.Parameter(torch.eye(3)),
                               **{k: torch.nn.Parameter(v.clone()) for k, v in t.items()})
    rf.function(rel, "prune", class_name="NeuralPoints")(me, 0.4)
    want = mp.prune(t, 0.4)
    assert 0 < want["xyz"].shape[0] < t["xyz"].shape[0]
    for k in t:
        assert torch.equal(getattr(me, k).data, want[k]), k
    g = torch.Generator().manual_seed(8)
    add = dict(add_xyz=torch.randn(40, 3, generator=g), add_embedding=torch.randn(40, 32, generator=g), add_color=torch.randn(40, 3, generator=g),
               add_dir=torch.randn(40, 3, generator=g), add_conf=torch.rand(40, 1, generator=g))
    rf.function(rel, "grow_points", class_name="NeuralPoints")(me, **add)
    want = mp.grow_points(want, **add)
    for k in t:
        assert torch.equal(getattr(me, k).data, want[k]), k
    assert me.points_embeding.requires_grad and not me.xyz.requires_grad


@needs_ref
def test_probe_and_unmask_restatements_equal_reference_block():
    """The ``opt.prob == 1`` block of neural_points_volumetric_model_ori.py (:351-372) and ``unmask`` (:131-137), executed from the
    reference's source on CPU tensors, against oracle/maintenance_port.py - bit for bit."""
    import types
    from oracle import maintenance_port as mp, ref_fragments as rf
    rel = "models/neural_points_volumetric_model_ori.py"
    g = torch.Generator().manual_seed(6)
    R, SR, K, F = 9, 7, 8, 32
    ins = dict(sample_pnt_mask=torch.rand(1, R, SR, K, generator=g) > 0.3, weight=torch.rand(1, R, SR, K, generator=g),
               conf_coefficient=torch.rand(1, R, SR, K, generator=g), sample_loc_w=torch.randn(1, R, SR, 3, generator=g),
               sampled_xyz=torch.randn(1, R, SR, K, 3, generator=g), sampled_color=torch.randn(1, R, SR, K, 3, generator=g),
               sampled_dir=torch.randn(1, R, SR, K, 3, generator=g), sampled_conf=torch.rand(1, R, SR, K, 1, generator=g),
               sampled_embedding=torch.randn(1, R, SR, K, F, generator=g))
    opac = torch.rand(1, R, SR, generator=g)
    opac[0, 2, 3] = opac[0, 2, 5] = 2.0                       # a tie: torch.max keeps the first index
    src, line = rf.if_block(rel, 'output["coarse_point_opacity"].shape[1] > 0')
    assert 300 < line < 400
    ns = dict(ins, output={"coarse_point_opacity": opac}, self=types.SimpleNamespace(opt=types.SimpleNamespace(prob=1)), torch=torch)
    exec(compile(src, rel + ":probe", "exec"), ns)
    got = ns["output"]
    want = mp.probe_outputs(opac, ins["sample_loc_w"], ins["weight"], ins["conf_coefficient"], ins["sampled_xyz"], ins["sampled_color"],
                            ins["sampled_dir"], ins["sampled_conf"], ins["sampled_embedding"])
    for k in ("ray_max_shading_opacity", "ray_max_sample_loc_w", "ray_max_far_dist", "shading_avg_color", "shading_avg_dir",
              "shading_avg_conf", "shading_avg_embedding"):
        assert got[k].shape == want[k].shape and torch.equal(got[k], want[k]), k
    assert int(want["ray_max_slot"][0, 2]) == 3
    # unmask: scatter the R'' rows back over all rays
    mask = torch.zeros(1, 20, dtype=torch.int8)
    mask[0, torch.randperm(20, generator=g)[:R].sort()[0]] = 1
    names = ["ray_max_sample_loc_w", "ray_max_shading_opacity", "shading_avg_color", "shading_avg_embedding", "ray_max_far_dist"]
    out_ref = rf.function(rel, "unmask", class_name="NeuralPointsVolumetricModel")(None, torch.nonzero(mask > 0), {k: got[k] for k in names}, names, 1, 20)
    out_port = mp.unmask(mask, {k: want[k] for k in names})
    for k in names:
        assert torch.equal(out_ref[k], out_port[k]), k


@needs_ref
def test_voxel_downsampling_restatement_equals_reference_call_site():
    """construct_vox_points_closest (models/mvs/mvs_utils.py:537-561) executed from the reference's source with torch_scatter's
    two calls restated (oracle/ref_fragments.py: version and algorithm named there)."""
    from oracle import maintenance_port as mp, ref_fragments as rf
    fn = rf.function("models/mvs/mvs_utils.py", "construct_vox_points_closest",
                     extra_globals={"scatter_mean": rf.scatter_mean, "scatter_min": rf.scatter_min})
    g = torch.Generator().manual_seed(12)
    xyz = torch.rand(3000, 3, generator=g) * torch.tensor([2.0, 1.0, 0.5]) - 0.3
    xyz[100:140] = xyz[60:100]                               # exact duplicates: equal residuals, the first index wins
    for res in (16, 40):
        c_ref, grid_ref, min_ref = fn(xyz.clone(), res)
        c, grid, mn, inv, _ = mp.construct_vox_points_closest(xyz.clone(), res)
        assert torch.equal(grid_ref, grid) and torch.equal(min_ref, mn)
        assert torch.equal(c_ref, c)
    # inside a box that filters nothing the reference's box branch gives the same voxels as the port's
    lo, hi = xyz.min(0)[0] - 0.01, xyz.max(0)[0] + 0.01
    c_ref, grid_ref, min_ref = fn(xyz.clone(), 16, space_min=lo, space_max=hi)
    c, grid, mn, _, _ = mp.construct_vox_points_closest(xyz.clone(), 16, space_min=lo, space_max=hi)
    assert torch.equal(grid_ref, grid) and torch.equal(min_ref, mn) and torch.equal(c_ref, c)


@needs_ref
def test_querier_host_glue_equals_reference_methods(tiny_scene):
    """get_hyperparameters / w2pers of the world-coordinate querier (worldcoords.py:48-81, 102-109) and get_hyperparameters /
    pers2w / uniform of the perspective-frustum one (query_point_indices.py:52-81, 110-123, 131-135), executed from the
    reference's source (oracle/ref_fragments.py; np_to_gpuarray - the pycuda upload - replaced by a pass-through), against
    the oracle's restatements and the product's host arithmetic."""
    import types
    from oracle import frustum_ref as fr, ref_fragments as rf
    from pointnerf_b200.query_frustum import frustum_hyperparameters
    glob = {"np": np, "np_to_gpuarray": lambda *a: list(a)}
    # ---- world-coordinate querier
    rel = "models/neural_points/query_point_indices_worldcoords.py"
    ghp = rf.function(rel, "get_hyperparameters", class_name="lighting_fast_querier", extra_globals=glob)
    cases = [(tiny_scene[0], tiny_scene[1]["xyz"])]
    for name in ("cfg2", "cfg3", "tiny4"):
        o = syn.config_opt(name)
        g = torch.Generator().manual_seed(len(name))
        lo, hi = torch.tensor(o.ranges[:3]), torch.tensor(o.ranges[3:])
        cases.append((o, lo + (hi - lo) * (torch.rand(200, 3, generator=g) * 1.1 - 0.05)))
    for o, xyz in cases:
        got = ghp(types.SimpleNamespace(opt=o), o.vsize, xyz[None], ranges=o.ranges)
        want = qr.get_hyperparameters(o, xyz[None])
        assert np.array_equal(got[0], want["radius_limit"]) and np.array_equal(got[2], want["ranges"])
        assert np.array_equal(got[5], want["scaled_vsize"]) and np.array_equal(got[6], want["scaled_vdim"])
        prod = syn.grid_hyperparameters(o, xyz.min(0)[0].numpy(), xyz.max(0)[0].numpy())
        assert np.array_equal(got[2], prod["ranges"]) and np.array_equal(got[6], prod["scaled_vdim"]) and np.array_equal(got[5], prod["scaled_vsize"])
    cam = tiny_scene[2]
    w2p = rf.function(rel, "w2pers", class_name="lighting_fast_querier", extra_globals=glob)
    pts = tiny_scene[1]["xyz"][:777]
    assert torch.equal(w2p(None, pts[None], cam["camrotc2w"], cam["campos"]), tp.w2pers_samples(pts[None], cam["camrotc2w"], cam["campos"]))
    # ---- perspective-frustum querier
    rel = "models/neural_points/query_point_indices.py"
    ghp = rf.function(rel, "get_hyperparameters", class_name="lighting_fast_querier", extra_globals=glob)
    for name in syn.FRUSTUM_CASES:
        o, _, cam, _ = syn.frustum_case(name)
        for conv in (lambda v: v, lambda v: np.asarray(v, np.float32)):
            near, far = conv(cam["near"]), conv(cam["far"])
            got = ghp(types.SimpleNamespace(opt=o, inverse=0), cam["h"], cam["w"], cam["intrinsic"], near, far)
            for want in (fr.get_hyperparameters(o, cam["h"], cam["w"], cam["intrinsic"], near, far),
                         frustum_hyperparameters(o, cam["h"], cam["w"], cam["intrinsic"], near, far)):
                for i, k in ((0, "radius_limit"), (1, "depth_limit"), (2, "ranges"), (3, "vsize"), (4, "vdim"), (5, "scaled_vsize"),
                             (6, "scaled_vdim"), (7, "vscale")):
                    assert np.array_equal(np.asarray(got[i]), np.asarray(want[k])) and np.asarray(got[i]).dtype == np.asarray(want[k]).dtype, (name, k)
    p2w = rf.function(rel, "pers2w", class_name="lighting_fast_querier", extra_globals=glob)
    g = torch.Generator().manual_seed(4)
    loc = torch.rand(1, 50, 6, 3, generator=g) + torch.tensor([-0.5, -0.5, 0.3])
    a, b = p2w(None, loc, cam["camrotc2w"], cam["campos"]), fr.pers2w(loc, cam["camrotc2w"], cam["campos"])
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    uni = rf.function(rel, "uniform", class_name="lighting_fast_querier", extra_globals=glob)
    vsize = np.array([0.01, 0.01, 0.006], np.float32)
    torch.manual_seed(77)
    u = torch.rand(1, 50, 6)
    torch.manual_seed(77)
    got = uni(None, loc.clone(), vsize)
    want = loc.clone()
    want[..., 2] = want[..., 2] + (u - 0.5) * vsize[2]                     # the line of oracle/frustum_ref.py:query_points
    assert torch.equal(got, want)


# ---- the reference's querier classes executed WHOLE on the CPU (their own Python glue + their own kernels) -----------------
@needs_ref
@pytest.mark.parametrize("train", [0, 1])
def test_reference_world_querier_class_equals_oracle_and_golden(tiny_scene, train):
    """lighting_fast_querier of query_point_indices_worldcoords.py: every method compiled from the reference's source, kernels =
    its CUDA string built for the host (oracle/ref_fragments.py).  Its 7-tuple equals the oracle's restated glue around the C
    port bit for bit, and the committed golden vector - so the restatement that travels to the GPU box is the reference."""
    from oracle import ref_fragments as rf
    opt, tables, cam = tiny_scene
    opt = copy.copy(opt)
    opt.is_train = train
    z = _load("querier_ref.npz")
    raydir, xyz = torch.from_numpy(z["raydir"]), tables["xyz"]
    R = raydir.shape[1]
    q = rf.reference_querier("world", opt)
    torch.manual_seed(31)
    uniform = torch.rand(1, R, opt.z_depth_dim) if train else None       # the draw of diff_ray_marching.py:373
    torch.manual_seed(31)
    got = q.query_points(torch.zeros(1, R, 2, dtype=torch.int32), None, xyz[None], torch.tensor([xyz.shape[0]], dtype=torch.int32),
                         cam["h"], cam["w"], cam["intrinsic"], cam["near"], cam["far"], raydir, cam["campos"], cam["camrotc2w"])
    want, _ = qr.query_points(opt, xyz[None], raydir, cam["campos"], cam["camrotc2w"], cam["near"], cam["far"], backend="port",
                              uniform=uniform)
    for i in range(5):
        assert got[i].shape == want[i].shape and got[i].dtype == want[i].dtype and torch.equal(got[i], want[i]), i
    assert np.array_equal(got[6], want[6]) and tuple(got[5]) == tuple(want[5])
    if not train:
        assert np.array_equal(got[0].numpy(), z["sample_pidx"]) and np.array_equal(got[2].numpy(), z["sample_loc_w"])
        assert np.array_equal(got[4].numpy(), z["ray_mask"])
    else:
        assert not np.array_equal(got[2].numpy(), z["sample_loc_w"][:, :got[2].shape[1]]) or got[2].shape[1] != z["sample_loc_w"].shape[1]


@needs_ref
@pytest.mark.parametrize("name", ["dtu", "crop"])
def test_reference_frustum_querier_class_equals_oracle_and_golden(name):
    """The same for the perspective-frustum lighting_fast_querier (query_point_indices.py), without and with its ``uniform``
    sample jitter: the golden vectors the GPU tests compare against ARE what the reference's class returns."""
    from oracle import ref_fragments as rf
    g = _load("frustum_%s_ref.npz" % name)
    opt, cam, pix, xyz_pers, hp = _frustum_inputs(name)
    n_act = torch.tensor([xyz_pers.shape[1]], dtype=torch.int32)
    args = (cam["h"], cam["w"], cam["intrinsic"], cam["near"], cam["far"], None, cam["campos"], cam["camrotc2w"])
    opt.is_train = 0
    out = rf.reference_querier("frustum", opt).query_points(pix, xyz_pers, None, n_act, *args)
    for i, k in enumerate(("sample_pidx", "sample_loc", "sample_loc_w", "ray_dirs", "ray_mask", "vsize", "ranges")):
        assert np.array_equal(np.asarray(out[i]), g[k]) and np.asarray(out[i]).dtype == g[k].dtype, k
    opt.is_train, opt.shpnt_jitter = 1, "uniform"
    torch.manual_seed(5)                                                  # oracle/make_golden.py drew ``uniform`` from this seed
    out = rf.reference_querier("frustum", opt).query_points(pix, xyz_pers, None, n_act, *args)
    assert np.array_equal(out[1].numpy(), g["jitter_sample_loc"]) and np.array_equal(out[2].numpy(), g["jitter_sample_loc_w"])


@needs_ref
@pytest.mark.parametrize("kind,kw", [("world", {}), ("frustum", {}), ("world", dict(agg_intrp_order=1)),
                                     ("frustum", dict(point_features_dim=63))])   # last: the DTU feed-forward trainer's shape
def test_reference_hot_path_whole_equals_pipeline_oracle(kind, kw):
    """The reference's per-ray driver assembled from ITS OWN code (NeuralPointsRayMarching.forward + fill_invalid over
    NeuralPoints.forward over lighting_fast_querier over its kernels, with the unmodified PointAggregator / ray_march;
    oracle/ref_fragments.reference_hot_path) against oracle/pipeline_ref.run - the oracle every GPU step test compares with:
    outputs bit for bit, autograd gradients to 1e-6."""
    from oracle import pipeline_ref, ref_fragments as rf
    R = 150
    if kind == "world":
        opt, tables = syn.make_scene("tiny")
        cam = syn.make_camera("tiny")
        raydir, pix = syn.make_rays(cam, syn.random_pixels(cam, R, seed=5))
        cam_o = cam
    else:
        opt, _, cam, pix = syn.frustum_case("dtu")
        _, tables = syn.make_scene("tiny")
        pix, raydir = pix[:, :R], torch.zeros(1, R, 3)
        # the reference hands near / far to the frustum querier as float32 0-d arrays (neural_points.py:786)
        cam_o = dict(cam, near=np.asarray(cam["near"], np.float32), far=np.asarray(cam["far"], np.float32))
    opt = copy.copy(opt)
    opt.is_train = 0
    for k, v in kw.items():
        setattr(opt, k, v)
    if opt.point_features_dim != tables["points_embeding"].shape[-1]:
        tables = dict(tables, points_embeding=torch.rand(1, tables["xyz"].shape[0], opt.point_features_dim,
                                                         generator=torch.Generator().manual_seed(9)) - 0.5)
    g = torch.Generator().manual_seed(1)
    agg = tp.AggregatorPort(opt)
    with torch.no_grad():
        agg.alpha_branch[0].bias.fill_(60.0)
        for n, p in agg.named_parameters():
            if p.dim() == 1 and not n.startswith("alpha"):
                p.normal_(0, 0.05, generator=g)
    bg = torch.rand(opt.shading_color_channel_num, generator=g)
    wt = torch.randn(1, R, opt.shading_color_channel_num, generator=g)
    drv, leaves = rf.reference_hot_path(kind, opt, tables, agg.state_dict())
    f32 = lambda v: torch.tensor([v], dtype=torch.float32)
    bg_r = bg.clone().requires_grad_(True)
    out = drv.forward(cam["campos"], raydir, bg_color=bg_r, camrotc2w=cam["camrotc2w"], pixel_idx=pix, near=f32(cam["near"]),
                      far=f32(cam["far"]), h=torch.tensor([cam["h"]]), w=torch.tensor([cam["w"]]),
                      intrinsic=torch.from_numpy(cam["intrinsic"])[None])
    loss_fn = lambda o: (o["coarse_raycolor"] * wt).sum() + (o["coarse_point_opacity"] ** 2).sum() * 0.1
    ref = pipeline_ref.run(opt, tables, cam_o, raydir, bg, agg.state_dict(), loss_fn=loss_fn, frustum_pix=pix if kind == "frustum" else None)
    assert 0 < int(ref["ray_mask"].sum()) < R and float(ref["coarse_is_background"].min()) < 0.5
    for k in ("coarse_raycolor", "coarse_point_opacity", "coarse_is_background", "coarse_mask", "queried_shading", "ray_mask"):
        assert out[k].shape == ref[k].shape and torch.equal(out[k].detach(), ref[k]), k
    params = dict(drv.aggregator.named_parameters())
    names = list(leaves) + ["bg_color"] + list(params)
    grads = torch.autograd.grad(loss_fn(out), list(leaves.values()) + [bg_r] + list(params.values()), allow_unused=True)
    for n, gr in zip(names, grads):
        assert gr is not None and rel_err(gr, ref["grads"][n]) < 1e-6, n


@pytest.mark.parametrize("kind", ["world", "frustum"])
def test_golden_reference_driver(kind):
    """tests/golden/driver_*_ref.npz hold what the reference's WHOLE per-ray driver returned in the build container
    (oracle/make_golden.py: driver); the pipeline oracle replays them bit for bit wherever the suite runs."""
    from oracle import frustum_ref, make_golden, pipeline_ref
    if kind == "frustum" and not frustum_ref.available():
        pytest.skip("oracle/_ref/libpn_ref_querier_frustum.so not built")
    z, wts = _load("driver_%s_ref.npz" % kind), _load("driver_weights.npz")
    opt, tables, cam, raydir, pix, bg = make_golden.driver_case(kind)
    assert np.array_equal(raydir.numpy(), z["raydir"]) and np.array_equal(pix.numpy(), z["pixel_idx"]) and np.array_equal(bg.numpy(), z["bg"])
    if kind == "frustum":
        cam = dict(cam, near=np.asarray(cam["near"], np.float32), far=np.asarray(cam["far"], np.float32))
    sd = {k: torch.from_numpy(wts[k]) for k in wts.files}
    out = pipeline_ref.run(opt, tables, cam, raydir, bg, sd, need_grads=False, frustum_pix=pix if kind == "frustum" else None)
    for k in ("coarse_raycolor", "coarse_point_opacity", "coarse_is_background", "coarse_mask", "queried_shading", "ray_mask"):
        assert np.array_equal(out[k].numpy(), z[k]), k


@needs_ref
def test_frustum_hyperparameters_with_the_data_loader_types():
    """h / w as int64 and near / far as float32 0-d arrays (what NeuralPoints.forward makes of the data loader's tensors,
    neural_points.py:786): numpy then evaluates the image-plane extents in float64 - the product follows."""
    import types
    from oracle import ref_fragments as rf
    from pointnerf_b200.neural_points import _as_reference_scalar as conv
    from pointnerf_b200.query_frustum import frustum_hyperparameters
    ghp = rf.function("models/neural_points/query_point_indices.py", "get_hyperparameters", class_name="lighting_fast_querier",
                      extra_globals={"np": np, "np_to_gpuarray": lambda *a: list(a)})
    for name in syn.FRUSTUM_CASES:
        o, _, cam, _ = syn.frustum_case(name)
        h, w = conv(torch.tensor([cam["h"]]), "max"), conv(torch.tensor([cam["w"]]), "max")
        near, far = conv(torch.tensor([cam["near"]]), "min"), conv(torch.tensor([cam["far"]]), "max")
        got = ghp(types.SimpleNamespace(opt=o, inverse=0), h, w, cam["intrinsic"], near, far)
        want = frustum_hyperparameters(o, h, w, cam["intrinsic"], near, far)
        for i, k in ((0, "radius_limit"), (1, "depth_limit"), (2, "ranges"), (3, "vsize"), (4, "vdim"), (5, "scaled_vsize"), (6, "scaled_vdim")):
            assert np.array_equal(np.asarray(got[i]), np.asarray(want[k])) and np.asarray(got[i]).dtype == np.asarray(want[k]).dtype, (name, k)


def test_frozen_activation_pattern_is_the_identity_on_the_oracles_own_pattern():
    """torch_port.LeakyReLUProbe: freezing the aggregator oracle's LeakyReLUs to ITS OWN activation pattern changes neither
    the forward nor any gradient bit (the conditioned backward is the ordinary backward wherever the patterns agree), and
    flipping one unit that is far from its kink is reported by pattern_disagreement with a large margin."""
    opt = make_opt()
    torch.manual_seed(1)
    port = tp.AggregatorPort(opt)
    ins = _agg_inputs(F=opt.point_features_dim)
    gk = ["sampled_color", "sampled_dir", "sampled_conf", "sampled_embedding"]
    for k in gk:
        ins[k].requires_grad_(True)
    wt = None

    def run():
        nonlocal wt
        o = port(*[ins[k] for k in ins], (0.004,) * 3, 0)
        if wt is None:
            wt = torch.randn(o[0].shape, generator=torch.Generator().manual_seed(3))
        return o, torch.autograd.grad((o[0] * wt).sum(), [ins[k] for k in gk] + list(port.parameters()))
    port.freeze_pattern(None, None, record=True)
    o1, g1 = run()
    pre = [a.last_pre.clone() for a in port.activations()]
    pair = torch.stack([p > 0 for p in pre[:4]]).to(torch.uint8)
    samp = torch.stack([p > 0 for p in pre[4:]]).to(torch.uint8)
    port.freeze_pattern(pair, samp)
    o2, g2 = run()
    assert torch.equal(o1[0], o2[0])
    for a, b in zip(g1, g2):
        assert torch.equal(a, b)
    assert port.pattern_disagreement()[0] == 0
    # one flipped unit far from zero: detected, with its margin
    j = int(pre[1].abs().argmax())
    pair[1].view(-1)[j] ^= 1
    port.freeze_pattern(pair, samp)
    run()
    n_diff, n_all, worst = port.pattern_disagreement()
    assert n_diff >= 1 and worst > 1.0 and n_all == sum(p.numel() for p in pre)   # (the flip also moves units downstream)
