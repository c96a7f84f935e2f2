"""PointAggregator on the GPU vs the torch oracle (AggregatorPort == the reference module, see test_oracle.py)."""
import copy

import pytest
import torch

from oracle import kinks, torch_port as tp
from pointnerf_b200.aggregator import PointAggregator
from pointnerf_b200.options import make_opt
from _util import assert_close, group_linear_grads

pytestmark = pytest.mark.gpu

GRAD_INPUTS = ("sampled_color", "sampled_dir", "sampled_conf", "sampled_embedding")


def make_inputs(R, SR, K, F, seed, p_mask=0.6):
    g = torch.Generator().manual_seed(seed)

    def rn(*s):
        return torch.randn(*s, generator=g)
    B = 1
    mask = torch.rand(B, R, SR, K, generator=g) < p_mask
    if R > 1:
        mask[0, 1] = False          # a ray without neighbours
    if SR > 3:
        mask[0, 0, 3] = False       # a sample without neighbours
    mask[0, 0, 0, :] = True
    ins = dict(sampled_color=rn(B, R, SR, K, 3), sampled_Rw2c=torch.linalg.qr(rn(3, 3))[0].contiguous(),
               sampled_dir=torch.nn.functional.normalize(rn(B, R, SR, K, 3), dim=-1),
               sampled_conf=torch.rand(B, R, SR, K, 1, generator=g) * 1.2 - 0.1,     # some outside the clamp range
               sampled_embedding=rn(B, R, SR, K, F) * 0.3,
               sampled_xyz_pers=rn(B, R, SR, K, 3) * 0.05 + torch.tensor([0, 0, 4.]),
               sampled_xyz=rn(B, R, SR, K, 3) * 0.01, sample_pnt_mask=mask,
               sample_loc=rn(B, R, SR, 3) * 0.05 + torch.tensor([0, 0, 4.]), sample_loc_w=rn(B, R, SR, 3) * 0.01,
               sample_ray_dirs=torch.nn.functional.normalize(rn(B, R, SR, 3), dim=-1))
    return ins


def run(mod, ins, dev, wt_dec, wt_conf, vsize=(0.004, 0.004, 0.004), dtype=torch.float32):
    x = {k: ((v.to(dev).to(dtype) if v.is_floating_point() else v.to(dev)) if torch.is_tensor(v) else v) for k, v in ins.items()}
    wt_dec, wt_conf = wt_dec.to(dtype), wt_conf.to(dtype)
    for k in GRAD_INPUTS:
        x[k] = x[k].clone().requires_grad_(True)
    out = mod(x["sampled_color"], x["sampled_Rw2c"], x["sampled_dir"], x["sampled_conf"], x["sampled_embedding"],
              x["sampled_xyz_pers"], x["sampled_xyz"], x["sample_pnt_mask"], x["sample_loc"], x["sample_loc_w"],
              x["sample_ray_dirs"], vsize, 0)
    loss = (out[0] * wt_dec.to(dev)).sum()
    if out[3] is not None:
        loss = loss + (out[3] * wt_conf.to(dev)).sum()
    params = list(mod.parameters())
    grads = torch.autograd.grad(loss, [x[k] for k in GRAD_INPUTS] + params)
    return out, grads


@pytest.mark.parametrize("R,SR,K,seed,kw", [(7, 5, 8, 1, {}), (33, 24, 8, 2, {}), (3, 4, 1, 3, {}), (5, 6, 16, 4, {}),
                                            (33, 24, 8, 5, dict(agg_intrp_order=1)),
                                            (9, 7, 8, 6, dict(agg_intrp_order=1, point_features_dim=63)),     # DTU feed-forward configuration
                                            (9, 7, 8, 7, dict(point_features_dim=63))])
def test_forward_backward_matches_oracle(cuda_dev, R, SR, K, seed, kw):
    opt = make_opt(K=K, SR=SR, **kw)
    torch.manual_seed(seed)
    port = tp.AggregatorPort(opt)
    with torch.no_grad():               # non-zero biases (xavier init zeroes them) and a live density head
        for p in port.parameters():
            if p.dim() == 1:
                p.normal_(0, 0.1)
    mod = PointAggregator(opt).to(cuda_dev)
    mod.load_state_dict(port.state_dict())
    assert [n for n, _ in mod.named_parameters()] == [n for n, _ in port.named_parameters()]
    ins = make_inputs(R, SR, K, opt.point_features_dim, seed)
    g = torch.Generator().manual_seed(seed + 100)
    wt_dec = torch.randn(1, R, SR, opt.shading_color_channel_num + 1, generator=g)
    wt_conf = torch.randn(1, R, SR, K, generator=g)
    # forward: against the unconditioned oracle.  backward: the oracle's autograd under the LeakyReLU pattern of OUR forward
    # (pn_agg_activation_signs; the MLP is piecewise linear, oracle/kinks.py) - every sample carries loss weight
    mod.export_signs = True
    o, gr = run(mod, ins, cuda_dev, wt_dec, wt_conf)
    signs = tuple(t.cpu() for t in mod.last_call.signs)
    o_ref, _ = run(port, ins, torch.device("cpu"), wt_dec, wt_conf)
    port64 = copy.deepcopy(port).double()
    o_64, _ = run(port64, ins, torch.device("cpu"), wt_dec, wt_conf, dtype=torch.float64)
    port.freeze_pattern(*signs)
    port64.freeze_pattern(*signs)
    o_cnd, g_ref = run(port, ins, torch.device("cpu"), wt_dec, wt_conf)
    _, g_64 = run(port64, ins, torch.device("cpu"), wt_dec, wt_conf, dtype=torch.float64)
    n_diff, n_all, worst = port64.pattern_disagreement()
    assert n_all > 0 and n_diff <= 1e-4 * n_all + 1 and worst < kinks.PATTERN_TOL, (n_diff, n_all, worst)
    assert_close(o_cnd[0], o_ref[0], "conditioned oracle forward")
    assert o[0].shape == o_ref[0].shape and o[0].dtype == torch.float32
    assert torch.equal(o[1].cpu(), o_ref[1]) and o[1].dtype == torch.bool
    assert_close(o[0], o_ref[0], "decoded_features", want64=o_64[0])
    assert_close(o[0][..., 0], o_ref[0][..., 0], "density", want64=o_64[0][..., 0])
    assert_close(o[2], o_ref[2], "weight")
    assert_close(o[3], o_ref[3], "conf_coefficient")
    names = list(GRAD_INPUTS) + [n for n, _ in port.named_parameters()]
    ga, gb, gb64 = (group_linear_grads(dict(zip(names, x))) for x in (gr, g_ref, g_64))
    for n in ga:
        assert_close(ga[n], gb[n], "grad " + n, want64=gb64[n])


def test_fused_tables_match_gathered(cuda_dev, tiny_scene):
    """forward_tables (indices + point tables, segmented gradient reduction) == gather + reference-API forward."""
    opt, tables, cam = tiny_scene
    d = cuda_dev
    torch.manual_seed(0)
    mod = PointAggregator(opt).to(d)
    N = tables["xyz"].shape[0]
    g = torch.Generator().manual_seed(3)
    R, SR, K = 20, opt.SR, opt.K
    pidx = torch.randint(0, N, (R, SR, K), generator=g, dtype=torch.int32)
    pidx[torch.rand(R, SR, K, generator=g) < 0.5] = -1
    pidx[3] = -1
    pidx[:, :, :] = torch.where(torch.arange(K)[None, None, :] < 6, pidx, pidx)  # keep dtype
    pidx[5, :, 0] = 17   # one point referenced many times: exercises the segmented reduction
    xyz = tables["xyz"]
    loc_w = xyz[torch.randint(0, N, (R, SR), generator=g)] + 0.003 * torch.randn(R, SR, 3, generator=g)
    campos, rot = cam["campos"], cam["camrotc2w"]
    xyz_pers = tp.w2pers_points(xyz, rot, campos)[0]
    loc_pers = tp.w2pers_samples(loc_w[None], rot, campos)[0]
    dirs = torch.nn.functional.normalize(torch.randn(R, 3, generator=g), dim=-1)
    Rw2c = torch.linalg.qr(torch.randn(3, 3, generator=g))[0].contiguous()
    leaves = {k: tables[k].to(d).clone().requires_grad_(True) for k in ("points_embeding", "points_color", "points_dir", "points_conf")}
    wt = torch.randn(R, SR, opt.shading_color_channel_num + 1, generator=g).to(d)

    dec, w = mod.forward_tables(xyz.to(d), xyz_pers.to(d), leaves["points_embeding"][0], leaves["points_color"][0],
                                leaves["points_dir"][0], leaves["points_conf"][0, :, 0], Rw2c.to(d), pidx.to(d), loc_pers.to(d),
                                loc_w.to(d), dirs.to(d), opt.vsize, dirs_per_ray=True, want_weight=True)
    grads = torch.autograd.grad((dec * wt).sum(), list(leaves.values()) + list(mod.parameters()))

    port = tp.AggregatorPort(opt)
    port.load_state_dict({k: v.cpu() for k, v in mod.state_dict().items()})
    leaves_c = {k: tables[k].clone().requires_grad_(True) for k in leaves}
    gt = tp.gather_point_tables(pidx[None].long().to(torch.int64), xyz, xyz_pers[None], leaves_c["points_embeding"],
                                leaves_c["points_color"], leaves_c["points_dir"], leaves_c["points_conf"])
    o_ref = port(gt["sampled_color"], Rw2c, gt["sampled_dir"], gt["sampled_conf"], gt["sampled_embedding"], gt["sampled_xyz_pers"],
                 gt["sampled_xyz"], gt["sample_pnt_mask"], loc_pers[None], loc_w[None],
                 dirs[None, :, None, :].expand(1, R, SR, 3).contiguous(), opt.vsize, 0)
    g_ref = torch.autograd.grad((o_ref[0][0] * wt.cpu()).sum(), list(leaves_c.values()) + list(port.parameters()))
    assert_close(dec, o_ref[0][0], "decoded (fused gather)")
    assert_close(w, o_ref[2][0], "weight (fused gather)")
    names = list(leaves) + [n for n, _ in port.named_parameters()]
    ga, gb = group_linear_grads(dict(zip(names, grads))), group_linear_grads(dict(zip(names, g_ref)))
    for n in ga:
        assert_close(ga[n], gb[n], "grad " + n)
    # deterministic: the segmented reduction has a fixed order, so two runs are bit-identical
    dec2, _ = mod.forward_tables(xyz.to(d), xyz_pers.to(d), leaves["points_embeding"][0], leaves["points_color"][0],
                                 leaves["points_dir"][0], leaves["points_conf"][0, :, 0], Rw2c.to(d), pidx.to(d), loc_pers.to(d),
                                 loc_w.to(d), dirs.to(d), opt.vsize, dirs_per_ray=True, want_weight=True)
    grads2 = torch.autograd.grad((dec2 * wt).sum(), list(leaves.values()) + list(mod.parameters()))
    assert torch.equal(dec, dec2)
    for a, b in zip(grads, grads2):
        assert torch.equal(a, b)
    # inference (no gradient wanted: pn_agg_params.forward_only = 1, the GEMM epilogues skip the sign-bit packing): same bits
    with torch.no_grad():
        dec3, _ = mod.forward_tables(xyz.to(d), xyz_pers.to(d), leaves["points_embeding"][0], leaves["points_color"][0],
                                     leaves["points_dir"][0], leaves["points_conf"][0, :, 0], Rw2c.to(d), pidx.to(d), loc_pers.to(d),
                                     loc_w.to(d), dirs.to(d), opt.vsize, dirs_per_ray=True, want_weight=True)
    assert mod.last_call.params.forward_only == 1
    assert torch.equal(dec, dec3)


def test_no_valid_samples_and_empty(cuda_dev):
    opt = make_opt(K=8, SR=4)
    mod = PointAggregator(opt).to(cuda_dev)
    ins = make_inputs(3, 4, 8, 32, seed=1)
    ins["sample_pnt_mask"][:] = False
    x = {k: (v.to(cuda_dev) if torch.is_tensor(v) else v) for k, v in ins.items()}
    out = mod(*[x[k] for k in ("sampled_color", "sampled_Rw2c", "sampled_dir", "sampled_conf", "sampled_embedding", "sampled_xyz_pers",
                               "sampled_xyz", "sample_pnt_mask", "sample_loc", "sample_loc_w", "sample_ray_dirs")], (0.004,) * 3, 0)
    assert tuple(out[0].shape) == (1, 3, 4, 129) and float(out[0].abs().max()) == 0.0 and not bool(out[1].any())
    ins0 = make_inputs(0, 4, 8, 32, seed=1) if False else {k: (v[:, :0] if torch.is_tensor(v) and v.dim() > 2 else v) for k, v in ins.items()}
    x0 = {k: (v.to(cuda_dev) if torch.is_tensor(v) else v) for k, v in ins0.items()}
    out0 = mod(*[x0[k] for k in ("sampled_color", "sampled_Rw2c", "sampled_dir", "sampled_conf", "sampled_embedding", "sampled_xyz_pers",
                                 "sampled_xyz", "sample_pnt_mask", "sample_loc", "sample_loc_w", "sample_ray_dirs")], (0.004,) * 3, 0)
    assert tuple(out0[0].shape) == (1, 0, 4, 129) and out0[2] is None


def test_state_dict_keys_match_reference_names():
    if not torch.cuda.is_available():
        pytest.skip("module constructor loads the CUDA library only; keys are checked on the GPU box too")
    mod = PointAggregator(make_opt())
    want = {"block1.0.weight": (256, 284), "block1.2.weight": (256, 256), "block3.0.weight": (256, 263), "block3.2.weight": (256, 256),
            "alpha_branch.0.weight": (1, 256), "color_branch.0.weight": (128, 280), "color_branch.2.weight": (128, 128),
            "color_branch.4.weight": (128, 128)}
    sd = mod.state_dict()
    for k, shp in want.items():
        assert tuple(sd[k].shape) == shp, k
        assert k.replace("weight", "bias") in sd
    assert len(sd) == 16


def test_unsupported_configuration_raises():
    with pytest.raises(NotImplementedError):
        PointAggregator(make_opt(agg_intrp_order=0))
    with pytest.raises(NotImplementedError):
        PointAggregator(make_opt(agg_distance_kernel="quadric"))
