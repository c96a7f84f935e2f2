"""ray_march / alpha_ray_march / ray_dist on the GPU vs the torch oracle (fp32, 1e-4 relative)."""
import pytest
import torch

from oracle import torch_port as tp
from pointnerf_b200 import ray_march as rm
from _util import assert_close

pytestmark = pytest.mark.gpu


def _inputs(R, SR, C, seed, sigma_scale=30.0, valid_p=0.7):
    g = torch.Generator().manual_seed(seed)
    feats = torch.randn(1, R, SR, C + 1, generator=g)
    feats[..., 0] = torch.rand(1, R, SR, generator=g) * sigma_scale
    dist = torch.rand(1, R, SR, generator=g) * 0.02
    valid = torch.rand(1, R, SR, generator=g) < valid_p
    valid[0, 0] = False            # a ray with no valid sample
    if R > 1:
        valid[0, 1] = True
    dist = dist * valid.float()
    bg = torch.rand(C, generator=g)
    return feats, dist, valid, bg


@pytest.mark.parametrize("R,SR,C,sig", [(37, 24, 128, 30.0), (50, 80, 128, 400.0), (9, 33, 3, 50.0), (5, 160, 48, 2000.0), (1, 1, 128, 10.0)])
def test_ray_march_forward_backward(cuda_dev, R, SR, C, sig):
    feats, dist, valid, bg = _inputs(R, SR, C, seed=R * 1000 + SR, sigma_scale=sig)
    g = torch.Generator().manual_seed(1)
    w_color, w_op, w_T, w_bw, w_bgT = (torch.randn(s, generator=g) for s in ((1, R, C), (1, R, SR), (1, R, SR), (1, R, SR, 1), (1, R, 1)))

    def loss(o):
        return (o[0] * w_color.to(o[0].device)).sum() + (o[2] * w_op.to(o[0].device)).sum() + (o[3] * w_T.to(o[0].device)).sum() + \
            (o[4] * w_bw.to(o[0].device)).sum() + (o[5] * w_bgT.to(o[0].device)).sum() + (o[1] ** 2).sum() * 0.01

    f_ref, bg_ref = feats.clone().requires_grad_(True), bg.clone().requires_grad_(True)
    o_ref = tp.ray_march(dist, valid, f_ref, bg_ref)
    gf_ref, gbg_ref = torch.autograd.grad(loss(o_ref), [f_ref, bg_ref])

    d = cuda_dev
    f_dev, bg_dev = feats.to(d).requires_grad_(True), bg.to(d).requires_grad_(True)
    o = rm.ray_march(dist.to(d), valid.to(d), f_dev, rm.radiance_render, rm.alpha_blend, bg_dev)
    assert len(o) == 7
    gf, gbg = torch.autograd.grad(loss(o), [f_dev, bg_dev])
    names = ("ray_color", "point_color", "opacity", "acc_transmission", "blend_weight", "background_transmission", "background_blend_weight")
    for n, a, b in zip(names, o, o_ref):
        assert a.shape == b.shape, n
        assert_close(a, b, n)
    assert_close(gf, gf_ref, "d/d ray_features")
    assert_close(gbg, gbg_ref, "d/d bg_color")


def test_early_termination_stays_within_tolerance(cuda_dev):
    """Opaque rays: transmittance collapses after a few samples; skipped rows must not change the result."""
    R, SR, C = 64, 80, 128
    feats, dist, valid, bg = _inputs(R, SR, C, seed=7, sigma_scale=5000.0, valid_p=1.0)
    o_ref = tp.ray_march(dist, valid, feats, bg)
    assert float(o_ref[5][0, 1:].max()) < 1e-8      # ray 0 has no valid sample by construction (_inputs)
    d = cuda_dev
    o = rm.ray_march(dist.to(d), valid.to(d), feats.to(d), rm.radiance_render, rm.alpha_blend, bg.to(d))
    assert_close(o[0], o_ref[0], "ray_color (terminated rays)")
    assert_close(o[3], o_ref[3], "acc_transmission")


def test_alpha_ray_march_and_no_bg(cuda_dev):
    feats, dist, valid, bg = _inputs(21, 24, 128, seed=3)
    d = cuda_dev
    o_ref = tp.ray_march(dist, valid, feats, None)
    o = rm.ray_march(dist.to(d), valid.to(d), feats.to(d), rm.radiance_render, rm.alpha_blend, None)
    assert_close(o[0], o_ref[0], "ray_color without bg")
    a = rm.alpha_ray_march(dist.to(d), valid.to(d), feats.to(d), rm.alpha_blend)
    assert len(a) == 5
    for got, want, n in zip(a, (o_ref[2], o_ref[3], o_ref[4], o_ref[5], o_ref[6]), ("opacity", "T", "blend", "bgT", "bgw")):
        assert_close(got, want, n)


def test_unsupported_functions_raise(cuda_dev):
    feats, dist, valid, bg = _inputs(3, 4, 16, seed=1)
    d = cuda_dev
    with pytest.raises(NotImplementedError):
        rm.ray_march(dist.to(d), valid.to(d), feats.to(d), lambda f: f, rm.alpha_blend)
    with pytest.raises(NotImplementedError):
        rm.find_blend_function("alpha2")


def test_zero_rays(cuda_dev):
    d = cuda_dev
    o = rm.ray_march(torch.zeros(1, 0, 24, device=d), torch.zeros(1, 0, 24, dtype=torch.bool, device=d),
                     torch.zeros(1, 0, 24, 129, device=d), rm.radiance_render, rm.alpha_blend, torch.rand(128, device=d))
    assert tuple(o[0].shape) == (1, 0, 128) and tuple(o[4].shape) == (1, 0, 24, 1)


@pytest.mark.parametrize("unit", [0, 1])
def test_ray_dist(cuda_dev, unit):
    g = torch.Generator().manual_seed(5)
    R, SR = 40, 24
    z = torch.cumsum(torch.rand(1, R, SR, generator=g) * 0.006, dim=-1) + 2.0
    z[0, 3, 5:] = 0.0      # unused slots are zeros in the reference
    z[0, 4, 7] = z[0, 4, 6]
    loc = torch.randn(1, R, SR, 3, generator=g)
    loc[..., 2] = z
    valid = torch.rand(1, R, SR, generator=g) < 0.8
    vsize = (0.004, 0.004, 0.004)
    want = tp.ray_dist_from_samples(loc, valid, vsize, unit)
    got = rm.ray_dist_from_samples(loc.to(cuda_dev), valid.to(cuda_dev), vsize, unit)
    assert torch.equal(got.cpu(), want)
