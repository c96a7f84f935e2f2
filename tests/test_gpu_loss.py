"""pn_color_mse (fused colour MSE + gradient) against torch.nn.MSELoss, the reference's loss
(base_rendering_model.py:543-550).  Tolerance: 1e-6 relative on the loss (fp32 summation order), gradient bit-exact."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from pointnerf_b200.ray_march import color_mse  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("R,C", [(1, 3), (37, 3), (4096, 3), (8192, 3), (1000, 1), (513, 5)])
def test_color_mse_matches_torch(R, C):
    g = torch.Generator(device="cuda").manual_seed(R * 7 + C)
    pred = torch.rand((1, R, C), device="cuda", generator=g, requires_grad=True)
    gt = torch.rand((1, R, C), device="cuda", generator=g)
    loss = color_mse(pred, gt)
    loss.backward()
    p2 = pred.detach().clone().requires_grad_(True)
    ref = torch.nn.MSELoss()(p2, gt)
    ref.backward()
    assert abs(float(loss) - float(ref)) <= 1e-6 * abs(float(ref))
    ref64 = float(((pred.detach().double() - gt.double()) ** 2).mean())
    assert abs(float(loss) - ref64) <= 1e-6 * ref64
    assert torch.equal(pred.grad, p2.grad) or torch.allclose(pred.grad, p2.grad, rtol=2e-7, atol=0)


def test_color_mse_masked_rays_and_upstream_scale():
    """ray_masked_coarse_raycolor: only rays with ray_mask > 0 enter, divisor = selected elements
    (base_rendering_model.py:520-541); an upstream gradient scales the stored gradient."""
    R, C = 3000, 3
    g = torch.Generator(device="cuda").manual_seed(5)
    pred = torch.rand((1, R, C), device="cuda", generator=g, requires_grad=True)
    gt = torch.rand((1, R, C), device="cuda", generator=g)
    mask = torch.rand((1, R), device="cuda", generator=g) > 0.4
    n_sel = int(mask.sum()) * C
    loss = color_mse(pred, gt, ray_weight=mask.float(), norm=n_sel)
    (loss * 0.25).backward()
    p2 = pred.detach().clone().requires_grad_(True)
    ref = torch.nn.MSELoss()(p2[mask], gt[mask])
    (ref * 0.25).backward()
    assert abs(float(loss) - float(ref)) <= 1e-6 * abs(float(ref))
    assert torch.allclose(pred.grad, p2.grad, rtol=1e-6, atol=0)
    assert float(pred.grad[~mask].abs().max()) == 0.0


def test_color_mse_is_deterministic_and_handles_empty():
    R, C = 8192, 3
    g = torch.Generator(device="cuda").manual_seed(9)
    pred = torch.rand((1, R, C), device="cuda", generator=g)
    gt = torch.rand((1, R, C), device="cuda", generator=g)
    a = float(color_mse(pred, gt))
    assert all(float(color_mse(pred, gt)) == a for _ in range(5))
    e = color_mse(torch.zeros((1, 0, 3), device="cuda"), torch.zeros((1, 0, 3), device="cuda"))
    assert float(e) == 0.0


def test_color_mse_rejects_mismatched_shapes():
    with pytest.raises(ValueError):
        color_mse(torch.zeros((1, 4, 3), device="cuda"), torch.zeros((1, 5, 3), device="cuda"))
    with pytest.raises(Exception):
        color_mse(torch.zeros((1, 4, 3)), torch.zeros((1, 4, 3)))


def test_conf_coefficient_gather_matches_reference_expression():
    """gather_clamp == gradiant_clamp(points_conf[sample_pidx.clamp(min=0)]) (point_aggregators.py:724-726, :809-811;
    neural_points.py index_select) in value (bit-exact) and in the straight-through gradient."""
    from pointnerf_b200.aggregator import PointAggregator, gather_clamp
    N = 5000
    g = torch.Generator(device="cuda").manual_seed(3)
    for shape in [(7, 5, 8), (1, 3, 1), (33, 24, 8), (0, 4, 8)]:
        conf = (torch.rand((1, N, 1), device="cuda", generator=g) * 1.2 - 0.1).requires_grad_(True)   # values outside [1e-4, 1] too
        idx = torch.randint(-1, N, shape, device="cuda", generator=g, dtype=torch.int32)
        idx[torch.rand(shape, device="cuda", generator=g) < 0.5] = -1
        up = torch.rand(shape, device="cuda", generator=g)
        out = gather_clamp(conf, idx)
        (out * up).sum().backward()
        c2 = conf.detach().clone().requires_grad_(True)
        ref = PointAggregator.gradiant_clamp(c2[0, :, 0][idx.clamp(min=0).long()], min=0.0001, max=1)
        (ref * up).sum().backward()
        assert out.shape == ref.shape and torch.equal(out, ref)
        assert torch.allclose(conf.grad, c2.grad, rtol=1e-5, atol=1e-6)
