"""Voxel down-sampling: pn_voxel_downsample vs the CPU restatement of construct_vox_points_closest."""
import numpy as np
import pytest
import torch

from oracle import maintenance_port as mp
from pointnerf_b200.voxelize import construct_vox_points_closest

pytestmark = pytest.mark.gpu


def _check(xyz, vox_res, dev, **box):
    want_c, want_g, want_m, inv, res = mp.construct_vox_points_closest(xyz, vox_res, **box)
    got_c, got_g, got_m, got_inv = construct_vox_points_closest(xyz.to(dev), vox_res, return_inverse=True, **{k: v.to(dev) for k, v in box.items()})
    assert got_g.dtype == torch.int32 and got_m.dtype == torch.int64
    assert torch.equal(got_g.cpu(), want_g), "voxel set / order differs"               # integer work: bit-exact
    assert torch.equal(got_inv.cpu().long(), inv)
    assert torch.equal(got_c.cpu(), want_c), "centroids (serial fp32 sums in point order) differ"
    # index work: bit-exact.  Tie-break: the smallest residual (fp32, the host's fma chain of torch.norm), then the smallest
    # point index (torch_scatter's scatter_min keeps the first minimum)
    assert torch.equal(got_m.cpu(), want_m), "nearest-point indices differ"


@pytest.mark.parametrize("n,vox_res", [(2, 4), (37, 3), (6000, 40), (50000, 100), (50000, 900)])
def test_matches_oracle(cuda_dev, n, vox_res):
    g = torch.Generator().manual_seed(n + vox_res)
    xyz = torch.rand(n, 3, generator=g) * torch.tensor([1.0, 0.6, 0.3]) - 0.2
    _check(xyz, vox_res, cuda_dev)


def test_duplicates_and_scene_points(tiny_scene, cuda_dev):
    _, tables, _ = tiny_scene
    xyz = torch.cat([tables["xyz"], tables["xyz"][:500]])          # exact duplicates: residual ties -> smallest index
    _check(xyz, 64, cuda_dev)


def test_bounding_box(cuda_dev):
    g = torch.Generator().manual_seed(8)
    xyz = torch.rand(20000, 3, generator=g) * 2 - 1
    _check(xyz, torch.tensor(32.0), cuda_dev, space_min=torch.tensor([-0.5, -0.6, -0.7]), space_max=torch.tensor([0.5, 0.6, 0.7]))


def test_empty(cuda_dev):
    c, gidx, m = construct_vox_points_closest(torch.zeros(0, 3, device=cuda_dev), 8, space_min=torch.zeros(3), space_max=torch.ones(3))
    assert c.shape == (0, 3) and gidx.shape == (0, 3) and m.shape == (0,)


def test_full_size_properties(cuda_dev):
    """3M points (cfg3 scale): voxel set = torch.unique on the device, voxels sorted, every point's voxel holds it, the
    selected point lies in its voxel, centroid = mean of the voxel's points."""
    g = torch.Generator(device="cuda").manual_seed(3)
    N, res = 3_000_000, 320
    xyz = torch.rand(N, 3, device=cuda_dev, generator=g)
    c, gidx, m, inv = construct_vox_points_closest(xyz, res, return_inverse=True)
    inv = inv.long()
    lo, hi = xyz.min(0)[0], xyz.max(0)[0]
    edge = (hi - lo).max() * 1.05
    smin = (hi + lo) / 2 - edge / 2
    coords = torch.floor((xyz - smin[None]) / (edge / res)).to(torch.int32)
    uniq = torch.unique(coords, dim=0)
    assert torch.equal(gidx, uniq)
    assert torch.equal(gidx[inv], coords)
    assert torch.equal(inv[m], torch.arange(len(m), device=cuda_dev))
    cnt = torch.bincount(inv, minlength=len(m)).float()
    mean = torch.zeros_like(c).index_add_(0, inv, xyz) / cnt[:, None]
    assert float((mean - c).abs().max()) < 1e-5
