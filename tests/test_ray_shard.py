"""Host logic of ray sharding (pointnerf_b200/ray_shard.py) on CPU: world_size-2 gloo.

The model is a CPU stand-in with the per-ray driver's interface (the real one needs the CUDA library); what is tested is
the sharding contract of SURVEY.md 8(e): contiguous ray slices, one flat all-reduce, and gradients equal to the
single-process gradient on the concatenated batch.
"""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

from pointnerf_b200.ray_shard import FlatGrads, RayShardedStep, shard_bounds


class _Points(nn.Module):
    def __init__(self, n=50, f=8):
        super().__init__()
        g = torch.Generator().manual_seed(1)
        self.points_embeding = nn.Parameter(torch.randn(1, n, f, generator=g))
        self.points_conf = nn.Parameter(torch.rand(1, n, 1, generator=g))
        self.points_dir = nn.Parameter(torch.randn(1, n, 3, generator=g))
        self.points_color = nn.Parameter(torch.randn(1, n, 3, generator=g))


class _Driver(nn.Module):
    """coarse_raycolor[r] = MLP(weighted mix of the point rows picked by the ray) + bg * miss(r)."""

    def __init__(self, C=6):
        super().__init__()
        torch.manual_seed(0)
        self.neural_points = _Points()
        self.aggregator = nn.Sequential(nn.Linear(8 + 1 + 3 + 3, 16), nn.LeakyReLU(), nn.Linear(16, C))

    def touched_point_ids(self):
        return self._idx

    def forward(self, campos, raydir, bg_color=None, camrotc2w=None, near=None, far=None, uniform=None):
        nps = self.neural_points
        n = nps.points_embeding.shape[1]
        idx = (raydir[0].abs().sum(-1) * 1000).long() % n
        self._idx = idx
        x = torch.cat([nps.points_embeding[0, idx], nps.points_conf[0, idx], nps.points_dir[0, idx], nps.points_color[0, idx]], -1)
        miss = (idx % 3 == 0).float()[:, None]
        return {"coarse_raycolor": (self.aggregator(x) * (1 - miss) + bg_color[None] * miss)[None]}


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _batch(R=37, C=6):
    g = torch.Generator().manual_seed(3)
    return torch.randn(1, R, 3, generator=g), torch.rand(1, R, C, generator=g)


def _worker(rank, world, port, out, sparse):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    model = _Driver()
    bg = nn.Parameter(torch.full((6,), 0.25))
    step = RayShardedStep(model, bg, sparse=sparse)
    step.BIG = 64          # the stand-in tables are tiny: exercise the in-place path as well
    raydir, gt = _batch()
    lo, hi = shard_bounds(raydir.shape[1], world, rank)
    cam = dict(campos=torch.zeros(1, 3), camrotc2w=torch.eye(3)[None], near=0.1, far=1.0)
    loss = step(cam, raydir[:, lo:hi], gt[:, lo:hi], global_rays=raydir.shape[1])
    tot = loss.clone()
    dist.all_reduce(tot)
    if rank == 0:
        torch.save({"flat": step.flat_grads().clone(), "loss": tot}, out)
    dist.destroy_process_group()


def test_shard_bounds_cover_and_balance():
    for n, w in ((10, 3), (4096, 8), (5, 8), (0, 2)):
        b = [shard_bounds(n, w, r) for r in range(w)]
        assert b[0][0] == 0 and b[-1][1] == n
        assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
        sizes = [h - l for l, h in b]
        assert max(sizes) - min(sizes) <= 1


def test_flat_grads_alias_and_zero():
    lin = nn.Linear(3, 2)
    fg = FlatGrads(lin.parameters())
    lin(torch.ones(4, 3)).sum().backward()
    fg.gather_from_params()
    assert fg.flat.numel() == 8 and float(fg.flat.abs().sum()) > 0
    assert torch.equal(fg.flat[:6].view(2, 3), lin.weight.grad)
    fg.zero()
    assert float(lin.weight.grad.abs().sum()) == 0.0


@pytest.mark.parametrize("sparse", [True, False])
def test_two_rank_gradients_equal_single_process(tmp_path, sparse):
    out = str(tmp_path / "g.pt")
    mp.spawn(_worker, args=(2, _free_port(), out, sparse), nprocs=2, join=True)
    got = torch.load(out)
    model = _Driver()
    bg = nn.Parameter(torch.full((6,), 0.25))
    step = RayShardedStep(model, bg)            # world = 1: no process group
    raydir, gt = _batch()
    cam = dict(campos=torch.zeros(1, 3), camrotc2w=torch.eye(3)[None], near=0.1, far=1.0)
    loss = step(cam, raydir, gt)
    assert torch.allclose(got["loss"], loss, rtol=1e-6, atol=1e-8)
    assert torch.allclose(got["flat"], step.flat_grads(), rtol=1e-5, atol=1e-7)
    assert float(step.flat_grads().abs().max()) > 0


def test_rank_local_step_issues_no_collective(monkeypatch):
    """A step with reduce=False (per-stage timing on one rank, single-process reference gradients) must not touch the
    process group: rank 0 once queued the capacity all-reduce alone there and every multi-GPU bench run hung."""
    import torch.distributed as dist
    from pointnerf_b200.ray_shard import RayShardedStep

    class Toy(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.neural_points = torch.nn.Module()
            for n, w in (("points_embeding", 4), ("points_conf", 1), ("points_dir", 3), ("points_color", 3)):
                setattr(self.neural_points, n, torch.nn.Parameter(torch.ones(1, 10, w)))
            self.aggregator = torch.nn.Linear(3, 2)

        def forward(self, campos, raydir, bg_color=None, **kw):
            s = sum(getattr(self.neural_points, n).sum() for n in ("points_embeding", "points_conf", "points_dir", "points_color"))
            return {"coarse_raycolor": self.aggregator(raydir) * s + bg_color}

    m = Toy()
    bg = torch.nn.Parameter(torch.zeros(2))
    st = RayShardedStep(m, bg, sparse=True)
    st.world = 2                                            # as on a 2-rank job, without a process group behind it

    def boom(*a, **k):
        raise AssertionError("collective issued by a rank-local step")
    for fn in ("all_reduce", "all_gather_into_tensor", "all_gather", "barrier"):
        monkeypatch.setattr(dist, fn, boom)
    monkeypatch.setattr(st, "_begin_cap_exchange", boom)
    cam = dict(campos=torch.zeros(1, 3), camrotc2w=torch.eye(3)[None], near=0.1, far=1.0)
    loss = st(cam, torch.randn(1, 5, 3), torch.zeros(1, 5, 2), global_rays=10, reduce=False)
    assert torch.isfinite(loss) and bg.grad is not None
    with pytest.raises(AssertionError):                     # and the same call WITH the exchange does reach the group
        st(cam, torch.randn(1, 5, 3), torch.zeros(1, 5, 2), global_rays=10, reduce=True)
