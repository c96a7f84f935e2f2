"""Sparse per-point gradient exchange kernels (pn_rows_pack / pn_rows_merge) and the touched-row list of the fused backward."""
import pytest
import torch

from pointnerf_b200 import _lib
from pointnerf_b200.aggregator import PointAggregator
from oracle import torch_port as tp

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("G", [2, 3, 8])
def test_pack_merge_equals_rank_ordered_sum(cuda_dev, G):
    lib = _lib.lib()
    d, N, F, cap = cuda_dev, 5000, 32, 2048
    g = torch.Generator().manual_seed(0)
    dense = []          # per-rank dense gradient tables, non-zero on that rank's touched rows only
    packed = []
    for r in range(G):
        n_t = [700, 1500, 33, 0, 2048, 1, 999, 1200][r]       # an empty rank, a full list, a single row
        ids = torch.sort(torch.randperm(N, generator=g)[:n_t])[0].to(torch.int32)
        tabs = dict(emb=torch.zeros(N, F), conf=torch.zeros(N), dir=torch.zeros(N, 3), color=torch.zeros(N, 3))
        for k, t in tabs.items():
            t[ids.long()] = torch.randn((n_t,) + tuple(t.shape[1:]), generator=g)
        dense.append(tabs)
        idp = torch.full((n_t + 5,), -1, dtype=torch.int32)   # ascending ids, -1 beyond n_touched, cap > list length
        idp[:n_t] = ids
        out = torch.empty(cap, F + 8, device=d)
        td = {k: v.to(d) for k, v in tabs.items()}
        _lib.check(lib.pn_rows_pack(_lib.ptr(idp.to(d)), idp.numel(), cap, F, _lib.ptr(td["emb"]), _lib.ptr(td["conf"]), _lib.ptr(td["dir"]),
                                    _lib.ptr(td["color"]), _lib.ptr(out), _lib.stream()), "pn_rows_pack")
        packed.append(out)
    gathered = torch.cat(packed, 0).contiguous()
    for r in range(G):   # every rank starts from ITS OWN dense table and must end with the same bits
        td = {k: v.to(d).clone() for k, v in dense[r].items()}
        n_points = int(td["emb"].shape[0])
        slots = torch.full((n_points * G,), -1, dtype=torch.int32, device=d)
        for _ in range(2):      # the slot table is left all -1, so a second merge over the same buffer is idempotent
            _lib.check(lib.pn_rows_merge(_lib.ptr(gathered), G, cap, F, n_points, _lib.ptr(slots), _lib.ptr(td["emb"]),
                                         _lib.ptr(td["conf"]), _lib.ptr(td["dir"]), _lib.ptr(td["color"]), _lib.stream()), "pn_rows_merge")
            assert bool((slots == -1).all())
        for k in td:
            want = dense[0][k].clone()
            for rr in range(1, G):
                want = want + dense[rr][k]          # rank order 0, 1, 2, ...: fp32 adds in the same order as the kernel
            # rows touched by no rank keep the local value (zero); rows touched only by other ranks are filled in
            assert torch.equal(td[k].cpu(), want), (r, k)


def test_backward_reports_touched_rows(cuda_dev, tiny_scene):
    opt, tables, cam = tiny_scene
    d = cuda_dev
    torch.manual_seed(0)
    mod = PointAggregator(opt).to(d)
    mod.collect_touched = True
    N = tables["xyz"].shape[0]
    g = torch.Generator().manual_seed(3)
    R, SR, K = 16, opt.SR, opt.K
    pidx = torch.randint(0, N, (R, SR, K), generator=g, dtype=torch.int32)
    pidx[torch.rand(R, SR, K, generator=g) < 0.6] = -1
    xyz = tables["xyz"]
    loc_w = xyz[torch.randint(0, N, (R, SR), generator=g)] + 0.003 * torch.randn(R, SR, 3, generator=g)
    xyz_pers = tp.w2pers_points(xyz, cam["camrotc2w"], cam["campos"])[0]
    loc_pers = tp.w2pers_samples(loc_w[None], cam["camrotc2w"], cam["campos"])[0]
    dirs = torch.nn.functional.normalize(torch.randn(R, 3, generator=g), dim=-1)
    leaves = {k: tables[k].to(d).clone().requires_grad_(True) for k in ("points_embeding", "points_color", "points_dir", "points_conf")}
    dec, _ = mod.forward_tables(xyz.to(d), xyz_pers.to(d), leaves["points_embeding"][0], leaves["points_color"][0],
                                leaves["points_dir"][0], leaves["points_conf"][0, :, 0], torch.eye(3, device=d), pidx.to(d),
                                loc_pers.to(d), loc_w.to(d), dirs.to(d), opt.vsize)
    dec.sum().backward()
    call = mod.last_call
    n = int(call.n_touched.item())
    want = torch.unique(pidx[pidx >= 0]).to(torch.int32)
    assert n == want.numel()
    assert torch.equal(call.touched_ids[:n].cpu(), want)
    assert bool((call.touched_ids[n:] == -1).all())
    nz = torch.nonzero(leaves["points_embeding"].grad[0].abs().sum(-1)).flatten().cpu()
    assert set(nz.tolist()) <= set(want.long().tolist())


def test_count_touched_rows(cuda_dev):
    lib = _lib.lib()
    d, N = cuda_dev, 50000
    g = torch.Generator().manual_seed(1)
    pidx = torch.randint(0, N, (3000, 8), generator=g, dtype=torch.int32)
    pidx[torch.rand(3000, 8, generator=g) < 0.7] = -1
    mark = torch.empty(N, dtype=torch.int32, device=d)
    cnt = torch.empty(1, dtype=torch.int32, device=d)
    pd = pidx.to(d)
    _lib.check(lib.pn_count_touched_rows(_lib.ptr(pd), pd.numel(), N, _lib.ptr(mark), _lib.ptr(cnt), _lib.stream()), "pn_count_touched_rows")
    assert int(cnt.item()) == torch.unique(pidx[pidx >= 0]).numel()
    _lib.check(lib.pn_count_touched_rows(None, 0, N, _lib.ptr(mark), _lib.ptr(cnt), _lib.stream()), "pn_count_touched_rows")
    assert int(cnt.item()) == 0
