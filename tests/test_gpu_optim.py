"""pn_adam_step against torch.optim.Adam on the same tensors."""
import pytest
import torch

from pointnerf_b200.optim import Adam

pytestmark = pytest.mark.gpu


def test_adam_matches_torch(cuda_dev):
    g = torch.Generator(device="cuda").manual_seed(0)
    shapes = [(1, 5000, 32), (1, 5000, 1), (1, 5000, 3), (256, 284), (256,), (1, 256), (1,), (128,), (7, 3)]
    ours = [torch.nn.Parameter(torch.randn(s, device=cuda_dev, generator=g)) for s in shapes]
    ref = [torch.nn.Parameter(p.detach().clone()) for p in ours]
    o1 = Adam(ours, lr=2e-3, betas=(0.9, 0.999))
    o2 = torch.optim.Adam(ref, lr=2e-3, betas=(0.9, 0.999), foreach=False, fused=False)
    for it in range(25):
        for a, b in zip(ours, ref):
            gr = torch.randn(a.shape, device=cuda_dev, generator=g) * (0.0 if it % 7 == 3 else 1.0)    # zero-gradient steps too
            if it % 5 == 2 and a.dim() == 3:
                gr[:, ::2] = 0          # untouched point rows keep moving on their momentum, as in torch
            a.grad, b.grad = gr, gr.clone()
        o1.step()
        o2.step()
    for a, b, s in zip(ours, ref, shapes):
        err = float((a.detach() - b.detach()).abs().max() / b.detach().abs().max())
        assert err <= 2e-6, (s, err)
        st = o2.state[b]
        assert float((o1.state[a]["exp_avg"] - st["exp_avg"]).abs().max()) <= 5e-6 * float(st["exp_avg"].abs().max())
        assert float((o1.state[a]["exp_avg_sq"] - st["exp_avg_sq"]).abs().max()) <= 5e-6 * float(st["exp_avg_sq"].abs().max())
        assert o1.state[a]["step"] == 25


def test_adam_skips_missing_grads_and_views(cuda_dev):
    flat = torch.zeros(1000, device=cuda_dev)
    p = torch.nn.Parameter(torch.ones(10, 10, device=cuda_dev))
    q = torch.nn.Parameter(torch.ones(3, device=cuda_dev))
    p.grad = flat[64:164].view(10, 10)          # a view into a flat gradient buffer (as the fused backward leaves them)
    p.grad.fill_(1.0)
    opt = Adam([p, q], lr=0.1)
    opt.step()
    assert torch.allclose(p, torch.full_like(p, 0.9), atol=1e-6) and bool((q == 1).all()) and q not in opt.state
    opt.zero_grad()
    assert p.grad is None


def test_row_adam_equals_torch_sparse_adam(cuda_dev):
    """pn_adam_rows (lazy, touched-row Adam: an OPTION, not the reference's optimiser) against torch.optim.SparseAdam on the
    same sparse gradients: parameters and both moments, rows outside the id list untouched, over several steps."""
    from pointnerf_b200.optim import RowAdam
    N = 5000
    g = torch.Generator().manual_seed(3)
    widths = (32, 1, 3, 3)
    mine = [torch.nn.Parameter(torch.randn(1, N, w, generator=g).to(cuda_dev)) for w in widths]
    ref = [torch.nn.Parameter(p.detach().clone().reshape(N, w)) for p, w in zip(mine, widths)]
    opt = RowAdam(mine, lr=2e-3, betas=(0.9, 0.999))
    ropt = torch.optim.SparseAdam(ref, lr=2e-3, betas=(0.9, 0.999))
    for it in range(6):
        ids = torch.unique(torch.randint(0, N, (400,), generator=g)).to(torch.int32)
        padded = torch.cat([ids, torch.full((37,), -1, dtype=torch.int32)]).to(cuda_dev)      # the backward's list is -1 padded
        for p, r, w in zip(mine, ref, widths):
            dense = torch.zeros(N, w)
            dense[ids.long()] = torch.randn(len(ids), w, generator=g)
            p.grad = dense.reshape(p.shape).to(cuda_dev)
            r.grad = torch.sparse_coo_tensor(ids.long()[None].to(cuda_dev), dense[ids.long()].to(cuda_dev), (N, w)).coalesce()
        before = [p.detach().clone() for p in mine]
        opt.step(padded)
        ropt.step()
        untouched = torch.ones(N, dtype=torch.bool)
        untouched[ids.long()] = False
        for p, r, b, w in zip(mine, ref, before, widths):
            assert torch.equal(p.detach().reshape(N, w)[untouched], b.reshape(N, w)[untouched])
            err = float((p.detach().reshape(N, w) - r.detach()).abs().max())
            assert err <= 2e-7 * float(r.detach().abs().max()), (it, w, err)
    for st, r in zip(opt.state, ref):
        rs = ropt.state[r]
        assert float((st["exp_avg"].reshape(rs["exp_avg"].shape) - rs["exp_avg"]).abs().max()) <= 1e-7 * float(rs["exp_avg"].abs().max())
        assert float((st["exp_avg_sq"].reshape(rs["exp_avg_sq"].shape) - rs["exp_avg_sq"]).abs().max()) <= 1e-7 * float(rs["exp_avg_sq"].abs().max())
