"""pn_compact_samples_static: the work-list builder with its scratch in a `static __device__` array inside the library, for
callers that cannot allocate (CUDA-graph capture, serving loops); no
counterpart in the reference.  Checked through the C ABI against pn_compact_samples (integer / index work: bit-identical),
on two calls in a row with different inputs (the second call finds the first call's contents in the shared scratch).

Collected LAST on purpose (file name)."""
import pytest
import torch

from pointnerf_b200 import _lib

pytestmark = pytest.mark.gpu


def _outputs(n_slots, K, dev):
    # poison values: every entry the contract defines must be written by the call itself
    return dict(sample_list=torch.full((max(n_slots, 1),), -7, dtype=torch.int32, device=dev),
                pair_start=torch.full((n_slots + 1,), -7, dtype=torch.int32, device=dev),
                pair_src=torch.full((max(n_slots * K, 1),), -7, dtype=torch.int32, device=dev),
                counts=torch.full((2,), -7, dtype=torch.int32, device=dev),
                slot_row=torch.full((max(n_slots, 1),), -7, dtype=torch.int32, device=dev))


def _reference(pidx, n_slots, K):
    lib = _lib.lib()
    o = _outputs(n_slots, K, pidx.device)
    ws_bytes = lib.pn_compact_workspace_bytes(n_slots)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=pidx.device)
    _lib.check(lib.pn_compact_samples(_lib.ptr(pidx), n_slots, K, _lib.ptr(o["sample_list"]), _lib.ptr(o["pair_start"]),
                                      _lib.ptr(o["pair_src"]), _lib.ptr(o["counts"]), _lib.ptr(o["slot_row"]), _lib.ptr(ws), ws_bytes,
                                      _lib.stream()), "pn_compact_samples")
    return o


def _static(pidx, n_slots, K, o):
    """Outputs are allocated by the caller BEFORE the call; the call itself allocates nothing."""
    lib = _lib.lib()
    return lib.pn_compact_samples_static(_lib.ptr(pidx), n_slots, K, _lib.ptr(o["sample_list"]), _lib.ptr(o["pair_start"]),
                                         _lib.ptr(o["pair_src"]), _lib.ptr(o["counts"]), _lib.ptr(o["slot_row"]), _lib.stream())


def _same(a, b, n_slots):
    S, M = (int(v) for v in b["counts"].cpu())
    assert torch.equal(a["counts"], b["counts"])
    assert torch.equal(a["sample_list"][:S], b["sample_list"][:S])
    assert torch.equal(a["pair_start"][:S + 1], b["pair_start"][:S + 1])
    assert torch.equal(a["pair_src"][:M], b["pair_src"][:M])
    assert torch.equal(a["slot_row"][:n_slots], b["slot_row"][:n_slots])
    return S, M


def test_static_scratch_variant_matches_on_two_calls_in_a_row(cuda_dev):
    lib = _lib.lib()
    d = cuda_dev
    assert lib.pn_compact_samples_static_max_slots() == 524288
    g = torch.Generator().manual_seed(77)
    # first call: a large batch (cfg2 shape: 4096 rays x 80 samples, K = 8); second: a smaller, denser one with another K -
    # it runs over the scratch the first call left behind
    cases = []
    for n_slots, K, p_empty in ((4096 * 80, 8, 0.93), (1500 * 24 + 5, 16, 0.4)):
        pidx = torch.randint(0, 1_000_000, (n_slots, K), generator=g, dtype=torch.int32)
        pidx[torch.rand(n_slots, K, generator=g) < p_empty] = -1
        cases.append((n_slots, K, pidx.to(d).contiguous()))
    outs = [_outputs(n, K, d) for n, K, _ in cases]
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats(d)["allocation.all.allocated"]
    for (n, K, pidx), o in zip(cases, outs):                      # two calls in a row, nothing in between
        rc = _static(pidx, n, K, o)
        assert rc == 0, lib.pn_last_error().decode()
    assert torch.cuda.memory_stats(d)["allocation.all.allocated"] == before, "the static variant must not allocate"
    torch.cuda.synchronize()
    for (n, K, pidx), o in zip(cases, outs):
        S, M = _same(o, _reference(pidx, n, K), n)
        assert S > 1000 and M >= S
        assert M == int((pidx >= 0).sum()) and S == int((pidx >= 0).any(-1).sum())
    # and once more on the first input: same bits again
    o2 = _outputs(cases[0][0], cases[0][1], d)
    assert _static(cases[0][2], cases[0][0], cases[0][1], o2) == 0
    torch.cuda.synchronize()
    _same(o2, outs[0], cases[0][0])


def test_static_scratch_variant_rejects_oversized_and_handles_empty(cuda_dev):
    lib = _lib.lib()
    d = cuda_dev
    K = 8
    n_big = lib.pn_compact_samples_static_max_slots() + 1
    pidx = torch.full((n_big, K), -1, dtype=torch.int32, device=d)
    o = _outputs(n_big, K, d)
    rc = _static(pidx, n_big, K, o)
    assert rc < 0
    msg = lib.pn_last_error().decode()
    assert "exceeds" in msg and "524288" in msg
    assert int(o["counts"][0]) == -7                               # rejected before anything was launched
    # exactly the maximum is accepted; no neighbours anywhere -> S = M = 0
    rc = _static(pidx, n_big - 1, K, o)
    assert rc == 0, lib.pn_last_error().decode()
    torch.cuda.synchronize()
    assert o["counts"].tolist() == [0, 0] and int(o["pair_start"][0]) == 0
    assert bool((o["slot_row"][:n_big - 1] == -1).all())
    # zero slots
    o0 = _outputs(0, K, d)
    assert _static(pidx, 0, K, o0) == 0
    torch.cuda.synchronize()
    assert o0["counts"].tolist() == [0, 0]
