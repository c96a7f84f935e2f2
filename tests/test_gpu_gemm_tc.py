"""tcgen05 3xTF32 GEMMs (csrc/gemm_tc.cu) vs float64 matmul, and vs the exact-fp32 SIMT path, through the C ABI test hooks."""
import pytest
import torch

from pointnerf_b200 import _lib

pytestmark = pytest.mark.gpu


def _nt(A, B, bias, leaky, slope, path, image=False):
    lib = _lib.lib()
    _lib.check(lib.pn_set_mlp_path(path), "pn_set_mlp_path")
    M, Kd = A.shape
    N = B.shape[0]
    C = torch.full((M, N), float("nan"), device=A.device)
    img = torch.empty(lib.pn_debug_nt_image_bytes(N, Kd), dtype=torch.uint8, device=A.device) if image else None
    try:
        _lib.check(lib.pn_debug_gemm_nt(_lib.ptr(A), A.stride(0), _lib.ptr(B), B.stride(0), _lib.ptr(C), C.stride(0), M, N, Kd,
                                        _lib.ptr(bias), int(leaky), float(slope), _lib.ptr(img), _lib.stream()), "pn_debug_gemm_nt")
        torch.cuda.synchronize()
    finally:
        lib.pn_set_mlp_path(1)
    return C


def _tn(A, B, Kout, path):
    lib = _lib.lib()
    _lib.check(lib.pn_set_mlp_path(path), "pn_set_mlp_path")
    M, N = A.shape
    Kb = B.shape[1]
    C = torch.full((N, Kout), float("nan"), device=A.device)
    cs = torch.full((N,), float("nan"), device=A.device)
    scratch = torch.empty(lib.pn_debug_gemm_tn_scratch_floats(M, N, Kb), device=A.device)
    try:
        _lib.check(lib.pn_debug_gemm_tn(_lib.ptr(A), A.stride(0), _lib.ptr(B), B.stride(0), _lib.ptr(C), C.stride(0), Kout,
                                        _lib.ptr(cs), M, N, Kb, _lib.ptr(scratch), _lib.stream()), "pn_debug_gemm_tn")
        torch.cuda.synchronize()
    finally:
        lib.pn_set_mlp_path(1)
    return C, cs


def _rel(a, b):
    return float((a.double() - b).abs().max() / b.abs().max())


@pytest.mark.parametrize("M,N,Kd", [(128, 256, 32), (300, 256, 288), (1000, 256, 272), (77, 128, 288), (513, 288, 256),
                                    (260, 272, 256), (129, 128, 128), (5000, 256, 256)])
def test_nt_matches_fp64(cuda_dev, M, N, Kd):
    g = torch.Generator().manual_seed(M + N + Kd)
    A = torch.randn(M, Kd, generator=g).to(cuda_dev)
    B = (torch.randn(N, Kd, generator=g) / Kd ** 0.5).to(cuda_dev)
    bias = torch.randn(N, generator=g).to(cuda_dev)
    ref = A.double() @ B.double().T + bias.double()
    ref = torch.where(ref > 0, ref, 0.01 * ref)
    for path, image in ((1, True), (1, False), (0, False)):
        C = _nt(A, B, bias, 1, 0.01, path, image)
        assert not torch.isnan(C).any(), "path %d left outputs unwritten" % path
        assert _rel(C, ref) < 5e-6, "path %d image %d: rel err %.3g" % (path, image, _rel(C, ref))
    # no bias / no activation
    C = _nt(A, B, None, 0, 0.01, 1)
    assert _rel(C, A.double() @ B.double().T) < 5e-6


@pytest.mark.parametrize("M,N,Kb,Kout", [(32, 128, 128, 128), (1000, 256, 288, 284), (4097, 256, 272, 263), (700, 128, 288, 280),
                                         (20000, 256, 256, 256), (5, 128, 128, 128)])
def test_tn_matches_fp64(cuda_dev, M, N, Kb, Kout):
    g = torch.Generator().manual_seed(M + N + Kb)
    A = torch.randn(M, N, generator=g).to(cuda_dev)
    B = torch.randn(M, Kb, generator=g).to(cuda_dev)
    ref = (A.double().T @ B.double())[:, :Kout]
    ref_cs = A.double().sum(0)
    for path in (1, 0):
        C, cs = _tn(A, B, Kout, path)
        assert not torch.isnan(C).any() and not torch.isnan(cs).any()
        assert _rel(C, ref) < 5e-6, "path %d: rel err %.3g" % (path, _rel(C, ref))
        assert _rel(cs, ref_cs) < 3e-6
    C2, cs2 = _tn(A, B, Kout, 1)
    C1, cs1 = _tn(A, B, Kout, 1)
    assert torch.equal(C1, C2) and torch.equal(cs1, cs2)   # fixed-order reduction: bit-reproducible
