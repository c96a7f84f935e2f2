"""tcgen05 3xTF32 GEMMs (csrc/gemm_tc.cu) vs float64 matmul through the C ABI (pn_gemm_nt / pn_gemm_tn), with the row count
given by the host and - the way the training step uses them - read from device memory under a larger capacity."""
import pytest
import torch

from pointnerf_b200 import _lib

pytestmark = pytest.mark.gpu


def _nt(A, B, bias, leaky, slope, rows_dev=None):
    """A may hold MORE rows than valid (capacity); rows_dev = device int32[1] with the valid count."""
    lib = _lib.lib()
    M, Kd = A.shape
    N = B.shape[0]
    C = torch.full((M, N), float("nan"), device=A.device)
    img = torch.empty(lib.pn_gemm_nt_image_bytes(N, Kd), dtype=torch.uint8, device=A.device)
    _lib.check(lib.pn_gemm_nt(_lib.ptr(A), A.stride(0), _lib.ptr(B), B.stride(0), _lib.ptr(C), C.stride(0), M, _lib.ptr(rows_dev), N, Kd,
                              _lib.ptr(bias), int(leaky), float(slope), _lib.ptr(img), _lib.stream()), "pn_gemm_nt")
    torch.cuda.synchronize()
    return C


def _tn(A, B, Kout, rows_dev=None):
    lib = _lib.lib()
    M, N = A.shape
    Kb = B.shape[1]
    C = torch.full((N, Kout), float("nan"), device=A.device)
    cs = torch.full((N,), float("nan"), device=A.device)
    scratch = torch.empty(lib.pn_gemm_tn_scratch_floats(M, N, Kb), device=A.device)
    _lib.check(lib.pn_gemm_tn(_lib.ptr(A), A.stride(0), _lib.ptr(B), B.stride(0), _lib.ptr(C), C.stride(0), Kout,
                              _lib.ptr(cs), M, _lib.ptr(rows_dev), N, Kb, _lib.ptr(scratch), _lib.stream()), "pn_gemm_tn")
    torch.cuda.synchronize()
    return C, cs


def _rel(a, b):
    return float((a.double() - b).abs().max() / b.abs().max())


@pytest.mark.parametrize("M,N,Kd", [(128, 256, 32), (300, 256, 288), (1000, 256, 272), (77, 128, 288), (513, 288, 256),
                                    (260, 272, 256), (129, 128, 128), (5000, 256, 256), (1, 128, 128), (40000, 256, 224)])
def test_nt_matches_fp64(cuda_dev, M, N, Kd):
    g = torch.Generator().manual_seed(M + N + Kd)
    A = torch.randn(M, Kd, generator=g).to(cuda_dev)
    B = (torch.randn(N, Kd, generator=g) / Kd ** 0.5).to(cuda_dev)
    bias = torch.randn(N, generator=g).to(cuda_dev)
    ref = A.double() @ B.double().T + bias.double()
    ref = torch.where(ref > 0, ref, 0.01 * ref)
    C = _nt(A, B, bias, 1, 0.01)
    assert not torch.isnan(C).any(), "outputs left unwritten"
    assert _rel(C, ref) < 5e-6, "rel err %.3g" % _rel(C, ref)
    # no bias / no activation
    C = _nt(A, B, None, 0, 0.01)
    assert _rel(C, A.double() @ B.double().T) < 5e-6


@pytest.mark.parametrize("cap,rows", [(5000, 0), (5000, 1), (5000, 257), (5000, 4999), (300000, 1234), (300000, 70001)])
def test_nt_device_row_count(cuda_dev, cap, rows):
    """The row count lives on the device: rows beyond it are neither read nor written, the result equals the host-count run."""
    g = torch.Generator().manual_seed(cap + rows)
    N, Kd = 256, 288
    A = torch.randn(cap, Kd, generator=g).to(cuda_dev)
    B = (torch.randn(N, Kd, generator=g) / Kd ** 0.5).to(cuda_dev)
    bias = torch.randn(N, generator=g).to(cuda_dev)
    cnt = torch.tensor([rows], dtype=torch.int32, device=cuda_dev)
    C = _nt(A, B, bias, 1, 0.01, rows_dev=cnt)
    assert torch.isnan(C[rows:]).all(), "rows beyond the device count were written"
    if rows:
        want = _nt(A[:rows].contiguous(), B, bias, 1, 0.01)
        assert torch.equal(C[:rows], want)


@pytest.mark.parametrize("M,N,Kb,Kout", [(32, 128, 128, 128), (1000, 256, 288, 284), (4097, 256, 272, 263), (700, 128, 288, 280),
                                         (20000, 256, 256, 256), (5, 128, 128, 128), (3000, 256, 512, 501), (140000, 256, 288, 284)])
def test_tn_matches_fp64(cuda_dev, M, N, Kb, Kout):
    g = torch.Generator().manual_seed(M + N + Kb)
    A = torch.randn(M, N, generator=g).to(cuda_dev)
    B = torch.randn(M, Kb, generator=g).to(cuda_dev)
    ref = (A.double().T @ B.double())[:, :Kout]
    ref_cs = A.double().sum(0)
    C, cs = _tn(A, B, Kout)
    assert not torch.isnan(C).any() and not torch.isnan(cs).any()
    # the fp32 accumulator in TMEM is a chain of M / splits / 8 additions per element (~700 at M = 140 000): its rounding
    # error grows with the chain length; 1.3e-5 of max measured at M = 140 000 (the step's own budget is 1e-4)
    tol = 5e-6 if M <= 20000 else 2e-5
    assert _rel(C, ref) < tol, "rel err %.3g" % _rel(C, ref)
    assert _rel(cs, ref_cs) < 3e-6
    C2, cs2 = _tn(A, B, Kout)
    assert torch.equal(C, C2) and torch.equal(cs, cs2)   # fixed-order reduction: bit-reproducible


@pytest.mark.parametrize("cap,rows,N,Kb", [(5000, 0, 256, 288), (5000, 1, 256, 288), (5000, 300, 128, 128), (300000, 1234, 256, 272),
                                           (300000, 134826, 256, 288), (20000, 19999, 128, 288)])
def test_tn_device_row_count(cuda_dev, cap, rows, N, Kb):
    """Split geometry cut on the device from the true row count."""
    g = torch.Generator().manual_seed(cap + rows)
    A = torch.randn(cap, N, generator=g).to(cuda_dev)
    B = torch.randn(cap, Kb, generator=g).to(cuda_dev)
    cnt = torch.tensor([rows], dtype=torch.int32, device=cuda_dev)
    C, cs = _tn(A, B, Kb, rows_dev=cnt)
    assert not torch.isnan(C).any() and not torch.isnan(cs).any()
    if rows == 0:
        assert float(C.abs().max()) == 0.0 and float(cs.abs().max()) == 0.0
        return
    ref = A[:rows].double().T @ B[:rows].double()
    assert _rel(C, ref) < (5e-6 if rows <= 20000 else 2e-5)
    assert _rel(cs, A[:rows].double().sum(0)) < 3e-6
