#!/usr/bin/env python
"""One launch of each MLP GEMM form at the training step's pair-level shape (M = 134 826 valid pairs of a cfg2 batch,
N = K = 256) through the C ABI (pn_gemm_nt / pn_gemm_tn): the target of `ncu --set full` captures
(profiles/r0x_ncu_full_pair_gemms.*) and of quick timing.  Usage: python tools/one_gemm.py [M] [reps]"""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch  # noqa: E402

from pointnerf_b200 import _lib  # noqa: E402


def main():
    M = int(sys.argv[1]) if len(sys.argv) > 1 else 134826
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    N = K = 256
    dev = torch.device("cuda", 0)
    lib = _lib.lib()
    A = torch.randn(M, K, device=dev)
    B = torch.randn(N, K, device=dev) / 16
    C = torch.empty(M, N, device=dev)
    bias = torch.randn(N, device=dev)
    img = torch.empty(lib.pn_gemm_nt_image_bytes(N, K), dtype=torch.uint8, device=dev)
    G = torch.randn(M, N, device=dev)
    W = torch.empty(N, K, device=dev)
    cs = torch.empty(N, device=dev)
    scratch = torch.empty(lib.pn_gemm_tn_scratch_floats(M, N, K), device=dev)
    for name, fn in (("nt", lambda: lib.pn_gemm_nt(_lib.ptr(A), K, _lib.ptr(B), K, _lib.ptr(C), N, M, None, N, K, _lib.ptr(bias), 1, 0.01, _lib.ptr(img), _lib.stream())),
                     ("tn", lambda: lib.pn_gemm_tn(_lib.ptr(G), N, _lib.ptr(A), K, _lib.ptr(W), K, K, _lib.ptr(cs), M, None, N, K, _lib.ptr(scratch), _lib.stream()))):
        for _ in range(2):
            _lib.check(fn(), name)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            _lib.check(fn(), name)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / reps * 1e3
        print("%s M=%d N=K=256: %.1f us per call (incl. pack / finish launches), %.1f TFLOP/s algorithmic" % (name, M, us, 2.0 * M * N * K / us / 1e6))


if __name__ == "__main__":
    main()
