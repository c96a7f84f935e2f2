"""The C-ABI library loads and exports every symbol include/pointnerf_b200.h declares (no compute calls)."""
import os
import re
import subprocess

from pointnerf_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "pointnerf_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pn_[a-z0-9_]+)\s*\(", text)))


def test_library_is_built():
    assert os.path.isfile(_lib.LIB_PATH), "run `make` / __graft_entry__.build()"


def test_exports_match_header():
    syms = _header_symbols()
    assert len(syms) >= 18
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r" T (pn_[a-z0-9_]+)", out))
    missing = [s for s in syms if s not in exported]
    assert not missing, "declared in the header but not exported: %s" % missing
    extra = sorted(exported - set(syms))
    assert not extra, "exported but not declared in the header: %s" % extra


def test_binding_covers_header():
    l = _lib.lib()
    assert l.pn_abi_version() == _lib.ABI_VERSION
    assert sorted(_lib.EXPORTS) == _header_symbols()
    assert l.pn_last_error() is not None


def test_size_queries_need_no_gpu():
    import ctypes
    l = _lib.lib()
    assert l.pn_grid_workspace_bytes(1000, 4096) > 0
    assert l.pn_compact_workspace_bytes(1000) > 0
    assert l.pn_composite_workspace_bytes(4096, 128, 20000) > 0
    assert l.pn_prune_workspace_bytes(1000) > 0 and l.pn_voxel_downsample_workspace_bytes(1000) > 0
    assert l.pn_color_mse_workspace_bytes() > 0
    assert l.pn_get_backward_schedule() == 1 and l.pn_set_backward_schedule(7) < 0 and b"pn_set_backward_schedule" in l.pn_last_error()
    assert l.pn_set_backward_schedule(0) == 0 and l.pn_get_backward_schedule() == 0 and l.pn_set_backward_schedule(1) == 0
    p = _lib.AggParams(32, 3, 5, 4, 256, 128, 8, 1, 80, 0.01, 1.0)
    assert l.pn_agg_saved_bytes(ctypes.byref(p), 100, 700) > 0
    assert l.pn_agg_workspace_bytes(ctypes.byref(p), 100, 700, 5000, 1) > l.pn_agg_workspace_bytes(ctypes.byref(p), 100, 700, 5000, 0)
    bad = _lib.AggParams(32, 3, 5, 4, 100, 128, 8, 1, 80, 0.01, 1.0)
    assert l.pn_agg_saved_bytes(ctypes.byref(bad), 1, 1) < 0
    assert b"shading_feature_num" in l.pn_last_error()


def test_no_cpu_fallback():
    import pytest
    import torch
    from pointnerf_b200.ray_march import ray_march, radiance_render, alpha_blend
    with pytest.raises(_lib.LibraryError):
        ray_march(torch.zeros(1, 2, 3), torch.ones(1, 2, 3, dtype=torch.bool), torch.zeros(1, 2, 3, 5), radiance_render, alpha_blend)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "pointnerf_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
                assert "from .. import oracle" not in src and "/root/reference" not in src, f


def test_frustum_host_arithmetic_matches_oracle():
    """get_hyperparameters of the perspective-frustum querier (query_point_indices.py:52-81) is host numpy in the product too:
    same ranges / voxel sizes / dims / limits as the oracle's restatement, for python floats and for the float32 0-d arrays the
    reference's caller passes (neural_points.py:786), whose promotion rules differ."""
    import numpy as np
    from oracle import frustum_ref
    from pointnerf_b200 import synthetic as syn
    from pointnerf_b200.query_frustum import frustum_hyperparameters
    for name in syn.FRUSTUM_CASES:
        opt, _, cam, _ = syn.frustum_case(name)
        for conv in (lambda v: v, lambda v: np.asarray(v, np.float32)):
            near, far = conv(cam["near"]), conv(cam["far"])
            a = frustum_hyperparameters(opt, cam["h"], cam["w"], cam["intrinsic"], near, far)
            b = frustum_ref.get_hyperparameters(opt, cam["h"], cam["w"], cam["intrinsic"], near, far)
            assert set(a) == set(b)
            for k in a:
                assert np.array_equal(np.asarray(a[k]), np.asarray(b[k])) and np.asarray(a[k]).dtype == np.asarray(b[k]).dtype, k
    l = _lib.lib()
    assert l.pn_frustum_grid_workspace_bytes(1000, 4096) > l.pn_grid_workspace_bytes(1000, 4096)
