"""ctypes binding of libpointnerf_b200.so (include/pointnerf_b200.h).

There is no fallback: if the shared library is missing or an entry point fails, the
caller gets an exception.  Tensors are passed as raw device pointers; every call is
enqueued on torch's current CUDA stream.
"""
import ctypes
import os
import subprocess
from ctypes import POINTER, Structure, c_double, c_float, c_int, c_int32, c_int64, c_void_p, c_char_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(_HERE)
LIB_PATH = os.path.join(_HERE, "lib", "libpointnerf_b200.so")
ABI_VERSION = 11

GRID_INFO_LEN = 8
GI_N_OCC, GI_FIRST_POINT, GI_SLOT0_CELL, GI_N_INSIDE, GI_MAX_COUNT, GI_OVER_P, GI_OVER_MAX_O = range(7)


class GridParams(Structure):
    _fields_ = [("shift", c_float * 3), ("vsize", c_float * 3), ("dims", c_int32 * 3), ("dilate", c_int32 * 3),
                ("P", c_int32), ("max_o", c_int32)]


class FrustumQuery(Structure):
    _fields_ = [("ray_vsize", c_float * 2), ("vscale", c_int32 * 2), ("radius_limit2", c_float), ("depth_limit2", c_float),
                ("n_layers", c_int32), ("z_layers", c_int32), ("nn", c_int32)]


class AggParams(Structure):
    _fields_ = [("F", c_int32), ("feat_freqs", c_int32), ("dist_freqs", c_int32), ("view_freqs", c_int32),
                ("W", c_int32), ("C", c_int32), ("K", c_int32), ("dirs_per_ray", c_int32), ("SR", c_int32),
                ("leaky_slope", c_float), ("dist_scale", c_float), ("compact_out", c_int32), ("intrp_order", c_int32), ("forward_only", c_int32)]


WEIGHT_FIELDS = ("block1_0_w", "block1_0_b", "block1_2_w", "block1_2_b", "block3_0_w", "block3_0_b",
                 "block3_2_w", "block3_2_b", "alpha_0_w", "alpha_0_b", "color_0_w", "color_0_b",
                 "color_2_w", "color_2_b", "color_4_w", "color_4_b")
# state_dict keys of the reference's PointAggregator, in the same order
WEIGHT_KEYS = ("block1.0.weight", "block1.0.bias", "block1.2.weight", "block1.2.bias", "block3.0.weight",
               "block3.0.bias", "block3.2.weight", "block3.2.bias", "alpha_branch.0.weight", "alpha_branch.0.bias",
               "color_branch.0.weight", "color_branch.0.bias", "color_branch.2.weight", "color_branch.2.bias",
               "color_branch.4.weight", "color_branch.4.bias")


class AggWeights(Structure):
    _fields_ = [(n, c_void_p) for n in WEIGHT_FIELDS]


class AggInputs(Structure):
    _fields_ = [(n, c_void_p) for n in ("xyz", "xyz_pers", "embedding", "color", "dir", "conf", "Rw2c", "sample_pidx",
                                        "sample_loc_w", "sample_loc", "ray_dirs", "sample_list", "pair_start", "pair_src")] + \
               [(n, c_int64) for n in ("n_slots", "S", "M", "T")] + [("counts", c_void_p)]


class AdamSegment(Structure):
    _fields_ = [("param", c_void_p), ("grad", c_void_p), ("exp_avg", c_void_p), ("exp_avg_sq", c_void_p), ("n", c_int64)]


class AdamRowsTable(Structure):
    _fields_ = [("param", c_void_p), ("grad", c_void_p), ("exp_avg", c_void_p), ("exp_avg_sq", c_void_p), ("width", c_int32)]


class LibraryError(RuntimeError):
    pass


_lib = None


def build(verbose=False):
    """Compiles the library in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
    out = subprocess.run(["make", "-C", REPO_ROOT, "-j8"], capture_output=True, text=True)
    if out.returncode != 0:
        raise LibraryError("building libpointnerf_b200.so failed:\n" + out.stdout[-4000:] + out.stderr[-4000:])
    if verbose:
        print(out.stdout[-2000:])
    return LIB_PATH


def _declare(lib):
    P = c_void_p
    sig = {
        "pn_last_error": (c_char_p, []),
        "pn_abi_version": (c_int, []),
        "pn_launch_count": (c_int64, []),
        "pn_launch_count_reset": (None, []),
        "pn_gemm_nt_image_bytes": (c_int64, [c_int32, c_int32]),
        "pn_gemm_nt": (c_int, [P, c_int64, P, c_int64, P, c_int64, c_int64, P, c_int32, c_int32, P, c_int32, c_float, P, P]),
        "pn_gemm_tn_scratch_floats": (c_int64, [c_int64, c_int32, c_int32]),
        "pn_gemm_tn": (c_int, [P, c_int64, P, c_int64, P, c_int64, c_int32, P, c_int64, P, c_int32, c_int32, P, P]),
        "pn_grid_workspace_bytes": (c_int64, [c_int64, c_int64]),
        "pn_grid_build": (c_int, [P, c_int64, POINTER(GridParams), P, P, P, P, P, c_int64, P]),
        "pn_sample_depths": (c_int, [P, P, c_int64, c_int32, c_float, c_float, P, P]),
        "pn_ray_sample": (c_int, [P, P, P, c_int64, c_int64, c_int32, c_int32, POINTER(GridParams), P, P, P, P, P]),
        "pn_knn_query": (c_int, [P, POINTER(GridParams), P, P, P, P, P, c_int64, c_int32, c_int32, c_int32, c_float,
                                 P, P, P, P]),
        "pn_frustum_grid_workspace_bytes": (c_int64, [c_int64, c_int64]),
        "pn_frustum_grid_build": (c_int, [P, c_int64, POINTER(GridParams), P, P, P, P, P, c_int64, P]),
        "pn_frustum_sample": (c_int, [P, c_int64, c_int32, POINTER(GridParams), POINTER(c_int32), P, P, P, P, P]),
        "pn_frustum_knn_query": (c_int, [P, POINTER(GridParams), POINTER(FrustumQuery), P, P, P, P, P, c_int64, c_int32,
                                         c_int32, P, P, P, P]),
        "pn_w2pers": (c_int, [P, c_int64, P, P, P, P]),
        "pn_compact_workspace_bytes": (c_int64, [c_int64]),
        "pn_compact_samples": (c_int, [P, c_int64, c_int32, P, P, P, P, P, P, c_int64, P]),
        "pn_compact_samples_capped": (c_int, [P, c_int64, c_int32, c_int64, c_int64, P, P, P, P, P, P, c_int64, P]),
        "pn_compact_samples_static_max_slots": (c_int64, []),
        "pn_compact_samples_static": (c_int, [P, c_int64, c_int32, P, P, P, P, P, P]),
        "pn_agg_saved_bytes": (c_int64, [POINTER(AggParams), c_int64, c_int64]),
        "pn_agg_activation_signs": (c_int, [POINTER(AggParams), P, c_int64, c_int64, c_int64, c_int64, c_int64, P, P, P]),
        "pn_agg_workspace_bytes": (c_int64, [POINTER(AggParams), c_int64, c_int64, c_int64, c_int]),
        "pn_agg_forward": (c_int, [POINTER(AggParams), POINTER(AggInputs), POINTER(AggWeights), P, P, P, c_int64, P,
                                   c_int64, P]),
        "pn_agg_backward": (c_int, [POINTER(AggParams), POINTER(AggInputs), POINTER(AggWeights), P, P, c_int64,
                                    POINTER(AggWeights), c_int, P, P, P, P, P, P, P, c_int64, P]),
        "pn_agg_backward_wait_rows": (c_int, [P]),
        "pn_set_backward_schedule": (c_int, [c_int]),
        "pn_get_backward_schedule": (c_int, []),
        "pn_count_touched_rows": (c_int, [P, c_int64, c_int64, P, P, P]),
        "pn_rows_pack": (c_int, [P, c_int64, c_int64, c_int32, P, P, P, P, P, P]),
        "pn_rows_zero": (c_int, [P, c_int64, c_int64, c_int32, P, P, P, P, P]),
        "pn_adam_rows": (c_int, [P, c_int64, c_int64, POINTER(AdamRowsTable), c_int32, c_double, c_double, c_double, c_double, c_int64, P]),
        "pn_rows_merge": (c_int, [P, c_int32, c_int64, c_int32, c_int64, P, P, P, P, P, P]),
        "pn_ray_dist": (c_int, [P, P, c_int64, c_int32, c_float, c_int32, P, P]),
        "pn_composite_forward": (c_int, [P, P, P, P, P, c_int64, c_int32, c_int32, P, P, P, P, P, P]),
        "pn_composite_workspace_bytes": (c_int64, [c_int64, c_int32, c_int64]),
        "pn_composite_backward": (c_int, [P, P, P, P, P, P, P, P, P, P, P, P, P, c_int64, c_int32, c_int32, c_int64, P, P, P, P,
                                          c_int64, P]),
        "pn_gather_clamp": (c_int, [P, c_int64, P, c_int64, c_float, c_float, P, P]),
        "pn_color_mse_workspace_bytes": (c_int64, []),
        "pn_color_mse": (c_int, [P, P, P, P, c_int64, c_int32, c_float, P, P, P, P, c_int64, P]),
        "pn_probe_rays": (c_int, [P, P, P, P, P, P, P, P, P, P, c_int64, c_int32, c_int32, c_int32, c_float, c_float,
                                  P, P, P, P, P, P, P, P, P]),
        "pn_prune_workspace_bytes": (c_int64, [c_int64]),
        "pn_adam_step": (c_int, [POINTER(AdamSegment), c_int32, c_double, c_double, c_double, c_double, c_int64, P]),
        "pn_voxel_downsample_workspace_bytes": (c_int64, [c_int64]),
        "pn_voxel_downsample": (c_int, [P, c_int64, POINTER(c_float), POINTER(c_float), POINTER(c_int32), c_int32, P, P, P, P, P, P, P, c_int64, P]),
        "pn_prune_mark": (c_int, [P, c_int64, c_float, P, P, c_int64, P]),
        "pn_points_compact": (c_int, [P, c_int64, c_int32, P, P, P, P, P, P, P, P, P, P, P]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing: loud by design
        fn.restype = res
        fn.argtypes = args
    return sorted(sig)


EXPORTS = None


def lib():
    """Loads (once) and returns the shared library; raises if it is absent or has the wrong ABI."""
    global _lib, EXPORTS
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise LibraryError("%s not found: run `make` (or __graft_entry__.build()); there is no CPU fallback" % LIB_PATH)
        l = ctypes.CDLL(LIB_PATH)
        EXPORTS = _declare(l)
        if l.pn_abi_version() != ABI_VERSION:
            raise LibraryError("ABI mismatch: library %d, binding %d" % (l.pn_abi_version(), ABI_VERSION))
        _lib = l
    return _lib


def check(rc, what):
    if rc != 0:
        raise LibraryError("%s failed (%d): %s" % (what, rc, lib().pn_last_error().decode()))


def ptr(t):
    """Device pointer of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise LibraryError("pointnerf_b200 kernels need CUDA tensors (got %s); there is no CPU path" % t.device)
    if not t.is_contiguous():
        raise LibraryError("non-contiguous tensor passed to the C ABI")
    return t.data_ptr()


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream():
    """cudaStream_t of torch's current stream on the current device (every launch goes there)."""
    if _raw_stream is not None:                   # ~0.3 us; torch.cuda.current_stream() builds a Stream object (~15 us)
        return _raw_stream(torch.cuda.current_device())
    return torch.cuda.current_stream().cuda_stream


def launch_count():
    return int(lib().pn_launch_count())


def reset_launch_count():
    lib().pn_launch_count_reset()
