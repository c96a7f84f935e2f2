"""Import shim for the UNMODIFIED reference modules (build container only).

``/root/reference`` does not exist on the GPU box, so nothing that runs there
may import this module; it is used by ``oracle/make_golden.py`` and by the
``not gpu`` tests that validate the restatements (skipped when the tree is
absent).

Two names removed from SciPy >= 1.15 (``sph_harm``, ``lpmn``) are imported at
module scope by ``utils/spherical.py:2`` of the reference although the shipped
configuration (``agg_distance_kernel=linear``) never calls them; they are
stubbed so that the import succeeds.  Nothing else is patched.
"""
import os
import sys
import types

REF_ROOT = os.environ.get("POINTNERF_REFERENCE", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "models", "aggregators", "point_aggregators.py"))


def _stub_scipy():
    import scipy.special as sp

    def _gone(*a, **k):  # pragma: no cover - never called in the shipped config
        raise NotImplementedError("scipy.special name removed; unused on the hot path")

    for name in ("sph_harm", "lpmn", "lpmv", "factorial"):
        if not hasattr(sp, name):
            setattr(sp, name, _gone)


_loaded = {}


def load():
    """Returns a namespace with the reference callables used by the oracle."""
    if "ns" in _loaded:
        return _loaded["ns"]
    if not available():
        raise RuntimeError("reference tree not present at %s" % REF_ROOT)
    _stub_scipy()
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    from models.aggregators.point_aggregators import PointAggregator
    from models.rendering.diff_ray_marching import (
        ray_march, alpha_ray_march, near_far_linear_ray_generation)
    from models.rendering.diff_render_func import radiance_render, alpha_blend
    from models.helpers.networks import positional_encoding
    ns = types.SimpleNamespace(
        PointAggregator=PointAggregator, ray_march=ray_march, alpha_ray_march=alpha_ray_march,
        near_far_linear_ray_generation=near_far_linear_ray_generation,
        radiance_render=radiance_render, alpha_blend=alpha_blend,
        positional_encoding=positional_encoding)
    _loaded["ns"] = ns
    return ns
