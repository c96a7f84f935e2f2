"""Host-side logic that needs no GPU: which querier NeuralPoints constructs, option defaults it applies, and the refusals of
the perspective-frustum querier (they are decided from the options before any kernel is launched)."""
import numpy as np
import pytest
import torch

from pointnerf_b200 import query, query_frustum, synthetic as syn
from pointnerf_b200.neural_points import NeuralPoints, NeuralPointsRayMarching
from pointnerf_b200.options import make_opt


def test_neural_points_picks_the_querier_like_the_reference():
    """neural_points.py:329-331: query_size (0,0,0) means kernel_size; wcoord_query > 0 -> world grid, else the frustum one."""
    cpu = torch.device("cpu")
    o = make_opt(query_size=(0, 0, 0), kernel_size=(5, 5, 3))
    nps = NeuralPoints(o.point_features_dim, 0, o, cpu)
    assert isinstance(nps.querier, query.lighting_fast_querier) and tuple(o.query_size) == (5, 5, 3)
    o = make_opt(wcoord_query=0)
    nps = NeuralPoints(o.point_features_dim, 0, o, cpu)
    assert isinstance(nps.querier, query_frustum.lighting_fast_querier)
    nps.reset_querier()
    assert isinstance(nps.querier, query_frustum.lighting_fast_querier)


def test_frustum_querier_refusals_are_decided_on_the_host():
    opt, _, cam, _ = syn.frustum_case("crop")
    for kw, word in ((dict(NN=0), "NN"), (dict(inverse=1), "inverse"), (dict(query_size=(3, 3, 3)), "query_size")):
        o = syn.config_opt("tiny", **{**syn.FRUSTUM_CASES["crop"]["opt"], **kw})
        q = query_frustum.lighting_fast_querier(torch.device("cpu"), o)
        hp = q.get_hyperparameters(cam["h"], cam["w"], cam["intrinsic"], cam["near"], cam["far"])
        with pytest.raises(NotImplementedError, match=word):
            q._check(hp)
    q = query_frustum.lighting_fast_querier(torch.device("cpu"), opt)
    q._check(q.get_hyperparameters(cam["h"], cam["w"], cam["intrinsic"], cam["near"], cam["far"]))   # the case itself is accepted
    assert q.grid_report() is None


def test_frustum_hyperparameters_shapes():
    """ceil([w, h, D] / vscale) cells, voxel sizes that tile the image plane exactly, limits as multiples of the UNSCALED size."""
    opt, _, cam, _ = syn.frustum_case("crop")
    hp = query_frustum.frustum_hyperparameters(opt, cam["h"], cam["w"], cam["intrinsic"], cam["near"], cam["far"])
    assert hp["scaled_vdim"].tolist() == [14, 15, 32] and hp["vdim"].tolist() == [40, 30, 64]
    assert np.allclose(hp["ranges"][:2] + hp["vsize"][:2] * hp["vdim"][:2], hp["ranges"][3:5], rtol=1e-6)
    assert np.isclose(hp["radius_limit"], 3.0 * max(hp["vsize"][0], hp["vsize"][1])) and np.isclose(hp["depth_limit"], 2.0 * hp["vsize"][2])


def test_frustum_driver_needs_pixels():
    """With wcoord_query = 0 rays are addressed by pixel: the ray-direction-only entry points say so instead of guessing."""
    o = make_opt(wcoord_query=0)
    cpu = torch.device("cpu")
    m = NeuralPointsRayMarching.__new__(NeuralPointsRayMarching)
    torch.nn.Module.__init__(m)
    m.opt = o
    m.neural_points = NeuralPoints(o.point_features_dim, 0, o, cpu)
    with pytest.raises(NotImplementedError, match="pixel"):
        m.sample(torch.zeros(1, 3), torch.zeros(1, 4, 3), torch.eye(3)[None], 0.1, 1.0)


def test_scalars_reach_the_frustum_querier_as_in_the_reference():
    """near / far / h / w arrive as tensors from the data loader; the reference turns them into 0-d numpy arrays
    (neural_points.py:786) whose dtypes (float32, int64) steer numpy's promotion in get_hyperparameters."""
    from pointnerf_b200.neural_points import _as_reference_scalar
    h = _as_reference_scalar(torch.tensor([48, 48]), "max")
    n = _as_reference_scalar(torch.tensor([0.3, 0.4]), "min")
    assert isinstance(h, np.ndarray) and h.shape == () and h.dtype == np.int64 and int(h) == 48
    assert isinstance(n, np.ndarray) and n.dtype == np.float32 and float(n) == float(np.float32(0.3))
    assert _as_reference_scalar(48, "max") == 48 and _as_reference_scalar(0.3, "min") == 0.3
    opt, _, cam, _ = syn.frustum_case("crop")
    a = query_frustum.frustum_hyperparameters(opt, h=_as_reference_scalar(torch.tensor([cam["h"]]), "max"),
                                              w=_as_reference_scalar(torch.tensor([cam["w"]]), "max"), intrinsic=cam["intrinsic"],
                                              near_depth=_as_reference_scalar(torch.tensor([cam["near"]]), "min"),
                                              far_depth=_as_reference_scalar(torch.tensor([cam["far"]]), "max"))
    b = query_frustum.frustum_hyperparameters(opt, cam["h"], cam["w"], cam["intrinsic"], cam["near"], cam["far"])
    assert a["scaled_vdim"].tolist() == b["scaled_vdim"].tolist() and np.allclose(a["vsize"], b["vsize"], rtol=1e-6)
