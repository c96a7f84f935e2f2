"""The CUDA path against files written by the reference's OWN per-ray driver (tests/golden/driver_*_ref.npz: the reference's
NeuralPointsRayMarching.forward + fill_invalid over NeuralPoints.forward over its querier class and kernels, with its
PointAggregator / ray_march, run whole on the CPU of the build container by oracle/make_golden.py: driver).  Forward outputs
within 1e-4 of the tensor's scale (fp32; LeakyReLU kinks adjudicated by the oracle's fp64 evaluation), ray mask exact; both
queriers, reference seams and fused driver.  (Gradients: tests/test_gpu_pipeline.py and test_gpu_wcoord0_frustum.py against
oracle/pipeline_ref.py, which tests/test_oracle.py pins to the same reference driver.)"""
import os

import numpy as np
import pytest
import torch

from oracle import frustum_ref, make_golden, pipeline_ref
from pointnerf_b200.neural_points import NeuralPointsRayMarching
from _util import assert_close

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("kind", ["world", "frustum"])
def test_forward_equals_reference_driver_output(kind, cuda_dev):
    if kind == "frustum" and not frustum_ref.available():
        pytest.skip("oracle/_ref/libpn_ref_querier_frustum.so not built")
    z, wts = np.load(os.path.join(GOLDEN, "driver_%s_ref.npz" % kind)), np.load(os.path.join(GOLDEN, "driver_weights.npz"))
    opt, tables, cam, raydir, pix, bg = make_golden.driver_case(kind)
    assert np.array_equal(raydir.numpy(), z["raydir"]) and np.array_equal(pix.numpy(), z["pixel_idx"])
    dev = cuda_dev
    model = NeuralPointsRayMarching(opt, device=dev)
    model.aggregator.load_state_dict({k: torch.from_numpy(wts[k]) for k in wts.files})
    model.neural_points.set_points(tables["xyz"], tables["points_embeding"], tables["points_color"], tables["points_dir"],
                                   tables["points_conf"], Rw2c=tables["Rw2c"])
    cam64 = dict(cam, near=np.asarray(cam["near"], np.float32), far=np.asarray(cam["far"], np.float32)) if kind == "frustum" else cam
    ref64 = pipeline_ref.run(opt, tables, cam64, raydir, bg, model.aggregator.state_dict(), need_grads=False, dtype=torch.float64,
                             frustum_pix=pix if kind == "frustum" else None)
    near, far = torch.tensor([cam["near"]], dtype=torch.float32), torch.tensor([cam["far"]], dtype=torch.float32)
    for fused in (False, True):
        with torch.no_grad():
            out = model(cam["campos"].to(dev), raydir.to(dev), bg_color=bg.to(dev), camrotc2w=cam["camrotc2w"].to(dev),
                        pixel_idx=pix.to(dev), near=near, far=far, h=torch.tensor([cam["h"]]), w=torch.tensor([cam["w"]]),
                        intrinsic=torch.from_numpy(cam["intrinsic"])[None], fused=fused)     # as the reference's data loader hands them
        assert np.array_equal(out["ray_mask"].cpu().numpy(), z["ray_mask"])
        for k in ("coarse_raycolor", "coarse_point_opacity", "coarse_is_background", "coarse_mask", "queried_shading"):
            assert_close(out[k], z[k], "%s %s %s" % (kind, "fused" if fused else "seams", k), want64=ref64[k])
