"""Timing + full-size parity of the perspective-frustum querier (not collected by pytest; run on the GPU box):

    python tests/run_frustum_timing.py > gpurun_out/frustum_timing.json

Workload: the DTU feed-forward settings (dev_scripts/ete/dtu_dgt_d012_img0123_conf_color_dir_agg2.sh: vscale 2 2 1, kernel 3 3 3,
SR 40, K 8, P 20, NN 2, no radius / depth limit, 400 depth bins) on the cfg1 cloud (200 K points, 400x400 view): a
200x200x400 frustum grid rebuilt per call, (a) a 4096-ray training batch, (b) the whole 160 000-pixel frame in one call.
The reference's own kernels (host-compiled, oracle/_ref, one thread) are run on the same inputs: results must be
bit-identical, and their time is the CPU baseline of this row."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import frustum_ref as fr, torch_port as tp  # noqa: E402
from pointnerf_b200 import _lib, query_frustum, synthetic as syn  # noqa: E402


def timed(fn, warm=3, reps=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(reps):
        fn()
    ev[1].record()
    torch.cuda.synchronize()
    return ev[0].elapsed_time(ev[1]) / reps


def main():
    dev = torch.device("cuda", 0)
    opt = syn.config_opt("cfg1", wcoord_query=0, max_o=None, vscale=(2, 2, 1), kernel_size=(3, 3, 3), query_size=(3, 3, 3), SR=40,
                         K=8, P=20, NN=2, radius_limit_scale=0.0, depth_limit_scale=0.0, z_depth_dim=400)
    _, tables = syn.make_scene("cfg1")
    cam = syn.make_camera("cfg1")
    xp = tp.w2pers_points(tables["xyz"], cam["camrotc2w"], cam["campos"])
    xp_d = xp.to(dev).reshape(-1, 3).contiguous()
    q = query_frustum.lighting_fast_querier(dev, opt)
    hp = q.get_hyperparameters(cam["h"], cam["w"], cam["intrinsic"], cam["near"], cam["far"])
    res = {"workload": "frustum querier, DTU feed-forward settings on the cfg1 cloud: %d points, grid %s, SR=40 K=8 P=20 NN=2"
           % (xp_d.shape[0], "x".join(str(int(v)) for v in hp["scaled_vdim"]))}
    for tag, pix in (("batch_4096", syn.random_pixels(cam, 4096)), ("frame_400x400", syn.all_pixels(cam))):
        pix_t = torch.as_tensor(pix, dtype=torch.int32)
        pix_d = pix_t.to(dev).contiguous()
        l0 = _lib.lib().pn_launch_count()
        raw = q.query_raw(pix_d, xp_d, hp, count_candidates=True)
        launches = _lib.lib().pn_launch_count() - l0
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        pidx, sloc, pc, mask = fr.query_grid_point_index(opt, hp, pix, xp.reshape(-1, 3).numpy())
        cpu_s = time.perf_counter() - t0
        sel = mask > 0
        same = (np.array_equal(raw["ray_mask"].cpu().numpy(), mask)
                and np.array_equal(raw["sample_pidx"].cpu().numpy()[sel], pidx)
                and np.array_equal(raw["sample_loc"].cpu().numpy()[sel], sloc))
        ms_raw = timed(lambda: q.query_raw(pix_d, xp_d, hp))
        ms_api = timed(lambda: q.query_points(pix_d[None], xp_d[None], None, None, cam["h"], cam["w"], cam["intrinsic"], cam["near"],
                                              cam["far"], None, cam["campos"].to(dev), cam["camrotc2w"].to(dev)))
        R = int(pix.shape[0])
        res[tag] = {"rays": R, "rays_kept": int(sel.sum()), "valid_samples": int((pidx >= 0).any(-1).sum()), "pairs": int((pidx >= 0).sum()),
                    "candidates": int(raw["candidates"].item()), "bit_identical_to_reference_kernels": bool(same),
                    "gpu_ms_kernels": round(ms_raw, 4), "gpu_ms_query_points": round(ms_api, 4), "launches": int(launches),
                    "rays_per_s": round(R / ms_api * 1e3, 1), "cpu_reference_s": round(cpu_s, 4),
                    "cpu_rays_per_s": round(R / cpu_s, 1), "cpu_kind": "reference kernels host-compiled, 1 thread"}
    # the per-view grid build alone (no rays)
    empty = torch.zeros((0, 2), dtype=torch.int32, device=dev)
    res["grid_build_ms"] = round(timed(lambda: q.query_raw(empty, xp_d, hp)), 4)
    V = int(np.prod(hp["scaled_vdim"]))
    N = int(xp_d.shape[0])
    # points read twice (mark, count) + cell / rank / tmp / bucket lists; cell_start: memset, scan read twice + write; two bitsets
    res["grid_build_algorithmic_bytes"] = 2 * 12 * N + 5 * 4 * N + 4 * 4 * (V + 1) + 2 * (V // 8)
    res["grid_build_GBps"] = round(res["grid_build_algorithmic_bytes"] / res["grid_build_ms"] / 1e6, 1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
