#!/usr/bin/env python
"""Per-stage device time of one full-frame render (cfg2 / cfg5 camera, 800x800, forward only): CUDA events around every
C-ABI call of NeuralPointsRayMarching.render_frame (bench.StageTimer), side stream off so that stages do not overlap.
Usage: python tools/frame_stages.py [workload] [chunk]"""
import copy
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch  # noqa: E402

import bench  # noqa: E402
from pointnerf_b200 import synthetic as syn  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
    chunk = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
    dev = torch.device("cuda", 0)
    opt, tables = syn.make_scene(name)
    opt = copy.copy(opt)
    opt.is_train = 0
    model = bench.make_model(opt, tables, dev)
    cam = syn.make_camera(name, view=1)
    raydir, _ = syn.make_rays(cam, syn.all_pixels(cam))
    rd, campos, rot = raydir.to(dev), cam["campos"].to(dev), cam["camrotc2w"].to(dev)
    bg = torch.rand(int(opt.shading_color_channel_num), device=dev)
    for side in (False, True):
        model.render_frame(campos, rd, bg, rot, cam["near"], cam["far"], chunk=chunk, side_stream=side)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            model.render_frame(campos, rd, bg, rot, cam["near"], cam["far"], chunk=chunk, side_stream=side)
        e1.record()
        torch.cuda.synchronize()
        print(json.dumps({"workload": name, "chunk": chunk, "side_stream": side, "frame_ms": round(e0.elapsed_time(e1) / 3, 3)}))
    with bench.StageTimer() as st:
        model.render_frame(campos, rd, bg, rot, cam["near"], cam["far"], chunk=chunk, side_stream=False)
        tot, cnt = st.totals_ms()
    print(json.dumps({"stages_ms_per_frame": {k: round(v, 3) for k, v in sorted(tot.items())}, "calls": cnt}))


if __name__ == "__main__":
    main()
