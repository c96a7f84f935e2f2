#!/usr/bin/env python
"""Renders ONE 65 536-ray chunk of the cfg2 800x800 frame (forward only) twice: the target of the ncu launch list
profiles/r0x_launches_frame_chunk_*.csv (the query and compositing kernels are not hidden under a backward here).
Usage: ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file raw.csv python tools/one_frame_chunk.py"""
import copy
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch  # noqa: E402

import bench  # noqa: E402
from pointnerf_b200 import synthetic as syn  # noqa: E402

dev = torch.device("cuda", 0)
opt, tables = syn.make_scene("cfg2")
opt = copy.copy(opt)
opt.is_train = 0
model = bench.make_model(opt, tables, dev)
cam = syn.make_camera("cfg2", view=1)
raydir, _ = syn.make_rays(cam, syn.all_pixels(cam))
rd = raydir[:, 287232:287232 + 65536].contiguous().to(dev)        # rows 359..440 of the image: the middle of the object
bg = torch.rand(int(opt.shading_color_channel_num), device=dev)
for _ in range(2):
    model.render_frame(cam["campos"].to(dev), rd, bg, cam["camrotc2w"].to(dev), cam["near"], cam["far"], chunk=65536, side_stream=False)
    torch.cuda.synchronize()
