#!/usr/bin/env python
"""bench.py - throughput of the Point-NeRF per-ray hot path on B200 (BASELINE.json metric: train rays/s fwd+bwd).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg2]

One "step" = one pass of the hot path over one batch of synthetic rays:
    neighbour query (ray sampling + K-NN on the cached voxel grid) -> fused gather + PointAggregator ->
    ray_dist -> ray_march -> colour MSE -> backward to the neural point tables, the MLP and bg_color
    (-> one NCCL all-reduce of the flat gradient buffer when N > 1).
Workload (N = 1 and per GPU for N > 1: weak scaling): BASELINE.json configs[1] - 1M points, 4096 rays/batch,
80 samples/ray, K = 8, on the synthetic lego-shaped scene of pointnerf_b200/synthetic.py.

Prints ONE JSON line (rank 0).  `value` = rays/s with the ray batches resident in HBM; `e2e` = the same step fed
from pinned HOST buffers (H2D of rays/jitter/targets and D2H of the loss inside the timed region).
`--impl reference` times the CPU oracle (the reference's algorithm restated, oracle/) on the host cores.
"""
import argparse
import copy
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np
import torch
import torch.distributed as dist

METRIC = "train_rays_per_s_fwd_bwd"
UNIT = "rays/s"



_REAL_STDOUT = None


def _quiet_stdout():
    """stdout carries the ONE JSON line only: libraries that print there from C (NCCL's version banner) go to stderr."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def _emit(line):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    if _REAL_STDOUT is None:
        os.write(1, data)
    else:
        os.write(_REAL_STDOUT, data)


def _phase(msg):
    """Progress marker on stderr (rank-tagged): a hung multi-GPU run shows where it stopped."""
    sys.stderr.write("[bench rank %s] %s\n" % (os.environ.get("RANK", "0"), msg))
    sys.stderr.flush()


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        d = json.load(open(p))
        return dict(hbm_gbs=float(d["hbm_gbs"]), bf16_tflops=float(d["bf16_tflops"]),
                    bf16_tflops_sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), source="measured")
    return dict(hbm_gbs=6650.0, bf16_tflops=1590.0, bf16_tflops_sustained=1400.0, source="fallback")


# ---------------------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# ---------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.rows, self.proc, self.th = index, [], None, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.th = threading.Thread(target=self._read, daemon=True)
        self.th.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------------------
# workload
# ---------------------------------------------------------------------------------------------------------
def build_workload(name, rays, n_batches, seed_base, train=1):
    """Scene + camera + a pool of ray batches (torch-CPU).  Same scene on every rank; batches differ per rank."""
    from pointnerf_b200 import synthetic as syn
    opt, tables = syn.make_scene(name)
    opt = copy.copy(opt)
    opt.is_train = train
    C = int(opt.shading_color_channel_num)
    batches = []
    for b in range(n_batches):
        cam = syn.make_camera(name, view=b % 4)
        raydir, _ = syn.make_rays(cam, syn.random_pixels(cam, rays, seed=seed_base + b))
        g = torch.Generator().manual_seed(seed_base * 7919 + b)
        uniform = torch.rand(1, rays, opt.z_depth_dim, generator=g) if train else None
        gt = torch.rand(1, rays, C, generator=g)
        batches.append(dict(cam=cam, raydir=raydir, uniform=uniform, gt=gt))
    g = torch.Generator().manual_seed(5)
    bg = torch.rand(C, generator=g)
    return opt, tables, batches, bg


def make_model(opt, tables, dev):
    from pointnerf_b200.neural_points import NeuralPointsRayMarching
    torch.manual_seed(0)
    model = NeuralPointsRayMarching(opt, device=dev)
    with torch.no_grad():
        # live density head (sigma*dist ~ 0.5): transmittance decays along the ray as in a trained scene
        model.aggregator.alpha_branch[0].bias.fill_(60.0)
    model.neural_points.set_points(tables["xyz"], tables["points_embeding"], tables["points_color"], tables["points_dir"],
                                   tables["points_conf"], Rw2c=tables["Rw2c"])
    return model


def to_dev(batch, dev, non_blocking=False):
    cam = batch["cam"]
    return dict(cam=dict(campos=cam["campos"].to(dev, non_blocking=non_blocking), camrotc2w=cam["camrotc2w"].to(dev, non_blocking=non_blocking),
                         near=cam["near"], far=cam["far"]),
                raydir=batch["raydir"].to(dev, non_blocking=non_blocking),
                uniform=batch["uniform"].to(dev, non_blocking=non_blocking) if batch["uniform"] is not None else None,
                gt=batch["gt"].to(dev, non_blocking=non_blocking))


def pack_pinned(batch):
    """One pinned fp32 buffer per batch [campos 3 | camrotc2w 9 | raydir R*3 | gt R*C | jitter draw R*D], segments padded to
    64 floats: the loader's upload is ONE host->device copy per step; ``unpack_dev`` slices the device copy (views only)."""
    parts = [batch["cam"]["campos"], batch["cam"]["camrotc2w"], batch["raydir"], batch["gt"]]
    if batch["uniform"] is not None:
        parts.append(batch["uniform"])
    offs, n = [], 0
    for t_ in parts:
        offs.append(n)
        n += (t_.numel() + 63) // 64 * 64
    buf = torch.zeros(n, dtype=torch.float32).pin_memory()
    for t_, o in zip(parts, offs):
        buf[o:o + t_.numel()].copy_(t_.reshape(-1).float())
    meta = dict(offs=offs, shapes=[tuple(t_.shape) for t_ in parts], near=batch["cam"]["near"], far=batch["cam"]["far"],
                bytes=sum(t_.numel() for t_ in parts) * 4)
    return buf, meta


def unpack_dev(buf, meta):
    v = [buf[o:o + int(np.prod(sh))].view(sh) for o, sh in zip(meta["offs"], meta["shapes"])]
    return dict(cam=dict(campos=v[0], camrotc2w=v[1], near=meta["near"], far=meta["far"]), raydir=v[2], gt=v[3],
                uniform=v[4] if len(v) > 4 else None)


def pin(batch):
    out = dict(batch)
    for k in ("raydir", "uniform", "gt"):
        if out[k] is not None:
            out[k] = out[k].pin_memory()
    return out


# ---------------------------------------------------------------------------------------------------------
# the other BASELINE configurations as extra legs of the same JSON line (north_star: rays/s at 1/2/4/8 GPUs on the
# 3 M-point and 20 M-point scenes; weak = fixed rays per GPU, strong = 65 536 rays per step split over the ranks)
# ---------------------------------------------------------------------------------------------------------
def build_scene_on_device(name, dev, rank, world):
    """Scene for a bench leg.  Rank 0 generates the point positions (the 20 M-point cloud takes ~45 s of host time) and
    broadcasts them; the per-point tables are drawn ON the device from a fixed seed (identical on every rank)."""
    from pointnerf_b200 import synthetic as syn
    opt = copy.copy(syn.config_opt(name))
    opt.is_train = 1
    if rank == 0:
        _, tables = syn.make_scene(name)
        xyz = tables["xyz"].to(dev)
        n = torch.tensor([xyz.shape[0]], dtype=torch.int64, device=dev)
    else:
        xyz, n = None, torch.zeros(1, dtype=torch.int64, device=dev)
    if world > 1:
        dist.broadcast(n, 0)
        if rank != 0:
            xyz = torch.empty((int(n.item()), 3), dtype=torch.float32, device=dev)
        dist.broadcast(xyz, 0)
    N, F = int(xyz.shape[0]), int(opt.point_features_dim)
    g = torch.Generator(device=dev).manual_seed(4321)
    d = torch.randn(1, N, 3, device=dev, generator=g)
    tables = dict(xyz=xyz, points_embeding=torch.rand(1, N, F, device=dev, generator=g) - 0.5,
                  points_conf=torch.rand(1, N, 1, device=dev, generator=g) * 0.9 + 0.1, points_dir=d / d.norm(dim=-1, keepdim=True),
                  points_color=torch.randn(1, N, 3, device=dev, generator=g), Rw2c=torch.eye(3, device=dev))
    return opt, tables


def leg_batches(name, opt, rays, n_batches, seed_base, dev):
    from pointnerf_b200 import synthetic as syn
    C = int(opt.shading_color_channel_num)
    out = []
    for b in range(n_batches):
        cam = syn.make_camera(name, view=1 + b % 3)      # (view 0 of the room scene looks at a corner: few rays meet geometry)
        raydir, _ = syn.make_rays(cam, syn.random_pixels(cam, rays, seed=seed_base + b))
        g = torch.Generator().manual_seed(seed_base * 7919 + b)
        out.append(to_dev(dict(cam=cam, raydir=raydir, uniform=torch.rand(1, rays, opt.z_depth_dim, generator=g), gt=torch.rand(1, rays, C, generator=g)), dev))
    return out


def run_leg(tag, name, model, opt, rays, dev, rank, world, steps, items, wts, sparse, adam=False):
    """One extra configuration through the product's training step (FusedTrainStep, pipelined graphs, touched-row exchange):
    device-timed, max over ranks."""
    from pointnerf_b200.train_step import FusedTrainStep
    n_pool = 4
    batches = leg_batches(name, opt, rays, n_pool, 500 + 1000 * rank, dev)
    C = int(opt.shading_color_channel_num)
    bg = torch.nn.Parameter(torch.rand(C, generator=torch.Generator().manual_seed(5)).to(dev))
    st = FusedTrainStep(model, bg, rays, batches[0]["cam"]["near"], batches[0]["cam"]["far"], color_loss_items=items, color_loss_weights=wts,
                        pipeline=True, sparse_exchange=sparse)
    caps = st.calibrate(batches)
    b0 = batches[0]
    st.prime(b0["cam"], b0["raydir"], b0["gt"], b0["uniform"])
    opts, row_opt = None, None
    if adam:
        from pointnerf_b200.optim import Adam, RowAdam
        nps_ = model.neural_points
        tabs_ = [t for t in (nps_.points_embeding, nps_.points_conf, nps_.points_dir, nps_.points_color) if t is not None and t.requires_grad]
        opts = [Adam(list(model.aggregator.parameters()), lr=5e-4, betas=(0.9, 0.999))]
        if adam == "rows":       # lazy Adam on the rows this step touched (an option: torch.optim.SparseAdam semantics)
            row_opt = RowAdam(tabs_, lr=2e-3, betas=(0.9, 0.999))
            opts.append(Adam([bg], lr=2e-3, betas=(0.9, 0.999)))
        else:                    # the reference's dense Adam over every point row
            opts.append(Adam(tabs_ + [bg], lr=2e-3, betas=(0.9, 0.999)))

    def one(i):
        nb = batches[(i + 1) % n_pool]
        st(nb["cam"], nb["raydir"], nb["gt"], nb["uniform"])
        if world > 1:
            st.reduce()
        if opts:
            for o_ in opts:
                o_.step()
        if row_opt is not None:
            if world > 1 and sparse:
                row_opt.step(st._gathered.view(torch.int32), stride=int(st._gathered.shape[1]))
            else:
                row_opt.step(st._touched_ids)
    for i in range(4):
        one(i)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        one(4 + i)
    e1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / steps
    c = st.counts_host.tolist()
    res = {"workload": "%s: %d points, %d rays/step/GPU x %d GPU(s), SR=%d K=%d" % (name, int(model.neural_points.xyz.shape[0]), rays, world, opt.SR, opt.K),
           "ms_per_step": round(ms, 4), "rays_per_s": round(rays * world / (ms * 1e-3), 1), "steps": steps,
           "valid_samples_last_batch": c[2], "valid_pairs_last_batch": c[3], "capacities": list(caps), "overflows": st.overflows,
           "with_adam": ("dense torch.optim.Adam semantics (every point row, %.1f GB of traffic per step)" % (int(model.neural_points.xyz.shape[0]) * 39 * 28 / 1e9)
                         if adam is True else "lazy: rows touched this step only (torch.optim.SparseAdam semantics)" if adam == "rows" else False)}
    if world > 1:
        res["exchange_bytes_per_rank_per_step"] = int(getattr(st, "exchange_bytes", 0))
        res["exchange"] = "touched rows (all-gather + rank-ordered merge)" if sparse else "dense all-reduce"
    del st
    return res


def extra_legs(dev, rank, world, steps, sparse, only=None):
    items, wts = ("ray_masked_coarse_raycolor", "ray_miss_coarse_raycolor", "coarse_raycolor"), (1.0, 0.0, 0.0)
    out = {}
    for name in ("cfg3", "cfg4"):
        if only and name not in only:
            continue
        _phase("leg scene " + name)
        opt, tables = build_scene_on_device(name, dev, rank, world)
        model = make_model(opt, tables, dev)
        del tables
        if rank == 0:       # grid (re)build of this cloud: what a prune / grow pass pays before the next query
            from pointnerf_b200 import query as pq
            nps_ = model.neural_points
            g0 = nps_.querier.grid_for(nps_.xyz.detach())
            ms_ = []
            for _ in range(4):
                a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a_.record()
                pq.VoxelGrid(nps_.xyz.detach(), g0.hp, opt.P, opt.max_o)
                b_.record()
                torch.cuda.synchronize()
                ms_.append(a_.elapsed_time(b_))
            N_, V_ = int(nps_.xyz.shape[0]), int(g0.V)
            algo = 32 * N_ + 16 * V_ + V_ // 8
            out[name + "_grid_build"] = {"points": N_, "cells": V_, "ms": round(min(ms_[1:]), 4), "algorithmic_bytes": algo,
                                         "GB/s": round(algo / (min(ms_[1:]) * 1e-3) / 1e9, 1),
                                         "note": "pn_grid_build incl. its allocations (kernel time by CUDA events); the reference rebuilds its grid every forward"}
        if name == "cfg3":
            out["cfg3_weak"] = run_leg("cfg3_weak", name, model, opt, 4096, dev, rank, world, steps, items, wts, sparse)
        else:
            out["cfg4_weak"] = run_leg("cfg4_weak", name, model, opt, 8192, dev, rank, world, steps, items, wts, sparse)
            out["cfg4_strong"] = run_leg("cfg4_strong", name, model, opt, 65536 // world, dev, rank, world, max(steps // 2, 4), items, wts, sparse)
            out["cfg4_weak_with_adam"] = run_leg("cfg4_weak_adam", name, model, opt, 8192, dev, rank, world, max(steps // 2, 4), items, wts, sparse, adam=True)
            out["cfg4_weak_with_row_adam"] = run_leg("cfg4_weak_row_adam", name, model, opt, 8192, dev, rank, world, max(steps // 2, 4), items, wts, sparse, adam="rows")
        del model
        torch.cuda.empty_cache()
    return out


# ---------------------------------------------------------------------------------------------------------
# stage timing (CUDA events on the launching stream) -> roofline of the dominant stage
# ---------------------------------------------------------------------------------------------------------
class StageTimer:
    """Wraps the C-ABI entry points with CUDA events (torch's current stream = the stream the kernels run on)."""
    STAGES = {"pn_ray_sample": "ray_sample", "pn_knn_query": "knn_query", "pn_w2pers": "w2pers", "pn_compact_samples": "compact",
              "pn_agg_forward": "agg_forward", "pn_agg_backward": "agg_backward", "pn_ray_dist": "ray_dist",
              "pn_composite_forward": "composite_forward", "pn_composite_backward": "composite_backward"}

    def __init__(self):
        from pointnerf_b200 import _lib
        self.lib = _lib.lib()
        self.orig, self.events = {}, []

    def __enter__(self):
        for name, stage in self.STAGES.items():
            fn = getattr(self.lib, name)
            self.orig[name] = fn

            def wrapped(*a, _fn=fn, _stage=stage):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                rc = _fn(*a)
                e1.record()
                self.events.append((_stage, e0, e1))
                return rc
            setattr(self.lib, name, wrapped)
        return self

    def __exit__(self, *exc):
        for name, fn in self.orig.items():
            setattr(self.lib, name, fn)

    def totals_ms(self):
        torch.cuda.synchronize()
        tot, cnt = {}, {}
        for stage, e0, e1 in self.events:
            tot[stage] = tot.get(stage, 0.0) + e0.elapsed_time(e1)
            cnt[stage] = cnt.get(stage, 0) + 1
        return tot, cnt


def work_counts(model, batch_dev, opt):
    """Algorithmic work of one batch: valid samples S, valid pairs M, kept samples, candidates."""
    nps = model.neural_points
    with torch.no_grad():
        raw = nps.querier.query_raw(nps.xyz.detach(), batch_dev["raydir"], batch_dev["cam"]["campos"], batch_dev["cam"]["near"],
                                    batch_dev["cam"]["far"], uniform=batch_dev["uniform"], count_candidates=True)
        pidx = raw["sample_pidx"]
        M = int((pidx >= 0).sum())
        S = int((pidx >= 0).any(-1).sum())
        Sq = int(raw["hit_count"].sum())
        cand = int(raw["candidates"].item())
        rays_hit = int((raw["ray_any"] > 0).sum())
    return dict(S=S, M=M, kept_samples=Sq, candidates=cand, rays_with_neighbours=rays_hit)


def agg_flops(opt, S, M, backward=True):
    F, W, C = int(opt.point_features_dim), int(opt.shading_feature_num), int(opt.shading_color_channel_num)
    in1 = F + 2 * opt.num_feat_freqs * F + 12 * opt.dist_xyz_freq
    in3, inc = W + 7, W + 6 * opt.num_viewdir_freqs
    per_pair = 2 * (in1 * W + W * W + in3 * W + W * W + W)
    per_sample = 2 * (inc * C + C * C + C * C)
    f = per_pair * M + per_sample * S
    return f * (3 if backward else 1)


# ---------------------------------------------------------------------------------------------------------
# CPU oracle arm (cpu_baseline / --impl reference)
# ---------------------------------------------------------------------------------------------------------
def cpu_oracle_steps(opt, tables, batches, bg, agg_state, rays, steps, warmup):
    """Times the CPU oracle (oracle/pipeline_ref.py: reference querier kernels compiled for the host when
    oracle/_ref is present, else the C port; reference torch aggregator + ray_march restated) fwd+bwd."""
    from oracle import pipeline_ref, querier_ref
    backend = "ref" if querier_ref.have_backend("ref") else "port"
    C = int(opt.shading_color_channel_num)

    def one(b):
        bb = batches[b % len(batches)]
        raydir = bb["raydir"][:, :rays].contiguous()
        gt = bb["gt"][:, :rays]
        uni = bb["uniform"][:, :rays].contiguous() if bb["uniform"] is not None else None
        pipeline_ref.run(opt, tables, bb["cam"], raydir, bg, agg_state, backend=backend, uniform=uni,
                         loss_fn=lambda o: ((o["coarse_raycolor"] - gt) ** 2).sum() / float(rays * C))
    for w in range(warmup):
        one(w)
    t0 = time.perf_counter()
    for s in range(steps):
        one(warmup + s)
    dt = time.perf_counter() - t0
    return rays * steps / dt, dt / steps, backend


# ---------------------------------------------------------------------------------------------------------
# full-frame inference (BASELINE.json metric, second item: ms per 800x800 frame) and PSNR against the oracle
# ---------------------------------------------------------------------------------------------------------
def frame_bench(model, opt, name, bg_dev, dev, rank, world, chunk, frames=3):
    """Forward-only render of the workload's full frame (cfg2/cfg5 camera: 800x800 = 640 000 rays) through
    NeuralPointsRayMarching.render_frame, is_train = 0 (no jitter).  N > 1: each rank renders every N-th pixel; the time is
    the max over ranks.  `ms` has the rays resident in HBM; `e2e_ms` uploads the ray directions from pinned host memory and
    reads the [R, C] image back to pinned host memory inside the timed region."""
    from pointnerf_b200 import synthetic as syn
    cam = syn.make_camera(name, view=1)
    raydir, _ = syn.make_rays(cam, syn.all_pixels(cam))
    R_all = raydir.shape[1]
    # N > 1: the pixels are dealt to the ranks round-robin (pixel p -> rank p mod N): every rank sees the same mix of
    # rays that hit the object and rays that miss it
    ray_host = raydir[:, rank::world].contiguous().pin_memory()
    campos, rot = cam["campos"].to(dev), cam["camrotc2w"].to(dev)
    C = int(opt.shading_color_channel_num)
    img_host = torch.empty((1, ray_host.shape[1], C), dtype=torch.float32).pin_memory()
    was = opt.is_train
    opt.is_train = 0
    try:
        ray_dev = ray_host.to(dev)
        warm = model.render_frame(campos, ray_dev, bg_dev, rot, cam["near"], cam["far"], chunk=chunk, keys=("coarse_raycolor", "ray_mask"))   # warm-up
        frac_hit = float((warm["ray_mask"] > 0).float().mean())
        del warm
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(frames):
            img = model.render_frame(campos, ray_dev, bg_dev, rot, cam["near"], cam["far"], chunk=chunk)["coarse_raycolor"]
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / frames
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for _ in range(frames):
            rd = ray_host.to(dev, non_blocking=True)
            img = model.render_frame(campos, rd, bg_dev, rot, cam["near"], cam["far"], chunk=chunk)["coarse_raycolor"]
            img_host.copy_(img, non_blocking=True)
        f1.record()
        torch.cuda.synchronize()
        e2e_ms = f0.elapsed_time(f1) / frames
    finally:
        opt.is_train = was
    t = torch.tensor([ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms = float(t[0]), float(t[1])
    assert bool(torch.isfinite(img_host).all())
    return {"ms": round(ms, 3), "e2e_ms": round(e2e_ms, 3), "h": cam["h"], "w": cam["w"], "rays": R_all, "SR": int(opt.SR), "K": int(opt.K),
            "chunk_rays": int(chunk), "rays_per_s": round(R_all / (ms * 1e-3), 1), "frac_rays_with_neighbours": round(frac_hit, 4),
            "h2d_bytes": int(ray_host.numel() * 4), "d2h_bytes": int(img_host.numel() * 4),
            "note": "forward only, grid cached, %d GPU(s), pixels dealt round-robin" % world}


def frame_sweep(model, opt, name, bg_dev, dev, chunk):
    """BASELINE.json configs[4]: full-frame inference over K = 8/16 and 40/80/160 samples per ray, with a prune (conf below
    a threshold that removes ~1 % of the points) + grow (1 % new points) between passes.  The maintenance step and the grid
    rebuild it forces are timed separately (wall clock around a device sync: both contain a host read-back)."""
    import warnings
    from pointnerf_b200 import synthetic as syn
    cam = syn.make_camera(name, view=1)
    raydir, _ = syn.make_rays(cam, syn.all_pixels(cam))
    ray_dev = raydir.to(dev)
    campos, rot = cam["campos"].to(dev), cam["camrotc2w"].to(dev)
    nps = model.neural_points
    keep = (opt.is_train, opt.K, opt.SR)
    gen = torch.Generator(device=dev).manual_seed(17)
    out = []
    try:
        opt.is_train = 0
        for K in (8, 16):
            for SR in (40, 80, 160):
                opt.K, opt.SR = K, SR
                with warnings.catch_warnings():
                    warnings.simplefilter("ignore")          # grown points may overfill a voxel beyond P (first P kept)
                    torch.cuda.synchronize()
                    t0 = time.perf_counter()
                    nps.querier.grid_for(nps.xyz.detach())   # rebuild after the previous pass's prune/grow
                    torch.cuda.synchronize()
                    grid_ms = (time.perf_counter() - t0) * 1e3
                model.render_frame(campos, ray_dev, bg_dev, rot, cam["near"], cam["far"], chunk=chunk)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(2):
                    model.render_frame(campos, ray_dev, bg_dev, rot, cam["near"], cam["far"], chunk=chunk)
                e1.record()
                torch.cuda.synchronize()
                N = int(nps.xyz.shape[0])
                n_add = N // 100
                pick = torch.randint(0, N, (n_add,), device=dev, generator=gen)
                add = dict(add_xyz=nps.xyz.detach()[pick] + 0.002 * torch.randn(n_add, 3, device=dev, generator=gen),
                           add_embedding=nps.points_embeding.detach()[0, pick], add_color=nps.points_color.detach()[0, pick],
                           add_dir=nps.points_dir.detach()[0, pick], add_conf=nps.points_conf.detach()[0, pick])
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                kept = nps.prune_and_grow(0.109, **add)
                torch.cuda.synchronize()
                maint_ms = (time.perf_counter() - t0) * 1e3
                out.append({"K": K, "SR": SR, "frame_ms": round(e0.elapsed_time(e1) / 2, 3), "points": N, "grid_build_ms": round(grid_ms, 3),
                            "prune_grow_ms": round(maint_ms, 3), "pruned": N - kept, "grown": n_add})
    finally:
        opt.is_train, opt.K, opt.SR = keep
    return out


def maintenance_bench(model, opt, batch_dev, bg_dev, dev):
    """The between-pass kernels of SURVEY.md 8(f) on the workload's cloud: voxel down-sampling (construct_vox_points_closest,
    wall clock: it ends in a host read-back of the voxel count) and the probe outputs of one batch (pn_probe_rays)."""
    from pointnerf_b200.voxelize import construct_vox_points_closest
    nps = model.neural_points
    xyz = nps.xyz.detach()
    N = int(xyz.shape[0])
    out = {}
    for res in (320, 1000):
        construct_vox_points_closest(xyz, res)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(3):
            c, _, _ = construct_vox_points_closest(xyz, res)
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) * 1e3 / 3
        passes = (3 * int(np.ceil(np.log2(res * 1.05 + 2))) + 7) // 8
        algo = N * (12 + passes * 48 + 24) + 36 * int(c.shape[0])
        out["voxel_downsample_res%d" % res] = {"ms": round(ms, 3), "points": N, "voxels": int(c.shape[0]), "radix_passes": passes,
                                               "GB/s": round(algo / (ms * 1e-3) / 1e9, 1),
                                               "note": "wall clock of construct_vox_points_closest: bbox reductions, 2 host read-backs, %d launches" % (6 + 4 * passes)}
    keep = (opt.prob, opt.is_train)
    try:
        opt.prob, opt.is_train = 1, 0
        with torch.no_grad():
            o = model(batch_dev["cam"]["campos"], batch_dev["raydir"], bg_color=bg_dev, camrotc2w=batch_dev["cam"]["camrotc2w"],
                      near=batch_dev["cam"]["near"], far=batch_dev["cam"]["far"])
            plan = model.sample(batch_dev["cam"]["campos"], batch_dev["raydir"], batch_dev["cam"]["camrotc2w"], batch_dev["cam"]["near"],
                                batch_dev["cam"]["far"])
            args = (o["coarse_point_opacity"][0], plan.raw["ray_any"], plan.loc_w, plan.pidx, o["weight"][0])
            model.probe_outputs(*args)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                model.probe_outputs(*args)
            e1.record()
            torch.cuda.synchronize()
        R, SR, K = (int(v) for v in plan.pidx.shape)
        F_ = int(opt.point_features_dim)
        hit = int((plan.raw["ray_any"] > 0).sum())
        algo = 4 * R * SR + hit * (16 + K * (8 + 4 * (F_ + 10))) + 4 * R * (F_ + 12)
        us = e0.elapsed_time(e1) / 20 * 1e3
        out["probe_rays"] = {"us": round(us, 2), "rays": R, "rays_with_neighbours": hit, "GB/s": round(algo / (us * 1e-6) / 1e9, 1),
                             "note": "per call of NeuralPointsRayMarching.probe_outputs (8 output allocations + one launch): host-bound at this ray count"}
    finally:
        opt.prob, opt.is_train = keep
    return out


def frustum_bench(dev):
    """the perspective-frustum querier (wcoord_query = 0) on the DTU feed-forward settings
    (dev_scripts/ete/dtu_dgt_d012_img0123_conf_color_dir_agg2.sh: vscale 2 2 1, kernel 3 3 3, SR 40, K 8, P 20, NN 2, no limits,
    400 depth bins) over the cfg1 cloud (200 K points, 400x400 view): a 200x200x400 frustum grid rebuilt per call, (a) a
    4096-ray batch, (b) the whole 160 000-pixel frame in one call.  The reference's own kernels (host-compiled, oracle/_ref,
    one thread) run on the same inputs as the checker (results must be bit-identical) and as this row's CPU baseline."""
    from oracle import frustum_ref as fr, torch_port as tp
    from pointnerf_b200 import _lib, query_frustum, synthetic as syn

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
    have_ref = fr.available()
    for tag, pix in (("batch_4096", syn.random_pixels(cam, 4096)), ("frame_400x400", syn.all_pixels(cam))):
        pix_d = torch.as_tensor(pix, dtype=torch.int32).to(dev).contiguous()
        l0 = _lib.lib().pn_launch_count()
        raw = q.query_raw(pix_d, xp_d, hp, count_candidates=True)
        launches = _lib.lib().pn_launch_count() - l0
        torch.cuda.synchronize()
        R = int(pix.shape[0])
        entry = {"rays": R, "rays_kept": int((raw["ray_mask"] > 0).sum()), "pairs": int((raw["sample_pidx"] >= 0).sum()),
                 "candidates": int(raw["candidates"].item()), "launches": int(launches)}
        if have_ref:
            t0 = time.perf_counter()
            pidx, sloc, _, mask = fr.query_grid_point_index(opt, hp, pix, xp.reshape(-1, 3).numpy())
            cpu_s = time.perf_counter() - t0
            sel = mask > 0
            entry.update(bit_identical_to_reference_kernels=bool(
                np.array_equal(raw["ray_mask"].cpu().numpy(), mask) and np.array_equal(raw["sample_pidx"].cpu().numpy()[sel], pidx)
                and np.array_equal(raw["sample_loc"].cpu().numpy()[sel], sloc)),
                cpu_reference_s=round(cpu_s, 4), cpu_rays_per_s=round(R / cpu_s, 1),
                cpu_kind="reference kernels host-compiled (oracle/_ref), 1 thread")
        ms_raw = timed(lambda: q.query_raw(pix_d, xp_d, hp))
        ms_api = timed(lambda: q.query_points(pix_d[None], xp_d[None], None, None, cam["h"], cam["w"], cam["intrinsic"], cam["near"],
                                              cam["far"], None, cam["campos"].to(dev), cam["camrotc2w"].to(dev)))
        entry.update(gpu_ms_kernels=round(ms_raw, 4), gpu_ms_query_points=round(ms_api, 4), rays_per_s=round(R / ms_api * 1e3, 1))
        res[tag] = entry
    # one training step on this querier: FusedTrainStep (two CUDA graphs; the per-view grid build sits in the query graph)
    try:
        import copy
        from pointnerf_b200.neural_points import NeuralPointsRayMarching
        from pointnerf_b200.train_step import FusedTrainStep
        topt = copy.copy(opt)
        topt.is_train, topt.shpnt_jitter = 1, "uniform"
        torch.manual_seed(0)
        model = NeuralPointsRayMarching(topt, device=dev)
        model.neural_points.set_points(tables["xyz"], tables["points_embeding"], tables["points_color"], tables["points_dir"],
                                       tables["points_conf"], Rw2c=tables["Rw2c"])
        R, C = 4096, int(topt.shading_color_channel_num)
        g = torch.Generator().manual_seed(11)
        bg = torch.nn.Parameter(torch.rand(C, generator=g).to(dev))
        camd = dict(campos=cam["campos"].to(dev), camrotc2w=cam["camrotc2w"].to(dev), near=cam["near"], far=cam["far"])
        pool = [dict(cam=camd, pixel_idx=torch.as_tensor(syn.random_pixels(cam, R, seed=50 + i), dtype=torch.int32).to(dev),
                     gt=torch.rand(1, R, C, generator=g).to(dev), uniform=torch.rand(1, R, int(topt.SR), generator=g).to(dev)) for i in range(4)]
        st = FusedTrainStep(model, bg, R, cam["near"], cam["far"], view=dict(h=cam["h"], w=cam["w"], intrinsic=cam["intrinsic"]))
        st.calibrate(pool)
        it = [0]

        def one():
            b = pool[it[0] % len(pool)]
            it[0] += 1
            st(b["cam"], None, b["gt"], b["uniform"], pixel_idx=b["pixel_idx"])
        ms = timed(one, reps=20)
        res["train_step_4096"] = {"ms_per_step": round(ms, 4), "rays_per_s": round(R / ms * 1e3, 1), "caps": list(st.caps), "overflows": st.overflows,
                                  "note": "fwd + bwd of 4096 pixel-indexed rays, uniform depth jitter, grid of the view rebuilt every step"}
    except Exception as e:      # a leg, not the headline: report and go on
        res["train_step_4096"] = {"error": repr(e)[:300]}
    empty = torch.zeros((0, 2), dtype=torch.int32, device=dev)
    res["grid_build_ms"] = round(timed(lambda: q.query_raw(empty, xp_d, hp)), 4)
    V, N = int(np.prod(hp["scaled_vdim"])), int(xp_d.shape[0])
    # points read twice (mark, count) + cell / rank / tmp / bucket lists; cell_start: memset, scan read twice + write; two bitsets
    res["grid_build_algorithmic_bytes"] = 2 * 12 * N + 5 * 4 * N + 4 * 4 * (V + 1) + 2 * (V // 8)
    res["grid_build_GBps"] = round(res["grid_build_algorithmic_bytes"] / res["grid_build_ms"] / 1e6, 1)
    return res


def reference_gpu_bench(model, opt, batch_dev, dev, reps=5):
    """The reference's own querier kernels (kernel string -> sm_100a cubin, driver API: oracle/ref_gpu.py) beside the
    product's query on the same batch, stage by stage (CUDA events, mean of `reps` after one warm-up).  The reference
    rebuilds its grid on every forward (:626) and materialises the [R, D, 3] candidate positions; both are in its column."""
    import ctypes
    from oracle import ref_gpu
    from pointnerf_b200 import query as pq
    if not ref_gpu.available(int(opt.K)):
        return {"unavailable": "oracle/_ref/ref_querier_sm100a_K%d.cubin not built" % int(opt.K)}
    nps = model.neural_points
    xyz = nps.xyz.detach()
    q = nps.querier
    raydir, campos, near, far, uni = batch_dev["raydir"], batch_dev["cam"]["campos"], batch_dev["cam"]["near"], batch_dev["cam"]["far"], batch_dev["uniform"]
    grid = q.grid_for(xyz)
    hp = grid.hp
    n_layers = (int(opt.kernel_size[0]) + 1) // 2

    def ev():
        return torch.cuda.Event(enable_timing=True)
    ours = {"grid_build_ms": 0.0, "depths_ms": 0.0, "ray_sample_ms": 0.0, "knn_query_ms": 0.0}
    for it in range(reps + 1):
        e = [ev() for _ in range(5)]
        e[0].record()
        g2 = pq.VoxelGrid(xyz, hp, opt.P, opt.max_o)
        e[1].record()
        ts = pq.middle_point_ts(raydir, opt.z_depth_dim, near, far, jitter=0.3 if opt.is_train > 0 else 0.0, uniform=uni)
        e[2].record()
        loc_w, hit = pq.ray_sample(g2, campos[0].contiguous(), raydir[0].contiguous(), ts.reshape(-1, ts.shape[-1]).contiguous(), opt.SR)
        e[3].record()
        pidx, ray_any = pq.knn_query(g2, xyz, loc_w, hit, opt.K, n_layers, float(hp["radius_limit2"]))
        e[4].record()
        torch.cuda.synchronize()
        if it:
            for k, a, b in (("grid_build_ms", 0, 1), ("depths_ms", 1, 2), ("ray_sample_ms", 2, 3), ("knn_query_ms", 3, 4)):
                ours[k] += e[a].elapsed_time(e[b]) / reps
    ref = ref_gpu.ReferenceGpuQuerier(opt, hp, dev)
    xyz1 = xyz[None].contiguous()
    rt = {}
    raypos_ms = 0.0
    for it in range(reps + 1):
        a, b = ev(), ev()
        a.record()
        raypos = (campos[:, None, None, :] + raydir[:, :, None, :] * ts[:, :, :, None]).contiguous()      # diff_ray_marching.py:387
        b.record()
        t = {} if it == 0 else rt
        pidx_r, loc_r, mask_r = ref.query(xyz1, raypos, timing=t)
        if it:
            raypos_ms += a.elapsed_time(b) / reps
    rt = {k: v / reps for k, v in rt.items()}
    rt["raypos_ms"] = raypos_ms
    # agreement (sets: the reference's bucket order and slot-0 voxel depend on the hardware's atomic order)
    both = (ray_any > 0) & (mask_r[0] > 0)
    rows = (torch.cumsum((mask_r[0] > 0).long(), 0) - 1)[both]
    so, sr = torch.sort(pidx[both], dim=-1)[0], torch.sort(pidx_r[0][rows], dim=-1)[0]
    samp = (so >= 0).any(-1) | (sr >= 0).any(-1)
    same = (so == sr).all(-1)
    o_total, r_total = sum(ours.values()), sum(rt.values())
    return {"rays": int(raydir.shape[1]), "points": int(xyz.shape[0]),
            "ours_ms": {k: round(v, 4) for k, v in ours.items()}, "ours_total_ms": round(o_total, 4),
            "ours_per_step_ms_grid_cached": round(o_total - ours["grid_build_ms"], 4),
            "reference_kernels_ms": {k: round(v, 4) for k, v in rt.items()}, "reference_total_ms": round(r_total, 4),
            "speedup_whole_query_reference_rebuilds_grid_each_forward": round(r_total / max(o_total - ours["grid_build_ms"], 1e-9), 2),
            "speedup_like_for_like_with_grid_build": round(r_total / max(o_total, 1e-9), 2),
            "knn_kernel_speedup": round(rt["knn_query_ms"] / max(ours["knn_query_ms"], 1e-9), 2),
            "ray_mask_agreement": round(float(((ray_any > 0) == (mask_r[0] > 0)).float().mean()), 6),
            "sample_positions_identical": bool(torch.equal(loc_w[both], loc_r[0][rows])),
            "neighbour_sets_identical_frac": round(float(same[samp].float().mean()), 6),
            "note": "reference = its own kernel string compiled by nvcc for sm_100a and launched with its geometry (%d threads per block) through the driver API; "
                    "sets, not slot order, are compared: its atomics resolve in hardware order" % ref.thr}


def psnr_vs_oracle(dev, name="cfg1", chunk=14400):
    """PSNR delta (BASELINE.json metric, third item) on a full frame of BASELINE configs[0] (`cfg1`: 200 K points, 400x400
    view, 40 samples per ray, K = 8 - the configuration whose CPU path BASELINE.json names): our render against the CPU
    oracle's render of the same scene, and both against a synthetic ground truth (the oracle's render of the scene with
    perturbed embeddings).  The oracle's frame time is configs[0]'s own number (forward render on the host cores, chunked
    like run/train_ft.py:260-261).  Part of the cpu_baseline leg (the oracle is the checker here, never the thing shipped)."""
    from oracle import pipeline_ref, querier_ref
    from pointnerf_b200 import synthetic as syn
    opt, tables = syn.make_scene(name)
    opt = copy.copy(opt)
    opt.is_train = 0
    cam = syn.make_camera(name)
    raydir, _ = syn.make_rays(cam, syn.all_pixels(cam))
    model = make_model(opt, tables, dev)
    g = torch.Generator().manual_seed(5)
    bg = torch.rand(int(opt.shading_color_channel_num), generator=g)
    sd = {k: v.detach().cpu() for k, v in model.aggregator.state_dict().items()}
    backend = "ref" if querier_ref.have_backend("ref") else "port"
    ours = model.render_frame(cam["campos"].to(dev), raydir.to(dev), bg.to(dev), cam["camrotc2w"].to(dev), cam["near"], cam["far"],
                              chunk=65536)["coarse_raycolor"].cpu()

    def oracle_frame(tb):
        t0 = time.perf_counter()
        parts = [pipeline_ref.run(opt, tb, cam, raydir[:, s:s + chunk], bg, sd, need_grads=False, backend=backend)["coarse_raycolor"]
                 for s in range(0, raydir.shape[1], chunk)]
        return torch.cat(parts, dim=1), time.perf_counter() - t0
    ref, cpu_s = oracle_frame(tables)
    t2 = dict(tables)
    t2["points_embeding"] = tables["points_embeding"] + 0.05 * torch.randn(tables["points_embeding"].shape, generator=g)
    gt, _ = oracle_frame(t2)

    def psnr(a, b):
        mse = float(((a.double() - b.double()) ** 2).mean())
        peak = float(gt.abs().max())
        return 10.0 * np.log10(peak * peak / max(mse, 1e-300))
    po, pr = psnr(ours, gt), psnr(ref, gt)
    del model
    return {"scene": "%s (BASELINE configs[0]): %d points, full %dx%d frame, SR=%d K=%d, C=%d feature channels, peak = max|gt|"
                     % (name, tables["xyz"].shape[0], cam["h"], cam["w"], opt.SR, opt.K, ref.shape[-1]),
            "ours_vs_oracle_db": round(psnr(ours, ref), 2), "ours_vs_gt_db": round(po, 4), "oracle_vs_gt_db": round(pr, 4),
            "delta_db": round(abs(po - pr), 6), "max_abs_diff": float((ours - ref).abs().max()),
            "max_rel_to_scale": float((ours - ref).abs().max() / ref.abs().max()),
            "cpu_forward_s_per_frame": round(cpu_s, 2), "cpu_cores": torch.get_num_threads(),
            "cpu_kind": "reference querier kernels (host build) + port of the torch aggregator / ray_march, %d-ray chunks" % chunk}


# ---------------------------------------------------------------------------------------------------------
def dominant_kernel_roofline(_lib, dev, M, peak, peaks, N=256, K=256, reps=20):
    """gemm_nt_img2_kernel at [M,K] x [N,K]^T through the C ABI's GEMM hook, CUDA events on the launching stream.
    ALGORITHMIC flops = 2*M*N*K (the kernel issues 3x that as TF32 MMAs: hi*lo, lo*hi, hi*hi); traffic = DRAM bytes of one
    launch from the committed ncu --set full capture (profiles/r02_ncu_full_gemm_nt_final.json), scaled by M."""
    lib = _lib.lib()
    A = torch.randn(M, K, device=dev)
    B = torch.randn(N, K, device=dev) / 16
    C = torch.empty(M, N, device=dev)
    bias = torch.randn(N, device=dev)
    img = torch.empty(lib.pn_gemm_nt_image_bytes(N, K), dtype=torch.uint8, device=dev)

    def fn():
        _lib.check(lib.pn_gemm_nt(_lib.ptr(A), K, _lib.ptr(B), K, _lib.ptr(C), N, M, None, N, K, _lib.ptr(bias), 1, 0.01, _lib.ptr(img),
                                  _lib.stream()), "pn_gemm_nt")
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / reps * 1e3                # includes the ~3 us weight-image pack the hook launches before each GEMM
    fl = 2.0 * M * N * K
    ach = fl / (us * 1e-6) / 1e12
    traffic = None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "r02_ncu_full_gemm_nt_final.json")))["metrics"]
        traffic = int((float(tr["dram__bytes_read.sum [Mbyte]"]) + float(tr["dram__bytes_write.sum [Mbyte]"])) * 1e6 * (M / 134826))
    except Exception:
        pass
    return {"kernel": "gemm_nt_img2_kernel (aggregator MLP forward / data-gradient GEMM, CTA pairs, 3xTF32)", "bound": "tensor",
            "achieved": round(ach, 3), "peak": peak, "unit": "TFLOP/s", "frac": round(ach / peak, 5), "traffic": traffic,
            "peak_source": "MEASURED_PEAKS.json bf16_tflops (burst: the kernel is timed alone) (%s); 3xTF32 ceiling = peak / 6" % peaks["source"],
            "traffic_source": "profiles/r02_ncu_full_gemm_nt_final.json (one ncu --set full capture of this kernel, this build, at M = 134 826), scaled by M: not a measurement of this run",
            "frac_of_3xtf32_ceiling": round(ach / (peak / 6.0), 4),
            "launch_us": round(us, 2), "flops_per_launch": fl, "algorithmic_bytes_per_launch": int(4 * (M * K + M * N + 2 * N * K)),
            "shape": [M, N, K], "mma_passes_per_product": 3}


def main():
    ap = argparse.ArgumentParser()
    _quiet_stdout()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-prefetch", dest="prefetch", action="store_false",
                    help="do not queue the next batch's sampling/query between this batch's forward and backward")
    ap.add_argument("--workload", default="cfg2")
    ap.add_argument("--rays", type=int, default=4096, help="rays per batch per GPU")
    ap.add_argument("--cpu-rays", type=int, default=0, help="rays per CPU step of the reference arm (0 = whole batches: the same workload as our arm)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-stages", action="store_true")
    ap.add_argument("--no-frame", action="store_true", help="skip the full-frame inference timing (ms per 800x800 frame)")
    ap.add_argument("--frame-chunk", type=int, default=65536, help="rays per render_frame chunk")
    ap.add_argument("--no-sweep", action="store_true", help="skip the cfg5 inference sweep (K x SR, prune/grow between passes)")
    ap.add_argument("--bwd-schedule", type=int, default=None, help="pn_set_backward_schedule: 0 one stream, 1 interleaved, 2 chain first")
    ap.add_argument("--no-adam", action="store_true", help="skip the secondary timing of the step WITH the Adam updates")
    ap.add_argument("--dense-allreduce", action="store_true", help="all-reduce the full N x 39 point-table gradients instead of the touched rows")
    ap.add_argument("--legs", default="auto", help="extra configurations in the same line: 'auto' = 'all' (cfg3 + cfg4 legs), 'none', or a comma list of cfg3,cfg4")
    ap.add_argument("--eager", action="store_true", help="drive the step through autograd from Python (RayShardedStep) instead of the captured CUDA graphs (FusedTrainStep)")
    ap.add_argument("--loss", default="lego", choices=["lego", "plain"],
                    help="colour loss: 'lego' = the shipped per-scene items ray_masked / ray_miss / coarse_raycolor with weights 1 0 0 (dev_scripts/w_n360/lego.sh:152-153); 'plain' = MSE over all rays")
    args = ap.parse_args()

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    n_pool = 8

    # ------------------------------------------------------------------ reference arm: CPU oracle on rank 0
    if args.impl == "reference":
        if rank != 0:
            return 0
        torch.set_num_threads(os.cpu_count() or 1)
        opt, tables, batches, bg = build_workload(args.workload, args.rays, 4, seed_base=100)
        from oracle import torch_port as tp
        torch.manual_seed(0)
        agg = tp.AggregatorPort(opt)
        with torch.no_grad():
            agg.alpha_branch[0].bias.fill_(60.0)
        cpu_rays = args.cpu_rays if args.cpu_rays > 0 else args.rays
        v, spstep, backend = cpu_oracle_steps(opt, tables, batches, bg, agg.state_dict(), cpu_rays, args.steps, args.warmup)
        sample = ("%s, fwd+bwd, %d steps; querier = %s, aggregator / ray_march = torch restatement of the reference modules (pinned to them bit for bit, tests/test_oracle.py)"
                  % ("whole %d-ray batches of the workload" % args.rays if cpu_rays == args.rays else "%d of the %d rays of each batch per step" % (cpu_rays, args.rays),
                     args.steps, "reference kernels compiled for the host (oracle/_ref)" if backend == "ref" else "C port"))
        line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": spstep * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic",
                "config": {"workload": "%s: %d-point neural cloud, %d rays/batch, SR=%d, K=%d, D=%d, fwd+bwd" %
                           (args.workload, tables["xyz"].shape[0], args.rays, opt.SR, opt.K, opt.z_depth_dim)},
                "cpu_baseline": {"value": v, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port" if backend == "port" else "reference",
                                 "kind_detail": "reference querier kernels (host build of its own CUDA string) + port of its torch aggregator / ray_march" if backend == "ref" else "port",
                                 "sample": sample},
                "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        _emit(line)
        return 0

    # ------------------------------------------------------------------ our arm
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    from pointnerf_b200 import _lib
    from pointnerf_b200.ray_shard import RayShardedStep
    peaks = load_peaks()
    if args.bwd_schedule is not None:
        _lib.check(_lib.lib().pn_set_backward_schedule(args.bwd_schedule), "pn_set_backward_schedule")

    opt, tables, batches, bg = build_workload(args.workload, args.rays, n_pool, seed_base=100 + 1000 * rank)
    model = make_model(opt, tables, dev)
    bg_p = torch.nn.Parameter(bg.to(dev))
    stepper = RayShardedStep(model, bg_p, sparse=not args.dense_allreduce)     # autograd-driven step: --eager, and the per-stage timing leg
    dev_batches = [to_dev(b, dev) for b in batches]
    global_rays = args.rays * world
    packed = [pack_pinned(b) for b in batches]
    fstep = None
    if not args.eager:
        # the product's training step: two CUDA graphs per step, counts on the device (pointnerf_b200/train_step.py)
        from pointnerf_b200.train_step import FusedTrainStep
        items = ("ray_masked_coarse_raycolor", "ray_miss_coarse_raycolor", "coarse_raycolor") if args.loss == "lego" else ("coarse_raycolor",)
        wts = (1.0, 0.0, 0.0) if args.loss == "lego" else (1.0,)
        fstep = FusedTrainStep(model, bg_p, args.rays, batches[0]["cam"]["near"], batches[0]["cam"]["far"], color_loss_items=items,
                               color_loss_weights=wts, pipeline=args.prefetch, sparse_exchange=not args.dense_allreduce)
        fstep.calibrate(dev_batches)
        flat_dev = [hb.to(dev) for hb, _ in packed]                    # inputs resident in HBM for the headline number
        assert flat_dev[0].numel() == fstep.inbuf.numel()
        if args.prefetch:
            fstep.prime(packed=flat_dev[0])
    shift = 1 if (fstep is not None and args.prefetch) else 0          # pipelined: step i is handed batch i + 1 (queried beside its backward)

    def step(i):
        if fstep is not None:
            loss = fstep(packed=flat_dev[(i + shift) % n_pool])        # one device-to-device copy of a batch, one or two graph launches
            if world > 1:
                fstep.reduce()
            return loss
        # the training loop hands the NEXT batch along: its sampling / neighbour query is queued between this batch's
        # forward and backward (RayShardedStep.prefetch), so every step still runs one full query + forward + backward
        b, nb = dev_batches[i % n_pool], (dev_batches[(i + 1) % n_pool] if args.prefetch else None)
        return stepper(b["cam"], b["raydir"], b["gt"], uniform=b["uniform"], global_rays=global_rays, next_batch=nb)

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    _phase("model built; warm-up")
    for w in range(3 * n_pool):      # every batch of the pool a few times: the caching allocator's pools (two streams) reach steady state
        step(w)
    for w in range(warmup):
        step(w)
    sync_all()
    _phase("timed loop")
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    _lib.reset_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    th0 = time.perf_counter()
    for s in range(args.steps):
        step(warmup + s)
    host_ms = (time.perf_counter() - th0) * 1e3 / args.steps      # host time to QUEUE a step (its only wait: the (S, M) counts)
    e1.record()
    sync_all()
    launches = _lib.launch_count()
    if fstep is not None:        # replays do not pass through the library's launch counter: kernels per replay (counted at capture) x steps
        launches += fstep.launches_per_step * args.steps
    ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    clocks = sampler.stop() if rank == 0 else None
    value = global_rays * args.steps / (ms_total * 1e-3)

    # ---- end to end: pinned host buffers -> H2D -> step -> D2H of the loss, every step
    _phase("timed loop done (%.3f ms/step); e2e loop" % (ms_total / args.steps))
    h2d = packed[0][1]["bytes"]

    e2e_next, e2e_pending = {}, []
    copy_stream = torch.cuda.Stream()
    loss_host = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(2)]
    e2e_losses, e2e_wait = [], [0.0]

    def upload(i):
        hbuf, meta = packed[i % n_pool]
        return unpack_dev(hbuf.to(dev, non_blocking=True), meta)

    stage = [torch.empty_like(fstep.inbuf) for _ in range(2)] if fstep is not None else None
    stage_ev = [None, None]

    def e2e_step_graph(i):
        # H2D of batch i + 1 from pinned host memory into a staging buffer on the copy stream (under this step's kernels);
        # this step copies ITS staging buffer into the graphs' static inputs (device to device) and replays; the loss goes
        # to pinned host memory every step and is read by the host one step later
        main = torch.cuda.current_stream()
        if stage_ev[i % 2] is None:
            with torch.cuda.stream(copy_stream):
                stage[i % 2].copy_(packed[(i + shift) % n_pool][0], non_blocking=True)
                stage_ev[i % 2] = torch.cuda.Event()
                stage_ev[i % 2].record(copy_stream)
        main.wait_event(stage_ev[i % 2])
        stage_ev[i % 2] = None
        fstep.load_packed(stage[i % 2])
        done = torch.cuda.Event()
        done.record(main)
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(done)                 # the other staging buffer was consumed by the previous step (stream order)
            j = (i + 1) % 2
            stage[j].copy_(packed[(i + 1 + shift) % n_pool][0], non_blocking=True)
            stage_ev[j] = torch.cuda.Event()
            stage_ev[j].record(copy_stream)
        loss = fstep()
        if world > 1:
            fstep.reduce()
        buf = loss_host[i % 2]
        buf.copy_(loss.reshape(1), non_blocking=True)          # D2H read of the step's result
        ev = torch.cuda.Event()
        ev.record()
        e2e_pending.append((ev, buf))
        if len(e2e_pending) > 1:
            e2e_drain(1)

    def e2e_step(i):
        if fstep is not None:
            return e2e_step_graph(i)
        # step i uses the device copy made (from pinned host memory, inside the timed region) during step i - 1 when
        # prefetching; the H2D of batch i + 1 is issued before this step's kernels so that its sampling can be queued.
        # The loss of every step is copied to pinned host memory inside the step and read by the host one step later
        # (asynchronous logging): the host never drains the GPU queue in the middle of the loop.
        b = e2e_next.pop(i, None) or upload(i)
        nb = None
        if args.prefetch:
            # upload of batch i + 1 on a side stream (overlaps this batch's kernels); the stepper waits for `ready`
            # right before it queues that batch's sampling
            main = torch.cuda.current_stream()
            with torch.cuda.stream(copy_stream):
                nb = upload(i + 1)
                nb["ready"] = torch.cuda.Event()
                nb["ready"].record(copy_stream)
            nb["raydir"].record_stream(main)         # all fields are views of the one uploaded buffer
            e2e_next[i + 1] = nb
        loss = stepper(b["cam"], b["raydir"], b["gt"], uniform=b["uniform"], global_rays=global_rays, next_batch=nb)
        buf = loss_host[i % 2]
        buf.copy_(loss.reshape(1), non_blocking=True)          # D2H read of the step's result
        ev = torch.cuda.Event()
        ev.record()
        e2e_pending.append((ev, buf))
        if len(e2e_pending) > 1:
            e2e_drain(1)

    def e2e_drain(keep):
        while len(e2e_pending) > keep:
            ev, buf = e2e_pending.pop(0)
            tw = time.perf_counter()
            ev.synchronize()
            e2e_wait[0] += time.perf_counter() - tw
            e2e_losses.append(float(buf[0]))

    E2E_WARM = 3
    for w in range(E2E_WARM):
        e2e_step(w)
    e2e_drain(0)
    sync_all()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    e2e_wait[0] = 0.0
    th0 = time.perf_counter()
    for s in range(args.steps):
        e2e_step(E2E_WARM + s)
    e2e_drain(0)                 # the last step's loss is on the host before the timer stops
    e2e_host_ms = (time.perf_counter() - th0) * 1e3 / args.steps
    e2e_wait_ms = e2e_wait[0] * 1e3 / args.steps
    f1.record()
    sync_all()
    assert len(e2e_losses) == args.steps + E2E_WARM and all(np.isfinite(v) for v in e2e_losses)
    t = torch.tensor([f0.elapsed_time(f1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = global_rays * args.steps / (float(t.item()) * 1e-3)

    # ---- multi-GPU: gradient equality of the sharded step, and the 3 M / 20 M-point configurations
    grad_check, legs = None, None
    if world > 1 and fstep is not None:
        _phase("sharded gradient check")
        from pointnerf_b200.train_step import sharded_gradient_check
        chk = to_dev(build_workload(args.workload, args.rays, 1, seed_base=77)[2][0], dev)     # the same batch on every rank
        grad_check = sharded_gradient_check(model, bg_p, chk, items, wts)
        fstep.invalidate()                 # the check's steps re-bound the parameters' .grad to their own buffers
        if args.prefetch:
            fstep.prime(packed=flat_dev[0])
    want_legs = "all" if args.legs == "auto" else args.legs
    if want_legs != "none" and fstep is not None:
        legs = extra_legs(dev, rank, world, max(10, min(args.steps, 20)), not args.dense_allreduce,
                          only=None if want_legs == "all" else want_legs.split(","))

    # ---- full-frame inference (all ranks: row blocks)
    _phase("e2e done; frame render")
    frame = None
    if not args.no_frame:
        frame = frame_bench(model, opt, args.workload, bg_p.detach(), dev, rank, world, args.frame_chunk)

    # ---- the same step followed by the optimiser updates (secondary number; the headline metric is fwd+bwd as BASELINE.json
    #      words it): the reference's two Adam optimisers (MLP lr, point tables + bg plr; mvs_points_volumetric_model.py:47-125)
    #      as one multi-tensor launch each (pn_adam_step)
    _phase("step + Adam")
    with_adam = None
    if not args.no_adam:
        from pointnerf_b200.optim import Adam
        nps_ = model.neural_points
        opt_mlp = Adam(list(model.aggregator.parameters()), lr=5e-4, betas=(0.9, 0.999))
        opt_pts = Adam([t for t in (nps_.points_embeding, nps_.points_conf, nps_.points_dir, nps_.points_color) if t is not None and t.requires_grad] + [bg_p],
                       lr=2e-3, betas=(0.9, 0.999))

        def step_adam(i):
            step(i)
            opt_mlp.step()
            opt_pts.step()
        for w in range(3):
            step_adam(w)
        sync_all()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        for s in range(args.steps):
            step_adam(3 + s)
        a1.record()
        sync_all()
        t = torch.tensor([a0.elapsed_time(a1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        n_el = sum(p.numel() for g_ in (opt_mlp.param_groups + opt_pts.param_groups) for p in g_["params"])
        with_adam = {"ms_per_step": round(float(t.item()) / args.steps, 4), "rays_per_s": round(global_rays * args.steps / (float(t.item()) * 1e-3), 1),
                     "adam_elements": int(n_el), "adam_bytes_per_step": int(n_el * 28),
                     "note": "dense Adam over every point row and MLP weight each step (exact torch.optim.Adam semantics), 2 launches"}

    # ---- per-stage device time (events around each C-ABI call) and the roofline of the dominant stage
    _phase("stages / roofline / cpu baseline (rank 0 only, no collectives)")
    stages, roof, counts = None, None, None
    if rank == 0 and not args.no_stages:
        n_stage_steps = min(args.steps, 8)
        cs = [work_counts(model, dev_batches[(warmup + s) % n_pool], opt) for s in range(n_stage_steps)]
        counts = {k: float(np.mean([c[k] for c in cs])) for k in cs[0]}
        torch.cuda.synchronize()
        with StageTimer() as st:
            for s in range(n_stage_steps):
                b = dev_batches[(warmup + s) % n_pool]
                stepper(b["cam"], b["raydir"], b["gt"], uniform=b["uniform"], global_rays=global_rays, reduce=False)
            tot, cnt = st.totals_ms()
        stages = {k: round(v / n_stage_steps, 4) for k, v in sorted(tot.items())}
        agg_ms = (tot.get("agg_forward", 0.0) + tot.get("agg_backward", 0.0)) / n_stage_steps
        fl = agg_flops(opt, counts["S"], counts["M"], backward=True)
        achieved = fl / (agg_ms * 1e-3) / 1e12 if agg_ms > 0 else 0.0
        peak = peaks["bf16_tflops_sustained"]       # whole aggregator, timed inside the step: sustained peak
        # dominant kernel: the pair-level NT GEMM of the MLP (14 of the 21 GEMM launches of a step), timed live on its own
        # at the workload's shape (M = mean valid pairs of the batches, N = K = 256; operands 2 x 138 MB > L2)
        dom = dominant_kernel_roofline(_lib, dev, int(round(counts["M"])), peaks["bf16_tflops"], peaks)     # timed alone: burst peak
        roof = dict(dom)
        roof["aggregator"] = {"kernels": "pn_agg_forward + pn_agg_backward (21 GEMM launches + gather / reduce / scatter kernels)",
                              "achieved": round(achieved, 3), "unit": "TFLOP/s", "frac": round(achieved / peak, 5),
                              "flops_per_step": fl, "ms_per_step": round(agg_ms, 4)}
        # HBM-bound stages, reported alongside (algorithmic bytes of SURVEY.md 8(d))
        C, SR, K, D = int(opt.shading_color_channel_num), int(opt.SR), int(opt.K), int(opt.z_depth_dim)
        R = args.rays
        hb = {}
        b_knn = counts["kept_samples"] * (12 + 27 * 8) + counts["candidates"] * 16 + counts["kept_samples"] * K * 4
        b_sample = R * 12 + R * D / 8 + counts["kept_samples"] * 16 + R
        # compact feature layout of the fused path: one [C+1] row per VALID sample + per-slot scalars (dist, valid, slot_row)
        b_cf = counts["S"] * (C + 1) * 4 + R * SR * 9 + R * (C + 1) * 4 + R * SR * 12
        b_cb = R * C * 4 + 2 * counts["S"] * (C + 1) * 4 + R * SR * (9 + 8)
        for nm, by in (("knn_query", b_knn), ("ray_sample", b_sample), ("composite_forward", b_cf), ("composite_backward", b_cb)):
            if tot.get(nm):
                gbs = by / (tot[nm] / n_stage_steps * 1e-3) / 1e9
                hb[nm] = {"GB/s": round(gbs, 2), "frac_of_hbm": round(gbs / peaks["hbm_gbs"], 5), "bytes": int(by)}
        roof["hbm_stages"] = hb

    # ---- CPU baseline (rank 0, N = 1 only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        torch.set_num_threads(os.cpu_count() or 1)
        sd = {k: v.detach().cpu() for k, v in model.aggregator.state_dict().items()}
        # bounded sample: ~10 s of CPU work - whole batches of the workload (the reference arm, `--impl reference`, takes
        # its step count from the command line and uses --cpu-rays rays of each batch instead)
        cpu_steps = 6
        v, spstep, backend = cpu_oracle_steps(opt, tables, batches, bg, sd, args.rays, cpu_steps, 1)
        cpu = {"value": v, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port" if backend == "port" else "reference",
               "kind_detail": "reference querier kernels (host build of its own CUDA string) + port of its torch aggregator / ray_march" if backend == "ref" else "port",
               "sample": "%d steps of %d rays (whole batches of the workload) after 1 warm-up step, fwd+bwd, %.2f s/step; querier = %s, aggregator/ray_march = torch restatement of the reference modules"
                         % (cpu_steps, args.rays, spstep,
                            "reference kernels compiled for the host (oracle/_ref)" if backend == "ref" else "C port")}
        cpu["psnr"] = psnr_vs_oracle(dev)

    maint = None
    if rank == 0 and world == 1 and not args.no_sweep:
        _phase("maintenance kernels")
        maint = maintenance_bench(model, opt, dev_batches[0], bg_p.detach(), dev)

    ref_gpu_res = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        _phase("reference querier kernels on the GPU")
        try:
            ref_gpu_res = reference_gpu_bench(model, opt, dev_batches[0], dev)
        except Exception as e:
            ref_gpu_res = {"error": repr(e)}

    frustum = None
    if rank == 0 and world == 1 and not args.no_sweep:
        _phase("frustum querier (wcoord_query = 0)")
        try:
            frustum = frustum_bench(dev)
        except Exception as e:   # an auxiliary row: report the failure in the line instead of losing the headline
            frustum = {"error": repr(e)}

    # ---- cfg5 sweep (N = 1, after everything else: it prunes / grows the point cloud)
    sweep = None
    if rank == 0 and world == 1 and not args.no_sweep and not args.no_frame:
        _phase("cfg5 inference sweep")
        sweep = frame_sweep(model, opt, args.workload, bg_p.detach(), dev, args.frame_chunk)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warmup,
                "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic",
                "config": {"workload": "%s: %d-point neural cloud (F=%d), %d rays/batch/GPU, SR=%d, K=%d, D=%d, C=%d, train jitter on, fwd+bwd"
                                       % (args.workload, tables["xyz"].shape[0], opt.point_features_dim, args.rays, opt.SR, opt.K,
                                          opt.z_depth_dim, opt.shading_color_channel_num),
                           "step": "query (cached grid) + fused gather/aggregator + ray_march + colour MSE, backward to point tables + MLP + bg"
                                   + (("; NCCL: " + ("dense all-reduce of %.1f MB gradients" % (stepper.nbytes / 1e6) if args.dense_allreduce else
                                                 "all-gather of the touched point-table gradient rows (ids + 39 floats, rank-ordered sum) + all-reduce of the 1.4 MB MLP gradients"))
                                      if world > 1 else ""),
                           "parallelism": "ray-sharded x%d, points+MLP replicated" % world,
                           "driver": ("autograd from Python (RayShardedStep, --eager)" if fstep is None else
                                      "FusedTrainStep: two CUDA graphs per step (query | forward + loss + backward), sample / pair counts stay on the device; "
                                      "capacities S_cap=%d M_cap=%d (calibrated on the batch pool x1.25), %d overflows; %d library kernels per replayed step"
                                      % (fstep.caps[0], fstep.caps[1], fstep.overflows, fstep.launches_per_step)),
                           "loss": ("color_loss_items ray_masked_coarse_raycolor ray_miss_coarse_raycolor coarse_raycolor, weights 1 0 0 (lego.sh:152-153)"
                                    if args.loss == "lego" else "coarse_raycolor MSE over all rays"),
                           "l2": "working set (point tables %.0f MB + per-step activations) exceeds the 126 MB L2; a pool of %d distinct ray batches is cycled"
                                 % (tables["xyz"].shape[0] * 42 * 4 / 1e6, n_pool),
                           "mean_valid_samples_per_batch": counts["S"] if counts else None,
                           "mean_valid_pairs_per_batch": counts["M"] if counts else None},
                "clocks": clocks, "gpu_launches": int(launches),
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4,
                        "host_queue_ms_per_step": round(e2e_host_ms - e2e_wait_ms, 4), "host_wait_ms_per_step": round(e2e_wait_ms, 4)},
                "host_queue_ms_per_step": round(host_ms, 4),
                "roofline": roof, "cpu_baseline": cpu, "stages_ms_per_step": stages, ("frame_%dx%d" % (frame["w"], frame["h"]) if frame else "frame_800x800"): frame,
                "sharded_gradient_check": grad_check, "legs": legs, "reference_gpu_kernels": ref_gpu_res,
                "train_step_with_adam": with_adam, "frame_sweep_cfg5": sweep, "maintenance": maint, "frustum_querier_wcoord0": frustum}
        _emit(line)
    if world > 1:
        # (no destroy_process_group: with collectives captured in CUDA graphs ProcessGroupNCCL's teardown was seen to wait
        #  forever on work handles recorded during the captures; the processes synchronise and leave)
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        _phase("done")
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)
    _phase("done")
    return 0


if __name__ == "__main__":
    sys.exit(main())
