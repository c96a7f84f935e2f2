"""``NeuralPoints`` (point tables + query + gather) and the per-ray driver ``NeuralPointsRayMarching``.

Mirrors, for the hot path only:
  * models/neural_points/neural_points.py - ``NeuralPoints``: the ``nn.Parameter`` tables ``xyz``,
    ``points_embeding``, ``points_conf``, ``points_dir``, ``points_color``, ``Rw2c`` (:246-326, 480-546, same
    names / shapes so reference checkpoints load), ``set_points`` (:480-546), ``w2pers`` (:687-693),
    ``get_point_indices`` (:634-660) and ``forward`` (:782-812, the 14-tuple of gathered tensors);
  * models/neural_points_volumetric_model.py - ``NeuralPointsRayMarching.forward`` (:272-352): query ->
    aggregate -> ray_dist -> ray_march -> ``fill_invalid`` (:354-389); the 2-D neural-render tail (:343-351)
    is out of scope.

``NeuralPointsRayMarching.forward`` runs the FUSED path by default: ``sample_pidx`` and the point tables go
straight to the aggregator kernels (no gathered ``[R,SR,K,*]`` tensors), rays are never compacted (rays
without neighbours composite to the background exactly as ``fill_invalid`` fills them), and the only host
synchronisation of a step is the read-back of the (valid samples, valid pairs) counts.  ``fused=False``
goes through the three reference seams one by one (query_points -> forward -> aggregator -> ray_march ->
fill_invalid) and produces the same numbers.
"""
import numpy as np
import torch
import torch.nn as nn

from . import _lib
from . import query as _query
from .aggregator import CompactSamples, PointAggregator, gather_clamp
from .query import lighting_fast_querier as lighting_fast_querier_w
from .query_frustum import lighting_fast_querier as lighting_fast_querier_p
from .ray_march import alpha_blend, radiance_render, ray_dist_from_samples, ray_march, ray_march_compact


def _as_reference_scalar(v, reduce):
    """How NeuralPoints.forward of the reference hands near / far / h / w to the frustum querier (neural_points.py:786):
    ``torch.min|max(v).cpu().numpy()``, a 0-d numpy array whose dtype (float32, int64) steers numpy's promotion in
    get_hyperparameters; python numbers and numpy values pass through."""
    if isinstance(v, (int, float, np.generic, np.ndarray)):
        return v
    t = torch.as_tensor(v)
    return (t.min() if reduce == "min" else t.max()).cpu().numpy()


class NeuralPoints(nn.Module):
    def __init__(self, num_channels, size, opt, device, checkpoint=None, feature_init_method='rand', reg_weight=0.,
                 feedforward=0):
        super().__init__()
        self.opt = opt
        self.device = device
        self.grid_vox_sz = 0
        self.points_conf = self.points_dir = self.points_color = None
        self.xyz = self.points_embeding = None
        self.Rw2c = nn.Parameter(torch.eye(3, device=device, dtype=torch.float32), requires_grad=False)
        # neural_points.py:329-331: query_size defaults to kernel_size; world-coordinate or perspective-frustum querier
        if opt.query_size[0] == 0:
            opt.query_size = opt.kernel_size
        self.lighting_fast_querier = lighting_fast_querier_w if getattr(opt, "wcoord_query", 1) > 0 else lighting_fast_querier_p
        self.querier = self.lighting_fast_querier(device, opt)
        # neural_points.py:240-300 (opt.load_points == 1): tables from a checkpoint's ``neural_points.*`` entries.  The
        # cloud-file initialisers of that block (load_blender_cloud, construct_grid_points, feature_init_method) belong
        # to scene set-up, outside the hot path: without a checkpoint the tables stay empty until set_points().
        if checkpoint:
            sd = torch.load(checkpoint, map_location=device) if isinstance(checkpoint, (str, bytes)) or hasattr(checkpoint, "read") else checkpoint
            if "neural_points.xyz" not in sd or "neural_points.points_embeding" not in sd:
                raise KeyError("checkpoint holds no neural_points.xyz / neural_points.points_embeding")
            get = lambda k: sd.get("neural_points." + k)
            self.set_points(get("xyz"), get("points_embeding"), points_color=get("points_color"), points_dir=get("points_dir"),
                            points_conf=get("points_conf"), Rw2c=get("Rw2c"))
        elif feature_init_method != 'rand' or size:
            raise NotImplementedError("NeuralPoints without a checkpoint: the cloud-file / feature_init_method initialisers "
                                      "(neural_points.py:246-292) are scene set-up, outside the hot path; call set_points()")

    def reset_querier(self):
        self.querier.clean_up()
        self.querier = self.lighting_fast_querier(self.device, self.opt)

    def set_points(self, points_xyz, points_embeding, points_color=None, points_dir=None, points_conf=None,
                   parameter=False, Rw2c=None, eulers=None):
        """neural_points.py:480-546: tensors in the reference's layouts (xyz [N,3]; the others [1,N,*])."""
        o = self.opt
        dev = self.device
        if points_embeding.shape[-1] > o.point_features_dim:
            points_embeding = points_embeding[..., :o.point_features_dim]
        self.xyz = nn.Parameter(points_xyz.to(dev).float().contiguous(), requires_grad=False)  # xyz_grad = 0 only
        if getattr(o, "xyz_grad", 0) > 0:
            raise NotImplementedError("xyz_grad=1 is outside the hot path")
        self.points_embeding = nn.Parameter(points_embeding.to(dev).float().contiguous(), requires_grad=o.feat_grad > 0)
        if points_conf is not None and "1" in list(o.point_conf_mode):
            self.points_conf = nn.Parameter(points_conf.to(dev).float().contiguous(), requires_grad=o.conf_grad > 0)
        if points_dir is not None and "1" in list(o.point_dir_mode):
            self.points_dir = nn.Parameter(points_dir.to(dev).float().contiguous(), requires_grad=o.dir_grad > 0)
        if points_color is not None and "1" in list(o.point_color_mode):
            self.points_color = nn.Parameter(points_color.to(dev).float().contiguous(), requires_grad=o.color_grad > 0)
        if Rw2c is not None:
            self.Rw2c = nn.Parameter(Rw2c.to(dev).float().contiguous(), requires_grad=False)
        self.querier.clean_up()  # the cached grid belongs to the old point set

    _TABLES = ("points_embeding", "points_conf", "points_dir", "points_color")

    def prune(self, thresh):
        """neural_points.py:350-373: keep points with conf >= thresh; invalidates the grid.  Returns the kept count."""
        return self.prune_and_grow(thresh)

    def grow_points(self, add_xyz, add_embedding, add_color, add_dir, add_conf):
        """neural_points.py:376-402: append points (add_* are [n,*] like the reference's); invalidates the grid."""
        self.prune_and_grow(None, add_xyz, add_embedding, add_color, add_dir, add_conf)

    def prune_and_grow(self, thresh=None, add_xyz=None, add_embedding=None, add_color=None, add_dir=None, add_conf=None):
        """``prune(thresh)`` followed by ``grow_points(add_*)`` (the pair run/train_ft.py:841-849 issues between passes)
        with ONE allocation and one pass per table: kept rows are moved by pn_points_compact (all five tables in one
        launch, order preserved like boolean-mask indexing) and the new rows are written behind them.  Either half may be
        omitted.  Same resulting tables / requires_grad flags as the reference's two methods.  Returns the kept count."""
        lib = _lib.lib()
        dev = self.xyz.device
        N, F_ = int(self.xyz.shape[0]), int(self.points_embeding.shape[-1])
        n_add = 0 if add_xyz is None else int(add_xyz.shape[0])
        dst = None
        kept = N
        if thresh is not None:
            if self.points_conf is None:
                raise RuntimeError("prune needs points_conf")
            dst = torch.empty(N + 1, dtype=torch.int32, device=dev)
            ws_bytes = lib.pn_prune_workspace_bytes(N)
            ws = torch.empty(max(ws_bytes, 1), dtype=torch.uint8, device=dev)
            _lib.check(lib.pn_prune_mark(_lib.ptr(self.points_conf.detach()), N, float(thresh), _lib.ptr(dst), _lib.ptr(ws), ws_bytes,
                                         _lib.stream()), "pn_prune_mark")
            kept = int(dst[N].item())            # host sync: the new tables' size (the reference syncs in every mask index)
        if dst is None and n_add == 0:
            return kept
        total = kept + n_add
        old = {n: getattr(self, n) for n in self._TABLES}
        new_xyz = torch.empty((total, 3), dtype=torch.float32, device=dev)
        new = {n: (None if t is None else torch.empty((1, total, t.shape[-1]), dtype=torch.float32, device=dev)) for n, t in old.items()}
        if dst is not None and kept > 0:
            _lib.check(lib.pn_points_compact(
                _lib.ptr(dst), N, F_, _lib.ptr(self.xyz.detach()), _lib.ptr(old["points_embeding"].detach()),
                _lib.ptr(old["points_conf"].detach()), _lib.ptr(None if old["points_dir"] is None else old["points_dir"].detach()),
                _lib.ptr(None if old["points_color"] is None else old["points_color"].detach()),
                _lib.ptr(new_xyz), _lib.ptr(new["points_embeding"]), _lib.ptr(new["points_conf"]), _lib.ptr(new["points_dir"]),
                _lib.ptr(new["points_color"]), _lib.stream()), "pn_points_compact")
        elif dst is None:
            new_xyz[:N].copy_(self.xyz.detach())
            for n, t in old.items():
                if t is not None:
                    new[n][0, :N].copy_(t.detach()[0])
        if n_add:
            adds = {"points_embeding": add_embedding, "points_conf": add_conf, "points_dir": add_dir, "points_color": add_color}
            new_xyz[kept:].copy_(add_xyz)
            for n, t in new.items():
                if t is not None:
                    t[0, kept:].copy_(adds[n].reshape(n_add, -1))
        self.xyz = nn.Parameter(new_xyz, requires_grad=self.xyz.requires_grad)
        for n, t in new.items():
            if t is not None:
                setattr(self, n, nn.Parameter(t, requires_grad=old[n].requires_grad))
        self.querier.clean_up()                  # the cached grid belongs to the old point set
        return kept

    def w2pers(self, point_xyz, camrotc2w, campos):
        """neural_points.py:687-693: [N,3] -> [1,N,3] (x/z, y/z, z) in the camera frame."""
        return _query.w2pers(point_xyz.detach(), camrotc2w[0], campos[0])[None]

    def get_point_indices(self, inputs, cam_rot_tensor, cam_pos_tensor, pixel_idx_tensor, near_plane, far_plane, h, w, intrinsic,
                          vox_query=False):
        """neural_points.py:634-660."""
        if vox_query:
            raise NotImplementedError("vox_query (NN<0) is outside the hot path")
        point_xyz_pers_tensor = self.w2pers(self.xyz, cam_rot_tensor, cam_pos_tensor)
        n = torch.full([1], self.xyz.shape[0], device=self.xyz.device, dtype=torch.int32)
        (sample_pidx_tensor, sample_loc_tensor, sample_loc_w_tensor, sample_ray_dirs_tensor, ray_mask_tensor, vsize,
         ranges) = self.querier.query_points(pixel_idx_tensor, point_xyz_pers_tensor, self.xyz[None, ...], n, h, w, intrinsic,
                                             near_plane, far_plane, inputs["raydir"], cam_pos_tensor, cam_rot_tensor,
                                             uniform=inputs.get("uniform"))
        return (sample_pidx_tensor, sample_loc_tensor, ray_mask_tensor, point_xyz_pers_tensor, sample_loc_w_tensor,
                sample_ray_dirs_tensor, vsize)

    def forward(self, inputs):
        """neural_points.py:782-812 (reference seam): the 14-tuple of gathered [1,R'',SR,K,*] tensors."""
        camrotc2w, campos, near, far = inputs["camrotc2w"], inputs["campos"], inputs["near"], inputs["far"]
        if getattr(self.opt, "wcoord_query", 1) > 0:
            near = float(near) if isinstance(near, (int, float)) else float(torch.as_tensor(near).min())
            far = float(far) if isinstance(far, (int, float)) else float(torch.as_tensor(far).max())
        else:   # the frustum grid's voxel size is numpy arithmetic on these: tensors become 0-d arrays like :786
            near, far = _as_reference_scalar(near, "min"), _as_reference_scalar(far, "max")
        pixel_idx = inputs.get("pixel_idx")
        h, w, intrinsic = inputs.get("h"), inputs.get("w"), inputs.get("intrinsic")
        if getattr(self.opt, "wcoord_query", 1) <= 0:
            h, w = _as_reference_scalar(h, "max"), _as_reference_scalar(w, "max")
            intrinsic = intrinsic.cpu().numpy()[0] if torch.is_tensor(intrinsic) else intrinsic        # :786
        sample_pidx, sample_loc, ray_mask_tensor, xyz_pers, sample_loc_w, sample_ray_dirs, vsize = self.get_point_indices(
            inputs, camrotc2w, campos, pixel_idx, near, far, h, w, intrinsic)
        sample_pnt_mask = sample_pidx >= 0
        B, R, SR, K = sample_pidx.shape
        flat = torch.clamp(sample_pidx, min=0).view(-1).long()

        def sel(t):
            return None if t is None else torch.index_select(t, 1, flat).view(B, R, SR, K, t.shape[2])
        return (sel(self.points_color), self.Rw2c, sel(self.points_dir), sel(self.points_conf), sel(self.points_embeding),
                sel(xyz_pers), sel(self.xyz[None, ...]), sample_pnt_mask, sample_loc, sample_loc_w, sample_ray_dirs,
                ray_mask_tensor, vsize, self.grid_vox_sz)


class RayPlan:
    """Weight-independent part of one batch (NeuralPointsRayMarching.sample): raw query results, perspective coordinates
    of samples and points, and the aggregator's work lists (CompactSamples)."""
    __slots__ = ("raw", "pidx", "loc_w", "loc_pers", "xyz_pers", "raydir", "work_lists", "vsize", "ray_mask", "sample_dirs")


class NeuralPointsRayMarching(nn.Module):
    """Per-ray driver (neural_points_volumetric_model.py:272-389) without the 2-D neural-render tail."""

    def __init__(self, opt, neural_points=None, aggregator=None, device=None):
        super().__init__()
        self.opt = opt
        device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.aggregator = aggregator if aggregator is not None else PointAggregator(opt).to(device)
        self.neural_points = neural_points if neural_points is not None else NeuralPoints(opt.point_features_dim, 0, opt, device)
        self.render_func, self.blend_func = radiance_render, alpha_blend
        self.return_color = True
        if getattr(opt, "which_tonemap_func", "off") != "off":
            raise NotImplementedError("tone mapping is outside the hot path (all scripts: which_tonemap_func='off')")

    # ---- reference seams, one by one ------------------------------------------------------------
    def forward_seams(self, campos, raydir, bg_color=None, camrotc2w=None, pixel_idx=None, near=None, far=None, h=None, w=None,
                      intrinsic=None, uniform=None, **kargs):
        o = self.opt
        (sampled_color, sampled_Rw2c, sampled_dir, sampled_conf, sampled_embedding, sampled_xyz_pers, sampled_xyz,
         sample_pnt_mask, sample_loc, sample_loc_w, sample_ray_dirs, ray_mask_tensor, vsize, grid_vox_sz) = self.neural_points(
            {"pixel_idx": pixel_idx, "camrotc2w": camrotc2w, "campos": campos, "near": near, "far": far, "h": h, "w": w,
             "intrinsic": intrinsic, "raydir": raydir, "uniform": uniform})
        decoded_features, ray_valid, weight, conf_coefficient = self.aggregator(
            sampled_color, sampled_Rw2c, sampled_dir, sampled_conf, sampled_embedding, sampled_xyz_pers, sampled_xyz,
            sample_pnt_mask, sample_loc, sample_loc_w, sample_ray_dirs, vsize, grid_vox_sz)
        ray_dist = ray_dist_from_samples(sample_loc, ray_valid, vsize, o.raydist_mode_unit)
        output = {"queried_shading": torch.logical_not(torch.any(ray_valid, dim=-1, keepdims=True)).repeat(1, 1, 3).float()}
        ray_color, _, opacity, _, blend_weight, background_transmission, _ = ray_march(
            ray_dist, ray_valid, decoded_features, self.render_func, self.blend_func, bg_color)
        output["coarse_raycolor"] = ray_color
        output["coarse_point_opacity"] = opacity
        output["coarse_is_background"] = background_transmission
        output["ray_mask"] = ray_mask_tensor
        if weight is not None:
            output["weight"] = weight.detach()
            output["blend_weight"] = blend_weight.detach()
            output["conf_coefficient"] = conf_coefficient
        return self.fill_invalid(output, bg_color)

    @staticmethod
    def fill_invalid(output, bg_color):
        """neural_points_volumetric_model.py:354-389 (tone map off)."""
        ray_mask = output["ray_mask"]
        B, OR = ray_mask.shape
        idx = torch.nonzero(ray_mask)
        dev = output["coarse_raycolor"].device
        is_bg = torch.ones([B, OR, 1], dtype=torch.float32, device=dev)
        is_bg[idx[..., 0], idx[..., 1], :] = output["coarse_is_background"]
        output["coarse_is_background"] = is_bg
        output["coarse_mask"] = 1 - is_bg
        colour = torch.ones([B, OR, output["coarse_raycolor"].shape[-1]], dtype=torch.float32, device=dev) * bg_color[None, ...]
        colour[idx[..., 0], idx[..., 1], :] = output["coarse_raycolor"]
        output["coarse_raycolor"] = colour
        opac = torch.zeros([B, OR, output["coarse_point_opacity"].shape[2]], dtype=torch.float32, device=dev)
        opac[idx[..., 0], idx[..., 1], :] = output["coarse_point_opacity"]
        output["coarse_point_opacity"] = opac
        shading = torch.ones([B, OR, output["queried_shading"].shape[2]], dtype=torch.float32, device=dev)
        shading[idx[..., 0], idx[..., 1], :] = output["queried_shading"]
        output["queried_shading"] = shading
        return output

    def touched_point_ids(self):
        """Point indices referenced by the most recent fused forward (one entry per valid (sample, neighbour) pair,
        duplicates included; no host sync) - the rows of the point tables that can carry gradient this step."""
        call = getattr(self.aggregator, "last_call", None)
        if call is None or call.M == 0:
            return None if call is None else torch.zeros(0, dtype=torch.long, device=call.sample_pidx.device)
        return call.sample_pidx.view(-1)[call.pair_src[:call.M].long()].long()

    # ---- fused path -----------------------------------------------------------------------------
    def sample(self, campos, raydir, camrotc2w, near, far, uniform=None, xyz_pers=None, pixel_idx=None, h=None, w=None,
               intrinsic=None):
        """Everything of one batch that does not depend on the trainable tensors: ray sampling, neighbour query,
        perspective coordinates and the work lists, queued on the current stream with NO host sync.  ``forward(plan=...)``
        consumes it; a training loop calls this for batch i+1 before the backward of batch i so that the (S, M)
        read-back the aggregator needs is complete long before it is waited for."""
        nps = self.neural_points
        if getattr(self.opt, "wcoord_query", 1) <= 0:
            return self._sample_frustum(campos, camrotc2w, near, far, pixel_idx, h, w, intrinsic, uniform, xyz_pers)
        near = float(near) if isinstance(near, (int, float)) else float(torch.as_tensor(near).min())
        far = float(far) if isinstance(far, (int, float)) else float(torch.as_tensor(far).max())
        xyz = nps.xyz.detach()
        raw = nps.querier.query_raw(xyz, raydir, campos, near, far, uniform=uniform)
        pidx, loc_w = raw["sample_pidx"], raw["sample_loc_w"]                     # [R,SR,K], [R,SR,3]
        R, SR, K = pidx.shape
        plan = RayPlan()
        plan.raw, plan.pidx, plan.loc_w, plan.raydir = raw, pidx, loc_w, raydir
        plan.loc_pers = _query.w2pers(loc_w, camrotc2w[0], campos[0])
        # the points' camera-frame coordinates depend on the camera only: a frame renderer passes them along
        plan.xyz_pers = xyz_pers if xyz_pers is not None else _query.w2pers(xyz, camrotc2w[0], campos[0])
        plan.work_lists = CompactSamples(pidx.contiguous(), R * SR, K)
        plan.vsize, plan.ray_mask, plan.sample_dirs = self.opt.vsize, None, None
        return plan

    def _sample_frustum(self, campos, camrotc2w, near, far, pixel_idx, h, w, intrinsic, uniform, xyz_pers):
        """``sample`` on the perspective-frustum querier (wcoord_query = 0): rays come as pixel indices, the grid is built in
        this view's perspective space, sample centres are perspective coordinates (jittered in z when training, then
        pers2w gives world positions and per-sample directions, query_point_indices.py:100-108).  Rays are not compacted:
        the ones the reference drops (ray_mask 0) simply keep no samples.  ``uniform``: [1,R,SR] draw of the jitter."""
        nps, o = self.neural_points, self.opt
        if pixel_idx is None or h is None or w is None or intrinsic is None:
            raise NotImplementedError("wcoord_query = 0 queries by pixel: pass pixel_idx, h, w and intrinsic (RayShardedStep, which "
                                      "knows rays by direction only, is world-coordinate)")
        q = nps.querier
        near, far = _as_reference_scalar(near, "min"), _as_reference_scalar(far, "max")
        h, w = _as_reference_scalar(h, "max"), _as_reference_scalar(w, "max")
        intrinsic = intrinsic.cpu().numpy()[0] if torch.is_tensor(intrinsic) else np.asarray(intrinsic)   # neural_points.py:786
        hp = q.get_hyperparameters(h, w, intrinsic, near, far)
        xp = xyz_pers if xyz_pers is not None else _query.w2pers(nps.xyz.detach(), camrotc2w[0], campos[0])
        raw = q.query_raw(pixel_idx.reshape(-1, 2).to(torch.int32).contiguous(), xp.reshape(-1, 3), hp)
        pidx, loc_pers = raw["sample_pidx"], raw["sample_loc"]                    # [R,SR,K], [R,SR,3]
        R, SR, K = pidx.shape
        if o.is_train:
            loc_pers = getattr(q, o.shpnt_jitter)(loc_pers[None], hp["vsize"], uniform=uniform)[0]
        loc_w, dirs = q.pers2w(loc_pers[None], camrotc2w, campos)
        plan = RayPlan()
        plan.raw, plan.pidx, plan.loc_w, plan.loc_pers, plan.raydir = raw, pidx, loc_w[0].contiguous(), loc_pers, None
        plan.xyz_pers = xp.reshape(-1, 3)
        plan.work_lists = CompactSamples(pidx.contiguous(), R * SR, K)
        plan.vsize, plan.ray_mask, plan.sample_dirs = hp["vsize"], raw["ray_mask"], dirs[0].contiguous()
        return plan

    @torch.no_grad()
    def render_frame(self, campos, raydir, bg_color, camrotc2w, near, far, chunk=65536, keys=("coarse_raycolor",), side_stream=True,
                     pixel_idx=None, h=None, w=None, intrinsic=None):
        """Full-frame inference: the chunk loop of run/train_ft.py:test (:284-340, ``chunk_size = random_sample_size**2``
        rays per ``model.test()`` call) over raydir [1,R,3], forward only.  The sampling / neighbour query / work lists of
        chunk i+1 run on a second stream under the aggregator of chunk i (they are small latency-bound kernels, ~20 % of a
        frame when serialised), so neither they nor the (S, M) read-back of a chunk sit on the critical path; chunks reuse
        the same activation buffers through the caching allocator (nothing is kept for a backward).
        Returns {key: [1,R,...]} for ``keys`` (any of the ``forward`` outputs).
        Perspective-frustum querier (``wcoord_query = 0``, the DTU feed-forward scripts): rays are addressed by
        ``pixel_idx`` int32 [1,R,2] (+ ``h``, ``w``, ``intrinsic``) instead of ``raydir``; the view's grid is a function of
        the camera only and is rebuilt per chunk by the querier (0.13 ms at DTU size)."""
        frustum = getattr(self.opt, "wcoord_query", 1) <= 0
        if frustum and (pixel_idx is None or h is None or w is None or intrinsic is None):
            raise NotImplementedError("wcoord_query = 0 renders by pixel: pass pixel_idx, h, w and intrinsic")
        R = int(pixel_idx.shape[1]) if frustum else int(raydir.shape[1])
        dev0 = pixel_idx.device if frustum else raydir.device
        if R == 0:
            return {k: torch.zeros((1, 0, 0), device=dev0) for k in keys}
        starts = list(range(0, R, int(chunk)))
        outs = {k: [] for k in keys}
        main = torch.cuda.current_stream()
        side = None
        if side_stream and dev0.type == "cuda":
            side = getattr(self, "_frame_stream", None)
            if side is None:
                side = self._frame_stream = torch.cuda.Stream()
        xp = _query.w2pers(self.neural_points.xyz.detach(), camrotc2w[0], campos[0])

        def one(s):
            if frustum:
                return self.sample(campos, None, camrotc2w, near, far, xyz_pers=xp, pixel_idx=pixel_idx[:, s:s + chunk].contiguous(), h=h, w=w,
                                   intrinsic=intrinsic)
            return self.sample(campos, raydir[:, s:s + chunk], camrotc2w, near, far, xyz_pers=xp)

        def sample(s):
            if side is None:
                return one(s), None
            # allocations from the side stream's pool only happen after this wait: every main-stream kernel that used the
            # plans freed so far has been queued before it
            side.wait_stream(main)
            with torch.cuda.stream(side):
                plan = one(s)
                done = torch.cuda.Event()
                done.record(side)
            return plan, done
        plan, done = sample(starts[0])
        for i, s in enumerate(starts):
            nxt, nxt_done = sample(starts[i + 1]) if i + 1 < len(starts) else (None, None)
            if done is not None:
                main.wait_event(done)
            o = self.forward(campos, None if frustum else raydir[:, s:s + chunk], bg_color=bg_color, camrotc2w=camrotc2w, near=near, far=far,
                             plan=plan)
            for k in keys:
                outs[k].append(o[k])
            plan, done = nxt, nxt_done
        return {k: (torch.cat(v, dim=1) if len(v) != 1 else v[0]) for k, v in outs.items()}

    def forward(self, campos, raydir, gt_image=None, bg_color=None, camrotc2w=None, pixel_idx=None, near=None, far=None,
                focal=None, h=None, w=None, style_code=None, intrinsic=None, uniform=None, fused=True, plan=None, **kargs):
        """Same keyword interface as the reference driver (:272-286).  Returns the ``fill_invalid`` dictionary:
        coarse_raycolor [1,R,C], coarse_is_background [1,R,1], coarse_mask, coarse_point_opacity [1,R,SR],
        queried_shading [1,R,3], ray_mask int8 [1,R] (+ weight / blend_weight / conf_coefficient per sample when the
        losses need them; in the fused path these are [1,R,SR,K] over ALL rays, zero on rays that miss)."""
        if not fused:
            return self.forward_seams(campos, raydir, bg_color=bg_color, camrotc2w=camrotc2w, pixel_idx=pixel_idx, near=near,
                                      far=far, h=h, w=w, intrinsic=intrinsic, uniform=uniform, **kargs)
        o, nps = self.opt, self.neural_points
        if plan is None:
            plan = self.sample(campos, raydir, camrotc2w, near, far, uniform=uniform, pixel_idx=pixel_idx, h=h, w=w,
                               intrinsic=intrinsic)
        raw, pidx, loc_w, loc_pers, xyz_pers = plan.raw, plan.pidx, plan.loc_w, plan.loc_pers, plan.xyz_pers
        R, SR, K = pidx.shape
        want_w = self.aggregator._want_weight()
        frustum = plan.sample_dirs is not None       # per-sample directions from pers2w, the view's own voxel size
        decoded, weight, slot_row = self.aggregator.forward_tables(
            nps.xyz, xyz_pers, nps.points_embeding, nps.points_color, nps.points_dir, nps.points_conf, nps.Rw2c,
            pidx, loc_pers, loc_w, plan.sample_dirs if frustum else plan.raydir[0], plan.vsize, dirs_per_ray=not frustum,
            want_weight=want_w, compact=True, work_lists=plan.work_lists)
        ray_valid = (slot_row >= 0)[None]                                         # [1,R,SR]
        ray_dist = ray_dist_from_samples(loc_pers[None], ray_valid, plan.vsize, o.raydist_mode_unit)
        ray_color, _, opacity, _, blend_weight, bg_T, _ = ray_march_compact(ray_dist, ray_valid, decoded, slot_row, bg_color)
        if frustum:
            # the frustum querier's ray mask says "the ray's column holds an occupied depth", not "a neighbour was found"
            ray_any = ray_valid[0].any(-1).to(torch.int32)
            ray_mask = (plan.ray_mask > 0).view(torch.int8)[None]
        else:
            ray_any = raw["ray_any"]
            ray_mask = (ray_any > 0).view(torch.int8)[None]
        out = {"coarse_raycolor": ray_color, "coarse_point_opacity": opacity, "coarse_is_background": bg_T,
               "coarse_mask": 1 - bg_T, "ray_mask": ray_mask,
               # a ray has a valid sample exactly when the query found it a neighbour (ray_any): same values as
               # logical_not(any(ray_valid)).repeat(1,1,3).float() of the reference (:162) in two launches
               "queried_shading": (ray_any.view(1, R, 1) == 0).expand(1, R, 3).float()}
        if want_w:
            out["weight"] = weight.detach()[None]
            out["blend_weight"] = blend_weight.detach()
            out["conf_coefficient"] = gather_clamp(nps.points_conf, pidx, 0.0001, 1.0)[None]
            if getattr(o, "prob", 0) == 1:
                out.update(self.probe_outputs(opacity, ray_any, loc_w, pidx, weight))
        return out

    @torch.no_grad()
    def probe_outputs(self, opacity, ray_any, sample_loc_w, sample_pidx, weight):
        """The ``opt.prob == 1`` block of neural_points_volumetric_model_ori.py:351-372 in one launch (pn_probe_rays), on
        the UNCOMPACTED ray layout: every output is [1,R,...] with zeros on rays the query dropped - what ``unmask``
        (:126-137) hands to probe_hole (run/train_ft.py:420-533).  Keys / shapes as the reference's:
        ray_max_shading_opacity [1,R,1], ray_max_sample_loc_w [1,R,3], ray_max_far_dist [1,R,1], shading_avg_color /
        shading_avg_dir [1,R,3], shading_avg_conf [1,R,1], shading_avg_embedding [1,R,F] (+ ray_max_slot int32 [1,R])."""
        nps = self.neural_points
        if nps.points_color is None or nps.points_dir is None or nps.points_conf is None:
            raise NotImplementedError("probe outputs need points_color / points_dir / points_conf (every shipped script has them)")
        R, SR, K = (int(v) for v in sample_pidx.shape)
        F_ = int(nps.points_embeding.shape[-1])
        dev = sample_pidx.device
        # one allocation carved into the eight outputs (sections of whole rows: every view is contiguous)
        widths = (1, 3, 1, 3, 3, 1, F_)
        flat = torch.empty(R * (sum(widths) + 1), dtype=torch.float32, device=dev)
        views, off = [], 0
        for wd in widths:
            views.append(flat[off:off + R * wd].view(1, R, wd))
            off += R * wd
        o_op, o_loc, o_far, o_col, o_dir, o_cnf, o_emb = views
        o_slot = flat[off:off + R].view(torch.int32).view(1, R)
        _lib.check(_lib.lib().pn_probe_rays(
            _lib.ptr(opacity.detach().reshape(R, SR).contiguous()), _lib.ptr(ray_any.contiguous()), _lib.ptr(sample_loc_w.contiguous()),
            _lib.ptr(sample_pidx.contiguous()), _lib.ptr(weight.detach().contiguous()), _lib.ptr(nps.xyz.detach()),
            _lib.ptr(nps.points_embeding.detach()), _lib.ptr(nps.points_color.detach()), _lib.ptr(nps.points_dir.detach()),
            _lib.ptr(nps.points_conf.detach()), R, SR, K, F_, 0.0001, 1.0, _lib.ptr(o_op), _lib.ptr(o_slot), _lib.ptr(o_loc),
            _lib.ptr(o_far), _lib.ptr(o_col), _lib.ptr(o_dir), _lib.ptr(o_cnf), _lib.ptr(o_emb), _lib.stream()), "pn_probe_rays")
        return {"ray_max_shading_opacity": o_op, "ray_max_sample_loc_w": o_loc, "ray_max_far_dist": o_far, "shading_avg_color": o_col,
                "shading_avg_dir": o_dir, "shading_avg_conf": o_cnf, "shading_avg_embedding": o_emb, "ray_max_slot": o_slot}
