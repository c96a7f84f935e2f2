"""The training step of the per-ray hot path as ONE replayable launch sequence (CUDA graphs), with no host read-back.

``NeuralPointsRayMarching.forward`` + ``loss.backward()`` (neural_points_volumetric_model.py:272-352 under
base_rendering_model.py:533-606 ``compute_losses`` / ``backward``) cost ~1.4 ms of host time per step when driven from
Python (~60 kernel launches, ~30 tensor allocations, the autograd graph, one wait for the (S, M) counts) - as much as the
kernels themselves take on a B200.  ``FusedTrainStep`` issues the same kernels in the same order through the same C-ABI
entry points, but

  * without autograd: the chain query -> gather + aggregator -> ray_dist -> ray_march -> colour loss and its backward is
    fixed, so the backward entry points are called directly (``agg_backward_raw``, ``_Composite.backward``);
  * without the host ever learning the number of valid samples / pairs: the work lists are built under CAPACITIES
    (pn_compact_samples_capped), every kernel reads the true counts from device memory (pn_agg_inputs.counts), buffers are
    sized by the capacities;
  * captured once into two CUDA graphs - ``query`` (ray sampling, neighbour query, work lists, per-ray loss weights) and
    ``step`` (forward, loss, backward) - and replayed: two graph launches per step instead of ~100 Python-level calls.
    Between the two a multi-GPU run all-reduces the masked-ray count (4 bytes) that normalises ``ray_masked_*`` losses.

Gradients land in static buffers that become the parameters' ``.grad`` (overwritten by the next step, the usual CUDA-graph
training contract).  Capacities come from ``calibrate()`` (exact counts of sample batches + headroom); a batch that
overflows them is truncated in slot order (later samples composite as empty space), flagged one step later through an
asynchronous copy of the counts, and the capacities are grown for the following steps.
"""
import warnings

import numpy as np
import torch
import torch.distributed as dist

from . import _lib
from . import query as _query
from .aggregator import CompactSamples, _AggCall, agg_backward_raw, agg_forward_raw
from .ray_march import _ColorMSE, _Composite, _as_u8, ray_dist_from_samples


class _Ctx:
    """Stand-in for an autograd context: the Function bodies are called directly (no graph is recorded)."""

    def __init__(self, needs_input_grad):
        self.needs_input_grad = needs_input_grad
        self.saved_tensors = ()

    def save_for_backward(self, *tensors):
        self.saved_tensors = tensors

    def set_materialize_grads(self, value):
        pass

    def mark_non_differentiable(self, *tensors):
        pass


def color_loss_ray_weights(ray_any, items, weights, channels, global_rays, masked_count=None):
    """Per-ray weights w_r such that  sum_i weights[i] * loss_i  =  sum_r w_r * sum_c (pred - gt)^2  for the colour-loss items
    of ``compute_losses`` (base_rendering_model.py:543-560, 598-606; the + 1e-6 per item is a constant):
        coarse_raycolor             MSELoss over all rays                          -> 1 / (R_global * C)
        ray_masked_coarse_raycolor  MSELoss over rays with ray_mask > 0            -> [mask] / (n_masked_global * C)
        ray_miss_coarse_raycolor    MSELoss over rays with ray_mask == 0, times their number  -> [1 - mask] / C
    ``masked_count``: float tensor [1] with the GLOBAL number of masked rays (ray sharding: summed over ranks); default: the
    local count.  Returns float32 [R]."""
    mask = (ray_any.reshape(-1) > 0).to(torch.float32)
    w = torch.zeros_like(mask)
    for name, wt in zip(items, weights):
        wt = float(wt)
        if wt == 0.0:
            continue
        if name.startswith("ray_masked"):
            cnt = masked_count if masked_count is not None else mask.sum().reshape(1)
            w = w + mask * (wt / channels) / cnt.clamp_min(1.0)
        elif name.startswith("ray_miss"):
            w = w + (1.0 - mask) * (wt / channels)
        elif name.startswith("ray_depth_masked"):
            raise NotImplementedError("ray_depth_masked_* needs the MVS depth mask, outside the hot path")
        else:
            w = w + wt / float(global_rays * channels)
    return w


class FusedTrainStep:
    """One training step (query -> forward -> colour loss -> backward) of ``NeuralPointsRayMarching`` as two CUDA graphs.

    model:     NeuralPointsRayMarching (the tables, grid, MLP it holds are used in place).  On the perspective-frustum querier
               (``opt.wcoord_query = 0``, the DTU feed-forward scripts) rays are addressed by ``pixel_idx`` int32 [R,2] instead
               of ``raydir``, and ``view=dict(h=, w=, intrinsic=)`` fixes the frustum the per-view grid is built in (like
               near / far it is baked into the captured step); sample jitter: ``shpnt_jitter`` passfunc or uniform
    bg_color:  nn.Parameter [C] (receives a gradient) or a plain tensor
    rays:      rays per step on this rank (static: the graphs are captured for it); near / far: the camera's depth range
    color_loss_items / color_loss_weights: as the reference's options of the same names (base_rendering_model.py:543-606)
    caps:      (S_cap, M_cap) valid samples / pairs the buffers are sized for; default: ``calibrate()`` must be called
    graph:     False runs the same launch sequence eagerly (tests; no capture)
    group:     torch.distributed group for ray sharding (gradients: summed over ranks by ``reduce()``; the masked-ray count
               is all-reduced between the query and the step)
    pipeline:  the query of the NEXT batch runs inside the step graph, on a forked branch beside the backward (its kernels
               are small and latency-bound, the backward's persistent GEMMs leave SMs idle at their tails).  Protocol:
               ``prime(batch 0)`` once, then ``step(batch i + 1)`` runs forward + loss + backward of batch i and returns its
               loss.  Every step still runs one full query + forward + backward.
    """

    def __init__(self, model, bg_color, rays, near, far, color_loss_items=("coarse_raycolor",), color_loss_weights=(1.0,),
                 caps=None, graph=True, group=None, pipeline=False, local_only=False, sparse_exchange=True, exchange_in_step=True,
                 view=None):
        o = model.opt
        self.frustum = getattr(o, "wcoord_query", 1) <= 0
        self.model, self.bg, self.R, self.near, self.far = model, bg_color, int(rays), float(near), float(far)
        self.hp = None
        if self.frustum:
            if view is None:
                raise ValueError("FusedTrainStep on the frustum querier (wcoord_query = 0) needs view=dict(h=, w=, intrinsic=)")
            if o.is_train > 0 and o.shpnt_jitter not in ("passfunc", "uniform"):
                raise NotImplementedError("FusedTrainStep: shpnt_jitter %r draws inside the query; passfunc / uniform take the draw as an input"
                                          % (o.shpnt_jitter,))
            intr = view["intrinsic"]
            intr = intr.cpu().numpy()[0] if torch.is_tensor(intr) else np.asarray(intr)
            # the frustum's hyper-parameters are host arithmetic on the view (query_point_indices.py:34-98): once, here
            self.hp = model.neural_points.querier.get_hyperparameters(view["h"], view["w"], intr, near, far)
        self.items, self.item_weights = tuple(color_loss_items), tuple(float(w) for w in color_loss_weights)
        assert len(self.items) == len(self.item_weights)
        self.caps, self.use_graph, self.group = (tuple(int(c) for c in caps) if caps is not None else None), bool(graph), group
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized() and not local_only) else 1
        self.dev = model.neural_points.xyz.device
        R, D, C = self.R, int(o.z_depth_dim), int(o.shading_color_channel_num)
        self.jitter = o.is_train > 0 and not (self.frustum and o.shpnt_jitter == "passfunc")
        # static inputs: ONE flat buffer [campos 3 | camrotc2w 9 | raydir R*3 | gt R*C | jitter draw R*D], segments padded to
        # 64 floats - a loader uploads a batch with one copy (load_packed); the graphs read the views.  Frustum querier: the
        # ray segment holds the pixel indices (R*2 int32 bit patterns) and the draw is per kept sample (R*SR)
        SRo = int(o.SR)
        sizes = [3, 9, R * (2 if self.frustum else 3), R * C] + ([R * (SRo if self.frustum else D)] if self.jitter else [])
        offs, n = [], 0
        for sz in sizes:
            offs.append(n)
            n += (sz + 63) // 64 * 64
        self.inbuf = torch.zeros(n, dtype=torch.float32, device=self.dev)
        self.in_offsets, self.in_sizes = offs, sizes
        v = [self.inbuf[a:a + sz] for a, sz in zip(offs, sizes)]
        self.campos, self.rot, self.gt = v[0].view(1, 3), v[1].view(1, 3, 3), v[3].view(1, R, C)
        self.raydir = None if self.frustum else v[2].view(1, R, 3)
        self.pix = v[2].view(torch.int32).view(R, 2) if self.frustum else None
        self.uniform = (v[4].view(1, R, SRo) if self.frustum else v[4].view(1, R, D)) if self.jitter else None
        self.rot.copy_(torch.eye(3, device=self.dev)[None])
        # the batch the step works on: the inputs and query results are PUBLISHED into these persistent buffers by the query
        # graph (or by the query branch at the end of a pipelined step), so that `inbuf` is free for the next batch
        self.cur_in = torch.zeros_like(self.inbuf)
        cv = [self.cur_in[a:a + sz] for a, sz in zip(offs, sizes)]
        self.c_campos, self.c_rot, self.c_gt = cv[0].view(1, 3), cv[1].view(1, 3, 3), cv[3].view(1, R, C)
        self.c_raydir = None if self.frustum else cv[2].view(1, R, 3)
        self.pipeline = bool(pipeline)
        self._primed = False
        self._side = None
        self.masked_count = torch.ones(1, dtype=torch.float32, device=self.dev)      # GLOBAL number of rays with neighbours
        self.counts_host = torch.zeros(4, dtype=torch.int32).pin_memory()           # written by every replay (async copy in the graph)
        self.touched_host = torch.zeros(1, dtype=torch.int32).pin_memory()          # distinct table rows the last backward touched
        self.row_cap = None          # ray sharding: rows per rank in the touched-row exchange (calibrate)
        self._gtab = None            # persistent per-point gradient tables (zero outside the rows the last step touched)
        self._touched_ids = self._n_touched = None
        self._gathered = None        # ray sharding: the all-gathered rows of the last exchange (their ids are cleared next step)
        self.sparse_exchange = bool(sparse_exchange)
        # ray sharding: the gradient exchange runs INSIDE the step (captured with it): the touched-row exchange starts where
        # the backward's data-gradient chain ends (pn_agg_backward_wait_rows), on its own stream, under the trailing
        # weight-gradient GEMMs; the MLP all-reduce follows them.  False: the caller runs reduce() after the step.
        self.exchange_in_step = bool(exchange_in_step) and self.world > 1
        self._comm = None
        self._full_zero = self.world > 1 and not self.sparse_exchange      # the dense all-reduce writes every row: whole-table clears
        self._slots = None
        self.launches_per_step = 0
        self._needs_mask = any(w != 0.0 and (n.startswith("ray_masked") or n.startswith("ray_miss")) for n, w in zip(self.items, self.item_weights))
        self._needs_count = any(w != 0.0 and n.startswith("ray_masked") for n, w in zip(self.items, self.item_weights))
        self._g_query = self._g_step = None
        self.loss = None
        self.overflows = 0
        self.q = {}          # tensors handed from the query graph to the step graph
        self.grads = {}

    # ---- inputs -------------------------------------------------------------------------------------------------------
    def set_inputs(self, cam, raydir, gt, uniform=None, pixel_idx=None):
        """Copies one batch into the static input buffers (device or pinned host tensors; asynchronous).  Frustum querier:
        ``pixel_idx`` int32 [.., R, 2] instead of ``raydir``, ``uniform`` [1,R,SR]."""
        if float(cam["near"]) != self.near or float(cam["far"]) != self.far:
            raise ValueError("near / far are baked into the captured step (%g, %g)" % (self.near, self.far))
        self.campos.copy_(cam["campos"].reshape(1, 3), non_blocking=True)
        self.rot.copy_(cam["camrotc2w"].reshape(1, 3, 3), non_blocking=True)
        if self.frustum:
            if pixel_idx is None:
                raise ValueError("the frustum querier addresses rays by pixel_idx")
            self.pix.copy_(pixel_idx.reshape(self.R, 2).to(torch.int32), non_blocking=True)
        else:
            self.raydir.copy_(raydir.reshape(1, self.R, 3), non_blocking=True)
        self.gt.copy_(gt.reshape(self.gt.shape), non_blocking=True)
        if self.jitter:
            if uniform is None:
                self.uniform.uniform_()
            else:
                self.uniform.copy_(uniform.reshape(self.uniform.shape), non_blocking=True)

    def load_packed(self, flat):
        """One copy of a whole batch laid out like ``inbuf`` (see ``in_offsets`` / ``in_sizes``)."""
        self.inbuf.copy_(flat, non_blocking=True)

    # ---- the launch sequence --------------------------------------------------------------------------------------------
    def _query_body(self):
        m, o = self.model, self.model.opt
        nps = m.neural_points
        xyz = nps.xyz.detach()
        q = {}
        if self.frustum:
            # per-view grid in perspective space -> column walk -> layered K-NN (no host sync), then the reference's torch tail:
            # jitter of the sample depth, pers2w -> world positions and per-sample directions (query_point_indices.py:100-123)
            qr = nps.querier
            raw = qr.query_raw(self.pix, _query.w2pers(xyz, self.rot[0], self.campos[0]), self.hp)
            loc_pers = raw["sample_loc"]
            if self.uniform is not None:
                loc_pers = qr.uniform(loc_pers[None], self.hp["vsize"], uniform=self.uniform)[0]
            loc_w, dirs = qr.pers2w(loc_pers[None], self.rot, self.campos)
            q["pidx"], q["loc_w"], q["loc_pers"], q["dirs"] = raw["sample_pidx"], loc_w[0].contiguous(), loc_pers.contiguous(), dirs[0].contiguous()
            q["ray_any"] = raw["ray_mask"]       # the reference's ray mask on this querier: the ray's column holds an occupied depth
        else:
            raw = nps.querier.query_raw(xyz, self.raydir, self.campos, self.near, self.far, uniform=self.uniform)
            q["pidx"], q["loc_w"], q["ray_any"] = raw["sample_pidx"], raw["sample_loc_w"], raw["ray_any"]
            q["loc_pers"] = _query.w2pers(q["loc_w"], self.rot[0], self.campos[0])
        R, SR, K = q["pidx"].shape
        cs = CompactSamples(q["pidx"], R * SR, K, caps=self.caps)
        q["sample_list"], q["pair_start"], q["pair_src"], q["slot_row"], q["counts"] = cs.sample_list, cs.pair_start, cs.pair_src, cs.slot_row, cs.counts_dev
        if self._needs_count:
            q["masked"] = (q["ray_any"] > 0).sum().to(torch.float32).reshape(1)
        return q

    def _publish(self, t):
        """Query results + the inputs they belong to -> the persistent buffers the step reads."""
        if not self.q or any(self.q[k].shape != v.shape for k, v in t.items() if k != "masked"):
            self.q = {k: torch.empty_like(v) for k, v in t.items() if k != "masked"}       # (first call: outside any capture)
        for k, v in t.items():
            if k == "masked":
                self.masked_count.copy_(v)
            else:
                self.q[k].copy_(v)
        self.cur_in.copy_(self.inbuf)

    def _step_body(self, prefetch=False):
        m, o, q = self.model, self.model.opt, self.q
        nps, agg = m.neural_points, m.aggregator
        pidx = q["pidx"]
        R, SR, K = pidx.shape
        xyz_pers = _query.w2pers(nps.xyz.detach(), self.c_rot[0], self.c_campos[0])
        C = int(o.shading_color_channel_num)
        tabs = (nps.points_embeding, nps.points_color, nps.points_dir, nps.points_conf)
        T = int(nps.points_embeding.shape[-2])
        vsize = self.hp["vsize"] if self.frustum else o.vsize       # the frustum view's own voxel size feeds the distances
        call = _AggCall(agg._params(K, SR, not self.frustum, vsize, True), nps.xyz.detach(), xyz_pers, nps.Rw2c.detach().contiguous(),
                        pidx, q["loc_w"], q["loc_pers"], q["dirs"] if self.frustum else self.c_raydir[0], R * SR, T, True, False)
        call.S, call.M = int(q["sample_list"].numel()), int(q["pair_src"].numel())       # the capacities
        call.counts_dev = q["counts"]
        call.sample_list, call.pair_start, call.pair_src, call.slot_row = q["sample_list"], q["pair_start"], q["pair_src"], q["slot_row"]
        call.collect_touched = True
        # persistent gradient tables, zero outside the rows the previous step (and, sharded, the previous exchange) wrote:
        # those rows are cleared here instead of four whole-table memsets per step (3.1 GB at 20 M points)
        gt_ = self._grad_tables(T, int(nps.points_embeding.shape[-1]), int(q["pair_src"].numel()))
        call.touched_ids, call.n_touched = self._touched_ids, self._n_touched
        lib = _lib.lib()
        F_ = int(nps.points_embeding.shape[-1])
        if self._full_zero:
            for t_ in gt_:
                t_.zero_()
        else:
            _lib.check(lib.pn_rows_zero(_lib.ptr(self._touched_ids), int(self._touched_ids.numel()), 1, F_, _lib.ptr(gt_[0]), _lib.ptr(gt_[3]),
                                        _lib.ptr(gt_[2]), _lib.ptr(gt_[1]), _lib.stream()), "pn_rows_zero")
            if self._gathered is not None:
                _lib.check(lib.pn_rows_zero(_lib.ptr(self._gathered.view(torch.int32)), int(self._gathered.shape[0]), F_ + 8, F_, _lib.ptr(gt_[0]),
                                            _lib.ptr(gt_[3]), _lib.ptr(gt_[2]), _lib.ptr(gt_[1]), _lib.stream()), "pn_rows_zero")
        if self.world > 1:                          # bg_color's gradient rides in the tail of the MLP gradient buffer: one all-reduce
            call.flat_tail = (C + 63) // 64 * 64
        weights = tuple(w.detach() for w in agg._weights())
        tables = tuple(t.detach() for t in tabs)
        # colour loss: per-ray weights realise the reference's three item kinds; the plain MSE needs none.  They depend on the
        # query (and the global masked count) only: eight small torch launches, forked here beside the forward instead of
        # sitting between compositing and the loss
        g_rays = self.R * self.world
        w_fork = None
        if self._needs_mask:
            main0 = torch.cuda.current_stream()
            self._side.wait_stream(main0)
            with torch.cuda.stream(self._side):
                w = color_loss_ray_weights(q["ray_any"], self.items, self.item_weights, C, g_rays, self.masked_count)
                w_fork = torch.cuda.Event()
                w_fork.record(self._side)
            inv_norm = 1.0
        else:
            w, inv_norm = None, sum(self.item_weights) / float(g_rays * C)
        decoded, _, saved, saved_bytes = agg_forward_raw(call, *tables, weights)
        slot_row = q["slot_row"].view(R, SR)
        ray_valid = slot_row >= 0
        ray_dist = ray_dist_from_samples(q["loc_pers"][None], ray_valid[None], vsize, o.raydist_mode_unit)[0]
        bg = self.bg.detach()
        bg_grad = bool(getattr(self.bg, "requires_grad", False))
        n_rows_dev = q["counts"][0:1]
        cctx = _Ctx((True, False, False, bg_grad, False, False, False))
        ray_color, opacity, acc_T, blend_w, bg_T = _Composite.forward(cctx, decoded, ray_dist, _as_u8(ray_valid), bg, True, slot_row, n_rows_dev)
        if w_fork is not None:
            torch.cuda.current_stream().wait_event(w_fork)
            w.record_stream(torch.cuda.current_stream())
        lctx = _Ctx((True, False, False, False, False, False))
        self.loss = _ColorMSE.forward(lctx, ray_color.view(1, R, C), self.c_gt, w, inv_norm)
        nxt = None
        if prefetch:        # the NEXT batch's query (inputs in `inbuf`) on a forked branch, beside this batch's backward
            main = torch.cuda.current_stream()
            self._side.wait_stream(main)
            with torch.cuda.stream(self._side):
                nxt = self._query_body()
        g_feat, _, _, g_bg, _, _, _ = _Composite.backward(cctx, lctx.grad, None, None, None, None)
        need = tuple(bool(t.requires_grad) for t in tabs)
        if not all(need):
            raise NotImplementedError("FusedTrainStep: all four per-point tables are trained (feat / color / dir / conf _grad = 1)")
        g_emb, g_col, g_dir, g_cnf, g_w = agg_backward_raw(call, tables, weights, saved, saved_bytes, g_feat, need, out=gt_)
        self.call = call
        if self.world > 1 and g_bg is not None:
            tail = call.grad_flat[call.flat_tail_off:call.flat_tail_off + C]
            tail.copy_(g_bg)
            g_bg = tail
        self.grads = {"points_embeding": g_emb, "points_color": g_col, "points_dir": g_dir, "points_conf": g_cnf, "bg_color": g_bg}
        if self.exchange_in_step:
            main = torch.cuda.current_stream()
            if self.sparse_exchange:
                if self._comm is None:
                    self._comm = torch.cuda.Stream()
                rc = _lib.lib().pn_agg_backward_wait_rows(self._comm.cuda_stream)     # fork at the end of the data-gradient chain
                if rc < 0:
                    _lib.check(rc, "pn_agg_backward_wait_rows")
                if rc != 1:
                    self._comm.wait_stream(main)
                with torch.cuda.stream(self._comm):
                    self._exchange_rows()
            else:
                self._exchange_dense()
            dist.all_reduce(call.grad_flat, op=dist.ReduceOp.SUM, group=self.group)    # MLP gradients + bg_color (after the weight-gradient GEMMs)
            if self.sparse_exchange:
                main.wait_stream(self._comm)
        self.counts_host.copy_(q["counts"], non_blocking=True)
        if call.n_touched is not None:
            self.touched_host.copy_(call.n_touched, non_blocking=True)
        self.out = {"coarse_raycolor": ray_color.view(1, R, C), "coarse_point_opacity": opacity.view(1, R, SR),
                    "coarse_is_background": bg_T.view(1, R, 1), "ray_any": q["ray_any"].clone(), "slot_row": slot_row.clone()}
        g = {"points_embeding": g_emb, "points_color": g_col, "points_dir": g_dir, "points_conf": g_cnf, "bg_color": g_bg}
        for name, gw in zip(("block1.0.weight", "block1.0.bias", "block1.2.weight", "block1.2.bias", "block3.0.weight", "block3.0.bias",
                             "block3.2.weight", "block3.2.bias", "alpha_branch.0.weight", "alpha_branch.0.bias", "color_branch.0.weight",
                             "color_branch.0.bias", "color_branch.2.weight", "color_branch.2.bias", "color_branch.4.weight",
                             "color_branch.4.bias"), g_w):
            g[name] = gw
        self.grads = g
        if prefetch:        # join; the next batch becomes the current one (after everything that read the current buffers)
            torch.cuda.current_stream().wait_stream(self._side)
            self._publish(nxt)

    def _grad_tables(self, T, F, m_cap):
        if self._gtab is None or self._gtab[0].shape[0] != T or self._touched_ids.numel() != max(m_cap, 1):
            dev = self.dev
            self._gtab = (torch.zeros((T, F), dtype=torch.float32, device=dev), torch.zeros((T, 3), dtype=torch.float32, device=dev),
                          torch.zeros((T, 3), dtype=torch.float32, device=dev), torch.zeros((T,), dtype=torch.float32, device=dev))
            self._touched_ids = torch.full((max(m_cap, 1),), -1, dtype=torch.int32, device=dev)
            self._n_touched = torch.zeros(1, dtype=torch.int32, device=dev)
        return self._gtab

    def _ensure_gathered(self):
        """Ray sharding, touched-row exchange: the persistent all-gather buffer [world * row_cap, F + 8] (id column = -1)."""
        if self.world <= 1 or not self.sparse_exchange:
            return
        F = int(self.model.neural_points.points_embeding.shape[-1])
        if F > 56 or not self.row_cap:
            raise RuntimeError("touched-row exchange: needs calibrate() (row capacity) and point_features_dim <= 56; "
                               "pass sparse_exchange=False for the dense all-reduce")
        rows = self.world * self.row_cap
        if self._gathered is None or tuple(self._gathered.shape) != (rows, F + 8):
            self._gathered = torch.zeros((rows, F + 8), dtype=torch.float32, device=self.dev)
            self._gathered.view(torch.int32)[:, 0] = -1
            self._packed = torch.zeros((self.row_cap, F + 8), dtype=torch.float32, device=self.dev)
        T = int(self.model.neural_points.points_embeding.shape[-2])
        if self._slots is None or self._slots.numel() != T * self.world:       # (allocated here: never inside a capture)
            self._slots = torch.full((T * self.world,), -1, dtype=torch.int32, device=self.dev)

    # ---- capacities -----------------------------------------------------------------------------------------------------
    def calibrate(self, batches, headroom=1.25, granule=4096):
        """Sets the capacities from the exact counts of sample batches (dicts with cam / raydir / uniform as for
        ``set_inputs``): max over the batches times ``headroom``, rounded up to ``granule``.  Host read-backs here only."""
        o, nps = self.model.opt, self.model.neural_points
        s_max = m_max = u_max = 1
        T = int(nps.points_embeding.shape[-2])
        mark = torch.empty(T, dtype=torch.int32, device=self.dev)
        cnt = torch.empty(1, dtype=torch.int32, device=self.dev)
        for b in batches:
            self.set_inputs(b["cam"], b.get("raydir"), b["gt"], b.get("uniform"), pixel_idx=b.get("pixel_idx"))
            if self.frustum:
                raw = {"sample_pidx": nps.querier.query_raw(self.pix, _query.w2pers(nps.xyz.detach(), self.rot[0], self.campos[0]), self.hp)["sample_pidx"]}
            else:
                raw = nps.querier.query_raw(nps.xyz.detach(), self.raydir, self.campos, self.near, self.far, uniform=self.uniform)
            R, SR, K = raw["sample_pidx"].shape
            S, M = CompactSamples(raw["sample_pidx"], R * SR, K).counts()
            s_max, m_max = max(s_max, S), max(m_max, M)
            if self.world > 1:
                _lib.check(_lib.lib().pn_count_touched_rows(_lib.ptr(raw["sample_pidx"]), raw["sample_pidx"].numel(), T, _lib.ptr(mark),
                                                            _lib.ptr(cnt), _lib.stream()), "pn_count_touched_rows")
                u_max = max(u_max, int(cnt.item()))
        up = lambda v: int((int(v * headroom) + granule - 1) // granule * granule)
        caps = [up(s_max), up(m_max), up(u_max)]
        if self.world > 1:                          # the same capacities on every rank (the exchange needs equal row counts)
            t = torch.tensor(caps, dtype=torch.int64, device=self.dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
            caps = [int(v) for v in t.tolist()]
        self.set_caps(caps[:2])
        self.row_cap = min(caps[2], self.caps[1], T)
        return self.caps

    def set_caps(self, caps):
        R, SR, K = self.R, int(self.model.opt.SR), int(self.model.opt.K)
        self.caps = (max(1, min(int(caps[0]), R * SR)), max(K, min(int(caps[1]), R * SR * K)))
        self._g_query = self._g_step = None          # buffers are sized by the capacities: capture again

    def invalidate(self):
        """Call after the point set, the MLP's tensors or bg_color were REPLACED (prune / grow, load_state_dict into new
        storage): the graphs hold their addresses."""
        self._g_query = self._g_step = None

    # ---- capture / replay -----------------------------------------------------------------------------------------------
    def _capture(self):
        if self.caps is None:
            raise RuntimeError("FusedTrainStep: call calibrate(batches) or pass caps=(S_cap, M_cap) first")
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        if self._side is None:
            self._side = torch.cuda.Stream()
        self._ensure_gathered()
        with torch.cuda.stream(side):               # warm-up outside capture: grid build, kernel attributes, library streams
            for _ in range(2):
                self._publish(self._query_body())
                self._step_body(prefetch=self.pipeline)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self._g_query, self._g_step = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        n0 = _lib.launch_count()
        with torch.cuda.graph(self._g_query):
            self._publish(self._query_body())
        n1 = _lib.launch_count()
        with torch.cuda.graph(self._g_step, pool=self._g_query.pool()):
            self._step_body(prefetch=self.pipeline)
        if self.pipeline:                           # a replayed step = forward + loss + backward + the next batch's query
            n0 = n1
        self.launches_per_step = int(_lib.launch_count() - n0)      # kernels of this library inside one replay of the two graphs
        self._bind_grads()

    def _bind_grads(self):
        nps, agg, g = self.model.neural_points, self.model.aggregator, self.grads
        for name in ("points_embeding", "points_color", "points_dir", "points_conf"):
            p = getattr(nps, name)
            if p is not None and p.requires_grad and g.get(name) is not None:
                p.grad = g[name].view(p.shape)
        for name, p in agg.named_parameters():
            p.grad = g[name]
        if isinstance(self.bg, torch.nn.Parameter) and self.bg.requires_grad and g.get("bg_color") is not None:
            self.bg.grad = g["bg_color"].view(self.bg.shape)

    def _check_overflow(self):
        c = self.counts_host              # written by the replays (a value from a step in flight is as good as any for this check)
        if int(c[2]) > int(c[0]) or int(c[3]) > int(c[1]):
            self.overflows += 1
            warnings.warn("FusedTrainStep: a batch held %d valid samples / %d pairs, above the capacities %s: it was truncated "
                          "in slot order; capacities grown" % (int(c[2]), int(c[3]), self.caps))
            self.set_caps((max(self.caps[0], int(int(c[2]) * 1.25) + 4096), max(self.caps[1], int(int(c[3]) * 1.25) + 4096)))
            self.counts_host.zero_()

    def __call__(self, cam=None, raydir=None, gt=None, uniform=None, packed=None, pixel_idx=None):
        """Runs one step on the given batch (or on whatever ``set_inputs`` / ``load_packed`` put into the static buffers).
        Returns the loss (0-d device tensor, static: copy it before the next step if it is read later)."""
        if packed is not None:
            self.load_packed(packed)
        elif raydir is not None or pixel_idx is not None:
            self.set_inputs(cam, raydir, gt, uniform, pixel_idx=pixel_idx)
        if self.pipeline and not self._primed:
            raise RuntimeError("FusedTrainStep(pipeline=True): call prime(first batch) before the first step")
        if not self.use_graph:
            if self.caps is None:
                raise RuntimeError("FusedTrainStep: call calibrate(batches) or pass caps=(S_cap, M_cap) first")
            if self._side is None:
                self._side = torch.cuda.Stream()
            self._ensure_gathered()
            if not self.pipeline:
                self._publish(self._query_body())
                self._sync_masked_count()
            self._step_body(prefetch=self.pipeline)
            if self.pipeline:
                self._sync_masked_count()
            self._bind_grads()
            return self.loss
        self._check_overflow()
        if self._g_step is None:
            self._recapture_keeping_current()
        if not self.pipeline:
            self._g_query.replay()
            self._sync_masked_count()
        self._g_step.replay()
        if self.pipeline:
            self._sync_masked_count()      # the count of the batch just queried normalises the NEXT step's loss
        return self.loss

    def prime(self, cam=None, raydir=None, gt=None, uniform=None, packed=None, pixel_idx=None):
        """pipeline=True: loads the FIRST batch and runs its query; the following ``step(next batch)`` calls work on it."""
        if packed is not None:
            self.load_packed(packed)
        elif raydir is not None or pixel_idx is not None:
            self.set_inputs(cam, raydir, gt, uniform, pixel_idx=pixel_idx)
        if self.use_graph:
            if self._g_step is None:
                self._capture()
            self._g_query.replay()
        else:
            if self.caps is None:
                raise RuntimeError("FusedTrainStep: call calibrate(batches) or pass caps=(S_cap, M_cap) first")
            self._ensure_gathered()
            self._publish(self._query_body())
        self._sync_masked_count()
        self._primed = True

    def _recapture_keeping_current(self):
        """(Re)capture.  The warm-up and the capture run the bodies on whatever `inbuf` holds and publish it; in pipeline mode
        the CURRENT batch (queried by the previous step) must survive that: its inputs are put back and queried again."""
        if not (self.pipeline and self._primed):
            return self._capture()
        keep_next, cur = self.inbuf.clone(), self.cur_in.clone()
        self.inbuf.copy_(cur)
        self._capture()
        self._g_query.replay()             # the current batch again, under the new capacities
        self._sync_masked_count()
        self.inbuf.copy_(keep_next)

    def _sync_masked_count(self):
        if self.world > 1 and self._needs_count:
            dist.all_reduce(self.masked_count, op=dist.ReduceOp.SUM, group=self.group)

    # ---- ray sharding: gradient exchange after the step ----------------------------------------------------------------
    def _exchange_rows(self):
        """Touched-row exchange of the per-point gradient tables on the current stream (see reduce())."""
        g, call, lib = self.grads, self.call, _lib.lib()
        F = int(g["points_embeding"].shape[-1])
        cap = self.row_cap
        packed, gathered = self._packed, self._gathered
        _lib.check(lib.pn_rows_pack(_lib.ptr(call.touched_ids), int(call.touched_ids.numel()), cap, F, _lib.ptr(g["points_embeding"]),
                                    _lib.ptr(g.get("points_conf")), _lib.ptr(g.get("points_dir")), _lib.ptr(g.get("points_color")),
                                    _lib.ptr(packed), _lib.stream()), "pn_rows_pack")
        dist.all_gather_into_tensor(gathered, packed, group=self.group)
        T = int(g["points_embeding"].shape[0])
        if self._slots is None or self._slots.numel() != T * self.world:
            self._slots = torch.full((T * self.world,), -1, dtype=torch.int32, device=self.dev)
        _lib.check(lib.pn_rows_merge(_lib.ptr(gathered), self.world, cap, F, T, _lib.ptr(self._slots), _lib.ptr(g["points_embeding"]),
                                     _lib.ptr(g.get("points_conf")), _lib.ptr(g.get("points_dir")), _lib.ptr(g.get("points_color")),
                                     _lib.stream()), "pn_rows_merge")
        self.exchange_bytes = (self.world - 1) * cap * (F + 8) * 4

    def _exchange_dense(self):
        g = self.grads
        for n in ("points_embeding", "points_color", "points_dir", "points_conf"):
            if g.get(n) is not None:
                dist.all_reduce(g[n], op=dist.ReduceOp.SUM, group=self.group)
        self.exchange_bytes = sum(g[n].numel() * 4 for n in ("points_embeding", "points_color", "points_dir", "points_conf") if g.get(n) is not None)

    def reduce(self):
        """Sum of every gradient over the ranks of ``group``.  Ranks normalise their losses by GLOBAL counts, so the sum is
        the gradient of the whole batch.  MLP gradients + bg_color: ONE all-reduce of the flat buffer the backward wrote
        them into.  Per-point tables: ``sparse_exchange`` - all-gather of the rows this step touched ([id | F + 7 floats],
        row_cap rows per rank from calibrate()) and a rank-ordered merge (pn_rows_pack / pn_rows_merge: identical bits on
        every replica); else a dense all-reduce of the four tables.  With ``exchange_in_step`` (default) all of this already
        ran inside the step - captured in its graph, the row exchange under the trailing weight-gradient GEMMs - and this
        call only checks the row capacity."""
        if self.world <= 1:
            return
        if self.sparse_exchange and int(self.touched_host[0]) > self.row_cap:
            # (a step that touched more rows than row_cap loses the surplus rows' gradients: flagged here, one step late;
            #  calibrate() sizes the capacity with headroom)
            self.overflows += 1
            warnings.warn("FusedTrainStep: %d touched rows exceed the exchange capacity %d" % (int(self.touched_host[0]), self.row_cap))
        if self.exchange_in_step:
            return
        h_flat = dist.all_reduce(self.call.grad_flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        if self.sparse_exchange:
            self._exchange_rows()
        else:
            self._exchange_dense()
        h_flat.wait()


def sharded_gradient_check(model, bg_color, batch, items, weights, group=None):
    """Ray sharding against the single-process step, on every rank of ``group`` (collective: all ranks call it with the SAME
    full batch - dict(cam, raydir [1,R,3], gt [1,R,C], uniform)).  Each rank steps on its contiguous slice of the rays.
    Returns a dict: ``replicas_identical`` (the exchanged gradients are the same bits on every rank), ``sparse_vs_dense``
    (touched-row exchange against the dense all-reduce, relative to the gradient's max; exactly 0 at two ranks, where a sum
    of two terms has one order), ``sharded_vs_single`` (against this rank's own single-process gradient of the WHOLE batch,
    masked-ray normaliser included; the sums run in a different order: <= 1e-4), ``ok``."""
    from .ray_shard import shard_bounds
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    R = int(batch["raydir"].shape[1])
    lo, hi = shard_bounds(R, world, rank)
    if (hi - lo) * world != R:
        raise ValueError("sharded_gradient_check: the ray count must divide by the world size")
    near, far = batch["cam"]["near"], batch["cam"]["far"]
    uni = batch.get("uniform")
    shard = dict(cam=batch["cam"], raydir=batch["raydir"][:, lo:hi].contiguous(), gt=batch["gt"][:, lo:hi].contiguous(),
                 uniform=uni[:, lo:hi].contiguous() if uni is not None else None)
    order = ("points_embeding", "points_color", "points_dir", "points_conf", "bg_color")

    def flat(step):
        g = step.grads
        names = [n for n in order if g.get(n) is not None] + sorted(n for n in g if n not in order)
        return torch.cat([g[n].reshape(-1) for n in names]).clone()

    single = FusedTrainStep(model, bg_color, R, near, far, items, weights, graph=False, local_only=True)
    single.calibrate([batch], headroom=1.0, granule=256)
    single(batch["cam"], batch["raydir"], batch["gt"], uni)
    whole = flat(single)
    res = {}
    for mode in ("sparse", "dense"):
        st = FusedTrainStep(model, bg_color, hi - lo, near, far, items, weights, graph=True, group=group, sparse_exchange=(mode == "sparse"))
        st.calibrate([shard], headroom=1.0, granule=256)
        for _ in range(2):          # twice: the second step runs on gradient tables the first one (and its exchange) wrote
            st(shard["cam"], shard["raydir"], shard["gt"], shard["uniform"])
            st.reduce()
        res[mode] = flat(st)
        if mode == "sparse":
            row_cap, overflows = int(st.row_cap or 0), st.overflows
    scale = float(whole.abs().max())
    e_sd = float((res["sparse"] - res["dense"]).abs().max()) / scale
    e_sw = float((res["sparse"] - whole).abs().max()) / scale
    gathered = [torch.empty_like(res["sparse"]) for _ in range(world)]
    dist.all_gather(gathered, res["sparse"], group=group)
    same = all(torch.equal(t, gathered[0]) for t in gathered)
    ok = same and e_sd <= (0.0 if world == 2 else 1e-6) and e_sw <= 1e-4 and overflows == 0
    flag = torch.tensor([1 if ok else 0], device=whole.device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
    return {"world": world, "rays": R, "replicas_identical": bool(same), "sparse_vs_dense": e_sd, "sharded_vs_single": e_sw,
            "touched_rows_per_rank_cap": row_cap, "loss_items": list(items), "ok": bool(int(flag))}
