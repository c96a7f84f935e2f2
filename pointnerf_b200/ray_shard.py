"""Ray-sharded data parallelism for the per-ray hot path.

One process per GPU.  The neural point tables, the voxel grid (a function of ``xyz`` only, built
redundantly per rank) and the aggregator MLP are REPLICATED; a batch of rays is split into contiguous
slices, one per rank; the forward needs no communication; the backward ends in the gradient
``all_reduce(SUM)``: the four per-point tables (``points_embeding | points_conf | points_dir |
points_color``, N x 39 floats) in place, and the MLP parameters + ``bg_color`` coalesced in one flat buffer.

The reference's loss is a mean over the rays of the local batch (base_rendering_model.py:543-550).
With ray sharding every rank divides its local SUM of squared errors by the GLOBAL number of rays
(times channels), so that the all-reduced gradient equals the single-GPU gradient on the concatenated
batch; nothing else is exchanged.

The reference itself has no working multi-GPU path (``DataParallel`` over one id,
neural_points_volumetric_model.py:173-176; ``train_ddp.py`` does not run) - this module is new.
It is backend-agnostic (``nccl`` on the B200s, ``gloo`` in the CPU tests of the host logic).
"""
import torch
import torch.distributed as dist

from .ray_march import color_mse


def shard_bounds(n_rays, world_size, rank):
    """Contiguous slice [lo, hi) of a batch of ``n_rays`` rays owned by ``rank`` (sizes differ by at most one)."""
    base, rem = divmod(int(n_rays), int(world_size))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class FlatGrads:
    """One contiguous fp32 buffer that aliases the ``.grad`` of every given parameter.

    ``all_reduce()`` is then a single collective; parameters keep reading their own gradient views, so the
    optimiser step needs no unflatten copy.  Parameter order is the caller's (it must be the same on every rank).
    """

    def __init__(self, params):
        self.params = [p for p in params if p is not None and p.requires_grad]
        if not self.params:
            raise ValueError("FlatGrads: no trainable parameters")
        dev = self.params[0].device
        sizes = [p.numel() for p in self.params]
        self.flat = torch.zeros(sum(sizes), dtype=torch.float32, device=dev)
        self.views, off = [], 0
        for p, n in zip(self.params, sizes):
            v = self.flat[off:off + n].view_as(p)
            p.grad = v
            self.views.append(v)
            off += n

    def zero(self):
        self.flat.zero_()
        for p, v in zip(self.params, self.views):  # autograd may have replaced .grad; re-alias
            p.grad = v

    def gather_from_params(self):
        """Copies gradients that autograd allocated outside the flat buffer back into it (no-op when aliased)."""
        for p, v in zip(self.params, self.views):
            if p.grad is None:
                v.zero_()
            elif p.grad.data_ptr() != v.data_ptr():
                v.copy_(p.grad)
                p.grad = v

    def all_reduce(self, group=None):
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)

    @property
    def nbytes(self):
        return self.flat.numel() * 4


class RayShardedStep:
    """Forward + loss + backward (+ gradient all-reduce) of one batch of rays on this rank's shard.

    model:      ``NeuralPointsRayMarching`` (replicated)
    bg_color:   ``nn.Parameter`` [C] or a plain tensor
    The loss is the reference's colour MSE (``compute_losses`` with ``color_loss_items=['coarse_raycolor']``,
    base_rendering_model.py:533-606) normalised by the GLOBAL ray count.

    Gradients: ``.grad`` is reset to None every step, so autograd adopts the buffers the backward kernels wrote (no
    zero-fill and no accumulate pass over the N x 39 point tables).  With more than one rank the large per-point
    tables are all-reduced in place, one collective each, and the ~341 k MLP / bg parameters travel in ONE flat buffer.
    """

    BIG = 1 << 18   # parameters at least this large are reduced in place; smaller ones are coalesced

    def __init__(self, model, bg_color, group=None, sparse=True, side_prefetch=True):
        self.model, self.bg, self.group, self.sparse = model, bg_color, group, sparse
        # the next batch's sampling / neighbour query runs on a SECOND stream, concurrently with this batch's backward: its
        # kernels are small and latency-bound (one block per ray, 27 dependent cell reads), the backward's GEMMs leave issue
        # slots and a few registers per SM free, and the GEMM tails leave whole SMs free
        self._side = torch.cuda.Stream() if (side_prefetch and torch.cuda.is_available()) else None
        self._side_done = None
        self._comm = None            # stream of the touched-row exchange (starts at the end of the backward's data-gradient chain)
        nps = model.neural_points
        params = [nps.points_embeding, nps.points_conf, nps.points_dir, nps.points_color]
        params += list(model.aggregator.parameters())
        if isinstance(bg_color, torch.nn.Parameter):
            params.append(bg_color)
        self.params = [p for p in params if p is not None and p.requires_grad]
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self.nbytes = sum(p.numel() for p in self.params) * 4
        self._cap_ev = self._cap_cpu = None
        self._plan = None
        self._mark = None
        self._slots = None           # slot table of pn_rows_merge (int32 [n_points * world], all -1 between calls)
        self.caps_piggybacked = 0    # exchanges whose capacity arrived with the previous step's all-reduce (introspection)
        self._next_cnt = None        # (plan, device count) of the NEXT batch's touched rows, queued by __call__ after prefetch
        self._cap_next = None        # (plan, event, pinned int32) capacity of the next exchange, produced by this step's all-reduce
        if self.world > 1 and sparse and hasattr(model.aggregator, "forward_tables"):
            model.aggregator.collect_touched = True     # the fused backward returns the sorted unique touched rows
            # tail of the MLP gradient buffer: [gradients outside tables and MLP (bg_color) | one row-count slot per rank]
            n_rest = sum(p.numel() for p in self.params if isinstance(bg_color, torch.nn.Parameter) and p is bg_color)
            self._tail_rest = (n_rest + 63) // 64 * 64
            model.aggregator.flat_tail_floats = self._tail_rest + (self.world + 63) // 64 * 64

    # ---- sparse exchange on the C ABI (CUDA tables) -------------------------------------------------------------
    def _begin_cap_exchange(self):
        """Right after the forward: the number of distinct table rows the batch references (= the touched rows of the
        coming backward) is counted on the device, max-reduced over ranks and copied to a pinned host word - all queued
        BEFORE the backward, so the exchange after it finds the capacity on the host without stalling anything."""
        call = getattr(self.model.aggregator, "last_call", None)
        if call is None or not call.sample_pidx.is_cuda:
            return
        dev = call.sample_pidx.device
        if self._cap_cpu is None:
            self._cap_cpu = torch.zeros(1, dtype=torch.int32).pin_memory()
        cnt = self._count_rows(call.sample_pidx, call.T)
        dist.all_reduce(cnt, op=dist.ReduceOp.MAX, group=self.group)
        self._cap_cpu.copy_(cnt, non_blocking=True)
        self._cap_ev = torch.cuda.Event()
        self._cap_ev.record(torch.cuda.current_stream(dev))

    def _count_rows(self, sample_pidx, T):
        """Device int32[1]: distinct table rows referenced by a query result (pn_count_touched_rows), no host sync."""
        from . import _lib
        dev = sample_pidx.device
        if self._mark is None or self._mark.numel() < T:
            self._mark = torch.empty(max(T, 1), dtype=torch.int32, device=dev)
        cnt = torch.empty(1, dtype=torch.int32, device=dev)
        _lib.check(_lib.lib().pn_count_touched_rows(_lib.ptr(sample_pidx), sample_pidx.numel(), T, _lib.ptr(self._mark),
                                                    _lib.ptr(cnt), _lib.stream()), "pn_count_touched_rows")
        return cnt

    def _sparse_tables_cuda(self, tables_by_name):
        from . import _lib
        lib = _lib.lib()
        call = self.model.aggregator.last_call
        self._cap_ev.synchronize()                          # queued before the backward: complete by now
        cap = max(256, (int(self._cap_cpu[0]) + 255) // 256 * 256)
        emb = tables_by_name["points_embeding"]
        dev, F = emb.grad.device, int(emb.shape[2])
        ids, n_ids = call.touched_ids, (int(call.M) if call.touched_ids is not None else 0)   # n_touched <= M entries, -1 beyond
        g = {k_: (t.grad.contiguous() if t is not None else None) for k_, t in tables_by_name.items()}
        packed = torch.empty((cap, F + 8), dtype=torch.float32, device=dev)
        _lib.check(lib.pn_rows_pack(_lib.ptr(ids), n_ids, cap, F, _lib.ptr(g["points_embeding"]), _lib.ptr(g["points_conf"]),
                                    _lib.ptr(g["points_dir"]), _lib.ptr(g["points_color"]), _lib.ptr(packed), _lib.stream()),
                   "pn_rows_pack")
        gathered = torch.empty((self.world * cap, F + 8), dtype=torch.float32, device=dev)
        dist.all_gather_into_tensor(gathered, packed, group=self.group)
        T = int(emb.shape[1])
        if self._slots is None or self._slots.numel() != T * self.world:
            self._slots = torch.full((T * self.world,), -1, dtype=torch.int32, device=dev)   # once: the merge leaves it all -1
        _lib.check(lib.pn_rows_merge(_lib.ptr(gathered), self.world, cap, F, T, _lib.ptr(self._slots),
                                     _lib.ptr(g["points_embeding"]), _lib.ptr(g["points_conf"]),
                                     _lib.ptr(g["points_dir"]), _lib.ptr(g["points_color"]), _lib.stream()), "pn_rows_merge")
        for k_, t in tables_by_name.items():
            if t is not None and t.grad.data_ptr() != g[k_].data_ptr():
                t.grad = g[k_]
        self.last_cap = cap

    def _sparse_tables(self, tables, pid):
        """Sparse exchange of the per-point gradients: only rows referenced this step are non-zero (<= valid pairs of
        the batch, ~10 % of a 1M-point cloud, ~0.5 % of a 20M-point one), so each rank all-gathers (row ids, rows) of ITS
        touched rows and every rank rebuilds the touched rows as the sum over ranks IN RANK ORDER (identical bits on every
        replica).  One host sync per step (the padded row count, max over ranks).  Bytes per GPU:
        (G-1) * U * 4*(39+1) against 2*(G-1)/G * 4*39*N for the dense all-reduce."""
        G, dev = self.world, tables[0].grad.device
        N = tables[0].shape[1]
        mark = torch.zeros(N, dtype=torch.int32, device=dev)
        if pid.numel():
            mark[pid] = 1
        pos = torch.cumsum(mark, 0, dtype=torch.int64) - 1
        umax = (pos[-1] + 1).reshape(1).clone()
        dist.all_reduce(umax, op=dist.ReduceOp.MAX, group=self.group)
        cap = max(1024, (int(umax.item()) + 1023) // 1024 * 1024)             # the host sync
        ids = torch.full((cap + 1,), -1, dtype=torch.int64, device=dev)
        ids.scatter_(0, torch.where(mark > 0, pos, torch.full_like(pos, cap)), torch.arange(N, device=dev))
        ids = ids[:cap]
        safe = ids.clamp(min=0)
        live = (ids >= 0).to(torch.float32)[:, None]
        widths = [t.shape[2] for t in tables]
        rows = torch.cat([t.grad[0].index_select(0, safe) for t in tables], dim=1) * live          # [cap, 39]
        rows_all = torch.empty((G * cap, rows.shape[1]), dtype=torch.float32, device=dev)
        ids_all = torch.empty((G * cap,), dtype=torch.int64, device=dev)
        h1 = dist.all_gather_into_tensor(rows_all, rows.contiguous(), group=self.group, async_op=True)
        h2 = dist.all_gather_into_tensor(ids_all, ids.contiguous(), group=self.group, async_op=True)
        for t in tables:                       # own contribution is re-added from the gathered buffer, in rank order
            t.grad[0].index_fill_(0, safe, 0.0)
        h1.wait()
        h2.wait()
        for r in range(G):
            idr = ids_all[r * cap:(r + 1) * cap].clamp(min=0)                  # unique inside a rank: deterministic
            blk = rows_all[r * cap:(r + 1) * cap]
            off = 0
            for t, wd in zip(tables, widths):
                t.grad[0].index_add_(0, idr, blk[:, off:off + wd])
                off += wd

    def all_reduce(self):
        if self.world <= 1:
            return
        for p in self.params:
            if p.grad is None:
                p.grad = torch.zeros_like(p)
        nps = self.model.neural_points
        names = ("points_embeding", "points_conf", "points_dir", "points_color")
        tables = [t for t in (getattr(nps, n) for n in names) if t is not None and t.requires_grad]
        tset = {id(t) for t in tables}
        # the fused backward leaves every MLP gradient inside one flat buffer: reduce that buffer in place
        flat = getattr(getattr(self.model.aggregator, "last_call", None), "grad_flat", None)
        lo, hi = (flat.data_ptr(), flat.data_ptr() + 4 * flat.numel()) if flat is not None else (0, 0)
        rest = [p for p in self.params if id(p) not in tset and not (lo <= p.grad.data_ptr() < hi)]
        mode, tail = None, None
        # (pn_rows_pack / pn_rows_merge carry one row in <= 64 columns: F + 8 <= 64; wider tables take the dense all-reduce)
        if (self.sparse and tables and tables[0].is_cuda and nps.points_embeding.requires_grad and self._cap_ev is not None
                and int(nps.points_embeding.shape[-1]) <= 56):
            mode = "cuda"        # ids from the fused backward, pack / merge kernels of the C ABI
            call = self.model.aggregator.last_call
            n_rest = sum(p.numel() for p in rest)
            if flat is not None and n_rest <= self._tail_rest and call.flat_tail >= self._tail_rest + self.world:
                # the few gradients outside the tables and the MLP (bg_color) and the NEXT batch's touched-row count ride in
                # the zeroed tail of the MLP gradient buffer: one all-reduce(SUM) instead of three small collectives
                tail = off = call.flat_tail_off
                for p in rest:
                    flat[off:off + p.numel()].copy_(p.grad.reshape(-1))
                    off += p.numel()
                if self._next_cnt is not None:          # one slot per rank: SUM leaves every rank's count in place
                    if self._side_done is not None:     # the count was produced on the prefetch stream
                        torch.cuda.current_stream().wait_event(self._side_done)
                    slot = tail + self._tail_rest + dist.get_rank(self.group)
                    flat[slot:slot + 1].copy_(self._next_cnt[1])
            elif rest:           # one small all-gather for them
                head = torch.cat([p.grad.reshape(-1) for p in rest])
                self._head_all = torch.empty((self.world, head.numel()), dtype=torch.float32, device=head.device)
                dist.all_gather_into_tensor(self._head_all, head, group=self.group)
        elif self.sparse and hasattr(self.model, "touched_point_ids") and tables and not tables[0].is_cuda:
            pid = self.model.touched_point_ids()     # device-agnostic torch formulation (CPU tests of the host logic)
            if pid is not None and all(t.dim() == 3 and t.shape[0] == 1 for t in tables):
                self._sparse_tables(tables, pid)
                mode = "torch"
        handles = []
        joined = None
        if mode == "cuda":
            # the touched rows are final at the end of the backward's data-gradient chain; the weight-gradient GEMMs still
            # trail on this stream: pack / all-gather / merge run on their own stream, under those GEMMs.  Issued BEFORE the
            # MLP all-reduce: the process group runs collectives in issue order, and that one has to wait for the GEMMs
            from . import _lib
            main = torch.cuda.current_stream()
            if self._comm is None:
                self._comm = torch.cuda.Stream()
            rc = _lib.lib().pn_agg_backward_wait_rows(self._comm.cuda_stream)
            if rc < 0:
                _lib.check(rc, "pn_agg_backward_wait_rows")
            if rc != 1:
                self._comm.wait_stream(main)
            with torch.cuda.stream(self._comm):
                self._sparse_tables_cuda({n: (getattr(nps, n) if getattr(nps, n) is not None and getattr(nps, n).requires_grad else None)
                                          for n in names})
            joined = (main, self._comm)
        if flat is not None and hi > lo:
            handles.append(dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True))
        if mode is None:
            handles += [dist.all_reduce(t.grad, op=dist.ReduceOp.SUM, group=self.group, async_op=True) for t in tables]
        if mode == "cuda":
            if rest and tail is None:   # rank-ordered sum of the gathered head columns: identical bits on every replica
                tot = self._head_all.sum(0)
                off = 0
                for p in rest:
                    p.grad.copy_(tot[off:off + p.numel()].view_as(p))
                    off += p.numel()
        elif rest:
            big = [p for p in rest if p.numel() >= self.BIG]
            small = [p for p in rest if p.numel() < self.BIG]
            handles += [dist.all_reduce(p.grad, op=dist.ReduceOp.SUM, group=self.group, async_op=True) for p in big]
            if small:
                buf = torch.cat([p.grad.reshape(-1) for p in small])
                dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=self.group)
                off = 0
                for p in small:
                    p.grad.copy_(buf[off:off + p.numel()].view_as(p))
                    off += p.numel()
        for h in handles:
            h.wait()
        if joined is not None:
            joined[0].wait_stream(joined[1])
        if tail is not None:
            off = tail
            for p in rest:       # the reduced values, in place in the tail (same bits on every replica)
                p.grad = flat[off:off + p.numel()].view_as(p)
                off += p.numel()
            if self._next_cnt is not None:
                # capacity of the NEXT exchange = max over ranks of the counts just summed into their slots: on its way to
                # the host a whole step before it is needed, and the 4-byte MAX all-reduce of that step is not issued
                a = tail + self._tail_rest
                if self._cap_cpu is None:
                    self._cap_cpu = torch.zeros(1, dtype=torch.int32).pin_memory()
                self._cap_cpu.copy_(flat[a:a + self.world].max().to(torch.int32).reshape(1), non_blocking=True)
                ev = torch.cuda.Event()
                ev.record()
                self._cap_next = (self._next_cnt[0], ev)
        self._next_cnt = None

    def prefetch(self, cam, raydir, uniform=None):
        """Queue the weight-independent part of a FUTURE batch (ray sampling, neighbour query, work lists) now.  Called by
        ``__call__(next_batch=...)`` between the forward and the backward of the current batch: the GPU runs it there, and
        the (S, M) read-back the next forward needs has completed long before the host asks for it - no idle gap."""
        if not hasattr(self.model, "sample"):
            return
        if self._side is not None and raydir.is_cuda:
            main = torch.cuda.current_stream()
            self._side.wait_stream(main)             # inputs are ready; every block freed so far is safe to reuse over there
            with torch.cuda.stream(self._side):
                plan = self.model.sample(cam["campos"], raydir, cam["camrotc2w"], cam["near"], cam["far"], uniform=uniform)
                self._side_done = torch.cuda.Event()
                self._side_done.record(self._side)
            # no record_stream on the plan's tensors: they are allocated from the side stream's pool, and the only later
            # allocations from that pool happen in a later prefetch, after its wait_stream(main) above - i.e. after every
            # main-stream kernel that was queued while those tensors were alive
        else:
            plan = self.model.sample(cam["campos"], raydir, cam["camrotc2w"], cam["near"], cam["far"], uniform=uniform)
        self._plan = (raydir, uniform, plan)

    def _take_plan(self, raydir, uniform):
        held, self._plan = self._plan, None
        done, self._side_done = self._side_done, None
        if done is not None:
            torch.cuda.current_stream().wait_event(done)      # whatever the side stream queued is ordered before this step
        if held is not None and held[0] is raydir and held[1] is uniform:
            return held[2]
        return None

    def __call__(self, cam, raydir, gt, uniform=None, global_rays=None, reduce=True, next_batch=None):
        """cam: dict(campos [1,3], camrotc2w [1,3,3], near, far); raydir [1,R_local,3]; gt [1,R_local,C].
        next_batch: optional dict(cam=, raydir=, uniform=[, ready=cuda event]) of the batch the NEXT call will get (see
        ``prefetch``); ``ready`` lets the caller upload it on a side stream while this batch computes.
        Returns (loss_local_sum / (global_rays*C)) as a 0-d tensor (the all-reduced SUM over ranks is the batch MSE)."""
        R_local = raydir.shape[1]
        global_rays = int(global_rays if global_rays is not None else R_local * self.world)
        for p in self.params:
            p.grad = None
        plan = self._take_plan(raydir, uniform)
        kw = dict(plan=plan) if plan is not None else {}
        out = self.model(cam["campos"], raydir, bg_color=self.bg, camrotc2w=cam["camrotc2w"], near=cam["near"], far=cam["far"],
                         uniform=uniform, **kw)
        self._cap_ev = None
        if self.world > 1 and self.sparse and reduce:   # (reduce=False: a rank-local step, e.g. per-stage timing on one rank, issues NO collective)
            held, self._cap_next = self._cap_next, None
            if held is not None and plan is not None and held[0] is plan:
                self._cap_ev = held[1]                  # agreed during the previous step's all-reduce
                self.caps_piggybacked += 1
            else:
                self._begin_cap_exchange()
        pred = out["coarse_raycolor"]
        if pred.is_cuda:               # one fused kernel pair: loss and d(loss)/d(colour) (pn_color_mse)
            loss = color_mse(pred, gt, norm=global_rays * gt.shape[-1])
        else:                          # gloo tests of this module's host logic drive it with a CPU toy model
            diff = pred - gt
            loss = (diff * diff).sum() / float(global_rays * gt.shape[-1])
        if next_batch is not None:
            if next_batch.get("ready") is not None:        # event of the side-stream H2D copy that produced the batch
                (self._side if self._side is not None else torch.cuda.current_stream()).wait_event(next_batch["ready"])
            self.prefetch(next_batch["cam"], next_batch["raydir"], next_batch.get("uniform"))
            if self.world > 1 and self.sparse and reduce and self._plan is not None and self._cap_ev is not None:
                nxt = self._plan[2]
                T = int(self.model.neural_points.points_embeding.shape[-2])
                if self._side is not None and self._side_done is not None:     # behind the query, on its stream
                    with torch.cuda.stream(self._side):
                        cnt = self._count_rows(nxt.pidx, T)
                        self._side_done = torch.cuda.Event()
                        self._side_done.record(self._side)
                else:
                    cnt = self._count_rows(nxt.pidx, T)
                self._next_cnt = (nxt, cnt)
        loss.backward()
        if reduce:
            self.all_reduce()
        return loss.detach()

    def flat_grads(self):
        """All gradients as one vector (parameter order of the constructor); for tests and checks."""
        return torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1) for p in self.params])
