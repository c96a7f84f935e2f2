"""Drop-in ``PointAggregator`` on the C ABI (pn_agg_forward / pn_agg_backward).

Mirrors models/aggregators/point_aggregators.py of the reference: ``PointAggregator(opt)`` is an
``nn.Module`` with the same parameter names and shapes (``block1.{0,2}``, ``block3.{0,2}``,
``alpha_branch.0``, ``color_branch.{0,2,4}``; :276-348), the same xavier initialisation (``init_seq``,
models/helpers/networks.py:163-172) and the same ``forward`` signature and return structure (:729-816),
so reference checkpoints load unchanged and ``NeuralPointsRayMarching.forward`` can call it as is.

Only the configurations the shipped scripts use are implemented (viewmlp, agg_intrp_order=2 or 1, linear
distance kernel, agg_dist_pers=20, apply_pnt_mask=1, LeakyReLU, act_super=1); anything else raises.

Two entry points share one autograd function:
  * ``PointAggregator.forward(...13 tensors...)``   - the reference API: already gathered
    ``[1,R,SR,K,*]`` tensors; rows are addressed with identity indices;
  * ``PointAggregator.forward_tables(...)``          - the fused gather path: the neural point tables and
    ``sample_pidx`` go straight to the kernels (no ``[R,SR,K,38]`` intermediate, no N x 38 ``torch.cat``),
    and the per-point gradients come back through a sorted segmented reduction instead of atomic
    ``index_add_`` (the backward of ``index_select`` in neural_points.py:791-797).
There is no CPU path: CPU tensors raise.
"""
import ctypes
import math

import numpy as np
import torch
import torch.nn as nn

from . import _lib
from ._lib import AggInputs, AggParams, AggWeights, WEIGHT_FIELDS


def _check_opt(opt):
    def need(name, ok, got=None):
        v = getattr(opt, name, None) if got is None else got
        if not ok(v):
            raise NotImplementedError("pointnerf_b200.PointAggregator: %s=%r is outside the hot path (SURVEY.md 2 #4)" % (name, v))

    need("which_agg_model", lambda v: v == "viewmlp")
    need("agg_intrp_order", lambda v: v in (1, 2))
    need("agg_distance_kernel", lambda v: v == "linear")
    need("agg_dist_pers", lambda v: v == 20)
    need("apply_pnt_mask", lambda v: v > 0)
    need("agg_weight_norm", lambda v: v > 0)
    need("act_type", lambda v: v == "LeakyReLU")
    need("act_super", lambda v: v > 0)
    need("shading_feature_mlp_layer0", lambda v: v in (0, 1))  # never built by viewmlp_init
    need("shading_feature_mlp_layer1", lambda v: v == 2)
    need("shading_feature_mlp_layer2", lambda v: v == 0)
    need("shading_feature_mlp_layer3", lambda v: v == 2)
    need("shading_alpha_mlp_layer", lambda v: v == 1)
    need("shading_color_mlp_layer", lambda v: v == 4)
    need("agg_feat_xyz_mode", lambda v: v == "None")
    need("agg_alpha_xyz_mode", lambda v: v == "None")
    need("agg_color_xyz_mode", lambda v: v == "None")
    need("point_color_mode", lambda v: "1" in list(v))
    need("point_dir_mode", lambda v: "1" in list(v))
    need("view_ori", lambda v: v == 0)
    need("num_feat_freqs", lambda v: v > 0)
    need("dist_xyz_freq", lambda v: v > 0)
    need("num_viewdir_freqs", lambda v: v > 0)
    aw = getattr(opt, "agg_axis_weight", None)
    if aw is not None and tuple(float(a) for a in aw) != (1.0, 1.0, 1.0):
        raise NotImplementedError("agg_axis_weight != (1,1,1) is outside the hot path")


def _xavier_like_reference(seq):
    """init_seq (helpers/networks.py:163-172): xavier_uniform with the gain of the FOLLOWING activation,
    zero bias; the last module of a Sequential gets gain 1 (it is the activation itself for block1/3 and
    the colour branch, a no-op there)."""
    mods = list(seq)
    for a, b in zip(mods[:-1], mods[1:]):
        if isinstance(a, nn.Linear):
            gain = nn.init.calculate_gain("leaky_relu", b.negative_slope) if isinstance(b, nn.LeakyReLU) else 1.0
            nn.init.xavier_uniform_(a.weight, gain=gain)
            nn.init.zeros_(a.bias)
    if isinstance(mods[-1], nn.Linear):
        nn.init.xavier_uniform_(mods[-1].weight, gain=1.0)
        nn.init.zeros_(mods[-1].bias)


class _AggCall:
    """Non-differentiable context of one aggregator call (plain Python object handed to the autograd function)."""

    def __init__(self, params, xyz, xyz_pers, Rw2c, sample_pidx, sample_loc_w, sample_loc, ray_dirs, n_slots, T,
                 scatter_rows, want_weight):
        self.params, self.xyz, self.xyz_pers, self.Rw2c = params, xyz, xyz_pers, Rw2c
        self.sample_pidx, self.sample_loc_w, self.sample_loc, self.ray_dirs = sample_pidx, sample_loc_w, sample_loc, ray_dirs
        self.n_slots, self.T, self.scatter_rows, self.want_weight = n_slots, T, scatter_rows, want_weight
        self.S = self.M = 0
        self.counts_dev = None                # device int32[2] (S, M): set when the host does not read the counts back
        self.sample_list = self.pair_start = self.pair_src = self.slot_row = None
        self.collect_touched = False          # ask the backward for the sorted unique list of table rows it touched
        self.touched_ids = self.n_touched = None
        self.grad_flat = None                 # backward: the MLP weight gradients, one contiguous buffer
        self.flat_tail, self.flat_tail_off = 0, 0   # extra zeroed floats appended to grad_flat for the caller, and where they start
        self.compact = None                   # optional CompactSamples built ahead of the call
        self.export_signs = False             # checker hook: keep the forward's LeakyReLU pattern (pn_agg_activation_signs)
        self.signs = None                     # (uint8 [4,M,W] pair layers, uint8 [3,S,C] colour layers)


def _weights_struct(tensors):
    w = AggWeights()
    for name, t in zip(WEIGHT_FIELDS, tensors):
        setattr(w, name, _lib.ptr(t))
    return w


def _inputs_struct(call, embedding, color, pdir, conf):
    a = AggInputs()
    a.xyz, a.xyz_pers, a.embedding = _lib.ptr(call.xyz), _lib.ptr(call.xyz_pers), _lib.ptr(embedding)
    a.color, a.dir, a.conf, a.Rw2c = _lib.ptr(color), _lib.ptr(pdir), _lib.ptr(conf), _lib.ptr(call.Rw2c)
    a.sample_pidx, a.sample_loc_w, a.sample_loc = _lib.ptr(call.sample_pidx), _lib.ptr(call.sample_loc_w), _lib.ptr(call.sample_loc)
    a.ray_dirs = _lib.ptr(call.ray_dirs)
    a.sample_list, a.pair_start, a.pair_src = _lib.ptr(call.sample_list), _lib.ptr(call.pair_start), _lib.ptr(call.pair_src)
    a.n_slots, a.S, a.M, a.T = call.n_slots, call.S, call.M, call.T
    a.counts = _lib.ptr(call.counts_dev)     # None: S / M are exact; else they are capacities and the counts stay on the device
    return a


class CompactSamples:
    """Work lists of valid samples / pairs (pn_compact_samples).  The kernels and the device->host copy of the two counts
    are queued at construction; ``counts()`` waits for that copy only - a caller that builds the lists early (the next
    batch's sampling, queued before this batch's backward) never stalls the GPU on the read-back.

    ``caps=(S_cap, M_cap)``: capacity mode (pn_compact_samples_capped) - nothing is copied to the host, ``counts_dev``
    (int32[4]: S kept, M kept, S total, M total) stays on the device, ``counts()`` returns the capacities."""

    def __init__(self, sample_pidx, n_slots, K, caps=None):
        lib = _lib.lib()
        dev = sample_pidx.device
        self.n_slots = n_slots
        self._sm = None
        self.counts_dev = None
        if n_slots == 0:
            z = torch.zeros(1, dtype=torch.int32, device=dev)
            self.sample_list, self.pair_start, self.pair_src = z, z.clone(), z.clone()
            self.slot_row = torch.zeros(0, dtype=torch.int32, device=dev)
            self._sm = (0, 0)
            return
        ws_bytes = lib.pn_compact_workspace_bytes(n_slots)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        self.slot_row = torch.empty(n_slots, dtype=torch.int32, device=dev)
        if caps is not None:
            S_cap, M_cap = min(int(caps[0]), n_slots), min(int(caps[1]), n_slots * K)
            self.sample_list = torch.empty(S_cap, dtype=torch.int32, device=dev)
            self.pair_start = torch.empty(S_cap + 1, dtype=torch.int32, device=dev)
            self.pair_src = torch.empty(M_cap, dtype=torch.int32, device=dev)
            self.counts_dev = torch.empty(4, dtype=torch.int32, device=dev)
            _lib.check(lib.pn_compact_samples_capped(_lib.ptr(sample_pidx), n_slots, K, S_cap, M_cap, _lib.ptr(self.sample_list),
                                                     _lib.ptr(self.pair_start), _lib.ptr(self.pair_src), _lib.ptr(self.counts_dev),
                                                     _lib.ptr(self.slot_row), _lib.ptr(ws), ws_bytes, _lib.stream()),
                       "pn_compact_samples_capped")
            self._sm = (S_cap, M_cap)
            return
        self.sample_list = torch.empty(max(n_slots, 1), dtype=torch.int32, device=dev)
        self.pair_start = torch.empty(n_slots + 1, dtype=torch.int32, device=dev)
        self.pair_src = torch.empty(max(n_slots * K, 1), dtype=torch.int32, device=dev)
        counts = torch.empty(2, dtype=torch.int32, device=dev)
        _lib.check(lib.pn_compact_samples(_lib.ptr(sample_pidx), n_slots, K, _lib.ptr(self.sample_list), _lib.ptr(self.pair_start),
                                          _lib.ptr(self.pair_src), _lib.ptr(counts), _lib.ptr(self.slot_row), _lib.ptr(ws), ws_bytes,
                                          _lib.stream()), "pn_compact_samples")
        self._host = torch.empty(2, dtype=torch.int32).pin_memory()
        self._host.copy_(counts, non_blocking=True)
        self._event = torch.cuda.Event()
        self._event.record()

    def counts(self):
        """(S, M): valid samples and valid (sample, neighbour) pairs - the one host read-back of the aggregator (capacity
        mode: the capacities, no read-back)."""
        if self._sm is None:
            self._event.synchronize()
            self._sm = (int(self._host[0]), int(self._host[1]))
        return self._sm


def compact_samples(sample_pidx, n_slots, K):
    """Synchronous form: (sample_list, pair_start, pair_src, S, M, slot_row)."""
    c = CompactSamples(sample_pidx, n_slots, K)
    S, M = c.counts()
    return c.sample_list, c.pair_start, c.pair_src, S, M, c.slot_row


def _shaped(grads, shapes):
    return tuple(None if g is None else g.view(sh) for g, sh in zip(grads, shapes))


def agg_forward_raw(call, embedding, color, pdir, conf, weights):
    """pn_agg_forward on already detached, contiguous tensors.  ``call.S`` / ``call.M`` are the exact counts, or - with
    ``call.counts_dev`` set - the capacities every buffer is sized by.  Returns (decoded, weight_out | None, saved, saved_bytes)."""
    lib = _lib.lib()
    p = call.params
    dev = embedding.device
    decoded = torch.empty((call.S if p.compact_out else call.n_slots, p.C + 1), dtype=torch.float32, device=dev)
    weight_out = torch.empty((call.n_slots, p.K), dtype=torch.float32, device=dev) if call.want_weight else None
    saved_bytes = lib.pn_agg_saved_bytes(ctypes.byref(p), call.S, call.M)
    ws_bytes = lib.pn_agg_workspace_bytes(ctypes.byref(p), call.S, call.M, call.T, 0)
    if saved_bytes < 0 or ws_bytes < 0:
        raise _lib.LibraryError("pn_agg_*_bytes: " + lib.pn_last_error().decode())
    saved = torch.empty(max(saved_bytes, 1), dtype=torch.uint8, device=dev)
    ws = torch.empty(max(ws_bytes, 1), dtype=torch.uint8, device=dev)
    ain = _inputs_struct(call, embedding, color, pdir, conf)
    wst = _weights_struct(weights)
    if call.n_slots > 0:
        _lib.check(lib.pn_agg_forward(ctypes.byref(p), ctypes.byref(ain), ctypes.byref(wst), _lib.ptr(decoded),
                                      _lib.ptr(weight_out), _lib.ptr(saved), saved_bytes, _lib.ptr(ws), ws_bytes,
                                      _lib.stream()), "pn_agg_forward")
        if call.export_signs:
            if call.counts_dev is not None:
                raise _lib.LibraryError("export_signs needs the counts on the host")
            ps = torch.empty((4, call.M, p.W), dtype=torch.uint8, device=dev)
            ss = torch.empty((3, call.S, p.C), dtype=torch.uint8, device=dev)
            _lib.check(lib.pn_agg_activation_signs(ctypes.byref(p), _lib.ptr(saved), saved_bytes, call.S, call.M, call.S, call.M,
                                                   _lib.ptr(ps), _lib.ptr(ss), _lib.stream()), "pn_agg_activation_signs")
            call.signs = (ps, ss)
    return decoded, weight_out, saved, saved_bytes


def agg_backward_raw(call, tables, weights, saved, saved_bytes, g_decoded, need=(True, True, True, True), out=None):
    """pn_agg_backward.  Returns (g_emb, g_col, g_dir, g_cnf, g_w): per-row gradients ([T,*] with scatter_rows, else
    [n_slots*K,*]; None where ``need`` is False) and the MLP gradients as views of ONE flat buffer (``call.grad_flat``).
    ``out`` = (g_emb, g_col, g_dir, g_cnf): persistent [T,*] tables the caller keeps zero outside the rows a step touches
    (pn_rows_zero) - they are written without the whole-table memsets (scatter_rows = 2); ``call.touched_ids`` /
    ``call.n_touched`` may then be preset persistent buffers too."""
    lib = _lib.lib()
    p = call.params
    embedding, color, pdir, conf = tables
    dev = embedding.device
    rows = call.T if call.scatter_rows else call.n_slots * p.K
    if out is not None:
        if not call.scatter_rows:
            raise ValueError("persistent gradient tables need scatter_rows")
        g_emb, g_col, g_dir, g_cnf = out
    else:
        g_emb = torch.empty((rows, p.F), dtype=torch.float32, device=dev) if need[0] else None
        g_col = torch.empty((rows, 3), dtype=torch.float32, device=dev) if need[1] else None
        g_dir = torch.empty((rows, 3), dtype=torch.float32, device=dev) if need[2] else None
        g_cnf = torch.empty((rows,), dtype=torch.float32, device=dev) if need[3] else None
    # all MLP gradients live in ONE flat buffer (256-byte aligned segments): ray_shard all-reduces it in place
    sizes = [(w.numel() + 63) // 64 * 64 for w in weights]
    call.flat_tail_off = sum(sizes)       # + a caller-owned tail (ray_shard: bg gradient and row counts ride the same all-reduce)
    call.grad_flat = torch.zeros(call.flat_tail_off + int(call.flat_tail), dtype=torch.float32, device=dev)
    g_w = tuple(seg[:w.numel()].view(w.shape) for seg, w in zip(call.grad_flat.split(sizes + [int(call.flat_tail)]), weights))
    if call.n_slots == 0:
        for t in (g_emb, g_col, g_dir, g_cnf):
            if t is not None:
                t.zero_()
        return g_emb, g_col, g_dir, g_cnf, g_w
    ws_bytes = lib.pn_agg_workspace_bytes(ctypes.byref(p), call.S, call.M, call.T, 1)
    ws = torch.empty(max(ws_bytes, 1), dtype=torch.uint8, device=dev)
    ain = _inputs_struct(call, embedding, color, pdir, conf)
    wst, gst = _weights_struct(weights), _weights_struct(g_w)
    if call.collect_touched and call.scatter_rows and call.touched_ids is None:
        call.touched_ids = torch.empty(max(call.M, 1), dtype=torch.int32, device=dev)
        call.n_touched = torch.empty(1, dtype=torch.int32, device=dev)
    _lib.check(lib.pn_agg_backward(ctypes.byref(p), ctypes.byref(ain), ctypes.byref(wst), _lib.ptr(g_decoded),
                                   _lib.ptr(saved), saved_bytes, ctypes.byref(gst), (2 if out is not None else 1) if call.scatter_rows else 0,
                                   _lib.ptr(g_emb), _lib.ptr(g_col), _lib.ptr(g_dir), _lib.ptr(g_cnf),
                                   _lib.ptr(call.touched_ids), _lib.ptr(call.n_touched), _lib.ptr(ws),
                                   ws_bytes, _lib.stream()), "pn_agg_backward")
    return g_emb, g_col, g_dir, g_cnf, g_w


class _AggFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, call, embedding, color, pdir, conf, *weights):
        p = call.params
        cs = call.compact if call.compact is not None else CompactSamples(call.sample_pidx, call.n_slots, p.K)
        call.S, call.M = cs.counts()
        call.sample_list, call.pair_start, call.pair_src, call.slot_row = cs.sample_list, cs.pair_start, cs.pair_src, cs.slot_row
        weights = tuple(w.detach().contiguous() for w in weights)
        ctx.table_shapes = tuple(t.shape for t in (embedding, color, pdir, conf))
        embedding, color, pdir, conf = (t.detach().contiguous() for t in (embedding, color, pdir, conf))
        decoded, weight_out, saved, saved_bytes = agg_forward_raw(call, embedding, color, pdir, conf, weights)
        ctx.call, ctx.saved, ctx.saved_bytes = call, saved, saved_bytes
        ctx.tables = (embedding, color, pdir, conf)
        ctx.weights = weights
        if weight_out is None:
            weight_out = torch.empty(0, device=embedding.device)
        ctx.mark_non_differentiable(weight_out)
        ctx.set_materialize_grads(False)      # no zero tensor for the non-differentiable weight output
        return decoded, weight_out

    @staticmethod
    def backward(ctx, g_decoded, _g_weight):
        call, p = ctx.call, ctx.call.params
        dev = ctx.tables[0].device
        if g_decoded is None:
            g_decoded = torch.zeros((call.S if p.compact_out else call.n_slots, p.C + 1), dtype=torch.float32, device=dev)
        need = ctx.needs_input_grad  # (call, embedding, color, dir, conf, *weights)
        g_emb, g_col, g_dir, g_cnf, g_w = agg_backward_raw(call, ctx.tables, ctx.weights, ctx.saved, ctx.saved_bytes,
                                                           g_decoded.contiguous(), need[1:5])
        if call.n_slots == 0:
            for t in g_w:
                t.zero_()
        return (None,) + _shaped((g_emb, g_col, g_dir, g_cnf), ctx.table_shapes) + tuple(g if n else None for g, n in zip(g_w, need[5:]))


class _GatherClamp(torch.autograd.Function):
    """conf_coefficient of the fused path: clamp(table[max(idx,0)], lo, hi) with the straight-through gradient of
    ``gradiant_clamp`` (point_aggregators.py:724-726, :809-811) scattered back to the table rows."""

    @staticmethod
    def forward(ctx, table, idx, lo, hi):
        lib = _lib.lib()
        flat = table.detach().reshape(-1).contiguous().float()
        idx = idx.contiguous()
        out = torch.empty(idx.shape, dtype=torch.float32, device=flat.device)
        _lib.check(lib.pn_gather_clamp(_lib.ptr(flat), int(flat.numel()), _lib.ptr(idx), int(idx.numel()), float(lo), float(hi),
                                       _lib.ptr(out), _lib.stream()), "pn_gather_clamp")
        ctx.idx, ctx.shape = idx, tuple(table.shape)
        return out

    @staticmethod
    def backward(ctx, g):
        n = 1
        for v in ctx.shape:
            n *= v
        gt = torch.zeros(n, dtype=torch.float32, device=g.device)
        gt.index_put_((ctx.idx.reshape(-1).clamp(min=0).long(),), g.reshape(-1).float(), accumulate=True)   # sort-based: deterministic
        return gt.view(ctx.shape), None, None, None


def gather_clamp(table, idx, lo=0.0001, hi=1.0):
    """table [..N..] (any shape with N elements), idx int32 [...] (-1 -> row 0) -> float32 [...] = clamp(table[idx], lo, hi)."""
    return _GatherClamp.apply(table, idx, lo, hi)


class PointAggregator(nn.Module):
    """Same constructor / parameters / forward as the reference class (point_aggregators.py:12-816)."""

    @staticmethod
    def modify_commandline_options(parser, is_train=True):
        """Hot-path subset of the reference's flags (point_aggregators.py:14-217), same names and defaults."""
        add = parser.add_argument
        add('--feat_grad', type=int, default=1)
        add('--conf_grad', type=int, default=1)
        add('--dir_grad', type=int, default=1)
        add('--color_grad', type=int, default=1)
        add('--num_feat_freqs', type=int, default=0)
        add('--dist_xyz_freq', type=int, default=2)
        add('--dist_xyz_deno', type=float, default=0)
        add('--agg_dist_pers', type=int, default=1)
        add('--agg_intrp_order', type=int, default=0)
        add('--agg_distance_kernel', type=str, default="quadric")
        add('--agg_weight_norm', type=int, default=1)
        add('--agg_axis_weight', type=float, nargs='+', default=None)
        add('--apply_pnt_mask', type=int, default=1)
        add('--which_agg_model', type=str, default="viewmlp")
        add('--shading_feature_mlp_layer0', type=int, default=0)
        add('--shading_feature_mlp_layer1', type=int, default=2)
        add('--shading_feature_mlp_layer2', type=int, default=0)
        add('--shading_feature_mlp_layer3', type=int, default=2)
        add('--shading_alpha_mlp_layer', type=int, default=1)
        add('--shading_color_mlp_layer', type=int, default=1)
        add('--shading_feature_num', type=int, default=256)
        add('--shading_color_channel_num', type=int, default=128)
        add('--act_type', type=str, default="ReLU")
        add('--act_super', type=int, default=1)
        add('--view_ori', type=int, default=0)
        add('--point_color_mode', type=str, default="0")
        add('--point_dir_mode', type=str, default="0")
        add('--point_conf_mode', type=str, default="0")

    def __init__(self, opt):
        super().__init__()
        _check_opt(opt)
        self.opt = opt
        F_, W = int(opt.point_features_dim), int(opt.shading_feature_num)
        self.dist_dim = 6
        in1 = F_ + 2 * opt.num_feat_freqs * F_ + 2 * opt.dist_xyz_freq * self.dist_dim  # :278-280
        in3 = W + 3 + 4                                                                  # :308-309
        self.viewdir_channels = 2 * opt.num_viewdir_freqs * 3                            # :252
        inc = W + self.viewdir_channels                                                  # :333
        C = int(W / 2)                                                                   # :335 (fork: no final Linear)
        if C != int(opt.shading_color_channel_num):
            raise NotImplementedError("shading_color_channel_num must equal shading_feature_num/2 in this fork "
                                      "(point_aggregators.py:335,343,644)")

        def act():
            return nn.LeakyReLU(inplace=True)

        self.block1 = nn.Sequential(nn.Linear(in1, W), act(), nn.Linear(W, W), act())
        self.block3 = nn.Sequential(nn.Linear(in3, W), act(), nn.Linear(W, W), act())
        self.alpha_branch = nn.Sequential(nn.Linear(W, 1))
        self.color_branch = nn.Sequential(nn.Linear(inc, C), act(), nn.Linear(C, C), act(), nn.Linear(C, C), act())
        for m in (self.block1, self.block3, self.alpha_branch, self.color_branch):  # creation order of :276-348
            _xavier_like_reference(m)
        _lib.lib()  # fail now, loudly, if the CUDA library is missing

    # ------------------------------------------------------------------------------------------
    def _weights(self):
        return (self.block1[0].weight, self.block1[0].bias, self.block1[2].weight, self.block1[2].bias,
                self.block3[0].weight, self.block3[0].bias, self.block3[2].weight, self.block3[2].bias,
                self.alpha_branch[0].weight, self.alpha_branch[0].bias,
                self.color_branch[0].weight, self.color_branch[0].bias, self.color_branch[2].weight,
                self.color_branch[2].bias, self.color_branch[4].weight, self.color_branch[4].bias)

    def _params(self, K, SR, dirs_per_ray, vsize, compact_out=False):
        o = self.opt
        p = AggParams()
        p.compact_out = int(compact_out)
        p.intrp_order = int(o.agg_intrp_order)
        p.F, p.feat_freqs, p.dist_freqs, p.view_freqs = int(o.point_features_dim), int(o.num_feat_freqs), int(o.dist_xyz_freq), int(o.num_viewdir_freqs)
        p.W, p.C, p.K, p.SR = int(o.shading_feature_num), int(o.shading_color_channel_num), int(K), int(SR)
        p.dirs_per_ray = int(dirs_per_ray)
        p.leaky_slope = 0.01
        p.forward_only = 0
        p.dist_scale = 1.0 if o.dist_xyz_deno == 0 else 1.0 / float(o.dist_xyz_deno * np.linalg.norm(vsize))
        return p

    def _want_weight(self):
        o = self.opt
        return (o.sparse_loss_weight > 0) or ("conf_coefficient" in o.zero_one_loss_items) or o.prob != 0  # :814

    @staticmethod
    def gradiant_clamp(x, min=0.0001, max=1):
        """Straight-through clamp (point_aggregators.py:724-726)."""
        return x - (x - torch.clamp(x, min=min, max=max)).detach()

    # ------------------------------------------------------------------------------------------
    def forward_tables(self, xyz, xyz_pers, embedding, color, pdir, conf, Rw2c, sample_pidx, sample_loc, sample_loc_w,
                       ray_dirs, vsize, dirs_per_ray=True, want_weight=False, compact=False, work_lists=None):
        """Fused gather + aggregate.  xyz/xyz_pers [N,3], embedding [N,F], color/pdir [N,3], conf [N] - the tables may also
        keep the reference's leading batch / trailing unit axes ([1,N,F], [1,N,3], [1,N,1]): passing the parameters
        themselves makes the backward's buffers their .grad with no select-backward fill + copy of the whole table;
        sample_pidx int32 [R,SR,K] (-1 empty); sample_loc / sample_loc_w [R,SR,3]; ray_dirs [R,3]
        (``dirs_per_ray``) or [R,SR,3].  Returns decoded [R,SR,C+1], weight [R,SR,K] or None.
        ``compact=True``: returns (decoded [S,C+1] with one row per VALID sample, weight, slot_row int32 [R,SR]) -
        the layout pn_composite_* consume directly, so the dense mostly-zero [R,SR,C+1] tensor is never written."""
        if Rw2c.dim() != 2:
            raise NotImplementedError("per-point Rw2c is outside the hot path (every shipped script uses one 3x3)")
        R, SR, K = (int(v) for v in sample_pidx.shape)
        T = int(embedding.shape[-2])
        call = _AggCall(self._params(K, SR, dirs_per_ray, vsize, compact), xyz.detach().contiguous(), xyz_pers.detach().contiguous(),
                        Rw2c.detach().contiguous().float(), sample_pidx.contiguous(), sample_loc_w.detach().contiguous(),
                        sample_loc.detach().contiguous(), ray_dirs.detach().contiguous(), R * SR, T, True, want_weight)
        call.collect_touched = bool(getattr(self, "collect_touched", False))
        call.flat_tail = int(getattr(self, "flat_tail_floats", 0))
        call.compact = work_lists             # CompactSamples of sample_pidx queued earlier (None: built here, with its sync)
        call.export_signs = bool(getattr(self, "export_signs", False))
        # inference (torch.no_grad, or nothing that wants a gradient): the forward skips what only a backward reads
        wants_grad = torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in (embedding, color, pdir, conf, *self._weights()))
        call.params.forward_only = 0 if wants_grad else 1
        decoded, weight = _AggFunction.apply(call, embedding, color, pdir, conf, *self._weights())
        self.last_call = call      # work lists of the most recent fused call (ray_shard: touched point rows)
        weight = weight.view(R, SR, K) if want_weight else None
        if compact:
            return decoded, weight, call.slot_row.view(R, SR)
        return decoded.view(R, SR, -1), weight

    def forward(self, sampled_color, sampled_Rw2c, sampled_dir, sampled_conf, sampled_embedding, sampled_xyz_pers, sampled_xyz,
                sample_pnt_mask, sample_loc, sample_loc_w, sample_ray_dirs, vsize, grid_vox_sz):
        """Reference signature (point_aggregators.py:729).  Returns (decoded_features [B,R,SR,C+1],
        ray_valid bool [B,R,SR], weight | None, conf_coefficient | None)."""
        o = self.opt
        in_shape = sample_loc_w.shape
        C1 = int(o.shading_color_channel_num) + 1
        ray_valid = torch.any(sample_pnt_mask, dim=-1)
        if ray_valid.numel() == 0:                                                      # :746-748
            return torch.zeros(in_shape[:-1] + (C1,), device=sample_loc_w.device), ray_valid.view(in_shape[:-1]), None, None
        if sampled_Rw2c.dim() != 2:
            raise NotImplementedError("per-point Rw2c is outside the hot path")
        if sampled_xyz.requires_grad or sampled_xyz_pers.requires_grad:
            raise NotImplementedError("xyz_grad=1 is outside the hot path (every shipped script sets xyz_grad 0)")
        B, R, SR, K = sample_pnt_mask.shape
        assert B == 1, "B = 1 only, like every caller of the reference"
        n_slots, T = R * SR, R * SR * K
        dev = sample_loc_w.device
        ident = torch.arange(T, dtype=torch.int32, device=dev)
        pidx = torch.where(sample_pnt_mask.reshape(-1), ident, torch.full_like(ident, -1))
        want_w = self._want_weight()
        call = _AggCall(self._params(K, SR, False, vsize), sampled_xyz.detach().reshape(T, 3).contiguous(),
                        sampled_xyz_pers.detach().reshape(T, 3).contiguous(), sampled_Rw2c.detach().contiguous().float(),
                        pidx, sample_loc_w.detach().reshape(n_slots, 3).contiguous(),
                        sample_loc.detach().reshape(n_slots, 3).contiguous(),
                        sample_ray_dirs.detach().reshape(n_slots, 3).contiguous(), n_slots, T, False, want_w)
        call.export_signs = bool(getattr(self, "export_signs", False))
        self.last_call = call
        decoded, weight = _AggFunction.apply(call, sampled_embedding.reshape(T, -1), sampled_color.reshape(T, 3),
                                             sampled_dir.reshape(T, 3), sampled_conf.reshape(T), *self._weights())
        if not want_w:
            return decoded.view(in_shape[:-1] + (C1,)), ray_valid.view(in_shape[:-1]), None, None
        conf_coefficient = self.gradiant_clamp(sampled_conf[..., 0], min=0.0001, max=1)   # :809-811
        return (decoded.view(in_shape[:-1] + (C1,)), ray_valid.view(in_shape[:-1]), weight.view(B, R, SR, K),
                conf_coefficient)
