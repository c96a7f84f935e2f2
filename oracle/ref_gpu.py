"""The reference's OWN querier kernels on the GPU (TEST / BENCH INFRASTRUCTURE, see oracle/__init__.py).

The reference launches its CUDA C string through pycuda (query_point_indices_worldcoords.py:112-537), which is absent here.
pycuda is a thin wrapper over the driver API, so the same string - extracted where it lies by oracle/build_ref.py and
compiled with ``nvcc -arch=sm_100a -cubin`` into oracle/_ref/ (a build output: git-ignored, travels to the GPU box) - is
loaded with ``cuModuleLoadData`` and launched with ``cuLaunchKernel`` (ctypes on libcuda) on buffers owned by torch, with
the reference's launch geometry (one thread per point / candidate sample / kept sample, ``gpu_maxthr`` threads per block,
:546-708) and the reference's fills (:549-555, :639, :657-658, :667).  The host glue between the launches (:639-721) is
restated with the torch ops the reference uses.

Two uses: (1) the GPU-side bar SURVEY.md 2.2 names ("beat the same kernel compiled for sm_100 by nvcc"): timed stage by
stage beside pn_grid_build / pn_ray_sample / pn_knn_query in bench.py; (2) a second oracle under the REAL schedule: the
atomics resolve in hardware order, so bucket order, voxel slot numbers (and WHICH voxel is the slot-0 voxel that never
receives points, :372) differ from the canonical serial schedule - neighbour SETS are compared, not slot order.
"""
import ctypes
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


def cubin_path(K):
    return os.path.join(HERE, "_ref", "ref_querier_sm100a_K%d.cubin" % int(K))


def available(K):
    return os.path.isfile(cubin_path(K))


class _Driver:
    _inst = None

    def __init__(self):
        self.cu = ctypes.CDLL("libcuda.so.1")
        torch.cuda.init()
        torch.zeros(1, device="cuda")                   # torch's primary context is current on this thread from here on
        self._check(self.cu.cuInit(0), "cuInit")
        self.mods = {}

    def _check(self, rc, what):
        if rc != 0:
            msg = ctypes.c_char_p()
            self.cu.cuGetErrorString(rc, ctypes.byref(msg))
            raise RuntimeError("%s failed: %d %s" % (what, rc, (msg.value or b"").decode()))

    @classmethod
    def get(cls):
        if cls._inst is None:
            cls._inst = cls()
        return cls._inst

    def module(self, K):
        if K not in self.mods:
            data = open(cubin_path(K), "rb").read()
            mod = ctypes.c_void_p()
            self._check(self.cu.cuModuleLoadData(ctypes.byref(mod), data), "cuModuleLoadData")
            fns = {}
            for name in ("claim_occ", "map_coor2occ", "fill_occ2pnts", "mask_raypos", "get_shadingloc", "query_neigh_along_ray_layered"):
                f = ctypes.c_void_p()
                self._check(self.cu.cuModuleGetFunction(ctypes.byref(f), mod, name.encode()), "cuModuleGetFunction " + name)
                fns[name] = f
            self.mods[K] = (mod, fns)
        return self.mods[K][1]

    def launch(self, fn, grid, block, args):
        """args: torch tensors (device pointers), python ints (int32), floats (float32), ('u64', v)."""
        store = []
        for a in args:
            if torch.is_tensor(a):
                assert a.is_cuda and a.is_contiguous()
                store.append(ctypes.c_void_p(a.data_ptr()))
            elif isinstance(a, tuple) and a[0] == "u64":
                store.append(ctypes.c_uint64(int(a[1])))
            elif isinstance(a, float):
                store.append(ctypes.c_float(a))
            else:
                store.append(ctypes.c_int32(int(a)))
        params = (ctypes.c_void_p * len(store))(*[ctypes.cast(ctypes.byref(s), ctypes.c_void_p) for s in store])
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        self._check(self.cu.cuLaunchKernel(fn, int(grid), 1, 1, int(block), 1, 1, 0, stream, params, None), "cuLaunchKernel")


class ReferenceGpuQuerier:
    """build_occ_vox + query_grid_point_index of the reference (:546-721) on the device, its kernels unmodified."""

    def __init__(self, opt, hp, dev):
        self.opt, self.dev = opt, dev
        self.K, self.SR, self.P, self.max_o = int(opt.K), int(opt.SR), int(opt.P), int(opt.max_o)
        self.thr = int(getattr(opt, "gpu_maxthr", 1024))
        self.fns = _Driver.get().module(self.K)
        self.vdim = [int(v) for v in hp["scaled_vdim"]]
        self.vol = self.vdim[0] * self.vdim[1] * self.vdim[2]
        self.shift = torch.tensor(np.asarray(hp["ranges"][:3], np.float32), device=dev)
        self.vsize = torch.tensor(np.asarray(hp["scaled_vsize"], np.float32), device=dev)
        self.vdim_gpu = torch.tensor(self.vdim, dtype=torch.int32, device=dev)
        self.kernel_size = torch.tensor(np.asarray(hp["kernel_size"], np.int32), device=dev)
        self.query_size = torch.tensor(np.asarray(hp["query_size"], np.int32), device=dev)
        self.radius2 = float(np.float32(hp["radius_limit"]) ** 2)

    def build_occ_vox(self, xyz):
        """xyz float32 [1,N,3] -> the six grid tensors (:546-611), fills included."""
        d, dev, B, N, P, max_o, thr = _Driver.get(), self.dev, 1, int(xyz.shape[1]), self.P, self.max_o, self.thr
        vd = self.vdim
        numpoints = torch.full([1], N, dtype=torch.int32, device=dev)
        coor_occ = torch.zeros([B, vd[0], vd[1], vd[2]], dtype=torch.int32, device=dev)
        occ_2_pnts = torch.full([B, max_o, P], -1, dtype=torch.int32, device=dev)
        occ_2_coor = torch.full([B, max_o, 3], -1, dtype=torch.int32, device=dev)
        occ_numpnts = torch.zeros([B, max_o], dtype=torch.int32, device=dev)
        coor_2_occ = torch.full([B, vd[0], vd[1], vd[2]], -1, dtype=torch.int32, device=dev)
        occ_idx = torch.zeros([B], dtype=torch.int32, device=dev)
        grid = (B * N + thr - 1) // thr
        d.launch(self.fns["claim_occ"], grid, thr, [xyz, numpoints, B, N, self.shift, self.vsize, self.vdim_gpu, self.vol, max_o, occ_idx,
                                                    coor_2_occ, occ_2_coor, ("u64", 0)])
        coor_2_occ = torch.full([B, vd[0], vd[1], vd[2]], -1, dtype=torch.int32, device=dev)
        d.launch(self.fns["map_coor2occ"], (B * max_o + thr - 1) // thr, thr, [B, self.vdim_gpu, self.query_size, self.vol, max_o, occ_idx, coor_occ,
                                                                               coor_2_occ, occ_2_coor])
        d.launch(self.fns["fill_occ2pnts"], grid, thr, [xyz, numpoints, B, N, P, self.shift, self.vsize, self.vdim_gpu, self.vol, max_o, coor_2_occ,
                                                        occ_2_pnts, occ_numpnts, ("u64", 0)])
        return coor_occ, occ_2_coor, coor_2_occ, occ_idx, occ_numpnts, occ_2_pnts

    def query(self, xyz, raypos, grid=None, timing=None):
        """raypos float32 [1,R,D,3] (campos + raydir * ts, built by the caller with the reference's torch ops) ->
        (sample_pidx [1,R'',SR,K], sample_loc_w [1,R'',SR,3], ray_mask int8 [1,R]).  ``timing``: dict receiving CUDA-event
        milliseconds per stage (grid build | mask + select + shading locations | neighbour query | compaction)."""
        d, dev, B, thr, SR, K = _Driver.get(), self.dev, 1, self.thr, self.SR, self.K
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
        ev[0].record()
        coor_occ, occ_2_coor, coor_2_occ, occ_idx, occ_numpnts, occ_2_pnts = grid if grid is not None else self.build_occ_vox(xyz)
        ev[1].record()
        R, D = int(raypos.shape[1]), int(raypos.shape[2])
        raypos_mask = torch.zeros([B, R, D], dtype=torch.int32, device=dev)
        gsz = (B * R * D + thr - 1) // thr
        d.launch(self.fns["mask_raypos"], gsz, thr, [raypos, coor_occ, B, R, D, self.vol, self.shift, self.vdim_gpu, self.vsize, raypos_mask])
        ray_mask = torch.max(raypos_mask, dim=-1)[0] > 0
        R2 = int(torch.max(torch.sum(ray_mask.to(torch.int32))).cpu().numpy())
        sample_loc = torch.zeros([B, R2, SR, 3], dtype=torch.float32, device=dev)
        sample_pidx = torch.full([B, R2, SR, K], -1, dtype=torch.int32, device=dev)
        if R2 > 0:
            raypos = torch.masked_select(raypos, ray_mask[..., None, None].expand(-1, -1, D, 3)).reshape(B, R2, D, 3)
            raypos_mask = torch.masked_select(raypos_mask, ray_mask[..., None].expand(-1, -1, D)).reshape(B, R2, D)
            cum = torch.cumsum(raypos_mask, dim=-1).to(torch.int32)
            raypos_mask = (raypos_mask * cum * (cum <= SR)) - 1
            sample_loc_mask = torch.zeros([B, R2, SR], dtype=torch.int32, device=dev)
            d.launch(self.fns["get_shadingloc"], gsz, thr, [raypos, raypos_mask, B, R2, D, SR, sample_loc, sample_loc_mask])
            ev[2].record()
            d.launch(self.fns["query_neigh_along_ray_layered"], (B * R2 * SR + thr - 1) // thr, thr,
                     [xyz, B, SR, R2, self.max_o, self.P, K, self.vol, self.radius2, self.shift, self.vdim_gpu, self.vsize, self.kernel_size,
                      occ_numpnts, occ_2_pnts, coor_2_occ, sample_loc, sample_loc_mask, sample_pidx, ("u64", 0), int(getattr(self.opt, "NN", 2))])
            ev[3].record()
            valid = torch.sum(sample_pidx.view(B, R2, -1) >= 0, dim=-1) > 0
            R3 = int(torch.max(torch.sum(valid.to(torch.int32), dim=-1)).cpu().numpy())
            ray_mask.masked_scatter_(ray_mask, valid)
            sample_pidx = torch.masked_select(sample_pidx, valid[..., None, None].expand(-1, -1, SR, K)).reshape(B, R3, SR, K)
            sample_loc = torch.masked_select(sample_loc, valid[..., None, None].expand(-1, -1, SR, 3)).reshape(B, R3, SR, 3)
        else:
            ev[2].record()
            ev[3].record()
        ev[4].record()
        if timing is not None:
            torch.cuda.synchronize()
            for k, a, b in (("grid_build_ms", 0, 1), ("mask_select_shadingloc_ms", 1, 2), ("knn_query_ms", 2, 3), ("compaction_ms", 3, 4)):
                timing[k] = timing.get(k, 0.0) + ev[a].elapsed_time(ev[b])
        return sample_pidx, sample_loc, ray_mask.to(torch.int8)
