// pn_voxel_downsample: voxel down-sampling of a point cloud.
// Reference: models/mvs/mvs_utils.py:537-561 `construct_vox_points_closest` - torch.unique(dim=0, return_inverse) of the
// integer voxel coordinates, torch_scatter.scatter_mean (centroid per voxel) and scatter_min (the point nearest to its
// voxel's centroid).  Here: one stable LSD radix sort of (packed voxel key, point index), head flags + scan for the
// voxel ids, and one pass per voxel over its (index-ascending) segment - no atomics, deterministic, and the output
// order is the lexicographic (x, y, z) order torch.unique produces.
#include "common.cuh"

namespace pn {

constexpr int kSortThreads = 256;
constexpr int kSortItems = 8;                       // per thread, in rounds: element = tile + round*256 + tid (index-ascending)
constexpr int kSortTile = kSortThreads * kSortItems;
constexpr int kSortWarps = kSortThreads / 32;

__global__ void voxel_key_kernel(const float* __restrict__ xyz, int64_t n, float3 smin, float3 vs, int3 cbase, int bits,
                                 unsigned long long* __restrict__ key, int* __restrict__ idx, int* __restrict__ bad) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // floor((p - space_min) / vox_size) in IEEE fp32, one rounding per operation like the reference's torch ops (:552-554)
    int cx = (int)floorf(__fdiv_rn(__fsub_rn(xyz[i * 3], smin.x), vs.x)) - cbase.x;
    int cy = (int)floorf(__fdiv_rn(__fsub_rn(xyz[i * 3 + 1], smin.y), vs.y)) - cbase.y;
    int cz = (int)floorf(__fdiv_rn(__fsub_rn(xyz[i * 3 + 2], smin.z), vs.z)) - cbase.z;
    const int lim = (1 << bits) - 1;
    if (cx < 0 || cy < 0 || cz < 0 || cx > lim || cy > lim || cz > lim) {
        atomicAdd(bad, 1);
        cx = min(max(cx, 0), lim); cy = min(max(cy, 0), lim); cz = min(max(cz, 0), lim);
    }
    key[i] = ((unsigned long long)cx << (2 * bits)) | ((unsigned long long)cy << bits) | (unsigned long long)cz;
    idx[i] = (int)i;
}

// digit histogram of one tile -> hist[digit * nblk + block]
__global__ void __launch_bounds__(kSortThreads) radix_hist_kernel(const unsigned long long* __restrict__ key, int64_t n, int shift,
                                                                  int nblk, int* __restrict__ hist) {
    __shared__ int h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * kSortTile;
#pragma unroll
    for (int r = 0; r < kSortItems; ++r) {
        int64_t i = base + r * kSortThreads + threadIdx.x;
        if (i < n) atomicAdd(&h[(int)((key[i] >> shift) & 255)], 1);
    }
    __syncthreads();
    hist[threadIdx.x * nblk + blockIdx.x] = h[threadIdx.x];
}

// stable scatter: rank of an element among equal digits of its tile = (equal digits in earlier (round, warp) groups) +
// (equal digits in lower lanes of its warp, by __match_any_sync)
__global__ void __launch_bounds__(kSortThreads) radix_scatter_kernel(
    const unsigned long long* __restrict__ key, const int* __restrict__ idx, int64_t n, int shift, int nblk,
    const int* __restrict__ offs /* exclusive scan of hist */, unsigned long long* __restrict__ key_out, int* __restrict__ idx_out) {
    __shared__ unsigned short cnt[kSortItems * kSortWarps][256];     // 32 KB
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int j = threadIdx.x; j < kSortItems * kSortWarps * 256; j += kSortThreads) (&cnt[0][0])[j] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * kSortTile;
    unsigned long long k[kSortItems];
    int v[kSortItems], dg[kSortItems], rk[kSortItems];
#pragma unroll
    for (int r = 0; r < kSortItems; ++r) {
        int64_t i = base + r * kSortThreads + threadIdx.x;
        bool ok = i < n;
        k[r] = ok ? key[i] : 0ull;
        v[r] = ok ? idx[i] : 0;
        dg[r] = ok ? (int)((k[r] >> shift) & 255) : 256 + lane;       // inactive lanes match nobody
        unsigned peers = __match_any_sync(kFull, dg[r]);
        rk[r] = __popc(peers & ((1u << lane) - 1));
        if (ok && rk[r] == 0) cnt[r * kSortWarps + warp][dg[r]] = (unsigned short)__popc(peers);
        if (!ok) dg[r] = -1;
    }
    __syncthreads();
    // exclusive prefix over the (round, warp) groups, per digit: thread d owns digit d
    {
        int run = offs[threadIdx.x * nblk + blockIdx.x];
        unsigned short carry_lo;
        int acc = 0;
        // stored back as offsets relative to `run` (fit 16 bits: a tile holds 2048 elements)
        for (int g = 0; g < kSortItems * kSortWarps; ++g) {
            carry_lo = cnt[g][threadIdx.x];
            cnt[g][threadIdx.x] = (unsigned short)acc;
            acc += carry_lo;
        }
        // the digit's global base goes to a second shared array
        __shared__ int gbase[256];
        gbase[threadIdx.x] = run;
        __syncthreads();
#pragma unroll
        for (int r = 0; r < kSortItems; ++r) {
            if (dg[r] >= 0) {
                int64_t pos = (int64_t)gbase[dg[r]] + cnt[r * kSortWarps + warp][dg[r]] + rk[r];
                key_out[pos] = k[r];
                idx_out[pos] = v[r];
            }
        }
    }
}

__global__ void voxel_head_kernel(const unsigned long long* __restrict__ key, int64_t n, int* __restrict__ flag) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flag[i] = (i == 0 || key[i] != key[i - 1]) ? 1 : 0;
}

// vid[i] = (inclusive scan of head flags) - 1 = exclusive scan + flag - 1; segment starts recorded per voxel
__global__ void voxel_start_kernel(const int* __restrict__ flag, const int* __restrict__ excl, int64_t n, int* __restrict__ seg_start) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (flag[i]) seg_start[excl[i]] = (int)i;
    if (i == n - 1) seg_start[excl[i] + flag[i]] = (int)n;
}

// one thread per voxel: its points are key-sorted with ascending original index inside the segment
__global__ void voxel_reduce_kernel(const float* __restrict__ xyz, const unsigned long long* __restrict__ key, const int* __restrict__ idx,
                                    const int* __restrict__ seg_start, const int* __restrict__ n_vox, int bits, int3 cbase,
                                    float* __restrict__ centroid, int* __restrict__ grid_idx, long long* __restrict__ min_idx,
                                    int* __restrict__ inv_idx) {
    int u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= *n_vox) return;
    const int a = seg_start[u], b = seg_start[u + 1];
    float sx = 0.f, sy = 0.f, sz = 0.f;
    for (int j = a; j < b; ++j) {            // sequential, index-ascending: the order of a serial scatter_mean
        int64_t p = idx[j];
        sx = __fadd_rn(sx, xyz[p * 3]); sy = __fadd_rn(sy, xyz[p * 3 + 1]); sz = __fadd_rn(sz, xyz[p * 3 + 2]);
        inv_idx[p] = u;
    }
    const float cnt = (float)(b - a);
    const float cx = __fdiv_rn(sx, cnt), cy = __fdiv_rn(sy, cnt), cz = __fdiv_rn(sz, cnt);
    float best = INFINITY;
    int bi = -1;
    for (int j = a; j < b; ++j) {
        int64_t p = idx[j];
        float dx = __fsub_rn(xyz[p * 3], cx), dy = __fsub_rn(xyz[p * 3 + 1], cy), dz = __fsub_rn(xyz[p * 3 + 2], cz);
        // torch.norm(., dim=-1) on the host (mvs_utils.py:556) accumulates its three squares as an FMA chain:
        // fma(dz, dz, fma(dy, dy, dx * dx)); the same chain here makes residual ties, and hence the arg-min, bit-identical
        float d = sqrtf(__fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx))));
        if (d < best) { best = d; bi = (int)p; }      // strict <: the smallest index wins ties
    }
    centroid[u * 3] = cx; centroid[u * 3 + 1] = cy; centroid[u * 3 + 2] = cz;
    const unsigned long long k = key[a];
    const unsigned long long m = (1ull << bits) - 1;
    grid_idx[u * 3] = (int)((k >> (2 * bits)) & m) + cbase.x;
    grid_idx[u * 3 + 1] = (int)((k >> bits) & m) + cbase.y;
    grid_idx[u * 3 + 2] = (int)(k & m) + cbase.z;
    min_idx[u] = bi;
}

struct VoxWs {
    unsigned long long *k0, *k1;
    int *i0, *i1, *hist, *hscan, *flag, *excl, *fscan, *seg, *bad;
    int nblk;
};
static int64_t carve_vox(int64_t n, void* p, int64_t bytes, VoxWs* out) {
    Carver c(p, bytes);
    int64_t m = n > 0 ? n : 1;
    VoxWs w;
    w.nblk = (int)cdiv(m, kSortTile);
    w.k0 = c.take<unsigned long long>(m);
    w.k1 = c.take<unsigned long long>(m);
    w.i0 = c.take<int>(m);
    w.i1 = c.take<int>(m);
    w.hist = c.take<int>((int64_t)256 * w.nblk + 1);
    w.hscan = c.take<int>(scan_scratch_elems_i32((int64_t)256 * w.nblk + 1));
    w.flag = c.take<int>(m + 1);
    w.excl = c.take<int>(m + 1);
    w.fscan = c.take<int>(scan_scratch_elems_i32(m + 1));
    w.seg = c.take<int>(m + 1);
    w.bad = c.take<int>(1);
    if (out) *out = w;
    return c.off;
}

}  // namespace pn

using namespace pn;

extern "C" int64_t pn_voxel_downsample_workspace_bytes(int64_t n_points) { return carve_vox(n_points, nullptr, 0, nullptr); }

extern "C" int pn_voxel_downsample(const float* xyz, int64_t n_points, const float* space_min, const float* vox_size,
                                   const int32_t* coord_base, int32_t bits, float* centroid, int32_t* grid_idx, int64_t* min_idx, int32_t* inv_idx, int32_t* n_voxels,
                                   int32_t* n_outside, void* workspace, int64_t workspace_bytes, void* stream) {
    PN_CHECK_ARG(n_points >= 0 && n_points < (int64_t)INT32_MAX - kSortTile, "pn_voxel_downsample: bad point count");
    PN_CHECK_ARG(bits >= 1 && bits <= 21, "pn_voxel_downsample: %d key bits per axis (1..21)", bits);
    PN_CHECK_ARG(space_min && vox_size && coord_base && n_voxels, "pn_voxel_downsample: null argument");
    const int3 cbase = make_int3(coord_base[0], coord_base[1], coord_base[2]);
    cudaStream_t st = (cudaStream_t)stream;
    if (n_points == 0) {
        PN_CUDA(cudaMemsetAsync(n_voxels, 0, sizeof(int32_t), st));
        if (n_outside) PN_CUDA(cudaMemsetAsync(n_outside, 0, sizeof(int32_t), st));
        return PN_OK;
    }
    PN_CHECK_ARG(xyz && centroid && grid_idx && min_idx && inv_idx, "pn_voxel_downsample: null argument");
    PN_CHECK_ARG(workspace_bytes >= carve_vox(n_points, nullptr, 0, nullptr), "pn_voxel_downsample: workspace too small");
    VoxWs w;
    carve_vox(n_points, workspace, workspace_bytes, &w);
    const int64_t n = n_points;
    PN_CUDA(cudaMemsetAsync(w.bad, 0, sizeof(int), st));
    voxel_key_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(xyz, n, make_float3(space_min[0], space_min[1], space_min[2]),
                                                             make_float3(vox_size[0], vox_size[1], vox_size[2]), cbase, bits, w.k0, w.i0, w.bad);
    PN_LAUNCH_CHECK("voxel_key");
    unsigned long long *ka = w.k0, *kb = w.k1;
    int *ia = w.i0, *ib = w.i1;
    for (int shift = 0; shift < 3 * bits; shift += 8) {
        radix_hist_kernel<<<w.nblk, kSortThreads, 0, st>>>(ka, n, shift, w.nblk, w.hist);
        PN_LAUNCH_CHECK("radix_hist");
        int rc = scan_exclusive_i32(w.hist, w.hist, (int64_t)256 * w.nblk, false, w.hscan, st);
        if (rc) return rc;
        radix_scatter_kernel<<<w.nblk, kSortThreads, 0, st>>>(ka, ia, n, shift, w.nblk, w.hist, kb, ib);
        PN_LAUNCH_CHECK("radix_scatter");
        unsigned long long* tk = ka; ka = kb; kb = tk;
        int* ti = ia; ia = ib; ib = ti;
    }
    voxel_head_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(ka, n, w.flag);
    PN_LAUNCH_CHECK("voxel_head");
    int rc = scan_exclusive_i32(w.flag, w.excl, n, true, w.fscan, st);
    if (rc) return rc;
    voxel_start_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(w.flag, w.excl, n, w.seg);
    PN_LAUNCH_CHECK("voxel_start");
    PN_CUDA(cudaMemcpyAsync(n_voxels, w.excl + n, sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
    if (n_outside) PN_CUDA(cudaMemcpyAsync(n_outside, w.bad, sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
    // the voxel count is only known on the device: launch for the upper bound, threads beyond *n_voxels exit
    voxel_reduce_kernel<<<(unsigned)cdiv(n, 128), 128, 0, st>>>(xyz, ka, ia, w.seg, w.excl + n, bits, cbase, centroid, grid_idx,
                                                                (long long*)min_idx, inv_idx);
    PN_LAUNCH_CHECK("voxel_reduce");
    return PN_OK;
}
