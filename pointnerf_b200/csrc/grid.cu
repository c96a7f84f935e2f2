// Voxel grid build: occupancy + per-voxel point buckets as a CSR over a dense cell index.
//
// Replaces claim_occ / map_coor2occ / fill_occ2pnts and the six dense fills of build_occ_vox
// (models/neural_points/query_point_indices_worldcoords.py:243-387, 546-611).  The reference
// claims table slots with atomicCAS races and fills fixed [max_o, P] buckets in arrival order; here
// a counting sort over the dense cell grid gives the same buckets in the order the reference would
// produce under the canonical (ascending thread index) schedule, deterministically:
//
//   count  : one thread per point -> cell id, atomicAdd(count[cell]) whose return value is an
//            arrival rank inside the cell (any permutation), warp-aggregated min of in-grid index
//   scan   : exclusive prefix over the V+1 cell counters (pn::scan_exclusive_i32)
//   place  : tmp[start[cell] + rank] = point
//   order  : bucket[start + #{members < point}] = point   (ascending index inside each cell); the thread of a cell's
//            first member ORs the cell's dilation box into the bitset
//
// All kernels are HBM/L2-bound integer work; algorithmic bytes: 12N (xyz) + 4N (bucket list)
// + 4(V+1) (cell_start) + V/8 (bitset).
#include <limits.h>

#include "common.cuh"

namespace pn {

struct GridDev {
    float sx, sy, sz, vx, vy, vz;
    int dx, dy, dz;
    int kx, ky, kz;
    int P, max_o;
};

static GridDev to_dev(const pn_grid_params* p) {
    GridDev g;
    g.sx = p->shift[0]; g.sy = p->shift[1]; g.sz = p->shift[2];
    g.vx = p->vsize[0]; g.vy = p->vsize[1]; g.vz = p->vsize[2];
    g.dx = p->dims[0]; g.dy = p->dims[1]; g.dz = p->dims[2];
    g.kx = p->dilate[0]; g.ky = p->dilate[1]; g.kz = p->dilate[2];
    g.P = p->P; g.max_o = p->max_o;
    return g;
}

__device__ __forceinline__ int cell_index(const GridDev& g, float x, float y, float z) {
    int cx = voxel_coord(x, g.sx, g.vx), cy = voxel_coord(y, g.sy, g.vy), cz = voxel_coord(z, g.sz, g.vz);
    if (cx < 0 || cx >= g.dx || cy < 0 || cy >= g.dy || cz < 0 || cz >= g.dz) return -1;
    return (cx * g.dy + cy) * g.dz + cz;
}

// 3 consecutive floats per point; a warp reads 384 contiguous bytes (coalesced through L1)
// Frustum grids (query_point_indices.py): get_occ_vox (:290-297) takes floor() of the scaled coordinate, insert_vox_points
// (:394-399) converts it with (int), i.e. toward zero - a point up to one voxel below a low face of the grid is not a
// source of occupancy but IS filed into the boundary voxel, provided that voxel was claimed by a floor()-valid point
// (only claimed voxels get a bucket, :345-349, :400).  So: a first pass sets `src_bits` for the cells that hold a
// floor()-valid point, and bucket membership follows the (int) rule restricted to those cells.
__device__ __forceinline__ int cell_index_trunc(const GridDev& g, float x, float y, float z) {
    float tx = __fdiv_rn(__fsub_rn(x, g.sx), g.vx), ty = __fdiv_rn(__fsub_rn(y, g.sy), g.vy), tz = __fdiv_rn(__fsub_rn(z, g.sz), g.vz);
    int cx = (int)tx, cy = (int)ty, cz = (int)tz;
    if (cx < 0 || cx >= g.dx || cy < 0 || cy >= g.dy || cz < 0 || cz >= g.dz) return -1;
    return (cx * g.dy + cy) * g.dz + cz;
}

__global__ void __launch_bounds__(256) grid_mark_src_kernel(const float* __restrict__ xyz, int n, GridDev g,
                                                            unsigned* __restrict__ src_bits) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int cell = cell_index(g, xyz[3 * (int64_t)i], xyz[3 * (int64_t)i + 1], xyz[3 * (int64_t)i + 2]);
    if (cell >= 0 && !((src_bits[cell >> 5] >> (cell & 31)) & 1u)) atomicOr(&src_bits[cell >> 5], 1u << (cell & 31));
}

__global__ void __launch_bounds__(256) grid_count_kernel(const float* __restrict__ xyz, int n, GridDev g,
                                                         int* __restrict__ cell_cnt, int* __restrict__ pt_cell,
                                                         int* __restrict__ pt_rank, int* __restrict__ info,
                                                         const unsigned* __restrict__ src_bits) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int cell = -1;
    if (i < n) {
        float x = xyz[3 * (int64_t)i], y = xyz[3 * (int64_t)i + 1], z = xyz[3 * (int64_t)i + 2];
        if (src_bits) {
            cell = cell_index_trunc(g, x, y, z);
            if (cell >= 0 && !((src_bits[cell >> 5] >> (cell & 31)) & 1u)) cell = -1;
        } else {
            cell = cell_index(g, x, y, z);
        }
        pt_cell[i] = cell;
        pt_rank[i] = cell >= 0 ? atomicAdd(&cell_cnt[cell], 1) : 0;
    }
    // lowest in-grid point index and number of in-grid points: one atomic per warp
    unsigned inside = __ballot_sync(kFull, cell >= 0);
    if (inside && (threadIdx.x & 31) == 0) {
        int first = (i & ~31) + (__ffs(inside) - 1);
        atomicMin(&info[PN_GI_FIRST_POINT], first);
        atomicAdd(&info[PN_GI_N_INSIDE], __popc(inside));
    }
}

__global__ void __launch_bounds__(256) grid_place_kernel(int n, const int* __restrict__ pt_cell,
                                                         const int* __restrict__ pt_rank,
                                                         const int* __restrict__ cell_start, int* __restrict__ tmp) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int cell = pt_cell[i];
    if (cell >= 0) tmp[cell_start[cell] + pt_rank[i]] = i;
}

// ascending point index inside each cell: position = number of smaller members (cells are small).  The thread of a
// cell's lowest-index member also ORs the cell's dilation box into the occupancy bitset and feeds the statistics, so the
// cost of the dilation follows the number of occupied cells (<= N), not the number of cells of the grid (a per-view
// frustum grid has 16-33 M cells of which ~1 % are occupied).
__global__ void __launch_bounds__(256) grid_order_kernel(int n, GridDev g, const int* __restrict__ pt_cell,
                                                         const int* __restrict__ cell_start,
                                                         const int* __restrict__ tmp, int* __restrict__ bucket,
                                                         unsigned* __restrict__ occ_bits, int* __restrict__ info,
                                                         int slot0_quirk) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int cell = i < n ? pt_cell[i] : -1;
    int cnt = 0;   // > 0 only on the thread of the cell's first member
    if (cell >= 0) {
        int s = cell_start[cell], e = cell_start[cell + 1];
        int smaller = 0;
        for (int j = s; j < e; ++j) smaller += (tmp[j] < i);
        bucket[s + smaller] = i;
        if (slot0_quirk && i == info[PN_GI_FIRST_POINT]) info[PN_GI_SLOT0_CELL] = cell;  // the reference's slot-0 voxel (:372)
        if (smaller == 0) {
            cnt = e - s;
            int cz = cell % g.dz, t = cell / g.dz;
            int cy = t % g.dy, cx = t / g.dy;
            int x0 = max(0, cx - g.kx / 2), x1 = min(g.dx, cx + (g.kx + 1) / 2);
            int y0 = max(0, cy - g.ky / 2), y1 = min(g.dy, cy + (g.ky + 1) / 2);
            int z0 = max(0, cz - g.kz / 2), z1 = min(g.dz, cz + (g.kz + 1) / 2);
            for (int x = x0; x < x1; ++x)
                for (int y = y0; y < y1; ++y) {
                    int64_t row = ((int64_t)x * g.dy + y) * g.dz;
                    for (int z = z0; z < z1; ++z) {
                        int64_t b = row + z;
                        unsigned bit = 1u << (b & 31);
                        if (!(occ_bits[b >> 5] & bit)) atomicOr(&occ_bits[b >> 5], bit);
                    }
                }
        }
    }
    // statistics: one atomic per warp
    unsigned occ = __ballot_sync(kFull, cnt > 0);
    unsigned over = __ballot_sync(kFull, cnt > g.P);
    int mx = cnt;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(kFull, mx, o));
    if ((threadIdx.x & 31) == 0 && occ) {
        atomicAdd(&info[PN_GI_N_OCC], __popc(occ));
        if (over) atomicAdd(&info[PN_GI_OVER_P], __popc(over));
        atomicMax(&info[PN_GI_MAX_COUNT], mx);
    }
}

__global__ void grid_finish_kernel(GridDev g, int* info) {
    info[PN_GI_OVER_MAX_O] = info[PN_GI_N_OCC] > g.max_o ? 1 : 0;
}

__global__ void grid_init_info_kernel(int* info) {
    int i = threadIdx.x;
    if (i < PN_GRID_INFO_LEN) info[i] = (i == PN_GI_FIRST_POINT) ? INT_MAX : (i == PN_GI_SLOT0_CELL ? -1 : 0);
}

}  // namespace pn

using namespace pn;

extern "C" int64_t pn_grid_workspace_bytes(int64_t n_points, int64_t n_cells) {
    int64_t n = n_points > 0 ? n_points : 1;
    return 3 * align_up(n * 4, 256) + align_up(scan_scratch_elems_i32(n_cells + 1) * 4, 256) + 256;
}

extern "C" int64_t pn_frustum_grid_workspace_bytes(int64_t n_points, int64_t n_cells) {
    return pn_grid_workspace_bytes(n_points, n_cells) + align_up(cdiv(n_cells, 32) * 4, 256);
}

static int grid_build_impl(const float* xyz, int64_t n_points, const pn_grid_params* params, int32_t* cell_start,
                           int32_t* bucket_pidx, uint32_t* occ_bits, int32_t* info, void* workspace,
                           int64_t workspace_bytes, void* stream, bool frustum) {
    PN_CHECK_ARG(params && cell_start && bucket_pidx && occ_bits && info, "pn_grid_build: null argument");
    PN_CHECK_ARG(n_points >= 0 && n_points < INT_MAX, "pn_grid_build: n_points out of range");
    int64_t V = (int64_t)params->dims[0] * params->dims[1] * params->dims[2];
    PN_CHECK_ARG(params->dims[0] > 0 && params->dims[1] > 0 && params->dims[2] > 0 && V < INT_MAX - 1,
                 "pn_grid_build: grid of %lld cells unsupported (dims %d %d %d)", (long long)V, params->dims[0],
                 params->dims[1], params->dims[2]);
    PN_CHECK_ARG(workspace_bytes >= (frustum ? pn_frustum_grid_workspace_bytes(n_points, V) : pn_grid_workspace_bytes(n_points, V)),
                 "pn_grid_build: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    GridDev g = to_dev(params);
    int n = (int)n_points;
    Carver ws(workspace, workspace_bytes);
    int* pt_cell = ws.take<int>(n > 0 ? n : 1);
    int* pt_rank = ws.take<int>(n > 0 ? n : 1);
    int* tmp = ws.take<int>(n > 0 ? n : 1);
    int* scratch = ws.take<int>(scan_scratch_elems_i32(V + 1));
    unsigned* src_bits = nullptr;
    if (frustum) {
        src_bits = ws.take<unsigned>(cdiv(V, 32));
        PN_CUDA(cudaMemsetAsync(src_bits, 0, cdiv(V, 32) * sizeof(unsigned), st));
    }

    PN_CUDA(cudaMemsetAsync(cell_start, 0, (V + 1) * sizeof(int), st));
    PN_CUDA(cudaMemsetAsync(occ_bits, 0, cdiv(V, 32) * sizeof(unsigned), st));
    grid_init_info_kernel<<<1, 32, 0, st>>>(info);
    PN_LAUNCH_CHECK("grid_init_info");
    if (n > 0 && frustum) {
        grid_mark_src_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(xyz, n, g, src_bits);
        PN_LAUNCH_CHECK("grid_mark_src");
    }
    if (n > 0) {
        grid_count_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(xyz, n, g, cell_start, pt_cell, pt_rank, info, src_bits);
        PN_LAUNCH_CHECK("grid_count");
    }
    int rc = scan_exclusive_i32(cell_start, cell_start, V, true, scratch, st);
    if (rc) return rc;
    if (n > 0) {
        grid_place_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(n, pt_cell, pt_rank, cell_start, tmp);
        PN_LAUNCH_CHECK("grid_place");
        grid_order_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(n, g, pt_cell, cell_start, tmp, bucket_pidx, occ_bits, info,
                                                                  frustum ? 0 : 1);
        PN_LAUNCH_CHECK("grid_order");
    }
    grid_finish_kernel<<<1, 1, 0, st>>>(g, info);
    PN_LAUNCH_CHECK("grid_finish");
    return PN_OK;
}

extern "C" int pn_grid_build(const float* xyz, int64_t n_points, const pn_grid_params* params, int32_t* cell_start,
                             int32_t* bucket_pidx, uint32_t* occ_bits, int32_t* info, void* workspace,
                             int64_t workspace_bytes, void* stream) {
    return grid_build_impl(xyz, n_points, params, cell_start, bucket_pidx, occ_bits, info, workspace, workspace_bytes, stream, false);
}

extern "C" int pn_frustum_grid_build(const float* xyz_pers, int64_t n_points, const pn_grid_params* params, int32_t* cell_start,
                                     int32_t* bucket_pidx, uint32_t* occ_bits, int32_t* info, void* workspace,
                                     int64_t workspace_bytes, void* stream) {
    return grid_build_impl(xyz_pers, n_points, params, cell_start, bucket_pidx, occ_bits, info, workspace, workspace_bytes, stream, true);
}
