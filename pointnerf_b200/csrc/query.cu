// Ray sampling against the dilated occupancy bitset and the layered K-neighbour query.
//
// Replaces mask_raypos / get_shadingloc / query_neigh_along_ray_layered and the host glue between
// them (models/neural_points/query_point_indices_worldcoords.py:390-528, 633-721).
#include "common.cuh"

namespace pn {

struct QGrid {
    float sx, sy, sz, vx, vy, vz;
    int dx, dy, dz, P;
};
static QGrid to_q(const pn_grid_params* p) {
    QGrid g;
    g.sx = p->shift[0]; g.sy = p->shift[1]; g.sz = p->shift[2];
    g.vx = p->vsize[0]; g.vy = p->vsize[1]; g.vz = p->vsize[2];
    g.dx = p->dims[0]; g.dy = p->dims[1]; g.dz = p->dims[2];
    g.P = p->P;
    return g;
}

// ------------------------------------------------------------------------------------------------
// Ray sampling.  One warp per ray, 32 candidate depths per step: position = campos + raydir*t with
// separately rounded mul/add (torch evaluates diff_ray_marching.py:387 as two kernels), voxel test
// against the bitset (1 bit per candidate instead of the reference's 12-byte raypos + 4-byte mask),
// ballot + popc give each hit its slot.  The walk stops at SR hits.
// Algorithmic bytes: 12R (dirs) + R*D/8 (bits) + 16 R'*SR (xyz + t out) + 4R.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ray_sample_kernel(const float* __restrict__ campos, const float* __restrict__ raydir,
                                                         const float* __restrict__ ts, int64_t ts_stride, int64_t n_rays,
                                                         int D, int SR, QGrid g, const unsigned* __restrict__ occ_bits,
                                                         float* __restrict__ loc_w, float* __restrict__ sample_t,
                                                         int* __restrict__ hit_count) {
    int64_t ray = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (ray >= n_rays) return;
    float ox = campos[0], oy = campos[1], oz = campos[2];
    float rx = raydir[3 * ray], ry = raydir[3 * ray + 1], rz = raydir[3 * ray + 2];
    const float* tr = ts + ray * ts_stride;
    float* out = loc_w + ray * SR * 3;
    int kept = 0;
    for (int d0 = 0; d0 < D && kept < SR; d0 += 32) {
        int d = d0 + lane;
        bool hit = false;
        float t = 0.f, px = 0.f, py = 0.f, pz = 0.f;
        if (d < D) {
            t = tr[d];
            px = __fadd_rn(ox, __fmul_rn(rx, t));
            py = __fadd_rn(oy, __fmul_rn(ry, t));
            pz = __fadd_rn(oz, __fmul_rn(rz, t));
            int cx = voxel_coord(px, g.sx, g.vx), cy = voxel_coord(py, g.sy, g.vy), cz = voxel_coord(pz, g.sz, g.vz);
            if (cx >= 0 && cx < g.dx && cy >= 0 && cy < g.dy && cz >= 0 && cz < g.dz) {
                int64_t b = ((int64_t)cx * g.dy + cy) * g.dz + cz;
                hit = (occ_bits[b >> 5] >> (b & 31)) & 1u;
            }
        }
        unsigned m = __ballot_sync(kFull, hit);
        int slot = kept + __popc(m & ((1u << lane) - 1u));
        if (hit && slot < SR) {
            out[3 * slot] = px;
            out[3 * slot + 1] = py;
            out[3 * slot + 2] = pz;
            if (sample_t) sample_t[ray * SR + slot] = t;
        }
        kept += __popc(m);
    }
    kept = min(kept, SR);
    // unused slots are zero, as the reference's torch.zeros sample_loc_tensor (:657)
    for (int s = kept + lane; s < SR; s += 32) {
        out[3 * s] = 0.f;
        out[3 * s + 1] = 0.f;
        out[3 * s + 2] = 0.f;
        if (sample_t) sample_t[ray * SR + s] = 0.f;
    }
    if (lane == 0) hit_count[ray] = kept;
}

// ------------------------------------------------------------------------------------------------
// Frustum sampling (perspective grid, query_point_indices.py).  One warp per ray walks the z column of the ray's
// (pixel / vscale) frustum cell through the dilated occupancy bitset, 32 depths per step, and keeps the first SR occupied
// depth indices - what near_vox_full (:316-365) stores per column in voxel_to_coorz_idx; ray_mask is its `far_id > 0`
// (:339: the deepest occupied index must be positive, so a column occupied only at depth 0 is masked out).
// Algorithmic bytes: 8R (pixels) + R*dz/8 (bits) + 4 R*SR (depth ids out) + 8R.
// ------------------------------------------------------------------------------------------------
struct FrustumArgs {
    const int* pixel_idx;   // [R,2] (x, y)
    const int* sample_z;    // [R,SR] depth index of every kept sample, -1 unused
    float* sample_loc;      // [R,SR,3] perspective coordinates of the sample centres (written for every slot, :532-540)
    float rvx, rvy;         // ray voxel size: scaled voxel size / vscale (:707)
    float depth_limit2;
    int vsx, vsy;           // vscale
    int nn;                 // opt.NN
    int z_layers;           // (kernel_size[2]+1)/2
};

__global__ void __launch_bounds__(256) frustum_sample_kernel(const int* __restrict__ pixel_idx, int64_t n_rays, int SR, QGrid g,
                                                             int vsx, int vsy, const unsigned* __restrict__ occ_bits,
                                                             int* __restrict__ sample_z, int* __restrict__ hit_count,
                                                             int* __restrict__ ray_mask) {
    int64_t ray = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (ray >= n_rays) return;
    int cx = pixel_idx[2 * ray] / vsx, cy = pixel_idx[2 * ray + 1] / vsy;
    int* out = sample_z + ray * SR;
    int kept = 0;
    bool deep = false;   // an occupied depth index > 0 exists
    if (cx >= 0 && cx < g.dx && cy >= 0 && cy < g.dy) {
        const int64_t base = ((int64_t)cx * g.dy + cy) * g.dz;
        for (int z0 = 0; z0 < g.dz && (kept < SR || !deep); z0 += 32) {
            int z = z0 + lane;
            bool hit = false;
            if (z < g.dz) {
                int64_t b = base + z;
                hit = (occ_bits[b >> 5] >> (b & 31)) & 1u;
            }
            unsigned m = __ballot_sync(kFull, hit);
            int slot = kept + __popc(m & ((1u << lane) - 1u));
            if (hit && slot < SR) out[slot] = z;
            kept += __popc(m);
            deep |= (z0 == 0 ? (m & ~1u) : m) != 0u;
        }
    }
    // a ray the reference drops (mask 0) keeps no samples: drivers that do not compact rays see it as a miss
    kept = deep ? min(kept, SR) : 0;
    for (int s = kept + lane; s < SR; s += 32) out[s] = -1;
    if (lane == 0) {
        hit_count[ray] = kept;
        ray_mask[ray] = deep ? 1 : 0;
    }
}

// ------------------------------------------------------------------------------------------------
// Neighbour query.  One warp per (ray, slot).  Per layer the (2l+1)^3 cube is enumerated in x,y,z
// order 32 cells at a time: each lane fetches one cell's [start, end) (non-ring / out-of-grid cells
// count 0), a warp prefix sum lays the cells' buckets out as one candidate sequence in the
// reference's visit order, and the sequence is consumed 32 candidates per step: every lane loads one
// point and evaluates d^2 in parallel (the loads are the cost), then the in-radius candidates are
// inserted ONE BY ONE in sequence order by warp-uniform code that mirrors :499-519 - lane k holds
// slot k - so the slot arrangement, the strict-'<' replacement rule and the first-maximum rescan are
// reproduced exactly.  Algorithmic bytes: S_q*(12 + 27*8) + 16*Cand + 4*S_q*K.
// ------------------------------------------------------------------------------------------------
//
// kFrustum = true is the perspective-grid variant (query_point_indices.py:495-600): the query centre comes from the ray's
// pixel and the sample's depth index (:532-536: float partial sums, the half-voxel offsets in double), the cube is
// (2l+1)^2 x (2*min(l, z_layers-1)+1) with the visit rule of :555, distances are taken in perspective coordinates (NN < 2) or
// after multiplying x and y by z (NN >= 2, :565-566), the radius limit applies to the xy part and a separate depth limit
// to the z part (:571), there is no early exit between layers, and the sample centre is written out for every slot.
template <bool kCount, bool kFrustum>
__global__ void __launch_bounds__(256) knn_kernel(const float* __restrict__ xyz, QGrid g, const int* __restrict__ cell_start,
                                                  const int* __restrict__ bucket, const int* __restrict__ grid_info,
                                                  const float* __restrict__ loc_w, const int* __restrict__ hit_count,
                                                  int64_t n_rays, int SR, int K, int n_layers, float r2,
                                                  int* __restrict__ sample_pidx, int* __restrict__ ray_any,
                                                  unsigned long long* __restrict__ n_cand, FrustumArgs fa) {
    // one block per ray: its warps take the hit slots round-robin, then the block blanks the unused slots
    const int64_t ray = blockIdx.x;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int hits = min(hit_count[ray], SR);
    for (int i = hits * K + (int)threadIdx.x; i < SR * K; i += blockDim.x) sample_pidx[ray * SR * K + i] = -1;
    bool any_seen = false;
    // frustum: every slot of the ray (kept or not) gets its centre written, so the loop covers all SR slots
    for (int slot = wib; slot < (kFrustum ? SR : hits); slot += nw) {
    const int64_t w = ray * SR + slot;
    int* out = sample_pidx + w * K;
    float qx, qy, qz, qxz = 0.f, qyz = 0.f;
    int fx, fy, fz, slot0_cell = -1;
    if (kFrustum) {
        const int px = fa.pixel_idx[2 * ray], py = fa.pixel_idx[2 * ray + 1];
        fx = px / fa.vsx; fy = py / fa.vsy; fz = fa.sample_z[w];
        qx = __double2float_rn(__fma_rn((double)(px % fa.vsx) + 0.5, (double)fa.rvx, (double)__fmaf_rn(g.vx, (float)fx, g.sx)));
        qy = __double2float_rn(__fma_rn((double)(py % fa.vsy) + 0.5, (double)fa.rvy, (double)__fmaf_rn(g.vy, (float)fy, g.sy)));
        qz = __double2float_rn(__fma_rn((double)fz + 0.5, (double)g.vz, (double)g.sz));
        if (lane == 0) { fa.sample_loc[3 * w] = qx; fa.sample_loc[3 * w + 1] = qy; fa.sample_loc[3 * w + 2] = qz; }
        if (slot >= hits) continue;   // fz < 0 (:541); the block-wide fill above has blanked the indices
        qxz = __fmul_rn(qz, qx); qyz = __fmul_rn(qz, qy);
    } else {
        qx = loc_w[3 * w]; qy = loc_w[3 * w + 1]; qz = loc_w[3 * w + 2];
        fx = voxel_coord(qx, g.sx, g.vx); fy = voxel_coord(qy, g.sy, g.vy); fz = voxel_coord(qz, g.sz, g.vz);
        slot0_cell = grid_info[PN_GI_SLOT0_CELL];
    }

    // lane k (< K; K <= 32 per register set, larger K loops) keeps slot k
    // K up to 64 is supported with two slots per lane
    int my_pid[2] = {-1, -1};
    float my_d2[2] = {0.f, 0.f};
    int seen = 0, far_slot = 0;
    float far2 = 0.f;
    unsigned long long visited = 0;

    for (int layer = 0; layer < n_layers; ++layer) {
        const int zlayer = kFrustum ? min(fa.z_layers - 1, layer) : layer;
        int side = 2 * layer + 1, zside = 2 * zlayer + 1, cube = side * side * zside;
        for (int c0 = 0; c0 < cube; c0 += 32) {
            // --- one cell per lane
            int ci = c0 + lane, start = 0, cnt = 0;
            if (ci < cube) {
                int ddx = ci / (side * zside) - layer, rem = ci % (side * zside);
                int ddy = rem / zside - layer, ddz = rem % zside - zlayer;
                bool visit;
                if (kFrustum) visit = max(abs(ddx), abs(ddy)) == layer || (zlayer == layer && abs(ddz) == zlayer);   // :555
                else visit = max(abs(ddx), max(abs(ddy), abs(ddz))) == layer;
                int cx = fx + ddx, cy = fy + ddy, cz = fz + ddz;
                if (visit && cx >= 0 && cx < g.dx && cy >= 0 && cy < g.dy && cz >= 0 && cz < g.dz) {
                    int cell = (cx * g.dy + cy) * g.dz + cz;
                    start = cell_start[cell];
                    cnt = min(cell_start[cell + 1] - start, g.P);
                    if (cell == slot0_cell) cnt = 0;  // `if (voxel_idx > 0)` of fill_occ2pnts (:372)
                }
            }
            int incl = warp_scan_incl(cnt, lane);
            int excl = incl - cnt;
            int total = __shfl_sync(kFull, incl, 31);
            if (kCount) visited += total;
            // --- candidates, 32 per step, in sequence order
            for (int b0 = 0; b0 < total; b0 += 32) {
                int c = b0 + lane;
                // owner cell of candidate c: the lane j with excl_j <= c < incl_j (binary search by shuffles)
                int lo = 0;
#pragma unroll
                for (int step = 16; step > 0; step >>= 1) {
                    int probe = __shfl_sync(kFull, excl, lo + step);  // excl is non-decreasing
                    if (probe <= c) lo += step;
                }
                // lanes with empty cells share excl with their successor: move to the last lane whose excl <= c
                // (binary search already lands on the LAST lane with excl <= c, whose cell is non-empty when c < total)
                int o_start = __shfl_sync(kFull, start, lo), o_excl = __shfl_sync(kFull, excl, lo);
                bool in = false;
                int pid = -1;
                float d2 = 0.f;
                if (c < total) {
                    pid = bucket[o_start + (c - o_excl)];
                    if (kFrustum) {
                        float px = xyz[3 * (int64_t)pid], py = xyz[3 * (int64_t)pid + 1], pz = xyz[3 * (int64_t)pid + 2];
                        float x = fa.nn < 2 ? __fsub_rn(px, qx) : __fsub_rn(__fmul_rn(px, pz), qxz);
                        float y = fa.nn < 2 ? __fsub_rn(py, qy) : __fsub_rn(__fmul_rn(py, pz), qyz);
                        float xy2 = __fmaf_rn(x, x, __fmul_rn(y, y));
                        float z = __fsub_rn(pz, qz);
                        float z2 = __fmul_rn(z, z);
                        d2 = __fadd_rn(xy2, z2);
                        in = (r2 == 0.f || xy2 <= r2) && (fa.depth_limit2 == 0.f || z2 <= fa.depth_limit2);
                    } else {
                        float x = __fsub_rn(xyz[3 * (int64_t)pid], qx);
                        float y = __fsub_rn(xyz[3 * (int64_t)pid + 1], qy);
                        float z = __fsub_rn(xyz[3 * (int64_t)pid + 2], qz);
                        d2 = __fmaf_rn(z, z, __fmaf_rn(x, x, __fmul_rn(y, y)));
                        in = (r2 == 0.f) || (d2 <= r2);
                    }
                }
                unsigned m = __ballot_sync(kFull, in);
                while (m) {
                    if (seen >= K) {   // slots full: only candidates closer than the current maximum can change anything,
                        m &= __ballot_sync(kFull, in && d2 < far2);   // and that maximum never grows - drop the rest at once
                        if (!m) break;
                    }
                    int src = __ffs(m) - 1;
                    m &= m - 1;
                    float cd = __shfl_sync(kFull, d2, src);
                    int cp = __shfl_sync(kFull, pid, src);
                    if (seen < K) {
                        if (lane == (seen & 31)) { my_pid[seen >> 5] = cp; my_d2[seen >> 5] = cd; }
                        if (cd > far2) { far2 = cd; far_slot = seen; }
                        ++seen;
                    } else {
                        ++seen;
                        if (cd < far2) {
                            if (lane == (far_slot & 31)) { my_pid[far_slot >> 5] = cp; my_d2[far_slot >> 5] = cd; }
                            // rescan: first slot holding the maximum, only if it exceeds the new value (:511-517)
                            float v0 = (lane < K) ? my_d2[0] : -1.f;
                            float v1 = (lane + 32 < K) ? my_d2[1] : -1.f;
                            float mx = fmaxf(v0, v1);
#pragma unroll
                            for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(kFull, mx, o));
                            far2 = cd;
                            if (mx > cd) {
                                unsigned b0m = __ballot_sync(kFull, v0 == mx);
                                unsigned b1m = __ballot_sync(kFull, v1 == mx);
                                far_slot = b0m ? (__ffs(b0m) - 1) : (32 + __ffs(b1m) - 1);
                                far2 = mx;
                            }
                        }
                    }
                }
            }
        }
        if (!kFrustum && seen >= K) break;  // tested only after a whole layer (:526); the frustum kernel has no such exit
    }
    for (int k = lane; k < K; k += 32) out[k] = my_pid[k >> 5];
    any_seen |= seen > 0;
    if (kCount && n_cand && lane == 0) atomicAdd(n_cand, visited);
    }
    if (!kFrustum && lane == 0 && any_seen) ray_any[ray] = 1;  // benign same-value race across the ray's warps
}

__global__ void __launch_bounds__(256) w2pers_kernel(const float* __restrict__ xyz, int64_t n, const float* __restrict__ rot,
                                                     const float* __restrict__ campos, float* __restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // cam_j = sum_i R[i][j] * (p_i - c_i), summed in i order like torch.sum over 3 products
    float s0 = __fsub_rn(xyz[3 * i], campos[0]), s1 = __fsub_rn(xyz[3 * i + 1], campos[1]), s2 = __fsub_rn(xyz[3 * i + 2], campos[2]);
    float c[3];
#pragma unroll
    for (int j = 0; j < 3; ++j)
        c[j] = __fadd_rn(__fadd_rn(__fmul_rn(rot[j], s0), __fmul_rn(rot[3 + j], s1)), __fmul_rn(rot[6 + j], s2));
    out[3 * i] = __fdiv_rn(c[0], c[2]);
    out[3 * i + 1] = __fdiv_rn(c[1], c[2]);
    out[3 * i + 2] = c[2];
}

// ---- work lists -----------------------------------------------------------------------------
__global__ void __launch_bounds__(256) slot_count_kernel(const int* __restrict__ pidx, int64_t n_slots, int K,
                                                         unsigned long long* __restrict__ packed) {
    int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_slots) return;
    int c = 0;
    for (int k = 0; k < K; ++k) c += pidx[s * K + k] >= 0;
    packed[s] = ((unsigned long long)(c > 0) << 32) | (unsigned)c;  // (valid sample, pairs)
}

__global__ void __launch_bounds__(256) slot_scatter_kernel(const int* __restrict__ pidx, int64_t n_slots, int K,
                                                           const unsigned long long* __restrict__ offs,
                                                           const unsigned long long* __restrict__ total,
                                                           int* __restrict__ sample_list, int* __restrict__ pair_start,
                                                           int* __restrict__ pair_src, int* __restrict__ counts,
                                                           int* __restrict__ slot_row, int64_t S_cap, int64_t M_cap, int capped) {
    // Capacities (capped != 0): the lists keep the longest PREFIX of valid samples (in slot order) that fits S_cap samples
    // and M_cap pairs; later samples are dropped whole (slot_row = -1), so every list the aggregator walks stays inside
    // the buffers sized by the capacities.  counts = {S kept, M kept, S total, M total}: kept < total flags the overflow.
    // (counts[0], counts[1] are zeroed by the host before the launch; offsets grow with the slot index, so the maxima
    // below are the kept prefix's totals and the kept samples' own pair_start[si + 1] writes close the list.)
    int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s == 0) {
        unsigned long long t = *total;
        if (capped) {
            counts[2] = (int)(t >> 32);
            counts[3] = (int)(t & 0xffffffffu);
            pair_start[0] = 0;
        } else {
            counts[0] = (int)(t >> 32);
            counts[1] = (int)(t & 0xffffffffu);
            pair_start[t >> 32] = (int)(t & 0xffffffffu);
        }
    }
    if (s >= n_slots) return;
    unsigned long long o = offs[s];
    int si = (int)(o >> 32), pi = (int)(o & 0xffffffffu);
    int c = 0;
    for (int k = 0; k < K; ++k) c += pidx[s * K + k] >= 0;
    const bool any = c > 0;
    const bool keep = any && (!capped || ((int64_t)si < S_cap && (int64_t)pi + c <= M_cap));
    if (keep) {
        const int p0 = pi;
        for (int k = 0; k < K; ++k)
            if (pidx[s * K + k] >= 0) pair_src[pi++] = (int)(s * K + k);
        sample_list[si] = (int)s;
        pair_start[si] = p0;
        if (capped) {
            pair_start[si + 1] = pi;          // the next kept sample writes the same value; the last kept one closes the list
            atomicMax(&counts[0], si + 1);
            atomicMax(&counts[1], pi);
        }
    }
    if (slot_row) slot_row[s] = keep ? si : -1;
}

}  // namespace pn

using namespace pn;

namespace pn {
// Jittered mid-point depths: the scalar chain of near_far_linear_ray_generation (diff_ray_marching.py:369-386) as the
// reference's torch ops evaluate it ON THE HOST - the canonical chain: every product and sum is rounded
// separately in fp32, except the running sum of torch.cumsum, which torch's CPU kernel accumulates in DOUBLE
// (at::acc_type<float, /*is_cuda=*/false>) left to right and rounds to fp32 per element.  (torch's CUDA cumsum is a
// tree scan in fp32 whose association depends on the launch geometry: not reproducible, not a parity target.)
__global__ void __launch_bounds__(256) sample_depths_kernel(const float* __restrict__ seg_base, const float* __restrict__ uniform,
                                                            int64_t n_rays, int D, float near, float jitter, float* __restrict__ ts) {
    // one warp per ray: lanes load / store 32 consecutive depths (coalesced) and form their segment lengths in parallel;
    // the running sum itself is taken strictly left to right - every lane walks the same 32-step chain over the
    // shuffled segment values (SIMT: the cost of one lane) and keeps the step that is its own
    int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (r >= n_rays) return;
    const float* u = uniform ? uniform + r * D : nullptr;
    float* out = ts + r * D;
    // The running sum is a chain of double additions of floats.  When every non-zero segment of the ray has its exponent
    // within 2^20 of the others (always, for a near/far range cut into D bins and jittered by +-15 %), every partial sum
    // of <= 2^9 segments is a multiple of the smallest segment's ulp and below 2^(emax + 10): it is EXACT in double, in any
    // association - so a warp-parallel scan gives the bits of the sequential chain.  Rays that fail the test (denormals,
    // wildly different segments) take the sequential chain.
    int emin = 255, emax = 0;
    bool odd = false;
    for (int d = lane; d < D; d += 32) {
        const float seg = __fmul_rn(seg_base[d], __fadd_rn(1.f, __fmul_rn(jitter, __fsub_rn(u ? u[d] : 0.5f, 0.5f))));
        const int e = (__float_as_int(seg) >> 23) & 0xff;
        if (seg != 0.f) { emin = min(emin, e); emax = max(emax, e); odd |= (e == 0 || e == 255); }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        emin = min(emin, __shfl_xor_sync(0xffffffffu, emin, o));
        emax = max(emax, __shfl_xor_sync(0xffffffffu, emax, o));
    }
    const bool exact = !__any_sync(0xffffffffu, odd) && emax - emin <= 20 && D <= 512;
    double c = 0.0;
    float prev_end = __fadd_rn(near, 0.f);
    if (exact) {
        for (int d0 = 0; d0 < D; d0 += 32) {
            const int d = d0 + lane;
            float seg = 0.f;
            if (d < D) seg = __fmul_rn(seg_base[d], __fadd_rn(1.f, __fmul_rn(jitter, __fsub_rn(u ? u[d] : 0.5f, 0.5f))));
            double incl = (double)seg;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const double t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            incl += c;                                               // exact: see above
            const float end = __fadd_rn(near, (float)incl);
            float before = __shfl_up_sync(0xffffffffu, end, 1);
            if (lane == 0) before = prev_end;
            if (d < D) out[d] = __fmul_rn(__fadd_rn(before, end), 0.5f);
            c = __shfl_sync(0xffffffffu, incl, 31);
            prev_end = __shfl_sync(0xffffffffu, end, 31);
        }
        return;
    }
    for (int d0 = 0; d0 < D; d0 += 32) {
        const int d = d0 + lane;
        float seg = 0.f;
        if (d < D) seg = __fmul_rn(seg_base[d], __fadd_rn(1.f, __fmul_rn(jitter, __fsub_rn(u ? u[d] : 0.5f, 0.5f))));
        float mine = 0.f;
        const int n = D - d0 < 32 ? D - d0 : 32;
        for (int j = 0; j < n; ++j) {
            const float v = __shfl_sync(0xffffffffu, seg, j);
            c = __dadd_rn(c, (double)v);
            const float end = __fadd_rn(near, (float)c);
            const float mid = __fmul_rn(__fadd_rn(prev_end, end), 0.5f);
            if (lane == j) mine = mid;
            prev_end = end;
        }
        if (d < D) out[d] = mine;
    }
}
}  // namespace pn

extern "C" int pn_sample_depths(const float* seg_base, const float* uniform, int64_t n_rays, int32_t D, float near_depth, float jitter,
                                float* ts, void* stream) {
    PN_CHECK_ARG(seg_base && ts && D > 0, "pn_sample_depths: bad argument");   // uniform == NULL: no jitter (every draw 0.5)
    if (n_rays <= 0) return PN_OK;
    sample_depths_kernel<<<(unsigned)cdiv(n_rays * 32, 256), 256, 0, (cudaStream_t)stream>>>(seg_base, uniform, n_rays, D, near_depth, jitter, ts);
    PN_LAUNCH_CHECK("sample_depths");
    return PN_OK;
}

extern "C" int pn_ray_sample(const float* campos, const float* raydir, const float* ts, int64_t ts_rows, int64_t n_rays,
                             int32_t D, int32_t SR, const pn_grid_params* params, const uint32_t* occ_bits,
                             float* sample_loc_w, float* sample_t, int32_t* hit_count, void* stream) {
    PN_CHECK_ARG(campos && raydir && ts && params && occ_bits && sample_loc_w && hit_count, "pn_ray_sample: null argument");
    PN_CHECK_ARG(ts_rows == 1 || ts_rows == n_rays, "pn_ray_sample: ts_rows must be 1 or n_rays");
    PN_CHECK_ARG(D >= 0 && SR > 0, "pn_ray_sample: bad D/SR");
    if (n_rays <= 0) return PN_OK;
    ray_sample_kernel<<<(unsigned)cdiv(n_rays * 32, 256), 256, 0, (cudaStream_t)stream>>>(
        campos, raydir, ts, ts_rows == 1 ? 0 : D, n_rays, D, SR, to_q(params), occ_bits, sample_loc_w, sample_t, hit_count);
    PN_LAUNCH_CHECK("ray_sample");
    return PN_OK;
}

namespace pn { int g_knn_threads = 256; }
extern "C" int pn_knn_query(const float* xyz, const pn_grid_params* params, const int32_t* cell_start,
                            const int32_t* bucket_pidx, const int32_t* grid_info, const float* sample_loc_w,
                            const int32_t* hit_count, int64_t n_rays, int32_t SR, int32_t K, int32_t n_layers,
                            float radius_limit2, int32_t* sample_pidx, int32_t* ray_any, unsigned long long* n_candidates,
                            void* stream) {
    PN_CHECK_ARG(xyz && params && cell_start && bucket_pidx && grid_info && sample_loc_w && hit_count && sample_pidx && ray_any,
                 "pn_knn_query: null argument");
    PN_CHECK_ARG(K >= 1 && K <= 64, "pn_knn_query: K=%d unsupported (1..64)", K);
    PN_CHECK_ARG(n_layers >= 1 && n_layers <= 8, "pn_knn_query: n_layers=%d unsupported", n_layers);
    if (n_rays <= 0) return PN_OK;
    cudaStream_t st = (cudaStream_t)stream;
    PN_CUDA(cudaMemsetAsync(ray_any, 0, n_rays * sizeof(int), st));
    unsigned blocks = (unsigned)n_rays;
    // (pn_debug_gemm_flags(2048) selects 128-thread blocks, which fit next to a resident GEMM CTA of 576 threads x 96
    // registers when the next batch's query runs under this batch's backward (ray_shard.py); measured: no difference there,
    // and 110 us instead of 74 us when the query runs alone)
    const int kKnnThreads = g_knn_threads;
    if (n_candidates)
        knn_kernel<true, false><<<blocks, kKnnThreads, 0, st>>>(xyz, to_q(params), cell_start, bucket_pidx, grid_info, sample_loc_w, hit_count,
                                                n_rays, SR, K, n_layers, radius_limit2, sample_pidx, ray_any, n_candidates, FrustumArgs{});
    else
        knn_kernel<false, false><<<blocks, kKnnThreads, 0, st>>>(xyz, to_q(params), cell_start, bucket_pidx, grid_info, sample_loc_w, hit_count,
                                                 n_rays, SR, K, n_layers, radius_limit2, sample_pidx, ray_any, nullptr, FrustumArgs{});
    PN_LAUNCH_CHECK("knn_query");
    return PN_OK;
}

extern "C" int pn_frustum_sample(const int32_t* pixel_idx, int64_t n_rays, int32_t SR, const pn_grid_params* params,
                                 const int32_t* vscale_xy, const uint32_t* occ_bits, int32_t* sample_z, int32_t* hit_count,
                                 int32_t* ray_mask, void* stream) {
    PN_CHECK_ARG(pixel_idx && params && vscale_xy && occ_bits && sample_z && hit_count && ray_mask, "pn_frustum_sample: null argument");
    PN_CHECK_ARG(SR > 0 && vscale_xy[0] > 0 && vscale_xy[1] > 0, "pn_frustum_sample: bad SR / vscale");
    if (n_rays <= 0) return PN_OK;
    frustum_sample_kernel<<<(unsigned)cdiv(n_rays * 32, 256), 256, 0, (cudaStream_t)stream>>>(
        pixel_idx, n_rays, SR, to_q(params), vscale_xy[0], vscale_xy[1], occ_bits, sample_z, hit_count, ray_mask);
    PN_LAUNCH_CHECK("frustum_sample");
    return PN_OK;
}

extern "C" int pn_frustum_knn_query(const float* xyz_pers, const pn_grid_params* params, const pn_frustum_query* q,
                                    const int32_t* cell_start, const int32_t* bucket_pidx, const int32_t* pixel_idx,
                                    const int32_t* sample_z, const int32_t* hit_count, int64_t n_rays, int32_t SR, int32_t K,
                                    int32_t* sample_pidx, float* sample_loc, unsigned long long* n_candidates, void* stream) {
    PN_CHECK_ARG(xyz_pers && params && q && cell_start && bucket_pidx && pixel_idx && sample_z && hit_count && sample_pidx && sample_loc,
                 "pn_frustum_knn_query: null argument");
    PN_CHECK_ARG(K >= 1 && K <= 64, "pn_frustum_knn_query: K=%d unsupported (1..64)", K);
    PN_CHECK_ARG(q->n_layers >= 1 && q->n_layers <= 8 && q->z_layers >= 1, "pn_frustum_knn_query: n_layers=%d z_layers=%d unsupported",
                 q->n_layers, q->z_layers);
    PN_CHECK_ARG(q->vscale[0] > 0 && q->vscale[1] > 0 && q->nn >= 1, "pn_frustum_knn_query: bad vscale / NN");
    if (n_rays <= 0) return PN_OK;
    cudaStream_t st = (cudaStream_t)stream;
    FrustumArgs fa;
    fa.pixel_idx = pixel_idx; fa.sample_z = sample_z; fa.sample_loc = sample_loc;
    fa.rvx = q->ray_vsize[0]; fa.rvy = q->ray_vsize[1];
    fa.depth_limit2 = q->depth_limit2;
    fa.vsx = q->vscale[0]; fa.vsy = q->vscale[1];
    fa.nn = q->nn; fa.z_layers = q->z_layers;
    unsigned blocks = (unsigned)n_rays;
    if (n_candidates)
        knn_kernel<true, true><<<blocks, 256, 0, st>>>(xyz_pers, to_q(params), cell_start, bucket_pidx, nullptr, nullptr, hit_count, n_rays,
                                                       SR, K, q->n_layers, q->radius_limit2, sample_pidx, nullptr, n_candidates, fa);
    else
        knn_kernel<false, true><<<blocks, 256, 0, st>>>(xyz_pers, to_q(params), cell_start, bucket_pidx, nullptr, nullptr, hit_count, n_rays,
                                                        SR, K, q->n_layers, q->radius_limit2, sample_pidx, nullptr, nullptr, fa);
    PN_LAUNCH_CHECK("frustum_knn_query");
    return PN_OK;
}

extern "C" int pn_w2pers(const float* xyz_w, int64_t n, const float* camrot, const float* campos, float* xyz_pers, void* stream) {
    PN_CHECK_ARG(xyz_w && camrot && campos && xyz_pers, "pn_w2pers: null argument");
    if (n <= 0) return PN_OK;
    w2pers_kernel<<<(unsigned)cdiv(n, 256), 256, 0, (cudaStream_t)stream>>>(xyz_w, n, camrot, campos, xyz_pers);
    PN_LAUNCH_CHECK("w2pers");
    return PN_OK;
}

extern "C" int64_t pn_compact_workspace_bytes(int64_t n_slots) {
    int64_t n = n_slots > 0 ? n_slots : 1;
    return 2 * align_up(n * 8, 256) + align_up(scan_scratch_elems(n) * 8, 256) + 512;
}

extern "C" int pn_compact_samples(const int32_t* sample_pidx, int64_t n_slots, int32_t K, int32_t* sample_list,
                                  int32_t* pair_start, int32_t* pair_src, int32_t* counts, int32_t* slot_row, void* workspace,
                                  int64_t workspace_bytes, void* stream) {
    PN_CHECK_ARG(sample_pidx && sample_list && pair_start && pair_src && counts, "pn_compact_samples: null argument");
    PN_CHECK_ARG(workspace_bytes >= pn_compact_workspace_bytes(n_slots), "pn_compact_samples: workspace too small");
    PN_CHECK_ARG(n_slots * (int64_t)K < 2147483647LL, "pn_compact_samples: n_slots*K exceeds int32");
    cudaStream_t st = (cudaStream_t)stream;
    if (n_slots <= 0) {
        PN_CUDA(cudaMemsetAsync(counts, 0, 2 * sizeof(int), st));
        PN_CUDA(cudaMemsetAsync(pair_start, 0, sizeof(int), st));
        return PN_OK;
    }
    Carver ws(workspace, workspace_bytes);
    unsigned long long* packed = ws.take<unsigned long long>(n_slots);
    unsigned long long* offs = ws.take<unsigned long long>(n_slots);
    unsigned long long* scratch = ws.take<unsigned long long>(scan_scratch_elems(n_slots));
    unsigned long long* total = ws.take<unsigned long long>(1);
    slot_count_kernel<<<(unsigned)cdiv(n_slots, 256), 256, 0, st>>>(sample_pidx, n_slots, K, packed);
    PN_LAUNCH_CHECK("slot_count");
    int rc = scan_exclusive_u64(packed, offs, n_slots, total, scratch, st);
    if (rc) return rc;
    slot_scatter_kernel<<<(unsigned)cdiv(n_slots, 256), 256, 0, st>>>(sample_pidx, n_slots, K, offs, total, sample_list,
                                                                      pair_start, pair_src, counts, slot_row, 0, 0, 0);
    PN_LAUNCH_CHECK("slot_scatter");
    return PN_OK;
}

extern "C" int pn_compact_samples_capped(const int32_t* sample_pidx, int64_t n_slots, int32_t K, int64_t S_cap, int64_t M_cap,
                                         int32_t* sample_list, int32_t* pair_start, int32_t* pair_src, int32_t* counts4,
                                         int32_t* slot_row, void* workspace, int64_t workspace_bytes, void* stream) {
    PN_CHECK_ARG(sample_pidx && sample_list && pair_start && pair_src && counts4, "pn_compact_samples_capped: null argument");
    PN_CHECK_ARG(S_cap >= 1 && M_cap >= K, "pn_compact_samples_capped: capacities must hold at least one sample (S_cap >= 1, M_cap >= K)");
    PN_CHECK_ARG(workspace_bytes >= pn_compact_workspace_bytes(n_slots), "pn_compact_samples_capped: workspace too small");
    PN_CHECK_ARG(n_slots * (int64_t)K < 2147483647LL, "pn_compact_samples_capped: n_slots*K exceeds int32");
    cudaStream_t st = (cudaStream_t)stream;
    PN_CUDA(cudaMemsetAsync(counts4, 0, 4 * sizeof(int), st));
    PN_CUDA(cudaMemsetAsync(pair_start, 0, sizeof(int), st));
    if (n_slots <= 0) return PN_OK;
    Carver ws(workspace, workspace_bytes);
    unsigned long long* packed = ws.take<unsigned long long>(n_slots);
    unsigned long long* offs = ws.take<unsigned long long>(n_slots);
    unsigned long long* scratch = ws.take<unsigned long long>(scan_scratch_elems(n_slots));
    unsigned long long* total = ws.take<unsigned long long>(1);
    slot_count_kernel<<<(unsigned)cdiv(n_slots, 256), 256, 0, st>>>(sample_pidx, n_slots, K, packed);
    PN_LAUNCH_CHECK("slot_count");
    int rc = scan_exclusive_u64(packed, offs, n_slots, total, scratch, st);
    if (rc) return rc;
    slot_scatter_kernel<<<(unsigned)cdiv(n_slots, 256), 256, 0, st>>>(sample_pidx, n_slots, K, offs, total, sample_list,
                                                                      pair_start, pair_src, counts4, slot_row, S_cap, M_cap, 1);
    PN_LAUNCH_CHECK("slot_scatter");
    return PN_OK;
}

// ---- pn_compact_samples without any allocation by anybody ---------------------------------------------------------------
// (no counterpart in the reference.)  Some callers cannot allocate device
// memory around a call - code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc - and so
// cannot hand pn_compact_samples its workspace.  This variant keeps the workspace in a file-scope device array of a fixed,
// modest size: every GPU that loads the library holds these 8.4 MB until the process exits.  In HOST code the array's name
// is a host-side stand-in, not a device address, so the address is asked for once per device (with that device current)
// through cudaGetSymbolAddress and that pointer is what the kernels get.
namespace pn {
constexpr int64_t kStaticCompactMaxSlots = 524288;                       // 4096 rays x 80 samples = 327 680; 8192 x 24 = 196 608
constexpr int64_t kStaticCompactBytes = 2 * kStaticCompactMaxSlots * 8 + 4096;   // packed + offs (uint64 per slot) + scan tile sums + total
static __device__ __align__(256) unsigned char g_compact_scratch[kStaticCompactBytes];
static void* g_compact_scratch_ptr[64] = {};                             // device address of the array, per device ordinal
}  // namespace pn

extern "C" int64_t pn_compact_samples_static_max_slots(void) { return kStaticCompactMaxSlots; }

extern "C" int pn_compact_samples_static(const int32_t* sample_pidx, int64_t n_slots, int32_t K, int32_t* sample_list,
                                         int32_t* pair_start, int32_t* pair_src, int32_t* counts, int32_t* slot_row, void* stream) {
    PN_CHECK_ARG(n_slots <= kStaticCompactMaxSlots,
                 "pn_compact_samples_static: n_slots = %lld exceeds the static scratch's maximum of %lld slots (use pn_compact_samples "
                 "with a caller-owned workspace)", (long long)n_slots, (long long)kStaticCompactMaxSlots);
    PN_CHECK_ARG(pn_compact_workspace_bytes(kStaticCompactMaxSlots) <= kStaticCompactBytes, "pn_compact_samples_static: scratch array undersized");
    int dev = -1;
    PN_CUDA(cudaGetDevice(&dev));
    PN_CHECK_ARG(dev >= 0 && dev < 64, "pn_compact_samples_static: device ordinal %d out of range", dev);
    if (!g_compact_scratch_ptr[dev]) {                                    // set-up, once per device, with that device current
        void* ptr = nullptr;
        PN_CUDA(cudaGetSymbolAddress(&ptr, g_compact_scratch));
        PN_CHECK_ARG(ptr != nullptr, "pn_compact_samples_static: cudaGetSymbolAddress returned a null address");
        g_compact_scratch_ptr[dev] = ptr;
    }
    // Every call shares the array and finds the previous call's contents in it.  Nothing has to start at zero: the per-slot
    // counts, the scanned offsets, the tile sums and the total are all written before they are read, for exactly n_slots
    // entries.  Calls that overlap on the GPU (two streams, two host threads) race on it.
    return pn_compact_samples(sample_pidx, n_slots, K, sample_list, pair_start, pair_src, counts, slot_row,
                              g_compact_scratch_ptr[dev], kStaticCompactBytes, stream);
}

