// Point-cloud maintenance between passes: probe outputs per ray, prune (stream compaction of
// the five point tables) and grow (append).  Reference:
//   probe  : models/neural_points_volumetric_model_ori.py:351-372 (per ray: max-opacity sample, its location, the
//            distance to the nearest of its K neighbours, weight*conf averages of the neighbours' colour/dir/conf/embedding)
//   prune  : models/neural_points/neural_points.py:350-373 (keep rows with points_conf >= thresh, boolean-mask indexing
//            of every table = five separate nonzero + gather passes)
//   grow   : models/neural_points/neural_points.py:376-402 (torch.cat of every table)
// All kernels are HBM-bound row movers: one warp per row, 128-bit accesses where the row width allows.
#include "common.cuh"

namespace pn {

// ------------------------------------------------------------------------------------------------------------
// probe: one warp per ray
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) probe_rays_kernel(
    const float* __restrict__ opacity, const int* __restrict__ ray_any, const float* __restrict__ loc_w,
    const int* __restrict__ pidx, const float* __restrict__ weight, const float* __restrict__ xyz,
    const float* __restrict__ emb, const float* __restrict__ color, const float* __restrict__ dir,
    const float* __restrict__ conf, int64_t R, int SR, int K, int F, float conf_lo, float conf_hi,
    float* __restrict__ o_opacity, int* __restrict__ o_slot, float* __restrict__ o_loc, float* __restrict__ o_far,
    float* __restrict__ o_color, float* __restrict__ o_dir, float* __restrict__ o_conf, float* __restrict__ o_emb) {
    const int lane = threadIdx.x & 31;
    const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= R) return;
    if (ray_any[r] <= 0) {  // rays the query dropped: `unmask` leaves zeros (:131-137)
        if (lane == 0) {
            o_opacity[r] = 0.f; o_slot[r] = -1; o_far[r] = 0.f; o_conf[r] = 0.f;
        }
        if (lane < 3) { o_loc[r * 3 + lane] = 0.f; o_color[r * 3 + lane] = 0.f; o_dir[r * 3 + lane] = 0.f; }
        for (int f = lane; f < F; f += 32) o_emb[r * F + f] = 0.f;
        return;
    }
    // torch.max(dim=-1): largest value, FIRST index among ties
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int s = lane; s < SR; s += 32) {
        float v = opacity[r * SR + s];
        if (v > best) { best = v; bi = s; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        float v2 = __shfl_xor_sync(kFull, best, o);
        int i2 = __shfl_xor_sync(kFull, bi, o);
        if (v2 > best || (v2 == best && i2 < bi)) { best = v2; bi = i2; }
    }
    const int64_t slot = r * SR + bi;
    const float lx = loc_w[slot * 3], ly = loc_w[slot * 3 + 1], lz = loc_w[slot * 3 + 2];
    // lane k < K holds neighbour k (empty slots read row 0 with weight 0, like the clamp(min=0) gather of the reference)
    int p = 0;
    float w = 0.f, d = INFINITY;
    if (lane < K) {
        int pi = pidx[slot * K + lane];
        p = pi < 0 ? 0 : pi;
        float c = conf[p];
        float cc = c - (c - fminf(fmaxf(c, conf_lo), conf_hi));      // gradiant_clamp forward value (point_aggregators.py:724-726)
        w = weight[slot * K + lane] * cc;
        float dx = xyz[(int64_t)p * 3] - lx, dy = xyz[(int64_t)p * 3 + 1] - ly, dz = xyz[(int64_t)p * 3 + 2] - lz;
        d = sqrtf(dx * dx + dy * dy + dz * dz);
    }
    float dmin = d;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dmin = fminf(dmin, __shfl_xor_sync(kFull, dmin, o));
    // weighted sums over k in ascending k (fixed order)
    float acc_c = 0.f, acc_d = 0.f, acc_f = 0.f;
    for (int k = 0; k < K; ++k) {
        int pk = __shfl_sync(kFull, p, k);
        float wk = __shfl_sync(kFull, w, k);
        if (lane < 3) {
            acc_c += color[(int64_t)pk * 3 + lane] * wk;
            acc_d += dir[(int64_t)pk * 3 + lane] * wk;
        }
        if (lane == 3) acc_f += conf[pk] * wk;
    }
    for (int f0 = 0; f0 < F; f0 += 32) {
        float a = 0.f;
        int f = f0 + lane;
        for (int k = 0; k < K; ++k) {
            int pk = __shfl_sync(kFull, p, k);
            float wk = __shfl_sync(kFull, w, k);
            if (f < F) a += emb[(int64_t)pk * F + f] * wk;
        }
        if (f < F) o_emb[r * F + f] = a;
    }
    if (lane < 3) { o_loc[r * 3 + lane] = lane == 0 ? lx : (lane == 1 ? ly : lz); o_color[r * 3 + lane] = acc_c; o_dir[r * 3 + lane] = acc_d; }
    if (lane == 3) o_conf[r] = acc_f;
    if (lane == 0) { o_opacity[r] = best; o_slot[r] = bi; o_far[r] = dmin; }
}

// ------------------------------------------------------------------------------------------------------------
// prune: keep flags -> exclusive scan (common.cu) -> one pass that moves every kept row of the five tables
// ------------------------------------------------------------------------------------------------------------
__global__ void prune_flag_kernel(const float* __restrict__ conf, int64_t n, float thresh, int* __restrict__ flag) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flag[i] = conf[i] >= thresh ? 1 : 0;
}

// one warp per source row; dst[i] = exclusive scan of the keep flags, dst[n] = kept count
__global__ void __launch_bounds__(256) points_compact_kernel(
    const int* __restrict__ dst, int64_t n, int F, const float* __restrict__ xyz, const float* __restrict__ emb,
    const float* __restrict__ conf, const float* __restrict__ dir, const float* __restrict__ color,
    float* __restrict__ o_xyz, float* __restrict__ o_emb, float* __restrict__ o_conf, float* __restrict__ o_dir,
    float* __restrict__ o_color) {
    const int lane = threadIdx.x & 31;
    const int64_t i = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= n) return;
    const int d0 = dst[i], d1 = dst[i + 1];
    if (d1 == d0) return;
    const int64_t j = d0;
    if ((F & 3) == 0) {
        const float4* s4 = reinterpret_cast<const float4*>(emb + i * F);
        float4* t4 = reinterpret_cast<float4*>(o_emb + j * F);
        for (int f = lane; f < (F >> 2); f += 32) t4[f] = s4[f];
    } else {
        for (int f = lane; f < F; f += 32) o_emb[j * F + f] = emb[i * F + f];
    }
    // lanes 16..18 xyz, 19..21 dir, 22..24 colour, 25 conf (at F = 32 the embedding row is lanes 0..7: disjoint lanes)
    if (lane >= 16 && lane < 19) o_xyz[j * 3 + lane - 16] = xyz[i * 3 + lane - 16];
    if (dir && lane >= 19 && lane < 22) o_dir[j * 3 + lane - 19] = dir[i * 3 + lane - 19];
    if (color && lane >= 22 && lane < 25) o_color[j * 3 + lane - 22] = color[i * 3 + lane - 22];
    if (conf && lane == 25) o_conf[j] = conf[i];
}

}  // namespace pn

using namespace pn;

extern "C" int pn_probe_rays(const float* opacity, const int32_t* ray_any, const float* sample_loc_w, const int32_t* sample_pidx,
                             const float* weight, const float* xyz, const float* embedding, const float* color, const float* dir,
                             const float* conf, int64_t n_rays, int32_t SR, int32_t K, int32_t F, float conf_lo, float conf_hi,
                             float* max_opacity, int32_t* max_slot, float* max_loc_w, float* far_dist, float* avg_color,
                             float* avg_dir, float* avg_conf, float* avg_embedding, void* stream) {
    PN_CHECK_ARG(n_rays >= 0 && SR >= 1 && K >= 1 && K <= 32 && F >= 1, "pn_probe_rays: bad extents (K <= 32)");
    if (n_rays == 0) return PN_OK;
    PN_CHECK_ARG(opacity && ray_any && sample_loc_w && sample_pidx && weight && xyz && embedding && color && dir && conf &&
                 max_opacity && max_slot && max_loc_w && far_dist && avg_color && avg_dir && avg_conf && avg_embedding,
                 "pn_probe_rays: null argument");
    probe_rays_kernel<<<(unsigned)cdiv(n_rays, 8), 256, 0, (cudaStream_t)stream>>>(
        opacity, ray_any, sample_loc_w, sample_pidx, weight, xyz, embedding, color, dir, conf, n_rays, SR, K, F, conf_lo, conf_hi,
        max_opacity, max_slot, max_loc_w, far_dist, avg_color, avg_dir, avg_conf, avg_embedding);
    PN_LAUNCH_CHECK("probe_rays");
    return PN_OK;
}

extern "C" int64_t pn_prune_workspace_bytes(int64_t n_points) {
    Carver c(nullptr, 0);
    c.take<int>(n_points + 1);
    c.take<int>(scan_scratch_elems_i32(n_points + 1));
    return c.off;
}

extern "C" int pn_prune_mark(const float* conf, int64_t n_points, float thresh, int32_t* dst_index, void* workspace,
                             int64_t workspace_bytes, void* stream) {
    PN_CHECK_ARG(n_points >= 0 && n_points < (int64_t)INT32_MAX, "pn_prune_mark: bad point count");
    PN_CHECK_ARG(dst_index && (conf || n_points == 0), "pn_prune_mark: null argument");
    PN_CHECK_ARG(workspace_bytes >= pn_prune_workspace_bytes(n_points), "pn_prune_mark: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    Carver c(workspace, workspace_bytes);
    int* flag = c.take<int>(n_points + 1);
    int* scratch = c.take<int>(scan_scratch_elems_i32(n_points + 1));
    if (n_points == 0) {
        PN_CUDA(cudaMemsetAsync(dst_index, 0, sizeof(int32_t), st));
        return PN_OK;
    }
    prune_flag_kernel<<<(unsigned)cdiv(n_points, 256), 256, 0, st>>>(conf, n_points, thresh, flag);
    PN_LAUNCH_CHECK("prune_flag");
    return scan_exclusive_i32(flag, dst_index, n_points, true, scratch, st);
}

extern "C" int pn_points_compact(const int32_t* dst_index, int64_t n_points, int32_t F, const float* xyz, const float* embedding,
                                 const float* conf, const float* dir, const float* color, float* out_xyz, float* out_embedding,
                                 float* out_conf, float* out_dir, float* out_color, void* stream) {
    PN_CHECK_ARG(n_points >= 0 && F >= 1, "pn_points_compact: bad extents");
    if (n_points == 0) return PN_OK;
    PN_CHECK_ARG(dst_index && xyz && embedding && out_xyz && out_embedding, "pn_points_compact: null argument");
    PN_CHECK_ARG((!conf || out_conf) && (!dir || out_dir) && (!color || out_color), "pn_points_compact: missing output table");
    points_compact_kernel<<<(unsigned)cdiv(n_points, 8), 256, 0, (cudaStream_t)stream>>>(
        dst_index, n_points, F, xyz, embedding, conf, dir, color, out_xyz, out_embedding, out_conf, out_dir, out_color);
    PN_LAUNCH_CHECK("points_compact");
    return PN_OK;
}
