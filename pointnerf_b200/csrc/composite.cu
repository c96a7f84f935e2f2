// Alpha compositing along rays, forward and backward, and the ray_dist block that feeds it.
//
// Replaces ray_march / alpha_ray_march (models/rendering/diff_ray_marching.py:509-573, with
// radiance_render / alpha_blend of diff_render_func.py:36-37,48-50) - about ten torch kernels and the
// generic cumprod backward - and the cummax/diff/clamp chain of
// models/neural_points_volumetric_model.py:293-301.
//
// One warp per ray.  Per-sample scalars live one per lane (32 samples per step, running product
// carried between steps); the transmittance is a warp prefix product by __shfl_up.  Feature rows
// ([C+1] floats, C = 128 in this fork) are streamed with lanes across channels (coalesced 128-byte
// segments) only for samples that can contribute: invalid samples have weight exactly 0, and once the
// transmittance falls below kTermT every later weight is below 1e-8 of the accumulated colour (early
// ray termination).  HBM-bound; algorithmic bytes forward 4*R*SR*(C+3) + 4*R*(C+1) + 12*R*SR.
#include "common.cuh"

namespace pn {

constexpr float kTermT = 1e-8f;

__device__ __forceinline__ float warp_prod_scan_incl(float v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        float t = __shfl_up_sync(kFull, v, o);
        if (lane >= o) v *= t;
    }
    return v;
}

template <int NC>  // channel registers per lane: C <= 32*NC
__global__ void __launch_bounds__(256) composite_fwd_kernel(const float* __restrict__ feat, const float* __restrict__ dist,
                                                            const uint8_t* __restrict__ valid, const int* __restrict__ slot_row,
                                                            const float* __restrict__ bg,
                                                            int64_t n_rays, int SR, int C, float* __restrict__ ray_color,
                                                            float* __restrict__ opacity, float* __restrict__ acc_T,
                                                            float* __restrict__ blend_w, float* __restrict__ bg_T) {
    int64_t ray = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (ray >= n_rays) return;
    const int C1 = C + 1;
    float acc[NC];
#pragma unroll
    for (int i = 0; i < NC; ++i) acc[i] = 0.f;
    float carry = 1.f;  // product of (1 - o + 1e-10) over all previous samples
    bool live = ray_color != nullptr;
    for (int s0 = 0; s0 < SR; s0 += 32) {
        int s = s0 + lane;
        float o = 0.f, a = 1.f;
        int64_t frow_i = -1;   // feature row of this lane's sample: dense (ray*SR+s) or compact (slot_row), -1 = no row
        if (s < SR) {
            frow_i = slot_row ? (int64_t)slot_row[ray * SR + s] : ray * SR + s;
            float sigma = (valid[ray * SR + s] && frow_i >= 0) ? feat[frow_i * C1] : 0.f;
            o = __fsub_rn(1.f, expf(-(sigma * dist[ray * SR + s])));
            a = __fadd_rn(__fsub_rn(1.f, o), 1e-10f);
        }
        float incl = warp_prod_scan_incl(a, lane);
        float prev = __shfl_up_sync(kFull, incl, 1);
        float T = carry * (lane == 0 ? 1.f : prev);
        float w = o * T;
        if (s < SR) {
            opacity[ray * SR + s] = o;
            acc_T[ray * SR + s] = T;
            blend_w[ray * SR + s] = w;
        }
        carry *= __shfl_sync(kFull, incl, 31);
        if (live) {
            unsigned m = __ballot_sync(kFull, s < SR && w != 0.f && T >= kTermT);
            // four feature rows in flight per round (a ray is one dependent chain of row loads otherwise: ~0.4 us per
            // sample, and the longest ray sets the kernel's time); the sums keep their sample order
            while (m) {
                float wj[4], v[4][NC];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int j = m ? __ffs(m) - 1 : 0;
                    wj[u] = m ? __shfl_sync(kFull, w, j) : 0.f;
                    const int64_t row = __shfl_sync(kFull, frow_i, j);
                    const bool on = m != 0;
                    m &= m - 1;
                    const float* f = feat + row * C1 + 1;
#pragma unroll
                    for (int i = 0; i < NC; ++i) {
                        int c = lane + 32 * i;
                        v[u][i] = (on && c < C) ? f[c] : 0.f;
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (wj[u] == 0.f) continue;          // unused slot of the round (a live sample never has w == 0)
#pragma unroll
                    for (int i = 0; i < NC; ++i) acc[i] = fmaf(wj[u], v[u][i], acc[i]);
                }
            }
            if (carry < kTermT) live = false;
        }
    }
    if (lane == 0) bg_T[ray] = carry;
    if (ray_color) {
#pragma unroll
        for (int i = 0; i < NC; ++i) {
            int c = lane + 32 * i;
            if (c < C) ray_color[ray * C + c] = bg ? fmaf(bg[c], carry, acc[i]) : acc[i];
        }
    }
}

// Backward.  Pass 1 (forward order): per live sample dot = <g_color, f_s>, and g_f[s,1:] = w_s*g_color.
// Pass 2 (reverse order): suffix sums give d/d(a_j); density gradient written to g_f[s,0].
// row_slot[row] = slot (ray * SR + s) of every compact feature row: inverse of slot_row
__global__ void __launch_bounds__(256) slot_invert_kernel(const int* __restrict__ slot_row, int64_t n_slots, int64_t n_rows,
                                                          int* __restrict__ row_slot) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_slots) return;
    int r = slot_row[i];
    if (r >= 0 && r < n_rows) row_slot[r] = (int)i;
}
// Pass 1 over the COMPACT layout, one warp per feature row (no walk along the ray: a ray crossing a dense region owns
// up to SR rows and a per-ray walk pays one DRAM round trip per row, in sequence).
template <int NC>
__global__ void __launch_bounds__(256) composite_bwd_rows_kernel(const float* __restrict__ feat, const uint8_t* __restrict__ valid,
                                                                 const int* __restrict__ row_slot, const float* __restrict__ opacity,
                                                                 const float* __restrict__ acc_T, const float* __restrict__ g_color,
                                                                 int64_t n_rows, const int* __restrict__ n_rows_dev, int SR, int C,
                                                                 float* __restrict__ g_feat) {
    const int lane = threadIdx.x & 31;
    if (n_rows_dev) { const int64_t nd = (int64_t)__ldg(n_rows_dev); n_rows = nd < n_rows ? nd : n_rows; }   // n_rows: then the capacity
    const int C1 = C + 1;
  for (int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; row < n_rows; row += ((int64_t)gridDim.x * blockDim.x) >> 5) {
    const int64_t slot = row_slot[row];
    float* grow = g_feat + row * C1;
    if (slot < 0) {                                     // row referenced by no slot: zero gradient
        for (int c = lane; c < C1; c += 32) grow[c] = 0.f;
        continue;
    }
    const int64_t ray = slot / SR;
    const float T = acc_T[slot], o = opacity[slot];
    const bool live = g_color && valid[slot] != 0 && T >= kTermT;
    const float w = o * T;
    const float* frow = feat + row * C1;
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < NC; ++i) {
        int c = lane + 32 * i;
        if (c < C) {
            float gc = g_color ? g_color[ray * C + c] : 0.f;
            if (live) dot = fmaf(gc, frow[1 + c], dot);
            grow[1 + c] = w * gc;
        }
    }
    if (live) dot = warp_sum(dot);
    if (lane == 0) grow[0] = live ? dot : 0.f;          // parked for pass 2, as in the per-ray kernel
  }
}

template <int NC, bool PASS1>
__global__ void __launch_bounds__(256) composite_bwd_kernel(
    const float* __restrict__ feat, const float* __restrict__ dist, const uint8_t* __restrict__ valid,
    const int* __restrict__ slot_row, const float* __restrict__ bg, const float* __restrict__ opacity, const float* __restrict__ acc_T,
    const float* __restrict__ bg_T, const float* __restrict__ g_color, const float* __restrict__ g_opacity,
    const float* __restrict__ g_acc_T, const float* __restrict__ g_blend_w, const float* __restrict__ g_bg_T, int64_t n_rays,
    int SR, int C, float* g_feat) {
    int64_t ray = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (ray >= n_rays) return;
    const int C1 = C + 1;
    float gc[NC];
    float gbg = 0.f;
#pragma unroll
    for (int i = 0; i < NC; ++i) {
        int c = lane + 32 * i;
        gc[i] = (g_color && c < C) ? g_color[ray * C + c] : 0.f;
        if (bg && c < C) gbg = fmaf(gc[i], bg[c], gbg);
    }
    gbg = warp_sum(gbg);
    float P_last = bg_T[ray];
    float G_P = (g_bg_T ? g_bg_T[ray] : 0.f) + gbg;

    // pass 1: lanes fetch the per-sample scalars of 32 samples at once, then the warp walks the samples that own a row
    for (int s0 = 0; PASS1 && s0 < SR; s0 += 32) {
        const int sl = s0 + lane;
        int64_t row_l = -1;
        float T_l = 0.f, o_l = 0.f;
        bool v_l = false;
        if (sl < SR) {
            const int64_t i = ray * SR + sl;
            row_l = slot_row ? (int64_t)slot_row[i] : i;
            T_l = acc_T[i];
            o_l = opacity[i];
            v_l = valid[i] != 0;
        }
        unsigned m = __ballot_sync(kFull, row_l >= 0);     // compact layout: samples without a row have no gradient row
        while (m) {
            const int j = __ffs(m) - 1;
            m &= m - 1;
            const int64_t row = __shfl_sync(kFull, row_l, j);
            const float T = __shfl_sync(kFull, T_l, j), o = __shfl_sync(kFull, o_l, j);
            const bool vj = __shfl_sync(kFull, (int)v_l, j) != 0;
            const float* frow = feat + row * C1;
            float* grow = g_feat + row * C1;
            const float w = o * T;
            const bool live = g_color && vj && T >= kTermT;  // warp-uniform
            float dot = 0.f;
#pragma unroll
            for (int i = 0; i < NC; ++i) {
                int c = lane + 32 * i;
                if (c < C) {
                    if (live) dot = fmaf(gc[i], frow[1 + c], dot);
                    grow[1 + c] = w * gc[i];
                }
            }
            // <g_color, f_s> is parked in the density column of this sample's gradient row; pass 2 (same warp)
            // consumes it and overwrites the column with the density gradient
            if (live) dot = warp_sum(dot);
            if (lane == 0) grow[0] = live ? dot : 0.f;
        }
    }
    __syncwarp();
    // pass 2: chunks of 32 samples from the end
    float carry = G_P * P_last;  // sum_{s > j} G_T[s]*T_s + G_P*P_last, for j = last sample of the chunk
    int nchunk = (SR + 31) / 32;
    for (int ch = nchunk - 1; ch >= 0; --ch) {
        int s = ch * 32 + lane;
        float term = 0.f, o = 0.f, T = 0.f, G_w = 0.f, d = 0.f, sigma = 0.f;
        bool ok = s < SR;
        bool v = false;
        int64_t row = -1;
        if (ok) {
            int64_t i = ray * SR + s;
            row = slot_row ? (int64_t)slot_row[i] : i;
            o = opacity[i];
            T = acc_T[i];
            d = dist[i];
            v = valid[i] != 0 && row >= 0;
            sigma = v ? feat[row * C1] : 0.f;
            G_w = (g_blend_w ? g_blend_w[i] : 0.f) + (row >= 0 ? g_feat[row * C1] : 0.f);
            float G_T = (g_acc_T ? g_acc_T[i] : 0.f) + G_w * o;
            term = G_T * T;
        }
        // exclusive suffix sum of term inside the chunk
        float incl = term;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            float t = __shfl_down_sync(kFull, incl, off);
            if (lane + off < 32) incl += t;
        }
        float excl = __shfl_down_sync(kFull, incl, 1);
        if (lane == 31) excl = 0.f;
        float suffix = carry + excl;
        carry += __shfl_sync(kFull, incl, 0);
        if (ok) {
            int64_t i = ray * SR + s;
            float a = __fadd_rn(__fsub_rn(1.f, o), 1e-10f);
            float G_o = (g_opacity ? g_opacity[i] : 0.f) + G_w * T - suffix / a;
            float e = expf(-(sigma * d));
            if (row >= 0) g_feat[row * C1] = v ? G_o * e * d : 0.f;
        }
    }
}

// g_bg[c] = sum_r g_color[r,c] * bg_T[r]; fixed-order two-stage reduction (no atomics)
constexpr int kBgRaysPerBlock = 16;    // short per-thread chains, many blocks: 128 rays per block left 32 blocks walking 128 dependent FMAs (19 us)
__global__ void __launch_bounds__(256) bg_grad_partial_kernel(const float* __restrict__ g_color, const float* __restrict__ bg_T,
                                                              int64_t n_rays, int C, float* __restrict__ partial) {
    int64_t r0 = (int64_t)blockIdx.x * kBgRaysPerBlock;
    int64_t r1 = r0 + kBgRaysPerBlock < n_rays ? r0 + kBgRaysPerBlock : n_rays;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float acc = 0.f;
        for (int64_t r = r0; r < r1; ++r) acc = fmaf(g_color[r * C + c], bg_T[r], acc);
        partial[(int64_t)blockIdx.x * C + c] = acc;
    }
}
// one warp per channel: lane l adds blocks l, l + 32, ... in order, then a fixed shuffle tree (deterministic)
__global__ void __launch_bounds__(256) bg_grad_final_kernel(const float* __restrict__ partial, int nblk, int C,
                                                            float* __restrict__ g_bg) {
    const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (c >= C) return;
    float acc = 0.f;
    for (int b = lane; b < nblk; b += 32) acc += partial[(int64_t)b * C + c];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) g_bg[c] = acc;
}

// Colour loss of the rendering model (base_rendering_model.py:543-550, torch.nn.MSELoss over the rays of the batch):
// loss = inv_norm * sum_r w_r * sum_c (pred-target)^2, grad = 2*inv_norm*w_r*(pred-target).  Fixed block count and a fixed
// summation tree: the value is a pure function of the inputs (run-to-run and rank-count independent).
constexpr int kLossBlocks = 128;
__global__ void __launch_bounds__(256) color_mse_partial_kernel(const float* __restrict__ pred, const float* __restrict__ target,
                                                                const float* __restrict__ ray_weight, const int* __restrict__ ray_mask,
                                                                int64_t n, int C, float inv_norm, const float* __restrict__ inv_norm_dev,
                                                                float* __restrict__ partial, float* __restrict__ grad) {
    __shared__ float red[8];
    if (inv_norm_dev) inv_norm *= __ldg(inv_norm_dev);
    const float g2 = 2.f * inv_norm;
    const bool weighted = ray_weight || ray_mask;
    float acc = 0.f;
    // ray index of element i, carried along instead of a 64-bit division per element (same elements in the same order)
    const int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (int64_t)gridDim.x * blockDim.x;
    int64_t r = i0 / C;
    int c = (int)(i0 - r * C);
    const int64_t dr = stride / C;
    const int dc = (int)(stride - dr * C);
    for (int64_t i = i0; i < n; i += stride) {
        float d = pred[i] - target[i];
        float w = ray_weight ? ray_weight[r] : 1.f;
        if (ray_mask) w = ray_mask[r] > 0 ? w : 0.f;
        acc = __fmaf_rn(w * d, d, acc);
        if (grad) grad[i] = weighted ? (g2 * w) * d : g2 * d;
        r += dr; c += dc;
        if (c >= C) { c -= C; ++r; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += red[w];
        partial[blockIdx.x] = t;
    }
}
__global__ void __launch_bounds__(kLossBlocks) color_mse_final_kernel(const float* __restrict__ partial, float inv_norm,
                                                                       const float* __restrict__ inv_norm_dev, float* __restrict__ loss) {
    __shared__ float red[kLossBlocks / 32];
    if (inv_norm_dev) inv_norm *= __ldg(inv_norm_dev);
    float acc = partial[threadIdx.x];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < kLossBlocks / 32; ++w) t += red[w];
        *loss = t * inv_norm;
    }
}

// ray_dist (neural_points_volumetric_model.py:293-301): running max of camera depth along the ray,
// forward difference, last = vsize_z, tiny / too large steps -> vsize_z, masked by sample validity.
__global__ void __launch_bounds__(256) ray_dist_kernel(const float* __restrict__ sample_loc, const uint8_t* __restrict__ valid,
                                                       int64_t n_rays, int SR, float vz, int unit_mode,
                                                       float* __restrict__ out) {
    extern __shared__ float sm[];
    int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int64_t ray = (int64_t)blockIdx.x * (blockDim.x >> 5) + wib;
    if (ray >= n_rays) return;
    float* cm = sm + (size_t)wib * SR;
    float carry = -INFINITY;
    for (int s0 = 0; s0 < SR; s0 += 32) {
        int s = s0 + lane;
        float z = s < SR ? sample_loc[(ray * SR + s) * 3 + 2] : -INFINITY;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            float t = __shfl_up_sync(kFull, z, o);
            if (lane >= o) z = fmaxf(z, t);
        }
        z = fmaxf(z, carry);
        if (s < SR) cm[s] = z;
        carry = __shfl_sync(kFull, z, 31);
    }
    __syncwarp();
    float two_vz = 2.f * vz;
    for (int s = lane; s < SR; s += 32) {
        float d = (s + 1 < SR) ? __fsub_rn(cm[s + 1], cm[s]) : vz;
        bool bad = d < 1e-8f;
        if (unit_mode > 0) bad = bad || d > two_vz;
        float m = bad ? 1.f : 0.f;
        d = __fadd_rn(__fmul_rn(d, __fsub_rn(1.f, m)), __fmul_rn(m, vz));
        out[ray * SR + s] = valid[ray * SR + s] ? d : __fmul_rn(d, 0.f);
    }
}

}  // namespace pn

using namespace pn;

#define PN_DISPATCH_NC(C, ...)                       \
    do {                                             \
        if ((C) <= 32) { constexpr int NC = 1; __VA_ARGS__; }       \
        else if ((C) <= 64) { constexpr int NC = 2; __VA_ARGS__; }  \
        else if ((C) <= 128) { constexpr int NC = 4; __VA_ARGS__; } \
        else { constexpr int NC = 8; __VA_ARGS__; }                 \
    } while (0)

extern "C" int pn_ray_dist(const float* sample_loc, const uint8_t* sample_valid, int64_t n_rays, int32_t SR, float vsize_z,
                           int32_t raydist_mode_unit, float* ray_dist, void* stream) {
    PN_CHECK_ARG(sample_loc && sample_valid && ray_dist, "pn_ray_dist: null argument");
    PN_CHECK_ARG(SR > 0 && SR <= 4096, "pn_ray_dist: SR=%d unsupported", SR);
    if (n_rays <= 0) return PN_OK;
    int warps = 8;
    size_t smem = (size_t)warps * SR * sizeof(float);
    PN_CHECK_ARG(smem <= 48 * 1024 * 2, "pn_ray_dist: SR too large");
    if (smem > 48 * 1024) cudaFuncSetAttribute(ray_dist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    ray_dist_kernel<<<(unsigned)cdiv(n_rays, warps), warps * 32, smem, (cudaStream_t)stream>>>(
        sample_loc, sample_valid, n_rays, SR, vsize_z, raydist_mode_unit, ray_dist);
    PN_LAUNCH_CHECK("ray_dist");
    return PN_OK;
}

extern "C" int pn_composite_forward(const float* ray_features, const float* ray_dist, const uint8_t* ray_valid,
                                    const int32_t* slot_row, const float* bg_color, int64_t n_rays, int32_t SR, int32_t C, float* ray_color,
                                    float* opacity, float* acc_T, float* blend_w, float* bg_T, void* stream) {
    PN_CHECK_ARG(ray_features && ray_dist && ray_valid && opacity && acc_T && blend_w && bg_T, "pn_composite_forward: null argument");
    PN_CHECK_ARG(C >= 1 && C <= 256 && SR > 0, "pn_composite_forward: C=%d SR=%d unsupported", C, SR);
    if (n_rays <= 0) return PN_OK;
    unsigned blocks = (unsigned)cdiv(n_rays * 32, 256);
    PN_DISPATCH_NC(C, (composite_fwd_kernel<NC><<<blocks, 256, 0, (cudaStream_t)stream>>>(
                          ray_features, ray_dist, ray_valid, slot_row, bg_color, n_rays, SR, C, ray_color, opacity, acc_T, blend_w, bg_T)));
    PN_LAUNCH_CHECK("composite_fwd");
    return PN_OK;
}

static int64_t bg_partial_bytes(int64_t n_rays, int32_t C) {
    int64_t r = n_rays > 0 ? n_rays : 1;
    return align_up(cdiv(r, kBgRaysPerBlock) * (int64_t)C * 4, 256) + 256;
}
extern "C" int64_t pn_composite_workspace_bytes(int64_t n_rays, int32_t C, int64_t n_rows) {
    return bg_partial_bytes(n_rays, C) + align_up((n_rows > 0 ? n_rows : 0) * 4, 256);
}

extern "C" int pn_composite_backward(const float* ray_features, const float* ray_dist, const uint8_t* ray_valid,
                                     const int32_t* slot_row, const float* bg_color, const float* opacity, const float* acc_T, const float* bg_T,
                                     const float* g_ray_color, const float* g_opacity, const float* g_acc_T,
                                     const float* g_blend_w, const float* g_bg_T, int64_t n_rays, int32_t SR, int32_t C,
                                     int64_t n_rows, const int32_t* n_rows_dev, float* g_features, float* g_bg_color, void* workspace,
                                     int64_t workspace_bytes, void* stream) {
    PN_CHECK_ARG(ray_features && ray_dist && ray_valid && opacity && acc_T && bg_T && g_features, "pn_composite_backward: null argument");
    PN_CHECK_ARG(C >= 1 && C <= 256 && SR > 0, "pn_composite_backward: C=%d SR=%d unsupported", C, SR);
    PN_CHECK_ARG(workspace_bytes >= pn_composite_workspace_bytes(n_rays, C, slot_row ? n_rows : 0), "pn_composite_backward: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    if (n_rays <= 0) {
        if (g_bg_color) PN_CUDA(cudaMemsetAsync(g_bg_color, 0, C * sizeof(float), st));
        return PN_OK;
    }
    unsigned blocks = (unsigned)cdiv(n_rays * 32, 256);
    if (slot_row && n_rows > 0) {   // compact layout: pass 1 runs one warp per feature row, the per-ray kernel keeps pass 2
        int* row_slot = (int*)((char*)workspace + bg_partial_bytes(n_rays, C));
        PN_CUDA(cudaMemsetAsync(row_slot, 0xFF, n_rows * sizeof(int), st));
        slot_invert_kernel<<<(unsigned)cdiv(n_rays * SR, 256), 256, 0, st>>>(slot_row, n_rays * SR, n_rows, row_slot);
        PN_LAUNCH_CHECK("slot_invert");
        const int64_t want = cdiv(n_rows * 32, 256), most = (int64_t)kNumSMs * 8;
        PN_DISPATCH_NC(C, (composite_bwd_rows_kernel<NC><<<(unsigned)(want < most ? want : most), 256, 0, st>>>(
                              ray_features, ray_valid, row_slot, opacity, acc_T, g_ray_color, n_rows, n_rows_dev, SR, C, g_features)));
        PN_LAUNCH_CHECK("composite_bwd_rows");
        PN_DISPATCH_NC(C, (composite_bwd_kernel<NC, false><<<blocks, 256, 0, st>>>(ray_features, ray_dist, ray_valid, slot_row, bg_color, opacity,
                                                                                 acc_T, bg_T, g_ray_color, g_opacity, g_acc_T, g_blend_w,
                                                                                 g_bg_T, n_rays, SR, C, g_features)));
    } else {
        PN_DISPATCH_NC(C, (composite_bwd_kernel<NC, true><<<blocks, 256, 0, st>>>(ray_features, ray_dist, ray_valid, slot_row, bg_color, opacity,
                                                                                acc_T, bg_T, g_ray_color, g_opacity, g_acc_T, g_blend_w,
                                                                                g_bg_T, n_rays, SR, C, g_features)));
    }
    PN_LAUNCH_CHECK("composite_bwd");
    if (g_bg_color) {
        if (g_ray_color && bg_color) {
            int nblk = (int)cdiv(n_rays, kBgRaysPerBlock);
            float* partial = (float*)workspace;
            bg_grad_partial_kernel<<<nblk, 128, 0, st>>>(g_ray_color, bg_T, n_rays, C, partial);
            PN_LAUNCH_CHECK("bg_grad_partial");
            bg_grad_final_kernel<<<(unsigned)cdiv((int64_t)C * 32, 256), 256, 0, st>>>(partial, nblk, C, g_bg_color);
            PN_LAUNCH_CHECK("bg_grad_final");
        } else {
            PN_CUDA(cudaMemsetAsync(g_bg_color, 0, C * sizeof(float), st));
        }
    }
    return PN_OK;
}

extern "C" int64_t pn_color_mse_workspace_bytes(void) { return (int64_t)kLossBlocks * sizeof(float); }

extern "C" int pn_color_mse(const float* pred, const float* target, const float* ray_weight, const int32_t* ray_mask, int64_t n_rays,
                            int32_t C, float inv_norm, const float* inv_norm_dev, float* loss, float* grad, void* workspace,
                            int64_t workspace_bytes, void* stream) {
    PN_CHECK_ARG(loss && workspace, "pn_color_mse: null argument");
    PN_CHECK_ARG(n_rays >= 0 && C >= 1, "pn_color_mse: n_rays=%lld C=%d", (long long)n_rays, C);
    PN_CHECK_ARG(n_rays == 0 || (pred && target), "pn_color_mse: null argument");
    PN_CHECK_ARG(workspace_bytes >= pn_color_mse_workspace_bytes(), "pn_color_mse: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    color_mse_partial_kernel<<<kLossBlocks, 256, 0, st>>>(pred, target, ray_weight, ray_mask, n_rays * C, C, inv_norm, inv_norm_dev, (float*)workspace, grad);
    PN_LAUNCH_CHECK("color_mse_partial");
    color_mse_final_kernel<<<1, kLossBlocks, 0, st>>>((const float*)workspace, inv_norm, inv_norm_dev, loss);
    PN_LAUNCH_CHECK("color_mse_final");
    return PN_OK;
}
