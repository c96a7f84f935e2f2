// pn_adam_step: one multi-tensor Adam step over every trainable tensor of the hot path.
// Reference: torch.optim.Adam(params, lr, betas=(0.9, 0.999)) as set up in models/mvs_points_volumetric_model.py:47-100
// (`optimizer` for the MLP, `neural_point_optimizer` for the point tables + bg_color; eps 1e-8, no weight decay, no
// amsgrad) and stepped in `backward` (:103-125).  The arithmetic follows torch's single-tensor Adam op by op in fp32:
//   m += (g - m) * (1 - beta1);  v = v * beta2 + (1 - beta2) * g * g;
//   denom = sqrt(v) / sqrt(1 - beta2^t) + eps;  p += (-lr / (1 - beta1^t)) * (m / denom)
// HBM-bound: 16 B read + 12 B written per element, one launch for up to 32 tensors (the reference launches ~10 elementwise
// kernels per parameter tensor, 21 tensors).
#include "common.cuh"

namespace pn {

constexpr int kAdamMaxSegs = 32;
constexpr int kAdamThreads = 256;
constexpr int kAdamPerBlock = kAdamThreads * 4 * 4;      // 4 float4 per thread

struct AdamBatch {
    float* p[kAdamMaxSegs];
    const float* g[kAdamMaxSegs];
    float* m[kAdamMaxSegs];
    float* v[kAdamMaxSegs];
    long long n[kAdamMaxSegs];
    int first_block[kAdamMaxSegs + 1];
    int n_segs;
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, float w1, float b2, float w2, float bc2_sqrt,
                                         float eps, float neg_step) {
    m = fmaf(w1, __fsub_rn(g, m), m);                                      // exp_avg.lerp_(grad, 1 - beta1): self + w * (end - self), contracted
    v = fmaf(__fmul_rn(w2, g), g, __fmul_rn(v, b2));                       // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
    float denom = __fadd_rn(__fdiv_rn(sqrtf(v), bc2_sqrt), eps);      // (exp_avg_sq.sqrt() / bias_correction2_sqrt).add_(eps)
    p = fmaf(neg_step, __fdiv_rn(m, denom), p);                            // param.addcdiv_(exp_avg, denom, value=-step_size)
}

__global__ void __launch_bounds__(kAdamThreads) adam_kernel(AdamBatch b, float w1, float b2, float w2, float bc2_sqrt, float eps,
                                                            float neg_step) {
    int s = 0;
    while (s + 1 < b.n_segs && (int)blockIdx.x >= b.first_block[s + 1]) ++s;
    const long long n = b.n[s];
    const long long base = (long long)(blockIdx.x - b.first_block[s]) * kAdamPerBlock;
    float* __restrict__ P = b.p[s];
    const float* __restrict__ G = b.g[s];
    float* __restrict__ M = b.m[s];
    float* __restrict__ V = b.v[s];
    const bool vec = ((((uintptr_t)P | (uintptr_t)G | (uintptr_t)M | (uintptr_t)V) & 15) == 0);
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        long long i = base + ((long long)r * kAdamThreads + threadIdx.x) * 4;
        if (i >= n) break;
        if (vec && i + 4 <= n) {
            float4 p = *reinterpret_cast<float4*>(P + i), m = *reinterpret_cast<float4*>(M + i), v = *reinterpret_cast<float4*>(V + i);
            const float4 g = *reinterpret_cast<const float4*>(G + i);
            adam_one(p.x, g.x, m.x, v.x, w1, b2, w2, bc2_sqrt, eps, neg_step);
            adam_one(p.y, g.y, m.y, v.y, w1, b2, w2, bc2_sqrt, eps, neg_step);
            adam_one(p.z, g.z, m.z, v.z, w1, b2, w2, bc2_sqrt, eps, neg_step);
            adam_one(p.w, g.w, m.w, v.w, w1, b2, w2, bc2_sqrt, eps, neg_step);
            *reinterpret_cast<float4*>(P + i) = p;
            *reinterpret_cast<float4*>(M + i) = m;
            *reinterpret_cast<float4*>(V + i) = v;
        } else {
            for (long long j = i; j < n && j < i + 4; ++j) {
                float p = P[j], m = M[j], v = V[j];
                adam_one(p, G[j], m, v, w1, b2, w2, bc2_sqrt, eps, neg_step);
                P[j] = p; M[j] = m; V[j] = v;
            }
        }
    }
}

}  // namespace pn

using namespace pn;

extern "C" int pn_adam_step(const pn_adam_segment* segs, int32_t n_segs, double lr, double beta1, double beta2, double eps,
                            int64_t step, void* stream) {
    PN_CHECK_ARG(n_segs >= 0 && (segs || n_segs == 0), "pn_adam_step: null segment list");
    PN_CHECK_ARG(step >= 1, "pn_adam_step: step counts from 1");
    // scalars exactly as torch's python side computes them (double), rounded to fp32 where the kernels take them
    // (the hyper-parameters arrive as doubles: 1 - (double)0.999f is off by 1.3e-5 relative)
    const double bc1 = 1.0 - pow(beta1, (double)step), bc2 = 1.0 - pow(beta2, (double)step);
    const float w1 = (float)(1.0 - beta1), w2 = (float)(1.0 - beta2);
    const float bc2_sqrt = (float)sqrt(bc2), neg_step = (float)(-(lr / bc1));
    for (int s0 = 0; s0 < n_segs; s0 += kAdamMaxSegs) {
        AdamBatch b;
        b.n_segs = 0;
        int blocks = 0;
        for (int s = s0; s < n_segs && s < s0 + kAdamMaxSegs; ++s) {
            if (segs[s].n <= 0) continue;
            PN_CHECK_ARG(segs[s].param && segs[s].grad && segs[s].exp_avg && segs[s].exp_avg_sq, "pn_adam_step: null tensor in segment %d", s);
            int k = b.n_segs++;
            b.p[k] = segs[s].param; b.g[k] = segs[s].grad; b.m[k] = segs[s].exp_avg; b.v[k] = segs[s].exp_avg_sq; b.n[k] = segs[s].n;
            b.first_block[k] = blocks;
            blocks += (int)cdiv(segs[s].n, kAdamPerBlock);
        }
        if (b.n_segs == 0) continue;
        b.first_block[b.n_segs] = blocks;
        adam_kernel<<<blocks, kAdamThreads, 0, (cudaStream_t)stream>>>(b, w1, (float)beta2, w2, bc2_sqrt, (float)eps, neg_step);
        PN_LAUNCH_CHECK("adam");
    }
    return PN_OK;
}

// ---- lazy (touched-row) Adam for the per-point tables ---------------------------------------------------------------------
// A training batch touches ~0.5 % of a 20 M-point cloud, but exact dense Adam moves EVERY row on its momentum each step:
// 28 B x 39 floats x N = 21.8 GB per step at 20 M points (3.4 ms at the HBM roof, measured 3.5 ms) against ~5 ms for the
// whole forward + backward.  pn_adam_rows updates only the rows that received gradient this step, with the arithmetic of
// torch.optim.SparseAdam (torch/optim/_functional.py sparse_adam: moments of untouched rows are left alone, the bias
// corrections use the global step) - an explicit OPTION with different semantics from the reference's dense Adam
// (mvs_points_volumetric_model.py:47-100), not a drop-in for it.
namespace pn {
struct AdamRowsBatch {
    float* p[4];
    const float* g[4];
    float* m[4];
    float* v[4];
    int width[4];
    int n_tabs;
};
__global__ void __launch_bounds__(256) adam_rows_kernel(const int* __restrict__ ids, int64_t n, int64_t stride, AdamRowsBatch b, float w1,
                                                        float w2, float eps, float neg_step) {
    const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (i >= n) return;
    const int row = ids[i * stride];
    if (row < 0) return;
    for (int t = 0; t < b.n_tabs; ++t) {
        const int wd = b.width[t];
        const int64_t base = (int64_t)row * wd;
        for (int c = lane; c < wd; c += 32) {
            const float g = b.g[t][base + c], m0 = b.m[t][base + c], v0 = b.v[t][base + c];
            const float mu = __fmul_rn(__fsub_rn(g, m0), w1);                    // grad.sub(old_exp_avg).mul_(1 - beta1)
            const float vu = __fmul_rn(__fsub_rn(__fmul_rn(g, g), v0), w2);      // grad.pow(2).sub_(old_exp_avg_sq).mul_(1 - beta2)
            const float m1 = __fadd_rn(mu, m0), v1 = __fadd_rn(vu, v0);         // update.add_(old): the new moments, as stored
            b.m[t][base + c] = m1;
            b.v[t][base + c] = v1;
            const float denom = __fadd_rn(sqrtf(v1), eps);
            b.p[t][base + c] = __fadd_rn(b.p[t][base + c], __fmul_rn(neg_step, __fdiv_rn(m1, denom)));
        }
    }
}
}  // namespace pn

extern "C" int pn_adam_rows(const int32_t* ids, int64_t n, int64_t stride, const pn_adam_rows_table* tabs, int32_t n_tabs, double lr,
                            double beta1, double beta2, double eps, int64_t step, void* stream) {
    PN_CHECK_ARG((ids || n == 0) && n >= 0 && stride >= 1 && tabs && n_tabs >= 1 && n_tabs <= 4, "pn_adam_rows: bad argument (1..4 tables)");
    PN_CHECK_ARG(step >= 1, "pn_adam_rows: step counts from 1");
    if (n == 0) return PN_OK;
    AdamRowsBatch b;
    b.n_tabs = n_tabs;
    for (int t = 0; t < n_tabs; ++t) {
        PN_CHECK_ARG(tabs[t].param && tabs[t].grad && tabs[t].exp_avg && tabs[t].exp_avg_sq && tabs[t].width >= 1, "pn_adam_rows: null tensor in table %d", t);
        b.p[t] = tabs[t].param; b.g[t] = tabs[t].grad; b.m[t] = tabs[t].exp_avg; b.v[t] = tabs[t].exp_avg_sq; b.width[t] = tabs[t].width;
    }
    const double bc1 = 1.0 - pow(beta1, (double)step), bc2 = 1.0 - pow(beta2, (double)step);
    const float neg_step = (float)(-(lr * sqrt(bc2) / bc1));
    adam_rows_kernel<<<(unsigned)cdiv(n * 32, 256), 256, 0, (cudaStream_t)stream>>>(ids, n, stride, b, (float)(1.0 - beta1), (float)(1.0 - beta2),
                                                                                  (float)eps, neg_step);
    PN_LAUNCH_CHECK("adam_rows");
    return PN_OK;
}
