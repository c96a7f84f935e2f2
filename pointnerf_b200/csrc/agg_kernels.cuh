// Device kernels of the point aggregator around its MLP layers.
//
// Reference semantics: PointAggregator.forward / linear / viewmlp (agg_intrp_order == 2),
// models/aggregators/point_aggregators.py:729-816, 421-429, 488-646, positional_encoding
// models/helpers/networks.py:175-189, and the gather of NeuralPoints.forward neural_points.py:788-799.
// Rows are COMPACT: one row per valid (sample, neighbour) pair m, one per valid sample s
// (work lists from pn_compact_samples), so no masked row is ever computed or stored.
#pragma once
#include "common.cuh"

namespace pn {

struct AggDims {
    int F, ff, df, vf, W, C, K, SR, dirs_per_ray, compact_out;
    int order;  // agg_intrp_order: 2 = a density per neighbour, interpolated; 1 = interpolate features, one density per sample, sigmoid colour
    int in1, in1p, in3, in3p, inc, incp;  // logical / padded (to 16) input widths of block1.0, block3.0, color.0
    float slope, dist_scale;
};

// Valid samples / pairs of the batch.  With in.counts (device int32[2], written by pn_compact_samples) the host never learns
// them: in.S / in.M are then the CAPACITIES the buffers and launches were sized for, and every kernel walks its items with
// a grid-stride loop bounded by the device value.
__device__ __forceinline__ int64_t n_samples(const pn_agg_inputs& in) {
    if (!in.counts) return in.S;
    const int64_t v = (int64_t)__ldg(in.counts);
    return v < in.S ? v : in.S;
}
__device__ __forceinline__ int64_t n_pairs(const pn_agg_inputs& in) {
    if (!in.counts) return in.M;
    const int64_t v = (int64_t)__ldg(in.counts + 1);
    return v < in.M ? v : in.M;
}
__device__ __forceinline__ int64_t warp_id_global() { return ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; }
__device__ __forceinline__ int64_t warps_total() { return ((int64_t)gridDim.x * blockDim.x) >> 5; }
__device__ __forceinline__ int64_t thread_id_global() { return (int64_t)blockIdx.x * blockDim.x + threadIdx.x; }
__device__ __forceinline__ int64_t threads_total() { return (int64_t)gridDim.x * blockDim.x; }

__device__ __forceinline__ void rot3(const float* __restrict__ R, float x, float y, float z, float& ox, float& oy, float& oz) {
    // row-vector times R^T  ==  R * v
    ox = R[0] * x + R[1] * y + R[2] * z;
    oy = R[3] * x + R[4] * y + R[5] * z;
    oz = R[6] * x + R[7] * y + R[8] * z;
}
__device__ __forceinline__ float softplus1(float x) { return x > 20.f ? x : log1pf(expf(x)); }  // torch Softplus(beta=1, threshold=20)
__device__ __forceinline__ float sigmoid1(float x) { return x > 20.f ? 1.f : 1.f / (1.f + expf(-x)); }
__device__ __forceinline__ float sigmoid_exact(float x) { return 1.f / (1.f + expf(-x)); }   // torch.sigmoid

__device__ __forceinline__ void load_view(const pn_agg_inputs& in, const AggDims& d, int slot, float& vx, float& vy, float& vz) {
    int64_t row = d.dirs_per_ray ? (slot / d.SR) : slot;
    rot3(in.Rw2c, in.ray_dirs[3 * row], in.ray_dirs[3 * row + 1], in.ray_dirs[3 * row + 2], vx, vy, vz);
}

// ---------------------------------------------------------------------------------------------
// prep: one warp per valid sample.
//   phase 1 (lane = pair): distance vector, inverse-distance weight, normalisation over the sample's
//            pairs, clamped confidence, the 7 extra inputs of block3 (colour, dir - view, <dir, view>)
//   phase 2 (lane = channel): block1 input row  [emb | PE_ff(emb) | PE_df(dist)]  written coalesced
// pairv[m] = (w*conf, w, -, -)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) agg_prep_kernel(pn_agg_inputs in, AggDims d, float* __restrict__ X1,
                                                       float* __restrict__ X3, float4* __restrict__ pairv,
                                                       float* __restrict__ weight_out) {
    const int lane = threadIdx.x & 31;
    const int64_t S = n_samples(in);
    const int64_t si = warp_id_global();                 // the launch covers the capacity: one sample per warp, no loop (registers)
    if (si >= S) return;
  {
    int slot = in.sample_list[si];
    int p0 = in.pair_start[si], np = in.pair_start[si + 1] - p0;
    float sx = in.sample_loc_w[3 * (int64_t)slot], sy = in.sample_loc_w[3 * (int64_t)slot + 1], sz = in.sample_loc_w[3 * (int64_t)slot + 2];
    float px = in.sample_loc[3 * (int64_t)slot], py = in.sample_loc[3 * (int64_t)slot + 1], pz = in.sample_loc[3 * (int64_t)slot + 2];
    float vx, vy, vz;
    load_view(in, d, slot, vx, vy, vz);

    // ---- phase 1: lane j < np owns pair j (np <= K <= 32)
    float dd[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float wl = 0.f;
    int g = 0, src = 0;
    bool act = lane < np;
    if (act) {
        int m = p0 + lane;
        src = in.pair_src[m];
        g = in.sample_pidx[src];
        const float* q = in.xyz + 3 * (int64_t)g;
        const float* qp = in.xyz_pers + 3 * (int64_t)g;
        float dx = __fsub_rn(q[0], sx), dy = __fsub_rn(q[1], sy), dz = __fsub_rn(q[2], sz);
        float nrm = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
        wl = 1.f / fmaxf(nrm, 1e-6f);
        // agg_dist_pers == 20 (:775-783): products rounded separately, as torch evaluates them
        float e0 = __fsub_rn(__fmul_rn(qp[0], qp[2]), __fmul_rn(px, pz));
        float e1 = __fsub_rn(__fmul_rn(qp[1], qp[2]), __fmul_rn(py, pz));
        float e2 = __fsub_rn(qp[2], pz);
        rot3(in.Rw2c, dx * d.dist_scale, dy * d.dist_scale, dz * d.dist_scale, dd[0], dd[1], dd[2]);
        dd[3] = e0 * d.dist_scale; dd[4] = e1 * d.dist_scale; dd[5] = e2 * d.dist_scale;
        // block3 extras (:556-570)
        float cdx, cdy, cdz;
        rot3(in.Rw2c, in.dir[3 * (int64_t)g], in.dir[3 * (int64_t)g + 1], in.dir[3 * (int64_t)g + 2], cdx, cdy, cdz);
        float* x3 = X3 + (int64_t)m * d.in3p + d.W;
        x3[0] = in.color[3 * (int64_t)g]; x3[1] = in.color[3 * (int64_t)g + 1]; x3[2] = in.color[3 * (int64_t)g + 2];
        x3[3] = cdx - vx; x3[4] = cdy - vy; x3[5] = cdz - vz;
        x3[6] = cdx * vx + cdy * vy + cdz * vz;
        for (int c = d.in3; c < d.in3p; ++c) X3[(int64_t)m * d.in3p + c] = 0.f;
    }
    float tot = warp_sum(wl);
    if (act) {
        float w = wl / fmaxf(tot, 1e-8f);
        float cf = fminf(fmaxf(in.conf[g], 1e-4f), 1.f);
        pairv[p0 + lane] = make_float4(w * cf, w, 0.f, 0.f);
        if (weight_out) weight_out[src] = w;
    }
    // the six scaled offsets of each pair are parked, raw, at the head of the row's distance-encoding block:
    // agg_prep_rows_kernel (one warp per PAIR) reads them and writes the whole block1 input row
    if (act) {
        float* park = X1 + (int64_t)(p0 + lane) * d.in1p + d.F + 2 * d.ff * d.F;
#pragma unroll
        for (int i = 0; i < 6; ++i) park[i] = dd[i];
    }
  }
}

// phase 2 of prep: block1 input row [emb | PE_ff(emb) | PE_df(dist) | 0-pad] of pair m; lane = channel.  The encoding
// interleaves (sin, cos) per (channel, frequency), so a lane owns short scattered runs of the row - stored directly,
// every instruction would touch every 32-byte sector of the row (6x the L2 write transactions).  The row is assembled
// in shared memory and leaves as whole float4s.
// The kernel is issue-bound (ncu: 76 % issue slots, ALU 45 %: index arithmetic of the scattered shared-memory stores and loop
// control on run-time F / ff / df), so the shipped shape (F = 32, 3 feature and 5 distance frequencies) is compiled with the
// three as constants - every loop unrolls and the store offsets fold; kF = 0 keeps the run-time form for other shapes.
template <int kF, int kFF, int kDF>
__global__ void __launch_bounds__(256, 8) agg_prep_rows_kernel(pn_agg_inputs in, AggDims d, float* __restrict__ X1) {
    extern __shared__ float prep_rows[];                 // blockDim.x / 32 rows of in1p floats
    const int lane = threadIdx.x & 31;
    const int64_t M = n_pairs(in);
    const int F = kF ? kF : d.F, ff = kF ? kFF : d.ff, df = kF ? kDF : d.df;
    const int in1 = kF ? kF + 2 * kFF * kF + 12 * kDF : d.in1, in1p = kF ? (in1 + 15) / 16 * 16 : d.in1p;
    float* srow = prep_rows + (threadIdx.x >> 5) * in1p;
    const int64_t m = warp_id_global();                  // the launch covers the capacity: one pair per warp, no loop (registers)
    if (m >= M) return;
  {
    const int gj = in.sample_pidx[in.pair_src[m]];
    float* row = X1 + m * in1p;
    const int pe0 = F + 2 * ff * F;
    float rv = lane < 6 ? row[pe0 + lane] : 0.f;         // lane = axis: the parked offset component
#pragma unroll
    for (int c = lane; c < F; c += 32) {
        float e = in.embedding[(int64_t)gj * F + c];
        srow[c] = e;
        // frequencies are powers of two: one accurate sincosf, then double-angle steps (each adds ~2 ulp of absolute
        // error on values in [-1, 1]; at most 15 steps, far inside the 1e-4 budget)
        float s, co;
        sincosf(e, &s, &co);
#pragma unroll
        for (int f = 0; f < ff; ++f) {
            srow[F + (c * ff + f) * 2] = s;
            srow[F + (c * ff + f) * 2 + 1] = co;
            const float s2 = 2.f * s * co, c2 = (co - s) * (co + s);
            s = s2; co = c2;
        }
    }
    if (lane < 6) {                                       // all df frequencies of one offset component
        float s, co;
        sincosf(rv, &s, &co);
#pragma unroll
        for (int f = 0; f < df; ++f) {
            const int q = lane * df + f;
            srow[pe0 + 2 * q] = s;
            srow[pe0 + 2 * q + 1] = co;
            const float s2 = 2.f * s * co, c2 = (co - s) * (co + s);
            s = s2; co = c2;
        }
    }
    for (int c = in1 + lane; c < in1p; c += 32) srow[c] = 0.f;
    __syncwarp();
    float4* row4 = reinterpret_cast<float4*>(row);       // in1p % 4 == 0 and X1 is 16-byte aligned
#pragma unroll
    for (int c4 = lane; c4 < (in1p >> 2); c4 += 32) row4[c4] = *reinterpret_cast<const float4*>(srow + 4 * c4);
  }
}

// ---------------------------------------------------------------------------------------------
// reduce: one warp per valid sample.  Density head (256 -> 1, softplus(x - 1)) per pair, then the
// K-weighted sums: alpha_s = sum_k alpha_k * w_k*conf_k, f_s = sum_k H4_k * w_k*conf_k, and the
// 24 view-direction encoding columns of the colour branch input.
// ---------------------------------------------------------------------------------------------
template <int NW>  // W == 32*NW
__global__ void __launch_bounds__(256) agg_reduce_kernel(pn_agg_inputs in, AggDims d, const float* __restrict__ H4,
                                                         float4* __restrict__ pairv, const float* __restrict__ wal,
                                                         const float* __restrict__ bal, float* __restrict__ FS,
                                                         float* __restrict__ alpha) {
    const int lane = threadIdx.x & 31;
    const int64_t S = n_samples(in);
    float wa[NW];
#pragma unroll
    for (int i = 0; i < NW; ++i) wa[i] = wal[lane + 32 * i];
    const float b = bal[0];
    const int64_t si = warp_id_global();                 // the launch covers the capacity: one sample per warp, no loop (registers)
    if (si >= S) return;
  {
    int slot = in.sample_list[si];
    int p0 = in.pair_start[si], np = in.pair_start[si + 1] - p0;
    float acc[NW];
#pragma unroll
    for (int i = 0; i < NW; ++i) acc[i] = 0.f;
    float a_s = 0.f;
    for (int j = 0; j < np; ++j) {
        int64_t m = p0 + j;
        float4 pv = pairv[m];
        float h[NW], dot = 0.f;
#pragma unroll
        for (int i = 0; i < NW; ++i) { h[i] = H4[m * d.W + lane + 32 * i]; dot = fmaf(h[i], wa[i], dot); }
        if (d.order != 1) {
            float araw = warp_sum(dot) + b;
            if (lane == 0) pairv[m] = make_float4(pv.x, pv.y, araw, 0.f);
            a_s = fmaf(softplus1(araw - 1.f), pv.x, a_s);
        }
#pragma unroll
        for (int i = 0; i < NW; ++i) acc[i] = fmaf(h[i], pv.x, acc[i]);
    }
    if (d.order == 1) {   // point_aggregators.py:573-588: the density head reads the INTERPOLATED feature
        float dot = 0.f;
#pragma unroll
        for (int i = 0; i < NW; ++i) dot = fmaf(acc[i], wa[i], dot);
        float araw = warp_sum(dot) + b;
        a_s = softplus1(araw - 1.f);
        if (lane == 0) {      // kept for the backward in the unused 4th lane of the sample's first pair
            float4 pv = pairv[p0];
            pairv[p0] = make_float4(pv.x, pv.y, pv.z, araw);
        }
    }
    float* row = FS + si * d.incp;
#pragma unroll
    for (int i = 0; i < NW; ++i) row[lane + 32 * i] = acc[i];
    if (lane == 0) alpha[si] = a_s;
    float vx, vy, vz;
    load_view(in, d, slot, vx, vy, vz);
    int nq = 3 * d.vf;  // positional_encoding(ori=True) minus its first 3 columns: [sin(a*vf+f) | cos(a*vf+f)]
    for (int q = lane; q < nq; q += 32) {
        int a = q / d.vf, f = q - a * d.vf;
        float v = a == 0 ? vx : a == 1 ? vy : vz;
        float s, co;
        sincosf(v * (float)(1 << f), &s, &co);
        row[d.W + q] = s;
        row[d.W + nq + q] = co;
    }
    for (int c = d.inc + lane; c < d.incp; c += 32) row[c] = 0.f;
  }
}

// decoded[slot] = [alpha_s | C3[s, :]]
__global__ void __launch_bounds__(256) agg_finish_kernel(pn_agg_inputs in, AggDims d, const float* __restrict__ alpha,
                                                         const float* __restrict__ C3, float* __restrict__ decoded) {
    // one warp per sample row, lanes over the C + 1 columns (a thread per element paid a 64-bit division each: 17 us for 10 MB)
    const int C1 = d.C + 1, lane = threadIdx.x & 31;
    const int64_t S = n_samples(in);
    for (int64_t si = warp_id_global(); si < S; si += warps_total()) {
        const int64_t orow = d.compact_out ? si : (int64_t)in.sample_list[si];
        for (int c = lane; c < C1; c += 32) {
            float v = c == 0 ? alpha[si] : C3[si * d.C + c - 1];
            if (c > 0 && d.order == 1) v = sigmoid_exact(v) * 1.002f - 0.001f;   // raw2out_color (:268-273), order 1 only in this fork
            decoded[orow * C1 + c] = v;
        }
    }
}

// ------------------------------------ backward ------------------------------------------------
// colour output: GC3[s,c] = g_decoded[slot,1+c] * lrelu'(C3[s,c]); galpha[s] = g_decoded[slot,0]
__global__ void __launch_bounds__(256) agg_gout_kernel(pn_agg_inputs in, AggDims d, const float* __restrict__ g_decoded,
                                                       const float* __restrict__ C3, float* __restrict__ GC3,
                                                       float* __restrict__ galpha) {
    const int C1 = d.C + 1, lane = threadIdx.x & 31;     // one warp per sample row (see agg_finish_kernel)
    const int64_t S = n_samples(in);
    for (int64_t si = warp_id_global(); si < S; si += warps_total()) {
        const int64_t irow = d.compact_out ? si : (int64_t)in.sample_list[si];
        for (int c = lane; c < C1; c += 32) {
            float gv = g_decoded[irow * C1 + c];
            if (c == 0) galpha[si] = gv;
            else {
                float c3 = C3[si * d.C + c - 1];
                if (d.order == 1) {
                    float o = sigmoid_exact(c3);
                    gv *= 1.002f * o * (1.f - o);
                }
                GC3[si * d.C + c - 1] = gv * (c3 > 0.f ? 1.f : d.slope);
            }
        }
    }
}

// reduce backward: CTA b owns the contiguous sample range [b*chunk, (b+1)*chunk); its warps stride over
// the range.  dPre4[m,:] = (wc*g_fs + g_araw*w_alpha) * lrelu'(H4), conf gradient per pair, and the
// density-head parameter gradients as per-CTA partials (summed in CTA order by agg_alpha_final_kernel).
template <int NW>
__global__ void __launch_bounds__(256, 4) agg_reduce_bwd_kernel(pn_agg_inputs in, AggDims d, const float* __restrict__ H4,
                                                             const float4* __restrict__ pairv, const float* __restrict__ wal,
                                                             const float* __restrict__ GFS, const float* __restrict__ galpha,
                                                             float* __restrict__ dPre4, int64_t ld4,
                                                             float* __restrict__ PG, int pgw, float* __restrict__ partial) {
    __shared__ float sm[8][32 * NW + 1];
    int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t S = n_samples(in);
    const int64_t chunk = (S + gridDim.x - 1) / gridDim.x;       // CTA b owns samples [b * chunk, (b + 1) * chunk)
    int64_t s_begin = (int64_t)blockIdx.x * chunk, s_end = s_begin + chunk < S ? s_begin + chunk : S;
    float wa[NW], gw[NW];
#pragma unroll
    for (int i = 0; i < NW; ++i) { wa[i] = wal[lane + 32 * i]; gw[i] = 0.f; }
    float gb = 0.f;
    for (int64_t si = s_begin + wib; si < s_end; si += 8) {
        int p0 = in.pair_start[si], np = in.pair_start[si + 1] - p0;
        float ga = galpha[si], gf[NW];
#pragma unroll
        for (int i = 0; i < NW; ++i) gf[i] = GFS[si * d.incp + lane + 32 * i];
        if (d.order == 1) {
            // one density per sample from the interpolated feature FS = sum_k wc_k H4_k: d/dFS gains g_araw * w_alpha,
            // the head's parameter gradients read FS (recomputed here), and the per-pair work is the plain weighted split
            float garaw = np > 0 ? ga * sigmoid1(pairv[p0].w - 1.f) : 0.f;
            float fs[NW];
#pragma unroll
            for (int i = 0; i < NW; ++i) { gf[i] = fmaf(garaw, wa[i], gf[i]); fs[i] = 0.f; }
            for (int j = 0; j < np; ++j) {
                int64_t m = p0 + j;
                float4 pv = pairv[m];
                float h[NW], dot = 0.f;
#pragma unroll
                for (int i = 0; i < NW; ++i) { h[i] = H4[m * d.W + lane + 32 * i]; dot = fmaf(gf[i], h[i], dot); }
                dot = warp_sum(dot);
                if (lane == 0) PG[m * pgw + d.F + 6] = dot * pv.y;
#pragma unroll
                for (int i = 0; i < NW; ++i) {
                    dPre4[m * ld4 + lane + 32 * i] = pv.x * gf[i] * (h[i] > 0.f ? 1.f : d.slope);
                    fs[i] = fmaf(h[i], pv.x, fs[i]);
                }
            }
#pragma unroll
            for (int i = 0; i < NW; ++i) gw[i] = fmaf(garaw, fs[i], gw[i]);
            gb += garaw;
            continue;
        }
        for (int j = 0; j < np; ++j) {
            int64_t m = p0 + j;
            float4 pv = pairv[m];  // (wc, wn, araw)
            float h[NW], dot = 0.f;
#pragma unroll
            for (int i = 0; i < NW; ++i) { h[i] = H4[m * d.W + lane + 32 * i]; dot = fmaf(gf[i], h[i], dot); }
            dot = warp_sum(dot);
            float x = pv.z - 1.f;
            float garaw = ga * pv.x * sigmoid1(x);
            if (lane == 0) PG[m * pgw + d.F + 6] = (ga * softplus1(x) + dot) * pv.y;  // d/d conf (straight-through clamp)
#pragma unroll
            for (int i = 0; i < NW; ++i) {
                float gh = fmaf(pv.x, gf[i], garaw * wa[i]);
                dPre4[m * ld4 + lane + 32 * i] = gh * (h[i] > 0.f ? 1.f : d.slope);
                gw[i] = fmaf(garaw, h[i], gw[i]);
            }
            gb += garaw;
        }
    }
#pragma unroll
    for (int i = 0; i < NW; ++i) sm[wib][lane + 32 * i] = gw[i];
    if (lane == 0) sm[wib][32 * NW] = gb;
    __syncthreads();
    for (int c = threadIdx.x; c <= 32 * NW; c += blockDim.x) {
        float a = 0.f;
#pragma unroll
        for (int w8 = 0; w8 < 8; ++w8) a += sm[w8][c];
        partial[(int64_t)blockIdx.x * (32 * NW + 1) + c] = a;
    }
}
__global__ void __launch_bounds__(256) agg_alpha_final_kernel(const float* __restrict__ partial, int nblk, int W,
                                                              float* __restrict__ g_wal, float* __restrict__ g_bal) {
    // one warp per column; lane l sums partials l, l+32, ... in order, then a fixed butterfly: deterministic
    int c = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    int lane = threadIdx.x & 31;
    if (c > W) return;
    float a = 0.f;
    for (int b = lane; b < nblk; b += 32) a += partial[(int64_t)b * (W + 1) + c];
    a = warp_sum(a);
    if (lane == 0) {
        if (c < W) g_wal[c] = a; else g_bal[0] = a;
    }
}

// block3 extras backward: thread per pair.  G3 [M, in3p] holds d/d(block3 input); columns W..W+6.
__global__ void __launch_bounds__(256) agg_extras_bwd_kernel(pn_agg_inputs in, AggDims d, const float* __restrict__ G3,
                                                             float* __restrict__ PG, int pgw) {
    const int64_t M = n_pairs(in);
  for (int64_t m = thread_id_global(); m < M; m += threads_total()) {
    int src = in.pair_src[m];
    int slot = src / d.K;
    float vx, vy, vz;
    load_view(in, d, slot, vx, vy, vz);
    const float* g = G3 + m * d.in3p + d.W;
    float* o = PG + m * pgw + d.F;
    o[0] = g[0]; o[1] = g[1]; o[2] = g[2];
    float px = g[3] + g[6] * vx, py = g[4] + g[6] * vy, pz = g[5] + g[6] * vz;  // d/d(R*dir)
    const float* R = in.Rw2c;
    o[3] = R[0] * px + R[3] * py + R[6] * pz;  // R^T
    o[4] = R[1] * px + R[4] * py + R[7] * pz;
    o[5] = R[2] * px + R[5] * py + R[8] * pz;
  }
}

// block1 input backward: warp per pair, lane = embedding channel.  d sin(e 2^f)/de = 2^f cos(e 2^f): the
// cos / sin values are read back from the saved input row, no trigonometry is recomputed.
__global__ void __launch_bounds__(256) agg_prep_bwd_kernel(pn_agg_inputs in, AggDims d, const float* __restrict__ X1,
                                                           const float* __restrict__ GX1, float* __restrict__ PG, int pgw) {
    const int lane = threadIdx.x & 31;
    const int64_t M = n_pairs(in);
  for (int64_t m = warp_id_global(); m < M; m += warps_total()) {
    const float* x = X1 + m * d.in1p;
    const float* g = GX1 + m * d.in1p;
    for (int c = lane; c < d.F; c += 32) {
        float ge = g[c], fr = 1.f;
        for (int f = 0; f < d.ff; ++f) {
            int col = d.F + (c * d.ff + f) * 2;
            ge = fmaf(fr, x[col + 1] * g[col] - x[col] * g[col + 1], ge);
            fr *= 2.f;
        }
        PG[m * pgw + c] = ge;
    }
  }
}

// per-pair gradients -> dense [n_slots*K, *] layout of the reference's sampled_* tensors
__global__ void __launch_bounds__(256) agg_scatter_dense_kernel(pn_agg_inputs in, AggDims d, const float* __restrict__ PG,
                                                                int pgw, float* __restrict__ g_emb, float* __restrict__ g_color,
                                                                float* __restrict__ g_dir, float* __restrict__ g_conf) {
    const int cols = d.F + 7;
    const int64_t n = n_pairs(in) * cols;
  for (int64_t idx = thread_id_global(); idx < n; idx += threads_total()) {
    int64_t m = idx / cols;
    int c = (int)(idx - m * cols);
    int64_t src = in.pair_src[m];
    float v = PG[m * pgw + c];
    if (c < d.F) { if (g_emb) g_emb[src * d.F + c] = v; }
    else if (c < d.F + 3) { if (g_color) g_color[src * 3 + (c - d.F)] = v; }
    else if (c < d.F + 6) { if (g_dir) g_dir[src * 3 + (c - d.F - 3)] = v; }
    else if (g_conf) g_conf[src] = v;
  }
}

// ---- per-point segmented reduction (replaces the atomic index_add_ of index_select's backward) -----
__global__ void __launch_bounds__(256) agg_key_count_kernel(pn_agg_inputs in, int* __restrict__ cnt, int* __restrict__ rank) {
    const int64_t M = n_pairs(in);
    for (int64_t m = thread_id_global(); m < M; m += threads_total()) {
        int g = in.sample_pidx[in.pair_src[m]];
        rank[m] = atomicAdd(&cnt[g], 1);
    }
}
__global__ void __launch_bounds__(256) agg_key_place_kernel(pn_agg_inputs in, const int* __restrict__ start,
                                                            const int* __restrict__ rank, int* __restrict__ tmp) {
    const int64_t M = n_pairs(in);
    for (int64_t m = thread_id_global(); m < M; m += threads_total()) {
        int g = in.sample_pidx[in.pair_src[m]];
        tmp[start[g] + rank[m]] = (int)m;
    }
}
__global__ void __launch_bounds__(256) agg_key_order_kernel(pn_agg_inputs in, const int* __restrict__ start,
                                                            const int* __restrict__ tmp, int* __restrict__ sorted,
                                                            int* __restrict__ skey) {
    const int64_t M = n_pairs(in);
    for (int64_t m = thread_id_global(); m < M; m += threads_total()) {
        int g = in.sample_pidx[in.pair_src[m]];
        int s = start[g], e = start[g + 1], smaller = 0;
        for (int j = s; j < e; ++j) smaller += (tmp[j] < (int)m);
        sorted[s + smaller] = (int)m;
        skey[s + smaller] = g;          // point index at each sorted position: segment boundaries without a gather chain
    }
}
// one warp per sorted position; only segment heads work.  Lanes own columns; rows are added in ascending
// pair order, so the sum is reproducible run to run.
__global__ void __launch_bounds__(256) agg_segment_reduce_kernel(pn_agg_inputs in, AggDims d, const int* __restrict__ skey,
                                                                 const int* __restrict__ sorted, const float* __restrict__ PG,
                                                                 int pgw, float* __restrict__ g_emb, float* __restrict__ g_color,
                                                                 float* __restrict__ g_dir, float* __restrict__ g_conf) {
    const int lane = threadIdx.x & 31;
    const int64_t M = n_pairs(in);
    const int cols = d.F + 7;
  for (int64_t i = warp_id_global(); i < M; i += warps_total()) {
    const int g = skey[i];
    if (i > 0 && skey[i - 1] == g) continue;  // not the head of its segment
    for (int c0 = 0; c0 < cols; c0 += 64) {               // 64 columns per walk of the segment (one walk when F <= 57)
        float a[2] = {0.f, 0.f};
        for (int64_t j0 = i; j0 < M; j0 += 32) {          // rows of the segment, 32 sorted positions per probe
            const int64_t j = j0 + lane;
            const bool mine = j < M && skey[j] == g;
            const int mrow = mine ? sorted[j] : -1;
            const unsigned mask = __ballot_sync(kFull, mine);
            const int cnt = mask == kFull ? 32 : __ffs(~mask) - 1;   // leading run that belongs to this segment
            for (int t = 0; t < cnt; ++t) {               // ascending pair order: reproducible sums
                const int64_t m = __shfl_sync(kFull, mrow, t);
                if (c0 + lane < cols) a[0] += PG[m * pgw + c0 + lane];
                if (c0 + lane + 32 < cols) a[1] += PG[m * pgw + c0 + lane + 32];
            }
            if (cnt < 32) break;
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int c = c0 + lane + 32 * h;
            if (c >= cols) continue;
            if (c < d.F) { if (g_emb) g_emb[(int64_t)g * d.F + c] = a[h]; }
            else if (c < d.F + 3) { if (g_color) g_color[(int64_t)g * 3 + (c - d.F)] = a[h]; }
            else if (c < d.F + 6) { if (g_dir) g_dir[(int64_t)g * 3 + (c - d.F - 3)] = a[h]; }
            else if (g_conf) g_conf[g] = a[h];
        }
    }
  }
}

// ---- touched rows (multi-GPU sparse gradient exchange) ----------------------------------------------------
// flag[i] = 1 iff sorted position i is the head of its per-point segment
// (flag has `cap` entries: positions >= the device pair count are 0, so the scan over the capacity leaves rank[cap] = heads)
__global__ void __launch_bounds__(256) agg_head_flag_kernel(pn_agg_inputs in, int64_t cap, const int* __restrict__ skey, int* __restrict__ flag) {
    const int64_t M = n_pairs(in);
    for (int64_t i = thread_id_global(); i < cap; i += threads_total())
        flag[i] = (i < M && (i == 0 || skey[i - 1] != skey[i])) ? 1 : 0;
}
// ids[rank of head i among heads] = point index; heads are in ascending point order, so ids is sorted and unique
__global__ void __launch_bounds__(256) agg_head_ids_kernel(pn_agg_inputs in, int64_t cap, const int* __restrict__ skey, const int* __restrict__ rank,
                                                           int* __restrict__ ids, int* __restrict__ n_touched) {
    const int64_t M = n_pairs(in);
    if (thread_id_global() == 0 && n_touched) n_touched[0] = rank[cap];
    for (int64_t i = thread_id_global(); i < M; i += threads_total())
        if (i == 0 || skey[i - 1] != skey[i]) ids[rank[i]] = skey[i];
}

}  // namespace pn
