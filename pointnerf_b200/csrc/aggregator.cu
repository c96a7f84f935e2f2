// pn_agg_forward / pn_agg_backward: the fused gather + PointAggregator (viewmlp, order 2) pipeline.
// Reference: models/aggregators/point_aggregators.py:729-816 (forward), 488-646 (viewmlp),
// neural_points.py:788-799 (gather); backward replaces autograd through those ops.
//
// Layer plan (compact rows; M pairs, S samples; all operand widths padded to multiples of 16):
//   prep    : X1[M,in1p] = [emb | PE(emb) | PE(dist)], X3[M, W..] = block3 extras, pairv = (w*conf, w)
//   block1  : H1 = lrelu(X1 W1a^T + b), X3[:, :W] = lrelu(H1 W1b^T + b)
//   block3  : H3 = lrelu(X3 W3a^T + b), H4 = lrelu(H3 W3b^T + b)
//   reduce  : alpha_s = sum_k softplus(H4 w_a + b_a - 1) wc,  FS[S, :W] = sum_k H4 wc,  FS[S, W:] = PE(view)
//   colour  : C1 = lrelu(FS Wc0^T + b), C2 = lrelu(C1 Wc2^T + b), C3 = lrelu(C2 Wc4^T + b)
//   finish  : decoded[slot] = [alpha_s | C3]
//
// Row counts.  S and M are data dependent (pn_compact_samples counts them on the device).  With in->counts == NULL the
// caller has read them back and in->S / in->M are exact.  With in->counts set nothing ever reaches the host: in->S / in->M
// are CAPACITIES (buffers and launches are sized for them), every kernel reads the true counts from in->counts, the GEMMs
// cut their tiles / reduction splits on the device, and the whole pass is a fixed sequence of launches.
#include "agg_kernels.cuh"
#include "gemm.cuh"

namespace pn {

static int make_dims(const pn_agg_params* p, AggDims& d) {
    PN_CHECK_ARG(p, "aggregator: null params");
    d.F = p->F; d.ff = p->feat_freqs; d.df = p->dist_freqs; d.vf = p->view_freqs;
    d.W = p->W; d.C = p->C; d.K = p->K; d.SR = p->SR; d.dirs_per_ray = p->dirs_per_ray;
    d.slope = p->leaky_slope; d.dist_scale = p->dist_scale; d.compact_out = p->compact_out;
    d.order = p->intrp_order == 0 ? 2 : p->intrp_order;
    PN_CHECK_ARG(d.order == 1 || d.order == 2, "aggregator: agg_intrp_order %d unsupported (1 or 2)", d.order);
    PN_CHECK_ARG(d.F >= 1 && d.F <= 1024 && d.ff >= 0 && d.df >= 1 && d.df <= 16 && d.vf >= 1 && d.vf <= 16,
                 "aggregator: unsupported feature/frequency counts");
    PN_CHECK_ARG(d.W == 128 || d.W == 256, "aggregator: shading_feature_num %d unsupported (128 or 256)", d.W);
    PN_CHECK_ARG(d.C % 16 == 0 && d.C >= 16 && d.C <= 256, "aggregator: shading_color_channel_num %d unsupported", d.C);
    PN_CHECK_ARG(d.K >= 1 && d.K <= 32, "aggregator: K=%d unsupported (1..32)", d.K);
    d.in1 = d.F + 2 * d.ff * d.F + 12 * d.df;
    d.in3 = d.W + 7;
    d.inc = d.W + 6 * d.vf;
    d.in1p = (int)align_up(d.in1, 16);
    d.in3p = (int)align_up(d.in3, 16);
    d.incp = (int)align_up(d.inc, 16);
    return PN_OK;
}

struct Saved {
    float *X1, *H1, *X3, *H3, *H4, *FS, *C1, *C2, *C3, *alpha;
    float4* pairv;
    // LeakyReLU sign bits of H1, H2 (= X3[:, :W]), H3 (W / 32 words per pair row) and C1, C2 (C / 32 words per sample row),
    // written by the forward GEMMs' epilogues for every full group of 32 rows: the data-gradient GEMMs then read 4 B per row
    // and sub-tile for their mask instead of 128 B of the activation
    uint32_t *sH1, *sH2, *sH3, *sC1, *sC2;
};
static int64_t carve_saved(const AggDims& d, int64_t S, int64_t M, void* p, int64_t bytes, Saved* sv) {
    Carver c(p, bytes);
    int64_t m = M > 0 ? M : 1, s = S > 0 ? S : 1;
    Saved t;
    t.X1 = c.take<float>(m * d.in1p);
    t.H1 = c.take<float>(m * d.W);
    t.X3 = c.take<float>(m * d.in3p);
    t.H3 = c.take<float>(m * d.W);
    t.H4 = c.take<float>(m * d.W);
    t.pairv = c.take<float4>(m);
    t.FS = c.take<float>(s * d.incp);
    t.C1 = c.take<float>(s * d.C);
    t.C2 = c.take<float>(s * d.C);
    t.C3 = c.take<float>(s * d.C);
    t.alpha = c.take<float>(s);
    t.sH1 = c.take<uint32_t>(m * cdiv(d.W, 32));
    t.sH2 = c.take<uint32_t>(m * cdiv(d.W, 32));
    t.sH3 = c.take<uint32_t>(m * cdiv(d.W, 32));
    t.sC1 = c.take<uint32_t>(s * cdiv(d.C, 32));
    t.sC2 = c.take<uint32_t>(s * cdiv(d.C, 32));
    if (sv) *sv = t;
    return c.off;
}

// launch geometry of the grid-stride kernels: one block per 8 warp items / 256 thread items OF THE CAPACITY (blocks beyond
// the device-side count leave at once; a capped, looping grid measured 10-25 % slower on the latency-bound kernels: every
// warp then walks its ~14 items one dependent gather chain after the other), bounded only by the grid limit
static unsigned grid_warp_items(int64_t items) {            // one warp per item, 8 warps per block
    int64_t b = cdiv((items > 0 ? items : 1) * 32, 256), mx = 1 << 30;
    return (unsigned)(b < mx ? b : mx);
}
static unsigned grid_thread_items(int64_t items) {          // one thread per item
    int64_t b = cdiv(items > 0 ? items : 1, 256), mx = 1 << 30;
    return (unsigned)(b < mx ? b : mx);
}

}  // namespace pn

using namespace pn;
#define PN_TRY(x) do { int _rc = (x); if (_rc) return _rc; } while (0)

extern "C" int64_t pn_agg_saved_bytes(const pn_agg_params* p, int64_t S, int64_t M) {
    AggDims d;
    if (make_dims(p, d)) return -1;
    return carve_saved(d, S, M, nullptr, 0, nullptr);
}

namespace pn {
struct FwdWs { char *I1a, *I1b, *I3a, *I3b, *Ic0, *Ic2, *Ic4; };
struct BwdWs {
    float *GA, *GB, *GA2, *GB2, *GA3, *GSa, *GSb, *GSc, *GSd, *galpha, *PG, *apart;
    float* tn[7];                                    // split partials of the seven weight-gradient GEMMs (finished by ONE launch)
    char *J1a, *J1b, *J3a, *J3b, *Jc0, *Jc2, *Jc4;   // operand images of the TRANSPOSED weights (data gradients)
    int *cnt, *rank, *tmp, *sorted, *skey, *scan, *hflag, *hscan;
    int nblk_reduce;
};
constexpr int kReduceBlocks = kNumSMs * 4;

static int64_t carve_fwd(const AggDims& d, void* p, int64_t bytes, FwdWs* w) {
    Carver c(p, bytes);
    FwdWs t;
    // operand images of the weights (gemm.cuh: pack_nt_image): pre-split, pre-swizzled, one TMA bulk copy per K chunk
    t.I1a = c.take<char>(nt_image_bytes(d.W, d.in1p));
    t.I1b = c.take<char>(nt_image_bytes(d.W, d.W));
    t.I3a = c.take<char>(nt_image_bytes(d.W, d.in3p));
    t.I3b = c.take<char>(nt_image_bytes(d.W, d.W));
    t.Ic0 = c.take<char>(nt_image_bytes(d.C, d.incp));
    t.Ic2 = c.take<char>(nt_image_bytes(d.C, d.C));
    t.Ic4 = c.take<char>(nt_image_bytes(d.C, d.C));
    if (w) *w = t;
    return c.off;
}
static int64_t carve_bwd(const AggDims& d, int64_t S, int64_t M, int64_t T, void* p, int64_t bytes, BwdWs* w) {
    Carver c(p, bytes);
    int64_t m = M > 0 ? M : 1, s = S > 0 ? S : 1;
    BwdWs t;
    t.J1a = c.take<char>(nt_image_bytes(d.in1p, d.W));
    t.J1b = c.take<char>(nt_image_bytes(d.W, d.W));
    t.J3a = c.take<char>(nt_image_bytes(d.in3p, d.W));
    t.J3b = c.take<char>(nt_image_bytes(d.W, d.W));
    t.Jc0 = c.take<char>(nt_image_bytes(d.incp, d.C));
    t.Jc2 = c.take<char>(nt_image_bytes(d.C, d.C));
    t.Jc4 = c.take<char>(nt_image_bytes(d.C, d.C));
    int64_t wide = d.in1p > d.in3p ? d.in1p : d.in3p;
    t.GA = c.take<float>(m * wide);
    t.GB = c.take<float>(m * d.W);
    // second pair of gradient buffers: every data gradient keeps its own buffer until the weight-gradient GEMM that reads
    // it has run, so the weight gradients can trail the data-gradient chain on another stream (pn_agg_backward)
    t.GA2 = c.take<float>(m * d.in3p);
    t.GB2 = c.take<float>(m * d.W);
    t.GA3 = c.take<float>(m * d.in1p);   // d/dX1 in the chain-first schedule: every other gradient buffer is still awaited by a held-back weight gradient
    t.GSa = c.take<float>(s * d.incp);
    t.GSb = c.take<float>(s * d.incp);
    t.GSc = c.take<float>(s * d.incp);
    t.GSd = c.take<float>(s * d.incp);
    t.galpha = c.take<float>(s);
    t.PG = c.take<float>(m * (d.F + 8));
    t.tn[0] = c.take<float>(gemm_tn_scratch_floats(s, d.C, d.C));       // color.4
    t.tn[1] = c.take<float>(gemm_tn_scratch_floats(s, d.C, d.C));       // color.2
    t.tn[2] = c.take<float>(gemm_tn_scratch_floats(s, d.C, d.incp));    // color.0
    t.tn[3] = c.take<float>(gemm_tn_scratch_floats(m, d.W, d.W));       // block3.2
    t.tn[4] = c.take<float>(gemm_tn_scratch_floats(m, d.W, d.in3p));    // block3.0
    t.tn[5] = c.take<float>(gemm_tn_scratch_floats(m, d.W, d.W));       // block1.2
    t.tn[6] = c.take<float>(gemm_tn_scratch_floats(m, d.W, d.in1p));    // block1.0
    t.nblk_reduce = (int)(s < kReduceBlocks ? s : kReduceBlocks);
    t.apart = c.take<float>((int64_t)kReduceBlocks * (d.W + 1));
    int64_t tt = T > 0 ? T : 1;
    t.cnt = c.take<int>(tt + 1);
    t.rank = c.take<int>(m);
    t.tmp = c.take<int>(m);
    t.sorted = c.take<int>(m);
    t.skey = c.take<int>(m);
    t.scan = c.take<int>(scan_scratch_elems_i32(tt + 1));
    t.hflag = c.take<int>(m + 1);
    t.hscan = c.take<int>(scan_scratch_elems_i32(m + 1));
    if (w) *w = t;
    return c.off;
}
}  // namespace pn

extern "C" int64_t pn_agg_workspace_bytes(const pn_agg_params* p, int64_t S, int64_t M, int64_t T, int backward) {
    AggDims d;
    if (make_dims(p, d)) return -1;
    return backward ? carve_bwd(d, S, M, T, nullptr, 0, nullptr) : carve_fwd(d, nullptr, 0, nullptr);
}

// ---- LeakyReLU activation signs of a saved forward ------------------------------------------------------------
// The MLP is piecewise linear: its backward is a function of the forward's activation pattern.  This entry exports the
// pattern the backward of THIS forward will use (unit > 0 -> slope 1, else leaky_slope), so that a checker can evaluate a
// reference backward conditioned on the same pattern: gradients are then comparable on every ray, including units whose
// pre-activation is within rounding of zero (oracle/kinks.py).
namespace pn {
__global__ void __launch_bounds__(256) act_signs_kernel(const float* __restrict__ src, int64_t rows, int ld, int cols,
                                                        uint8_t* __restrict__ out) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < rows * cols; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / cols;
        const int c = (int)(i - r * cols);
        out[i] = src[r * ld + c] > 0.f ? 1 : 0;
    }
}
}  // namespace pn
extern "C" int pn_agg_activation_signs(const pn_agg_params* p, const void* saved, int64_t saved_bytes, int64_t S_cap, int64_t M_cap,
                                       int64_t S, int64_t M, uint8_t* pair_signs, uint8_t* sample_signs, void* stream) {
    AggDims d;
    PN_TRY(make_dims(p, d));
    PN_CHECK_ARG(saved && pair_signs && sample_signs && S <= S_cap && M <= M_cap, "pn_agg_activation_signs: bad argument");
    PN_CHECK_ARG(saved_bytes >= carve_saved(d, S_cap, M_cap, nullptr, 0, nullptr), "pn_agg_activation_signs: saved buffer too small");
    if (S <= 0 || M <= 0) return PN_OK;
    Saved sv;
    carve_saved(d, S_cap, M_cap, const_cast<void*>(saved), saved_bytes, &sv);
    cudaStream_t st = (cudaStream_t)stream;
    const float* pl[4] = {sv.H1, sv.X3, sv.H3, sv.H4};
    const int pld[4] = {d.W, d.in3p, d.W, d.W};
    for (int l = 0; l < 4; ++l) {
        act_signs_kernel<<<grid_thread_items(M * d.W), 256, 0, st>>>(pl[l], M, pld[l], d.W, pair_signs + (int64_t)l * M * d.W);
        PN_LAUNCH_CHECK("act_signs");
    }
    const float* sl[3] = {sv.C1, sv.C2, sv.C3};
    for (int l = 0; l < 3; ++l) {
        act_signs_kernel<<<grid_thread_items(S * d.C), 256, 0, st>>>(sl[l], S, d.C, d.C, sample_signs + (int64_t)l * S * d.C);
        PN_LAUNCH_CHECK("act_signs");
    }
    return PN_OK;
}

// C[rows, N] = epilogue(A[rows, Kd] x image(B)^T); rows = min(*rows_dev, rows_cap) when the count lives on the device
static GemmNT nt(const float* A, int64_t lda, const char* Bimg, float* C, int64_t ldc, int64_t rows_cap, const int32_t* rows_dev, int N, int Kd,
                 const float* bias, int leaky, float slope) {
    GemmNT g;
    g.A = A; g.lda = lda; g.B = nullptr; g.ldb = 0; g.C = C; g.ldc = ldc; g.M = rows_cap; g.N = N; g.Kd = Kd;
    g.bias = bias; g.leaky = leaky; g.slope = slope; g.aux = nullptr; g.ldaux = 0; g.aux_cols = 0; g.rowmap = nullptr; g.Bimg = Bimg;
    g.Mdev = rows_dev;
    g.sign_out = nullptr; g.sign_in = nullptr; g.sign_ld = 0;
    return g;
}
// layers whose width is whole 32-bit words (and at least one TMA box) keep their activation signs as bits
static bool sign_ok(int width) { return width % 32 == 0 && width >= 32; }
static GemmNT with_sign_out(GemmNT g, uint32_t* bits, int width, bool wanted) {
    if (wanted && sign_ok(width)) { g.sign_out = bits; g.sign_ld = width / 32; }
    return g;
}
static GemmNT with_mask(GemmNT g, const uint32_t* bits, const float* act, int64_t ldact, int width) {
    g.aux = act; g.ldaux = ldact; g.aux_cols = width;
    if (sign_ok(width)) { g.sign_in = bits; g.sign_ld = width / 32; }
    return g;
}

extern "C" int pn_agg_forward(const pn_agg_params* p, const pn_agg_inputs* in, const pn_agg_weights* w, float* decoded,
                              float* weight_out, void* saved, int64_t saved_bytes, void* workspace, int64_t workspace_bytes,
                              void* stream) {
    AggDims d;
    PN_TRY(make_dims(p, d));
    PN_CHECK_ARG(in && w && saved && (decoded || (p->compact_out && in->S == 0)), "pn_agg_forward: null argument");
    const int64_t S = in->S, M = in->M;           // exact counts, or capacities when in->counts is set
    PN_CHECK_ARG(saved_bytes >= carve_saved(d, S, M, nullptr, 0, nullptr), "pn_agg_forward: saved buffer too small");
    PN_CHECK_ARG(workspace_bytes >= carve_fwd(d, nullptr, 0, nullptr), "pn_agg_forward: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    if (!d.compact_out) PN_CUDA(cudaMemsetAsync(decoded, 0, in->n_slots * (int64_t)(d.C + 1) * sizeof(float), st));
    if (weight_out) PN_CUDA(cudaMemsetAsync(weight_out, 0, in->n_slots * (int64_t)d.K * sizeof(float), st));
    if (S <= 0 || M <= 0) return PN_OK;
    const int32_t* Sdev = in->counts;
    const int32_t* Mdev = in->counts ? in->counts + 1 : nullptr;
    Saved sv;
    carve_saved(d, S, M, saved, saved_bytes, &sv);
    FwdWs ws;
    carve_fwd(d, workspace, workspace_bytes, &ws);
    const PackJob jobs[7] = {{w->block1_0_w, d.in1, 1, d.W, d.in1, d.W, d.in1p, ws.I1a}, {w->block1_2_w, d.W, 1, d.W, d.W, d.W, d.W, ws.I1b},
                             {w->block3_0_w, d.in3, 1, d.W, d.in3, d.W, d.in3p, ws.I3a}, {w->block3_2_w, d.W, 1, d.W, d.W, d.W, d.W, ws.I3b},
                             {w->color_0_w, d.inc, 1, d.C, d.inc, d.C, d.incp, ws.Ic0},  {w->color_2_w, d.C, 1, d.C, d.C, d.C, d.C, ws.Ic2},
                             {w->color_4_w, d.C, 1, d.C, d.C, d.C, d.C, ws.Ic4}};
    PN_TRY(pack_nt_images(jobs, 7, st));
    const bool bwd_follows = !p->forward_only;
    agg_prep_kernel<<<grid_warp_items(S), 256, 0, st>>>(*in, d, sv.X1, sv.X3, sv.pairv, weight_out);
    PN_LAUNCH_CHECK("agg_prep");
    if (d.F == 32 && d.ff == 3 && d.df == 5)
        agg_prep_rows_kernel<32, 3, 5><<<grid_warp_items(M), 256, 8 * d.in1p * sizeof(float), st>>>(*in, d, sv.X1);
    else
        agg_prep_rows_kernel<0, 0, 0><<<grid_warp_items(M), 256, 8 * d.in1p * sizeof(float), st>>>(*in, d, sv.X1);
    PN_LAUNCH_CHECK("agg_prep_rows");
    PN_TRY(gemm_nt(with_sign_out(nt(sv.X1, d.in1p, ws.I1a, sv.H1, d.W, M, Mdev, d.W, d.in1p, w->block1_0_b, 1, d.slope), sv.sH1, d.W, bwd_follows), st));
    PN_TRY(gemm_nt(with_sign_out(nt(sv.H1, d.W, ws.I1b, sv.X3, d.in3p, M, Mdev, d.W, d.W, w->block1_2_b, 1, d.slope), sv.sH2, d.W, bwd_follows), st));
    PN_TRY(gemm_nt(with_sign_out(nt(sv.X3, d.in3p, ws.I3a, sv.H3, d.W, M, Mdev, d.W, d.in3p, w->block3_0_b, 1, d.slope), sv.sH3, d.W, bwd_follows), st));
    PN_TRY(gemm_nt(nt(sv.H3, d.W, ws.I3b, sv.H4, d.W, M, Mdev, d.W, d.W, w->block3_2_b, 1, d.slope), st));
    if (d.W == 256)
        agg_reduce_kernel<8><<<grid_warp_items(S), 256, 0, st>>>(*in, d, sv.H4, sv.pairv, w->alpha_0_w, w->alpha_0_b, sv.FS, sv.alpha);
    else
        agg_reduce_kernel<4><<<grid_warp_items(S), 256, 0, st>>>(*in, d, sv.H4, sv.pairv, w->alpha_0_w, w->alpha_0_b, sv.FS, sv.alpha);
    PN_LAUNCH_CHECK("agg_reduce");
    PN_TRY(gemm_nt(with_sign_out(nt(sv.FS, d.incp, ws.Ic0, sv.C1, d.C, S, Sdev, d.C, d.incp, w->color_0_b, 1, d.slope), sv.sC1, d.C, bwd_follows), st));
    PN_TRY(gemm_nt(with_sign_out(nt(sv.C1, d.C, ws.Ic2, sv.C2, d.C, S, Sdev, d.C, d.C, w->color_2_b, 1, d.slope), sv.sC2, d.C, bwd_follows), st));
    PN_TRY(gemm_nt(nt(sv.C2, d.C, ws.Ic4, sv.C3, d.C, S, Sdev, d.C, d.C, w->color_4_b, 1, d.slope), st));
    agg_finish_kernel<<<grid_warp_items(S), 256, 0, st>>>(*in, d, sv.alpha, sv.C3, decoded);
    PN_LAUNCH_CHECK("agg_finish");
    return PN_OK;
}

// ---- two-stream backward -------------------------------------------------------------------------------------------
// The backward is a CHAIN of data-gradient kernels (each needs the previous one's output) plus seven weight-gradient
// GEMMs that only read what the chain left behind.  Every GEMM is a persistent kernel on all SMs whose last wave leaves
// most SMs idle (527 tiles on 74 CTA pairs: 9 pairs run an 8th tile) and whose prologue (barriers, TMEM, first loads) is
// exposed.  The chain runs on a library-owned high-priority stream, the weight gradients on the caller's stream: CTAs of
// whichever kernel is queued fill the SMs the other one's tail frees.  Events order the two (RAW per gradient buffer, one
// WAR before the last chain GEMM reuses GA); the caller's stream joins the chain at the end, so for the caller the call is
// still "enqueued on the given stream" (and, under stream capture, the fork / join becomes two branches of the graph).
namespace pn {
struct BwdStreams {
    cudaStream_t chain = nullptr;
    cudaEvent_t ev[12] = {};
    cudaEvent_t chain_end = nullptr;     // end of the data-gradient chain of the most recent backward (pn_agg_backward_wait_rows)
    bool chain_end_valid = false;
    int device = -1;
    bool ok = false;
};
static BwdStreams* bwd_streams() {
    static BwdStreams per_dev[16];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) return nullptr;
    BwdStreams& b = per_dev[dev];
    if (!b.ok) {
        int lo = 0, hi = 0;
        if (cudaDeviceGetStreamPriorityRange(&lo, &hi) != cudaSuccess) return nullptr;
        if (cudaStreamCreateWithPriority(&b.chain, cudaStreamNonBlocking, hi) != cudaSuccess) return nullptr;
        for (auto& e : b.ev)
            if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) return nullptr;
        if (cudaEventCreateWithFlags(&b.chain_end, cudaEventDisableTiming) != cudaSuccess) return nullptr;
        b.device = dev;
        b.ok = true;
    }
    return &b;
}
static int g_bwd_two_streams = 1;
static int g_bwd_chain_first = 0;    // pn_set_backward_schedule(2): all weight-gradient GEMMs after the data-gradient chain
}  // namespace pn

extern "C" int pn_agg_backward(const pn_agg_params* p, const pn_agg_inputs* in, const pn_agg_weights* w,
                               const float* grad_decoded, const void* saved, int64_t saved_bytes, const pn_agg_weights* gw,
                               int scatter_rows, float* g_emb, float* g_color, float* g_dir, float* g_conf,
                               int32_t* touched_ids, int32_t* n_touched, void* workspace, int64_t workspace_bytes,
                               void* stream) {
    AggDims d;
    PN_TRY(make_dims(p, d));
    PN_CHECK_ARG(in && w && saved && gw && (grad_decoded || (p->compact_out && in->S == 0)), "pn_agg_backward: null argument");
    const int64_t S = in->S, M = in->M, T = in->T;       // S, M: exact counts, or capacities when in->counts is set
    PN_CHECK_ARG(saved_bytes >= carve_saved(d, S, M, nullptr, 0, nullptr), "pn_agg_backward: saved buffer too small");
    PN_CHECK_ARG(workspace_bytes >= carve_bwd(d, S, M, T, nullptr, 0, nullptr), "pn_agg_backward: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    const int32_t* Sdev = in->counts;
    const int32_t* Mdev = in->counts ? in->counts + 1 : nullptr;
    // row gradients default to zero (empty slots / unreferenced table rows); scatter_rows == 2: the caller keeps the [T, *]
    // buffers zero outside the touched rows (pn_rows_zero on the previous step's rows) - at 20 M points the four memsets
    // are 3.1 GB per step
    int64_t rows = scatter_rows ? T : in->n_slots * (int64_t)d.K;
    if (scatter_rows != 2) {
        if (g_emb) PN_CUDA(cudaMemsetAsync(g_emb, 0, rows * d.F * sizeof(float), st));
        if (g_color) PN_CUDA(cudaMemsetAsync(g_color, 0, rows * 3 * sizeof(float), st));
        if (g_dir) PN_CUDA(cudaMemsetAsync(g_dir, 0, rows * 3 * sizeof(float), st));
        if (g_conf) PN_CUDA(cudaMemsetAsync(g_conf, 0, rows * sizeof(float), st));
    }
    if (touched_ids && scatter_rows) PN_CUDA(cudaMemsetAsync(touched_ids, 0xFF, (M > 0 ? M : 1) * sizeof(int), st));
    if (n_touched) PN_CUDA(cudaMemsetAsync(n_touched, 0, sizeof(int), st));
    if (S <= 0 || M <= 0) {                               // (capacities are never 0 for a batch with slots)
        struct { float* ptr; int64_t n; } items[] = {
            {gw->block1_0_w, (int64_t)d.W * d.in1}, {gw->block1_0_b, d.W}, {gw->block1_2_w, (int64_t)d.W * d.W}, {gw->block1_2_b, d.W},
            {gw->block3_0_w, (int64_t)d.W * d.in3}, {gw->block3_0_b, d.W}, {gw->block3_2_w, (int64_t)d.W * d.W}, {gw->block3_2_b, d.W},
            {gw->alpha_0_w, d.W}, {gw->alpha_0_b, 1}, {gw->color_0_w, (int64_t)d.C * d.inc}, {gw->color_0_b, d.C},
            {gw->color_2_w, (int64_t)d.C * d.C}, {gw->color_2_b, d.C}, {gw->color_4_w, (int64_t)d.C * d.C}, {gw->color_4_b, d.C}};
        for (auto& it : items)
            if (it.ptr) PN_CUDA(cudaMemsetAsync(it.ptr, 0, it.n * sizeof(float), st));
        return PN_OK;
    }

    Saved sv;
    carve_saved(d, S, M, const_cast<void*>(saved), saved_bytes, &sv);
    BwdWs ws;
    carve_bwd(d, S, M, T, workspace, workspace_bytes, &ws);
    const int pgw = d.F + 8;
    // block1.0's data gradient is only needed for the embedding-derived input columns [emb | PE(emb)]: the distance
    // encoding behind them has no trainable input (xyz_grad = 0), so the GEMM stops at column ng1 (224 of 288)
    const int ng1 = (int)align_up(d.F + 2 * d.ff * d.F, 16) < d.in1p ? (int)align_up(d.F + 2 * d.ff * d.F, 16) : d.in1p;
    {   // images of W^T: B[n = input column, k = output unit] = W[k, n]
        const PackJob jobs[7] = {{w->block1_0_w, 1, d.in1, ng1 < d.in1 ? ng1 : d.in1, d.W, ng1, d.W, ws.J1a}, {w->block1_2_w, 1, d.W, d.W, d.W, d.W, d.W, ws.J1b},
                                 {w->block3_0_w, 1, d.in3, d.in3, d.W, d.in3p, d.W, ws.J3a}, {w->block3_2_w, 1, d.W, d.W, d.W, d.W, d.W, ws.J3b},
                                 {w->color_0_w, 1, d.inc, d.inc, d.C, d.incp, d.C, ws.Jc0},  {w->color_2_w, 1, d.C, d.C, d.C, d.C, d.C, ws.Jc2},
                                 {w->color_4_w, 1, d.C, d.C, d.C, d.C, d.C, ws.Jc4}};
        PN_TRY(pack_nt_images(jobs, 7, st));
    }

    // ---- streams: hs = data-gradient chain, st = weight gradients (see BwdStreams); one stream when disabled
    const bool chain_first = g_bwd_chain_first != 0;
    BwdStreams* bs = (g_bwd_two_streams || chain_first) ? bwd_streams() : nullptr;
    if (!bs) {
        BwdStreams* any = bwd_streams();
        if (any) any->chain_end_valid = false;
    }
    // chain_first: one stream, but the weight-gradient GEMMs are held back until the chain has finished and its end is
    // marked (chain_end) - a multi-GPU host exchanges the per-row gradients on another stream under those GEMMs
    cudaStream_t hs = (bs && !chain_first) ? bs->chain : st;
    struct Held { GemmTN t; float* scratch; int slot; } held[8];
    TnFinish fin[8];
    int n_held = 0, n_fin = 0;
    auto wgrad = [&](GemmTN t, float* scratch) -> int {
        const int slot = n_fin++;
        if (chain_first) { held[n_held++] = Held{t, scratch, slot}; return PN_OK; }
        // finished right behind the GEMM: on the weight-gradient stream the split sum hides under the data-gradient chain (one
        // batched finish after the join measured +50 us on the critical path)
        (void)fin;
        return gemm_tn(t, scratch, nullptr, st);
    };
    int evi = 0;
    auto order = [&](cudaStream_t from, cudaStream_t to) -> int {   // work queued on `to` from now on sees everything queued on `from` so far
        if (!bs || chain_first) return PN_OK;
        cudaEvent_t e = bs->ev[evi++];
        PN_CUDA(cudaEventRecord(e, from));
        PN_CUDA(cudaStreamWaitEvent(to, e, 0));
        return PN_OK;
    };
    PN_TRY(order(st, hs));                      // inputs, zero-fills and weight images are ready
    // ---- colour branch (S rows): gradient buffers GSa -> GSb -> GSc -> GSd
    agg_gout_kernel<<<grid_warp_items(S), 256, 0, hs>>>(*in, d, grad_decoded, sv.C3, ws.GSa, ws.galpha);
    PN_LAUNCH_CHECK("agg_gout");
    PN_TRY(order(hs, st));
    {   // color.4
        PN_TRY(wgrad(GemmTN{ws.GSa, d.C, sv.C2, d.C, gw->color_4_w, d.C, d.C, gw->color_4_b, S, d.C, d.C, Sdev}, ws.tn[0]));
        PN_TRY(gemm_nt(with_mask(nt(ws.GSa, d.C, ws.Jc4, ws.GSb, d.C, S, Sdev, d.C, d.C, nullptr, 0, d.slope), sv.sC2, sv.C2, d.C, d.C), hs));
        PN_TRY(order(hs, st));
    }
    {   // color.2
        PN_TRY(wgrad(GemmTN{ws.GSb, d.C, sv.C1, d.C, gw->color_2_w, d.C, d.C, gw->color_2_b, S, d.C, d.C, Sdev}, ws.tn[1]));
        PN_TRY(gemm_nt(with_mask(nt(ws.GSb, d.C, ws.Jc2, ws.GSc, d.C, S, Sdev, d.C, d.C, nullptr, 0, d.slope), sv.sC1, sv.C1, d.C, d.C), hs));
        PN_TRY(order(hs, st));
    }
    {   // color.0: d/dFS (only the first W columns are consumed)
        PN_TRY(wgrad(GemmTN{ws.GSc, d.C, sv.FS, d.incp, gw->color_0_w, d.inc, d.inc, gw->color_0_b, S, d.C, d.incp, Sdev}, ws.tn[2]));
        PN_TRY(gemm_nt(nt(ws.GSc, d.C, ws.Jc0, ws.GSd, d.incp, S, Sdev, d.incp, d.C, nullptr, 0, d.slope), hs));
    }
    // ---- K-reduction and density head
    if (d.W == 256)
        agg_reduce_bwd_kernel<8><<<ws.nblk_reduce, 256, 0, hs>>>(*in, d, sv.H4, sv.pairv, w->alpha_0_w, ws.GSd, ws.galpha,
                                                                 ws.GA, d.W, ws.PG, pgw, ws.apart);
    else
        agg_reduce_bwd_kernel<4><<<ws.nblk_reduce, 256, 0, hs>>>(*in, d, sv.H4, sv.pairv, w->alpha_0_w, ws.GSd, ws.galpha,
                                                                 ws.GA, d.W, ws.PG, pgw, ws.apart);
    PN_LAUNCH_CHECK("agg_reduce_bwd");
    agg_alpha_final_kernel<<<(unsigned)cdiv((int64_t)(d.W + 1) * 32, 256), 256, 0, hs>>>(ws.apart, ws.nblk_reduce, d.W, gw->alpha_0_w, gw->alpha_0_b);
    PN_LAUNCH_CHECK("agg_alpha_final");
    PN_TRY(order(hs, st));
    // ---- block3 (M rows).  GA = dPre4 [M,W]
    {   // block3.2
        PN_TRY(wgrad(GemmTN{ws.GA, d.W, sv.H3, d.W, gw->block3_2_w, d.W, d.W, gw->block3_2_b, M, d.W, d.W, Mdev}, ws.tn[3]));
        PN_TRY(gemm_nt(with_mask(nt(ws.GA, d.W, ws.J3b, ws.GB, d.W, M, Mdev, d.W, d.W, nullptr, 0, d.slope), sv.sH3, sv.H3, d.W, d.W), hs));  // GB = dPre3
        PN_TRY(order(hs, st));
    }
    cudaEvent_t ga_free = nullptr;              // recorded on st once block3.2's weight gradient has read GA (dPre4)
    {   // block3.0: GA2[M,in3p] = d/dX3, first W columns masked by lrelu'(H2) (H2 = X3[:, :W])
        if (bs && !chain_first) {               // st so far: ... TN(block3.2).  Mark the point where GA becomes reusable.
            ga_free = bs->ev[evi++];
            PN_CUDA(cudaEventRecord(ga_free, st));
        }
        PN_TRY(wgrad(GemmTN{ws.GB, d.W, sv.X3, d.in3p, gw->block3_0_w, d.in3, d.in3, gw->block3_0_b, M, d.W, d.in3p, Mdev}, ws.tn[4]));
        PN_TRY(gemm_nt(with_mask(nt(ws.GB, d.W, ws.J3a, ws.GA2, d.in3p, M, Mdev, d.in3p, d.W, nullptr, 0, d.slope), sv.sH2, sv.X3, d.in3p, d.W), hs));
    }
    agg_extras_bwd_kernel<<<grid_thread_items(M), 256, 0, hs>>>(*in, d, ws.GA2, ws.PG, pgw);
    PN_LAUNCH_CHECK("agg_extras_bwd");
    PN_TRY(order(hs, st));
    // ---- block1.  GA2[:, :W] = dPre2 (row stride in3p)
    {   // block1.2
        PN_TRY(wgrad(GemmTN{ws.GA2, d.in3p, sv.H1, d.W, gw->block1_2_w, d.W, d.W, gw->block1_2_b, M, d.W, d.W, Mdev}, ws.tn[5]));
        PN_TRY(gemm_nt(with_mask(nt(ws.GA2, d.in3p, ws.J1b, ws.GB2, d.W, M, Mdev, d.W, d.W, nullptr, 0, d.slope), sv.sH1, sv.H1, d.W, d.W), hs));  // GB2 = dPre1
        PN_TRY(order(hs, st));
    }
    {   // block1.0: GA[M,in1p] = d/dX1 (GA is free again once block3.2's weight gradient has read dPre4)
        PN_TRY(wgrad(GemmTN{ws.GB2, d.W, sv.X1, d.in1p, gw->block1_0_w, d.in1, d.in1, gw->block1_0_b, M, d.W, d.in1p, Mdev}, ws.tn[6]));
        if (ga_free) PN_CUDA(cudaStreamWaitEvent(hs, ga_free, 0));
        // (chain_first: all four earlier gradient buffers are still awaited by held-back weight gradients -> GA3)
        PN_TRY(gemm_nt(nt(ws.GB2, d.W, ws.J1a, chain_first ? ws.GA3 : ws.GA, d.in1p, M, Mdev, ng1, d.W, nullptr, 0, d.slope), hs));
    }
    agg_prep_bwd_kernel<<<grid_warp_items(M), 256, 0, hs>>>(*in, d, sv.X1, chain_first ? ws.GA3 : ws.GA, ws.PG, pgw);
    PN_LAUNCH_CHECK("agg_prep_bwd");
    // ---- per-row gradients out
    if (!scatter_rows) {
        agg_scatter_dense_kernel<<<grid_thread_items(M * (d.F + 7)), 256, 0, hs>>>(*in, d, ws.PG, pgw, g_emb, g_color, g_dir, g_conf);
        PN_LAUNCH_CHECK("agg_scatter_dense");
    } else {
        PN_CUDA(cudaMemsetAsync(ws.cnt, 0, (T + 1) * sizeof(int), hs));
        const unsigned mb = grid_thread_items(M);
        agg_key_count_kernel<<<mb, 256, 0, hs>>>(*in, ws.cnt, ws.rank);
        PN_LAUNCH_CHECK("agg_key_count");
        PN_TRY(scan_exclusive_i32(ws.cnt, ws.cnt, T, true, ws.scan, hs));
        agg_key_place_kernel<<<mb, 256, 0, hs>>>(*in, ws.cnt, ws.rank, ws.tmp);
        PN_LAUNCH_CHECK("agg_key_place");
        agg_key_order_kernel<<<mb, 256, 0, hs>>>(*in, ws.cnt, ws.tmp, ws.sorted, ws.skey);
        PN_LAUNCH_CHECK("agg_key_order");
        agg_segment_reduce_kernel<<<grid_warp_items(M), 256, 0, hs>>>(*in, d, ws.skey, ws.sorted, ws.PG, pgw, g_emb, g_color,
                                                                      g_dir, g_conf);
        PN_LAUNCH_CHECK("agg_segment_reduce");
        if (touched_ids) {   // sorted unique list of the table rows that received gradient
            agg_head_flag_kernel<<<mb, 256, 0, hs>>>(*in, M, ws.skey, ws.hflag);
            PN_LAUNCH_CHECK("agg_head_flag");
            PN_TRY(scan_exclusive_i32(ws.hflag, ws.hflag, M, true, ws.hscan, hs));
            agg_head_ids_kernel<<<mb, 256, 0, hs>>>(*in, M, ws.skey, ws.hflag, touched_ids, n_touched);
            PN_LAUNCH_CHECK("agg_head_ids");
        }
    }
    if (bs) {                                   // per-row gradients and touched ids are final here, before the trailing weight gradients
        PN_CUDA(cudaEventRecord(bs->chain_end, hs));
        bs->chain_end_valid = true;
    }
    PN_TRY(order(hs, st));                      // the caller's stream joins the chain
    for (int i = 0; i < n_held; ++i) PN_TRY(gemm_tn(held[i].t, held[i].scratch, &fin[held[i].slot], st));
    if (n_held) PN_TRY(tn_finish_batch(fin, n_fin, st));    // chain-first schedule: the held-back split reductions in ONE launch
    return PN_OK;
}

extern "C" int pn_set_backward_schedule(int schedule) {
    PN_CHECK_ARG(schedule >= 0 && schedule <= 2, "pn_set_backward_schedule: 0 (one stream), 1 (interleaved on two streams), 2 (chain first)");
    g_bwd_two_streams = schedule == 1;
    g_bwd_chain_first = schedule == 2;
    return PN_OK;
}
extern "C" int pn_get_backward_schedule(void) { return g_bwd_chain_first ? 2 : (g_bwd_two_streams ? 1 : 0); }

extern "C" int pn_agg_backward_wait_rows(void* stream) {
    BwdStreams* bs = (g_bwd_two_streams || g_bwd_chain_first) ? bwd_streams() : nullptr;
    if (!bs || !bs->chain_end_valid) return 0;
    PN_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, bs->chain_end, 0));
    return 1;
}

// ---- the two GEMM building blocks of the MLP as entry points (kernel-level tests and the roofline leg of bench.py) --------
extern "C" int64_t pn_gemm_nt_image_bytes(int32_t N, int32_t Kd) { return nt_image_bytes(N, Kd); }
extern "C" int pn_gemm_nt(const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc, int64_t M, const int32_t* M_dev,
                          int32_t N, int32_t Kd, const float* bias, int32_t leaky, float slope, void* image_scratch, void* stream) {
    PN_CHECK_ARG(image_scratch, "pn_gemm_nt: image_scratch (pn_gemm_nt_image_bytes bytes) is required");
    PN_TRY(pack_nt_image(B, ldb, 1, N, Kd, N, Kd, (char*)image_scratch, (cudaStream_t)stream));
    return gemm_nt(nt(A, lda, (const char*)image_scratch, C, ldc, M, M_dev, N, Kd, bias, leaky, slope), (cudaStream_t)stream);
}
extern "C" int64_t pn_gemm_tn_scratch_floats(int64_t M, int32_t N, int32_t Kb) { return gemm_tn_scratch_floats(M, N, Kb); }
extern "C" int pn_gemm_tn(const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc, int32_t Kout,
                          float* colsum, int64_t M, const int32_t* M_dev, int32_t N, int32_t Kb, float* scratch, void* stream) {
    return gemm_tn(GemmTN{A, lda, B, ldb, C, ldc, Kout, colsum, M, N, Kb, M_dev}, scratch, nullptr, (cudaStream_t)stream);
}

// ---- sparse exchange of per-point gradient rows between ranks (ray_shard.py) ---------------------------------
namespace pn {
// out[u] = [id bits | emb(F) | conf(1) | dir(3) | color(3)] of table row ids[u]; u >= n_ids or id < 0 -> id = -1, zeros
__global__ void __launch_bounds__(256) rows_pack_kernel(const int* __restrict__ ids, int64_t n_ids, int64_t cap, int F,
                                                        const float* __restrict__ g_emb,
                                                        const float* __restrict__ g_conf, const float* __restrict__ g_dir,
                                                        const float* __restrict__ g_color, float* __restrict__ out) {
    const int W = F + 8;
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= cap * W) return;
    int64_t u = idx / W;
    int c = (int)(idx - u * W);
    int g = u < n_ids ? ids[u] : -1;
    float v = 0.f;
    if (c == 0) v = __int_as_float(g);
    else if (g >= 0) {
        int k = c - 1;
        if (k < F) v = g_emb[(int64_t)g * F + k];
        else if (k == F) v = g_conf ? g_conf[g] : 0.f;
        else if (k < F + 4) v = g_dir ? g_dir[(int64_t)g * 3 + (k - F - 1)] : 0.f;
        else v = g_color ? g_color[(int64_t)g * 3 + (k - F - 4)] : 0.f;
    }
    out[idx] = v;
}
// pn_rows_merge: every table row named in the gathered buffer becomes the sum over ranks IN RANK ORDER of the rows carrying
// its id (identical bits on every replica, and the plain v0 + v1 of a dense all-reduce at two ranks).  One pass: a slot table
// [n_points, n_ranks] (persistent, -1 outside this call) records where each rank holds each id; the FIRST rank that holds an id
// owns it: its warp walks the ranks in order, reads each present row once, and WRITES the table row (no read-modify-write, no
// zero pass: rows touched only by this rank's own backward are among the gathered rows and get overwritten).
__global__ void __launch_bounds__(256) rows_slot_kernel(const float* __restrict__ rows, int64_t cap, int n_ranks, int W, int set,
                                                        int* __restrict__ slots) {
    int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= cap * n_ranks) return;
    const int g = __float_as_int(rows[e * W]);
    if (g < 0) return;
    const int r = (int)(e / cap);
    slots[(int64_t)g * n_ranks + r] = set ? (int)(e - (int64_t)r * cap) : -1;
}
__global__ void __launch_bounds__(256) rows_sum_kernel(const float* __restrict__ rows, int64_t cap, int n_ranks, int F,
                                                       const int* __restrict__ slots, float* __restrict__ g_emb,
                                                       float* __restrict__ g_conf, float* __restrict__ g_dir,
                                                       float* __restrict__ g_color) {
    // eight lanes per gathered entry (four entries per warp: the chain id -> slots -> rows -> store is latency bound, so
    // entries in flight are what counts); lane j of a group owns columns j, j + 8, ... of the W - 1 <= 63 value columns
    const int W = F + 8;
    const int lane = threadIdx.x & 31, j = lane & 7, grp = lane >> 3;
    const unsigned gmask = 0xffu << (grp * 8);
    const int64_t e = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 4 + grp;
    const bool in_range = e < cap * n_ranks;
    const int g = in_range ? __float_as_int(rows[e * W]) : -1;
    const int r = in_range ? (int)(e / cap) : 0;
    int slotv[4];
    unsigned present = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {            // n_ranks <= 32: four rounds of eight ranks; whole-warp ballots, so no early exit above
        const int rr = j + 8 * k;
        slotv[k] = (g >= 0 && rr < n_ranks) ? slots[(int64_t)g * n_ranks + rr] : -1;
        const unsigned b = __ballot_sync(0xffffffffu, slotv[k] >= 0);
        present |= ((b >> (grp * 8)) & 0xffu) << (8 * k);
    }
    if (g < 0 || __ffs(present) - 1 != r) return;      // padding, or an earlier rank owns this id (uniform inside the group)
    float acc[8];
    bool first = true;
    while (present) {
        const int rr = __ffs(present) - 1;
        present &= present - 1;
        const int k = rr >> 3;
        const int sv = k == 0 ? slotv[0] : k == 1 ? slotv[1] : k == 2 ? slotv[2] : slotv[3];
        const int u = __shfl_sync(gmask, sv, (grp << 3) | (rr & 7));
        const float* src = rows + ((int64_t)rr * cap + u) * W + 1;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int c = j + 8 * i;
            const float v = c < W - 1 ? src[c] : 0.f;
            acc[i] = first ? v : acc[i] + v;
        }
        first = false;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int c = j + 8 * i;
        if (c >= W - 1) continue;
        float* dst;
        if (c < F) dst = g_emb + (int64_t)g * F + c;
        else if (c == F) dst = g_conf ? g_conf + g : nullptr;
        else if (c < F + 4) dst = g_dir ? g_dir + (int64_t)g * 3 + (c - F - 1) : nullptr;
        else dst = g_color ? g_color + (int64_t)g * 3 + (c - F - 4) : nullptr;
        if (dst) *dst = acc[i];
    }
}
}  // namespace pn

namespace pn {
__global__ void __launch_bounds__(256) rows_mark_kernel(const int* __restrict__ pidx, int64_t n, int64_t n_points, int* __restrict__ mark) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int g = pidx[i];
    if (g >= 0 && g < n_points) mark[g] = 1;      // benign same-value race
}
__global__ void __launch_bounds__(256) rows_count_kernel(const int* __restrict__ mark, int64_t n_points, int* __restrict__ count) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int v = i < n_points ? mark[i] : 0;
    v = __reduce_add_sync(0xffffffffu, v);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(count, v);   // integer sum: order-independent
}
// conf_coefficient of the aggregator (point_aggregators.py:809-811 on the gathered points_conf of neural_points.py:
// index_select with sample_pidx.clamp(min=0)): out[i] = clamp(table[max(idx[i], 0)], lo, hi), one pass over the query result.
// The reference's straight-through clamp evaluates x - (x - clamp(x)) (:724-726), which is NOT always clamp(x) in fp32 when
// the clamp is active; the same two subtractions are done here so the values are bit-identical.
__device__ __forceinline__ float st_clamp(float x, float lo, float hi) {
    return __fsub_rn(x, __fsub_rn(x, fminf(fmaxf(x, lo), hi)));
}
__global__ void __launch_bounds__(256) gather_clamp_kernel(const float* __restrict__ table, const int32_t* __restrict__ idx, int64_t n,
                                                           float lo, float hi, float* __restrict__ out) {
    int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i + 3 < n) {
        int4 p = *reinterpret_cast<const int4*>(idx + i);
        float4 v;
        v.x = st_clamp(__ldg(table + max(p.x, 0)), lo, hi);
        v.y = st_clamp(__ldg(table + max(p.y, 0)), lo, hi);
        v.z = st_clamp(__ldg(table + max(p.z, 0)), lo, hi);
        v.w = st_clamp(__ldg(table + max(p.w, 0)), lo, hi);
        *reinterpret_cast<float4*>(out + i) = v;
    } else {
        for (; i < n; ++i) out[i] = st_clamp(__ldg(table + max(idx[i], 0)), lo, hi);
    }
}
}  // namespace pn

extern "C" int pn_count_touched_rows(const int32_t* sample_pidx, int64_t n, int64_t n_points, int32_t* mark, int32_t* count,
                                     void* stream) {
    PN_CHECK_ARG(mark && count && n_points > 0 && (sample_pidx || n == 0), "pn_count_touched_rows: bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    PN_CUDA(cudaMemsetAsync(mark, 0, n_points * sizeof(int), st));
    PN_CUDA(cudaMemsetAsync(count, 0, sizeof(int), st));
    if (n > 0) {
        rows_mark_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(sample_pidx, n, n_points, mark);
        PN_LAUNCH_CHECK("rows_mark");
    }
    rows_count_kernel<<<(unsigned)cdiv(n_points, 256), 256, 0, st>>>(mark, n_points, count);
    PN_LAUNCH_CHECK("rows_count");
    return PN_OK;
}

extern "C" int pn_gather_clamp(const float* table, int64_t n_points, const int32_t* idx, int64_t n, float lo, float hi, float* out,
                               void* stream) {
    PN_CHECK_ARG(table && n_points > 0 && (n == 0 || (idx && out)) && n >= 0, "pn_gather_clamp: bad argument");
    PN_CHECK_ARG((((uintptr_t)idx | (uintptr_t)out) & 15) == 0, "pn_gather_clamp: idx / out must be 16-byte aligned");
    if (n == 0) return PN_OK;
    gather_clamp_kernel<<<(unsigned)cdiv(cdiv(n, 4), 256), 256, 0, (cudaStream_t)stream>>>(table, idx, n, lo, hi, out);
    PN_LAUNCH_CHECK("gather_clamp");
    return PN_OK;
}

namespace pn {
// rows named by ids[i * stride] (>= 0) of the four gradient tables become zero: a step's gradients live in persistent
// [n_points, *] buffers of which only the rows the PREVIOUS step touched are non-zero
__global__ void __launch_bounds__(256) rows_zero_kernel(const int* __restrict__ ids, int64_t n, int64_t stride, int F, float* __restrict__ g_emb,
                                                        float* __restrict__ g_conf, float* __restrict__ g_dir, float* __restrict__ g_color) {
    const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;      // eight lanes per row
    const int j = threadIdx.x & 7;
    if (i >= n) return;
    const int g = ids[i * stride];
    if (g < 0) return;
    for (int c = j; c < F; c += 8) g_emb[(int64_t)g * F + c] = 0.f;
    if (j < 3) {
        if (g_dir) g_dir[(int64_t)g * 3 + j] = 0.f;
        if (g_color) g_color[(int64_t)g * 3 + j] = 0.f;
    } else if (j == 3 && g_conf) g_conf[g] = 0.f;
}
}  // namespace pn
extern "C" int pn_rows_zero(const int32_t* ids, int64_t n, int64_t stride, int32_t F, float* g_emb, float* g_conf, float* g_dir, float* g_color,
                            void* stream) {
    PN_CHECK_ARG((ids || n == 0) && n >= 0 && stride >= 1 && g_emb && F >= 1, "pn_rows_zero: bad argument");
    if (n == 0) return PN_OK;
    rows_zero_kernel<<<(unsigned)cdiv(n * 8, 256), 256, 0, (cudaStream_t)stream>>>(ids, n, stride, F, g_emb, g_conf, g_dir, g_color);
    PN_LAUNCH_CHECK("rows_zero");
    return PN_OK;
}

extern "C" int pn_rows_pack(const int32_t* ids, int64_t n_ids, int64_t cap, int32_t F, const float* g_emb, const float* g_conf, const float* g_dir,
                            const float* g_color, float* out, void* stream) {
    PN_CHECK_ARG((ids || n_ids == 0) && n_ids >= 0 && g_emb && out && F >= 1 && F <= 56, "pn_rows_pack: bad argument (F <= 56)");
    if (cap <= 0) return PN_OK;
    rows_pack_kernel<<<(unsigned)cdiv(cap * (F + 8), 256), 256, 0, (cudaStream_t)stream>>>(ids, n_ids, cap, F, g_emb, g_conf, g_dir, g_color, out);
    PN_LAUNCH_CHECK("rows_pack");
    return PN_OK;
}
extern "C" int pn_rows_merge(const float* gathered, int32_t n_ranks, int64_t cap, int32_t F, int64_t n_points, int32_t* slots,
                             float* g_emb, float* g_conf, float* g_dir, float* g_color, void* stream) {
    PN_CHECK_ARG(gathered && g_emb && slots && n_points > 0 && n_ranks >= 1 && n_ranks <= 32 && F >= 1 && F <= 56,
                 "pn_rows_merge: bad argument (F <= 56, n_ranks <= 32)");
    if (cap <= 0) return PN_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t all_rows = (int64_t)n_ranks * cap;
    rows_slot_kernel<<<(unsigned)cdiv(all_rows, 256), 256, 0, st>>>(gathered, cap, n_ranks, F + 8, 1, slots);
    PN_LAUNCH_CHECK("rows_slot_set");
    rows_sum_kernel<<<(unsigned)cdiv(cdiv(all_rows, 4) * 32, 256), 256, 0, st>>>(gathered, cap, n_ranks, F, slots, g_emb, g_conf, g_dir, g_color);
    PN_LAUNCH_CHECK("rows_sum");
    rows_slot_kernel<<<(unsigned)cdiv(all_rows, 256), 256, 0, st>>>(gathered, cap, n_ranks, F + 8, 0, slots);
    PN_LAUNCH_CHECK("rows_slot_clear");
    return PN_OK;
}
