// fp32 SIMT GEMM building blocks for the aggregator MLP (128x128x16 tiles, 8x8 outputs per thread,
// register-prefetched double-buffered shared memory).  This is the exact-fp32 path: it is the
// in-library reference the tensor-core path is validated against on the device, and the path used
// when the caller asks for strict fp32 (PN_MLP=simt).
#include "gemm.cuh"

namespace pn {

constexpr int BM = 128, BN = 128, BK = 16, LDS = BM + 4;

__device__ __forceinline__ void tile_fma(const float* __restrict__ As, const float* __restrict__ Bs, float (&acc)[8][8], int ty,
                                         int tx) {
#pragma unroll
    for (int k = 0; k < BK; ++k) {
        float4 a0 = *reinterpret_cast<const float4*>(&As[k * LDS + ty * 4]);
        float4 a1 = *reinterpret_cast<const float4*>(&As[k * LDS + 64 + ty * 4]);
        float4 b0 = *reinterpret_cast<const float4*>(&Bs[k * LDS + tx * 4]);
        float4 b1 = *reinterpret_cast<const float4*>(&Bs[k * LDS + 64 + tx * 4]);
        float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
        float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
}

// ---------------------------------------------------------------------------------------------
// NT: both operands K-contiguous.  Thread t loads rows (t>>2) and (t>>2)+64, k quad (t&3).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) gemm_nt_kernel(GemmNT g) {
    __shared__ __align__(16) float As[2][BK * LDS];
    __shared__ __align__(16) float Bs[2][BK * LDS];
    const int t = threadIdx.x, ty = t >> 4, tx = t & 15;
    const int64_t m0 = (int64_t)blockIdx.x * BM;
    const int n0 = blockIdx.y * BN;
    const int lr = t >> 2, kq = t & 3;
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 ra[2], rb[2];
    auto load = [&](int k0) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            int64_t m = m0 + lr + 64 * h;
            ra[h] = (m < g.M) ? *reinterpret_cast<const float4*>(g.A + m * g.lda + k0 + 4 * kq) : z4;
            int n = n0 + lr + 64 * h;
            rb[h] = (n < g.N) ? *reinterpret_cast<const float4*>(g.B + (int64_t)n * g.ldb + k0 + 4 * kq) : z4;
        }
    };
    auto store = [&](int buf) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            int r = lr + 64 * h;
            As[buf][(4 * kq + 0) * LDS + r] = ra[h].x;
            As[buf][(4 * kq + 1) * LDS + r] = ra[h].y;
            As[buf][(4 * kq + 2) * LDS + r] = ra[h].z;
            As[buf][(4 * kq + 3) * LDS + r] = ra[h].w;
            Bs[buf][(4 * kq + 0) * LDS + r] = rb[h].x;
            Bs[buf][(4 * kq + 1) * LDS + r] = rb[h].y;
            Bs[buf][(4 * kq + 2) * LDS + r] = rb[h].z;
            Bs[buf][(4 * kq + 3) * LDS + r] = rb[h].w;
        }
    };
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

    const int nk = g.Kd / BK;
    load(0);
    store(0);
    __syncthreads();
    for (int kt = 0; kt < nk; ++kt) {
        if (kt + 1 < nk) load((kt + 1) * BK);
        tile_fma(As[kt & 1], Bs[kt & 1], acc, ty, tx);
        if (kt + 1 < nk) {
            store((kt + 1) & 1);
            __syncthreads();
        }
    }
    // epilogue
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        int64_t m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        if (m >= g.M) continue;
        int64_t orow = g.rowmap ? (int64_t)g.rowmap[m] : m;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            int n = n0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
            if (n >= g.N) continue;
            float v = acc[i][j];
            if (g.bias) v += g.bias[n];
            if (g.leaky) v = v > 0.f ? v : g.slope * v;
            if (g.aux && n < g.aux_cols) v *= (g.aux[m * g.ldaux + n] > 0.f) ? 1.f : g.slope;
            g.C[orow * g.ldc + n] = v;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// TN (weight gradient): reduction index m is the slow dimension of both operands, so tiles go to
// shared memory untransposed.  Thread t loads reduction rows (t>>5) and (t>>5)+8, column quad (t&31).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) gemm_tn_kernel(GemmTN g, int64_t rows_per_split, float* __restrict__ partial) {
    __shared__ __align__(16) float As[2][BK * LDS];
    __shared__ __align__(16) float Bs[2][BK * LDS];
    const int t = threadIdx.x, ty = t >> 4, tx = t & 15;
    const int j0 = blockIdx.x * BN;  // column of B (k of the weight)
    const int i0 = blockIdx.y * BM;  // column of A (n of the weight)
    const int64_t r_begin = (int64_t)blockIdx.z * rows_per_split;
    const int64_t r_end = r_begin + rows_per_split < g.M ? r_begin + rows_per_split : g.M;
    const int rr = t >> 5, cq = t & 31;
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 ra[2], rb[2];
    auto load = [&](int64_t r0) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            int64_t r = r0 + rr + 8 * h;
            bool ok = r < r_end;
            ra[h] = (ok && i0 + 4 * cq < g.N) ? *reinterpret_cast<const float4*>(g.A + r * g.lda + i0 + 4 * cq) : z4;
            rb[h] = (ok && j0 + 4 * cq < g.Kb) ? *reinterpret_cast<const float4*>(g.B + r * g.ldb + j0 + 4 * cq) : z4;
        }
    };
    auto store = [&](int buf) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            *reinterpret_cast<float4*>(&As[buf][(rr + 8 * h) * LDS + 4 * cq]) = ra[h];
            *reinterpret_cast<float4*>(&Bs[buf][(rr + 8 * h) * LDS + 4 * cq]) = rb[h];
        }
    };
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    const int64_t nk = (r_end - r_begin + BK - 1) / BK;
    if (nk > 0) {
        load(r_begin);
        store(0);
        __syncthreads();
        for (int64_t kt = 0; kt < nk; ++kt) {
            if (kt + 1 < nk) load(r_begin + (kt + 1) * BK);
            tile_fma(As[kt & 1], Bs[kt & 1], acc, ty, tx);
            if (kt + 1 < nk) {
                store((int)((kt + 1) & 1));
                __syncthreads();
            }
        }
    }
    float* out = partial + (int64_t)blockIdx.z * g.N * g.Kb;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        int n = i0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        if (n >= g.N) continue;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            int k = j0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
            if (k < g.Kb) out[(int64_t)n * g.Kb + k] = acc[i][j];
        }
    }
}

__global__ void __launch_bounds__(256) tn_reduce_kernel(const float* __restrict__ partial, int splits, int N, int Kb, int Kout,
                                                        float* __restrict__ C, int64_t ldc) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= N * Kb) return;
    int n = idx / Kb, k = idx - n * Kb;
    if (k >= Kout) return;
    float acc = 0.f;
    for (int z = 0; z < splits; ++z) acc += partial[(int64_t)z * N * Kb + idx];
    C[(int64_t)n * ldc + k] = acc;
}

// tensor-core path: matrix and (optional) column-sum partials reduced in ONE launch, fixed order over the splits
__global__ void __launch_bounds__(256) tn_finish_kernel(const float* __restrict__ partial, const float* __restrict__ cpartial, int splits,
                                                        int N, int Kb, int Kout, float* __restrict__ C, int64_t ldc,
                                                        float* __restrict__ colsum) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx < N * Kb) {
        int n = idx / Kb, k = idx - n * Kb;
        if (k >= Kout) return;
        float acc = 0.f;
        for (int z = 0; z < splits; ++z) acc += partial[(int64_t)z * N * Kb + idx];
        C[(int64_t)n * ldc + k] = acc;
    } else if (colsum && idx - N * Kb < N) {
        int n = idx - N * Kb;
        float acc = 0.f;
        for (int z = 0; z < splits; ++z) acc += cpartial[(int64_t)z * N + n];
        colsum[n] = acc;
    }
}

__global__ void __launch_bounds__(128) colsum_partial_kernel(const float* __restrict__ A, int64_t lda, int64_t M, int N,
                                                             int64_t rows_per_split, float* __restrict__ partial) {
    int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    int64_t r0 = (int64_t)blockIdx.y * rows_per_split, r1 = r0 + rows_per_split < M ? r0 + rows_per_split : M;
    float acc = 0.f;
    for (int64_t r = r0; r < r1; ++r) acc += A[r * lda + n];
    partial[(int64_t)blockIdx.y * N + n] = acc;
}
__global__ void __launch_bounds__(128) colsum_final_kernel(const float* __restrict__ partial, int splits, int N,
                                                           float* __restrict__ out) {
    int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    float acc = 0.f;
    for (int z = 0; z < splits; ++z) acc += partial[(int64_t)z * N + n];
    out[n] = acc;
}

__global__ void __launch_bounds__(256) transpose_pad_kernel(const float* __restrict__ src, int rows, int cols, int64_t lds,
                                                            float* __restrict__ dst, int64_t ldd, int dst_rows_padded) {
    // dst is [dst_rows_padded, ldd]; element (c, r) = src(r, c) when in range else 0
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)dst_rows_padded * ldd) return;
    int c = (int)(idx / ldd), r = (int)(idx - (int64_t)c * ldd);
    dst[idx] = (c < cols && r < rows) ? src[(int64_t)r * lds + c] : 0.f;
}
__global__ void __launch_bounds__(256) copy_pad_kernel(const float* __restrict__ src, int rows, int cols, int64_t lds,
                                                       float* __restrict__ dst, int64_t ldd) {
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)rows * ldd) return;
    int r = (int)(idx / ldd), c = (int)(idx - (int64_t)r * ldd);
    dst[idx] = c < cols ? src[(int64_t)r * lds + c] : 0.f;
}

static int64_t tn_rows_per_split(int64_t M) {
    int64_t rps = align_up(cdiv(M > 0 ? M : 1, 256), BK);
    return rps < 1024 ? 1024 : rps;
}

static int g_mlp_path = 1;   // 0 = exact-fp32 SIMT, 1 = tcgen05 3xTF32 (default)
int mlp_path() { return g_mlp_path; }
void set_mlp_path(int p) { g_mlp_path = p; }
int gemm_nt_tc(const GemmNT& g, cudaStream_t st);
int gemm_tn_tc(const GemmTN& g, float* scratch, cudaStream_t st);
int64_t gemm_tn_tc_scratch_floats(int64_t M, int N, int Kb);

int gemm_nt(const GemmNT& g, cudaStream_t st) {
    if (g_mlp_path == 1) return gemm_nt_tc(g, st);
    PN_CHECK_ARG(g.Kd % BK == 0 && g.lda % 4 == 0 && g.ldb % 4 == 0, "gemm_nt: operands must be padded (Kd %d lda %lld ldb %lld)",
                 g.Kd, (long long)g.lda, (long long)g.ldb);
    PN_CHECK_ARG(((uintptr_t)g.A & 15) == 0 && ((uintptr_t)g.B & 15) == 0, "gemm_nt: operands must be 16-byte aligned");
    if (g.M <= 0 || g.N <= 0) return PN_OK;
    dim3 grid((unsigned)cdiv(g.M, BM), (unsigned)cdiv(g.N, BN));
    gemm_nt_kernel<<<grid, 256, 0, st>>>(g);
    PN_LAUNCH_CHECK("gemm_nt");
    return PN_OK;
}

int64_t gemm_tn_scratch_floats(int64_t M, int N, int Kb) {
    int64_t splits = cdiv(M > 0 ? M : 1, tn_rows_per_split(M));
    int64_t a = splits * ((int64_t)N * Kb + N) + 64, b = gemm_tn_tc_scratch_floats(M, N, Kb);
    return a > b ? a : b;   // either path may be selected at run time
}

// second stage shared by both paths: partials are combined in split order (deterministic)
int tn_finish(const GemmTN& g, const float* partial, const float* cpartial, int splits, cudaStream_t st) {
    float* cs = cpartial ? g.colsum : nullptr;
    tn_finish_kernel<<<(unsigned)cdiv((int64_t)g.N * g.Kb + (cs ? g.N : 0), 256), 256, 0, st>>>(partial, cpartial, splits, g.N, g.Kb, g.Kout,
                                                                                             g.C, g.ldc, cs);
    PN_LAUNCH_CHECK("tn_finish");
    return PN_OK;
}

int gemm_tn(const GemmTN& g, float* scratch, cudaStream_t st) {
    if (g_mlp_path == 1) return gemm_tn_tc(g, scratch, st);
    PN_CHECK_ARG(g.lda % 4 == 0 && g.ldb % 4 == 0 && g.N % 4 == 0 && g.Kb % 4 == 0, "gemm_tn: operands must be padded");
    PN_CHECK_ARG(((uintptr_t)g.A & 15) == 0 && ((uintptr_t)g.B & 15) == 0, "gemm_tn: operands must be 16-byte aligned");
    int64_t rps = tn_rows_per_split(g.M);
    int splits = (int)cdiv(g.M > 0 ? g.M : 1, rps);
    float* partial = scratch;
    float* cpartial = scratch + (int64_t)splits * g.N * g.Kb;
    dim3 grid((unsigned)cdiv(g.Kb, BN), (unsigned)cdiv(g.N, BM), (unsigned)splits);
    gemm_tn_kernel<<<grid, 256, 0, st>>>(g, rps, partial);
    PN_LAUNCH_CHECK("gemm_tn");
    tn_reduce_kernel<<<(unsigned)cdiv((int64_t)g.N * g.Kb, 256), 256, 0, st>>>(partial, splits, g.N, g.Kb, g.Kout, g.C, g.ldc);
    PN_LAUNCH_CHECK("tn_reduce");
    if (g.colsum) {
        dim3 cg((unsigned)cdiv(g.N, 128), (unsigned)splits);
        colsum_partial_kernel<<<cg, 128, 0, st>>>(g.A, g.lda, g.M, g.N, rps, cpartial);
        PN_LAUNCH_CHECK("colsum_partial");
        colsum_final_kernel<<<(unsigned)cdiv(g.N, 128), 128, 0, st>>>(cpartial, splits, g.N, g.colsum);
        PN_LAUNCH_CHECK("colsum_final");
    }
    return PN_OK;
}

int transpose_pad(const float* src, int rows, int cols, int64_t lds, float* dst, int64_t ldd, int dst_rows_padded,
                  cudaStream_t st) {
    int64_t n = (int64_t)dst_rows_padded * ldd;
    transpose_pad_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(src, rows, cols, lds, dst, ldd, dst_rows_padded);
    PN_LAUNCH_CHECK("transpose_pad");
    return PN_OK;
}
int copy_pad(const float* src, int rows, int cols, int64_t lds, float* dst, int64_t ldd, cudaStream_t st) {
    int64_t n = (int64_t)rows * ldd;
    copy_pad_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(src, rows, cols, lds, dst, ldd);
    PN_LAUNCH_CHECK("copy_pad");
    return PN_OK;
}

}  // namespace pn
