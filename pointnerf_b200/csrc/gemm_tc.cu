// tcgen05 GEMMs of the aggregator MLP: 3xTF32 split products, fp32 accumulation in TMEM (sm_100a).
//
// Why 3xTF32: north_star asks for 1e-4 relative in fp32 through four 256-wide layers and their gradients; one TF32 or
// BF16 pass (~1e-3 per product) does not hold that, and LeakyReLU kinks amplify forward error into gradient flips
// (DESIGN.md 2).  Each fp32 operand is split v = hi + lo (both exactly representable in TF32) while it is staged into
// shared memory, and every K-step issues three MMAs: hi*lo, lo*hi, hi*hi.  The dropped lo*lo term is ~2^-22.
//
// gemm_nt_tc:  C[m, n] = epilogue( sum_k A[m, k] * B[n, k] )   (forward layers, data gradients)
//   CTA tile 128 x BN (BN <= 256, multiple of 16), K chunks of 32 (one 128-byte swizzle atom per row), 2-stage ring.
//   warps 0-7: stage operands (global fp32 -> split -> swizzled smem) then run the epilogue (TMEM -> regs -> global);
//   warp 8: one elected thread issues tcgen05.mma and commits to mbarriers.
// gemm_tn_tc:  C[n, k] = sum_m A[m, n] * B[m, k]                (weight gradients; both operands MN-major)
//   CTA tile 128(n) x BK(k) over a slice of the reduction rows m; fixed-order second-stage reduction.
#include "gemm.cuh"
#include "tc.cuh"

namespace pn {

namespace {
constexpr int TBM = 128;          // rows of A per CTA = TMEM lanes
constexpr int TBK = 32;           // tf32 elements per 128-byte swizzle row
constexpr int TMAXN = 256;
constexpr int kLoadThreads = 256;
constexpr int kThreads = kLoadThreads + 32;
constexpr int A_BYTES = TBM * 128;              // 16 KB per (hi | lo)
constexpr int B_BYTES = TMAXN * 128;            // 32 KB per (hi | lo)
constexpr int STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;   // 96 KB
constexpr int NT_SMEM = 2 * STAGE_BYTES + 1024 + 256;

__device__ __forceinline__ uint32_t sw128_off(int row, int chunk) {
    return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((chunk ^ (row & 7)) << 4));
}
// MN-major tf32 (SWIZZLE_128B_BASE32B): rows of 128 B, the 32-byte chunk index is XORed with (row & 3)
__device__ __forceinline__ uint32_t sw128b32_off(int row, int chunk16) {
    return (uint32_t)(row * 128 + ((((chunk16 >> 1) ^ (row & 3)) << 5) | ((chunk16 & 1) << 4)));
}
__device__ __forceinline__ void split_store(char* hi_base, char* lo_base, uint32_t off, float4 v) {
    float4 h, l;
    tc::split_tf32(v.x, h.x, l.x);
    tc::split_tf32(v.y, h.y, l.y);
    tc::split_tf32(v.z, h.z, l.z);
    tc::split_tf32(v.w, h.w, l.w);
    *reinterpret_cast<float4*>(hi_base + off) = h;
    *reinterpret_cast<float4*>(lo_base + off) = l;
}
}  // namespace

__global__ void __launch_bounds__(kThreads, 1) gemm_nt_tc_kernel(GemmNT g, int BN, int tmem_cols) {
    extern __shared__ char smem_raw[];
    char* smem = (char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * STAGE_BYTES);  // full[2], empty[2], done
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int64_t m0 = (int64_t)blockIdx.x * TBM;
    const int n0 = blockIdx.y * BN;
    const int nk = (g.Kd + TBK - 1) / TBK;

    if (t == 0) {
        tc::mbar_init(&bars[0], kLoadThreads);
        tc::mbar_init(&bars[1], kLoadThreads);
        tc::mbar_init(&bars[2], 1);
        tc::mbar_init(&bars[3], 1);
        tc::mbar_init(&bars[4], 1);
        tc::mbar_fence_init();
    }
    if (warp == 8) tc::tmem_alloc(tmem_slot, (uint32_t)tmem_cols);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_d = *tmem_slot;

    if (warp < 8) {
        // ------------------------------------------------ operand staging
        const int c = t & 7, r0 = t >> 3;            // 16-byte chunk, first row
        const int nb = (BN + 31) >> 5;               // row groups of 32 in the B tile
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        float4 ca[4], cb[8], na[4], nb4[8];
        auto load = [&](int kc, float4 (&ra)[4], float4 (&rb)[8]) {
            const int gk = kc * TBK + 4 * c;
            const bool kok = gk < g.Kd;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                int64_t m = m0 + r0 + 32 * i;
                ra[i] = (kok && m < g.M) ? __ldg(reinterpret_cast<const float4*>(g.A + m * g.lda + gk)) : z4;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                int row = r0 + 32 * i;
                int n = n0 + row;
                rb[i] = (i < nb && kok && row < BN && n < g.N) ? __ldg(reinterpret_cast<const float4*>(g.B + (int64_t)n * g.ldb + gk)) : z4;
            }
        };
        load(0, ca, cb);
        for (int kc = 0; kc < nk; ++kc) {
            const int s = kc & 1;
            if (kc + 1 < nk) load(kc + 1, na, nb4);
            tc::mbar_wait(&bars[2 + s], ((kc >> 1) & 1) ^ 1);
            char* st = smem + s * STAGE_BYTES;
#pragma unroll
            for (int i = 0; i < 4; ++i) split_store(st, st + A_BYTES, sw128_off(r0 + 32 * i, c), ca[i]);
#pragma unroll
            for (int i = 0; i < 8; ++i)
                if (i < nb) split_store(st + 2 * A_BYTES, st + 2 * A_BYTES + B_BYTES, sw128_off(r0 + 32 * i, c), cb[i]);
            tc::fence_proxy_async();
            tc::mbar_arrive(&bars[s]);
#pragma unroll
            for (int i = 0; i < 4; ++i) ca[i] = na[i];
#pragma unroll
            for (int i = 0; i < 8; ++i) cb[i] = nb4[i];
        }
        // ------------------------------------------------ epilogue: TMEM -> registers -> global
        tc::mbar_wait(&bars[4], 0);
        tc::fence_after_sync();
        const int q = warp & 3, half = warp >> 2;
        const int64_t m = m0 + q * 32 + lane;
        const int nchunks = BN >> 4;
        for (int j = half; j < nchunks; j += 2) {
            float v[16];
            tc::tmem_ld16(tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)(j * 16), v);
            if (m < g.M) {
                const int nb0 = n0 + j * 16;
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    int n = nb0 + i;
                    if (n < g.N) {
                        float x = v[i];
                        if (g.bias) x += __ldg(g.bias + n);
                        if (g.leaky) x = x > 0.f ? x : g.slope * x;
                        if (g.aux && n < g.aux_cols) x *= (__ldg(g.aux + m * g.ldaux + n) > 0.f) ? 1.f : g.slope;
                        v[i] = x;
                    }
                }
                float* crow = g.C + m * g.ldc + nb0;
                if (nb0 + 16 <= g.N && ((((uintptr_t)crow) & 15) == 0)) {
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        *reinterpret_cast<float4*>(crow + 4 * i) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
                } else {
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        if (nb0 + i < g.N) crow[i] = v[i];
                }
            }
        }
        tc::fence_before_sync();
    } else {
      if (lane == 0) {
        // ------------------------------------------------ MMA issue (one thread)
        const uint32_t idesc = tc::idesc_tf32(TBM, BN, 0, 0);
        const uint32_t sbase = tc::smem_u32(smem);
        for (int kc = 0; kc < nk; ++kc) {
            const int s = kc & 1;
            tc::mbar_wait(&bars[s], (kc >> 1) & 1);
            tc::fence_after_sync();
            const uint32_t a_hi = sbase + s * STAGE_BYTES, a_lo = a_hi + A_BYTES;
            const uint32_t b_hi = a_hi + 2 * A_BYTES, b_lo = b_hi + B_BYTES;
            int ksteps = (g.Kd - kc * TBK + 7) >> 3;
            if (ksteps > 4) ksteps = 4;
            for (int k = 0; k < ksteps; ++k) {
                const uint32_t ko = 32u * k;
                uint64_t dah = tc::smem_desc_sw128(a_hi + ko, 16, 1024), dal = tc::smem_desc_sw128(a_lo + ko, 16, 1024);
                uint64_t dbh = tc::smem_desc_sw128(b_hi + ko, 16, 1024), dbl = tc::smem_desc_sw128(b_lo + ko, 16, 1024);
                tc::mma_tf32(tmem_d, dah, dbl, idesc, (kc | k) != 0);
                tc::mma_tf32(tmem_d, dal, dbh, idesc, 1);
                tc::mma_tf32(tmem_d, dah, dbh, idesc, 1);
            }
            tc::mma_commit(&bars[2 + s]);   // frees the stage when these MMAs have read it
        }
        tc::mma_commit(&bars[4]);           // accumulator complete
      }
      __syncwarp();
    }
    __syncthreads();
    if (warp == 8) {
        tc::fence_after_sync();
        tc::tmem_dealloc(tmem_d, (uint32_t)tmem_cols);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// TN: C[n, k] = sum_m A[m, n] * B[m, k].  The reduction index m is the slow dimension of both operands in global
// memory, i.e. both are MN-major for the tensor core: a chunk of 32 reduction rows is staged as 128-byte panels
// (32 consecutive n or k per row), 4 reduction rows per 32-byte-base swizzle atom (tc.cuh).  grid = (k tiles, n tiles of 128, splits of m).
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 1) gemm_tn_tc_kernel(GemmTN g, int BKt, int tmem_cols, int64_t rows_per_split,
                                                                 float* __restrict__ partial, float* __restrict__ cpartial) {
    extern __shared__ char smem_raw[];
    char* smem = (char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * STAGE_BYTES);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int j0 = blockIdx.x * BKt;       // first column of B (k of the weight)
    const int i0 = blockIdx.y * TBM;       // first column of A (n of the weight)
    const int64_t r_begin = (int64_t)blockIdx.z * rows_per_split;
    const int64_t r_end = r_begin + rows_per_split < g.M ? r_begin + rows_per_split : g.M;
    const int nk = (int)((r_end - r_begin + TBK - 1) / TBK);

    if (t == 0) {
        tc::mbar_init(&bars[0], kLoadThreads);
        tc::mbar_init(&bars[1], kLoadThreads);
        tc::mbar_init(&bars[2], 1);
        tc::mbar_init(&bars[3], 1);
        tc::mbar_init(&bars[4], 1);
        tc::mbar_fence_init();
    }
    if (warp == 8) tc::tmem_alloc(tmem_slot, (uint32_t)tmem_cols);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_d = *tmem_slot;
    constexpr int PANEL = TBK * 128;       // bytes of one 32-wide panel over a 32-row chunk

    if (warp < 8) {
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        const int ac4 = t & 31, ar0 = t >> 5;      // A: 32 float4 per row (128 columns), rows ar0 + 8 i
        const int bc4 = t & 63, br0 = t >> 6;      // B: up to 64 float4 per row (256 columns), rows br0 + 4 i
        const bool a_ok = i0 + 4 * ac4 < g.N;
        const bool b_ok = 4 * bc4 < BKt && j0 + 4 * bc4 < g.Kb;
        float4 ca[4], cb[8], na[4], nb4[8];
        float4 csum = z4;
        auto load = [&](int kc, float4 (&ra)[4], float4 (&rb)[8]) {
            const int64_t rb0 = r_begin + (int64_t)kc * TBK;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                int64_t r = rb0 + ar0 + 8 * i;
                ra[i] = (a_ok && r < r_end) ? __ldg(reinterpret_cast<const float4*>(g.A + r * g.lda + i0 + 4 * ac4)) : z4;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                int64_t r = rb0 + br0 + 4 * i;
                rb[i] = (b_ok && r < r_end) ? __ldg(reinterpret_cast<const float4*>(g.B + r * g.ldb + j0 + 4 * bc4)) : z4;
            }
        };
        if (nk > 0) load(0, ca, cb);
        for (int kc = 0; kc < nk; ++kc) {
            const int s = kc & 1;
            if (kc + 1 < nk) load(kc + 1, na, nb4);
            tc::mbar_wait(&bars[2 + s], ((kc >> 1) & 1) ^ 1);
            char* st = smem + s * STAGE_BYTES;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                int row = ar0 + 8 * i;
                split_store(st, st + A_BYTES, (uint32_t)((ac4 >> 3) * PANEL) + sw128b32_off(row, ac4 & 7), ca[i]);
                csum.x += ca[i].x; csum.y += ca[i].y; csum.z += ca[i].z; csum.w += ca[i].w;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                int row = br0 + 4 * i;
                split_store(st + 2 * A_BYTES, st + 2 * A_BYTES + B_BYTES, (uint32_t)((bc4 >> 3) * PANEL) + sw128b32_off(row, bc4 & 7), cb[i]);
            }
            tc::fence_proxy_async();
            tc::mbar_arrive(&bars[s]);
#pragma unroll
            for (int i = 0; i < 4; ++i) ca[i] = na[i];
#pragma unroll
            for (int i = 0; i < 8; ++i) cb[i] = nb4[i];
        }
        // ---- epilogue: D[n, k] -> partial[split][n][k]
        if (nk > 0) {
            tc::mbar_wait(&bars[4], 0);
            tc::fence_after_sync();
        }
        const int q = warp & 3, half = warp >> 2;
        const int n = i0 + q * 32 + lane;
        float* out = partial + (int64_t)blockIdx.z * g.N * g.Kb;
        const int nchunks = BKt >> 4;
        for (int j = half; j < nchunks; j += 2) {
            float v[16];
            if (nk > 0) {
                tc::tmem_ld16(tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)(j * 16), v);
            } else {
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = 0.f;
            }
            if (n < g.N) {
                const int k0 = j0 + j * 16;
                float* orow = out + (int64_t)n * g.Kb + k0;
#pragma unroll
                for (int i = 0; i < 16; ++i)
                    if (k0 + i < g.Kb) orow[i] = v[i];
            }
        }
        tc::fence_before_sync();
        // ---- bias gradient: column sums of A over this split (k tile 0 only), fixed order
        if (cpartial && blockIdx.x == 0) {
            float4* red = reinterpret_cast<float4*>(smem);      // stage memory is free once the accumulator is complete
            __syncwarp();
            asm volatile("bar.sync 1, %0;" ::"n"(kLoadThreads));
            red[t] = csum;
            asm volatile("bar.sync 1, %0;" ::"n"(kLoadThreads));
            if (t < 32) {
                float4 a = red[t];
#pragma unroll
                for (int w8 = 1; w8 < 8; ++w8) {
                    float4 b = red[t + 32 * w8];
                    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
                }
                int nn = i0 + 4 * t;
                if (nn < g.N) *reinterpret_cast<float4*>(cpartial + (int64_t)blockIdx.z * g.N + nn) = a;
            }
        }
    } else {
      if (lane == 0) {
        const uint32_t idesc = tc::idesc_tf32(TBM, BKt, 1, 1);
        const uint32_t sbase = tc::smem_u32(smem);
        for (int kc = 0; kc < nk; ++kc) {
            const int s = kc & 1;
            tc::mbar_wait(&bars[s], (kc >> 1) & 1);
            tc::fence_after_sync();
            const uint32_t a_hi = sbase + s * STAGE_BYTES, a_lo = a_hi + A_BYTES;
            const uint32_t b_hi = a_hi + 2 * A_BYTES, b_lo = b_hi + B_BYTES;
            int64_t rows = r_end - (r_begin + (int64_t)kc * TBK);
            int ksteps = rows >= TBK ? 4 : (int)((rows + 7) >> 3);
            for (int k = 0; k < ksteps; ++k) {
                const uint32_t ko = 1024u * k;
                uint64_t dah = tc::smem_desc_sw128(a_hi + ko, PANEL, 512, 1), dal = tc::smem_desc_sw128(a_lo + ko, PANEL, 512, 1);
                uint64_t dbh = tc::smem_desc_sw128(b_hi + ko, PANEL, 512, 1), dbl = tc::smem_desc_sw128(b_lo + ko, PANEL, 512, 1);
                tc::mma_tf32(tmem_d, dah, dbl, idesc, (kc | k) != 0);
                tc::mma_tf32(tmem_d, dal, dbh, idesc, 1);
                tc::mma_tf32(tmem_d, dah, dbh, idesc, 1);
            }
            tc::mma_commit(&bars[2 + s]);
        }
        tc::mma_commit(&bars[4]);
      }
      __syncwarp();
    }
    __syncthreads();
    if (warp == 8) {
        tc::fence_after_sync();
        tc::tmem_dealloc(tmem_d, (uint32_t)tmem_cols);
    }
}

static int pick_bn(int N, int* n_tiles) {
    int nt = (N + TMAXN - 1) / TMAXN;
    int bn = (int)align_up(cdiv(N, nt), 16);
    *n_tiles = (int)cdiv(N, bn);
    return bn;
}
static int tmem_cols_for(int bn) { return bn <= 32 ? 32 : bn <= 64 ? 64 : bn <= 128 ? 128 : 256; }

int gemm_nt_tc(const GemmNT& g, cudaStream_t st) {
    PN_CHECK_ARG(g.Kd % 4 == 0 && g.lda % 4 == 0 && g.ldb % 4 == 0, "gemm_nt_tc: operands must be padded to 4 (Kd %d lda %lld ldb %lld)",
                 g.Kd, (long long)g.lda, (long long)g.ldb);
    PN_CHECK_ARG(((uintptr_t)g.A & 15) == 0 && ((uintptr_t)g.B & 15) == 0, "gemm_nt_tc: operands must be 16-byte aligned");
    PN_CHECK_ARG(g.rowmap == nullptr, "gemm_nt_tc: rowmap unsupported");
    if (g.M <= 0 || g.N <= 0) return PN_OK;
    static bool attr_set = false;
    if (!attr_set) {
        PN_CUDA(cudaFuncSetAttribute(gemm_nt_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, NT_SMEM));
        attr_set = true;
    }
    int n_tiles;
    int bn = pick_bn(g.N, &n_tiles);
    dim3 grid((unsigned)cdiv(g.M, TBM), (unsigned)n_tiles);
    gemm_nt_tc_kernel<<<grid, kThreads, NT_SMEM, st>>>(g, bn, tmem_cols_for(bn));
    PN_LAUNCH_CHECK("gemm_nt_tc");
    return PN_OK;
}

// second stage of the TN reduction (kernels in gemm.cu)
int tn_finish(const GemmTN& g, const float* partial, const float* cpartial, int splits, cudaStream_t st);
int64_t tn_rows_per_split_tc(int64_t M) {
    int64_t rps = align_up(cdiv(M > 0 ? M : 1, 128), TBK);
    return rps < 512 ? 512 : rps;
}

int gemm_tn_tc(const GemmTN& g, float* scratch, cudaStream_t st) {
    PN_CHECK_ARG(g.lda % 4 == 0 && g.ldb % 4 == 0 && g.N % 4 == 0 && g.Kb % 4 == 0, "gemm_tn_tc: operands must be padded");
    PN_CHECK_ARG(((uintptr_t)g.A & 15) == 0 && ((uintptr_t)g.B & 15) == 0, "gemm_tn_tc: operands must be 16-byte aligned");
    static bool attr_set = false;
    if (!attr_set) {
        PN_CUDA(cudaFuncSetAttribute(gemm_tn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, NT_SMEM));
        attr_set = true;
    }
    int64_t rps = tn_rows_per_split_tc(g.M);
    int splits = (int)cdiv(g.M > 0 ? g.M : 1, rps);
    float* partial = scratch;
    float* cpartial = scratch + (int64_t)splits * g.N * g.Kb;
    int k_tiles;
    int bkt = pick_bn(g.Kb, &k_tiles);
    dim3 grid((unsigned)k_tiles, (unsigned)cdiv(g.N, TBM), (unsigned)splits);
    gemm_tn_tc_kernel<<<grid, kThreads, NT_SMEM, st>>>(g, bkt, tmem_cols_for(bkt), rps, partial, g.colsum ? cpartial : nullptr);
    PN_LAUNCH_CHECK("gemm_tn_tc");
    return tn_finish(g, partial, g.colsum ? cpartial : nullptr, splits, st);
}
int64_t gemm_tn_tc_scratch_floats(int64_t M, int N, int Kb) {
    int64_t splits = cdiv(M > 0 ? M : 1, tn_rows_per_split_tc(M));
    return splits * ((int64_t)N * Kb + N) + 64;
}

}  // namespace pn
