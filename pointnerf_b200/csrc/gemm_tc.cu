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
#include <cuda.h>

#include "gemm.cuh"
#include "tc.cuh"

namespace pn {

namespace {
constexpr int TBM = 128;          // rows of A per CTA = TMEM lanes
constexpr int TBK = 32;           // tf32 elements per 128-byte swizzle row
constexpr int TMAXN = 256;
constexpr int kLoadThreads = 256;
constexpr int A_BYTES = TBM * 128;              // 16 KB per (hi | lo)
constexpr int B_BYTES = TMAXN * 128;            // 32 KB per (hi | lo)
constexpr int STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;   // 96 KB

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
__device__ __forceinline__ void split_store_s(uint32_t hi_addr, uint32_t lo_addr, float4 v) {
    float4 h, l;
    tc::split_tf32(v.x, h.x, l.x);
    tc::split_tf32(v.y, h.y, l.y);
    tc::split_tf32(v.z, h.z, l.z);
    tc::split_tf32(v.w, h.w, l.w);
    tc::sts128(hi_addr, h);
    tc::sts128(lo_addr, l);
}
// Epilogue of a 128 x BN accumulator tile.  TMEM hands every thread ONE ROW (lane = row), which would make global stores
// 32 rows wide per instruction; the tile is therefore staged through shared memory (operand stages are free once the
// accumulator is complete) and leaves as whole 512-byte row segments: emit(row, col, float4) with consecutive threads
// on consecutive columns.  t in [0, 256): the eight staging warps.  stage: >= 128 * (BN + 4) floats.
template <class Emit>
__device__ __forceinline__ void epilogue_via_smem(uint32_t tmem_d, int BN, float* stage, int t, bool have_acc, Emit emit) {
    const int warp = t >> 5, lane = t & 31, q = warp & 3, half = warp >> 2;
    const int ld = BN + 4;
    const uint32_t sbase = tc::smem_u32(stage);
    const uint32_t myrow = sbase + (uint32_t)((q * 32 + lane) * ld) * 4u;
    for (int j = half; j < (BN >> 4); j += 2) {
        float v[16];
        if (have_acc) {
            tc::tmem_ld16(tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)(j * 16), v);
        } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = 0.f;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
            tc::sts128(myrow + (uint32_t)(j * 16 + 4 * i) * 4u, make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]));
    }
    tc::fence_before_sync();
    asm volatile("bar.sync 1, 256;" ::: "memory");
    const int c4n = BN >> 2;
    for (int idx = t; idx < TBM * c4n; idx += 256) {
        int row = idx / c4n, c4 = idx - row * c4n;
        emit(row, 4 * c4, tc::lds128(sbase + (uint32_t)(row * ld + 4 * c4) * 4u));
    }
}
}  // namespace

// ---------------------------------------------------------------------------------------------------------------
// TN: C[n, k] = sum_m A[m, n] * B[m, k].  The reduction index m is the slow dimension of both operands in global
// memory, i.e. both are MN-major for the tensor core: a chunk of 32 reduction rows is staged as 128-byte panels
// (32 consecutive n or k per row), 4 reduction rows per 32-byte-base swizzle atom (tc.cuh).  grid = (k tiles, n tiles of 128, splits of m).
// ---------------------------------------------------------------------------------------------------------------
// BKt <= 288 columns of B per CTA: above 256 the tile is issued as two MMAs per K-step (N1 = 160 and N2 = BKt - 160 columns,
// both panel-aligned) into adjacent TMEM columns, so that A (the dPre tile) is staged ONCE for the whole weight row.
// Split geometry of the TN reduction, a function of the row count M and the number of split slots the launch provides:
// every slot reduces a contiguous slice of `rps` rows (a multiple of the 32-row chunk, at least 256); only the first
// cdiv(M, rps) slots hold work.  Evaluated on the DEVICE when the row count lives there (GemmTN::Mdev).
__host__ __device__ __forceinline__ int64_t tn_rows_per_split(int64_t M, int slots) {
    int64_t r = ((M > 0 ? M : 1) + slots - 1) / slots;
    r = (r + TBK - 1) / TBK * TBK;
    return r < 256 ? 256 : r;
}
__device__ __forceinline__ int64_t tn_rows(const GemmTN& g) {
    if (!g.Mdev) return g.M;
    const int64_t md = (int64_t)__ldg(g.Mdev);
    return md < g.M ? md : g.M;
}
constexpr int kTnLoad = 512;                 // staging threads of the TN kernel (16 warps) + 1 MMA warp
constexpr int kTnThreads = kTnLoad + 32;
__global__ void __launch_bounds__(kTnThreads, 1) gemm_tn_tc_kernel(GemmTN g, int BKt, int tmem_cols,
                                                                   float* __restrict__ partial, float* __restrict__ cpartial) {
    extern __shared__ char smem_raw[];
    char* smem = (char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    constexpr int PANEL = TBK * 128;                     // bytes of one 32-wide panel over a 32-row chunk
    const int npan = (BKt + 31) >> 5;                    // B panels per stage
    const int b_bytes = npan * PANEL;                    // per (hi | lo)
    const int stage_bytes = 2 * A_BYTES + 2 * b_bytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * stage_bytes);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int j0 = blockIdx.x * BKt;       // first column of B (k of the weight)
    const int i0 = blockIdx.y * TBM;       // first column of A (n of the weight)
    g.M = tn_rows(g);
    const int64_t rows_per_split = tn_rows_per_split(g.M, (int)gridDim.z);
    const int64_t r_begin = (int64_t)blockIdx.z * rows_per_split;
    const int64_t r_end = r_begin + rows_per_split < g.M ? r_begin + rows_per_split : g.M;
    const int nk = r_end > r_begin ? (int)((r_end - r_begin + TBK - 1) / TBK) : 0;
    if (nk == 0 && blockIdx.z > 0) return;      // an empty split slot (the finish reads only the first cdiv(M, rps) slots)

    if (t == 0) {
        tc::mbar_init(&bars[0], kTnLoad / 64);
        tc::mbar_init(&bars[1], kTnLoad / 64);
        tc::mbar_init(&bars[2], 1);
        tc::mbar_init(&bars[3], 1);
        tc::mbar_init(&bars[4], 1);
        tc::mbar_fence_init();
    }
    if (warp == 16) tc::tmem_alloc(tmem_slot, (uint32_t)tmem_cols);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_d = *tmem_slot;

    if (warp < 16) {
        // Two groups of eight warps alternate the 32-row chunks (group 0: even, group 1: odd): a thread reloads its ONE
        // register set right after its arrive, so the loads have two chunk periods to land before the thread's next
        // fence.proxy.async (whose membar waits for them) and before they are needed.
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        const int grp = t >> 8, tt = t & 255;
        const int ac4 = tt & 31, ar0 = tt >> 5;    // A: 32 float4 per row (128 columns), rows ar0 + 8 i
        const bool a_ok = i0 + 4 * ac4 < g.N;
        const int ncb = BKt >> 2;                  // float4 per B row (<= 72)
        constexpr int NB = 9;                      // 256 threads x 9 >= 32 rows x 72 float4
        int brow[NB], bc4[NB];
#pragma unroll
        for (int i = 0; i < NB; ++i) {
            int idx = tt + 256 * i;
            const int br = idx / ncb;
            bc4[i] = idx - br * ncb;
            // -1: no shared-memory slot; bit 8: a slot (zero-filled) with no global column behind it
            brow[i] = br >= TBK ? -1 : (br | (j0 + 4 * bc4[i] >= g.Kb ? 0x100 : 0));
        }
        float4 ra[4], rb[NB];
        float4 csum = z4;
        const uint32_t smem_u = tc::smem_u32(smem);
        auto load = [&](int kc) {
            const int64_t rb0 = r_begin + (int64_t)kc * TBK;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                int64_t rr = rb0 + ar0 + 8 * i;
                ra[i] = (a_ok && rr < r_end) ? __ldg(reinterpret_cast<const float4*>(g.A + rr * g.lda + i0 + 4 * ac4)) : z4;
            }
#pragma unroll
            for (int i = 0; i < NB; ++i) {
                int64_t rr = rb0 + (brow[i] & 0xFF);
                rb[i] = (brow[i] >= 0 && !(brow[i] & 0x100) && rr < r_end) ? __ldg(reinterpret_cast<const float4*>(g.B + rr * g.ldb + j0 + 4 * bc4[i])) : z4;
            }
        };
        if (grp < nk) load(grp);
        for (int kc = grp; kc < nk; kc += 2) {
            const int s = kc & 1;
            tc::mbar_wait(&bars[2 + s], ((kc >> 1) & 1) ^ 1);
            const uint32_t st = smem_u + (uint32_t)(s * stage_bytes);
            {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                int row = ar0 + 8 * i;
                const uint32_t off = (uint32_t)((ac4 >> 3) * PANEL) + sw128b32_off(row, ac4 & 7);
                split_store_s(st + off, st + A_BYTES + off, ra[i]);
                csum.x += ra[i].x; csum.y += ra[i].y; csum.z += ra[i].z; csum.w += ra[i].w;
            }
#pragma unroll
            for (int i = 0; i < NB; ++i)
                if (brow[i] >= 0) {
                    const uint32_t off = (uint32_t)((bc4[i] >> 3) * PANEL) + sw128b32_off(brow[i] & 0xFF, bc4[i] & 7);
                    split_store_s(st + 2 * A_BYTES + off, st + 2 * A_BYTES + (uint32_t)b_bytes + off, rb[i]);
                }
            }
            tc::fence_proxy_async();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&bars[s]);
            if (kc + 2 < nk) load(kc + 2);
        }
      if (warp < 8) {
        // ---- epilogue: D[n, k] -> partial[split][n][k]
        if (nk > 0) {
            tc::mbar_wait(&bars[4], 0);
            tc::fence_after_sync();
        }
        float* out = partial + (int64_t)blockIdx.z * g.N * g.Kb;
        epilogue_via_smem(tmem_d, BKt, reinterpret_cast<float*>(smem), t, nk > 0, [&](int row, int col, float4 v) {
            const int n = i0 + row, k = j0 + col;
            if (n >= g.N || k >= g.Kb) return;
            *reinterpret_cast<float4*>(out + (int64_t)n * g.Kb + k) = v;     // Kb % 4 == 0, partial 256-byte aligned
        });
        tc::fence_before_sync();
      }
        // ---- bias gradient: column sums of A over this split (k tile 0 only), fixed order
        if (cpartial && blockIdx.x == 0) {
            float4* red = reinterpret_cast<float4*>(smem);      // stage memory is free once the tile has left
            __syncwarp();
            asm volatile("bar.sync 2, %0;" ::"n"(kTnLoad));
            red[t] = csum;
            asm volatile("bar.sync 2, %0;" ::"n"(kTnLoad));
            if (t < 32) {
                float4 a = red[t];
#pragma unroll
                for (int w16 = 1; w16 < 16; ++w16) {
                    float4 b = red[t + 32 * w16];
                    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
                }
                int nn = i0 + 4 * t;
                if (nn < g.N) *reinterpret_cast<float4*>(cpartial + (int64_t)blockIdx.z * g.N + nn) = a;
            }
        }
    } else {
      if (lane == 0) {
        const int N1 = BKt > 256 ? 160 : BKt, N2 = BKt - N1;
        const uint32_t idesc1 = tc::idesc_tf32(TBM, N1, 1, 1);
        const uint32_t idesc2 = N2 > 0 ? tc::idesc_tf32(TBM, N2, 1, 1) : 0u;
        const uint32_t sbase = tc::smem_u32(smem);
        const uint32_t off2 = (uint32_t)((N1 >> 5) * PANEL);      // second MMA starts at a panel boundary of B
        for (int kc = 0; kc < nk; ++kc) {
            const int s = kc & 1;
            tc::mbar_wait(&bars[s], (kc >> 1) & 1);
            tc::fence_after_sync();
            const uint32_t a_hi = sbase + s * stage_bytes, a_lo = a_hi + A_BYTES;
            const uint32_t b_hi = a_hi + 2 * A_BYTES, b_lo = b_hi + (uint32_t)b_bytes;
            int64_t rows = r_end - (r_begin + (int64_t)kc * TBK);
            int ksteps = rows >= TBK ? 4 : (int)((rows + 7) >> 3);
            for (int k = 0; k < ksteps; ++k) {
                const uint32_t ko = 1024u * k;
                uint64_t dah = tc::smem_desc_sw128(a_hi + ko, PANEL, 512, 1), dal = tc::smem_desc_sw128(a_lo + ko, PANEL, 512, 1);
                uint64_t dbh = tc::smem_desc_sw128(b_hi + ko, PANEL, 512, 1), dbl = tc::smem_desc_sw128(b_lo + ko, PANEL, 512, 1);
                tc::mma_tf32(tmem_d, dah, dbl, idesc1, (kc | k) != 0);
                tc::mma_tf32(tmem_d, dal, dbh, idesc1, 1);
                tc::mma_tf32(tmem_d, dah, dbh, idesc1, 1);
                if (N2 > 0) {
                    uint64_t dbh2 = tc::smem_desc_sw128(b_hi + off2 + ko, PANEL, 512, 1), dbl2 = tc::smem_desc_sw128(b_lo + off2 + ko, PANEL, 512, 1);
                    tc::mma_tf32(tmem_d + (uint32_t)N1, dah, dbl2, idesc2, (kc | k) != 0);
                    tc::mma_tf32(tmem_d + (uint32_t)N1, dal, dbh2, idesc2, 1);
                    tc::mma_tf32(tmem_d + (uint32_t)N1, dah, dbh2, idesc2, 1);
                }
            }
            tc::mma_commit(&bars[2 + s]);
        }
        tc::mma_commit(&bars[4]);
      }
      __syncwarp();
    }
    __syncthreads();
    if (warp == 16) {
        tc::fence_after_sync();
        tc::tmem_dealloc(tmem_d, (uint32_t)tmem_cols);
    }
}

// Constants shared by the persistent NT kernel below.
namespace {
constexpr int A_STAGE = 2 * A_BYTES;                 // 32 KB (hi | lo)
constexpr int B_STAGE = 2 * B_BYTES;                 // 64 KB (hi | lo) at BN = 256
constexpr int kEpiWarps = 8;                         // two per TMEM lane quadrant, alternating 32-column chunks
constexpr int SLAB_BYTES = 32 * 32 * 4;              // per epilogue warp: 32 rows x 32 floats, 16-byte chunks XOR-swizzled
constexpr int kThreadsImg = kLoadThreads + 64 + 32 * kEpiWarps;
}  // namespace

// ---------------------------------------------------------------------------------------------------------------
// NT on CTA PAIRS (cta_group::2).  The two CTAs of a cluster sit on the two SMs of one TPC and compute one 256 x BN tile:
// each stages ITS 128 rows of A and loads HALF of the weight block (BN / 2 rows); the leader (rank 0) issues one MMA of
// M = 256 that reads both halves.  Per SM this halves the weight bytes copied into and read from shared memory - the
// resource that bounds the single-CTA kernel - and leaves room for a third A / B stage.
//   fullA / fullB   per CTA, local producers          emptyA / emptyB / accFull   per CTA, multicast tcgen05.commit of the leader
//   peerFull        on the leader: the peer's relay thread (its otherwise idle MMA warp) forwards "my stage s is full"
//   accEmpty        on the leader: 8 local + 8 remote epilogue warps
// ---------------------------------------------------------------------------------------------------------------
namespace {
constexpr int NS2 = 3;
constexpr int B2_STAGE = B_BYTES;                    // 2 x (128 rows x 128 B): half of a 256-row block, hi | lo
constexpr int IMG2_SMEM = NS2 * A_STAGE + NS2 * B2_STAGE + kEpiWarps * SLAB_BYTES + 1024 + 256;
}  // namespace

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreadsImg, 1)
    gemm_nt_img2_kernel(GemmNT g, int BN, int n_tiles_n, const __grid_constant__ CUtensorMap tmC, int tma_store) {
    extern __shared__ char smem_raw[];
    char* smem = (char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    char* smA = smem;
    char* smB = smem + NS2 * A_STAGE;
    char* slabs = smB + NS2 * B2_STAGE;
    uint64_t* bars = reinterpret_cast<uint64_t*>(slabs + kEpiWarps * SLAB_BYTES);
    uint64_t *fullA = bars, *emptyA = bars + 3, *fullB = bars + 6, *emptyB = bars + 9, *peerFull = bars + 12, *accFull = bars + 15,
             *accEmpty = bars + 17;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 20);
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const uint32_t rank = tc::cluster_ctarank();
    const int pair = (int)blockIdx.x >> 1, n_pairs = (int)gridDim.x >> 1;
    const int nk = (g.Kd + TBK - 1) / TBK;
    // rows of A / C: a host value, or read from device memory (g.Mdev: the valid-pair / valid-sample count of the batch, which
    // the host never learns - g.M is then only the capacity the launch was sized for)
    if (g.Mdev) { const int64_t md = (int64_t)__ldg(g.Mdev); g.M = md < g.M ? md : g.M; }
    const int n_ptiles_total = (int)((g.M + 2 * TBM - 1) / (2 * TBM)) * n_tiles_n;
    const uint32_t b_block = (uint32_t)(2 * BN * 128);    // image block of one (n tile, K chunk)
    // Work items.  Tile i of the (row tile, n tile) grid goes to pair i % n_pairs; the tiles of the LAST, partial wave (rem of
    // them, the other pairs would idle for a whole tile time) are cut along n into ns column slices of BN / ns, one per pair:
    // the tail then costs 1 / ns of a tile.  At S = 19 K sample rows (76 row tiles on 74 pairs) that is 1.25 tile times
    // instead of 2; at M = 135 K pair rows 7.25 instead of 8.
    const int full_waves = n_ptiles_total / n_pairs, rem = n_ptiles_total - full_waves * n_pairs;
    int ns = 1;
    if (n_tiles_n == 1 && rem > 0) {
        if (rem * 4 <= n_pairs && BN % 128 == 0) ns = 4;
        else if (rem * 2 <= n_pairs && BN % 64 == 0) ns = 2;
    }
    const int my_tiles = full_waves + (pair < rem * ns ? 1 : 0);
    const int total_chunks = my_tiles * nk;
    struct Item { int64_t m0; int bn, n0, nt, img_row; };   // rows m0.. of this pair | slice width | first column | n tile | first image row
    auto item = [&](int it) {
        Item w;
        if (it < full_waves) {
            const int ptile = pair + it * n_pairs;
            w.m0 = (int64_t)(ptile / n_tiles_n) * (2 * TBM); w.nt = ptile % n_tiles_n; w.bn = BN; w.n0 = w.nt * BN; w.img_row = 0;
        } else {
            const int ptile = full_waves * n_pairs + pair / ns, sl = pair % ns;
            w.m0 = (int64_t)(ptile / n_tiles_n) * (2 * TBM); w.nt = ptile % n_tiles_n; w.bn = BN / ns; w.img_row = sl * w.bn;
            w.n0 = w.nt * BN + w.img_row;
        }
        return w;
    };

    if (t == 0) {
        for (int i = 0; i < NS2; ++i) {
            tc::mbar_init(&fullA[i], kLoadThreads / 64); tc::mbar_init(&emptyA[i], 1);
            tc::mbar_init(&fullB[i], 1); tc::mbar_init(&emptyB[i], 1);
            tc::mbar_init(&peerFull[i], 1);
        }
        for (int i = 0; i < 2; ++i) { tc::mbar_init(&accFull[i], 1); tc::mbar_init(&accEmpty[i], 2 * kEpiWarps); }
        tc::mbar_fence_init();
    }
    if (warp == 8) tc::tmem_alloc2(tmem_slot, 512u);
    tc::fence_before_sync();
    __syncthreads();
    tc::cluster_sync();                                   // both CTAs' barriers exist before any remote arrive / multicast commit
    tc::fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp < 8) {
        // ---------------------------------------------------------------- A staging (as the single-CTA kernel; own 128 rows)
        const int grp = warp >> 2, tt = t & 127;
        const int c = tt & 7, r0 = tt >> 3;
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        const uint32_t sA_u32 = tc::smem_u32(smA);
        float4 ra[8], rb[8];
        auto load = [&](int L, float4 (&r)[8]) {
            const int it = L / nk, kc = L - it * nk;
            const int64_t m0 = item(it).m0 + (int64_t)rank * TBM;
            const int gk = kc * TBK + 4 * c;
            const bool kok = gk < g.Kd;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                int64_t m = m0 + r0 + 16 * i;
                r[i] = (kok && m < g.M) ? __ldg(reinterpret_cast<const float4*>(g.A + m * g.lda + gk)) : z4;
            }
        };
        auto step = [&](int L, float4 (&cur)[8]) {
            const int sa = L % NS2;
            tc::mbar_wait(&emptyA[sa], ((L / NS2) & 1) ^ 1);
            const uint32_t st = sA_u32 + (uint32_t)(sa * A_STAGE);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const uint32_t off = sw128_off(r0 + 16 * i, c);
                split_store_s(st + off, st + A_BYTES + off, cur[i]);
            }
            tc::fence_proxy_async();
            __syncwarp();                                   // one arrival per warp: arrivals on one barrier word serialise
            if (lane == 0) tc::mbar_arrive(&fullA[sa]);
            if (L + 4 < total_chunks) load(L + 4, cur);
        };
        if (grp < total_chunks) load(grp, ra);
        if (grp + 2 < total_chunks) load(grp + 2, rb);
        for (int L = grp; L < total_chunks; L += 4) {
            step(L, ra);
            if (L + 2 < total_chunks) step(L + 2, rb);
        }
    } else if (warp == 8) {
        if (lane == 0 && rank == 0) {
            // ---------------------------------------------------------------- MMA issue (leader)
            const uint32_t sA = tc::smem_u32(smA), sB = tc::smem_u32(smB);
            int L = 0;
            for (int it = 0; it < my_tiles; ++it) {
                const int ab = it & 1;
                tc::mbar_wait_cluster(&accEmpty[ab], ((it >> 1) & 1) ^ 1);
                tc::fence_after_sync();
                const uint32_t tmem_d = tmem_base + (uint32_t)(ab * 256);
                const int bn = item(it).bn;
                const uint32_t idesc = tc::idesc_tf32(2 * TBM, bn, 0, 0);
                const uint32_t b_half = (uint32_t)(bn * 64);  // bytes of (bn / 2 rows x 128 B): one of hi | lo in this CTA
                for (int kc = 0; kc < nk; ++kc, ++L) {
                    const int s = L % NS2;
                    const uint32_t ph = (uint32_t)((L / NS2) & 1);
                    tc::mbar_wait(&fullB[s], ph);
                    tc::mbar_wait(&fullA[s], ph);
                    tc::mbar_wait_cluster(&peerFull[s], ph);
                    tc::fence_after_sync();
                    const uint32_t a_hi = sA + s * A_STAGE, a_lo = a_hi + A_BYTES;
                    const uint32_t b_hi = sB + s * B2_STAGE, b_lo = b_hi + b_half;
                    int ksteps = (g.Kd - kc * TBK + 7) >> 3;
                    if (ksteps > 4) ksteps = 4;
                    for (int k = 0; k < ksteps; ++k) {
                        const uint32_t ko = 32u * k;
                        uint64_t dah = tc::smem_desc_sw128(a_hi + ko, 16, 1024), dal = tc::smem_desc_sw128(a_lo + ko, 16, 1024);
                        uint64_t dbh = tc::smem_desc_sw128(b_hi + ko, 16, 1024), dbl = tc::smem_desc_sw128(b_lo + ko, 16, 1024);
                        tc::mma_tf32_2(tmem_d, dah, dbl, idesc, (kc | k) != 0);
                        tc::mma_tf32_2(tmem_d, dal, dbh, idesc, 1);
                        tc::mma_tf32_2(tmem_d, dah, dbh, idesc, 1);
                    }
                    tc::mma_commit_2(&emptyA[s]);
                    tc::mma_commit_2(&emptyB[s]);
                }
                tc::mma_commit_2(&accFull[ab]);
            }
        } else if (lane == 0) {
            // ---------------------------------------------------------------- relay (peer): forward "stage full" to the leader
            for (int L = 0; L < total_chunks; ++L) {
                const int s = L % NS2;
                const uint32_t ph = (uint32_t)((L / NS2) & 1);
                tc::mbar_wait(&fullB[s], ph);
                tc::mbar_wait(&fullA[s], ph);
                tc::mbar_arrive_cluster(tc::mapa(tc::smem_u32(&peerFull[s]), 0));
            }
        }
        __syncwarp();
    } else if (warp == 9) {
        // ---------------------------------------------------------------- B image producer: this CTA's half of every block
        if (lane == 0) {
            int L = 0;
            for (int it = 0; it < my_tiles; ++it) {
                const Item w = item(it);
                const uint32_t b_half = (uint32_t)(w.bn * 64);
                const char* src = g.Bimg + (int64_t)w.nt * nk * b_block + (int64_t)w.img_row * 128 + (int64_t)rank * b_half;
                for (int kc = 0; kc < nk; ++kc, ++L) {
                    const int s = L % NS2;
                    tc::mbar_wait(&emptyB[s], ((L / NS2) & 1) ^ 1);
                    tc::mbar_arrive_expect_tx(&fullB[s], 2 * b_half);
                    const char* blk = src + (int64_t)kc * b_block;
                    tc::bulk_g2s(smB + s * B2_STAGE, blk, b_half, &fullB[s]);                               // hi rows of this CTA
                    tc::bulk_g2s(smB + s * B2_STAGE + b_half, blk + (int64_t)BN * 128, b_half, &fullB[s]);  // lo rows
                }
            }
        }
        __syncwarp();
    } else {
        // ---------------------------------------------------------------- epilogue (warps 10..17), own 128 rows
        const int q = warp & 3;
        const int half = (warp - 10) >> 2;
        const uint32_t slab = tc::smem_u32(slabs + (warp - 10) * SLAB_BYTES);
        const bool vec_ok = (g.ldc % 4 == 0) && ((((uintptr_t)g.C) & 15) == 0) && (g.N % 4 == 0) &&
                            (!g.aux || (g.ldaux % 4 == 0 && (((uintptr_t)g.aux) & 15) == 0 && g.aux_cols % 4 == 0)) &&
                            (!g.bias || (((uintptr_t)g.bias) & 15) == 0);
        const int rsub = lane >> 3, c8 = lane & 7;
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        const uint32_t accEmpty0 = tc::mapa(tc::smem_u32(&accEmpty[0]), 0);     // the leader's barriers
        auto release = [&](int ab) {
            tc::fence_before_sync();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive_cluster(accEmpty0 + 8u * (uint32_t)ab);
        };
        auto prefetch_aux = [&](int it) {
            if (!g.aux) return;
            const Item w = item(it);
            const int64_t m = w.m0 + (int64_t)rank * TBM + q * 32 + lane;
            const int n0 = w.n0, nch = (w.bn + 31) >> 5;
            if (m >= g.M) return;
            for (int j = half; j < nch; j += 2) {
                const int n = n0 + j * 32;
                if (n < g.aux_cols) asm volatile("prefetch.global.L2 [%0];" ::"l"(g.aux + m * g.ldaux + n));
            }
        };
        if (my_tiles > 0) prefetch_aux(0);
        for (int it = 0; it < my_tiles; ++it) {
            const int ab = it & 1;
            const Item w = item(it);
            const int BN = w.bn;                              // this item's width (a column slice in the last wave)
            const int64_t m0 = w.m0 + (int64_t)rank * TBM + q * 32;
            const int n0 = w.n0;
            const int nch = (BN + 31) >> 5;
            const int my_last = ((nch - 1 - half) & ~1) + half;
            if (it + 1 < my_tiles) prefetch_aux(it + 1);
            tc::mbar_wait(&accFull[ab], (it >> 1) & 1);
            tc::fence_after_sync();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(ab * 256);
            // does this warp's 32-row group lie inside the valid rows?  (then whole 32 x 32 chunks leave as ONE TMA store)
            const bool rows_full = tma_store && vec_ok && m0 + 32 <= g.M;
            for (int j = half; j < nch; j += 2) {
                uint32_t lo16[16], hi16[16];
                tc::tmem_ld16_issue(taddr + (uint32_t)(j * 32), lo16);
                tc::tmem_ld16_issue(taddr + (uint32_t)(j * 32 + 16), hi16);
                const int n = n0 + j * 32 + 4 * c8;
                const bool ncol_ok = (j * 32 + 4 * c8 < BN) && n < g.N;
                const bool by_tma = rows_full && (!g.aux || g.sign_in) && j * 32 + 32 <= BN && n0 + j * 32 + 32 <= g.N;
                float4 bias4 = z4;
                if (!by_tma && vec_ok && ncol_ok && g.bias) bias4 = __ldg(reinterpret_cast<const float4*>(g.bias + n));
                // the slab is free once the previous chunk's bulk store has READ it
                if (tma_store) {
                    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                    __syncwarp();
                }
                tc::tmem_ld_wait();
                if (j == my_last) release(ab);
                if (by_tma) {
                    // Whole sub-tile inside the matrix, no activation mask: bias + LeakyReLU in the TMEM layout (a thread owns one
                    // row; the 32 bias values are warp-uniform loads, four floats at a time so that nothing else stays live), ONE
                    // swizzled store pass, and the 32 x 32 box leaves through TMA: 8 + 8 LSU instructions per sub-tile instead of
                    // 8 STS + 8 LDS + 8 STG - the LSU queue is what the latency-bound roles of this kernel wait in
                    const float* bj = g.bias ? g.bias + n0 + j * 32 : nullptr;
                    const int64_t mrow = m0 + lane;
                    // LeakyReLU mask of a data gradient: the 32 sign bits of this row's columns (two words when the sub-tile
                    // does not start on a word; columns past the masked ones read as positive)
                    uint32_t mbits = 0xffffffffu;
                    if (g.sign_in) {
                        const int c0 = n0 + j * 32, wi = c0 >> 5;
                        const uint32_t w0 = wi < g.sign_ld ? __ldg(g.sign_in + mrow * g.sign_ld + wi) : 0xffffffffu;
                        const uint32_t w1 = ((c0 & 31) && wi + 1 < g.sign_ld) ? __ldg(g.sign_in + mrow * g.sign_ld + wi + 1) : 0xffffffffu;
                        mbits = __funnelshift_r(w0, w1, c0 & 31);
                    }
                    uint32_t obits = 0;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        float4 b0 = z4, b1 = z4;
                        if (bj) { b0 = __ldg(reinterpret_cast<const float4*>(bj + 4 * i)); b1 = __ldg(reinterpret_cast<const float4*>(bj + 16 + 4 * i)); }
                        float4 y = make_float4(__uint_as_float(lo16[4 * i]) + b0.x, __uint_as_float(lo16[4 * i + 1]) + b0.y,
                                               __uint_as_float(lo16[4 * i + 2]) + b0.z, __uint_as_float(lo16[4 * i + 3]) + b0.w);
                        float4 z = make_float4(__uint_as_float(hi16[4 * i]) + b1.x, __uint_as_float(hi16[4 * i + 1]) + b1.y,
                                               __uint_as_float(hi16[4 * i + 2]) + b1.z, __uint_as_float(hi16[4 * i + 3]) + b1.w);
                        if (g.leaky) {
                            y.x = y.x > 0.f ? y.x : g.slope * y.x; y.y = y.y > 0.f ? y.y : g.slope * y.y;
                            y.z = y.z > 0.f ? y.z : g.slope * y.z; y.w = y.w > 0.f ? y.w : g.slope * y.w;
                            z.x = z.x > 0.f ? z.x : g.slope * z.x; z.y = z.y > 0.f ? z.y : g.slope * z.y;
                            z.z = z.z > 0.f ? z.z : g.slope * z.z; z.w = z.w > 0.f ? z.w : g.slope * z.w;
                        }
                        if (g.sign_in) {
                            const uint32_t a = mbits >> (4 * i), b = mbits >> (16 + 4 * i);
                            y.x *= (a & 1u) ? 1.f : g.slope; y.y *= (a & 2u) ? 1.f : g.slope; y.z *= (a & 4u) ? 1.f : g.slope; y.w *= (a & 8u) ? 1.f : g.slope;
                            z.x *= (b & 1u) ? 1.f : g.slope; z.y *= (b & 2u) ? 1.f : g.slope; z.z *= (b & 4u) ? 1.f : g.slope; z.w *= (b & 8u) ? 1.f : g.slope;
                        }
                        if (g.sign_out) {
                            obits |= ((y.x > 0.f ? 1u : 0u) | (y.y > 0.f ? 2u : 0u) | (y.z > 0.f ? 4u : 0u) | (y.w > 0.f ? 8u : 0u)) << (4 * i);
                            obits |= ((z.x > 0.f ? 1u : 0u) | (z.y > 0.f ? 2u : 0u) | (z.z > 0.f ? 4u : 0u) | (z.w > 0.f ? 8u : 0u)) << (16 + 4 * i);
                        }
                        tc::sts128(slab + (uint32_t)(lane * 32 + 4 * (i ^ (lane & 7))) * 4u, y);
                        tc::sts128(slab + (uint32_t)(lane * 32 + 4 * ((i + 4) ^ (lane & 7))) * 4u, z);
                    }
                    if (g.sign_out) g.sign_out[mrow * g.sign_ld + ((n0 + j * 32) >> 5)] = obits;
                    tc::fence_proxy_async();
                    __syncwarp();
                    if (lane == 0) {
                        asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(reinterpret_cast<uint64_t>(&tmC)),
                                     "r"(n0 + j * 32), "r"((int)m0), "r"(slab)
                                     : "memory");
                        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                    }
                    continue;
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    tc::sts128(slab + (uint32_t)(lane * 32 + 4 * (i ^ (lane & 7))) * 4u,
                               make_float4(__uint_as_float(lo16[4 * i]), __uint_as_float(lo16[4 * i + 1]), __uint_as_float(lo16[4 * i + 2]),
                                           __uint_as_float(lo16[4 * i + 3])));
                    tc::sts128(slab + (uint32_t)(lane * 32 + 4 * ((i + 4) ^ (lane & 7))) * 4u,
                               make_float4(__uint_as_float(hi16[4 * i]), __uint_as_float(hi16[4 * i + 1]), __uint_as_float(hi16[4 * i + 2]),
                                           __uint_as_float(hi16[4 * i + 3])));
                }
                __syncwarp();
                const bool use_aux = vec_ok && ncol_ok && g.aux && n < g.aux_cols;
                float4 aux4[8];
                if (use_aux) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int64_t m = m0 + 4 * i + rsub;
                        aux4[i] = m < g.M ? __ldg(reinterpret_cast<const float4*>(g.aux + m * g.ldaux + n)) : z4;
                    }
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int row = 4 * i + rsub;
                    const int64_t m = m0 + row;
                    const uint32_t saddr = slab + (uint32_t)(row * 32 + 4 * (c8 ^ (row & 7))) * 4u;
                    float4 x4 = tc::lds128(saddr);
                    if (m >= g.M || !ncol_ok) continue;
                    float x[4] = {x4.x, x4.y, x4.z, x4.w};
                    if (vec_ok) {
                        x[0] += bias4.x; x[1] += bias4.y; x[2] += bias4.z; x[3] += bias4.w;
                        if (g.leaky) {
#pragma unroll
                            for (int e = 0; e < 4; ++e) x[e] = x[e] > 0.f ? x[e] : g.slope * x[e];
                        }
                        if (use_aux) {
                            const float4 a = aux4[i];
                            x[0] *= a.x > 0.f ? 1.f : g.slope; x[1] *= a.y > 0.f ? 1.f : g.slope;
                            x[2] *= a.z > 0.f ? 1.f : g.slope; x[3] *= a.w > 0.f ? 1.f : g.slope;
                        }
                        *reinterpret_cast<float4*>(g.C + m * g.ldc + n) = make_float4(x[0], x[1], x[2], x[3]);
                    } else {
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            if (n + e >= g.N) break;
                            float y = x[e];
                            if (g.bias) y += __ldg(g.bias + n + e);
                            if (g.leaky) y = y > 0.f ? y : g.slope * y;
                            if (g.aux && n + e < g.aux_cols) y *= (__ldg(g.aux + m * g.ldaux + n + e) > 0.f) ? 1.f : g.slope;
                            g.C[m * g.ldc + n + e] = y;
                        }
                    }
                }
                __syncwarp();
            }
            if (my_last < half) release(ab);
        }
        if (tma_store && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::cluster_sync();                                   // the peer's smem / TMEM / barriers stay alive until the leader is done
    if (warp == 8) {
        tc::fence_after_sync();
        tc::tmem_dealloc2(tmem_base, 512u);
    }
}

// image block (tile j, chunk kc) = [hi: BN rows x 128 B swizzled | lo: same]; one thread per (row, 16-byte chunk)
__global__ void __launch_bounds__(256) pack_nt_image_kernel(const float* __restrict__ src, int64_t s_n, int64_t s_k, int N, int K,
                                                            int BN, int n_tiles, int nk, char* __restrict__ img) {
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t total = (int64_t)n_tiles * nk * BN * 8;
    if (idx >= total) return;
    int c = (int)(idx & 7);
    int64_t rest = idx >> 3;
    int r = (int)(rest % BN);
    rest /= BN;
    int kc = (int)(rest % nk);
    int j = (int)(rest / nk);
    int n = j * BN + r;
    float v[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        int k = kc * TBK + 4 * c + e;
        v[e] = (n < N && k < K) ? src[(int64_t)n * s_n + (int64_t)k * s_k] : 0.f;
    }
    char* blk = img + ((int64_t)j * nk + kc) * (2 * BN * 128);
    split_store(blk, blk + BN * 128, sw128_off(r, c), make_float4(v[0], v[1], v[2], v[3]));
}

static int pick_bn(int N, int* n_tiles) {
    int nt = (N + TMAXN - 1) / TMAXN;
    int bn = (int)align_up(cdiv(N, nt), 16);
    *n_tiles = (int)cdiv(N, bn);
    return bn;
}
static int tmem_cols_for(int bn) { return bn <= 32 ? 32 : bn <= 64 ? 64 : bn <= 128 ? 128 : 256; }

// cuTensorMapEncodeTiled through the runtime's driver entry point: the library keeps no link-time dependency on libcuda
// (it must load, and answer its size queries, on a machine without a driver)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}

int gemm_nt(const GemmNT& g, cudaStream_t st) {
    PN_CHECK_ARG(g.Kd % 4 == 0 && g.lda % 4 == 0, "gemm_nt: operands must be padded to 4 (Kd %d lda %lld)", g.Kd, (long long)g.lda);
    PN_CHECK_ARG(((uintptr_t)g.A & 15) == 0 && g.Bimg, "gemm_nt: A must be 16-byte aligned and B given as an operand image");
    if (g.M <= 0 || g.N <= 0) return PN_OK;
    static bool attr_img2 = false;
    if (!attr_img2) {
        PN_CUDA(cudaFuncSetAttribute(gemm_nt_img2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, IMG2_SMEM));
        attr_img2 = true;
    }
    int n_tiles;
    int bn = pick_bn(g.N, &n_tiles);
    int64_t ptotal = cdiv(g.M, 2 * TBM) * n_tiles;        // g.M: the row count, or its capacity when g.Mdev is given
    // few tiles: the kernel cuts them into column slices (see its work-item comment), so up to 4 pairs share a tile
    const int max_slices = n_tiles == 1 ? (bn % 128 == 0 ? 4 : bn % 64 == 0 ? 2 : 1) : 1;
    int pairs = ptotal * max_slices < kNumSMs / 2 ? (int)(ptotal * max_slices) : kNumSMs / 2;
    // the C tile leaves through TMA (one 32 x 32 box per epilogue warp and chunk) when the layout allows a tensor map
    CUtensorMap tmC;
    memset(&tmC, 0, sizeof(tmC));
    int tma_store = 0;
    if (g.N >= 32 && g.ldc % 4 == 0 && (((uintptr_t)g.C) & 15) == 0 && g.M < (int64_t)1 << 31) {
        const cuuint64_t dims[2] = {(cuuint64_t)g.N, (cuuint64_t)g.M};
        const cuuint64_t strides[1] = {(cuuint64_t)g.ldc * 4};
        const cuuint32_t box[2] = {32, 32}, estr[2] = {1, 1};
        EncodeTiledFn enc = encode_tiled();
        if (enc) {
            CUresult r = enc(&tmC, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)g.C, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            PN_CHECK_ARG(r == CUDA_SUCCESS, "gemm_nt: cuTensorMapEncodeTiled failed (%d)", (int)r);
            tma_store = 1;
        }
    }
    // the sign-bit forms live on the TMA path only: writer and reader must both take it, or the bits are never written
    PN_CHECK_ARG(!(g.sign_out || g.sign_in) || (tma_store && g.sign_ld > 0 && (!g.sign_out || (g.N % 32 == 0 && bn % 32 == 0 && g.sign_ld * 32 >= g.N)) &&
                                                (!g.sign_in || (g.aux && g.aux_cols == g.sign_ld * 32)) && (!g.bias || (((uintptr_t)g.bias) & 15) == 0)),
                 "gemm_nt: sign bits need the TMA epilogue (N %d ldc %lld sign_ld %d)", g.N, (long long)g.ldc, g.sign_ld);
    gemm_nt_img2_kernel<<<2 * pairs, kThreadsImg, IMG2_SMEM, st>>>(g, bn, n_tiles, tmC, tma_store);
    PN_LAUNCH_CHECK("gemm_nt_img2");
    return PN_OK;
}

int64_t nt_image_bytes(int Np, int Kp) {
    int n_tiles;
    int bn = pick_bn(Np, &n_tiles);
    int nk = (Kp + TBK - 1) / TBK;
    return align_up((int64_t)n_tiles * nk * 2 * bn * 128, 256);
}
int pack_nt_image(const float* src, int64_t s_n, int64_t s_k, int N, int K, int Np, int Kp, char* img, cudaStream_t st) {
    int n_tiles;
    int bn = pick_bn(Np, &n_tiles);
    int nk = (Kp + TBK - 1) / TBK;
    int64_t total = (int64_t)n_tiles * nk * bn * 8;
    pack_nt_image_kernel<<<(unsigned)cdiv(total, 256), 256, 0, st>>>(src, s_n, s_k, N, K, bn, n_tiles, nk, img);
    PN_LAUNCH_CHECK("pack_nt_image");
    return PN_OK;
}

// several images in ONE launch (blockIdx.y = job): the aggregator packs its seven weight matrices per pass
struct PackJobDev { const float* src; int64_t s_n, s_k; int N, K, BN, n_tiles, nk; char* img; };
struct PackBatchDev { PackJobDev j[kMaxPackJobs]; };
__global__ void __launch_bounds__(256) pack_nt_image_batch_kernel(PackBatchDev b) {
    const PackJobDev& d = b.j[blockIdx.y];
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t total = (int64_t)d.n_tiles * d.nk * d.BN * 8;
    if (idx >= total) return;
    int c = (int)(idx & 7);
    int64_t rest = idx >> 3;
    int r = (int)(rest % d.BN);
    rest /= d.BN;
    int kc = (int)(rest % d.nk);
    int j = (int)(rest / d.nk);
    int n = j * d.BN + r;
    float v[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        int k = kc * TBK + 4 * c + e;
        v[e] = (n < d.N && k < d.K) ? d.src[(int64_t)n * d.s_n + (int64_t)k * d.s_k] : 0.f;
    }
    char* blk = d.img + ((int64_t)j * d.nk + kc) * (2 * d.BN * 128);
    split_store(blk, blk + d.BN * 128, sw128_off(r, c), make_float4(v[0], v[1], v[2], v[3]));
}
int pack_nt_images(const PackJob* jobs, int n, cudaStream_t st) {
    PN_CHECK_ARG(n >= 0 && n <= kMaxPackJobs, "pack_nt_images: at most %d jobs", kMaxPackJobs);
    if (n == 0) return PN_OK;
    PackBatchDev b;
    int64_t max_total = 0;
    for (int i = 0; i < n; ++i) {
        PackJobDev& d = b.j[i];
        d.src = jobs[i].src; d.s_n = jobs[i].s_n; d.s_k = jobs[i].s_k; d.N = jobs[i].N; d.K = jobs[i].K; d.img = jobs[i].img;
        d.BN = pick_bn(jobs[i].Np, &d.n_tiles);
        d.nk = (jobs[i].Kp + TBK - 1) / TBK;
        int64_t total = (int64_t)d.n_tiles * d.nk * d.BN * 8;
        if (total > max_total) max_total = total;
    }
    for (int i = n; i < kMaxPackJobs; ++i) b.j[i] = b.j[0];
    pack_nt_image_batch_kernel<<<dim3((unsigned)cdiv(max_total, 256), (unsigned)n), 256, 0, st>>>(b);
    PN_LAUNCH_CHECK("pack_nt_images");
    return PN_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// TN on CTA PAIRS (cta_group::2) for N = 129..256 output rows: the two n tiles of one row split form a cluster.  Each CTA
// stages ITS 128 columns of A (dY) and HALF of the columns of B (X); the leader issues MMAs of M = 256 that read both
// halves.  Per CTA that is 272 instead of 416 staged columns per chunk and 68 KB instead of 104 KB per stage, which makes
// room for a third stage (the two-stage kernel is bound by the producer -> MMA -> producer round trip).
// BKt2 = Kb rounded up to 64 (<= 320): columns 0..255 go through an MMA of N = 256 (128 per CTA), the rest through a second
// one of N = 64 (32 per CTA); CTA r therefore stages k columns [128 r, 128 r + 128) and [256 + 32 r, 256 + 32 r + 32).
// ---------------------------------------------------------------------------------------------------------------
constexpr int TNP_STAGES = 3;
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kTnThreads, 1)
    gemm_tn_pair_kernel(GemmTN g, int BKt2, float* __restrict__ partial, float* __restrict__ cpartial) {
    extern __shared__ char smem_raw[];
    char* smem = (char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    constexpr int PANEL = TBK * 128;
    const int N2 = BKt2 - 256 > 0 ? BKt2 - 256 : 0, N1 = BKt2 - N2;    // 256 | 0..64, or BKt2 | 0 below 256
    const int npan = (N1 >> 6) + (N2 >> 6);                            // B panels of this CTA: half of each MMA's columns
    const int b_bytes = npan * PANEL;
    const int stage_bytes = 2 * A_BYTES + 2 * b_bytes;
    const int region = TNP_STAGES * stage_bytes > TBM * (BKt2 + 4) * 4 ? TNP_STAGES * stage_bytes : TBM * (BKt2 + 4) * 4;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + ((region + 15) & ~15));
    uint64_t *full = bars, *empty = bars + 3, *peerFull = bars + 6, *done = bars + 9;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const uint32_t rank = tc::cluster_ctarank();
    const int split = (int)blockIdx.x >> 1;
    const int i0 = (int)rank * TBM;        // first column of A (n of the weight) of this CTA
    g.M = tn_rows(g);
    const int64_t rows_per_split = tn_rows_per_split(g.M, (int)gridDim.x >> 1);
    const int64_t r_begin = (int64_t)split * rows_per_split;
    const int64_t r_end = r_begin + rows_per_split < g.M ? r_begin + rows_per_split : g.M;
    const int nk = r_end > r_begin ? (int)((r_end - r_begin + TBK - 1) / TBK) : 0;
    if (nk == 0 && split > 0) return;           // an empty split slot: both CTAs of the pair leave (same M, same decision)

    if (t == 0) {
        for (int i = 0; i < TNP_STAGES; ++i) { tc::mbar_init(&full[i], kTnLoad / 64); tc::mbar_init(&empty[i], 1); tc::mbar_init(&peerFull[i], 1); }
        tc::mbar_init(done, 1);
        tc::mbar_fence_init();
    }
    if (warp == 16) tc::tmem_alloc2(tmem_slot, 512u);
    tc::fence_before_sync();
    __syncthreads();
    tc::cluster_sync();
    tc::fence_after_sync();
    const uint32_t tmem_d = *tmem_slot;

    if (warp < 16) {
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        const int grp = t >> 8, tt = t & 255;
        const int ac4 = tt & 31, ar0 = tt >> 5;    // A: 32 float4 per row (128 columns), rows ar0 + 8 i
        const bool a_ok = i0 + 4 * ac4 < g.N;
        const int ncb = npan * 8;                  // float4 per local B row (<= 40)
        constexpr int NB = 5;                      // 256 threads x 5 >= 32 rows x 40 float4
        int brow[NB], bc4[NB], bgc[NB];            // row | flags, local float4 column, global k column
#pragma unroll
        for (int i = 0; i < NB; ++i) {
            int idx = tt + 256 * i;
            const int br = idx / ncb;
            bc4[i] = idx - br * ncb;
            const int cl = 4 * bc4[i];             // local column
            bgc[i] = cl < (N1 >> 1) ? (int)rank * (N1 >> 1) + cl : N1 + (int)rank * (N2 >> 1) + (cl - (N1 >> 1));
            brow[i] = br >= TBK ? -1 : (br | (bgc[i] >= g.Kb ? 0x100 : 0));
        }
        float4 ra[4], rb[NB];
        float4 csum = z4;
        const uint32_t smem_u = tc::smem_u32(smem);
        auto load = [&](int kc) {
            const int64_t rb0 = r_begin + (int64_t)kc * TBK;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                int64_t rr = rb0 + ar0 + 8 * i;
                ra[i] = (a_ok && rr < r_end) ? __ldg(reinterpret_cast<const float4*>(g.A + rr * g.lda + i0 + 4 * ac4)) : z4;
            }
#pragma unroll
            for (int i = 0; i < NB; ++i) {
                int64_t rr = rb0 + (brow[i] & 0xFF);
                rb[i] = (brow[i] >= 0 && !(brow[i] & 0x100) && rr < r_end) ? __ldg(reinterpret_cast<const float4*>(g.B + rr * g.ldb + bgc[i])) : z4;
            }
        };
        if (grp < nk) load(grp);
        for (int kc = grp; kc < nk; kc += 2) {
            const int s = kc % TNP_STAGES;
            tc::mbar_wait(&empty[s], ((kc / TNP_STAGES) & 1) ^ 1);
            const uint32_t st = smem_u + (uint32_t)(s * stage_bytes);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                int row = ar0 + 8 * i;
                const uint32_t off = (uint32_t)((ac4 >> 3) * PANEL) + sw128b32_off(row, ac4 & 7);
                split_store_s(st + off, st + A_BYTES + off, ra[i]);
                csum.x += ra[i].x; csum.y += ra[i].y; csum.z += ra[i].z; csum.w += ra[i].w;
            }
#pragma unroll
            for (int i = 0; i < NB; ++i)
                if (brow[i] >= 0) {
                    const uint32_t off = (uint32_t)((bc4[i] >> 3) * PANEL) + sw128b32_off(brow[i] & 0xFF, bc4[i] & 7);
                    split_store_s(st + 2 * A_BYTES + off, st + 2 * A_BYTES + (uint32_t)b_bytes + off, rb[i]);
                }
            tc::fence_proxy_async();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&full[s]);
            if (kc + 2 < nk) load(kc + 2);
        }
      if (warp < 8) {
        // ---- epilogue: D[n, k] -> partial[split][n][k] (own 128 rows n; TMEM column = k)
        if (nk > 0) {
            tc::mbar_wait(done, 0);
            tc::fence_after_sync();
        }
        float* out = partial + (int64_t)split * g.N * g.Kb;
        epilogue_via_smem(tmem_d, BKt2, reinterpret_cast<float*>(smem), t, nk > 0, [&](int row, int col, float4 v) {
            const int n = i0 + row, k = col;
            if (n >= g.N || k >= g.Kb) return;
            *reinterpret_cast<float4*>(out + (int64_t)n * g.Kb + k) = v;
        });
        tc::fence_before_sync();
      }
        if (cpartial) {      // bias gradient: column sums of A over this split, fixed order
            float4* red = reinterpret_cast<float4*>(smem);
            __syncwarp();
            asm volatile("bar.sync 2, %0;" ::"n"(kTnLoad));
            red[t] = csum;
            asm volatile("bar.sync 2, %0;" ::"n"(kTnLoad));
            if (t < 32) {
                float4 a = red[t];
#pragma unroll
                for (int w16 = 1; w16 < 16; ++w16) {
                    float4 b = red[t + 32 * w16];
                    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
                }
                int nn = i0 + 4 * t;
                if (nn < g.N) *reinterpret_cast<float4*>(cpartial + (int64_t)split * g.N + nn) = a;
            }
        }
    } else {
      if (lane == 0 && rank == 0) {
        const uint32_t idesc1 = tc::idesc_tf32(2 * TBM, N1, 1, 1);
        const uint32_t idesc2 = N2 > 0 ? tc::idesc_tf32(2 * TBM, N2, 1, 1) : 0u;
        const uint32_t sbase = tc::smem_u32(smem);
        const uint32_t off2 = (uint32_t)((N1 >> 6) * PANEL);      // this CTA's columns of the second MMA follow those of the first
        for (int kc = 0; kc < nk; ++kc) {
            const int s = kc % TNP_STAGES;
            const uint32_t ph = (uint32_t)((kc / TNP_STAGES) & 1);
            tc::mbar_wait(&full[s], ph);
            tc::mbar_wait_cluster(&peerFull[s], ph);
            tc::fence_after_sync();
            const uint32_t a_hi = sbase + s * stage_bytes, a_lo = a_hi + A_BYTES;
            const uint32_t b_hi = a_hi + 2 * A_BYTES, b_lo = b_hi + (uint32_t)b_bytes;
            int64_t rows = r_end - (r_begin + (int64_t)kc * TBK);
            int ksteps = rows >= TBK ? 4 : (int)((rows + 7) >> 3);
            for (int k = 0; k < ksteps; ++k) {
                const uint32_t ko = 1024u * k;
                uint64_t dah = tc::smem_desc_sw128(a_hi + ko, PANEL, 512, 1), dal = tc::smem_desc_sw128(a_lo + ko, PANEL, 512, 1);
                uint64_t dbh = tc::smem_desc_sw128(b_hi + ko, PANEL, 512, 1), dbl = tc::smem_desc_sw128(b_lo + ko, PANEL, 512, 1);
                tc::mma_tf32_2(tmem_d, dah, dbl, idesc1, (kc | k) != 0);
                tc::mma_tf32_2(tmem_d, dal, dbh, idesc1, 1);
                tc::mma_tf32_2(tmem_d, dah, dbh, idesc1, 1);
                if (N2 > 0) {
                    uint64_t dbh2 = tc::smem_desc_sw128(b_hi + off2 + ko, PANEL, 512, 1), dbl2 = tc::smem_desc_sw128(b_lo + off2 + ko, PANEL, 512, 1);
                    tc::mma_tf32_2(tmem_d + (uint32_t)N1, dah, dbl2, idesc2, (kc | k) != 0);
                    tc::mma_tf32_2(tmem_d + (uint32_t)N1, dal, dbh2, idesc2, 1);
                    tc::mma_tf32_2(tmem_d + (uint32_t)N1, dah, dbh2, idesc2, 1);
                }
            }
            tc::mma_commit_2(&empty[s]);
        }
        tc::mma_commit_2(done);
      } else if (lane == 0) {
        for (int kc = 0; kc < nk; ++kc) {       // relay: this CTA's stage is full -> the leader's peerFull
            const int s = kc % TNP_STAGES;
            tc::mbar_wait(&full[s], (uint32_t)((kc / TNP_STAGES) & 1));
            tc::mbar_arrive_cluster(tc::mapa(tc::smem_u32(&peerFull[s]), 0));
        }
      }
      __syncwarp();
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::cluster_sync();
    if (warp == 16) {
        tc::fence_after_sync();
        tc::tmem_dealloc2(tmem_d, 512u);
    }
}

// ---- second stage of the TN reductions: partials are combined in split order (deterministic).  All weight-gradient GEMMs
// of a backward pass are finished by ONE launch (blockIdx.y = job): a launch per GEMM cost ~12 us each for microseconds of work.
struct TnFinishBatch { TnFinish j[kMaxTnJobs]; };
__global__ void __launch_bounds__(256) tn_finish_batch_kernel(TnFinishBatch b) {
    const TnFinish& f = b.j[blockIdx.y];
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    int64_t M = f.M;
    if (f.Mdev) { const int64_t md = (int64_t)__ldg(f.Mdev); M = md < M ? md : M; }
    const int64_t rps = tn_rows_per_split(M, f.slots);
    const int splits = M > 0 ? (int)((M + rps - 1) / rps) : 0;
    if (idx < f.N * f.Kb) {
        const int n = idx / f.Kb, k = idx - n * f.Kb;
        if (k >= f.Kout) return;
        // eight loads in flight per thread (the kernel is latency-bound: ~74 K threads, 74 strided reads each), summed in
        // split order - the same bits as a plain loop
        float acc = 0.f;
        const int64_t stride = (int64_t)f.N * f.Kb;
        const float* src = f.partial + idx;
        int z = 0;
        for (; z + 8 <= splits; z += 8) {
            float v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = __ldg(src + (int64_t)(z + i) * stride);
#pragma unroll
            for (int i = 0; i < 8; ++i) acc += v[i];
        }
        for (; z < splits; ++z) acc += __ldg(src + (int64_t)z * stride);
        f.C[(int64_t)n * f.ldc + k] = acc;
    } else if (f.colsum && idx - f.N * f.Kb < f.N) {
        const int n = idx - f.N * f.Kb;
        float acc = 0.f;
        for (int z = 0; z < splits; ++z) acc += f.cpartial[(int64_t)z * f.N + n];
        f.colsum[n] = acc;
    }
}
int tn_finish_batch(const TnFinish* jobs, int n, cudaStream_t st) {
    PN_CHECK_ARG(n >= 0 && n <= kMaxTnJobs, "tn_finish_batch: at most %d jobs", kMaxTnJobs);
    if (n == 0) return PN_OK;
    TnFinishBatch b;
    int64_t mx = 0;
    for (int i = 0; i < n; ++i) {
        b.j[i] = jobs[i];
        int64_t e = (int64_t)jobs[i].N * jobs[i].Kb + (jobs[i].colsum ? jobs[i].N : 0);
        if (e > mx) mx = e;
    }
    for (int i = n; i < kMaxTnJobs; ++i) b.j[i] = jobs[0];
    tn_finish_batch_kernel<<<dim3((unsigned)cdiv(mx, 256), (unsigned)n), 256, 0, st>>>(b);
    PN_LAUNCH_CHECK("tn_finish_batch");
    return PN_OK;
}

// one CTA per SM: k tiles x n tiles x split slots <= 148; the slices themselves are cut on the device (tn_rows_per_split)
static void tn_plan(int64_t Mcap, int N, int Kb, int* bkt, int* k_tiles, int* n_tiles, int* slots) {
    if (Kb <= 288) { *bkt = (int)align_up(Kb, 16); *k_tiles = 1; }
    else *bkt = pick_bn(Kb, k_tiles);
    *n_tiles = (int)cdiv(N, TBM);
    int s = kNumSMs / (*k_tiles * *n_tiles);
    if (s < 1) s = 1;
    int64_t most = cdiv(Mcap > 0 ? Mcap : 1, 256);      // a slice holds at least 256 rows
    *slots = (int)(most < s ? most : s);
}

int gemm_tn(const GemmTN& g, float* scratch, TnFinish* fin, cudaStream_t st) {
    PN_CHECK_ARG(g.lda % 4 == 0 && g.ldb % 4 == 0 && g.N % 4 == 0 && g.Kb % 4 == 0, "gemm_tn: operands must be padded");
    PN_CHECK_ARG(((uintptr_t)g.A & 15) == 0 && ((uintptr_t)g.B & 15) == 0, "gemm_tn: operands must be 16-byte aligned");
    int bkt, k_tiles, n_tiles, slots;
    tn_plan(g.M, g.N, g.Kb, &bkt, &k_tiles, &n_tiles, &slots);
    float* partial = scratch;
    float* cpartial = g.colsum ? scratch + (int64_t)slots * g.N * g.Kb : nullptr;
    TnFinish f{partial, cpartial, slots, g.N, g.Kb, g.Kout, g.C, g.ldc, g.colsum, g.M, g.Mdev};
    const int bkt2 = (int)align_up(g.Kb, 64);
    if (n_tiles == 2 && k_tiles == 1 && bkt2 <= 320) {   // CTA pairs: the two n tiles of a split
        const int n2 = bkt2 > 256 ? bkt2 - 256 : 0, n1 = bkt2 - n2;
        const int np2 = (n1 >> 6) + (n2 >> 6);
        const int stages_b = TNP_STAGES * (2 * A_BYTES + 2 * np2 * TBK * 128), tile_b = TBM * (bkt2 + 4) * 4;
        const int smem2 = (stages_b > tile_b ? stages_b : tile_b) + 16 + 1024 + 256;
        static int attr2 = 0;
        if (smem2 > attr2) {
            PN_CUDA(cudaFuncSetAttribute(gemm_tn_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem2));
            attr2 = smem2;
        }
        gemm_tn_pair_kernel<<<2 * slots, kTnThreads, smem2, st>>>(g, bkt2, partial, cpartial);
        PN_LAUNCH_CHECK("gemm_tn_pair");
    } else {
        const int npan = (bkt + 31) / 32;
        const int smem_bytes = 2 * (2 * A_BYTES + 2 * npan * TBK * 128) + 1024 + 256;
        static int attr_bytes = 0;
        if (smem_bytes > attr_bytes) {
            PN_CUDA(cudaFuncSetAttribute(gemm_tn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
            attr_bytes = smem_bytes;
        }
        dim3 grid((unsigned)k_tiles, (unsigned)n_tiles, (unsigned)slots);
        gemm_tn_tc_kernel<<<grid, kTnThreads, smem_bytes, st>>>(g, bkt, bkt > 256 ? 512 : tmem_cols_for(bkt), partial, cpartial);
        PN_LAUNCH_CHECK("gemm_tn_tc");
    }
    if (fin) { *fin = f; return PN_OK; }
    return tn_finish_batch(&f, 1, st);
}
int64_t gemm_tn_scratch_floats(int64_t Mcap, int N, int Kb) {
    int bkt, k_tiles, n_tiles, slots;
    tn_plan(Mcap, N, Kb, &bkt, &k_tiles, &n_tiles, &slots);
    return align_up((int64_t)slots * ((int64_t)N * Kb + N) + 64, 64);
}

}  // namespace pn
