// Internal GEMM building blocks of the aggregator MLP (not part of the C ABI).
#pragma once
#include "common.cuh"

namespace pn {

// C[m, n] = epilogue( sum_k A[m*lda + k] * B[n*ldb + k] ),  m < M, n < N, k < Kd  (Kd % 16 == 0,
// lda % 4 == 0, ldb % 4 == 0, A/B 16-byte aligned: the callers pad their operands).
// epilogue, in this order:
//   v += bias[n]                                 (bias != null)
//   v  = v > 0 ? v : slope * v                   (leaky != 0)
//   sign_out[m*sign_ld + n/32] bit n%32 = v > 0  (sign_out != null; N % 32 == 0; written for every FULL group of 32 rows: what a
//                                                 later data-gradient GEMM needs of this layer's output, 1 bit per value)
//   v *= aux[m*ldaux + n] > 0 ? 1 : slope        (aux != null and n < aux_cols: LeakyReLU backward mask)
//                                                 with sign_in != null (bits of the same aux, sign_ld = aux_cols / 32) the full
//                                                 32 x 32 sub-tiles read the bits instead of the values
//   C[rowmap ? rowmap[m] : m][n] = v             (row stride ldc)
struct GemmNT {
    const float* A; int64_t lda;
    const float* B; int64_t ldb;
    float* C; int64_t ldc;
    int64_t M; int N; int Kd;
    const float* bias;
    int leaky; float slope;
    const float* aux; int64_t ldaux; int aux_cols;
    const int* rowmap;  // unused (kept so that positional initialisers stay valid)
    const char* Bimg;   // pre-split, pre-swizzled operand image of B (pack_nt_image)
    const int32_t* Mdev;  // optional DEVICE row count: rows = min(*Mdev, M); M is then the capacity the launch is sized for
    uint32_t* sign_out; const uint32_t* sign_in; int sign_ld;   // activation sign bits, sign_ld words per row
};
int gemm_nt(const GemmNT& g, cudaStream_t st);
// Operand image of a B matrix for the tensor-core NT GEMM: B[n, k] = src[n*s_n + k*s_k] for n < N, k < K, zero beyond,
// logically padded to Np rows x Kp columns; split into TF32 hi/lo and stored tile by tile in the exact shared-memory
// layout (128-byte swizzle) so that one TMA bulk copy per K chunk stages it.
int64_t nt_image_bytes(int Np, int Kp);
int pack_nt_image(const float* src, int64_t s_n, int64_t s_k, int N, int K, int Np, int Kp, char* img, cudaStream_t st);
constexpr int kMaxPackJobs = 8;
struct PackJob { const float* src; int64_t s_n, s_k; int N, K, Np, Kp; char* img; };
int pack_nt_images(const PackJob* jobs, int n, cudaStream_t st);   // same images, one launch

// Weight gradient: C[n, k] = sum_m A[m*lda + n] * B[m*ldb + k]  for n < N, k < Kb, reduced over M rows in
// fixed-size splits; partial sums are combined in split order by a second kernel (deterministic).
// Output is written UNPADDED: C[n*ldc + k] for k < Kout (<= Kb).  colsum (optional, [N]) receives
// sum_m A[m, n] (the bias gradient) through the same fixed-order reduction.
// scratch: gemm_tn_scratch_floats(M, N, Kb) floats.
struct GemmTN {
    const float* A; int64_t lda;   // [M, N]  (dPre)
    const float* B; int64_t ldb;   // [M, Kb] (layer input, padded)
    float* C; int64_t ldc; int Kout;
    float* colsum;
    int64_t M; int N; int Kb;
    const int32_t* Mdev;  // optional DEVICE row count (see GemmNT::Mdev)
};
// second stage (fixed-order sum of the split partials) of one TN GEMM; several are finished by one launch
struct TnFinish {
    const float* partial; const float* cpartial; int slots, N, Kb, Kout;
    float* C; int64_t ldc; float* colsum; int64_t M; const int32_t* Mdev;
};
constexpr int kMaxTnJobs = 8;
int64_t gemm_tn_scratch_floats(int64_t Mcap, int N, int Kb);
// scratch: gemm_tn_scratch_floats floats, private to this GEMM until its finish has run.  fin == NULL: finished right away.
int gemm_tn(const GemmTN& g, float* scratch, TnFinish* fin, cudaStream_t st);
int tn_finish_batch(const TnFinish* jobs, int n, cudaStream_t st);

}  // namespace pn
