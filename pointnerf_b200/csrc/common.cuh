// Shared host/device helpers of libpointnerf_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/pointnerf_b200.h"

namespace pn {

constexpr int kWarp = 32;
constexpr unsigned kFull = 0xffffffffu;
constexpr int kNumSMs = 148;  // B200

void set_error(const char* fmt, ...);
void count_launch(int n = 1);

#define PN_CHECK_ARG(cond, ...)              \
    do {                                     \
        if (!(cond)) {                       \
            pn::set_error(__VA_ARGS__);      \
            return PN_ERR_ARG;               \
        }                                    \
    } while (0)

// call after a kernel launch: records the launch and converts launch errors into a status
#define PN_LAUNCH_CHECK(name)                                                            \
    do {                                                                                 \
        pn::count_launch();                                                              \
        cudaError_t _e = cudaGetLastError();                                             \
        if (_e != cudaSuccess) {                                                         \
            pn::set_error("%s: launch failed: %s", name, cudaGetErrorString(_e));       \
            return PN_ERR_CUDA;                                                          \
        }                                                                                \
    } while (0)

#define PN_CUDA(call)                                                                    \
    do {                                                                                 \
        cudaError_t _e = (call);                                                         \
        if (_e != cudaSuccess) {                                                         \
            pn::set_error("%s failed: %s", #call, cudaGetErrorString(_e));              \
            return PN_ERR_CUDA;                                                          \
        }                                                                                \
    } while (0)

static inline int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }
static inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }

// bump allocator over a caller-provided workspace blob (256-byte aligned slices)
struct Carver {
    char* base;
    int64_t off, cap;
    Carver(void* p, int64_t bytes) : base((char*)p), off(0), cap(bytes) {}
    template <typename T>
    T* take(int64_t n) {
        int64_t b = align_up(n * (int64_t)sizeof(T), 256);
        T* r = (T*)(base + off);
        off += b;
        return r;
    }
    bool ok() const { return off <= cap && (base != nullptr || off == 0); }
};

// ---- device helpers -------------------------------------------------------------------------
// voxel coordinate exactly as the reference's PTX: sub.f32, div.rn.f32, cvt.rmi, cvt.rzi
__device__ __forceinline__ int voxel_coord(float p, float shift, float vsize) {
    return (int)floorf(__fdiv_rn(__fsub_rn(p, shift), vsize));
}

__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    return v;
}
// inclusive prefix sum across the warp
__device__ __forceinline__ int warp_scan_incl(int v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(kFull, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

// ---- device-wide exclusive scan of uint64 (three phases, deterministic) ----------------------
// out[i] = sum_{j<i} in[j]; total written to *total (may be null).  scratch: scan_scratch_elems(n) uint64.
int64_t scan_scratch_elems(int64_t n);
int scan_exclusive_u64(const unsigned long long* in, unsigned long long* out, int64_t n,
                       unsigned long long* total, unsigned long long* scratch, cudaStream_t st);
// same for int32 in place-capable (out may alias in); out has n+1 entries when write_total_at_end
int scan_exclusive_i32(const int* in, int* out, int64_t n, bool write_total_at_end, int* scratch,
                       cudaStream_t st);
int64_t scan_scratch_elems_i32(int64_t n);

}  // namespace pn
