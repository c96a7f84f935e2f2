// Error state, launch counter and the device-wide scan used by the grid build and the work lists.
#include <stdarg.h>
#include <stdint.h>

#include <atomic>

#include "common.cuh"

namespace pn {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// ------------------------------------------------------------------------------------------------
// scan: tile = 1024 threads x 8 items.  Phase 1 reduces tiles, phase 2 scans the tile sums in one
// block, phase 3 rescans each tile with its offset.  Reads the input twice, writes once: HBM-bound,
// 12 B/element for int32.
// ------------------------------------------------------------------------------------------------
constexpr int kScanThreads = 1024;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;

template <typename T>
__device__ __forceinline__ T block_scan_excl(T v, T* total_out, T* smem /* 32 */) {
    // exclusive scan of one value per thread across a 1024-thread block
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    T incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        T t = __shfl_up_sync(kFull, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) smem[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        T w = smem[lane];
        T wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            T t = __shfl_up_sync(kFull, wi, o);
            if (lane >= o) wi += t;
        }
        smem[lane] = wi - w;  // exclusive warp offsets
        if (lane == 31 && total_out) *total_out = wi;
    }
    __syncthreads();
    T r = smem[wid] + incl - v;
    __syncthreads();
    return r;
}

template <typename T>
__global__ void __launch_bounds__(kScanThreads) scan_reduce_kernel(const T* __restrict__ in, int64_t n, T* __restrict__ bsum) {
    __shared__ T sm[32];
    __shared__ T tot;
    int64_t base = (int64_t)blockIdx.x * kScanTile;
    T acc = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
        int64_t idx = base + (int64_t)i * kScanThreads + threadIdx.x;
        if (idx < n) acc += in[idx];
    }
    block_scan_excl<T>(acc, &tot, sm);
    if (threadIdx.x == 0) bsum[blockIdx.x] = tot;
}

template <typename T>
__global__ void __launch_bounds__(kScanThreads) scan_bsum_kernel(T* __restrict__ bsum, int64_t nb, T* __restrict__ total) {
    __shared__ T sm[32];
    __shared__ T tot;
    // each thread owns a contiguous chunk of the tile sums
    int64_t per = (nb + kScanThreads - 1) / kScanThreads;
    int64_t lo = (int64_t)threadIdx.x * per, hi = lo + per < nb ? lo + per : nb;
    T acc = 0;
    for (int64_t i = lo; i < hi; ++i) acc += bsum[i];
    T off = block_scan_excl<T>(acc, &tot, sm);
    for (int64_t i = lo; i < hi; ++i) {
        T v = bsum[i];
        bsum[i] = off;
        off += v;
    }
    if (threadIdx.x == 0 && total) *total = tot;
}

template <typename T>
__global__ void __launch_bounds__(kScanThreads) scan_apply_kernel(const T* in, T* out, int64_t n, const T* __restrict__ bsum) {
    __shared__ T sm[32];
    // thread t owns kScanItems CONSECUTIVE elements so that one block scan suffices
    int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
    // a thread's 8 elements are 32 / 64 contiguous bytes: moved as 16-byte vectors when the tile is whole and the
    // pointers are aligned (element-wise loads touch 32 sectors per instruction and ran at a quarter of HBM speed)
    constexpr int kVec = (int)(sizeof(T) * kScanItems / 16);
    const bool vec = base + kScanItems <= n && (((uintptr_t)in | (uintptr_t)out) & 15) == 0;
    union { uint4 q[kVec]; T v[kScanItems]; } u;
    T acc = 0;
    if (vec) {
        const uint4* src = reinterpret_cast<const uint4*>(in + base);
#pragma unroll
        for (int i = 0; i < kVec; ++i) u.q[i] = src[i];
#pragma unroll
        for (int i = 0; i < kScanItems; ++i) acc += u.v[i];
    } else {
#pragma unroll
        for (int i = 0; i < kScanItems; ++i) {
            u.v[i] = (base + i < n) ? in[base + i] : (T)0;
            acc += u.v[i];
        }
    }
    T off = block_scan_excl<T>(acc, (T*)nullptr, sm) + bsum[blockIdx.x];
    if (vec) {
#pragma unroll
        for (int i = 0; i < kScanItems; ++i) {
            T t = u.v[i];
            u.v[i] = off;
            off += t;
        }
        uint4* dst = reinterpret_cast<uint4*>(out + base);
#pragma unroll
        for (int i = 0; i < kVec; ++i) dst[i] = u.q[i];
    } else {
#pragma unroll
        for (int i = 0; i < kScanItems; ++i) {
            if (base + i < n) out[base + i] = off;
            off += u.v[i];
        }
    }
}

template <typename T>
static int scan_impl(const T* in, T* out, int64_t n, T* total, T* scratch, cudaStream_t st, const char* what) {
    if (n <= 0) {
        if (total) cudaMemsetAsync(total, 0, sizeof(T), st);
        return PN_OK;
    }
    int64_t nb = cdiv(n, kScanTile);
    scan_reduce_kernel<T><<<(unsigned)nb, kScanThreads, 0, st>>>(in, n, scratch);
    PN_LAUNCH_CHECK(what);
    scan_bsum_kernel<T><<<1, kScanThreads, 0, st>>>(scratch, nb, total);
    PN_LAUNCH_CHECK(what);
    scan_apply_kernel<T><<<(unsigned)nb, kScanThreads, 0, st>>>(in, out, n, scratch);
    PN_LAUNCH_CHECK(what);
    return PN_OK;
}

int64_t scan_scratch_elems(int64_t n) { return cdiv(n > 0 ? n : 1, kScanTile) + 1; }
int64_t scan_scratch_elems_i32(int64_t n) { return cdiv(n > 0 ? n : 1, kScanTile) + 1; }

int scan_exclusive_u64(const unsigned long long* in, unsigned long long* out, int64_t n, unsigned long long* total,
                       unsigned long long* scratch, cudaStream_t st) {
    return scan_impl<unsigned long long>(in, out, n, total, scratch, st, "scan_u64");
}

int scan_exclusive_i32(const int* in, int* out, int64_t n, bool write_total_at_end, int* scratch, cudaStream_t st) {
    return scan_impl<int>(in, out, n, write_total_at_end ? out + n : nullptr, scratch, st, "scan_i32");
}

}  // namespace pn

extern "C" {
const char* pn_last_error(void) { return pn::g_err; }
int pn_abi_version(void) { return 11; }
int64_t pn_launch_count(void) { return pn::g_launches.load(); }
void pn_launch_count_reset(void) { pn::g_launches.store(0); }
}
