// Host stand-in for <cuda_runtime.h>: just enough of the CUDA dialect for the six kernels of
// models/neural_points/query_point_indices_worldcoords.py:114-530 and the five of query_point_indices.py:135-600
// (the perspective-frustum querier) to compile with g++ and run
// one "thread" at a time.  Atomics are plain read-modify-write because execution is serial
// (ascending global thread index = the canonical schedule of DESIGN.md).
#pragma once
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <algorithm>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline

struct pn_uint3 { unsigned int x, y, z; };
static pn_uint3 blockIdx = {0, 0, 0}, threadIdx = {0, 0, 0}, blockDim = {1, 1, 1}, gridDim = {1, 1, 1};

static inline int max(int a, int b) { return a > b ? a : b; }
static inline int min(int a, int b) { return a < b ? a : b; }

static inline int atomicCAS(int* p, int cmp, int val) { int o = *p; if (o == cmp) *p = val; return o; }
static inline unsigned int atomicCAS(unsigned int* p, unsigned int cmp, unsigned int val) { unsigned int o = *p; if (o == cmp) *p = val; return o; }
static inline unsigned long long atomicCAS(unsigned long long* p, unsigned long long cmp, unsigned long long val) { unsigned long long o = *p; if (o == cmp) *p = val; return o; }
static inline int atomicAdd(int* p, int v) { int o = *p; *p = o + v; return o; }
static inline unsigned int atomicAdd(unsigned int* p, unsigned int v) { unsigned int o = *p; *p = o + v; return o; }
static inline unsigned int atomicSub(unsigned int* p, unsigned int v) { unsigned int o = *p; *p = o - v; return o; }
static inline int atomicMin(int* p, int v) { int o = *p; if (v < o) *p = v; return o; }
static inline int atomicMax(int* p, int v) { int o = *p; if (v > o) *p = v; return o; }

static inline unsigned int __byte_perm(unsigned int x, unsigned int y, unsigned int s) {
    unsigned long long src = ((unsigned long long)y << 32) | x; unsigned int r = 0;
    for (int i = 0; i < 4; i++) { unsigned int sel = (s >> (4 * i)) & 7; r |= (unsigned int)((src >> (8 * sel)) & 0xff) << (8 * i); }
    return r;
}
static inline double __longlong_as_double(long long v) { double d; memcpy(&d, &v, 8); return d; }
static inline long long __double_as_longlong(double d) { long long v; memcpy(&v, &d, 8); return v; }
static inline float __uint_as_float(unsigned int v) { float f; memcpy(&f, &v, 4); return f; }
static inline unsigned int __float_as_uint(float f) { unsigned int v; memcpy(&v, &f, 4); return v; }
