// Host stand-in for <curand_kernel.h>.  The time-seeded reservoir paths of claim_occ /
// fill_occ2pnts (worldcoords.py:290-298, 378-383) are not a parity target (overflow is
// a documented divergence); parity scenes never reach them, and the stub aborts if one does.
//
// This is the LAST header the kernel string includes, so it also switches on the arithmetic model:
// nvcc compiles the string with its default -fmad=true, which contracts
//     x*x + y*y + z*z   into   fma(z, z, fma(x, x, y*y))
// (checked with `nvcc -arch=sm_100a -ptx`, CUDA 12.9).  g++ would not contract, so `float` in the
// kernel bodies is replaced by a one-float struct whose operator* defers the rounding of a product
// and whose operator+ applies exactly that contraction.  Everything else converts back to plain
// float, i.e. IEEE single precision sub / div.rn / floor as in the PTX.
#pragma once
#include <cstdio>
#include <cmath>

struct curandState { int unused; };
static inline void curand_init(unsigned long long, unsigned long long, unsigned long long, curandState*) {}
static inline float curand_uniform(curandState*) {
    fprintf(stderr, "pn_ref_querier: reservoir overflow path reached (max_o or P exceeded)\n");
    abort();
}

struct pn_gfloat;
struct pn_gprod { float a, b; };
struct pn_gfloat {
    float v;
    pn_gfloat() = default;
    pn_gfloat(float f) : v(f) {}
    pn_gfloat(double d) : v((float)d) {}
    pn_gfloat(int i) : v((float)i) {}
    pn_gfloat(pn_gprod p) : v(p.a * p.b) {}
    operator float() const { return v; }
};
static inline pn_gprod operator*(pn_gfloat a, pn_gfloat b) { return pn_gprod{a.v, b.v}; }
static inline pn_gfloat operator+(pn_gprod p, pn_gprod q) { return pn_gfloat(fmaf(p.a, p.b, q.a * q.b)); }
static inline pn_gfloat operator+(pn_gfloat s, pn_gprod q) { return pn_gfloat(fmaf(q.a, q.b, s.v)); }
static inline pn_gfloat operator+(pn_gprod p, pn_gfloat s) { return pn_gfloat(fmaf(p.a, p.b, s.v)); }

// The frustum querier (query_point_indices.py:532-553) adds: int * float (voxel index times voxel size, contracted into the
// following add), (int + 0.5) * float evaluated in double and contracted into a double fma with the float partial sum, a
// difference of two products that nvcc leaves as mul, mul, sub, and a ?: whose arms must have one type (so float - float
// is declared to return the wrapper).  All checked against `nvcc -arch=sm_100a -ptx` of the kernel string, CUDA 12.9.
struct pn_gdprod { double a, b; };
static inline pn_gprod operator*(int a, pn_gfloat b) { return pn_gprod{(float)a, b.v}; }
static inline pn_gdprod operator*(double a, pn_gfloat b) { return pn_gdprod{a, (double)b.v}; }
static inline double operator+(pn_gfloat s, pn_gdprod q) { return fma(q.a, q.b, (double)s.v); }
static inline pn_gfloat operator-(pn_gfloat a, pn_gfloat b) { return pn_gfloat(a.v - b.v); }
static inline pn_gfloat operator-(pn_gprod p, pn_gprod q) { return pn_gfloat(p.a * p.b - q.a * q.b); }
static inline bool operator<=(pn_gprod p, pn_gfloat s) { return p.a * p.b <= s.v; }   // only query_rand_along_ray (:416-492, not a parity target)

#define float pn_gfloat
