// common.cuh — shared types and helpers for the pebblesdr_b200 kernels.
// The core math lives in PB_HD functions: the same source compiles as sm_100a device code and, in the host-only
// translation units (host_design.cpp), with plain g++.
#pragma once
#include <stdint.h>
#include <math.h>

#if defined(__CUDACC__)
#include <cuda_runtime.h>
#define PB_HD __host__ __device__ __forceinline__
#define PB_D __device__ __forceinline__
#else
#define PB_HD inline
struct alignas(16) double2 { double x, y; };  // same layout as CUDA's vector type
static inline double2 make_double2(double x, double y) { double2 r; r.x = x; r.y = y; return r; }
#endif

// CPX == std::complex<double> (pebblelib/cpx.h:96): 16 bytes, re then im.
typedef double2 cpx;

PB_HD cpx cpx_make(double re, double im) { return make_double2(re, im); }
PB_HD cpx cpx_add(cpx a, cpx b) { return make_double2(a.x + b.x, a.y + b.y); }
PB_HD cpx cpx_sub(cpx a, cpx b) { return make_double2(a.x - b.x, a.y - b.y); }
PB_HD cpx cpx_mul(cpx a, cpx b) { return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
// a * conj(b)
PB_HD cpx cpx_mulc(cpx a, cpx b) { return make_double2(a.x * b.x + a.y * b.y, a.y * b.x - a.x * b.y); }
PB_HD cpx cpx_scale(cpx a, double s) { return make_double2(a.x * s, a.y * s); }
// multiply by -j (forward) or +j (inverse)
template <bool INV> PB_HD cpx cpx_mulj(cpx a) {
    return INV ? make_double2(-a.y, a.x) : make_double2(a.y, -a.x);
}
static_assert(sizeof(cpx) == 16 && alignof(cpx) == 16, "cpx must match CUDA double2 in host-only translation units");

#if defined(__CUDACC__)
// sqrt(x), x >= 0, without the library routine's branches: float rsqrt seed (2^-22), two Newton steps on 1/sqrt (2^-85), one
// correction of x * r: within 1 ulp of the correctly rounded root. Outside the float range (never for IQ envelopes) and for
// 0 it falls back to sqrt().
__device__ __forceinline__ double pb_sqrt_newton(double x) {
    if (!(x > 1e-36 && x < 1e36)) return sqrt(x);
    double r = (double)rsqrtf((float)x);
    const double hx = 0.5 * x;
    r = fma(r, fma(-hx, r * r, 0.5), r);
    r = fma(r, fma(-hx, r * r, 0.5), r);
    double s = x * r;
    s = fma(fma(-s, s, x), 0.5 * r, s);
    return s;
}
#endif
