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
// correction of x * r: within 1 ulp of the correctly rounded root.
__device__ __forceinline__ double pb_sqrt_newton(double x) {
    // no branch at all (a fallback branch to sqrt() kept the 8 evaluations of a block from interleaving: ncu of the AM walk, DFMA /
    // DMUL waits inside this routine were a third of all stall samples): inputs outside the float range are rescaled by an even
    // power of two with selects, zero is patched at the end
    const bool tiny = x < 1e-30, huge = x > 1e30;
    const double xs = x * (tiny ? 0x1p200 : (huge ? 0x1p-200 : 1.0));
    double r = (double)rsqrtf((float)xs);
    const double hx = 0.5 * xs;
    r = fma(r, fma(-hx, r * r, 0.5), r);
    r = fma(r, fma(-hx, r * r, 0.5), r);
    double s = xs * r;
    s = fma(fma(-s, s, xs), 0.5 * r, s);
    s *= tiny ? 0x1p-100 : (huge ? 0x1p100 : 1.0);
    // 0 -> 0, NaN -> NaN, +inf -> +inf by selects (negative arguments do not occur: the callers pass |z|^2)
    return x > 0.0 ? (x > 1.7e308 ? x : s) : x * 0.0;
}
#endif
