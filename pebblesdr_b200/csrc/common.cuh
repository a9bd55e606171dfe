// common.cuh — shared types and helpers for the pebblesdr_b200 kernels.
// The core math lives in PB_HD functions so the same source compiles (a) as sm_100a device code and
// (b) with plain g++ for the thread-by-thread host emulation in tests/host_emul/ (no GPU in CI).
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
