// dc_input.cuh — device sample formats at the mixer's load.
//
// DeviceInterfaceBase::normalizeIQ (pebblelib/deviceinterfacebase.cpp:648-878) turns what a device delivers — CPX8 (HackRF),
// CPXU8 (RTL2832, offset binary), CPX16 (SDRplay / sound cards), CPXFLOAT, CPX (pebblelib/cpx.h:43-96) — into CPX on the
// host, one multiply per component:  out = (code [- 128]) * scale,  scale = 1/128 (8-bit) or 1/32768 (16-bit) or 1, times
// userIQGain * normalizeIQGain. Here the raw samples are what crosses PCIe and HBM (2..8 bytes instead of 16) and the
// conversion happens in registers where the mixer loads its input:
//   * integer code -> double without the FP64 conversion unit (I2F.F64 issues at a quarter of the DADD rate): the code is
//     OR-ed into the mantissa of 2^52 (offset binary, so the value is non-negative) and one DADD subtracts 2^52 + offset.
//     Every code is an exact double, so this equals the reference's int -> double promotion bit for bit;
//   * `scale` does NOT cost a multiply: mixing is linear, so it rides on the oscillator seed (x * scale) * osc = x * (scale *
//     osc). For the default gains scale is a power of two and the product is exact; otherwise the chain output differs from
//     "normalizeIQ, then CDownConvert" by one rounding (1e-16 relative), far inside the chain's 1e-10 tolerance.
// The stand-alone conversion (iq_convert_kernel, pebble_normalize_iq*) multiplies exactly like the reference and is
// bit-identical to it for every code, order and gain.
#pragma once
#include "common.cuh"

// public numbering (include/pebble_cuda.h PEBBLE_IQ_*)
#define PB_IQ_CPX 0
#define PB_IQ_CS16 1
#define PB_IQ_CU8 2
#define PB_IQ_CS8 3
#define PB_IQ_CF32 4

PB_HD int iq_bytes_per_sample(int fmt) { return fmt == PB_IQ_CPX ? 16 : fmt == PB_IQ_CS16 ? 4 : fmt == PB_IQ_CF32 ? 8 : 2; }

#if defined(__CUDACC__)
// (2^52 + u) - (2^52 + off): u in [0, 2^16)
__device__ __forceinline__ double iq_code_to_double(unsigned u, double magic_minus) {
    return __hiloint2double(0x43300000, (int)u) - magic_minus;
}
#define PB_IQ_MAGIC16 4503599627403264.0   /* 2^52 + 32768 */
#define PB_IQ_MAGIC8 4503599627370624.0    /* 2^52 + 128 */

// One PAIR of consecutive samples from raw words.
template <int FMT> struct IqPair;
template <> struct IqPair<PB_IQ_CS16> {   // w0 = sample 0 (I | Q << 16), w1 = sample 1
    static __device__ __forceinline__ void cvt(unsigned w0, unsigned w1, cpx &x0, cpx &x1) {
        w0 ^= 0x80008000u; w1 ^= 0x80008000u;                       // two's complement -> offset binary
        x0 = cpx_make(iq_code_to_double(w0 & 0xffffu, PB_IQ_MAGIC16), iq_code_to_double(w0 >> 16, PB_IQ_MAGIC16));
        x1 = cpx_make(iq_code_to_double(w1 & 0xffffu, PB_IQ_MAGIC16), iq_code_to_double(w1 >> 16, PB_IQ_MAGIC16));
    }
};
template <> struct IqPair<PB_IQ_CU8> {    // w = I0 | Q0 << 8 | I1 << 16 | Q1 << 24
    static __device__ __forceinline__ void cvt(unsigned w, cpx &x0, cpx &x1) {
        x0 = cpx_make(iq_code_to_double(w & 0xffu, PB_IQ_MAGIC8), iq_code_to_double((w >> 8) & 0xffu, PB_IQ_MAGIC8));
        x1 = cpx_make(iq_code_to_double((w >> 16) & 0xffu, PB_IQ_MAGIC8), iq_code_to_double(w >> 24, PB_IQ_MAGIC8));
    }
};
template <> struct IqPair<PB_IQ_CS8> {
    static __device__ __forceinline__ void cvt(unsigned w, cpx &x0, cpx &x1) { IqPair<PB_IQ_CU8>::cvt(w ^ 0x80808080u, x0, x1); }
};

// The raw registers of one thread's 2K consecutive samples (K pairs), loaded with one or two wide global loads, and the
// in-register conversion of pair q. `p` is the BYTE address of the thread's first sample.
template <int FMT, int K> struct IqRaw;

template <int K> struct IqRaw<PB_IQ_CPX, K> {
    static constexpr int BYTES = 16;
    cpx a[K], b[K];
    __device__ __forceinline__ void zero() {
#pragma unroll
        for (int k = 0; k < K; ++k) a[k] = b[k] = cpx_make(0.0, 0.0);
    }
    __device__ __forceinline__ void load(const char *p) {
#pragma unroll
        for (int k = 0; k < K; ++k)
            asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];"
                         : "=d"(a[k].x), "=d"(a[k].y), "=d"(b[k].x), "=d"(b[k].y) : "l"(p + 32 * k));
    }
    __device__ __forceinline__ void get(int q, cpx &x0, cpx &x1) const { x0 = a[q]; x1 = b[q]; }
};

template <int K> struct IqRaw<PB_IQ_CS16, K> {   // 4 bytes per sample: one 128-bit load per two pairs
    static constexpr int BYTES = 4;
    static_assert(K % 2 == 0, "pairs come in twos (128-bit loads)");
    unsigned w[2 * K];
    __device__ __forceinline__ void zero() {
#pragma unroll
        for (int k = 0; k < 2 * K; ++k) w[k] = 0u;
    }
    __device__ __forceinline__ void load(const char *p) {
#pragma unroll
        for (int k = 0; k < K / 2; ++k)
            asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(w[4 * k]), "=r"(w[4 * k + 1]), "=r"(w[4 * k + 2]), "=r"(w[4 * k + 3]) : "l"(p + 16 * k));
    }
    __device__ __forceinline__ void get(int q, cpx &x0, cpx &x1) const { IqPair<PB_IQ_CS16>::cvt(w[2 * q], w[2 * q + 1], x0, x1); }
};

template <int FMT8, int K> struct IqRaw8 {       // 2 bytes per sample: one 64-bit load per two pairs
    static constexpr int BYTES = 2;
    static_assert(K % 2 == 0, "pairs come in twos (64-bit loads)");
    unsigned w[K];
    __device__ __forceinline__ void zero() {
#pragma unroll
        for (int k = 0; k < K; ++k) w[k] = FMT8 == PB_IQ_CU8 ? 0x80808080u : 0u;   // the code of 0.0
    }
    __device__ __forceinline__ void load(const char *p) {
#pragma unroll
        for (int k = 0; k < K / 2; ++k)
            asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(w[2 * k]), "=r"(w[2 * k + 1]) : "l"(p + 8 * k));
    }
    __device__ __forceinline__ void get(int q, cpx &x0, cpx &x1) const { IqPair<FMT8>::cvt(w[q], x0, x1); }
};
template <int K> struct IqRaw<PB_IQ_CU8, K> : IqRaw8<PB_IQ_CU8, K> {};
template <int K> struct IqRaw<PB_IQ_CS8, K> : IqRaw8<PB_IQ_CS8, K> {};

template <int K> struct IqRaw<PB_IQ_CF32, K> {   // 8 bytes per sample: one 128-bit load per pair
    static constexpr int BYTES = 8;
    float f[4 * K];
    __device__ __forceinline__ void zero() {
#pragma unroll
        for (int k = 0; k < 4 * K; ++k) f[k] = 0.f;
    }
    __device__ __forceinline__ void load(const char *p) {
#pragma unroll
        for (int k = 0; k < K; ++k)
            asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                         : "=f"(f[4 * k]), "=f"(f[4 * k + 1]), "=f"(f[4 * k + 2]), "=f"(f[4 * k + 3]) : "l"(p + 16 * k));
    }
    __device__ __forceinline__ void get(int q, cpx &x0, cpx &x1) const {
        x0 = cpx_make((double)f[4 * q], (double)f[4 * q + 1]);
        x1 = cpx_make((double)f[4 * q + 2], (double)f[4 * q + 3]);
    }
};

// One halo pair (the two samples before a warp's first): a single narrow load.
template <int FMT> struct IqHalo;
template <> struct IqHalo<PB_IQ_CPX> {
    cpx a, b;
    __device__ __forceinline__ void zero() { a = b = cpx_make(0.0, 0.0); }
    __device__ __forceinline__ void load(const char *p) {
        asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a.x), "=d"(a.y), "=d"(b.x), "=d"(b.y) : "l"(p));
    }
    __device__ __forceinline__ void get(cpx &x0, cpx &x1) const { x0 = a; x1 = b; }
};
template <> struct IqHalo<PB_IQ_CS16> {
    unsigned w0, w1;
    __device__ __forceinline__ void zero() { w0 = w1 = 0u; }
    __device__ __forceinline__ void load(const char *p) { asm volatile("ld.global.nc.v2.u32 {%0,%1}, [%2];" : "=r"(w0), "=r"(w1) : "l"(p)); }
    __device__ __forceinline__ void get(cpx &x0, cpx &x1) const { IqPair<PB_IQ_CS16>::cvt(w0, w1, x0, x1); }
};
template <int FMT8> struct IqHalo8 {
    unsigned w;
    __device__ __forceinline__ void zero() { w = FMT8 == PB_IQ_CU8 ? 0x80808080u : 0u; }
    __device__ __forceinline__ void load(const char *p) { asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(w) : "l"(p)); }
    __device__ __forceinline__ void get(cpx &x0, cpx &x1) const { IqPair<FMT8>::cvt(w, x0, x1); }
};
template <> struct IqHalo<PB_IQ_CU8> : IqHalo8<PB_IQ_CU8> {};
template <> struct IqHalo<PB_IQ_CS8> : IqHalo8<PB_IQ_CS8> {};
template <> struct IqHalo<PB_IQ_CF32> {
    float f0, f1, f2, f3;
    __device__ __forceinline__ void zero() { f0 = f1 = f2 = f3 = 0.f; }
    __device__ __forceinline__ void load(const char *p) {
        asm volatile("ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(f0), "=f"(f1), "=f"(f2), "=f"(f3) : "l"(p));
    }
    __device__ __forceinline__ void get(cpx &x0, cpx &x1) const { x0 = cpx_make((double)f0, (double)f1); x1 = cpx_make((double)f2, (double)f3); }
};
#endif  // __CUDACC__
