// fft_core.cuh — register-radix butterflies and in-place shared-memory FFT passes (FP64).
//
// Replaces the FFT backend behind pebblelib/fft.h:37-42 (FFTfftw::fftForward/fftInverse,
// fftw.cpp:31-67: forward = e^{-j2pi nk/N}, inverse = e^{+j...}, both unnormalised like FFTW).
//
// Layout: one transform lives in shared memory as N cpx (16 B each) and is transformed IN PLACE by
// radix-R decimation-in-frequency passes; each thread pulls R points into registers, runs the
// R-point DFT, applies the inter-pass twiddle and writes back to the same slots. After the last
// pass, slot p holds X[freq_of_pos(p)] (mixed-radix digit reversal). The inverse runs the mirrored
// decimation-in-time passes in reverse order, so forward -> pointwise multiply -> inverse needs no
// reordering at all (that is the CFastFIR hot loop, fastfir.cpp:305-307).
// Bank conflicts: slot index i is stored at fft_swz<SH>(i) (two XOR terms, see below); every pass and the natural-order
// epilogues then touch 8 distinct 16-byte lanes per quarter warp.
#pragma once
#include "common.cuh"

#define PB_SQRT1_2 0.70710678118654752440
#define PB_COS_PI_8 0.92387953251128675613
#define PB_SIN_PI_8 0.38268343236508977173

template <bool INV> PB_HD void dft2(cpx &a, cpx &b) {
    cpx t = cpx_sub(a, b);
    a = cpx_add(a, b);
    b = t;
}

// 4-point DFT, natural order in and out.
template <bool INV> PB_HD void dft4(cpx &a0, cpx &a1, cpx &a2, cpx &a3) {
    cpx s02 = cpx_add(a0, a2), d02 = cpx_sub(a0, a2);
    cpx s13 = cpx_add(a1, a3), d13 = cpx_mulj<INV>(cpx_sub(a1, a3));
    a0 = cpx_add(s02, s13);
    a2 = cpx_sub(s02, s13);
    a1 = cpx_add(d02, d13);
    a3 = cpx_sub(d02, d13);
}

// multiply by W8^1 = (1 -/+ j)/sqrt2 and W8^3 = (-1 -/+ j)/sqrt2
template <bool INV> PB_HD cpx mul_w8_1(cpx a) {
    return INV ? make_double2((a.x - a.y) * PB_SQRT1_2, (a.x + a.y) * PB_SQRT1_2)
               : make_double2((a.x + a.y) * PB_SQRT1_2, (a.y - a.x) * PB_SQRT1_2);
}
template <bool INV> PB_HD cpx mul_w8_3(cpx a) {
    return INV ? make_double2((-a.x - a.y) * PB_SQRT1_2, (a.x - a.y) * PB_SQRT1_2)
               : make_double2((a.y - a.x) * PB_SQRT1_2, (-a.x - a.y) * PB_SQRT1_2);
}

// 8-point DFT, natural order in and out (v[q] = X[q]).
template <bool INV> PB_HD void dft8(cpx (&v)[8]) {
    cpx e0 = v[0], e1 = v[2], e2 = v[4], e3 = v[6];
    cpx o0 = v[1], o1 = v[3], o2 = v[5], o3 = v[7];
    dft4<INV>(e0, e1, e2, e3);
    dft4<INV>(o0, o1, o2, o3);
    o1 = mul_w8_1<INV>(o1);
    o2 = cpx_mulj<INV>(o2);
    o3 = mul_w8_3<INV>(o3);
    v[0] = cpx_add(e0, o0); v[4] = cpx_sub(e0, o0);
    v[1] = cpx_add(e1, o1); v[5] = cpx_sub(e1, o1);
    v[2] = cpx_add(e2, o2); v[6] = cpx_sub(e2, o2);
    v[3] = cpx_add(e3, o3); v[7] = cpx_sub(e3, o3);
}

// W16^k = cos(k pi/8) -/+ j sin(k pi/8)
template <bool INV> PB_HD cpx mul_w16(cpx a, double c, double s) {
    // forward: a * (c - j s); inverse: a * (c + j s)
    return INV ? make_double2(a.x * c - a.y * s, a.y * c + a.x * s)
               : make_double2(a.x * c + a.y * s, a.y * c - a.x * s);
}

// 16-point DFT, natural order in and out.
template <bool INV> PB_HD void dft16(cpx (&v)[16]) {
    cpx e[8], o[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { e[k] = v[2 * k]; o[k] = v[2 * k + 1]; }
    dft8<INV>(e);
    dft8<INV>(o);
    o[1] = mul_w16<INV>(o[1], PB_COS_PI_8, PB_SIN_PI_8);
    o[2] = mul_w8_1<INV>(o[2]);
    o[3] = mul_w16<INV>(o[3], PB_SIN_PI_8, PB_COS_PI_8);
    o[4] = cpx_mulj<INV>(o[4]);
    o[5] = mul_w16<INV>(o[5], -PB_SIN_PI_8, PB_COS_PI_8);
    o[6] = mul_w8_3<INV>(o[6]);
    o[7] = mul_w16<INV>(o[7], -PB_COS_PI_8, PB_SIN_PI_8);
#pragma unroll
    for (int k = 0; k < 8; ++k) { v[k] = cpx_add(e[k], o[k]); v[k + 8] = cpx_sub(e[k], o[k]); }
}

template <int R, bool INV> struct DftR;
template <bool INV> struct DftR<4, INV> {
    static PB_HD void run(cpx (&v)[4]) { dft4<INV>(v[0], v[1], v[2], v[3]); }
};
template <bool INV> struct DftR<8, INV> { static PB_HD void run(cpx (&v)[8]) { dft8<INV>(v); } };
template <bool INV> struct DftR<16, INV> { static PB_HD void run(cpx (&v)[16]) { dft16<INV>(v); } };

// SH encodes two shifts, SH = s1 + 100 * s2: slot i is stored at i ^ ((i >> s1) & 7) ^ ((i >> s2) & 7). s1 = log2(last radix)
// makes the last pass (R contiguous slots per thread) conflict-free; s2 = log2(N / first radix) does the same for reads in
// NATURAL bin order (consecutive bins sit N / R0 slots apart after digit reversal: the spectrum epilogue and the
// natural-order copy-out were 8-way conflicted without it, ncu: 41 % of all shared wavefronts). Both terms are constant
// inside every aligned group of 8 consecutive slots, so the passes that walk consecutive slots stay conflict-free.
template <int SH> PB_HD int fft_swz(int i) {
    constexpr int s1 = SH % 100, s2 = SH / 100;
    return s2 ? (i ^ ((i >> s1) & 7) ^ ((i >> s2) & 7)) : (i ^ ((i >> s1) & 7));
}

// Twiddle powers w^1..w^(R-1) of one butterfly from FOUR table reads (w^1, w^2, w^4, w^8) and R-5 complex
// multiplies (depth <= 3), instead of R-1 scattered 16-byte gathers: the gathers kept the L1/LSU data pipe (which the
// shared-memory passes already saturate) and the long-scoreboard busy; FP64 had room (ncu: 23 % of the pipe).
// tw[m] = exp(-j 2 pi m / N); `step` = i * (N / M). Powers beyond R-1 are not formed.
template <int R> PB_HD void fft_twiddle_powers(const cpx *tw, int step, int nmask, cpx (&w)[16]) {
    w[1] = tw[step & nmask];
    w[2] = tw[(2 * step) & nmask];
    w[3] = cpx_mul(w[1], w[2]);
    if (R > 4) {
        w[4] = tw[(4 * step) & nmask];
        w[5] = cpx_mul(w[1], w[4]);
        w[6] = cpx_mul(w[2], w[4]);
        w[7] = cpx_mul(w[3], w[4]);
    }
    if (R > 8) {
        w[8] = tw[(8 * step) & nmask];
#pragma unroll
        for (int q = 9; q < 16; ++q) w[q] = cpx_mul(w[q - 8], w[8]);
    }
}

// One radix-R butterfly of the in-place pass whose current sub-transform length is M (stride M/R).
// tw[m] = exp(-j 2 pi m / N), m < N. Forward: DFT then twiddle; inverse: conj twiddle then DFT.
template <int N, int R, int M, bool INV, int SH>
PB_HD void fft_butterfly(cpx *a, const cpx *tw, int bf) {
    constexpr int S = M / R;
    const int b = bf / S, i = bf - b * S;
    const int base = b * M + i;
    cpx v[R];
    cpx w[16];
#pragma unroll
    for (int k = 0; k < R; ++k) v[k] = a[fft_swz<SH>(base + k * S)];
    if (S > 1) fft_twiddle_powers<R>(tw, i * (N / M), N - 1, w);
    if (INV && S > 1) {
#pragma unroll
        for (int q = 1; q < R; ++q) v[q] = cpx_mulc(v[q], w[q]);
    }
    DftR<R, INV>::run(v);
    if (!INV && S > 1) {
#pragma unroll
        for (int q = 1; q < R; ++q) v[q] = cpx_mul(v[q], w[q]);
    }
#pragma unroll
    for (int k = 0; k < R; ++k) a[fft_swz<SH>(base + k * S)] = v[k];
}

// All butterflies of one pass, strided over the CTA's threads. Caller puts a barrier between passes.
template <int N, int R, int M, bool INV, int SH>
PB_HD void fft_pass(cpx *a, const cpx *tw, int tid, int nthreads) {
    for (int bf = tid; bf < N / R; bf += nthreads) fft_butterfly<N, R, M, INV, SH>(a, tw, bf);
}

// Per-size plans. pass<P,INV>() runs forward pass P (P = 0..NP-1); the inverse transform calls the
// same P in the order NP-1..0 with INV=true. pos2freq/freq2pos give the digit-reversed slot <-> bin map.
template <int N> struct FftPlan;

template <> struct FftPlan<1024> {   // 16 x 8 x 8 (the polyphase channelizer's branch DFT, chz_poly.cu)
    static constexpr int NP = 3, SH = 3 + 100 * 6, THREADS = 128;
    static constexpr int R0 = 16;   // radix of pass 0 (sub-transform length N)
    template <int P, bool INV> static PB_HD void pass(cpx *a, const cpx *tw, int tid, int nt) {
        if (P == 0) fft_pass<1024, 16, 1024, INV, SH>(a, tw, tid, nt);
        if (P == 1) fft_pass<1024, 8, 64, INV, SH>(a, tw, tid, nt);
        if (P == 2) fft_pass<1024, 8, 8, INV, SH>(a, tw, tid, nt);
    }
    static PB_HD int pos2freq(int p) { return (p >> 6) + 16 * ((p >> 3) & 7) + 128 * (p & 7); }
    static PB_HD int freq2pos(int f) { return ((f & 15) << 6) + (((f >> 4) & 7) << 3) + (f >> 7); }
};

template <> struct FftPlan<2048> {
    static constexpr int NP = 3, SH = 3 + 100 * 7, THREADS = 128;
    static constexpr int R0 = 16;   // radix of pass 0 (sub-transform length N)
    template <int P, bool INV> static PB_HD void pass(cpx *a, const cpx *tw, int tid, int nt) {
        if (P == 0) fft_pass<2048, 16, 2048, INV, SH>(a, tw, tid, nt);
        if (P == 1) fft_pass<2048, 16, 128, INV, SH>(a, tw, tid, nt);
        if (P == 2) fft_pass<2048, 8, 8, INV, SH>(a, tw, tid, nt);
    }
    static PB_HD int pos2freq(int p) { return (p >> 7) + 16 * ((p >> 3) & 15) + 256 * (p & 7); }
    static PB_HD int freq2pos(int f) { return ((f & 15) << 7) + (((f >> 4) & 15) << 3) + (f >> 8); }
};

template <> struct FftPlan<4096> {
    static constexpr int NP = 3, SH = 4 + 100 * 8, THREADS = 256;
    static constexpr int R0 = 16;   // radix of pass 0 (sub-transform length N)
    template <int P, bool INV> static PB_HD void pass(cpx *a, const cpx *tw, int tid, int nt) {
        if (P == 0) fft_pass<4096, 16, 4096, INV, SH>(a, tw, tid, nt);
        if (P == 1) fft_pass<4096, 16, 256, INV, SH>(a, tw, tid, nt);
        if (P == 2) fft_pass<4096, 16, 16, INV, SH>(a, tw, tid, nt);
    }
    static PB_HD int pos2freq(int p) { return (p >> 8) + 16 * ((p >> 4) & 15) + 256 * (p & 15); }
    static PB_HD int freq2pos(int f) { return ((f & 15) << 8) + (((f >> 4) & 15) << 4) + (f >> 8); }
};

template <> struct FftPlan<8192> {
    static constexpr int NP = 4, SH = 4 + 100 * 10, THREADS = 512;
    static constexpr int R0 = 8;   // radix of pass 0 (sub-transform length N)
    template <int P, bool INV> static PB_HD void pass(cpx *a, const cpx *tw, int tid, int nt) {
        if (P == 0) fft_pass<8192, 8, 8192, INV, SH>(a, tw, tid, nt);
        if (P == 1) fft_pass<8192, 8, 1024, INV, SH>(a, tw, tid, nt);
        if (P == 2) fft_pass<8192, 8, 128, INV, SH>(a, tw, tid, nt);
        if (P == 3) fft_pass<8192, 16, 16, INV, SH>(a, tw, tid, nt);
    }
    static PB_HD int pos2freq(int p) {
        return (p >> 10) + 8 * ((p >> 7) & 7) + 64 * ((p >> 4) & 7) + 512 * (p & 15);
    }
    static PB_HD int freq2pos(int f) {
        return ((f & 7) << 10) + (((f >> 3) & 7) << 7) + (((f >> 6) & 7) << 4) + (f >> 9);
    }
};
