// ff_fused.cuh — building blocks of the register-fused 2048-point pass schedule (8 x 16 x 16), shared by
// fastfir2048_kernel / fastfir4096x3_kernel (fastfir.cu) and the long filter's row kernel (fastfir_long.cu).
// Slot p = 256 q0 + 16 q1 + q2 holds bin q0 + 8 q1 + 128 q2 (ff2048_pos2freq).
#pragma once
#include "fft_core.cuh"

__device__ __forceinline__ int ff_swz(int p) { return p ^ ((p >> 4) & 7); }
PB_HD int ff2048_pos2freq(int p) { return (p >> 8) + 8 * ((p >> 4) & 15) + 128 * (p & 15); }

template <bool INV> __device__ __forceinline__ void ff_tw8(cpx (&u)[8], cpx w1, cpx w2, cpx w4) {
    const cpx w3 = cpx_mul(w1, w2), w5 = cpx_mul(w1, w4), w6 = cpx_mul(w2, w4), w7 = cpx_mul(w3, w4);
    if (INV) {
        u[1] = cpx_mulc(u[1], w1); u[2] = cpx_mulc(u[2], w2); u[3] = cpx_mulc(u[3], w3); u[4] = cpx_mulc(u[4], w4);
        u[5] = cpx_mulc(u[5], w5); u[6] = cpx_mulc(u[6], w6); u[7] = cpx_mulc(u[7], w7);
    } else {
        u[1] = cpx_mul(u[1], w1); u[2] = cpx_mul(u[2], w2); u[3] = cpx_mul(u[3], w3); u[4] = cpx_mul(u[4], w4);
        u[5] = cpx_mul(u[5], w5); u[6] = cpx_mul(u[6], w6); u[7] = cpx_mul(u[7], w7);
    }
}

// v[q] *= w^q (forward) or conj(w^q) (inverse), q = 1..15, from w, w^2, w^4, w^8: 11 extra complex multiplies instead
// of 11 more table reads (the L1/shared data pipe is the busier one, see above).
template <bool INV> __device__ __forceinline__ cpx ff_twmul(cpx v, cpx w) { return INV ? cpx_mulc(v, w) : cpx_mul(v, w); }
template <bool INV> __device__ __forceinline__ void ff_tw16(cpx (&v)[16], cpx w1, cpx w2, cpx w4, cpx w8) {
    const cpx w3 = cpx_mul(w1, w2), w5 = cpx_mul(w1, w4), w6 = cpx_mul(w2, w4), w7 = cpx_mul(w3, w4);
    v[1] = ff_twmul<INV>(v[1], w1); v[2] = ff_twmul<INV>(v[2], w2); v[3] = ff_twmul<INV>(v[3], w3);
    v[4] = ff_twmul<INV>(v[4], w4); v[5] = ff_twmul<INV>(v[5], w5); v[6] = ff_twmul<INV>(v[6], w6);
    v[7] = ff_twmul<INV>(v[7], w7); v[8] = ff_twmul<INV>(v[8], w8);
    v[9] = ff_twmul<INV>(v[9], cpx_mul(w1, w8)); v[10] = ff_twmul<INV>(v[10], cpx_mul(w2, w8));
    v[11] = ff_twmul<INV>(v[11], cpx_mul(w3, w8)); v[12] = ff_twmul<INV>(v[12], cpx_mul(w4, w8));
    v[13] = ff_twmul<INV>(v[13], cpx_mul(w5, w8)); v[14] = ff_twmul<INV>(v[14], cpx_mul(w6, w8));
    v[15] = ff_twmul<INV>(v[15], cpx_mul(w7, w8));
}

// The three middle phases, shared by the 2048 (T = 128 threads) and 4096 (T = 256) kernels: every thread owns one
// radix-16 butterfly per phase. a: T*16 slots; t1[16 b + i] = exp(-j 2 pi i 2^b / 256); ht[T k + t] = H of slot 16 t + k.
// All three phases of thread t stay inside the 256-slot block t/16, which is written and read by the SAME half warp
// (F1 butterflies of block j are threads 16 j .. 16 j + 15, the MID groups of those threads are slots 256 j .. 256 j + 255),
// so the phases are separated by __syncwarp() only: the CTA-wide barriers are the ones around F0 and I0.
template <int T, bool NAMED>
__device__ __forceinline__ void ff_mid_phases(cpx *a, const cpx *t1, const cpx *__restrict__ ht, int t, int bar) {
    const int i1 = t & 15, base1 = (t >> 4) * 256 + i1;
    cpx v[16];
    // ---- F1
#pragma unroll
    for (int k = 0; k < 16; ++k) v[k] = a[ff_swz(base1 + 16 * k)];
    dft16<false>(v);
    ff_tw16<false>(v, t1[i1], t1[16 + i1], t1[32 + i1], t1[48 + i1]);
#pragma unroll
    for (int k = 0; k < 16; ++k) a[ff_swz(base1 + 16 * k)] = v[k];
    __syncwarp();
    // ---- MID: forward radix-16, times H (CpxMpy, fastfir.cpp:306,324-334), inverse radix-16 on 16 contiguous slots
#pragma unroll
    for (int k = 0; k < 16; ++k) v[k] = a[ff_swz(16 * t + k)];
    dft16<false>(v);
#pragma unroll
    for (int k = 0; k < 16; ++k) v[k] = cpx_mul(ht[T * k + t], v[k]);
    dft16<true>(v);
#pragma unroll
    for (int k = 0; k < 16; ++k) a[ff_swz(16 * t + k)] = v[k];
    __syncwarp();
    // ---- I1
#pragma unroll
    for (int k = 0; k < 16; ++k) v[k] = a[ff_swz(base1 + 16 * k)];
    ff_tw16<true>(v, t1[i1], t1[16 + i1], t1[32 + i1], t1[48 + i1]);
    dft16<true>(v);
#pragma unroll
    for (int k = 0; k < 16; ++k) a[ff_swz(base1 + 16 * k)] = v[k];
    if (NAMED) asm volatile("bar.sync %0, %1;" ::"r"(bar), "n"(T) : "memory");
    else __syncthreads();
}

