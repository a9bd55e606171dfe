// demod_kernels.cuh — device kernels of the demodulator tail (CFir, affine-scan AM, FM2, PLL, AGC), shared by
// demod.cu (single-stream operators) and channelizer.cu (batched channels). `static`: each translation unit gets its
// own copy (the library is built without relocatable device code).
#pragma once
#include "common.cuh"
#include "host_design.h"
#include <math_constants.h>
#define PB_FIR_MAX 75
#define PB_SCAN_THREADS 256
#define PB_SCAN_PER 8
#define PB_SCAN_TILE (PB_SCAN_THREADS * PB_SCAN_PER)

// ---- CFir ---------------------------------------------------------------------------------------
// virtual stream = hist (ntaps-1 samples) ++ in; out[i] = sum_k h[k] * xv[i + ntaps-1 - k]   (fir.cpp:85-130)
// One CTA filters a tile of PB_FIR_TILE outputs: the tile's window (tile + ntaps-1 samples) is staged in shared memory
// once (one pad slot per 8 so that threads 8 outputs apart read distinct banks), every thread produces PB_FIR_R
// CONSECUTIVE outputs from a register window that slides by one sample per tap (one 16-byte shared load per tap for
// 8 outputs); the taps are kernel parameters, read from the constant bank.
// The accumulation order over k is the reference's. CTA 0 also stores the carried history (ping-pong buffer).
#define PB_FIR_R 8
#define PB_FIR_THREADS 128
#define PB_FIR_TILE (PB_FIR_R * PB_FIR_THREADS)
struct FirTaps { double hI[PB_FIR_MAX], hQ[PB_FIR_MAX]; };
__device__ __forceinline__ int fir_slot(int p) { return p + (p >> 3); }

static __global__ void __launch_bounds__(PB_FIR_THREADS)
cfir_kernel(const cpx *__restrict__ hist, const cpx *__restrict__ in, int n, const __grid_constant__ FirTaps taps, int ntaps,
            cpx *__restrict__ out, cpx *__restrict__ new_hist) {
    __shared__ cpx win[PB_FIR_TILE + PB_FIR_MAX + (PB_FIR_TILE + PB_FIR_MAX) / 8 + 1];
    const int h = ntaps - 1;
    if (blockIdx.x == 0 && new_hist) {
        for (int i = threadIdx.x; i < h; i += PB_FIR_THREADS) {
            const int v = n + i;  // last h samples of hist ++ in
            new_hist[i] = v < h ? hist[v] : in[v - h];
        }
    }
    const int ntiles = (n + PB_FIR_TILE - 1) / PB_FIR_TILE;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int t0 = tile * PB_FIR_TILE;
        const int valid = min(PB_FIR_TILE, n - t0);
        __syncthreads();
        for (int p = threadIdx.x; p < valid + h; p += PB_FIR_THREADS) {
            const int v = t0 + p;
            win[fir_slot(p)] = v < h ? hist[v] : in[v - h];
        }
        __syncthreads();
        const int i0 = threadIdx.x * PB_FIR_R;
        if (i0 >= valid) continue;
        double re[PB_FIR_R], im[PB_FIR_R];
        cpx w[PB_FIR_R];
#pragma unroll
        for (int j = 0; j < PB_FIR_R; ++j) { re[j] = 0.0; im[j] = 0.0; w[j] = win[fir_slot(i0 + h + j)]; }
        // taps in groups of PB_FIR_R: after PB_FIR_R one-sample shifts the register window is back in place, so the
        // group body is unrolled with static register names while the loop over groups stays rolled (a full unroll of
        // 75 taps x 16 FMAs is ~25 KB of code and runs out of instruction cache)
        for (int k0 = 0; k0 < ntaps; k0 += PB_FIR_R) {
#pragma unroll
            for (int kk = 0; kk < PB_FIR_R; ++kk) {
                const int k = k0 + kk;
                if (k < ntaps) {
                    const double cI = taps.hI[k], cQ = taps.hQ[k];
#pragma unroll
                    for (int j = 0; j < PB_FIR_R; ++j) { re[j] += cI * w[j].x; im[j] += cQ * w[j].y; }
#pragma unroll
                    for (int j = PB_FIR_R - 1; j > 0; --j) w[j] = w[j - 1];
                    if (k + 1 < ntaps) w[0] = win[fir_slot(i0 + h - k - 1)];
                }
            }
        }
#pragma unroll
        for (int j = 0; j < PB_FIR_R; ++j)
            if (i0 + j < valid) out[t0 + i0 + j] = cpx_make(re[j], im[j]);
    }
}

// The same filter when the caller guarantees re == im for every input sample and the I and Q tap sets are equal (the AM
// detector's output (y, y) through its low-pass, demod_am.cpp:53-62): one real convolution, written to both components.
// Half the FP64 work and half the shared-memory window of cfir_kernel; bit-identical to it under those two conditions.
static __global__ void __launch_bounds__(PB_FIR_THREADS)
cfir_dup_kernel(const cpx *__restrict__ hist, const cpx *__restrict__ in, int n, const __grid_constant__ FirTaps taps, int ntaps,
                cpx *__restrict__ out, cpx *__restrict__ new_hist) {
    __shared__ double win[PB_FIR_TILE + PB_FIR_MAX + (PB_FIR_TILE + PB_FIR_MAX) / 8 + 1];
    const int h = ntaps - 1;
    if (blockIdx.x == 0 && new_hist) {
        for (int i = threadIdx.x; i < h; i += PB_FIR_THREADS) {
            const int v = n + i;
            new_hist[i] = v < h ? hist[v] : in[v - h];
        }
    }
    const int ntiles = (n + PB_FIR_TILE - 1) / PB_FIR_TILE;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int t0 = tile * PB_FIR_TILE;
        const int valid = min(PB_FIR_TILE, n - t0);
        __syncthreads();
        for (int p = threadIdx.x; p < valid + h; p += PB_FIR_THREADS) {
            const int v = t0 + p;
            win[fir_slot(p)] = v < h ? hist[v].x : in[v - h].x;
        }
        __syncthreads();
        const int i0 = threadIdx.x * PB_FIR_R;
        if (i0 >= valid) continue;
        double re[PB_FIR_R], w[PB_FIR_R];
#pragma unroll
        for (int j = 0; j < PB_FIR_R; ++j) { re[j] = 0.0; w[j] = win[fir_slot(i0 + h + j)]; }
        for (int k0 = 0; k0 < ntaps; k0 += PB_FIR_R) {
#pragma unroll
            for (int kk = 0; kk < PB_FIR_R; ++kk) {
                const int k = k0 + kk;
                if (k < ntaps) {
                    const double cI = taps.hI[k];
#pragma unroll
                    for (int j = 0; j < PB_FIR_R; ++j) re[j] += cI * w[j];
#pragma unroll
                    for (int j = PB_FIR_R - 1; j > 0; --j) w[j] = w[j - 1];
                    if (k + 1 < ntaps) w[0] = win[fir_slot(i0 + h - k - 1)];
                }
            }
        }
#pragma unroll
        for (int j = 0; j < PB_FIR_R; ++j)
            if (i0 + j < valid) out[t0 + i0 + j] = cpx_make(re[j], re[j]);
    }
}

// ---- affine scan helpers (d <- A*d + B composition) ---------------------------------------------------
struct Aff { double a, b; };
__device__ __forceinline__ Aff aff_then(Aff first, Aff second) {  // apply `first`, then `second`
    Aff r;
    r.a = first.a * second.a;
    r.b = second.a * first.b + second.b;
    return r;
}
// Exclusive scan of per-thread maps over the CTA (in thread order). Returns the composition of all
// lower-numbered threads' maps; *total (valid in every thread) = composition of all threads' maps.
template <int NT = PB_SCAN_THREADS>
__device__ __forceinline__ Aff block_aff_exclusive(Aff mine, Aff *total) {
    __shared__ Aff warp_tot[NT / 32];
    __shared__ Aff warp_pre[NT / 32 + 1];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    Aff inc = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        Aff o;
        o.a = __shfl_up_sync(0xffffffffu, inc.a, d);
        o.b = __shfl_up_sync(0xffffffffu, inc.b, d);
        if (lane >= d) inc = aff_then(o, inc);
    }
    if (lane == 31) warp_tot[w] = inc;
    __syncthreads();
    if (threadIdx.x == 0) {
        Aff run; run.a = 1.0; run.b = 0.0;
        for (int i = 0; i < NT / 32; ++i) { warp_pre[i] = run; run = aff_then(run, warp_tot[i]); }
        warp_pre[NT / 32] = run;
    }
    __syncthreads();
    Aff ex;
    ex.a = __shfl_up_sync(0xffffffffu, inc.a, 1);
    ex.b = __shfl_up_sync(0xffffffffu, inc.b, 1);
    if (lane == 0) { ex.a = 1.0; ex.b = 0.0; }
    *total = warp_pre[NT / 32];
    return aff_then(warp_pre[w], ex);
}

// ---- Demod_AM::processBlockFiltered, demod_am.cpp:40-59 ------------------------------------------------
// d_n = alpha*d_{n-1} + |x_n| ; y_n = d_n - d_{n-1} ; out = (y, y)
// pass 1: per-tile aggregate map; pass 2: serial carry over tiles; pass 3: apply.
static __global__ void __launch_bounds__(PB_SCAN_THREADS) am_tile_kernel(const cpx *__restrict__ in, int n, double alpha,
                                                                  Aff *__restrict__ tile_map) {
    const int base = blockIdx.x * PB_SCAN_TILE + threadIdx.x * PB_SCAN_PER;
    Aff m; m.a = 1.0; m.b = 0.0;
#pragma unroll
    for (int k = 0; k < PB_SCAN_PER; ++k) {
        if (base + k < n) {
            const cpx x = in[base + k];
            m.b = alpha * m.b + sqrt(x.x * x.x + x.y * x.y);  // DB::amplitude, db.h:28-30
            m.a *= alpha;
        }
    }
    Aff tot;
    block_aff_exclusive(m, &tot);
    if (threadIdx.x == 0) tile_map[blockIdx.x] = tot;
}
// Parallel carry (streams of more than PB_AM_DIRECT tiles): one CTA composes the tile maps chunk by chunk.
static __global__ void __launch_bounds__(PB_SCAN_THREADS) am_carry_kernel(const Aff *__restrict__ tile_map, int ntiles,
                                                                   const double *__restrict__ d_last_in,
                                                                   double *__restrict__ carry) {
    const int per = (ntiles + PB_SCAN_THREADS - 1) / PB_SCAN_THREADS;
    const int j0 = threadIdx.x * per, j1 = min(ntiles, j0 + per);
    Aff m; m.a = 1.0; m.b = 0.0;
    for (int j = j0; j < j1; ++j) m = aff_then(m, tile_map[j]);
    Aff tot;
    const Aff ex = block_aff_exclusive(m, &tot);
    double d = ex.a * (*d_last_in) + ex.b;
    for (int j = j0; j < j1; ++j) {
        carry[j] = d;
        d = tile_map[j].a * d + tile_map[j].b;
    }
}
// carry == nullptr: the CTA folds the maps of all earlier tiles itself (at most PB_AM_DIRECT of them), which
// removes the carry launch from the short streams the receiver tail sees (2^19 samples = 256 tiles).
#define PB_AM_DIRECT 2048   // 2^21 decimated samples = one 2^28-sample call of the 10 MSPS chain; <= 16 maps folded per thread
// NT threads x PB_SCAN_PER samples per tile: 256 x 8 = PB_SCAN_TILE for the stand-alone demodulator, 128 x 8 = one
// FastFIR block when the tile maps come from the FastFIR kernel's epilogue (receiver chain, AGC off).
template <int NT>
static __global__ void __launch_bounds__(NT) am_apply_kernel(const cpx *__restrict__ in, int n, double alpha,
                                                                   const double *__restrict__ carry,
                                                                   const Aff *__restrict__ tile_map,
                                                                   const double *__restrict__ d_last_in,
                                                                   cpx *__restrict__ out, double *__restrict__ d_last_out) {
    const int base = blockIdx.x * (NT * PB_SCAN_PER) + threadIdx.x * PB_SCAN_PER;
    double mag[PB_SCAN_PER];
    Aff m; m.a = 1.0; m.b = 0.0;
#pragma unroll
    for (int k = 0; k < PB_SCAN_PER; ++k) {
        mag[k] = 0.0;
        if (base + k < n) {
            const cpx x = in[base + k];
            mag[k] = sqrt(x.x * x.x + x.y * x.y);
            m.b = alpha * m.b + mag[k];
            m.a *= alpha;
        }
    }
    double d_tile;  // m_amDcLast entering this tile
    if (carry) {
        d_tile = carry[blockIdx.x];
    } else {
        const int nt = blockIdx.x, per = (nt + NT - 1) / NT;
        const int j0 = threadIdx.x * per, j1 = min(nt, j0 + per);
        Aff pm; pm.a = 1.0; pm.b = 0.0;
        for (int j = j0; j < j1; ++j) pm = aff_then(pm, tile_map[j]);
        Aff ptot;
        block_aff_exclusive<NT>(pm, &ptot);
        d_tile = ptot.a * (*d_last_in) + ptot.b;
    }
    Aff tot;
    const Aff ex = block_aff_exclusive<NT>(m, &tot);
    double d = ex.a * d_tile + ex.b;  // m_amDcLast entering this thread's first sample
#pragma unroll
    for (int k = 0; k < PB_SCAN_PER; ++k) {
        if (base + k < n) {
            const double dn = (alpha * d) + mag[k];  // demod_am.cpp:52
            const double y = dn - d;                 // demod_am.cpp:53
            d = dn;
            out[base + k] = cpx_make(y, y);
            if (base + k == n - 1) *d_last_out = d;
        }
    }
}

// ---- AGC_OFF / gain: out = g * in (agc.cpp:86-94, receiver.cpp:935-938) ---------------------------------
static __global__ void scale_kernel(const cpx *__restrict__ in, int n, double g, cpx *__restrict__ out) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[i] = cpx_scale(in[i], g);
}

// ---- Demod_NFM::processBlockFM2, demod_nfm.cpp:124-140 (phase difference) --------------------------------
__device__ __forceinline__ double phase_cpx(cpx c) {  // CpxUtil::phaseCpx, cpx.cpp:5-21
    double tmp = atan(c.y / ((c.x == 0) ? 1e-200 : c.x));
    if (c.x < 0 && c.y < 0) tmp -= 3.14159265358979323846;
    else if (c.x < 0 && c.y >= 0) tmp += 3.14159265358979323846;
    return tmp;
}
static __global__ void fm2_kernel(const cpx *__restrict__ in, int n, const cpx *__restrict__ last_in, cpx *__restrict__ out,
                           cpx *__restrict__ last_out) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const cpx prev = i ? in[i - 1] : *last_in;
        const cpx prod = cpx_mulc(in[i], prev);
        const double v = phase_cpx(prod) * .0005;
        out[i] = cpx_make(v, v);
        if (i == n - 1) *last_out = in[i];
    }
}

// ---- Demod_NFM::processBlockNCO, demod_nfm.cpp:225-257 (PLL; state and constants are `float`) ----------
struct PllConst { float alpha, beta, dc_alpha, lo, hi, out_gain; };
struct PllState { float phase, freq, dc; };
// The PLL recurrence over one stream (demod_nfm.cpp:225-257); state and constants are `float` like the reference's.
//
// The recurrence is a serial dependency chain, so a stream's throughput is that chain's LATENCY and nothing else (one thread
// per stream; ncu of the first version: 1 warp per SM, 975 cycles per sample, FP64 pipe 2.5 % busy). The reference's chain is
//     sin/cos(phase) -> rotate x by it (double) -> phzerror = -atan2(im, re) (double) -> freq += beta * phzerror -> clamp ->
//     phase += freq + alpha * phzerror,
// i.e. two transcendental evaluations per sample ON the chain. But arg(x e^{j phase}) = arg(x) + phase (mod 2 pi), and arg(x)
// does not depend on the loop state: theta_n = atan2f(x.im, x.re) is computed for a whole register block up front (independent
// instructions, pipelined), and the chain keeps only
//     r = phase reduced to [-pi, pi] (Cody-Waite, two FMAs: exact to the float's own ulp, like sinf/cosf's own reduction),
//     phzerror = -wrap(theta_n + r),  freq = clamp(fma(beta, phzerror, freq)),  phase += fma(alpha, phzerror, freq)
// — about a dozen dependent FP32 operations (~60 cycles) instead of ~700 cycles. The float phase itself is accumulated and
// wrapped exactly as the reference does (its growing rounding error inside a frame is part of the reference's behaviour), and
// the phase error agrees with the reference's to a few 1e-7 rad (float atan2f and the reduction, against float sinf/cosf +
// double rotation there). Measured relative RMS of the demodulated audio against the compiled reference: 6e-6 for the
// straight double transcription, the same for this form (both set by sinf/cosf implementation differences and the float
// state); tolerance 1e-4. atan2(0, 0) = 0 is kept for an exactly-zero sample.
#define PB_SERIAL_BLK 16
#define PB_TWO_PI_HI 6.28125f                  /* 9 trailing zero bits: k * HI is exact for k < 2^15 */
#define PB_TWO_PI_LO 1.9353071795864769e-3f    /* 2 pi - HI */
// arg(x) in float, branch-free (atan2f's special-case branches would serialise the block's 16 independent evaluations):
// odd minimax polynomial for atan on [0, 1] (max abs error 6.4e-8) + octant fix-up; 2-ulp division.
__device__ __forceinline__ float pll_theta(const cpx xi) {
    const float xr = (float)xi.x, xq = (float)xi.y;
    const float ax = fabsf(xr), ay = fabsf(xq);
    const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
    const float a = __fdividef(mn, mx);
    const float s = a * a;
    float r = 0.0027856871f;
    r = fmaf(r, s, -0.0158660002f);
    r = fmaf(r, s, 0.0424722321f);
    r = fmaf(r, s, -0.0749753043f);
    r = fmaf(r, s, 0.106448799f);
    r = fmaf(r, s, -0.142070308f);
    r = fmaf(r, s, 0.199934542f);
    r = fmaf(r, s, -0.333331466f);
    r = r * s;
    r = fmaf(r, a, a);
    r = ay > ax ? 1.57079632679489661923f - r : r;
    r = xr < 0.0f ? 3.14159265358979323846f - r : r;
    r = copysignf(r, xq);
    return mx == 0.0f ? CUDART_NAN_F : r;   // NaN marks "phzerror = 0" (atan2(0, 0) = 0 in the reference)
}
__device__ __forceinline__ void pll_step(const float theta, cpx &yo, const PllConst &c, PllState &v) {
    // k = nearest integer to phase / 2 pi by the 1.5 * 2^23 trick (|k| < 2^22): two FP32-pipe operations instead of FMUL + FRND
    const float k = fmaf(v.phase, 0.15915494309189535f, 12582912.0f) - 12582912.0f;
    float r = fmaf(k, -PB_TWO_PI_HI, v.phase);
    r = fmaf(k, -PB_TWO_PI_LO, r);
    float e = theta + r;                                   // arg(x e^{j phase}) in (-2 pi, 2 pi)
    e = e > 3.14159265358979323846f ? e - 6.28318530717958647692f : e;
    e = e < -3.14159265358979323846f ? e + 6.28318530717958647692f : e;
    const float phzerror = theta != theta ? 0.0f : -e;
    float f = fmaf(c.beta, phzerror, v.freq);
    f = fminf(fmaxf(f, c.lo), c.hi);
    v.freq = f;
    v.phase = v.phase + fmaf(c.alpha, phzerror, f);
    v.dc = fmaf(c.dc_alpha, f, (1.0f - c.dc_alpha) * v.dc);   // off the chain
    const float o = (f - v.dc) * c.out_gain;
    yo = cpx_make((double)o, 0.0);  // CPX = float: imaginary part is zero
}
// The serial recurrence walks its stream in register blocks of PB_SERIAL_BLK samples: the loads of a block (uncoalesced, one
// row per thread) and its arg(x) evaluations are independent and issued back to back.
// `frame`: the reference calls processBlockNCO once per receiver frame (framesPerBuffer samples, receiver.cpp:987) and wraps
// the float phase with fmod at the END of each call (demod_nfm.cpp:252); between wraps the phase keeps growing and the float
// rounding of the accumulation grows with it, so the wrap positions are part of the arithmetic. frame <= 0: one wrap at n.
__device__ __forceinline__ void pll_run(const cpx *__restrict__ x, cpx *__restrict__ y, int n, const PllConst c, PllState &v, int frame) {
    if (frame <= 0) frame = n;
    for (int f0 = 0; f0 < n; f0 += frame) {
        const int f1 = f0 + frame < n ? f0 + frame : n;
        int i = f0;
        // register double buffer: block b+1's loads (one row per thread: uncoalesced, L2 / DRAM latency) are in flight while
        // block b is computed (ncu of the un-prefetched form: 41 % of the stall samples on the first use of the loaded data)
        cpx cur[PB_SERIAL_BLK], nxt[PB_SERIAL_BLK];
        if (i + PB_SERIAL_BLK <= f1) {
#pragma unroll
            for (int j = 0; j < PB_SERIAL_BLK; ++j) cur[j] = x[i + j];
        }
        for (; i + PB_SERIAL_BLK <= f1; i += PB_SERIAL_BLK) {
            if (i + 2 * PB_SERIAL_BLK <= f1) {
#pragma unroll
                for (int j = 0; j < PB_SERIAL_BLK; ++j) nxt[j] = x[i + PB_SERIAL_BLK + j];
            }
            float th[PB_SERIAL_BLK];
#pragma unroll
            for (int j = 0; j < PB_SERIAL_BLK; ++j) th[j] = pll_theta(cur[j]);
#pragma unroll
            for (int j = 0; j < PB_SERIAL_BLK; ++j) {
                cpx yo;
                pll_step(th[j], yo, c, v);
                y[i + j] = yo;
            }
#pragma unroll
            for (int j = 0; j < PB_SERIAL_BLK; ++j) cur[j] = nxt[j];
        }
        for (; i < f1; ++i) {
            cpx yo;
            pll_step(pll_theta(x[i]), yo, c, v);
            y[i] = yo;
        }
        v.phase = (float)fmod((double)v.phase, 6.28318530717958647692528676656);
    }
}
// one thread per stream; stream s occupies in[s*stride .. s*stride + n)
static __global__ void pll_kernel(const cpx *__restrict__ in, int n, long long stride, int nstreams, PllConst c,
                                  PllState *__restrict__ st, cpx *__restrict__ out, int frame) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= nstreams) return;
    PllState v = st[s];
    pll_run(in + (long long)s * stride, out + (long long)s * stride, n, c, v, frame);
    st[s] = v;
}

// ---- Demod_SAM::processBlock, demod_sam.cpp:41-112 (synchronous AM: dttsp PLL + DC block per component) ---------------
// pll() (:41-79): z = (cos, sin)(phase) [float], pllSample = z * sig (double), difference = |sig| * phaseCpx(pllSample)
// -> float, frequency += beta * difference (clamped), phase += frequency + alpha * difference, wrapped into [0, 2 pi).
//
// Unlike dttsp's loop, the reference multiplies by z, NOT by conj(z): arg(pllSample) = arg(sig) + phase, and the phase is
// then pushed FURTHER in the direction of that error. The error therefore runs away from 0 and the loop settles in a
// chattering regime around +-pi (the atan2 cut), where every sample's sign decision kicks the phase by +-alpha*|sig|*pi.
// Between kicks the map phase -> phase*(1 + alpha*|sig|) is expanding, so two evaluations that differ in ONE float state
// bit decorrelate sample by sample within a few thousand samples: the reference's output is only reproducible by
// reproducing every rounding of the state. Hence no arg(x)+phase shortcut here (the NFM loop, a contracting map, has it):
// the chain is the reference's, operation by operation —
//   * cos/sin(float): the host libm's algorithm (glibc >= 2.28 = ARM optimized-routines sinf/cosf: one-step reduction by
//     pi/2 and two double polynomials, result rounded to float), restated in sam_sincosf below; that restatement was
//     checked bit for bit against the libm of this image for EVERY float in [0, 2 pi] (1 086 918 620 values, 0 differences);
//   * z * sig, |sig|, phaseCpx (atan of the quotient + quadrant fix-up, cpx.cpp:5-21) in double with separately rounded
//     products and sums (no FMA contraction, like the reference's x86-64 build); CUDA's atan is within 2 ulp of the libm's,
//     which changes the float `difference` about once in 3e8 samples;
//   * the float updates with their own roundings (a float product, then a float sum), the wraps in double (TWOPI is a double).
// Cost: ~1 us per sample on one thread (13x real time at 78 125 S/s) — this demodulator is one stream of audio-rate work.
struct SamConst { float alpha, beta, lo, hi; };
struct SamState { float phase, freq; float pad0, pad1; double dc_re_last, dc_im_last; };
__device__ __forceinline__ double sam_sin_poly(double x, double x2) {
    const double x3 = __dmul_rn(x, x2);
    const double s1 = __dadd_rn(0x1.1107605230bc4p-7, __dmul_rn(x2, -0x1.994eb3774cf24p-13));
    const double x7 = __dmul_rn(x3, x2);
    const double sv = __dadd_rn(x, __dmul_rn(x3, -0x1.555545995a603p-3));
    return __dadd_rn(sv, __dmul_rn(x7, s1));
}
__device__ __forceinline__ double sam_cos_poly(double x2) {
    const double x4 = __dmul_rn(x2, x2);
    const double c2 = __dadd_rn(-0x1.6c087e89a359dp-10, __dmul_rn(x2, 0x1.99343027bf8c3p-16));
    const double c1 = __dadd_rn(1.0, __dmul_rn(x2, -0x1.ffffffd0c621cp-2));
    const double x6 = __dmul_rn(x4, x2);
    const double cv = __dadd_rn(c1, __dmul_rn(x4, 0x1.55553e1068f19p-5));
    return __dadd_rn(cv, __dmul_rn(x6, c2));
}
// (float)cos(y), (float)sin(y) for 0 <= y < 120 exactly as the host libm returns them
__device__ __forceinline__ void sam_sincosf(float y, float &sn, float &cs) {
    const double x = (double)y;
    const unsigned top = (__float_as_uint(y) >> 20) & 0x7ff;
    if (top < 0x3f4) {                      // |y| < pi/4
        const double x2 = __dmul_rn(x, x);
        if (top < 0x398) { sn = y; cs = 1.0f; return; }     // |y| < 2^-12
        sn = (float)sam_sin_poly(x, x2);
        cs = (float)sam_cos_poly(x2);
        return;
    }
    const double r = __dmul_rn(x, 0x1.45F306DC9C883p+23);                    // 2/pi * 2^24
    const int n = (__double2int_rz(r) + 0x800000) >> 24;
    const double xr = __dsub_rn(x, __dmul_rn((double)n, 0x1.921FB54442D18p0));
    const double x2 = __dmul_rn(xr, xr);
    const double sg = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;           // sign of sine in quadrant n
    const double sp = sam_sin_poly(__dmul_rn(xr, sg), x2), cp = sam_cos_poly(x2);
    // sinf: even n -> sine polynomial of (x * sign), odd n -> cosine polynomial, negated in quadrants 2, 3
    sn = (n & 1) ? (float)((n & 2) ? -cp : cp) : (float)sp;
    // cosf: odd n -> sine polynomial of (x * sign), even n -> cosine polynomial, negated when (n + 1) & 2
    cs = (n & 1) ? (float)sp : (float)(((n + 1) & 2) ? -cp : cp);
}
// CpxUtil::phaseCpx (cpx.cpp:5-21): atan(im / re) with re == 0 replaced by 1e-200, +-pi in the left half plane
__device__ __forceinline__ double sam_phase(double re, double im) {
    double t = atan(__ddiv_rn(im, re == 0.0 ? 1e-200 : re));
    if (re < 0.0 && im < 0.0) t = __dsub_rn(t, 3.14159265358979323846);
    else if (re < 0.0 && im >= 0.0) t = __dadd_rn(t, 3.14159265358979323846);
    return t;
}
static __global__ void sam_kernel(const cpx *__restrict__ in, int n, SamConst c, double dc_alpha, SamState *__restrict__ st,
                                  cpx *__restrict__ out) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    SamState v = *st;
    cpx cur[PB_SERIAL_BLK], nxt[PB_SERIAL_BLK];
#pragma unroll
    for (int j = 0; j < PB_SERIAL_BLK; ++j) cur[j] = j < n ? in[j] : cpx_make(0.0, 0.0);
    for (int i0 = 0; i0 < n; i0 += PB_SERIAL_BLK) {
#pragma unroll
        for (int j = 0; j < PB_SERIAL_BLK; ++j) nxt[j] = i0 + PB_SERIAL_BLK + j < n ? in[i0 + PB_SERIAL_BLK + j] : cpx_make(0.0, 0.0);
        double mg[PB_SERIAL_BLK];
#pragma unroll
        for (int j = 0; j < PB_SERIAL_BLK; ++j)      // magCpx (cpx.h:228-238), off the chain
            mg[j] = sqrt(__dadd_rn(__dmul_rn(cur[j].x, cur[j].x), __dmul_rn(cur[j].y, cur[j].y)));
        cpx pl[PB_SERIAL_BLK];
#pragma unroll
        for (int j = 0; j < PB_SERIAL_BLK; ++j) {
            if (i0 + j >= n) break;
            float sn, cs;
            sam_sincosf(v.phase, sn, cs);
            const double zr = (double)cs, zi = (double)sn;
            const double pr = __dsub_rn(__dmul_rn(zr, cur[j].x), __dmul_rn(zi, cur[j].y));      // z * sig
            const double pi = __dadd_rn(__dmul_rn(zr, cur[j].y), __dmul_rn(zi, cur[j].x));
            pl[j] = cpx_make(pr, pi);
            const float diff = (float)__dmul_rn(mg[j], sam_phase(pr, pi));                      // :58
            float f = __fadd_rn(v.freq, __fmul_rn(c.beta, diff));                               // :61
            f = f < c.lo ? c.lo : f;                                                            // :64-67
            f = f > c.hi ? c.hi : f;
            v.freq = f;
            float p = __fadd_rn(v.phase, __fadd_rn(f, __fmul_rn(c.alpha, diff)));               // :69
            while ((double)p >= 6.28318530717958647692528676656) p = (float)((double)p - 6.28318530717958647692528676656);   // :72-75
            while ((double)p < 0.0) p = (float)((double)p + 6.28318530717958647692528676656);
            v.phase = p;
        }
#pragma unroll
        for (int j = 0; j < PB_SERIAL_BLK; ++j) {
            if (i0 + j < n) {
                const double dre = __dadd_rn(__dmul_rn(dc_alpha, v.dc_re_last), pl[j].x);       // :97-98
                const double dim = __dadd_rn(__dmul_rn(dc_alpha, v.dc_im_last), pl[j].y);
                out[i0 + j] = cpx_make(__dsub_rn(dre, v.dc_re_last), __dsub_rn(dim, v.dc_im_last));
                v.dc_re_last = dre; v.dc_im_last = dim;
            }
        }
#pragma unroll
        for (int j = 0; j < PB_SERIAL_BLK; ++j) cur[j] = nxt[j];
    }
    *st = v;
}
// out = (re + im, re - im): lower sideband left, upper sideband right (demod_sam.cpp:105-111)
static __global__ void sam_mix_kernel(cpx *__restrict__ io, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { const cpx t = io[i]; io[i] = cpx_make(t.x + t.y, t.x - t.y); }
}

// ---- AGC::processBlock, agc.cpp:84-207 (modes other than OFF) --------------------------------------------
#define PB_AGC_BUF 2048  // MAX_DELAY_BUF, agc.h:27
struct AgcState {
    int sig_ptr, mag_pos, hang_timer, pad;
    double peak, attack_avg, decay_avg;
};
static __global__ void agc_kernel(const cpx *__restrict__ in, int n, long long stride, int nstreams, pebble_host::AgcParams p,
                           AgcState *__restrict__ st, cpx *__restrict__ sig_buf, double *__restrict__ mag_buf,
                           cpx *__restrict__ out) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= nstreams) return;
    AgcState v = st[s];
    cpx *sig = sig_buf + (long long)s * PB_AGC_BUF;
    double *mb = mag_buf + (long long)s * PB_AGC_BUF;
    const cpx *x = in + (long long)s * stride;
    cpx *y = out + (long long)s * stride;
    const double min_const = (double)1e-8f;   // MIN_CONSTANT is a float constant, agc.h:58
    const double outscale = (double)0.7f;     // AGC_OUTSCALE, agc.h:49
    for (int i = 0; i < n; ++i) {
        const cpx xin = x[i];
        const cpx delayed = sig[v.sig_ptr];
        sig[v.sig_ptr++] = xin;
        if (v.sig_ptr >= p.delay_samples) v.sig_ptr = 0;
        double mag = fabs(xin.x);
        const double mim = fabs(xin.y);
        if (mim > mag) mag = mim;
        mag = log10(mag + min_const);  // - log10(MAX_AMPLITUDE) == 0
        double tmp = mb[v.mag_pos];
        mb[v.mag_pos++] = mag;
        if (v.mag_pos >= p.window_samples) v.mag_pos = 0;
        if (mag > v.peak) {
            v.peak = mag;
        } else if (tmp == v.peak) {
            v.peak = -8.0;
            for (int k = 0; k < p.window_samples; ++k) {
                tmp = mb[k];
                if (tmp > v.peak) v.peak = tmp;
            }
        }
        if (v.peak > v.attack_avg) v.attack_avg = (1.0 - p.attack_rise) * v.attack_avg + p.attack_rise * v.peak;
        else v.attack_avg = (1.0 - p.attack_fall) * v.attack_avg + p.attack_fall * v.peak;
        if (p.use_hang) {
            if (v.peak > v.decay_avg) {
                v.decay_avg = (1.0 - p.decay_rise) * v.decay_avg + p.decay_rise * v.peak;
                v.hang_timer = 0;
            } else if (v.hang_timer < p.hang_time) {
                v.hang_timer++;
            } else {
                v.decay_avg = (1.0 - p.decay_fall) * v.decay_avg + p.decay_fall * v.peak;
            }
        } else {
            if (v.peak > v.decay_avg) v.decay_avg = (1.0 - p.decay_rise) * v.decay_avg + p.decay_rise * v.peak;
            else v.decay_avg = (1.0 - p.decay_fall) * v.decay_avg + p.decay_fall * v.peak;
        }
        mag = v.attack_avg > v.decay_avg ? v.attack_avg : v.decay_avg;
        double gain;
        if (mag <= p.knee) gain = p.fixed_gain;
        else gain = outscale * pow(10.0, mag * (p.gain_slope - 1.0));
        y[i] = cpx_make(delayed.x * gain, delayed.y * gain);
    }
    st[s] = v;
}

// ---- WFM (demod_wfm.cpp), shared by demod.cu (pilot PLL) and rds.cu (RDS PLL) -----------------------------------------
// Demod_WFM::arctan2 (:792-818), the reference's own approximation ("|error| < 0.005"; its |z| >= 1 branch returns angles near
// 2 pi — kept as written: the loop's behaviour far from lock depends on it)
__device__ __forceinline__ double wfm_arctan2(double y, double x) {
    const double TWOPI = 6.28318530717958647692528676656, PI = 3.14159265358979323846264338328;
    if (x == 0.0) return y > 0.0 ? TWOPI : (y == 0.0 ? 0.0 : -TWOPI);
    const double z = __ddiv_rn(y, x);
    double angle;
    if (fabs(z) < 1.0) {
        angle = __ddiv_rn(z, __dadd_rn(1.0, __dmul_rn(__dmul_rn(0.2854, z), z)));
        if (x < 0.0) return y < 0.0 ? __dsub_rn(angle, PI) : __dadd_rn(angle, PI);
    } else {
        angle = __dsub_rn(TWOPI, __ddiv_rn(z, __dadd_rn(__dmul_rn(z, z), 0.2854)));
        if (y < 0.0) return __dsub_rn(angle, PI);
    }
    return angle;
}
