// fastfir.cu — CFastFIR (overlap-save FFT band-pass) on the GPU + its C ABI.
// Reference: pebblelib/fastfir.cpp:77-334. One CTA per overlap-save block: 2L = N samples (L old + L new)
// are staged into shared memory, forward FFT, multiplied by the filter spectrum H (kept in the same
// digit-reversed slot order the forward passes produce), inverse FFT, and the upper L samples written.
// The 1/N of the unnormalised FFTW inverse is folded into H by the reference itself (fastfir.cpp:243).
#include "lib_common.h"
#include "fft_kernels.cuh"
#include "host_design.h"
#include "fastfir_long.cuh"
#include "ff_fused.cuh"
#include <cstdlib>
#include <cstring>
#include <vector>

// The "virtual stream" of a call is state ++ input, state = overlap (L) + not yet processed samples.
__device__ __forceinline__ cpx vs_load(const cpx *__restrict__ state, int s0, const cpx *__restrict__ in, long long v) {
    return v < s0 ? state[v] : in[v - s0];
}

// Streaming variants for the fused kernels: samples are read once (twice with the overlap) and must not evict the filter
// spectrum H from L1 (ncu: 18 % L1 hit rate on the H loads when the samples allocate); outputs are write-once.
__device__ __forceinline__ cpx ff_ld_stream(const cpx *p) {
    cpx r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
    return r;
}
__device__ __forceinline__ cpx vs_load_stream(const cpx *__restrict__ state, int s0, const cpx *__restrict__ in, long long v) {
    return ff_ld_stream(v < s0 ? state + v : in + (v - s0));
}
__device__ __forceinline__ void ff_st_stream(cpx *p, cpx v) {
    asm volatile("st.global.cs.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(v.x), "d"(v.y) : "memory");
}
// L2 prefetch of `lines` 128-byte lines of the virtual stream starting at sample v (T threads cooperate).
template <int T>
__device__ __forceinline__ void vs_prefetch_l2(const cpx *__restrict__ state, int s0, const cpx *__restrict__ in, long long v,
                                               int samples, int t) {
    for (int i = 8 * t; i < samples; i += 8 * T) {
        const long long u = v + i;
        const cpx *p = u < s0 ? state + u : in + (u - s0);
        asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
    }
}

// Carried state of the NEXT call, new_state[i] = V[start + i] (V = state ++ in): written by CTA 0 of the last kernel a
// process call launches (inputs only, so it does not depend on the filtering), instead of by a launch of its own.
struct FfCarry { cpx *new_state; long long start; int count; };
__device__ __forceinline__ void ff_carry_state(const cpx *__restrict__ state, int s0, const cpx *__restrict__ in, FfCarry c,
                                               int state_stride) {
    if (blockIdx.x != 0 || c.count <= 0) return;
    cpx *ns = c.new_state + (long long)blockIdx.y * state_stride;
    for (int i = threadIdx.x; i < c.count; i += blockDim.x) ns[i] = vs_load(state, s0, in, c.start + i);
}

template <int N>
__global__ void __launch_bounds__(FftPlan<N>::THREADS)
fastfir_kernel(const cpx *__restrict__ state, int s0, const cpx *__restrict__ in, cpx *__restrict__ out,
               const cpx *__restrict__ hpos, const cpx *__restrict__ tw, int nblocks, double in_scale,
               double out_scale, long long in_stride, long long out_stride, int state_stride, FfCarry carry) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cpx *a = reinterpret_cast<cpx *>(smem_raw);
    constexpr int L = N / 2;
    constexpr int SH = FftPlan<N>::SH;
    // batched form: blockIdx.y = independent stream (own overlap state, own rows of in/out), same filter
    state += (long long)blockIdx.y * state_stride;
    in += blockIdx.y * in_stride;
    out += blockIdx.y * out_stride;
    ff_carry_state(state, s0, in, carry, state_stride);
    for (int blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
        const long long v0 = (long long)blk * L;
        for (int i = threadIdx.x; i < N; i += blockDim.x)
            a[fft_swz<SH>(i)] = cpx_scale(vs_load(state, s0, in, v0 + i), in_scale);
        __syncthreads();
        fft_forward_smem<N>(a, tw);
        // CpxMpy(N, m_pFilterCoef, m_pFFTBuf, m_pFFTBuf), fastfir.cpp:306,324-334
        for (int p = threadIdx.x; p < N; p += blockDim.x) a[fft_swz<SH>(p)] = cpx_mul(hpos[p], a[fft_swz<SH>(p)]);
        __syncthreads();
        fft_inverse_smem<N>(a, tw);
        // fastfir.cpp:308-311: output = samples FIR-1 .. N-1 of the inverse transform
        for (int i = threadIdx.x; i < L; i += blockDim.x) out[v0 + i] = cpx_scale(a[fft_swz<SH>(L + i)], out_scale);
        __syncthreads();
    }
}

// ---- N = 2048 (the reference's CONV_FFT_SIZE, fastfir.cpp:65): register-fused pass schedule --------------------------
// The generic kernel above makes 16 shared-memory accesses per point (load, 3 forward passes, multiply, 3 inverse
// passes, store) and ncu showed its L1/LSU data pipe 92 % busy: for 16-byte FP64 samples shared-memory bandwidth, not
// HBM or FP64, is the first wall. This schedule factors 2048 = 8 x 16 x 16 and makes 8 accesses per point:
//   F0   global -> registers -> radix-8 (stride 256) -> twiddle -> smem         (input never staged untransformed)
//   F1   smem -> radix-16 (stride 16 inside each 256-block) -> twiddle (smem table) -> smem
//   MID  smem -> radix-16 (contiguous) -> multiply by H (CpxMpy, fastfir.cpp:306,324-334) -> inverse radix-16 -> smem
//   I1   smem -> conj twiddle -> inverse radix-16 -> smem
//   I0   smem -> conj twiddle -> inverse radix-8 -> registers -> global         (upper L samples only, fastfir.cpp:308-311)
// Register budget <= 128 so that four CTAs (16 warps) share an SM: the F0/I0 twiddles are rebuilt from w, w^2, w^4 per
// butterfly (4 complex multiplies), the F1/I1 twiddles come from a 4 KB shared table laid out [power][butterfly].
// Slot p = 256 q0 + 16 q1 + q2 holds bin q0 + 8 q1 + 128 q2; H is pre-permuted to that order and transposed to
// [k][thread] so the MID loads are coalesced (ff2048_permute_kernel).
// `scale` = in_scale * out_scale of the handle: the fused kernels are linear, so the two gains ride on H.
__global__ void ff2048_permute_kernel(const cpx *__restrict__ hpos, cpx *__restrict__ ht, double scale) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;   // idx = k * 128 + t  <->  slot 16 t + k
    if (idx < 2048) ht[idx] = cpx_scale(hpos[FftPlan<2048>::freq2pos(ff2048_pos2freq(16 * (idx & 127) + (idx >> 7)))], scale);
}

// Optional epilogue for the receiver chain (AM, AGC off): the AM detector's recurrence d <- alpha d + |y| over one output
// block is the affine map {alpha^1024, sum_k alpha^(1023-k) |y_k|}; all 1024 terms share alpha, so the sum is a plain
// weighted reduction of values the I0 phase already holds in registers. wts[k] = alpha^(1023-k).
struct FfAgg { double2 *tile_map; const double *wts; double a_pow; };

template <int MINB>
__global__ void __launch_bounds__(128, MINB)
fastfir2048_kernel(const cpx *__restrict__ state, int s0, const cpx *__restrict__ in, cpx *__restrict__ out,
                   const cpx *__restrict__ ht, const cpx *__restrict__ tw, int blk_first, int nblocks, long long in_stride,
                   long long out_stride, int state_stride, FfCarry carry, FfAgg agg) {
    __shared__ double agg_red[4];
    constexpr int L = 1024;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cpx *a = reinterpret_cast<cpx *>(smem_raw);
    cpx *t1 = a + 2048;
    state += (long long)blockIdx.y * state_stride;
    in += blockIdx.y * in_stride;
    out += blockIdx.y * out_stride;
    const int t = threadIdx.x;
    ff_carry_state(state, s0, in, carry, state_stride);
    if (t < 64) t1[t] = tw[8 * ((t & 15) << (t >> 4))];   // t = 16 b + i -> exp(-j 2 pi i 2^b / 256)
    const cpx wa1 = tw[t], wa2 = tw[2 * t], wa4 = tw[4 * t];
    const cpx wb1 = tw[t + 128], wb2 = tw[2 * t + 256], wb4 = tw[4 * t + 512];
    __syncthreads();
    for (int blk = blk_first + blockIdx.x; blk < nblocks; blk += gridDim.x) {
        const long long v0 = (long long)blk * L;
        // ---- F0
#pragma unroll
        for (int g = 0; g < 2; ++g) {
            const int i = t + 128 * g;
            cpx u[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) u[k] = vs_load_stream(state, s0, in, v0 + i + 256 * k);
            dft8<false>(u);
            ff_tw8<false>(u, g ? wb1 : wa1, g ? wb2 : wa2, g ? wb4 : wa4);
#pragma unroll
            for (int k = 0; k < 8; ++k) a[ff_swz(i + 256 * k)] = u[k];
        }
        __syncthreads();
        ff_mid_phases<128, false>(a, t1, ht, t, 0);
        // ---- I0 -> global
        double agg_sum = 0.0;
#pragma unroll
        for (int g = 0; g < 2; ++g) {
            const int i = t + 128 * g;
            cpx u[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) u[k] = a[ff_swz(i + 256 * k)];
            ff_tw8<true>(u, g ? wb1 : wa1, g ? wb2 : wa2, g ? wb4 : wa4);
            dft8<true>(u);
#pragma unroll
            for (int k = 4; k < 8; ++k) ff_st_stream(out + v0 + i + 256 * (k - 4), u[k]);
            if (agg.tile_map) {
#pragma unroll
                for (int k = 4; k < 8; ++k)
                    agg_sum = fma(agg.wts[i + 256 * (k - 4)], sqrt(u[k].x * u[k].x + u[k].y * u[k].y), agg_sum);   // DB::amplitude, db.h:28-30
            }
        }
        if (agg.tile_map) {   // (uniform) block reduction of the 128 partial sums
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) agg_sum += __shfl_xor_sync(0xffffffffu, agg_sum, d);
            if ((t & 31) == 0) agg_red[t >> 5] = agg_sum;
            __syncthreads();
            if (t == 0) agg.tile_map[(long long)blockIdx.y * nblocks + blk] = make_double2(agg.a_pow, (agg_red[0] + agg_red[1]) + (agg_red[2] + agg_red[3]));
            __syncthreads();
        }
        // no barrier here: the next F0 of thread t overwrites exactly the slots i + 256 k that only thread t read in I0
    }
}

// ---- Triple block: the same linear convolution from 4096-point transforms ---------------------------------------------
// Overlap-save with N = 2048 and 1025 taps transforms 2 points per output sample; FP64 issue and shared-memory wavefronts
// (ncu: 53 % and 78 % of peak in the kernel above) both scale with the points transformed. A 4096-point transform with
// the SAME 1025 taps yields 3072 valid samples (4096 - 1024), i.e. 1.33 points per output, and 4096 = 16^3 needs no more
// passes than 2048 = 8*16*16. The circular convolution of length 4096 holds the linear convolution of 3 consecutive
// reference blocks exactly (3072 + 1025 - 1 = 4096), so outputs equal the reference's up to FP64 rounding; the
// reference's 1/2048 normalisation folded into the taps (fastfir.cpp:243) becomes 1/4096 by an exact factor 0.5 in H.
// Big block b produces outputs [3072 b, 3072 b + 3072) from virtual-stream samples [3072 b, 3072 b + 4096).
// tw4[m] = exp(-j 2 pi m / 4096); ht4[256 k + t] = 0.5 * FFT4096(taps)[bin of slot 16 t + k].
__global__ void ff4096_permute_kernel(const cpx *__restrict__ hpos4, cpx *__restrict__ ht4, double scale) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;   // idx = 256 k + t  <->  slot 16 t + k (FftPlan<4096> order)
    if (idx < 4096) ht4[idx] = cpx_scale(hpos4[16 * (idx & 255) + (idx >> 8)], 0.5 * scale);
}

// One CTA per SM, 512 threads = two independent groups of 256, each transforming its own big block in its own 64 KB of
// shared memory; H (64 KB) and the twiddle tables are loaded into shared memory ONCE per CTA and shared by both groups.
// ncu on the two-CTAs-per-SM form with H read through L1: 31 % L1 hit rate (the register spills of 2 x 256 threads and
// the streaming samples evict H), long-scoreboard the top stall. Here the MID phase has no global access at all, no
// twiddle lives in a register across phases (no spills), and the groups synchronise on their own named barrier.
#define PB_FF4_SMEM_CPX (4096 + 2 * 4096 + 64 + 4 * 256)
__global__ void __launch_bounds__(512, 1)
fastfir4096x3_kernel(const cpx *__restrict__ state, int s0, const cpx *__restrict__ in, cpx *__restrict__ out,
                     const cpx *__restrict__ ht4, const cpx *__restrict__ tw4, int nbig, long long in_stride,
                     long long out_stride, int state_stride, FfCarry carry) {
    constexpr int LB = 3072;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cpx *hs = reinterpret_cast<cpx *>(smem_raw);      // hs[256 k + t] = H of slot 16 t + k
    const int g = threadIdx.x >> 8, t = threadIdx.x & 255;
    cpx *a = hs + 4096 + 4096 * g;
    cpx *t1 = hs + 3 * 4096;                            // t1[16 b + i] = exp(-j 2 pi i 2^b / 256)
    cpx *f0 = t1 + 64;                                  // f0[256 b + t] = exp(-j 2 pi t 2^b / 4096)
    state += (long long)blockIdx.y * state_stride;
    in += blockIdx.y * in_stride;
    out += blockIdx.y * out_stride;
    ff_carry_state(state, s0, in, carry, state_stride);
    for (int i = threadIdx.x; i < 4096; i += 512) hs[i] = ht4[i];
    if (threadIdx.x < 64) t1[threadIdx.x] = tw4[16 * ((threadIdx.x & 15) << (threadIdx.x >> 4))];
    for (int i = threadIdx.x; i < 1024; i += 512) f0[i] = tw4[(i & 255) << (i >> 8)];
    __syncthreads();
    const int step = 2 * gridDim.x;
    for (int blk = 2 * blockIdx.x + g; blk < nbig; blk += step) {
        const long long v0 = (long long)blk * LB;
        {
            cpx v[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) v[k] = vs_load_stream(state, s0, in, v0 + t + 256 * k);
            if (blk + step < nbig) vs_prefetch_l2<256>(state, s0, in, v0 + (long long)step * LB + 1024, LB, t);
            dft16<false>(v);
            ff_tw16<false>(v, f0[t], f0[256 + t], f0[512 + t], f0[768 + t]);
#pragma unroll
            for (int k = 0; k < 16; ++k) a[ff_swz(t + 256 * k)] = v[k];
        }
        asm volatile("bar.sync %0, 256;" ::"r"(g + 1) : "memory");
        ff_mid_phases<256, true>(a, t1, hs, t, g + 1);
        {
            cpx v[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) v[k] = a[ff_swz(t + 256 * k)];
            ff_tw16<true>(v, f0[t], f0[256 + t], f0[512 + t], f0[768 + t]);
            dft16<true>(v);
            // samples 1024 .. 4095 of the inverse transform are the valid ones
#pragma unroll
            for (int k = 4; k < 16; ++k) ff_st_stream(out + v0 + t + 256 * (k - 4), v[k]);
        }
        // no barrier here: the next F0 of thread t overwrites exactly the slots t + 256 k that only thread t read in I0
    }
}

// Forward FFT of the zero-padded taps, stored in slot order (what fastfir_kernel multiplies by).
template <int N>
__global__ void __launch_bounds__(FftPlan<N>::THREADS)
fastfir_filter_kernel(const cpx *__restrict__ taps, cpx *__restrict__ hpos, const cpx *__restrict__ tw) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cpx *a = reinterpret_cast<cpx *>(smem_raw);
    constexpr int SH = FftPlan<N>::SH;
    for (int i = threadIdx.x; i < N; i += blockDim.x) a[fft_swz<SH>(i)] = taps[i];
    __syncthreads();
    fft_forward_smem<N>(a, tw);
    for (int p = threadIdx.x; p < N; p += blockDim.x) hpos[p] = a[fft_swz<SH>(p)];
}

// new_state[i] = V[start + i], V = state ++ in
__global__ void fastfir_state_kernel(const cpx *__restrict__ state, int s0, const cpx *__restrict__ in,
                                     cpx *__restrict__ new_state, long long start, int count, long long in_stride,
                                     int state_stride) {
    state += (long long)blockIdx.y * state_stride;
    new_state += (long long)blockIdx.y * state_stride;
    in += blockIdx.y * in_stride;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) new_state[i] = vs_load(state, s0, in, start + i);
}

typedef void (*FfKern2048)(const cpx *, int, const cpx *, cpx *, const cpx *, const cpx *, int, int, long long, long long, int, FfCarry,
                           FfAgg);
struct pebble_fastfir {
    int N, fir, L;
    // launch configuration of the fused 2048 kernels, resolved at the handle's first process call on ITS device
    int cfg_occ2 = 0, cfg_waves = 2;
    bool cfg_triple = true;
    FfKern2048 cfg_kern = nullptr;
    // reference members (fastfir.cpp:140-143): "unset" sentinels so the first setup always runs
    double lo, hi, offset, fs;
    cpx *d_tw, *d_hpos, *d_ht, *d_taps, *d_state[2];   // d_ht: H in the fused 2048 kernel's order
    double2 *agg_map; double *d_aggw; double agg_alpha, agg_apow;   // AM aggregate epilogue (one call), its weight table
    bool h_scale_valid; double h_scale;               // gain currently folded into d_ht / d_ht4
    cpx *d_tw4, *d_hpos4, *d_ht4;                      // N == 2048 only: 4096-point twiddles and H for the triple-block kernel
    int cur, fill;
    LongFir *lf;          // N > 8192: four-step transform through global memory (fastfir_long.cu)
    int streams;          // independent streams sharing this filter (batched form); state is streams x N cpx
    bool have_filter;
    double in_scale, out_scale;
    cudaStream_t stream;
    pebble::Staging stg;
};

template <int N> static cudaError_t ff_launch_filter(pebble_fastfir *h, cudaStream_t st) {
    const size_t smem = sizeof(cpx) * N;
    cudaError_t e = cudaFuncSetAttribute(fastfir_filter_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    fastfir_filter_kernel<N><<<1, FftPlan<N>::THREADS, smem, st>>>(h->d_taps, h->d_hpos, h->d_tw);
    pebble::count_launch();
    if (N == 2048) {
        // the triple-block kernel's H: 4096-point transform of the same taps (d_taps is zero beyond N)
        const size_t smem4 = sizeof(cpx) * 4096;
        e = cudaFuncSetAttribute(fastfir_filter_kernel<4096>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem4);
        if (e != cudaSuccess) return e;
        fastfir_filter_kernel<4096><<<1, FftPlan<4096>::THREADS, smem4, st>>>(h->d_taps, h->d_hpos4, h->d_tw4);
        pebble::count_launch();
        h->h_scale_valid = false;   // d_ht / d_ht4 are rebuilt (with the fused gains) by the next process call
    }
    return cudaGetLastError();
}
template <int N>
static cudaError_t ff_launch_process(pebble_fastfir *h, const cpx *in, cpx *out, int nblocks, cudaStream_t st, FfCarry carry,
                                     int streams = 1, long long in_stride = 0, long long out_stride = 0) {
    const FfCarry none = {nullptr, 0, 0};
    const size_t smem = sizeof(cpx) * N;
    cudaError_t e = cudaFuncSetAttribute(fastfir_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int occ = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fastfir_kernel<N>, FftPlan<N>::THREADS, smem);
    if (occ < 1) occ = 1;
    int grid = pebble::num_sms() * occ * 4;
    if (streams > 1) grid = (grid + streams - 1) / streams;
    if (grid > nblocks) grid = nblocks;
    if (grid < 1) grid = 1;
    if (N == 2048 && !getenv("PEBBLE_FF_GENERIC")) {
        const size_t smem2 = sizeof(cpx) * (2048 + 64), smem4 = sizeof(cpx) * PB_FF4_SMEM_CPX;
        // per HANDLE, not function-static: the shared-memory opt-in below is a per-device attribute, and a process may hold
        // handles on several devices (pebble_set_device); a handle is used by one thread at a time
        int &occ2 = h->cfg_occ2, &waves = h->cfg_waves;
        bool &triple = h->cfg_triple;
        FfKern2048 &kern = h->cfg_kern;
        if (!occ2) {
            const char *mb = getenv("PEBBLE_FF_MINB");   // registers per thread: 4 -> 128, 3 -> 168, 2 -> 255
            const int minb = mb ? atoi(mb) : 4;
            kern = minb <= 2 ? fastfir2048_kernel<2> : minb == 3 ? fastfir2048_kernel<3> : fastfir2048_kernel<4>;
            e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
            if (e != cudaSuccess) return e;
            e = cudaFuncSetAttribute(fastfir4096x3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem4);
            if (e != cudaSuccess) return e;
            const char *wv = getenv("PEBBLE_FF_WAVES");
            if (wv) waves = atoi(wv) > 0 ? atoi(wv) : 1;
            triple = !getenv("PEBBLE_FF_NO3");
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ2, kern, 128, smem2);
            if (occ2 < 1) occ2 = 1;
        }
        // the fused kernels carry in_scale * out_scale on H; rebuild the permuted copies when the filter or the gain changed
        const double hs = h->in_scale * h->out_scale;
        if (!h->h_scale_valid || hs != h->h_scale) {
            ff2048_permute_kernel<<<8, 256, 0, st>>>(h->d_hpos, h->d_ht, hs);
            ff4096_permute_kernel<<<16, 256, 0, st>>>(h->d_hpos4, h->d_ht4, hs);
            pebble::count_launch(); pebble::count_launch();
            h->h_scale = hs; h->h_scale_valid = true;
        }
        const int state_stride = streams > 1 ? h->N : 0;
        // The triple-block kernel wins on throughput (fewer points transformed per output) but one big block is a long
        // dependent chain and every CTA first loads 64 KB of H: take it only when each 256-thread group gets at least
        // `big_min` big blocks, otherwise the call is latency-bound and ONE launch of the 2048 kernel is faster
        // (the receiver tail: 512 reference blocks per step, 29 us as big + rest vs ~10 us).
        const int big_min = getenv("PEBBLE_FF_BIGMIN") ? atoi(getenv("PEBBLE_FF_BIGMIN")) : 4;
        const int ctas_big = streams > 1 ? (pebble::num_sms() + streams - 1) / streams : pebble::num_sms();
        // (the AM aggregate epilogue lives in the 2048 kernel only: such calls are the receiver tail, always latency-bound)
        const int nbig = (triple && !h->agg_map && nblocks / 3 >= 2 * ctas_big * big_min) ? nblocks / 3 : 0;
        const FfAgg agg = {h->agg_map, h->d_aggw, h->agg_apow};
        h->agg_map = nullptr;
        const int rest = nblocks - 3 * nbig;
        if (nbig > 0) {
            grid = pebble::num_sms();                       // one 512-thread CTA per SM, two big blocks in flight each
            if (streams > 1) grid = (grid + streams - 1) / streams;
            if (grid > (nbig + 1) / 2) grid = (nbig + 1) / 2;
            fastfir4096x3_kernel<<<dim3(grid, streams), 512, smem4, st>>>(
                h->d_state[h->cur], h->L + h->fill, in, out, h->d_ht4, h->d_tw4, nbig, in_stride, out_stride, state_stride,
                rest > 0 ? none : carry);
            pebble::count_launch();
            e = cudaGetLastError();
            if (e != cudaSuccess) return e;
        }
        if (rest > 0) {
            grid = pebble::num_sms() * occ2 * waves;
            if (streams > 1) grid = (grid + streams - 1) / streams;
            if (grid > rest) grid = rest;
            kern<<<dim3(grid, streams), 128, smem2, st>>>(
                h->d_state[h->cur], h->L + h->fill, in, out, h->d_ht, h->d_tw, 3 * nbig, nblocks, in_stride, out_stride,
                state_stride, carry, agg);
            pebble::count_launch();
        }
        return cudaGetLastError();
    }
    fastfir_kernel<N><<<dim3(grid, streams), FftPlan<N>::THREADS, smem, st>>>(
        h->d_state[h->cur], h->L + h->fill, in, out, h->d_hpos, h->d_tw, nblocks, h->in_scale, h->out_scale, in_stride,
        out_stride, streams > 1 ? h->N : 0, carry);
    pebble::count_launch();
    return cudaGetLastError();
}

// Long filters: the four-step kernels, then the carried state by its own small launch.
static cudaError_t ff_launch_long(pebble_fastfir *h, const cpx *in, cpx *out, int nblocks, cudaStream_t st, FfCarry carry,
                                  long long in_stride, long long out_stride) {
    const int state_stride = h->streams > 1 ? h->N : 0;
    cudaError_t e = lf_process(h->lf, h->d_hpos, h->d_state[h->cur], h->L + h->fill, state_stride, in, in_stride, out, out_stride,
                               nblocks, h->streams, h->in_scale * h->out_scale, st);
    if (e != cudaSuccess) return e;
    fastfir_state_kernel<<<dim3((carry.count + 255) / 256, h->streams), 256, 0, st>>>(
        h->d_state[h->cur], h->L + h->fill, in, carry.new_state, carry.start, carry.count, in_stride, state_stride);
    pebble::count_launch();
    return cudaGetLastError();
}

int fastfir_am_aggregate_next(pebble_fastfir *h, void *tile_map, double alpha, int n) {
    if (!h || h->N != 2048 || h->streams != 1 || getenv("PEBBLE_FF_GENERIC") || getenv("PEBBLE_FF_NOAGG")) return PEBBLE_ERR_STATE;
    if (h->fill != 0 || n <= 0 || n % h->L) return PEBBLE_ERR_STATE;   // output block b must be samples [1024 b, 1024 b + 1024)
    if (!h->d_aggw || alpha != h->agg_alpha) {
        std::vector<double> w(1024);
        w[1023] = 1.0;
        for (int k = 1022; k >= 0; --k) w[k] = w[k + 1] * alpha;
        if (!h->d_aggw) PB_CUDA(cudaMalloc(&h->d_aggw, sizeof(double) * 1024), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMemcpy(h->d_aggw, w.data(), sizeof(double) * 1024, cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
        h->agg_alpha = alpha;
        h->agg_apow = w[0] * alpha;   // alpha^1024
    }
    h->agg_map = (double2 *)tile_map;
    return 0;
}

#define FF_DISPATCH(h, expr2048, expr4096, expr8192)                 \
    ((h)->N == 2048 ? (expr2048) : (h)->N == 4096 ? (expr4096) : (expr8192))

extern "C" {

pebble_fastfir_t *pebble_fastfir_create_sized(int fft_size, int fir_size) {
    if ((fft_size != 2048 && fft_size != 4096 && fft_size != 8192 && !lf_size_supported(fft_size)) || fir_size != fft_size / 2 + 1) {
        pebble::set_error("fastfir: unsupported size %d/%d (fft in {2048,4096,8192,32768,131072}, fir = fft/2+1)", fft_size, fir_size);
        return nullptr;
    }
    pebble_fastfir *h = new pebble_fastfir();
    h->N = fft_size; h->fir = fir_size; h->L = fft_size / 2;
    h->lo = -1.0; h->hi = 1.0; h->offset = 1.0; h->fs = 1.0;   // fastfir.cpp:140-143
    h->d_tw = h->d_hpos = h->d_ht = h->d_taps = nullptr; h->d_tw4 = h->d_hpos4 = h->d_ht4 = nullptr; h->h_scale_valid = false; h->h_scale = 1.0; h->agg_map = nullptr; h->d_aggw = nullptr; h->agg_alpha = 0.0; h->agg_apow = 0.0; h->d_state[0] = h->d_state[1] = nullptr;
    h->lf = nullptr; h->cur = 0; h->fill = 0; h->streams = 1; h->have_filter = false; h->in_scale = h->out_scale = 1.0; h->stream = nullptr;
    const size_t nb = sizeof(cpx) * (size_t)fft_size;
    std::vector<double> tw;
    pebble_host::twiddle_table(fft_size, tw);
    bool ok = pebble::cuda_ok(cudaMalloc(&h->d_tw, nb), "cudaMalloc") && pebble::cuda_ok(cudaMalloc(&h->d_hpos, nb), "cudaMalloc") &&
              pebble::cuda_ok(cudaMalloc(&h->d_ht, nb), "cudaMalloc") && pebble::cuda_ok(cudaMemset(h->d_ht, 0, nb), "cudaMemset") &&
              pebble::cuda_ok(cudaMalloc(&h->d_taps, 2 * nb), "cudaMalloc") && pebble::cuda_ok(cudaMemset(h->d_taps, 0, 2 * nb), "cudaMemset") && pebble::cuda_ok(cudaMalloc(&h->d_state[0], nb), "cudaMalloc") &&
              pebble::cuda_ok(cudaMalloc(&h->d_state[1], nb), "cudaMalloc") &&
              pebble::cuda_ok(cudaMemcpy(h->d_tw, tw.data(), nb, cudaMemcpyHostToDevice), "cudaMemcpy(tw)") &&
              pebble::cuda_ok(cudaMemset(h->d_state[0], 0, nb), "cudaMemset") && pebble::cuda_ok(cudaMemset(h->d_state[1], 0, nb), "cudaMemset") &&
              // until SetupParameters runs the reference multiplies by an uninitialised buffer; we use zeros
              pebble::cuda_ok(cudaMemset(h->d_hpos, 0, nb), "cudaMemset") &&
              pebble::cuda_ok(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "cudaStreamCreate");
    if (ok && fft_size == 2048) {
        std::vector<double> tw4;
        pebble_host::twiddle_table(4096, tw4);
        ok = pebble::cuda_ok(cudaMalloc(&h->d_tw4, 2 * nb), "cudaMalloc") && pebble::cuda_ok(cudaMalloc(&h->d_hpos4, 2 * nb), "cudaMalloc") &&
             pebble::cuda_ok(cudaMalloc(&h->d_ht4, 2 * nb), "cudaMalloc") && pebble::cuda_ok(cudaMemset(h->d_ht4, 0, 2 * nb), "cudaMemset") &&
             pebble::cuda_ok(cudaMemcpy(h->d_tw4, tw4.data(), 2 * nb, cudaMemcpyHostToDevice), "cudaMemcpy(tw4)");
    }
    if (ok && lf_size_supported(fft_size)) ok = (h->lf = lf_create(fft_size, h->d_tw)) != nullptr;
    if (!ok) { pebble_fastfir_destroy(h); return nullptr; }
    return h;
}
pebble_fastfir_t *pebble_fastfir_create(void) { return pebble_fastfir_create_sized(2048, 1025); }  // fastfir.cpp:65-66

void pebble_fastfir_destroy(pebble_fastfir_t *h) {
    if (!h) return;
    lf_destroy(h->lf);
    cudaFree(h->d_tw); cudaFree(h->d_hpos); cudaFree(h->d_ht); cudaFree(h->d_tw4); cudaFree(h->d_hpos4); cudaFree(h->d_ht4); cudaFree(h->d_aggw); cudaFree(h->d_taps); cudaFree(h->d_state[0]); cudaFree(h->d_state[1]);
    if (h->stream) cudaStreamDestroy(h->stream);
    h->stg.release();
    delete h;
}

int pebble_fastfir_setup(pebble_fastfir_t *h, double lo, double hi, double offset, double fs) {
    if (!h) return PEBBLE_ERR_ARG;
    if (lo == h->lo && hi == h->hi && offset == h->offset && fs == h->fs) return 1;  // fastfir.cpp:195-199
    h->lo = lo; h->hi = hi; h->offset = offset; h->fs = fs;
    std::vector<double> taps;
    if (!pebble_host::fastfir_design(h->N, h->fir, lo, hi, offset, fs, taps)) return 2;  // fastfir.cpp:208-216
    // The filter tables are rewritten below. Process calls enqueue on caller streams (non-blocking ones included) without
    // synchronising, so wait for everything in flight first: the reference serialises the same pair with m_Mutex
    // (fastfir.cpp:218,290).
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(h->d_taps, taps.data(), sizeof(cpx) * (size_t)h->N, cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    cudaError_t e = h->lf ? lf_build_filter(h->lf, h->d_taps, h->d_hpos, 0)
                          : FF_DISPATCH(h, ff_launch_filter<2048>(h, 0), ff_launch_filter<4096>(h, 0), ff_launch_filter<8192>(h, 0));
    PB_CUDA(e, PEBBLE_ERR_CUDA);
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    h->have_filter = true;
    return 0;
}

int pebble_fastfir_set_fused_scale(pebble_fastfir_t *h, double in_scale, double out_scale) {
    if (!h) return PEBBLE_ERR_ARG;
    h->in_scale = in_scale; h->out_scale = out_scale;
    return 0;
}

int pebble_fastfir_out_count(const pebble_fastfir_t *h, int n) {
    if (!h || n < 0) return PEBBLE_ERR_ARG;
    return (int)((((long long)h->fill + n) / h->L) * h->L);
}

int pebble_fastfir_set_streams(pebble_fastfir_t *h, int streams) {
    if (!h || streams < 1) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    const size_t nb = sizeof(cpx) * (size_t)h->N * (size_t)streams;
    for (int b = 0; b < 2; ++b) {
        cudaFree(h->d_state[b]); h->d_state[b] = nullptr;
        PB_CUDA(cudaMalloc(&h->d_state[b], nb), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMemset(h->d_state[b], 0, nb), PEBBLE_ERR_CUDA);
    }
    h->streams = streams; h->cur = 0; h->fill = 0;
    return 0;
}

// The kernels read a block's input (and its L-sample overlap) while other CTAs / lanes write other blocks' outputs: the
// input and output of a device call must not overlap (the reference's caller, BandPassFilter::process, never aliases them
// either: bandpassfilter.cpp:47-53 writes ProcessStep::out). Refused with a clear error instead of racing.
int pebble_fastfir_process_batch_device(pebble_fastfir_t *h, int n, const double *d_in, long long in_stride,
                                        double *d_out, long long out_stride, void *stream_v) {
    if (!h || n < 0 || (n && !d_in)) { pebble::set_error("fastfir: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream_v;
    const long long total = (long long)h->fill + n;
    const int nblocks = (int)(total / h->L);
    // carry: the last L samples before the first unprocessed one, plus the unprocessed remainder
    const int new_fill = (int)(total % h->L);
    const int count = h->L + new_fill;
    const long long vlen = (long long)h->L + total;  // length of state ++ in
    const FfCarry carry = {h->d_state[h->cur ^ 1], vlen - count, count};
    if (nblocks > 0) {
        if (!d_out) { pebble::set_error("fastfir: NULL output"); return PEBBLE_ERR_ARG; }
        if (pebble::ranges_overlap(d_in, sizeof(cpx) * ((size_t)(h->streams - 1) * (size_t)in_stride + (size_t)n), d_out,
                              sizeof(cpx) * ((size_t)(h->streams - 1) * (size_t)out_stride + (size_t)nblocks * h->L))) {
            pebble::set_error("fastfir: d_in and d_out overlap (in-place device calls are not supported)");
            return PEBBLE_ERR_ARG;
        }
        cudaError_t e = h->lf ? ff_launch_long(h, (const cpx *)d_in, (cpx *)d_out, nblocks, st, carry, in_stride, out_stride) : FF_DISPATCH(h,
            ff_launch_process<2048>(h, (const cpx *)d_in, (cpx *)d_out, nblocks, st, carry, h->streams, in_stride, out_stride),
            ff_launch_process<4096>(h, (const cpx *)d_in, (cpx *)d_out, nblocks, st, carry, h->streams, in_stride, out_stride),
            ff_launch_process<8192>(h, (const cpx *)d_in, (cpx *)d_out, nblocks, st, carry, h->streams, in_stride, out_stride));
        PB_CUDA(e, PEBBLE_ERR_CUDA);
    } else {
        fastfir_state_kernel<<<dim3((count + 255) / 256, h->streams), 256, 0, st>>>(
            h->d_state[h->cur], h->L + h->fill, (const cpx *)d_in, carry.new_state, carry.start, count, in_stride,
            h->streams > 1 ? h->N : 0);
        pebble::count_launch();
        PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    }
    h->cur ^= 1;
    h->fill = new_fill;
    return nblocks * h->L;
}

int pebble_fastfir_process_device(pebble_fastfir_t *h, int n, const double *d_in, double *d_out, void *stream_v) {
    if (!h || n < 0 || (n && !d_in)) { pebble::set_error("fastfir: bad argument"); return PEBBLE_ERR_ARG; }
    if (h->streams != 1) { pebble::set_error("fastfir: handle is batched, use process_batch_device"); return PEBBLE_ERR_STATE; }
    if (n == 0) return 0;  // fastfir.cpp:287-288
    cudaStream_t st = (cudaStream_t)stream_v;
    const long long total = (long long)h->fill + n;
    const int nblocks = (int)(total / h->L);
    // carry: the last L samples before the first unprocessed one, plus the unprocessed remainder
    const int new_fill = (int)(total % h->L);
    const int count = h->L + new_fill;
    const long long vlen = (long long)h->L + total;  // length of state ++ in
    const FfCarry carry = {h->d_state[h->cur ^ 1], vlen - count, count};
    if (nblocks > 0) {
        if (!d_out) { pebble::set_error("fastfir: NULL output"); return PEBBLE_ERR_ARG; }
        if (pebble::ranges_overlap(d_in, sizeof(cpx) * (size_t)n, d_out, sizeof(cpx) * (size_t)nblocks * h->L)) {
            pebble::set_error("fastfir: d_in and d_out overlap (in-place device calls are not supported)");
            return PEBBLE_ERR_ARG;
        }
        cudaError_t e = h->lf ? ff_launch_long(h, (const cpx *)d_in, (cpx *)d_out, nblocks, st, carry, 0, 0) :
                        FF_DISPATCH(h, ff_launch_process<2048>(h, (const cpx *)d_in, (cpx *)d_out, nblocks, st, carry),
                                    ff_launch_process<4096>(h, (const cpx *)d_in, (cpx *)d_out, nblocks, st, carry),
                                    ff_launch_process<8192>(h, (const cpx *)d_in, (cpx *)d_out, nblocks, st, carry));
        PB_CUDA(e, PEBBLE_ERR_CUDA);
    } else {
        fastfir_state_kernel<<<(count + 255) / 256, 256, 0, st>>>(h->d_state[h->cur], h->L + h->fill, (const cpx *)d_in,
                                                                 carry.new_state, carry.start, count, 0, 0);
        pebble::count_launch();
        PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    }
    h->cur ^= 1;
    h->fill = new_fill;
    return nblocks * h->L;
}

int pebble_fastfir_process(pebble_fastfir_t *h, int n, const double *in, double *out) {
    if (!h || n < 0 || (n && (!in || !out))) { pebble::set_error("fastfir: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    const int k = pebble_fastfir_out_count(h, n);
    const size_t bi = (size_t)n * sizeof(cpx), bo = (size_t)k * sizeof(cpx);
    if (!h->stg.reserve(bi, bo ? bo : 16)) return PEBBLE_ERR_CUDA;
    memcpy(h->stg.h_in, in, bi);
    PB_CUDA(cudaMemcpyAsync(h->stg.d_in, h->stg.h_in, bi, cudaMemcpyHostToDevice, h->stream), PEBBLE_ERR_CUDA);
    int r = pebble_fastfir_process_device(h, n, (const double *)h->stg.d_in, (double *)h->stg.d_out, h->stream);
    if (r < 0) return r;
    if (bo) PB_CUDA(cudaMemcpyAsync(h->stg.h_out, h->stg.d_out, bo, cudaMemcpyDeviceToHost, h->stream), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamSynchronize(h->stream), PEBBLE_ERR_CUDA);
    if (bo) memcpy(out, h->stg.h_out, bo);
    return r;
}

int64_t pebble_fastfir_state_size(const pebble_fastfir_t *h) {
    if (!h) return PEBBLE_ERR_ARG;
    return 8 + (int64_t)sizeof(cpx) * (h->L + h->fill);
}
int pebble_fastfir_get_state(pebble_fastfir_t *h, void *buf, int64_t bytes) {
    if (!h || !buf || bytes < pebble_fastfir_state_size(h)) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    int64_t f = h->fill;
    memcpy(buf, &f, 8);
    PB_CUDA(cudaMemcpy((char *)buf + 8, h->d_state[h->cur], sizeof(cpx) * (h->L + h->fill), cudaMemcpyDeviceToHost), PEBBLE_ERR_CUDA);
    return 0;
}
int pebble_fastfir_set_state(pebble_fastfir_t *h, const void *buf, int64_t bytes) {
    if (!h || !buf || bytes < 8) return PEBBLE_ERR_ARG;
    int64_t f;
    memcpy(&f, buf, 8);
    if (f < 0 || f >= h->L || bytes < 8 + (int64_t)sizeof(cpx) * (h->L + f)) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(h->d_state[h->cur], (const char *)buf + 8, sizeof(cpx) * (h->L + f), cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    h->fill = (int)f;
    return 0;
}
int pebble_fastfir_get_filter(pebble_fastfir_t *h, double *coef_out) {
    if (!h || !coef_out) return PEBBLE_ERR_ARG;
    std::vector<cpx> hp(h->N);
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(hp.data(), h->d_hpos, sizeof(cpx) * h->N, cudaMemcpyDeviceToHost), PEBBLE_ERR_CUDA);
    for (int f = 0; f < h->N; ++f) {
        int p = h->lf ? lf_freq2index(h->N, f)
                      : FF_DISPATCH(h, FftPlan<2048>::freq2pos(f), FftPlan<4096>::freq2pos(f), FftPlan<8192>::freq2pos(f));
        coef_out[2 * f] = hp[p].x;
        coef_out[2 * f + 1] = hp[p].y;
    }
    return 0;
}

}  // extern "C"
