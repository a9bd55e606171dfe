// spectrum.cu — the FFT operator (pebblelib/fft.h:29-62, fftw.cpp:17-81) and the SignalSpectrum hot
// path (application/signalspectrum.cpp:63-113 -> FFT::fftSpectrum) on the GPU.
// One CTA per frame: window x input (+ overload flag) -> shared memory -> in-place FP64 FFT -> unfold
// (fftshift) -> |X|/coherentGain/maxBinPower -> 2-tap average with the previous frame's UNAVERAGED value
// (finding F14) -> 20*log10 -> clip [-120, 0]. In a batch a CTA walks a run of consecutive frames and
// keeps the previous frame's amplitudes in registers; the first frame of a run recomputes its
// predecessor (one extra FFT per run) instead of exchanging data with the neighbouring CTA.
#include "lib_common.h"
#include "fft_kernels.cuh"
#include "host_design.h"
#include <cstdlib>
#include <cstring>
#include <vector>

template <int N>
__device__ __forceinline__ bool spec_load_frame(cpx *a, const cpx *__restrict__ in, int n, const double *__restrict__ win,
                                                bool windowed, double over_limit) {
    constexpr int SH = FftPlan<N>::SH;
    int over = 0;
    for (int i = threadIdx.x; i < N; i += blockDim.x) {
        cpx v = cpx_make(0.0, 0.0);
        if (i < n) {
            v = in[i];
            if (windowed) {
                // fft.cpp:136-142: overload test on the raw sample, then in[i] * windowCpx[i] (= (w, 0))
                if (fabs(v.x) > over_limit || fabs(v.y) > over_limit) over = 1;
                v = cpx_scale(v, win[i]);
            }
        }
        a[fft_swz<SH>(i)] = v;
    }
    return __syncthreads_or(over) != 0;  // also the barrier that completes a[]
}

// amplitude (already / coherentGain / maxBinPower) of unfolded bin j
template <int N> __device__ __forceinline__ double spec_asd(const cpx *a, int j, double inv_cg, double inv_maxbin) {
    const int f = (j + N / 2) & (N - 1);  // m_unfoldInOrder, fft.cpp:183-225
    const cpx X = a[fft_swz<FftPlan<N>::SH>(FftPlan<N>::freq2pos(f))];
    // DB::amplitude / coherentGain (fft.cpp:350) / maxBinPower (fft.cpp:358). cg and maxbin arrive as RECIPROCALS: two
    // FP64 divisions per bin were ~60 instructions of the ~300 per sample; the products differ from the quotients by
    // <= 1 ulp each (1e-15 dB), far inside the 1e-6 dB tolerance; 1/maxbin is exact for power-of-two buffer sizes.
    // sqrt: the branch-free Newton form (common.cuh), within 1 ulp of the correctly rounded root: ~2e-15 dB
    return pb_sqrt_newton(X.x * X.x + X.y * X.y) * inv_cg * inv_maxbin;
}

// DB::amplitudeTodB (db.h:44-48) + DB::clip (db.h:24-26): 20 log10(bin) clipped to [-120, 0]. The library log10 was 31 % of
// the kernel's instructions (ncu source page, ~90 per bin). Here x = m 2^e, m in [1, 2) is reduced with a 32-entry table
// of interval centres c_i: r = m / c_i - 1, |r| <= 1/64, and log(1 + r) is a degree-6 series (truncation r^7 / 7 < 7e-14,
// i.e. < 1e-12 dB against the 1e-6 dB tolerance of the parity tests). tab[i] = 1 / c_i, tab[32 + i] = 20 log10(c_i).
// Values outside [1e-7, 1e300] (zero, results that clip anyway, inf, NaN) take the library path.
__device__ __forceinline__ double spec_db(double bin, const double *tab) {
    double db;
    if (!(bin >= 1e-7 && bin <= 1e300)) {
        db = bin == 0 ? -120.0 : 20 * log10(bin);
    } else {
        const long long b = __double_as_longlong(bin);
        const int e = (int)(b >> 52) - 1023;
        const int i = (int)(b >> 47) & 31;
        const double m = __longlong_as_double((b & 0x000fffffffffffffLL) | 0x3ff0000000000000LL);
        const double r = fma(m, tab[i], -1.0);
        double p = fma(r, -1.0 / 6.0, 0.2);
        p = fma(p, r, -0.25);
        p = fma(p, r, 1.0 / 3.0);
        p = fma(p, r, -0.5);
        p = fma(p, r, 1.0);
        // 20 log10(x) = 20 log10(2) e + 20 log10(c_i) + (20 / ln 10) log(1 + r)
        db = fma((double)e, 6.020599913279623904, fma(r * p, 8.685889638065036553, tab[32 + i]));
    }
    return db < -120.0 ? -120.0 : (db > 0.0 ? 0.0 : db);
}

template <int N>
__global__ void __launch_bounds__(FftPlan<N>::THREADS)
spectrum_kernel(const cpx *__restrict__ in, int frames, int n, const double *__restrict__ win, int windowed,
                double cg, double maxbin /* both reciprocals, see spec_asd */, double over_limit, const double *__restrict__ prev_in,
                double *__restrict__ prev_out, const int *__restrict__ flag_in, int *__restrict__ flag_out,
                double *__restrict__ out_db, int *__restrict__ out_flags, int run, const cpx *__restrict__ tw) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cpx *a = reinterpret_cast<cpx *>(smem_raw);
    constexpr int T = FftPlan<N>::THREADS;
    constexpr int PER = N / T;
    __shared__ double db_tab[64];
    if (threadIdx.x < 32) {
        const double c = 1.0 + (threadIdx.x + 0.5) / 32.0;
        db_tab[threadIdx.x] = 1.0 / c;
        db_tab[32 + threadIdx.x] = 20 * log10(c);
    }
    double prev[PER];
    const int f0 = blockIdx.x * run;
    int f1 = f0 + run;
    if (f1 > frames) f1 = frames;
    const int carried_flag = *flag_in;
    if (f0 == 0) {
#pragma unroll
        for (int k = 0; k < PER; ++k) prev[k] = prev_in[threadIdx.x + k * T];
    } else {
        // pre-roll: unaveraged amplitudes of frame f0-1
        spec_load_frame<N>(a, in + (long long)(f0 - 1) * n, n, win, windowed != 0, over_limit);
        fft_forward_smem<N>(a, tw);
#pragma unroll
        for (int k = 0; k < PER; ++k) prev[k] = spec_asd<N>(a, threadIdx.x + k * T, cg, maxbin);
        __syncthreads();
    }
    for (int f = f0; f < f1; ++f) {
        const bool over = spec_load_frame<N>(a, in + (long long)f * n, n, win, windowed != 0, over_limit);
        if (f + 1 < f1) {   // next frame -> L2 while this one is transformed (one 128-byte line per 8 samples)
            const cpx *nx = in + (long long)(f + 1) * n;
            for (int i = 8 * threadIdx.x; i < n; i += 8 * T) asm volatile("prefetch.global.L2 [%0];" ::"l"(nx + i));
        }
        fft_forward_smem<N>(a, tw);
        double *o = out_db + (long long)f * N;
#pragma unroll
        for (int k = 0; k < PER; ++k) {
            const int j = threadIdx.x + k * T;
            const double asd = spec_asd<N>(a, j, cg, maxbin);
            const double bin = (asd + prev[k]) / 2;   // fft.cpp:380 (m_isAveraged is always true, fft.cpp:103)
            prev[k] = asd;                            // fft.cpp:388: keep the UNAVERAGED value
            o[j] = spec_db(bin, db_tab);              // DB::amplitudeTodB + DB::clip
        }
        if (threadIdx.x == 0) {
            const int fl = windowed ? (over ? 1 : 0) : carried_flag;
            if (out_flags) out_flags[f] = fl;
            if (f == frames - 1) *flag_out = fl;
        }
        __syncthreads();
    }
    if (f1 == frames && f0 < f1) {
#pragma unroll
        for (int k = 0; k < PER; ++k) prev_out[threadIdx.x + k * T] = prev[k];
    }
}

// ---- full frames (n == N): the frame never sits in shared memory as raw input ------------------------------------------
// spectrum_kernel stages a frame (global -> window -> shared), then pass 0 reads it back: one shared-memory round trip, one
// barrier, and the whole DRAM latency of the frame exposed once per frame (one CTA per SM at N = 8192: nothing else runs
// meanwhile; ncu r1: DRAM 17.8 % busy, LSU data pipe 63 %). Pass 0's butterfly bf takes the points bf + k N / R0 — consecutive
// bf are consecutive samples, so the butterfly can take its inputs straight from global memory, coalesced, window and overload
// test applied in registers. The loads of frame f+1 are issued BEFORE the epilogue of frame f (sqrt, dB, stores: thousands of
// cycles) and consumed after it, so they are in flight under work that needs no shared memory bandwidth.
template <int N> struct SpecFetch {
    static constexpr int T = FftPlan<N>::THREADS, R0 = FftPlan<N>::R0, S0 = N / R0, NB = S0 / T;   // butterflies per thread in pass 0
    cpx x[NB][R0];
    __device__ __forceinline__ void fetch(const cpx *__restrict__ in) {
#pragma unroll
        for (int b = 0; b < NB; ++b)
#pragma unroll
            for (int k = 0; k < R0; ++k) x[b][k] = __ldg(in + threadIdx.x + b * T + k * S0);
    }
    // window + overload test + pass 0 (forward: DFT, then twiddle), written to the swizzled slots of a[]
    __device__ __forceinline__ int pass0(cpx *a, const cpx *__restrict__ tw, const double *__restrict__ win, bool windowed, double over_limit) {
        constexpr int SH = FftPlan<N>::SH;
        int over = 0;
#pragma unroll
        for (int b = 0; b < NB; ++b) {
            const int i = threadIdx.x + b * T;
            cpx v[R0];
            cpx w[16];
#pragma unroll
            for (int k = 0; k < R0; ++k) {
                v[k] = x[b][k];
                if (windowed) {   // fft.cpp:136-142: overload test on the raw sample, then in[i] * windowCpx[i]
                    if (fabs(v[k].x) > over_limit || fabs(v[k].y) > over_limit) over = 1;
                    v[k] = cpx_scale(v[k], __ldg(win + i + k * S0));
                }
            }
            fft_twiddle_powers<R0>(tw, i, N - 1, w);
            DftR<R0, false>::run(v);
#pragma unroll
            for (int q = 1; q < R0; ++q) v[q] = cpx_mul(v[q], w[q]);
#pragma unroll
            for (int k = 0; k < R0; ++k) a[fft_swz<SH>(i + k * S0)] = v[k];
        }
        return over;
    }
};
template <int N, int P> struct FftRunFrom1 {   // passes 1 .. P, a barrier after each
    static PB_D void fwd(cpx *a, const cpx *tw) {
        FftRunFrom1<N, P - 1>::fwd(a, tw);
        FftPlan<N>::template pass<P, false>(a, tw, threadIdx.x, blockDim.x);
        __syncthreads();
    }
};
template <int N> struct FftRunFrom1<N, 0> { static PB_D void fwd(cpx *, const cpx *) {} };

template <int N>
__global__ void __launch_bounds__(FftPlan<N>::THREADS)
spectrum_full_kernel(const cpx *__restrict__ in, int frames, const double *__restrict__ win, int windowed,
                     double cg, double maxbin /* both reciprocals, see spec_asd */, double over_limit, const double *__restrict__ prev_in,
                     double *__restrict__ prev_out, const int *__restrict__ flag_in, int *__restrict__ flag_out,
                     double *__restrict__ out_db, int *__restrict__ out_flags, int run, const cpx *__restrict__ tw) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cpx *a = reinterpret_cast<cpx *>(smem_raw);
    constexpr int T = FftPlan<N>::THREADS;
    constexpr int PER = N / T;
    __shared__ double db_tab[64];
    if (threadIdx.x < 32) {
        const double c = 1.0 + (threadIdx.x + 0.5) / 32.0;
        db_tab[threadIdx.x] = 1.0 / c;
        db_tab[32 + threadIdx.x] = 20 * log10(c);
    }
    // the previous frame's unaveraged amplitudes sit behind the frame in shared memory (8 N bytes): in registers they and the
    // prefetched frame (64 + 32 registers at N = 8192, 512 threads = 128 registers each) spilled
    double *prev = reinterpret_cast<double *>(a + N) + threadIdx.x;   // element k of this thread: prev[k * T]
    const int f0 = blockIdx.x * run;
    int f1 = f0 + run;
    if (f1 > frames) f1 = frames;
    if (f0 >= f1) return;
    const int carried_flag = *flag_in;
    SpecFetch<N> sf;
    if (f0 == 0) {
#pragma unroll
        for (int k = 0; k < PER; ++k) prev[k * T] = prev_in[threadIdx.x + k * T];
    } else {
        // pre-roll: unaveraged amplitudes of frame f0-1
        sf.fetch(in + (long long)(f0 - 1) * N);
        sf.pass0(a, tw, win, windowed != 0, over_limit);
        __syncthreads();
        FftRunFrom1<N, FftPlan<N>::NP - 1>::fwd(a, tw);
#pragma unroll
        for (int k = 0; k < PER; ++k) prev[k * T] = spec_asd<N>(a, threadIdx.x + k * T, cg, maxbin);
        __syncthreads();
    }
    sf.fetch(in + (long long)f0 * N);
    for (int f = f0; f < f1; ++f) {
        const bool over = __syncthreads_or(sf.pass0(a, tw, win, windowed != 0, over_limit)) != 0;   // also the barrier after pass 0
        FftRunFrom1<N, FftPlan<N>::NP - 1>::fwd(a, tw);
        if (f + 1 < f1) sf.fetch(in + (long long)(f + 1) * N);      // in flight during the epilogue below
        double *o = out_db + (long long)f * N;
#pragma unroll
        for (int k = 0; k < PER; ++k) {
            const int j = threadIdx.x + k * T;
            const double asd = spec_asd<N>(a, j, cg, maxbin);
            const double bin = (asd + prev[k * T]) / 2;   // fft.cpp:380 (m_isAveraged is always true, fft.cpp:103)
            prev[k * T] = asd;                            // fft.cpp:388: keep the UNAVERAGED value
            __stcs(o + j, spec_db(bin, db_tab));      // DB::amplitudeTodB + DB::clip; streaming store: never re-read here
        }
        if (threadIdx.x == 0) {
            const int fl = windowed ? (over ? 1 : 0) : carried_flag;
            if (out_flags) out_flags[f] = fl;
            if (f == frames - 1) *flag_out = fl;
        }
        __syncthreads();
    }
    if (f1 == frames) {
#pragma unroll
        for (int k = 0; k < PER; ++k) prev_out[threadIdx.x + k * T] = prev[k * T];
    }
}

// plain transforms with natural (FFTW) bin order at the interface
template <int N, bool INV>
__global__ void __launch_bounds__(FftPlan<N>::THREADS)
fft_c2c_kernel(const cpx *__restrict__ in, int n, const double *__restrict__ win, int windowed, double over_limit,
               cpx *__restrict__ out, const int *__restrict__ flag_in, int *__restrict__ flag_out,
               const cpx *__restrict__ tw) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cpx *a = reinterpret_cast<cpx *>(smem_raw);
    constexpr int SH = FftPlan<N>::SH;
    if (!INV) {
        const bool over = spec_load_frame<N>(a, in, n, win, windowed != 0, over_limit);
        if (threadIdx.x == 0) *flag_out = windowed ? (over ? 1 : 0) : *flag_in;
        fft_forward_smem<N>(a, tw);
        for (int f = threadIdx.x; f < N; f += blockDim.x) out[f] = a[fft_swz<SH>(FftPlan<N>::freq2pos(f))];
    } else {
        // fftw.cpp:55-60: bins beyond numSamples are zero
        for (int f = threadIdx.x; f < N; f += blockDim.x)
            a[fft_swz<SH>(FftPlan<N>::freq2pos(f))] = f < n ? in[f] : cpx_make(0.0, 0.0);
        __syncthreads();
        fft_inverse_smem<N>(a, tw);
        for (int i = threadIdx.x; i < N; i += blockDim.x) out[i] = a[fft_swz<SH>(i)];
    }
}

struct pebble_fft {
    int N = 0, spb = 0, wtype = 14;
    int backend = 0;                          // PEBBLE_FFT_FFTW / PEBBLE_FFT_OOURA: which reference class's contract
    double cg = 1.0, maxbin = 1.0, fs = 0.0;
    bool params_set = false;
    cpx *d_tw = nullptr;
    double *d_win = nullptr;
    double *d_prev[2] = {nullptr, nullptr};   // unaveraged amplitudes of the last frame (m_fftAmplitude)
    int *d_flag[2] = {nullptr, nullptr};      // m_isOverload
    int cur = 0;
    cudaStream_t stream = nullptr;
    pebble::Staging stg;
    int *h_flag = nullptr;                    // pinned
};

static void fft_free(pebble_fft *h) {
    cudaFree(h->d_tw); cudaFree(h->d_win); cudaFree(h->d_prev[0]); cudaFree(h->d_prev[1]);
    cudaFree(h->d_flag[0]); cudaFree(h->d_flag[1]);
    h->d_tw = nullptr; h->d_win = nullptr; h->d_prev[0] = h->d_prev[1] = nullptr; h->d_flag[0] = h->d_flag[1] = nullptr;
}

template <int N> static cudaError_t spec_launch(pebble_fft *h, int frames, int n, const cpx *in, double *out, int *flags,
                                                cudaStream_t st) {
    const size_t smem = sizeof(cpx) * N;
    cudaError_t e = cudaFuncSetAttribute(spectrum_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int occ = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, spectrum_kernel<N>, FftPlan<N>::THREADS, smem);
    if (occ < 1) occ = 1;
    // runs of consecutive frames per CTA: >= 8 frames (pre-roll <= 12.5 %) once there is enough work
    // ONE wave: with a single resident CTA per SM (128 KB of shared memory at N = 8192) a grid of 2 x SMs quantised to
    // 256 CTAs = 1.73 waves (ncu), i.e. 13 % of the machine idle in the second wave.
    int target = pebble::num_sms() * occ;
    int run = (frames + target - 1) / target;
    if (run < 8) run = 8;
    if (run > frames) run = frames;
    const int grid = (frames + run - 1) / run;
    const bool windowed = h->wtype != PEBBLE_WINDOW_NONE && n == h->spb;
    if constexpr (N == 8192) {   // one CTA per SM either way at this size; the smaller sizes keep their higher occupancy
      if (n == N && !getenv("PEBBLE_SPEC_STAGED")) {     // full frames: pass 0 fed from global memory (spectrum_full_kernel)
        const size_t smem2 = smem + sizeof(double) * N;
        e = cudaFuncSetAttribute(spectrum_full_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
        if (e != cudaSuccess) return e;
        spectrum_full_kernel<N><<<grid, FftPlan<N>::THREADS, smem2, st>>>(in, frames, h->d_win, windowed ? 1 : 0, 1.0 / h->cg, 1.0 / h->maxbin, 0.9,
                                                                        h->d_prev[h->cur], h->d_prev[h->cur ^ 1], h->d_flag[h->cur],
                                                                        h->d_flag[h->cur ^ 1], out, flags, run, h->d_tw);
        pebble::count_launch();
        return cudaGetLastError();
      }
    }
    spectrum_kernel<N><<<grid, FftPlan<N>::THREADS, smem, st>>>(in, frames, n, h->d_win, windowed ? 1 : 0, 1.0 / h->cg, 1.0 / h->maxbin, 0.9,
                                                               h->d_prev[h->cur], h->d_prev[h->cur ^ 1], h->d_flag[h->cur],
                                                               h->d_flag[h->cur ^ 1], out, flags, run, h->d_tw);
    pebble::count_launch();
    return cudaGetLastError();
}
template <int N, bool INV> static cudaError_t c2c_launch(pebble_fft *h, int n, const cpx *in, cpx *out, cudaStream_t st, bool allow_window = true) {
    const size_t smem = sizeof(cpx) * N;
    cudaError_t e = cudaFuncSetAttribute(fft_c2c_kernel<N, INV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    const bool windowed = allow_window && !INV && h->wtype != PEBBLE_WINDOW_NONE && n == h->spb;
    fft_c2c_kernel<N, INV><<<1, FftPlan<N>::THREADS, smem, st>>>(in, n, h->d_win, windowed ? 1 : 0, 0.9, out, h->d_flag[h->cur],
                                                                h->d_flag[h->cur ^ 1], h->d_tw);
    pebble::count_launch();
    return cudaGetLastError();
}

#define FFT_DISPATCH(h, X) ((h)->N == 2048 ? X(2048) : (h)->N == 4096 ? X(4096) : X(8192))

extern "C" {

pebble_fft_t *pebble_fft_create(void) {
    pebble_fft *h = new pebble_fft();
    if (!pebble::cuda_ok(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "cudaStreamCreate") ||
        !pebble::cuda_ok(cudaMallocHost(&h->h_flag, sizeof(int)), "cudaMallocHost")) {
        delete h;
        return nullptr;
    }
    return h;
}
void pebble_fft_destroy(pebble_fft_t *h) {
    if (!h) return;
    fft_free(h);
    if (h->stream) cudaStreamDestroy(h->stream);
    if (h->h_flag) cudaFreeHost(h->h_flag);
    h->stg.release();
    delete h;
}

int pebble_fft_params(pebble_fft_t *h, unsigned fft_size, double db_comp, double fs, int spb, int wtype) {
    (void)db_comp;  // unused by the reference too (fft.cpp:70)
    if (!h) return PEBBLE_ERR_ARG;
    if (fft_size == 0) return 0;  // fft.cpp:72-73: silently ignored
    // fft.cpp:74-79 clamps to [2048, 65535]; this build implements the power-of-two sizes below
    if (fft_size < 2048) fft_size = 2048;
    if (fft_size != 2048 && fft_size != 4096 && fft_size != 8192) {
        pebble::set_error("fft: size %u not supported (2048, 4096, 8192)", fft_size);
        return PEBBLE_ERR_ARG;
    }
    if (spb <= 0 || spb > (int)fft_size || wtype == PEBBLE_WINDOW_BLACKMAN || wtype < 1 || wtype > 14) {
        pebble::set_error("fft: bad samples_per_buffer/window (BLACKMAN reads uninitialised memory in the reference)");
        return PEBBLE_ERR_ARG;
    }
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    fft_free(h);
    h->N = (int)fft_size; h->spb = spb; h->wtype = wtype; h->fs = fs;
    h->maxbin = 1.0 * spb;  // m_maxBinPower = m_ampMax * m_samplesPerBuffer, fft.cpp:84
    std::vector<double> tw, win;
    pebble_host::twiddle_table(h->N, tw);
    // WindowFunction(NONE) leaves coherentGain uninitialised in the reference; 1.0 is used here
    h->cg = wtype == PEBBLE_WINDOW_NONE ? 1.0 : pebble_host::window_table(wtype, spb, win);
    if (win.empty()) win.assign((size_t)spb, 1.0);
    const size_t nb = sizeof(cpx) * (size_t)h->N;
    PB_CUDA(cudaMalloc(&h->d_tw, nb), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMalloc(&h->d_win, sizeof(double) * (size_t)spb), PEBBLE_ERR_CUDA);
    for (int b = 0; b < 2; ++b) {
        PB_CUDA(cudaMalloc(&h->d_prev[b], sizeof(double) * (size_t)h->N), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMemset(h->d_prev[b], 0, sizeof(double) * (size_t)h->N), PEBBLE_ERR_CUDA);  // reference: new double[] (uninitialised)
        PB_CUDA(cudaMalloc(&h->d_flag[b], sizeof(int)), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMemset(h->d_flag[b], 0, sizeof(int)), PEBBLE_ERR_CUDA);  // m_isOverload = false, fft.cpp:26
    }
    PB_CUDA(cudaMemcpy(h->d_tw, tw.data(), nb, cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(h->d_win, win.data(), sizeof(double) * (size_t)spb, cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    h->cur = 0;
    h->params_set = true;
    return 0;
}
int pebble_fft_size(const pebble_fft_t *h) { return h ? h->N : PEBBLE_ERR_ARG; }
int pebble_fft_set_backend(pebble_fft_t *h, int backend) {
    if (!h || (backend != PEBBLE_FFT_FFTW && backend != PEBBLE_FFT_OOURA)) return PEBBLE_ERR_ARG;
    h->backend = backend;
    return 0;
}

int pebble_fft_spectrum_batch_device(pebble_fft_t *h, int frames, int n, const double *d_in, double *d_out, int *d_flags,
                                     void *stream_v) {
    if (!h || frames < 0 || n <= 0 || !d_in || !d_out) { pebble::set_error("fft: bad argument"); return PEBBLE_ERR_ARG; }
    if (!h->params_set) { pebble::set_error("fft: fftParams not called"); return PEBBLE_ERR_STATE; }  // fftw.cpp:72-73
    if (n > h->N) { pebble::set_error("fft: numSamples %d > fftSize %d", n, h->N); return PEBBLE_ERR_ARG; }
    if (frames == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream_v;
#define X(NN) spec_launch<NN>(h, frames, n, (const cpx *)d_in, d_out, d_flags, st)
    cudaError_t e = FFT_DISPATCH(h, X);
#undef X
    PB_CUDA(e, PEBBLE_ERR_CUDA);
    h->cur ^= 1;
    return frames;
}

int pebble_fft_spectrum(pebble_fft_t *h, const double *in, double *out_db, int n) {
    if (!h || !in || !out_db || n <= 0) { pebble::set_error("fft: bad argument"); return PEBBLE_ERR_ARG; }
    if (!h->params_set) return 0;  // fftw.cpp:72-73 returns false
    if (n > h->N) { pebble::set_error("fft: numSamples %d > fftSize %d", n, h->N); return PEBBLE_ERR_ARG; }
    const size_t bi = sizeof(cpx) * (size_t)n, bo = sizeof(double) * (size_t)h->N;
    if (!h->stg.reserve(bi, bo)) return PEBBLE_ERR_CUDA;
    memcpy(h->stg.h_in, in, bi);
    PB_CUDA(cudaMemcpyAsync(h->stg.d_in, h->stg.h_in, bi, cudaMemcpyHostToDevice, h->stream), PEBBLE_ERR_CUDA);
    int r = pebble_fft_spectrum_batch_device(h, 1, n, (const double *)h->stg.d_in, (double *)h->stg.d_out, nullptr, h->stream);
    if (r < 0) return r;
    PB_CUDA(cudaMemcpyAsync(h->stg.h_out, h->stg.d_out, bo, cudaMemcpyDeviceToHost, h->stream), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpyAsync(h->h_flag, h->d_flag[h->cur], sizeof(int), cudaMemcpyDeviceToHost, h->stream), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamSynchronize(h->stream), PEBBLE_ERR_CUDA);
    memcpy(out_db, h->stg.h_out, bo);
    return *h->h_flag ? 1 : 0;
}

static int fft_c2c_host(pebble_fft_t *h, const double *in, double *out, int n, bool inverse) {
    if (!h || !in || !out || n <= 0) {
        pebble::set_error("fft: in==NULL/out==NULL internal-buffer forms are not part of this ABI");
        return PEBBLE_ERR_ARG;
    }
    if (!h->params_set) return 0;  // fftw.cpp:33-34: silent return
    if (n > h->N) { pebble::set_error("fft: numSamples %d > fftSize %d", n, h->N); return PEBBLE_ERR_ARG; }
    const size_t bi = sizeof(cpx) * (size_t)n, bo = sizeof(cpx) * (size_t)h->N;
    if (!h->stg.reserve(bi, bo)) return PEBBLE_ERR_CUDA;
    memcpy(h->stg.h_in, in, bi);
    PB_CUDA(cudaMemcpyAsync(h->stg.d_in, h->stg.h_in, bi, cudaMemcpyHostToDevice, h->stream), PEBBLE_ERR_CUDA);
    cudaError_t e;
    if (inverse && h->backend == PEBBLE_FFT_OOURA) {
        // CFft::fftInverse = bitrv2conj + cftbsub (fftcute.cpp:262-263): Ooura's "backward" transform is the e^{-j} one, i.e.
        // the same sign as FFTW_FORWARD; no window on this path
#define X(NN) c2c_launch<NN, false>(h, n, (const cpx *)h->stg.d_in, (cpx *)h->stg.d_out, h->stream, false)
        e = FFT_DISPATCH(h, X);
#undef X
        h->cur ^= 1;
        cudaMemcpyAsync(h->d_prev[h->cur], h->d_prev[h->cur ^ 1], sizeof(double) * (size_t)h->N, cudaMemcpyDeviceToDevice, h->stream);
    } else if (inverse) {
#define X(NN) c2c_launch<NN, true>(h, n, (const cpx *)h->stg.d_in, (cpx *)h->stg.d_out, h->stream)
        e = FFT_DISPATCH(h, X);
#undef X
    } else {
#define X(NN) c2c_launch<NN, false>(h, n, (const cpx *)h->stg.d_in, (cpx *)h->stg.d_out, h->stream)
        e = FFT_DISPATCH(h, X);
#undef X
        h->cur ^= 1;  // the forward kernel wrote the flag into the other slot; amplitudes follow it
        cudaMemcpyAsync(h->d_prev[h->cur], h->d_prev[h->cur ^ 1], sizeof(double) * (size_t)h->N, cudaMemcpyDeviceToDevice, h->stream);
    }
    PB_CUDA(e, PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpyAsync(h->stg.h_out, h->stg.d_out, bo, cudaMemcpyDeviceToHost, h->stream), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamSynchronize(h->stream), PEBBLE_ERR_CUDA);
    if (!inverse && h->backend == PEBBLE_FFT_OOURA) {
        // CFft::fftForward swaps I and Q of the windowed input before Ooura's transform (fftcute.cpp:194-201); for that
        // transform's sign this equals the FFTW_FORWARD result with re and im exchanged, bin for bin
        const double *src = (const double *)h->stg.h_out;
        for (int i = 0; i < h->N; ++i) { out[2 * i] = src[2 * i + 1]; out[2 * i + 1] = src[2 * i]; }
    } else {
        memcpy(out, h->stg.h_out, bo);
    }
    return h->N;
}
int pebble_fft_forward(pebble_fft_t *h, const double *in, double *out, int n) { return fft_c2c_host(h, in, out, n, false); }
int pebble_fft_inverse(pebble_fft_t *h, const double *in, double *out, int n) { return fft_c2c_host(h, in, out, n, true); }

double pebble_window(int wtype, int n, double *w_out) {
    std::vector<double> w;
    double cg = pebble_host::window_table(wtype, n, w);
    if (w_out && !w.empty()) memcpy(w_out, w.data(), sizeof(double) * (size_t)n);
    return cg;
}

// state: m_fftPower[N] then m_fftAmplitude[N] (fft.cpp:387-388); power is derived: psd = asd^2 * maxBinPower
int pebble_fft_get_state(pebble_fft_t *h, double *buf, int64_t count) {
    if (!h || !buf || !h->params_set || count < 2 * (int64_t)h->N) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(buf + h->N, h->d_prev[h->cur], sizeof(double) * (size_t)h->N, cudaMemcpyDeviceToHost), PEBBLE_ERR_CUDA);
    for (int i = 0; i < h->N; ++i) buf[i] = buf[h->N + i] * buf[h->N + i] * h->maxbin;
    return 0;
}
int pebble_fft_set_state(pebble_fft_t *h, const double *buf, int64_t count) {
    if (!h || !buf || !h->params_set || count < 2 * (int64_t)h->N) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(h->d_prev[h->cur], buf + h->N, sizeof(double) * (size_t)h->N, cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    return 0;
}


// ---- SignalSpectrum (application/signalspectrum.cpp:6-131): two FFT objects behind the display-rate gate -------------------
struct pebble_spectrum {
    pebble_fft *fft_un = nullptr, *fft_hi = nullptr;    // m_fftUnprocessed, m_fftHiRes
    unsigned sample_rate = 0, hires_rate = 0, num_samples = 0;
    int bins = 0, hires_bins = 0;
    int updates_per_sec = 0;
    long long timer_update = 0, hires_timer_update = 0;   // ms between spectra (1000 / updatesPerSec, integer division)
    bool timer_valid = false, hires_timer_valid = false;
    long long timer_t0 = 0, hires_timer_t0 = 0;
    bool use_hires = false, overload = false, update_complete = true;
    int overrun = 0;
    std::vector<double> un, hi;                           // m_unprocessedSpectrum, m_hiResSpectrum
};

pebble_spectrum_t *pebble_spectrum_create(unsigned sample_rate, unsigned hires_sample_rate, unsigned buffer_size, int num_bins,
                                          int num_hires_bins, int updates_per_sec) {
    if (updates_per_sec <= 0 || num_bins <= 0 || num_hires_bins <= 0 || buffer_size == 0) {
        pebble::set_error("spectrum: bad argument (the reference divides 1000 by updatesPerSec in its constructor)");
        return nullptr;
    }
    pebble_spectrum *h = new pebble_spectrum();
    h->fft_un = pebble_fft_create();
    h->fft_hi = pebble_fft_create();
    if (!h->fft_un || !h->fft_hi) { pebble_spectrum_destroy(h); return nullptr; }
    h->num_samples = buffer_size; h->bins = num_bins; h->hires_bins = num_hires_bins;
    h->updates_per_sec = updates_per_sec;
    h->timer_update = h->hires_timer_update = 1000 / updates_per_sec;        // signalspectrum.cpp:32-33
    if (pebble_spectrum_set_sample_rate(h, sample_rate, hires_sample_rate) < 0) { pebble_spectrum_destroy(h); return nullptr; }
    return h;
}
void pebble_spectrum_destroy(pebble_spectrum_t *h) {
    if (!h) return;
    pebble_fft_destroy(h->fft_un);
    pebble_fft_destroy(h->fft_hi);
    delete h;
}
// SignalSpectrum::setSampleRate (signalspectrum.cpp:54-61): both objects BLACKMANHARRIS over numSamples
int pebble_spectrum_set_sample_rate(pebble_spectrum_t *h, unsigned sample_rate, unsigned hires_sample_rate) {
    if (!h) return PEBBLE_ERR_ARG;
    h->sample_rate = sample_rate; h->hires_rate = hires_sample_rate;
    int r = pebble_fft_params(h->fft_un, (unsigned)h->bins, 0.0, (double)sample_rate, (int)h->num_samples, PEBBLE_WINDOW_BLACKMANHARRIS);
    if (r < 0) return r;
    r = pebble_fft_params(h->fft_hi, (unsigned)h->hires_bins, 0.0, (double)hires_sample_rate, (int)h->num_samples, PEBBLE_WINDOW_BLACKMANHARRIS);
    if (r < 0) return r;
    h->bins = pebble_fft_size(h->fft_un);            // fftParams clamps to >= 2048 (fft.cpp:74-79)
    h->hires_bins = pebble_fft_size(h->fft_hi);
    h->un.assign((size_t)h->bins, 0.0);
    h->hi.assign((size_t)h->hires_bins, 0.0);
    return 0;
}
void pebble_spectrum_set_hires(pebble_spectrum_t *h, int on) { if (h) h->use_hires = on != 0; }
// SignalSpectrum::setUpdatesPerSec (signalspectrum.cpp:124-133): 0 stops the updates and keeps the old interval
void pebble_spectrum_set_updates_per_sec(pebble_spectrum_t *h, int u) {
    if (!h) return;
    h->updates_per_sec = u;
    if (u > 0) h->timer_update = h->hires_timer_update = 1000 / u;
}
// SignalSpectrum::unprocessed (signalspectrum.cpp:63-86). now_ms = the caller's monotonic clock in milliseconds (the
// reference reads a QElapsedTimer). Returns 1 when a new spectrum was made (the reference emits newFftData()), 0 when the
// gate skipped the call, < 0 on error.
int pebble_spectrum_unprocessed(pebble_spectrum_t *h, const double *in, int n, long long now_ms) {
    if (!h || !in) return PEBBLE_ERR_ARG;
    if (!h->timer_valid) { h->timer_valid = true; h->timer_t0 = now_ms; return 0; }       // first call only starts the timer
    if (h->updates_per_sec == 0 || now_ms - h->timer_t0 < h->timer_update) return 0;
    h->timer_t0 = now_ms;
    if (!h->update_complete) h->overrun++;
    const int r = pebble_fft_spectrum(h->fft_un, in, h->un.data(), n);                     // makeSpectrum, :108-112
    if (r < 0) return r;
    h->overload = r == 1;
    h->update_complete = false;
    return 1;
}
// SignalSpectrum::zoomed (signalspectrum.cpp:89-106)
int pebble_spectrum_zoomed(pebble_spectrum_t *h, const double *in, int n, long long now_ms) {
    if (!h || !in) return PEBBLE_ERR_ARG;
    if (!h->use_hires) return 0;
    if (!h->hires_timer_valid) { h->hires_timer_valid = true; h->hires_timer_t0 = now_ms; return 0; }
    if (h->updates_per_sec == 0 || now_ms - h->hires_timer_t0 < h->hires_timer_update) return 0;
    h->hires_timer_t0 = now_ms;
    const int r = pebble_fft_spectrum(h->fft_hi, in, h->hi.data(), n);
    if (r < 0) return r;
    h->overload = r == 1;
    return 1;
}
int pebble_spectrum_bin_count(const pebble_spectrum_t *h) { return h ? h->bins : PEBBLE_ERR_ARG; }
int pebble_spectrum_hires_bin_count(const pebble_spectrum_t *h) { return h ? h->hires_bins : PEBBLE_ERR_ARG; }
int pebble_spectrum_overload(const pebble_spectrum_t *h) { return h ? (h->overload ? 1 : 0) : PEBBLE_ERR_ARG; }
int pebble_spectrum_overrun(const pebble_spectrum_t *h) { return h ? h->overrun : PEBBLE_ERR_ARG; }
// m_displayUpdateComplete: set false by unprocessed(), set true again by the consumer (the widget) through ack
void pebble_spectrum_ack_update(pebble_spectrum_t *h) { if (h) h->update_complete = true; }
const double *pebble_spectrum_get_unprocessed(const pebble_spectrum_t *h) { return h ? h->un.data() : nullptr; }
const double *pebble_spectrum_get_hires(const pebble_spectrum_t *h) { return h ? h->hi.data() : nullptr; }

}  // extern "C"
