// fastfir_long.cu — CFastFIR for long filters: N-point overlap-save with N = N1 x 2048 too large for shared memory
// (BASELINE config 5: CONV_FFT_SIZE 131072 / CONV_FIR_SIZE 65537, fastfir.cpp:65-66 with the sizes lifted; algorithm
// fastfir.cpp:281-334 unchanged: FFT of L old + L new samples, x H, unnormalised inverse, keep samples L..N-1).
//
// Four-step transform, index split n = 2048 n1 + n2, k = k1 + N1 k2 (N1 = R x R, R = 8 for 131072, 4 for 32768):
//   cols fwd : for every column n2 an N1-point DFT over n1 (stride 2048)                 -> scratch[k1][n2]
//   rows     : x W_N^(n2 k1), 2048-point DFT over n2 (shared memory, fft_core.cuh), x H (CpxMpy, fastfir.cpp:324-334),
//              inverse 2048-point DFT, x conj W_N^(n2 k1)                                -> scratch[k1][n2] in place
//   cols inv : inverse N1-point DFT over k1; only n1 >= N1/2 (the L valid samples) is formed and written.
// H is kept in the rows kernel's own [k1][slot] order (slot = digit-reversed k2), so nothing is ever reordered.
//
// HBM: algorithmic 32 B per input sample (16 read + 16 written). The scratch round trips (3 x 32 B per sample) are kept
// out of DRAM by processing `chunk` blocks at a time with chunk x N x 16 B well inside the 126 MB L2: the three launches
// of a chunk run back to back on one stream, so the rows and the inverse columns find their input in L2.
#include "fastfir_long.cuh"
#include "fft_kernels.cuh"
#include "ff_fused.cuh"
#include <cstdlib>
#include <vector>

namespace {
constexpr int LF_N2 = 2048;
constexpr int LF_TC = 16;   // columns per CTA of the column kernels: 16 x 16 B = 256 B contiguous per row

__device__ __forceinline__ cpx lf_ld_stream(const cpx *p) {
    cpx r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
    return r;
}
__device__ __forceinline__ void lf_st_stream(cpx *p, cpx v) {
    asm volatile("st.global.cs.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(v.x), "d"(v.y) : "memory");
}

struct LfGeom {
    int nblocks;            // blocks per stream in this call
    int L;                  // N / 2
    int s0;                 // samples of the virtual stream that come from `state`
    int state_stride;       // cpx between the states of consecutive streams
    long long in_stride, out_stride;
    long long job0;         // first job (stream * nblocks + block) of this launch
};

// Column transforms of LF_TC adjacent columns of one block. Thread (c, i): column c, butterfly i of each radix-R pass.
// Forward: pass 0 on rows i + R k (twiddle W_N1^(i q) after the DFT), pass 1 on rows R i + k; slot R i + q = bin i + R q.
// Inverse: the mirror image (pass 1, conj twiddle, pass 0).
template <int R, bool INV, int MINB>
__global__ void __launch_bounds__(LF_TC *R, MINB)
lf_cols_kernel(const cpx *__restrict__ state, const cpx *__restrict__ in, cpx *__restrict__ out, cpx *__restrict__ scratch,
               const cpx *__restrict__ twN, LfGeom g, double scale) {
    constexpr int N1 = R * R;
    __shared__ __align__(16) cpx a[N1 * LF_TC];
    __shared__ __align__(16) cpx w[N1];
    const int t = threadIdx.x, c = t % LF_TC, i = t / LF_TC;
    const long long job = g.job0 + blockIdx.y;
    const int s = (int)(job / g.nblocks), b = (int)(job % g.nblocks);
    const int col = blockIdx.x * LF_TC + c;
    cpx *sc = scratch + (size_t)blockIdx.y * ((size_t)N1 * LF_N2) + col;
    if (t < N1) w[t] = twN[(size_t)t * LF_N2];   // W_N1^t = W_N^(2048 t)
    cpx v[R];
    if (!INV) {
        const cpx *st = state + (long long)s * g.state_stride;
        const cpx *ip = in + s * g.in_stride;
        const long long v0 = (long long)b * g.L + col;
#pragma unroll
        for (int k = 0; k < R; ++k) {
            const long long u = v0 + (long long)LF_N2 * (i + R * k);
            v[k] = lf_ld_stream(u < g.s0 ? st + u : ip + (u - g.s0));
        }
        __syncthreads();   // w[]
        DftR<R, false>::run(v);
#pragma unroll
        for (int q = 0; q < R; ++q) a[(i + R * q) * LF_TC + c] = q ? cpx_mul(v[q], w[i * q]) : v[q];
        __syncthreads();
#pragma unroll
        for (int k = 0; k < R; ++k) v[k] = a[(R * i + k) * LF_TC + c];
        DftR<R, false>::run(v);
#pragma unroll
        for (int q = 0; q < R; ++q) sc[(size_t)(i + R * q) * LF_N2] = v[q];
    } else {
#pragma unroll
        for (int k = 0; k < R; ++k) v[k] = sc[(size_t)(i + R * k) * LF_N2];
        DftR<R, true>::run(v);
#pragma unroll
        for (int q = 0; q < R; ++q) a[(R * i + q) * LF_TC + c] = v[q];
        __syncthreads();
#pragma unroll
        for (int q = 0; q < R; ++q) {
            v[q] = a[(i + R * q) * LF_TC + c];
            if (q) v[q] = cpx_mulc(v[q], w[i * q]);
        }
        DftR<R, true>::run(v);
        // v[k] = sample n1 = i + R k of column col; the valid half is n1 >= N1/2 (fastfir.cpp:308-311)
        cpx *op = out + s * g.out_stride + (long long)b * g.L + col;
#pragma unroll
        for (int k = R / 2; k < R; ++k) lf_st_stream(op + (long long)LF_N2 * (i + R * k - N1 / 2), cpx_scale(v[k], scale));
    }
}

// One row k1 of one block: twiddle, 2048-point forward, x H, inverse, conj twiddle, in place. FILTER: forward only -> H.
template <bool FILTER>
__global__ void __launch_bounds__(FftPlan<LF_N2>::THREADS)
lf_rows_kernel(cpx *__restrict__ scratch, const cpx *__restrict__ hpos, cpx *__restrict__ hout, const cpx *__restrict__ tw2,
               const cpx *__restrict__ twN, int N1) {
    __shared__ __align__(128) cpx a[LF_N2];
    constexpr int SH = FftPlan<LF_N2>::SH;
    const int k1 = blockIdx.x;
    cpx *row = scratch + ((size_t)blockIdx.y * N1 + k1) * LF_N2;
    for (int n2 = threadIdx.x; n2 < LF_N2; n2 += blockDim.x) a[fft_swz<SH>(n2)] = cpx_mul(row[n2], twN[n2 * k1]);
    __syncthreads();
    fft_forward_smem<LF_N2>(a, tw2);
    if (FILTER) {
        for (int p = threadIdx.x; p < LF_N2; p += blockDim.x) hout[(size_t)k1 * LF_N2 + p] = a[fft_swz<SH>(p)];
        return;
    }
    for (int p = threadIdx.x; p < LF_N2; p += blockDim.x)
        a[fft_swz<SH>(p)] = cpx_mul(hpos[(size_t)k1 * LF_N2 + p], a[fft_swz<SH>(p)]);
    __syncthreads();
    fft_inverse_smem<LF_N2>(a, tw2);
    for (int n2 = threadIdx.x; n2 < LF_N2; n2 += blockDim.x) row[n2] = cpx_mulc(a[fft_swz<SH>(n2)], twN[n2 * k1]);
}

// H of every row from the generic kernel's slot order to the fused schedule's [k][thread] order (ff_fused.cuh).
__global__ void lf_permute_kernel(const cpx *__restrict__ hpos, cpx *__restrict__ ht) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;   // idx = k * 128 + t  <->  slot 16 t + k
    const size_t r = (size_t)blockIdx.y * LF_N2;
    ht[r + idx] = hpos[r + FftPlan<LF_N2>::freq2pos(ff2048_pos2freq(16 * (idx & 127) + (idx >> 7)))];
}

// The process-path row kernel on the register-fused 8 x 16 x 16 schedule of fastfir2048_kernel (8 shared-memory accesses
// per point instead of 16): F0 reads the row from scratch (L2) and applies the inter-step twiddle W_N^(n2 k1) on the way
// in, I0 applies its conjugate on the way out. n2 = i + 256 k: W_N^(n2 k1) = W_N^(i k1) * W_N^(256 k k1), one gather per
// (thread, g) and eight CTA-uniform factors, both exact table entries.
template <int MINB>
__global__ void __launch_bounds__(128, MINB)
lf_rows_fused_kernel(cpx *__restrict__ scratch, const cpx *__restrict__ ht_all, const cpx *__restrict__ tw2,
                     const cpx *__restrict__ twN, int N1, int nmask) {
    __shared__ __align__(128) cpx a[LF_N2 + 64 + 8];
    cpx *t1 = a + LF_N2, *pk = a + LF_N2 + 64;
    const int t = threadIdx.x, k1 = blockIdx.x;
    cpx *row = scratch + ((size_t)blockIdx.y * N1 + k1) * LF_N2;
    const cpx *ht = ht_all + (size_t)k1 * LF_N2;
    if (t < 64) t1[t] = tw2[8 * ((t & 15) << (t >> 4))];   // t = 16 b + i -> exp(-j 2 pi i 2^b / 256)
    if (t >= 64 && t < 72) pk[t - 64] = twN[(256 * (t - 64) * k1) & nmask];
    const cpx wa1 = tw2[t], wa2 = tw2[2 * t], wa4 = tw2[4 * t];
    const cpx wb1 = tw2[t + 128], wb2 = tw2[2 * t + 256], wb4 = tw2[4 * t + 512];
    const cpx ba = twN[t * k1], bb = twN[(t + 128) * k1];
    __syncthreads();
#pragma unroll
    for (int g = 0; g < 2; ++g) {
        const int i = t + 128 * g;
        const cpx bg = g ? bb : ba;
        cpx u[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) u[k] = __ldcg(row + i + 256 * k);
#pragma unroll
        for (int k = 0; k < 8; ++k) u[k] = cpx_mul(u[k], k ? cpx_mul(bg, pk[k]) : bg);
        dft8<false>(u);
        ff_tw8<false>(u, g ? wb1 : wa1, g ? wb2 : wa2, g ? wb4 : wa4);
#pragma unroll
        for (int k = 0; k < 8; ++k) a[ff_swz(i + 256 * k)] = u[k];
    }
    __syncthreads();
    ff_mid_phases<128, false>(a, t1, ht, t, 0);
#pragma unroll
    for (int g = 0; g < 2; ++g) {
        const int i = t + 128 * g;
        const cpx bg = g ? bb : ba;
        cpx u[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) u[k] = a[ff_swz(i + 256 * k)];
        ff_tw8<true>(u, g ? wb1 : wa1, g ? wb2 : wa2, g ? wb4 : wa4);
        dft8<true>(u);
#pragma unroll
        for (int k = 0; k < 8; ++k) __stcg(row + i + 256 * k, cpx_mulc(u[k], k ? cpx_mul(bg, pk[k]) : bg));
    }
}
}  // namespace

struct LongFir {
    int N, N1, R;
    const cpx *d_twN;
    cpx *d_tw2;        // exp(-2 pi j m / 2048)
    cpx *d_scratch;
    cpx *d_ht;         // H in the fused row kernel's order (lf_permute_kernel)
    bool fused;        // PEBBLE_LF_GENERIC=1 keeps the generic in-place row kernel (cross-check)
    int chunk;         // blocks per launch triple
    // Consecutive chunks rotate over `lanes` (default 3) internal streams (own scratch each), forked from and joined to the
    // caller's stream by events: the column kernels are L2-bandwidth-bound and the row kernel FP64/shared-memory-bound,
    // so chunk c's rows overlap chunk c+1's forward columns.
    int lanes;
    bool cols12;       // column kernels compiled for 12 CTAs per SM (40 registers) instead of 9 (54)
    cudaStream_t lane[3];
    cudaEvent_t fork, join[3];
};

bool lf_size_supported(int fft_size) { return fft_size == 32768 || fft_size == 131072; }

LongFir *lf_create(int fft_size, const cpx *d_twN) {
    if (!lf_size_supported(fft_size)) return nullptr;
    LongFir *p = new LongFir();
    p->N = fft_size; p->N1 = fft_size / LF_N2; p->R = p->N1 == 64 ? 8 : 4; p->d_twN = d_twN;
    p->d_tw2 = nullptr; p->d_scratch = nullptr; p->d_ht = nullptr; p->fused = !getenv("PEBBLE_LF_GENERIC");
    // default (N = 131072): 9 blocks per chunk and three lanes = 54 MB of scratch in flight, inside the 126 MB L2 together
    // with the streaming input/output. 9 x 64 = 576 row CTAs = one wave of 148 SMs x 4; 1152 column CTAs = 0.86 wave of
    // 148 x 9: every launch is a single wave and three of them (one per lane) are in flight. Sweep in
    // profiles/r1_longfir_history.md: lanes x chunk = 2x18 9.51 ms, 3x12 9.48, 3x9 9.12-9.15, 3x6 9.82, 3x18 12.3 (L2 spill).
    const char *e = getenv("PEBBLE_LF_CHUNK_MB");
    const int mb = e && atoi(e) > 0 ? atoi(e) : 18;
    p->chunk = (int)(((size_t)mb << 20) / (sizeof(cpx) * (size_t)fft_size));
    if ((e = getenv("PEBBLE_LF_CHUNK_BLOCKS")) && atoi(e) > 0) p->chunk = atoi(e);
    if (p->chunk < 1) p->chunk = 1;
    e = getenv("PEBBLE_LF_COLS_MINB");
    p->cols12 = e && atoi(e) == 12;
    e = getenv("PEBBLE_LF_LANES");
    p->lanes = e && atoi(e) >= 1 && atoi(e) <= 3 ? atoi(e) : 3;
    for (int l = 0; l < 3; ++l) { p->lane[l] = nullptr; p->join[l] = nullptr; }
    p->fork = nullptr;
    bool ok = pebble::cuda_ok(cudaMalloc(&p->d_tw2, sizeof(cpx) * LF_N2), "cudaMalloc") &&
              pebble::cuda_ok(cudaMalloc(&p->d_ht, sizeof(cpx) * (size_t)fft_size), "cudaMalloc") &&
              pebble::cuda_ok(cudaMemset(p->d_ht, 0, sizeof(cpx) * (size_t)fft_size), "cudaMemset") &&
              pebble::cuda_ok(cudaMalloc(&p->d_scratch, sizeof(cpx) * (size_t)fft_size * p->chunk * p->lanes), "cudaMalloc") &&
              pebble::cuda_ok(cudaEventCreateWithFlags(&p->fork, cudaEventDisableTiming), "cudaEventCreate") &&
              // W_2048^m = W_N^(N1 m): strided copy out of the caller's table
              pebble::cuda_ok(cudaMemcpy2D(p->d_tw2, sizeof(cpx), d_twN, sizeof(cpx) * p->N1, sizeof(cpx), LF_N2,
                                           cudaMemcpyDeviceToDevice), "cudaMemcpy2D(tw2)");
    for (int l = 0; ok && l < p->lanes && p->lanes > 1; ++l)
        ok = pebble::cuda_ok(cudaStreamCreateWithFlags(&p->lane[l], cudaStreamNonBlocking), "cudaStreamCreate") &&
             pebble::cuda_ok(cudaEventCreateWithFlags(&p->join[l], cudaEventDisableTiming), "cudaEventCreate");
    if (!ok) { lf_destroy(p); return nullptr; }
    // Keep the scratch in L2: a persisting-access window over each lane's scratch on that lane's stream (the set-aside is
    // a device-wide limit, sized to the scratch and capped by the device maximum). Everything else the kernels touch
    // (input, output, H, twiddles) lies outside the windows and keeps the normal policy. PEBBLE_LF_NOPERSIST=1 disables.
    if (p->lanes > 1 && !getenv("PEBBLE_LF_NOPERSIST")) {
        int dev = 0, max_persist = 0, max_win = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, dev);
        cudaDeviceGetAttribute(&max_win, cudaDevAttrMaxAccessPolicyWindowSize, dev);
        const size_t lane_bytes = sizeof(cpx) * (size_t)fft_size * p->chunk;
        size_t want = lane_bytes * p->lanes;
        if (want > (size_t)max_persist) want = (size_t)max_persist;
        if (want > 0 && max_win > 0 && cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want) == cudaSuccess) {
            for (int l = 0; l < p->lanes; ++l) {
                cudaStreamAttrValue v = {};
                v.accessPolicyWindow.base_ptr = p->d_scratch + (size_t)l * p->chunk * (size_t)fft_size;
                v.accessPolicyWindow.num_bytes = lane_bytes < (size_t)max_win ? lane_bytes : (size_t)max_win;
                const double share = (double)(want / p->lanes) / (double)v.accessPolicyWindow.num_bytes;
                v.accessPolicyWindow.hitRatio = share < 1.0 ? (float)share : 1.0f;
                v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
                v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
                cudaStreamSetAttribute(p->lane[l], cudaStreamAttributeAccessPolicyWindow, &v);
            }
        }
        cudaGetLastError();   // the policy is an optimisation: a refusal is not an error
    }
    return p;
}

void lf_destroy(LongFir *p) {
    if (!p) return;
    cudaFree(p->d_tw2); cudaFree(p->d_scratch); cudaFree(p->d_ht);
    for (int l = 0; l < 3; ++l) {
        if (p->lane[l]) cudaStreamDestroy(p->lane[l]);
        if (p->join[l]) cudaEventDestroy(p->join[l]);
    }
    if (p->fork) cudaEventDestroy(p->fork);
    delete p;
}

int lf_freq2index(int fft_size, int f) {
    const int N1 = fft_size / LF_N2;
    return (f % N1) * LF_N2 + FftPlan<LF_N2>::freq2pos(f / N1);
}

template <int R>
static cudaError_t lf_run(LongFir *p, const cpx *hpos, cpx *hout, const cpx *state, const cpx *in, cpx *out, LfGeom g,
                          long long jobs, double scale, cudaStream_t st) {
    const dim3 cgrid(LF_N2 / LF_TC, 1), cblock(LF_TC * R);
    const int lanes = (hout || jobs <= p->chunk) ? 1 : p->lanes;
    if (lanes > 1) {
        cudaEventRecord(p->fork, st);
        for (int l = 0; l < lanes; ++l) cudaStreamWaitEvent(p->lane[l], p->fork, 0);
    }
    int c = 0;
    for (long long j0 = 0; j0 < jobs; j0 += p->chunk, ++c) {
        const unsigned nj = (unsigned)(jobs - j0 < p->chunk ? jobs - j0 : p->chunk);
        cudaStream_t cs = lanes > 1 ? p->lane[c % lanes] : st;
        cpx *scr = p->d_scratch + (size_t)(c % lanes) * p->chunk * (size_t)p->N;
        g.job0 = j0;
        if (p->cols12) lf_cols_kernel<R, false, 12><<<dim3(cgrid.x, nj), cblock, 0, cs>>>(state, in, nullptr, scr, p->d_twN, g, 1.0);
        else lf_cols_kernel<R, false, 1><<<dim3(cgrid.x, nj), cblock, 0, cs>>>(state, in, nullptr, scr, p->d_twN, g, 1.0);
        if (hout) {
            lf_rows_kernel<true><<<dim3(p->N1, nj), FftPlan<LF_N2>::THREADS, 0, cs>>>(scr, nullptr, hout, p->d_tw2, p->d_twN, p->N1);
            pebble::count_launch(2);
            continue;
        }
        if (p->fused)
            lf_rows_fused_kernel<4><<<dim3(p->N1, nj), 128, 0, cs>>>(scr, p->d_ht, p->d_tw2, p->d_twN, p->N1, p->N - 1);
        else
            lf_rows_kernel<false><<<dim3(p->N1, nj), FftPlan<LF_N2>::THREADS, 0, cs>>>(scr, hpos, nullptr, p->d_tw2, p->d_twN, p->N1);
        if (p->cols12) lf_cols_kernel<R, true, 12><<<dim3(cgrid.x, nj), cblock, 0, cs>>>(nullptr, nullptr, out, scr, p->d_twN, g, scale);
        else lf_cols_kernel<R, true, 1><<<dim3(cgrid.x, nj), cblock, 0, cs>>>(nullptr, nullptr, out, scr, p->d_twN, g, scale);
        pebble::count_launch(3);
    }
    if (lanes > 1) {
        for (int l = 0; l < lanes; ++l) {
            cudaEventRecord(p->join[l], p->lane[l]);
            cudaStreamWaitEvent(st, p->join[l], 0);
        }
    }
    return cudaGetLastError();
}

cudaError_t lf_build_filter(LongFir *p, const cpx *d_taps, cpx *d_hpos, cudaStream_t st) {
    // the taps as a one-block stream with no carried state: V = taps
    LfGeom g = {1, p->N / 2, 0, 0, 0, 0, 0};
    cudaError_t e = p->R == 8 ? lf_run<8>(p, nullptr, d_hpos, d_taps, d_taps, nullptr, g, 1, 1.0, st)
                              : lf_run<4>(p, nullptr, d_hpos, d_taps, d_taps, nullptr, g, 1, 1.0, st);
    if (e != cudaSuccess) return e;
    lf_permute_kernel<<<dim3(LF_N2 / 256, p->N1), 256, 0, st>>>(d_hpos, p->d_ht);
    pebble::count_launch();
    return cudaGetLastError();
}

cudaError_t lf_process(LongFir *p, const cpx *d_hpos, const cpx *state, int s0, int state_stride, const cpx *in,
                       long long in_stride, cpx *out, long long out_stride, int nblocks, int streams, double scale,
                       cudaStream_t st) {
    LfGeom g = {nblocks, p->N / 2, s0, state_stride, in_stride, out_stride, 0};
    const long long jobs = (long long)nblocks * streams;
    return p->R == 8 ? lf_run<8>(p, d_hpos, nullptr, state, in, out, g, jobs, scale, st)
                     : lf_run<4>(p, d_hpos, nullptr, state, in, out, g, jobs, scale, st);
}
