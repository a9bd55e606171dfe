// decimator.cu — the reference's ALTERNATIVE decimation chain (`Decimator`, pebblelib/decimator.cpp) on the GPU:
// kernel + host planner + C ABI (include/pebble_cuda.h, pebble_decimator_*).
//
// What the reference does (decimator.cpp:64-149, 152-226, 593-659, 695-740):
//   * buildDecimationChain walks the rate down by 2 per step, picking CIC3 / hb11 ... hb59 by `rate >= protectBw / wPass`
//     (float rate, double threshold), and — m_combineStages is true — MERGES consecutive steps that picked the same
//     design into ONE stage that filters once and keeps every D-th sample (D = 2^merged). That is not the same
//     arithmetic as a cascade and is reproduced as is.
//   * a half-band stage computes y[m] = sum_{k<taps} h[k] * X[m*D + k], X = (taps-1 carried samples) ++ x, k ascending;
//   * a CIC3 stage computes y[m] = .125*(x[mD+1] + x[(m-1)D] + 3*(x[(m-1)D+1] + x[mD])) (combined: samples mD+2.. unused).
// Every stage is a pure FIR (no recursion), so for calls whose length is a multiple of the total decimation the chain is
// a function of the input STREAM alone: y_s[m] depends on x_s[m*D - back .. m*D + fwd - 1] with zeros before the stream
// start. The carried state of all stages is therefore re-derivable from the last `hist` RAW input samples, which is the
// only state this implementation keeps (bit-identical: the recomputed values go through the same operations in the
// same order; additions and multiplications are done with __dadd_rn/__dmul_rn, never contracted to FMA).
//
// Kernel: one CTA per tile of `To` final outputs. The FRONT (stage 0, or CIC3 fused with stage 1 so the full-rate CIC3
// output never exists anywhere) reads the raw stream straight from global memory with 256-bit loads (every lane a full
// 32-byte sector), later stages ping-pong between two shared-memory buffers, the last stage writes global memory.
// Halo samples a tile needs from before its start are recomputed (pre-roll), so tiles are independent.
#include "lib_common.h"
#include "common.cuh"
#include "dec_coefs.h"
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#define PB_DEC_MAX_STAGES 16
#define PB_DEC_THREADS 256

struct DecPlan {
    int S;                         // stages after combining
    int front;                     // 0: stage 0 is a half-band read from global; 1: CIC3 alone; 2: CIC3 fused with stage 1
    int taps[PB_DEC_MAX_STAGES];   // 0 = CIC3
    int D[PB_DEC_MAX_STAGES];
    int back[PB_DEC_MAX_STAGES];   // stage-input samples needed before index m*D
    int fwd[PB_DEC_MAX_STAGES];    // stage-input samples needed from index m*D on
    int off[PB_DEC_MAX_STAGES];    // offset into c_dec_taps
    int To;                        // final outputs per tile
    int hist;                      // raw input samples of history carried between calls (even)
    int buf_len[2];                // shared-memory buffer lengths (cpx)
};

__constant__ double c_dec_taps[407];

// raw stream with virtual history: index i < 0 -> carried samples of the previous calls
struct DecIn {
    const cpx *in;
    const cpx *hist;
    int hist_len;
};
__device__ __forceinline__ void dec_ld_pair(const DecIn &s, long long i, bool vec, cpx &a, cpx &b) {
    // i is even (all stage decimations are even), so a pair never straddles the history boundary (hist_len is even)
    const cpx *p = i < 0 ? s.hist + (s.hist_len + i) : s.in + i;
    if (vec) {
        asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a.x), "=d"(a.y), "=d"(b.x), "=d"(b.y) : "l"(p));
    } else {
        a = __ldg(p);
        b = __ldg(p + 1);
    }
}
__device__ __forceinline__ cpx dec_ld(const DecIn &s, long long i) {
    return __ldg(i < 0 ? s.hist + (s.hist_len + i) : s.in + i);
}
// decimator.cpp:695-716, operation order kept: .125*( (odd + xEven) + 3.0*(xOdd + even) )
__device__ __forceinline__ double dec_cic(double odd, double x_even, double x_odd, double even) {
    return __dmul_rn(.125, __dadd_rn(__dadd_rn(odd, x_even), __dmul_rn(3.0, __dadd_rn(x_odd, even))));
}
__device__ __forceinline__ cpx dec_cic_global(const DecIn &s, long long j, int D, bool vec) {
    cpx even, odd, xe, xo;
    dec_ld_pair(s, j * D, vec, even, odd);
    dec_ld_pair(s, (j - 1) * D, vec, xe, xo);
    return make_double2(dec_cic(odd.x, xe.x, xo.x, even.x), dec_cic(odd.y, xe.y, xo.y, even.y));
}
// acc += x*h exactly as the scalar loop of decimator.cpp:637-647 / 366-376 (zero taps add nothing)
__device__ __forceinline__ void dec_mac(cpx &acc, cpx x, double h) {
    acc.x = __dadd_rn(acc.x, __dmul_rn(x.x, h));
    acc.y = __dadd_rn(acc.y, __dmul_rn(x.y, h));
}

__global__ void __launch_bounds__(PB_DEC_THREADS) dec_chain_kernel(const DecPlan p, const cpx *__restrict__ in,
                                                                   const cpx *__restrict__ hist, cpx *__restrict__ out,
                                                                   long long n_out, int vec) {
    extern __shared__ cpx dec_smem[];
    __shared__ long long s_a[PB_DEC_MAX_STAGES], s_b[PB_DEC_MAX_STAGES];
    const int S = p.S;
    if (threadIdx.x == 0) {
        long long a = (long long)blockIdx.x * p.To, b = a + p.To;
        if (b > n_out) b = n_out;
        s_a[S - 1] = a; s_b[S - 1] = b;
        for (int s = S - 1; s >= 1; --s) {
            b = (b - 1) * p.D[s] + p.fwd[s];
            a = a * p.D[s] - p.back[s];
            s_a[s - 1] = a; s_b[s - 1] = b;
        }
    }
    __syncthreads();
    const DecIn src = {in, hist, p.hist};
    cpx *const buf0 = dec_smem, *const buf1 = dec_smem + p.buf_len[0];
    const int f = p.front == 2 ? 1 : 0;      // index of the stage the front produces
    {
        const long long a = s_a[f];
        const int cnt = (int)(s_b[f] - a);
        cpx *dst = f == S - 1 ? out + a : buf0;
        if (p.front == 1) {
            const int D = p.D[0];
            for (int m = threadIdx.x; m < cnt; m += PB_DEC_THREADS) dst[m] = dec_cic_global(src, a + m, D, vec);
        } else if (p.front == 2 && p.taps[1] == 11) {
            // CIC3 + hb11 (the front of every chain planned from a rate >= 16x the protected bandwidth). The 11 CIC3 values
            // one output needs are built from 12 consecutive raw sample PAIRS, each shared by two of them; all 12 are
            // requested first, as independent 256-bit loads (12 x 32 B in flight per thread), and only then combined.
            // Tiles alternate direction: an even tile walks its range upwards, an odd one downwards, so the raw samples two
            // neighbours both need (the later tile's pre-roll = the end of the earlier tile's range) are asked for by
            // both at the same end of their walks and the second request is served from L2 instead of DRAM.
            const int D0 = p.D[0], D1 = p.D[1], bk = p.back[1];
            const double *h = c_dec_taps + p.off[1];
            const int iters = (cnt + PB_DEC_THREADS - 1) / PB_DEC_THREADS;
            const bool down = blockIdx.x & 1;
            for (int it = 0; it < iters; ++it) {
                const int m = (down ? iters - 1 - it : it) * PB_DEC_THREADS + threadIdx.x;
                if (m >= cnt) continue;
                const long long j0 = (a + m) * D1 - bk;
                cpx pa[12], pb[12];
#pragma unroll
                for (int q = 0; q < 12; ++q) dec_ld_pair(src, (j0 - 1 + q) * D0, vec, pa[q], pb[q]);
                cpx acc = make_double2(0.0, 0.0);
#pragma unroll
                for (int k = 0; k < 11; ++k) {
                    const double hk = h[k];
                    if (hk != 0.0) {     // CIC3 value j0+k: (even, odd) = pair k+1, (xEven, xOdd) = pair k
                        const cpx c = make_double2(dec_cic(pb[k + 1].x, pa[k].x, pb[k].x, pa[k + 1].x),
                                                   dec_cic(pb[k + 1].y, pa[k].y, pb[k].y, pa[k + 1].y));
                        dec_mac(acc, c, hk);
                    }
                }
                dst[m] = acc;
            }
        } else if (p.front == 2) {
            const int D0 = p.D[0], D1 = p.D[1], taps = p.taps[1], bk = p.back[1];
            const double *h = c_dec_taps + p.off[1];
            for (int m = threadIdx.x; m < cnt; m += PB_DEC_THREADS) {
                const long long j0 = (a + m) * D1 - bk;
                cpx acc = make_double2(0.0, 0.0);
                for (int k = 0; k < taps; ++k) {
                    const double hk = h[k];
                    if (hk != 0.0) dec_mac(acc, dec_cic_global(src, j0 + k, D0, vec), hk);
                }
                dst[m] = acc;
            }
        } else {
            const int D = p.D[0], taps = p.taps[0], bk = p.back[0];
            const double *h = c_dec_taps + p.off[0];
            for (int m = threadIdx.x; m < cnt; m += PB_DEC_THREADS) {
                const long long i0 = (a + m) * D - bk;
                cpx acc = make_double2(0.0, 0.0);
                for (int k = 0; k < taps; ++k) {
                    const double hk = h[k];
                    if (hk != 0.0) dec_mac(acc, dec_ld(src, i0 + k), hk);
                }
                dst[m] = acc;
            }
        }
    }
    for (int s = f + 1; s < S; ++s) {
        __syncthreads();
        const cpx *x = ((s - f - 1) & 1) ? buf1 : buf0;
        const long long xa = s_a[s - 1], a = s_a[s];
        const int cnt = (int)(s_b[s] - a);
        cpx *dst = s == S - 1 ? out + a : (((s - f) & 1) ? buf1 : buf0);
        const int D = p.D[s], taps = p.taps[s], bk = p.back[s];
        if (taps == 0) {
            for (int m = threadIdx.x; m < cnt; m += PB_DEC_THREADS) {
                const int i = (int)((a + m) * D - xa);
                const cpx even = x[i], odd = x[i + 1], xe = x[i - D], xo = x[i - D + 1];
                dst[m] = make_double2(dec_cic(odd.x, xe.x, xo.x, even.x), dec_cic(odd.y, xe.y, xo.y, even.y));
            }
        } else {
            const double *h = c_dec_taps + p.off[s];
            for (int m = threadIdx.x; m < cnt; m += PB_DEC_THREADS) {
                const int i0 = (int)((a + m) * D - bk - xa);
                cpx acc = make_double2(0.0, 0.0);
                for (int k = 0; k < taps; ++k) {
                    const double hk = h[k];
                    if (hk != 0.0) dec_mac(acc, x[i0 + k], hk);
                }
                dst[m] = acc;
            }
        }
    }
}

// new_hist = last H samples of (old_hist ++ in[0..n))
__global__ void dec_hist_kernel(const cpx *__restrict__ old_hist, const cpx *__restrict__ in, long long n,
                                cpx *__restrict__ new_hist, int H) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= H) return;
    const long long i = (long long)j + n - H;
    new_hist[j] = i >= 0 ? in[i] : old_hist[j + n];
}

namespace {
struct Stage { int taps, D; };

// Decimator::buildDecimationChain (decimator.cpp:64-149). The running rate is a float, the protected bandwidth a
// double, exactly as the reference declares them (decimator.h:250-251).
float dec_plan(unsigned rate_in, unsigned protect_bw, unsigned rate_out, std::vector<Stage> &chain, int *dec_by2) {
    using namespace pebble_dec;
    chain.clear();
    float rate = (float)rate_in;
    const double bw = (double)protect_bw;
    const unsigned min_out = rate_out > 0 ? rate_out : 15000u;   // minDecimatedSampleRate, decimator.h:245
    int stages = 0;
    while (rate > (float)min_out) {
        int taps = -1;
        if (rate >= (bw / kDecCicWpass)) {
            taps = 0;
        } else {
            for (int k = 1; k < kDecNumHb; ++k)       // k = 0 is hb7, which the reference leaves out of the chain (:81-86)
                if (rate >= (bw / kDecHbWpass[k])) { taps = kDecHbLen[k]; break; }
        }
        if (taps < 0) break;                          // "Ran out of filters before minimum sample rate"
        ++stages;
        if (chain.empty() || chain.back().taps != taps) chain.push_back({taps, 2});
        else chain.back().D *= 2;                     // m_combineStages (:130-143)
        rate /= 2;
    }
    if (dec_by2) *dec_by2 = stages;
    return rate;
}
}  // namespace

struct pebble_decimator {
    unsigned sample_rate, buffer_size;
    std::vector<Stage> chain;
    int dec_by2 = 0;
    float out_rate = 0.f;
    long long total_dec = 1;
    DecPlan plan;
    size_t smem = 0;
    cpx *d_hist[2] = {nullptr, nullptr};
    int cur = 0;
    cudaStream_t stream = nullptr;
    pebble::Staging stg;
};

static int dec_make_plan(pebble_decimator *h, long long n_hint) {
    using namespace pebble_dec;
    DecPlan &p = h->plan;
    memset(&p, 0, sizeof(p));
    const int S = (int)h->chain.size();
    p.S = S;
    h->total_dec = 1;
    for (int s = 0; s < S; ++s) {
        const Stage &st = h->chain[s];
        p.taps[s] = st.taps;
        p.D[s] = st.D;
        p.back[s] = st.taps ? st.taps - 1 : st.D;
        p.fwd[s] = st.taps ? 1 : 2;
        for (int k = 0; k < kDecNumHb; ++k)
            if (kDecHbLen[k] == st.taps) p.off[s] = kDecHbOff[k];
        h->total_dec *= st.D;
    }
    if (S == 0) return 0;
    p.front = h->chain[0].taps ? 0 : (S >= 2 && h->chain[1].taps ? 2 : 1);
    long long need = 0;
    for (int s = S - 1; s >= 0; --s) need = need * p.D[s] + p.back[s];
    p.hist = (int)((need + 1) & ~1LL);
    // tile size: as large as shared memory allows, but keep >= 4 tiles per SM when the call is long enough
    const int f = p.front == 2 ? 1 : 0;
    int To = 1024;
    for (;; To >>= 1) {
        long long len = To, l0 = 0, l1 = 0;
        for (int s = S - 1; s > f; --s) {
            len = (len - 1) * p.D[s] + p.back[s] + p.fwd[s];      // length of stage s-1's output range
            if (((s - 1 - f) & 1) == 0) { if (len > l0) l0 = len; } else { if (len > l1) l1 = len; }
        }
        const size_t bytes = (size_t)(l0 + l1) * sizeof(cpx);
        const long long tiles = n_hint > 0 ? (n_hint / h->total_dec + To - 1) / To : 1 << 30;
        if (To == 1 || (bytes <= 96 * 1024 && (tiles >= 4LL * pebble::num_sms() || To <= 32))) {
            if (bytes > 200 * 1024) { pebble::set_error("decimator: chain needs %zu B of shared memory", bytes); return PEBBLE_ERR_STATE; }
            p.To = To; p.buf_len[0] = (int)l0; p.buf_len[1] = (int)l1; h->smem = bytes;
            break;
        }
    }
    return 0;
}

extern "C" {

int pebble_decimator_plan(unsigned rate_in, unsigned protect_bw, unsigned rate_out, int *taps, int *decimate, int max,
                          float *out_rate, int *dec_by2_stages) {
    std::vector<Stage> c;
    const float r = dec_plan(rate_in, protect_bw, rate_out, c, dec_by2_stages);
    if (out_rate) *out_rate = r;
    for (int i = 0; i < (int)c.size() && i < max; ++i) {
        if (taps) taps[i] = c[i].taps;
        if (decimate) decimate[i] = c[i].D;
    }
    return (int)c.size();
}

pebble_decimator_t *pebble_decimator_create(unsigned sample_rate, unsigned buffer_size) {
    int ndev = 0;
    if (!pebble::cuda_ok(cudaGetDeviceCount(&ndev), "cudaGetDeviceCount") || ndev < 1) {
        pebble::set_error("decimator: no CUDA device (this library has no CPU path)");
        return nullptr;
    }
    pebble_decimator *h = new pebble_decimator();
    h->sample_rate = sample_rate;
    h->buffer_size = buffer_size;
    h->out_rate = (float)sample_rate;     // decimator.cpp:23
    memset(&h->plan, 0, sizeof(h->plan));
    if (!pebble::cuda_ok(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "cudaStreamCreate") ||
        !pebble::cuda_ok(cudaMemcpyToSymbol(c_dec_taps, pebble_dec::kDecHbTaps, sizeof(double) * 407), "cudaMemcpyToSymbol(c_dec_taps)")) {
        pebble_decimator_destroy(h);
        return nullptr;
    }
    return h;
}

void pebble_decimator_destroy(pebble_decimator_t *h) {
    if (!h) return;
    cudaFree(h->d_hist[0]); cudaFree(h->d_hist[1]);
    if (h->stream) cudaStreamDestroy(h->stream);
    h->stg.release();
    delete h;
}

float pebble_decimator_build_chain(pebble_decimator_t *h, unsigned rate_in, unsigned protect_bw, unsigned rate_out) {
    if (!h) return 0.f;
    cudaDeviceSynchronize();
    h->out_rate = dec_plan(rate_in, protect_bw, rate_out, h->chain, &h->dec_by2);
    if (dec_make_plan(h, 0) != 0) { h->chain.clear(); h->plan.S = 0; return 0.f; }
    // new HalfbandFilter objects: every delay line restarts from zero (decimator.cpp:233-263)
    cudaFree(h->d_hist[0]); cudaFree(h->d_hist[1]);
    h->d_hist[0] = h->d_hist[1] = nullptr;
    h->cur = 0;
    const size_t hb = sizeof(cpx) * (size_t)(h->plan.hist > 0 ? h->plan.hist : 2);
    if (!pebble::cuda_ok(cudaMalloc(&h->d_hist[0], hb), "cudaMalloc(hist)") ||
        !pebble::cuda_ok(cudaMalloc(&h->d_hist[1], hb), "cudaMalloc(hist)") ||
        !pebble::cuda_ok(cudaMemset(h->d_hist[0], 0, hb), "cudaMemset") ||
        !pebble::cuda_ok(cudaMemset(h->d_hist[1], 0, hb), "cudaMemset"))
        return 0.f;
    return h->out_rate;
}

int pebble_decimator_dec_by2_stages(const pebble_decimator_t *h) { return h ? h->dec_by2 : PEBBLE_ERR_ARG; }
int pebble_decimator_total_decimation(const pebble_decimator_t *h) { return h ? (int)h->total_dec : PEBBLE_ERR_ARG; }
int pebble_decimator_chain(const pebble_decimator_t *h, int *taps, int *decimate, int max) {
    if (!h) return PEBBLE_ERR_ARG;
    for (int i = 0; i < (int)h->chain.size() && i < max; ++i) {
        if (taps) taps[i] = h->chain[i].taps;
        if (decimate) decimate[i] = h->chain[i].D;
    }
    return (int)h->chain.size();
}

int pebble_decimator_out_count(const pebble_decimator_t *h, int n) {
    if (!h || n < 0) return PEBBLE_ERR_ARG;
    if (n == 0) return 0;
    if (n % h->total_dec) return PEBBLE_ERR_ARG;
    long long m = n;
    for (const Stage &st : h->chain) {
        if (st.taps && m < st.taps) return PEBBLE_ERR_ARG;
        m /= st.D;
    }
    return (int)m;
}

int pebble_decimator_process_device(pebble_decimator_t *h, int n, const double *d_in, double *d_out, void *stream_v) {
    if (!h || n < 0 || (n && (!d_in || !d_out))) { pebble::set_error("decimator: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    cudaStream_t stream = (cudaStream_t)stream_v;
    if (h->chain.empty()) {   // decimator.cpp:155-160: no chain -> copy
        PB_CUDA(cudaMemcpyAsync(d_out, d_in, sizeof(cpx) * (size_t)n, cudaMemcpyDeviceToDevice, stream), PEBBLE_ERR_CUDA);
        return n;
    }
    const int n_out = pebble_decimator_out_count(h, n);
    if (n_out > 0 && pebble::ranges_overlap(d_in, sizeof(cpx) * (size_t)n, d_out, sizeof(cpx) * (size_t)n_out)) {
        pebble::set_error("decimator: d_in and d_out overlap (in-place device calls are not supported)");
        return PEBBLE_ERR_ARG;
    }
    if (n_out < 0) {
        // The reference degrades here: a stage that sees fewer samples than taps plain-decimates and refills its delay
        // line from beyond the valid input (decimator.cpp:605-626), and a call that is not a multiple of the total
        // decimation loses the sampling phase. Neither is reproduced.
        pebble::set_error("decimator: n=%d must be a multiple of %lld and give every half-band stage >= taps samples", n, h->total_dec);
        return PEBBLE_ERR_ARG;
    }
    // re-tile for this call length (cheap; host only)
    if (dec_make_plan(h, n) != 0) return PEBBLE_ERR_STATE;
    const DecPlan &p = h->plan;
    if (h->smem > 48 * 1024)
        PB_CUDA(pebble::ensure_dynamic_smem((const void *)dec_chain_kernel, h->smem), PEBBLE_ERR_CUDA);
    const int tiles = (n_out + p.To - 1) / p.To;
    const int vec = (((uintptr_t)d_in) & 31) == 0;
    dec_chain_kernel<<<tiles, PB_DEC_THREADS, h->smem, stream>>>(p, (const cpx *)d_in, h->d_hist[h->cur], (cpx *)d_out, n_out, vec);
    dec_hist_kernel<<<(p.hist + 255) / 256, 256, 0, stream>>>(h->d_hist[h->cur], (const cpx *)d_in, n, h->d_hist[h->cur ^ 1], p.hist);
    pebble::count_launch(2);
    PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    h->cur ^= 1;
    return n_out;
}

int pebble_decimator_process(pebble_decimator_t *h, int n, const double *in, double *out) {
    if (!h || n < 0 || (n && (!in || !out))) { pebble::set_error("decimator: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    const int n_out = h->chain.empty() ? n : pebble_decimator_out_count(h, n);
    if (n_out < 0) { pebble::set_error("decimator: unsupported call length %d", n); return PEBBLE_ERR_ARG; }
    const size_t bi = (size_t)n * sizeof(cpx), bo = (size_t)n_out * sizeof(cpx);
    if (!h->stg.reserve(bi, bo)) return PEBBLE_ERR_CUDA;
    memcpy(h->stg.h_in, in, bi);
    PB_CUDA(cudaMemcpyAsync(h->stg.d_in, h->stg.h_in, bi, cudaMemcpyHostToDevice, h->stream), PEBBLE_ERR_CUDA);
    const int k = pebble_decimator_process_device(h, n, (const double *)h->stg.d_in, (double *)h->stg.d_out, h->stream);
    if (k < 0) return k;
    PB_CUDA(cudaMemcpyAsync(h->stg.h_out, h->stg.d_out, bo, cudaMemcpyDeviceToHost, h->stream), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamSynchronize(h->stream), PEBBLE_ERR_CUDA);
    memcpy(out, h->stg.h_out, bo);
    return k;
}

// state blob: `hist` CPX = the last raw input samples (oldest first); everything the stages carry derives from them
int64_t pebble_decimator_state_size(const pebble_decimator_t *h) {
    if (!h) return PEBBLE_ERR_ARG;
    return (int64_t)sizeof(cpx) * h->plan.hist;
}
int pebble_decimator_get_state(pebble_decimator_t *h, void *buf, int64_t bytes) {
    if (!h || !buf || bytes < pebble_decimator_state_size(h)) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    if (h->plan.hist) PB_CUDA(cudaMemcpy(buf, h->d_hist[h->cur], sizeof(cpx) * h->plan.hist, cudaMemcpyDeviceToHost), PEBBLE_ERR_CUDA);
    return 0;
}
int pebble_decimator_set_state(pebble_decimator_t *h, const void *buf, int64_t bytes) {
    if (!h || !buf || bytes < pebble_decimator_state_size(h)) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    if (h->plan.hist) PB_CUDA(cudaMemcpy(h->d_hist[h->cur], buf, sizeof(cpx) * h->plan.hist, cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    return 0;
}

}  // extern "C"
