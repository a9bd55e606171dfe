// downconvert.cu — CDownConvert on the GPU: kernel wrapper + C ABI (include/pebble_cuda.h).
// Kernel body: dc_fused.cuh (fused NCO mix + decimate-by-2 cascade, one segment per CTA).
#include "lib_common.h"
#include "dc_pipe.cuh"
#include "host_design.h"
#include "hb_coefs.h"
#include "chz_poly.h"
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

struct pebble_downconvert {
    DcChain chain;          // host copy
    DcChain *d_chain;       // device copy
    bool dirty;             // host copy newer than device copy
    cpx *d_state[2];
    int cur;
    uint64_t phase_hi, phase_lo;
    long long stream_pos;
    // reference members (downconvert.cpp:64-77)
    double in_rate, max_bw, nco_freq, cw_offset, out_rate;
    std::vector<int> lens;
    int threads, chunk, front, group;
    size_t smem_set = 0;
    // batched (multi-channel) form: nch NCOs sharing this object's stage chain
    int nch = 0;
    std::vector<DcChan> chans;      // host copy (phase = before sample 0 of the next call)
    std::vector<double> chan_freq;  // m_NcoFreq per channel (already negated, downconvert.cpp:105-108)
    DcChan *d_chans = nullptr;
    cpx *d_cstate[2] = {nullptr, nullptr};
    int ccur = 0;
    ChzPoly *poly = nullptr;        // polyphase-DFT front end (chz_poly.cu) when the channels sit on the grid fs / 2^S
    int batch_path = -1;            // -1 undecided, 0 per-channel kernel, 1 polyphase: fixed by the first call of a stream
    const void *fn_set = nullptr;
    // device sample format of `in` (DeviceInterfaceBase::normalizeIQ): see dc_input.cuh
    int in_fmt = PB_IQ_CPX, in_order = 0;
    double user_gain = 1.0, norm_gain = 1.0, in_scale = 1.0;
    cpx *d_conv = nullptr; size_t conv_cap = 0;   // CPX scratch of the un-fused path
    bool strict_lengths = false;                  // refuse calls the reference would not filter (downconvert.cpp:362-363)
    cudaStream_t stream;    // for the host-pointer entry point
    pebble::Staging stg;
};

// The specialised kernel when the three front stages are one of the instantiated combinations (every chain
// SetDataRate plans for an input rate >= ~300x the output bandwidth starts with one of these), else the generic one.
template <int C, int MINB> static const void *dc_pick_cic(int l1, int l2) {
    if (l1 == 11 && l2 == 11) return (const void *)dc_cic_kernel<C, 11, 11, MINB>;
    if (l1 == 3 && l2 == 11) return (const void *)dc_cic_kernel<C, 3, 11, MINB>;
    if (l1 == 3 && l2 == 3) return (const void *)dc_cic_kernel<C, 3, 3, MINB>;
    return nullptr;
}
template <int C> static const void *dc_pick_fast(int l0, int l1, int l2) {
    if (l0 == 3 && l1 == 11 && l2 == 11) return (const void *)dc_fast_kernel<C, 3, 11, 11>;
    if (l0 == 3 && l1 == 3 && l2 == 11) return (const void *)dc_fast_kernel<C, 3, 3, 11>;
    if (l0 == 3 && l1 == 3 && l2 == 3) return (const void *)dc_fast_kernel<C, 3, 3, 3>;
    if (l0 == 11 && l1 == 11 && l2 == 11) return (const void *)dc_fast_kernel<C, 11, 11, 11>;
    if (l0 == 11 && l1 == 11 && l2 == 15) return (const void *)dc_fast_kernel<C, 11, 11, 15>;
    return nullptr;
}
static const void *dc_pick_kernel(int C, int front, const DcChain &ch, bool *cic) {
    const void *fn = nullptr;
    *cic = false;
    if (ch.nstages >= 3 && front == 3 && ch.st[0].len == 3 && !getenv("PEBBLE_DC_NOCIC") && !getenv("PEBBLE_DC_GENERIC")) {
        const bool m4 = getenv("PEBBLE_DC_MINB") && atoi(getenv("PEBBLE_DC_MINB")) == 4;
        if (C == 1024) fn = m4 ? dc_pick_cic<1024, 4>(ch.st[1].len, ch.st[2].len) : dc_pick_cic<1024, 3>(ch.st[1].len, ch.st[2].len);
        if (C == 2048) fn = dc_pick_cic<2048, 2>(ch.st[1].len, ch.st[2].len);
        if (fn) { *cic = true; return fn; }
    }
    if (ch.nstages >= 3 && front == 3 && !getenv("PEBBLE_DC_GENERIC")) {
        if (C == 1024) fn = dc_pick_fast<1024>(ch.st[0].len, ch.st[1].len, ch.st[2].len);
        if (C == 2048) fn = dc_pick_fast<2048>(ch.st[0].len, ch.st[1].len, ch.st[2].len);
    }
    if (fn) return fn;
    return C == 512 ? (const void *)dc_fused_kernel<512> : C == 1024 ? (const void *)dc_fused_kernel<1024>
         : C == 2048 ? (const void *)dc_fused_kernel<2048> : (const void *)dc_fused_kernel<4096>;
}

// The fully compile-time, chunk-pipelined kernel (dc_pipe.cuh) for the 7-stage chains it is instantiated for.
// fmt != PB_IQ_CPX: only the default instantiation converts raw device samples in its load stage; nullptr otherwise (the
// caller then converts into a CPX scratch buffer first).
static const void *dc_pick_pipe(int C, const DcChain &ch, size_t *smem, int *threads, int fmt = PB_IQ_CPX) {
    if (getenv("PEBBLE_DC_NOPIPE") || getenv("PEBBLE_DC_NOCIC") || getenv("PEBBLE_DC_GENERIC")) return nullptr;
    if (ch.nstages != 7 || ch.st[0].len != 3) return nullptr;
    const int l[6] = {ch.st[1].len, ch.st[2].len, ch.st[3].len, ch.st[4].len, ch.st[5].len, ch.st[6].len};
    *threads = PB_DC_THREADS;
    if (l[0] == 11 && l[1] == 11 && l[2] == 11 && l[3] == 11 && l[4] == 15 && l[5] == 27) {   // 10 MSPS / 15 kHz: BASELINE config 2
        const int minb = getenv("PEBBLE_DC_MINB") ? atoi(getenv("PEBBLE_DC_MINB")) : 3;
        const int thr = getenv("PEBBLE_DC_THREADS") ? atoi(getenv("PEBBLE_DC_THREADS")) : 256;
        if (fmt != PB_IQ_CPX) {
            if (C != 1024) return nullptr;
            *smem = sizeof(cpx) * DcPipe<1024, 16, 256, 11, 11, 11, 11, 15, 27>::total();
            switch (fmt) {
                case PB_IQ_CS16: return (const void *)dc_pipe_kernel<1024, 16, 256, 3, 11, 11, 11, 11, 15, 27, PB_IQ_CS16>;
                case PB_IQ_CU8: return (const void *)dc_pipe_kernel<1024, 16, 256, 3, 11, 11, 11, 11, 15, 27, PB_IQ_CU8>;
                case PB_IQ_CS8: return (const void *)dc_pipe_kernel<1024, 16, 256, 3, 11, 11, 11, 11, 15, 27, PB_IQ_CS8>;
                case PB_IQ_CF32: return (const void *)dc_pipe_kernel<1024, 16, 256, 3, 11, 11, 11, 11, 15, 27, PB_IQ_CF32>;
                default: return nullptr;
            }
        }
        if (C == 2048) { *smem = sizeof(cpx) * DcPipe<2048, 32, 256, 11, 11, 11, 11, 15, 27>::total(); return (const void *)dc_pipe_kernel<2048, 32, 256, 2, 11, 11, 11, 11, 15, 27>; }
        if (C == 1024 && thr == 128) {
            *threads = 128;
            *smem = sizeof(cpx) * DcPipe<1024, 16, 128, 11, 11, 11, 11, 15, 27>::total();
            return (const void *)dc_pipe_kernel<1024, 16, 128, 3, 11, 11, 11, 11, 15, 27>;
        }
        if (C == 1024 && minb == 2) { *smem = sizeof(cpx) * DcPipe<1024, 32, 256, 11, 11, 11, 11, 15, 27>::total(); return (const void *)dc_pipe_kernel<1024, 32, 256, 2, 11, 11, 11, 11, 15, 27>; }
        if (C == 1024) { *smem = sizeof(cpx) * DcPipe<1024, 16, 256, 11, 11, 11, 11, 15, 27>::total(); return (const void *)dc_pipe_kernel<1024, 16, 256, 3, 11, 11, 11, 11, 15, 27>; }
    }
    return nullptr;
}

static int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return v ? atoi(v) : dflt;
}

static void dc_update_osc(pebble_downconvert *h) {
    // downconvert.cpp:108-110
    double inc = 6.28318530717958647692528676656 * h->nco_freq / h->in_rate;
    pebble_host::dc_set_osc(h->chain, cos(inc), sin(inc), PB_DC_THREADS, h->chunk);
    h->dirty = true;
}

extern "C" {

pebble_downconvert_t *pebble_downconvert_create(void) {
    pebble_downconvert *h = new pebble_downconvert();
    memset(&h->chain, 0, sizeof(h->chain));
    h->d_chain = nullptr;
    h->d_state[0] = h->d_state[1] = nullptr;
    h->cur = 0;
    h->phase_hi = h->phase_lo = 0;   // m_Osc1 = (1,0), downconvert.cpp:75-76
    h->stream_pos = 0;
    h->in_rate = 100000.0; h->max_bw = 10000.0; h->nco_freq = 0.0; h->cw_offset = 0.0; h->out_rate = 0.0;
    h->threads = PB_DC_THREADS;
    h->chunk = 1024;
    h->front = env_int("PEBBLE_DC_FRONT", 3);
    h->group = env_int("PEBBLE_DC_GROUP", 4);
    if (h->front < 1) h->front = 1;
    if (h->group < 1) h->group = 1;
    h->stream = nullptr;
    const size_t sbytes = sizeof(cpx) * PB_DC_MAX_STAGES * PB_DC_PREFIX;
    if (!pebble::cuda_ok(cudaMalloc(&h->d_chain, sizeof(DcChain)), "cudaMalloc(chain)") ||
        !pebble::cuda_ok(cudaMalloc(&h->d_state[0], sbytes), "cudaMalloc(state)") ||
        !pebble::cuda_ok(cudaMalloc(&h->d_state[1], sbytes), "cudaMalloc(state)") ||
        !pebble::cuda_ok(cudaMemset(h->d_state[0], 0, sbytes), "cudaMemset") ||
        !pebble::cuda_ok(cudaMemset(h->d_state[1], 0, sbytes), "cudaMemset") ||
        !pebble::cuda_ok(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "cudaStreamCreate") ||
        !pebble::cuda_ok(cudaMemcpyToSymbol(c_hb, pebble_hb::kHalfBandTaps, sizeof(double) * 341), "cudaMemcpyToSymbol(c_hb)")) {
        pebble_downconvert_destroy(h);
        return nullptr;
    }
    dc_update_osc(h);
    return h;
}

void pebble_downconvert_destroy(pebble_downconvert_t *h) {
    if (!h) return;
    if (h->d_chain) cudaFree(h->d_chain);
    if (h->d_state[0]) cudaFree(h->d_state[0]);
    if (h->d_state[1]) cudaFree(h->d_state[1]);
    if (h->stream) cudaStreamDestroy(h->stream);
    cudaFree(h->d_chans); cudaFree(h->d_cstate[0]); cudaFree(h->d_cstate[1]); cudaFree(h->d_conv);
    chz_poly_destroy(h->poly);
    h->stg.release();
    delete h;
}

static void dc_chan_set_freq(pebble_downconvert *h, int k) {
    // same rounding path as SetFrequency: inc -> cos/sin doubles -> effective angle of the recursive oscillator
    const double inc = 6.28318530717958647692528676656 * h->chan_freq[k] / h->in_rate;
    static thread_local DcChain tmp;
    pebble_host::dc_set_osc(tmp, cos(inc), sin(inc), PB_DC_THREADS, h->chunk);
    DcChan &c = h->chans[k];
    c.turn_hi = tmp.turn_hi; c.turn_lo = tmp.turn_lo; c.w1 = tmp.w1; c.wchunk = tmp.wchunk;
}

static double dc_set_rate(pebble_downconvert *h, double in_rate, double max_bw, bool simple) {
    if (!h) return 0.0;
    // downconvert.cpp:143-146 / 224-227: rebuild only when a parameter changed
    if (h->in_rate != in_rate || h->max_bw != max_bw) {
        h->in_rate = in_rate;
        h->max_bw = max_bw;
        h->out_rate = pebble_host::dc_plan(in_rate, max_bw, simple, h->lens);
        pebble_host::dc_fill_chain(h->chain, h->lens);
        // chunk length: the CIC3-fused kernel amortises its per-chunk barriers best at 2048 (measured);
        // the others at 1024. PEBBLE_DC_CHUNK overrides.
        h->chunk = (h->chain.nstages >= 3 && h->chain.st[0].len == 3) ? 2048 : 1024;
        {
            // the chunk-pipelined kernel (one barrier per chunk) runs best at 1024: 0.267 ms vs 0.291 ms per 2^26 samples
            size_t unused = 0;
            int unused_t = 0;
            if (dc_pick_pipe(1024, h->chain, &unused, &unused_t)) h->chunk = 1024;
        }
        {
            const int c = env_int("PEBBLE_DC_CHUNK", 0);
            if (c == 512 || c == 1024 || c == 2048 || c == 4096) h->chunk = c;
        }
        // DeleteFilters + new stages: the delay lines restart from zero (downconvert.cpp:150). A process call may still be in
        // flight on a caller's (non-blocking) stream: wait for it, as the reference's m_Mutex does (downconvert.cpp:148,322)
        cudaDeviceSynchronize();
        const size_t sbytes = sizeof(cpx) * PB_DC_MAX_STAGES * PB_DC_PREFIX;
        cudaMemset(h->d_state[0], 0, sbytes);
        cudaMemset(h->d_state[1], 0, sbytes);
        // SetFrequency(m_NcoFreq) (downconvert.cpp:210): NOTE it negates the stored value again
        h->nco_freq = -h->nco_freq + h->cw_offset;
        dc_update_osc(h);
        for (int k = 0; k < h->nch; ++k) dc_chan_set_freq(h, k);
        if (h->poly) chz_poly_invalidate(h->poly);
        h->batch_path = -1;
        if (h->nch) {
            const size_t cb = sizeof(cpx) * PB_DC_MAX_STAGES * PB_DC_PREFIX * (size_t)h->nch;
            cudaMemset(h->d_cstate[0], 0, cb);
            cudaMemset(h->d_cstate[1], 0, cb);
        }
    }
    return h->out_rate;
}
double pebble_downconvert_set_data_rate(pebble_downconvert_t *h, double in_rate, double max_bw) {
    return dc_set_rate(h, in_rate, max_bw, false);
}
double pebble_downconvert_set_data_rate_simple(pebble_downconvert_t *h, double in_rate, double max_bw) {
    return dc_set_rate(h, in_rate, max_bw, true);
}
void pebble_downconvert_set_frequency(pebble_downconvert_t *h, double f) {
    if (!h) return;
    h->nco_freq = -f + h->cw_offset;  // downconvert.cpp:105-108; phase is NOT reset
    dc_update_osc(h);
}
void pebble_downconvert_set_cw_offset(pebble_downconvert_t *h, double off) {
    if (h) h->cw_offset = off;  // downconvert.h:34: takes effect at the next SetFrequency
}
int pebble_downconvert_plan(double in_rate, double max_bw, int simple, int *lens, int max, double *out_rate) {
    std::vector<int> v;
    const double r = pebble_host::dc_plan(in_rate, max_bw, simple != 0, v);
    if (out_rate) *out_rate = r;
    for (int i = 0; i < (int)v.size() && i < max && lens; ++i) lens[i] = v[i];
    return (int)v.size();
}
// Host-only: the cascade SetDataRate plans for (in_rate, max_bw) as one decimate-by-2^S FIR, y[n] = sum_T h[T] x[2^S n + c - T]
// (the polyphase channelizer's prototype filter). Writes min(len, max) taps, *c_out = c, returns len.
int pebble_downconvert_composite_response(double in_rate, double max_bw, int simple, double *h, int max, int *c_out) {
    std::vector<int> lens;
    pebble_host::dc_plan(in_rate, max_bw, simple != 0, lens);
    static thread_local DcChain tmp;
    pebble_host::dc_fill_chain(tmp, lens);
    std::vector<double> hv;
    int c = 0;
    const int len = pebble_host::dc_composite(tmp, hv, c);
    if (c_out) *c_out = c;
    for (int i = 0; i < len && i < max && h; ++i) h[i] = hv[i];
    return len;
}
int pebble_downconvert_num_stages(const pebble_downconvert_t *h) { return h ? h->chain.nstages : PEBBLE_ERR_ARG; }
int pebble_downconvert_stage_lengths(const pebble_downconvert_t *h, int *lens, int max) {
    if (!h) return PEBBLE_ERR_ARG;
    for (int i = 0; i < h->chain.nstages && i < max; ++i) lens[i] = h->chain.st[i].len;
    return h->chain.nstages;
}

}  // extern "C"

// Mix only (no decimation stages): out[i] = in[i] * Osc_i. Used for a 0-stage chain
// (downconvert.cpp:323 loop never runs) and by the Mixer operator (mixer.cpp:48-81).
__global__ void __launch_bounds__(256) dc_mix_kernel(const DcChain *__restrict__ ch, const cpx *__restrict__ in,
                                                     cpx *__restrict__ out, long long n, uint64_t phase_hi,
                                                     uint64_t phase_lo, long long stream_pos) {
    // each thread: 8 consecutive-in-stride samples (i = base + tid + k*T), one exact seed
    const int T = blockDim.x;
    const long long base = (long long)blockIdx.x * T * 8;
    const long long i0 = base + threadIdx.x;
    if (i0 >= n) return;
    const cpx seed = dc_phase_to_osc(dc_phase_at(phase_hi, phase_lo, ch->turn_hi, ch->turn_lo, (uint64_t)(i0 + 1)));
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const long long i = i0 + (long long)k * T;
        if (i < n) {
            const long long g = stream_pos + i;
            const double r = g >= PB_DC_RTAB ? ch->r_inf : ch->rtab[g];
            cpx o = cpx_scale(cpx_mul(seed, ch->wtab[k]), r);
            out[i] = cpx_mul(in[i], o);
        }
    }
}

// DeviceInterfaceBase::normalizeIQ, block forms (pebblelib/deviceinterfacebase.cpp:648-878): out = code * scale with
// the reference's operation order, so the result is bit-identical for every code. 4 samples per thread.
template <int FMT>
__global__ void __launch_bounds__(256) iq_convert_kernel(const void *__restrict__ in, cpx *__restrict__ out, long long n, int order,
                                                         double scale) {
    const long long i0 = ((long long)blockIdx.x * 256 + threadIdx.x) * 4;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const long long i = i0 + k;
        if (i >= n) return;
        double I, Q;
        if (FMT == PB_IQ_CS16) { const short2 v = ((const short2 *)in)[i]; I = (double)v.x; Q = (double)v.y; }
        else if (FMT == PB_IQ_CU8) { const uchar2 v = ((const uchar2 *)in)[i]; I = (double)v.x - 128.0; Q = (double)v.y - 128.0; }
        else if (FMT == PB_IQ_CS8) { const char2 v = ((const char2 *)in)[i]; I = (double)v.x; Q = (double)v.y; }
        else if (FMT == PB_IQ_CF32) { const float2 v = ((const float2 *)in)[i]; I = (double)v.x; Q = (double)v.y; }
        else { const cpx v = ((const cpx *)in)[i]; I = v.x; Q = v.y; }
        cpx o;
        switch (order) {   // DeviceInterface::IQOrder (device_interfaces.h:140-145)
            case 0: o = cpx_make(__dmul_rn(I, scale), __dmul_rn(Q, scale)); break;
            case 1: o = cpx_make(__dmul_rn(Q, scale), __dmul_rn(I, scale)); break;
            case 2: o.x = __dmul_rn(I, scale); o.y = o.x; break;
            default: o.y = __dmul_rn(Q, scale); o.x = o.y; break;
        }
        out[i] = o;
    }
}

// scale exactly as the reference forms it (left to right): 1/128.0 * user * norm, 1/32768.0 * user * norm, user * norm
static double iq_scale(int fmt, double user_gain, double norm_gain) {
    if (fmt == PB_IQ_CU8 || fmt == PB_IQ_CS8) return 1 / 128.0 * user_gain * norm_gain;
    if (fmt == PB_IQ_CS16) return 1 / 32768.0 * user_gain * norm_gain;
    return user_gain * norm_gain;
}

static int iq_convert_launch(int fmt, int order, double scale, long long n, const void *d_in, cpx *d_out, cudaStream_t st) {
    const unsigned blocks = (unsigned)((n + 1023) / 1024);
    switch (fmt) {
        case PB_IQ_CPX: iq_convert_kernel<PB_IQ_CPX><<<blocks, 256, 0, st>>>(d_in, d_out, n, order, scale); break;
        case PB_IQ_CS16: iq_convert_kernel<PB_IQ_CS16><<<blocks, 256, 0, st>>>(d_in, d_out, n, order, scale); break;
        case PB_IQ_CU8: iq_convert_kernel<PB_IQ_CU8><<<blocks, 256, 0, st>>>(d_in, d_out, n, order, scale); break;
        case PB_IQ_CS8: iq_convert_kernel<PB_IQ_CS8><<<blocks, 256, 0, st>>>(d_in, d_out, n, order, scale); break;
        case PB_IQ_CF32: iq_convert_kernel<PB_IQ_CF32><<<blocks, 256, 0, st>>>(d_in, d_out, n, order, scale); break;
        default: pebble::set_error("normalize_iq: unknown format %d", fmt); return PEBBLE_ERR_ARG;
    }
    pebble::count_launch();
    PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    return 0;
}

extern "C" {

int pebble_iq_bytes_per_sample(int fmt) { return (fmt < 0 || fmt > PB_IQ_CF32) ? PEBBLE_ERR_ARG : iq_bytes_per_sample(fmt); }

int pebble_normalize_iq_device(int fmt, int order, double user_gain, double norm_gain, int reverse, long long n, const void *d_in,
                               double *d_out, void *stream) {
    if (fmt < 0 || fmt > PB_IQ_CF32 || order < 0 || order > 3 || n < 0 || (n && (!d_in || !d_out))) {
        pebble::set_error("normalize_iq: bad argument");
        return PEBBLE_ERR_ARG;
    }
    if (n == 0) return 0;
    if (reverse && order < 2) order ^= 1;   // deviceinterfacebase.cpp:653-657
    return iq_convert_launch(fmt, order, iq_scale(fmt, user_gain, norm_gain), n, d_in, (cpx *)d_out, (cudaStream_t)stream);
}

int pebble_normalize_iq(int fmt, int order, double user_gain, double norm_gain, int reverse, long long n, const void *in, double *out) {
    if (fmt < 0 || fmt > PB_IQ_CF32 || n < 0 || (n && (!in || !out))) { pebble::set_error("normalize_iq: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    void *d_in = nullptr; double *d_out = nullptr;
    const size_t bi = (size_t)n * iq_bytes_per_sample(fmt), bo = (size_t)n * sizeof(cpx);
    PB_CUDA(cudaMalloc(&d_in, bi), PEBBLE_ERR_CUDA);
    if (!pebble::cuda_ok(cudaMalloc(&d_out, bo), "cudaMalloc")) { cudaFree(d_in); return PEBBLE_ERR_CUDA; }
    int r = PEBBLE_ERR_CUDA;
    if (pebble::cuda_ok(cudaMemcpy(d_in, in, bi, cudaMemcpyHostToDevice), "cudaMemcpy")) {
        r = pebble_normalize_iq_device(fmt, order, user_gain, norm_gain, reverse, n, d_in, d_out, nullptr);
        if (r == 0 && !pebble::cuda_ok(cudaMemcpy(out, d_out, bo, cudaMemcpyDeviceToHost), "cudaMemcpy")) r = PEBBLE_ERR_CUDA;
    }
    cudaFree(d_in); cudaFree(d_out);
    return r;
}

int pebble_downconvert_set_input_format(pebble_downconvert_t *h, int fmt, int order, double user_gain, double norm_gain) {
    if (!h || fmt < 0 || fmt > PB_IQ_CF32 || order < 0 || order > 3) { pebble::set_error("downconvert: bad input format"); return PEBBLE_ERR_ARG; }
    h->in_fmt = fmt; h->in_order = order; h->user_gain = user_gain; h->norm_gain = norm_gain;
    h->in_scale = iq_scale(fmt, user_gain, norm_gain);
    return 0;
}
int pebble_downconvert_set_strict_lengths(pebble_downconvert_t *h, int on) {
    if (!h) return PEBBLE_ERR_ARG;
    h->strict_lengths = on != 0;
    return 0;
}
int pebble_downconvert_input_bytes_per_sample(const pebble_downconvert_t *h) { return h ? iq_bytes_per_sample(h->in_fmt) : PEBBLE_ERR_ARG; }

int pebble_downconvert_process_device(pebble_downconvert_t *h, int n, const void *d_in_v, double *d_out,
                                      void *stream_v) {
    if (!h || n < 0 || (n && (!d_in_v || !d_out))) { pebble::set_error("downconvert: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    cudaStream_t stream = (cudaStream_t)stream_v;
    const int S = h->chain.nstages;
    // the reference decimates in place (downconvert.cpp:309-325); a grid cannot: segments are read while others are written
    if (pebble::ranges_overlap(d_in_v, (size_t)pebble_downconvert_input_bytes_per_sample(h) * (size_t)n, d_out,
                               sizeof(cpx) * (size_t)((n >> S) > 0 ? (n >> S) : 1))) {
        pebble::set_error("downconvert: d_in and d_out overlap (in-place device calls are not supported)");
        return PEBBLE_ERR_ARG;
    }
    if (h->strict_lengths) {
        // the reference's per-stage preconditions: a generic half-band filters only for InLength >= taps (downconvert.cpp:
        // 362-363) and, run in place as CDownConvert runs it (:325), keeps a correct history only for InLength >= 2*(taps-1)
        // (:387-390 re-read the caller's buffer after the outputs were written); the 11-tap one reads pInData[0..17] (:426-480)
        for (int s = 0; s < S; ++s) {
            const int need = h->chain.st[s].len == 11 ? 18 : h->chain.st[s].len == 3 ? 2 : 2 * (h->chain.st[s].len - 1);
            if ((n >> s) < need) {
                pebble::set_error("downconvert: stage %d (%d taps) would receive %d samples (strict lengths)", s, h->chain.st[s].len, n >> s);
                return PEBBLE_ERR_ARG;
            }
        }
    }
    // Raw device samples: fused into the load stage where the chain has such a kernel (dc_pipe_kernel<..., FMT>, IQ order);
    // otherwise normalizeIQ runs as its own kernel into a CPX scratch buffer and the chain reads that.
    const double *d_in = (const double *)d_in_v;
    int fused_fmt = PB_IQ_CPX;
    double fused_scale = 1.0;
    if (h->in_fmt != PB_IQ_CPX || h->in_order != 0 || h->in_scale != 1.0) {
        size_t sm = 0; int th = 0;
        const bool can_fuse = h->in_order == 0 && S > 0 && !(n & ((1 << S) - 1)) &&
                              (h->in_fmt == PB_IQ_CPX ? dc_pick_pipe(h->chunk, h->chain, &sm, &th) != nullptr
                                                      : dc_pick_pipe(h->chunk, h->chain, &sm, &th, h->in_fmt) != nullptr);
        if (can_fuse) {
            fused_fmt = h->in_fmt;
            fused_scale = h->in_scale;
        } else {
            if ((size_t)n > h->conv_cap) {
                PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
                cudaFree(h->d_conv); h->d_conv = nullptr; h->conv_cap = 0;
                PB_CUDA(cudaMalloc(&h->d_conv, sizeof(cpx) * (size_t)n), PEBBLE_ERR_CUDA);
                h->conv_cap = (size_t)n;
            }
            const int r = iq_convert_launch(h->in_fmt, h->in_order, h->in_scale, n, d_in_v, h->d_conv, stream);
            if (r < 0) return r;
            d_in = (const double *)h->d_conv;
        }
    }
    if (n & ((1 << S) - 1)) {
        pebble::set_error("downconvert: n=%d is not a multiple of 2^%d", n, S);
        return PEBBLE_ERR_ARG;
    }
    if (h->dirty) {
        PB_CUDA(cudaMemcpyAsync(h->d_chain, &h->chain, sizeof(DcChain), cudaMemcpyHostToDevice, stream), PEBBLE_ERR_CUDA);
        h->dirty = false;
    }
    if (S == 0) {
        if (h->chain.wtab_stride != 256) { pebble::set_error("downconvert: mix kernel needs 256-thread table"); return PEBBLE_ERR_STATE; }
        const int blocks = (int)((n + 2047) / 2048);
        dc_mix_kernel<<<blocks, 256, 0, stream>>>(h->d_chain, (const cpx *)d_in, (cpx *)d_out, n, h->phase_hi,
                                                  h->phase_lo, h->stream_pos);
        pebble::count_launch();
    } else {
        const int C = h->chunk;
        if (S > 12 || (C >> S) < 1) {
            pebble::set_error("downconvert: unsupported chain (S=%d, chunk=%d)", S, C);
            return PEBBLE_ERR_STATE;
        }
        DcCall call;
        call.in = (const cpx *)d_in;
        call.out = (cpx *)d_out;
        call.state_in = h->d_state[h->cur];
        call.state_out = h->d_state[h->cur ^ 1];
        call.n = n;
        call.phase_hi = h->phase_hi;
        call.phase_lo = h->phase_lo;
        call.stream_pos = h->stream_pos;
        call.front = h->front < S ? h->front : S;
        call.group = h->group;
        call.chan = nullptr;
        call.nch = 1;
        call.state_stride = 0;
        call.out_stride = 0;
        call.w1 = h->chain.w1; call.wchunk = h->chain.wchunk; call.r_inf = h->chain.r_inf * fused_scale;
        call.turn_hi = h->chain.turn_hi; call.turn_lo = h->chain.turn_lo;
        bool cic = false;
        const void *fn = dc_pick_kernel(C, call.front, h->chain, &cic);
        // the CIC3-fused kernel keeps the full-rate stage in registers: no stage-0 arrays
        size_t smem = (size_t)(dc_smem_cpx(h->chain, C, call.front, call.group) -
                               (cic ? 2 * dc_stage_arr(h->chain.st[0], C) : 0)) * sizeof(cpx);
        size_t smem_pipe = 0;
        int threads = PB_DC_THREADS;
        if (const void *pf = dc_pick_pipe(C, h->chain, &smem_pipe, &threads, fused_fmt)) { fn = pf; smem = smem_pipe; }
        if (smem > h->smem_set || fn != h->fn_set) {
            PB_CUDA(pebble::ensure_dynamic_smem(fn, smem), PEBBLE_ERR_CUDA);
            h->smem_set = smem;
            h->fn_set = fn;
        }
        int occ = 1;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, threads, smem);
        if (occ < 1) occ = 1;
        // segments: one wave of resident CTAs (PEBBLE_DC_WAVES to change), never shorter than 8x the pre-roll or
        // 16 chunks, and a whole number of chunks
        long long min_seg = (long long)h->chain.preroll * 8;
        if (min_seg < 16LL * C) min_seg = 16LL * C;
        long long target = (long long)pebble::num_sms() * occ * env_int("PEBBLE_DC_WAVES", 1);
        long long seg = (n + target - 1) / target;
        if (seg < min_seg) seg = min_seg;
        seg = ((seg + C - 1) / C) * C;
        call.seg_len = seg;
        const int nseg = (int)((n + seg - 1) / seg);
        void *args[2] = {(void *)&h->d_chain, (void *)&call};
        PB_CUDA(cudaLaunchKernel(fn, dim3(nseg), dim3(threads), args, smem, stream), PEBBLE_ERR_CUDA);
        pebble::count_launch();
    }
    PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    h->cur ^= (S > 0);
    pebble_host::phase_advance(h->phase_hi, h->phase_lo, h->chain.turn_hi, h->chain.turn_lo, (uint64_t)n);
    h->stream_pos += n;
    return n >> S;
}

int pebble_downconvert_process(pebble_downconvert_t *h, int n, const void *in, double *out) {
    if (!h || n < 0 || (n && (!in || !out))) { pebble::set_error("downconvert: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    const int S = h->chain.nstages;
    const size_t bi = (size_t)n * iq_bytes_per_sample(h->in_fmt), bo = (size_t)(n >> S) * sizeof(cpx);
    if (!h->stg.reserve(bi, bo)) return PEBBLE_ERR_CUDA;
    memcpy(h->stg.h_in, in, bi);
    PB_CUDA(cudaMemcpyAsync(h->stg.d_in, h->stg.h_in, bi, cudaMemcpyHostToDevice, h->stream), PEBBLE_ERR_CUDA);
    int k = pebble_downconvert_process_device(h, n, (const double *)h->stg.d_in, (double *)h->stg.d_out, h->stream);
    if (k < 0) return k;
    PB_CUDA(cudaMemcpyAsync(h->stg.h_out, h->stg.d_out, bo, cudaMemcpyDeviceToHost, h->stream), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamSynchronize(h->stream), PEBBLE_ERR_CUDA);
    memcpy(out, h->stg.h_out, bo);
    return k;
}

// state blob: uint64 phase_hi, phase_lo, int64 stream_pos, then nstages * PB_DC_PREFIX cpx
int64_t pebble_downconvert_state_size(const pebble_downconvert_t *h) {
    if (!h) return PEBBLE_ERR_ARG;
    return 24 + (int64_t)sizeof(cpx) * PB_DC_PREFIX * h->chain.nstages;
}
int pebble_downconvert_get_state(pebble_downconvert_t *h, void *buf, int64_t bytes) {
    if (!h || !buf || bytes < pebble_downconvert_state_size(h)) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    uint64_t hdr[3] = {h->phase_hi, h->phase_lo, (uint64_t)h->stream_pos};
    memcpy(buf, hdr, 24);
    PB_CUDA(cudaMemcpy((char *)buf + 24, h->d_state[h->cur], sizeof(cpx) * PB_DC_PREFIX * h->chain.nstages,
                       cudaMemcpyDeviceToHost), PEBBLE_ERR_CUDA);
    return 0;
}
int pebble_downconvert_set_state(pebble_downconvert_t *h, const void *buf, int64_t bytes) {
    if (!h || !buf || bytes < pebble_downconvert_state_size(h)) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    uint64_t hdr[3];
    memcpy(hdr, buf, 24);
    h->phase_hi = hdr[0]; h->phase_lo = hdr[1]; h->stream_pos = (long long)hdr[2];
    PB_CUDA(cudaMemcpy(h->d_state[h->cur], (const char *)buf + 24, sizeof(cpx) * PB_DC_PREFIX * h->chain.nstages,
                       cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    return 0;
}


// ---- batched form: nch channels = nch CDownConvert objects with the same (InRate, MaxBW) and their own NCO ---------
int pebble_downconvert_set_channels(pebble_downconvert_t *h, int nch) {
    if (!h || nch < 1) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    cudaFree(h->d_chans); cudaFree(h->d_cstate[0]); cudaFree(h->d_cstate[1]);
    h->d_chans = nullptr; h->d_cstate[0] = h->d_cstate[1] = nullptr; h->nch = 0;
    const size_t cb = sizeof(cpx) * PB_DC_MAX_STAGES * PB_DC_PREFIX * (size_t)nch;
    PB_CUDA(cudaMalloc(&h->d_chans, sizeof(DcChan) * (size_t)nch), PEBBLE_ERR_CUDA);
    for (int b = 0; b < 2; ++b) {
        PB_CUDA(cudaMalloc(&h->d_cstate[b], cb), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMemset(h->d_cstate[b], 0, cb), PEBBLE_ERR_CUDA);
    }
    h->nch = nch; h->ccur = 0;
    if (!h->poly) h->poly = chz_poly_create();
    chz_poly_invalidate(h->poly);
    h->batch_path = -1;
    h->chans.assign(nch, DcChan());
    h->chan_freq.assign(nch, 0.0);
    for (int k = 0; k < nch; ++k) {
        h->chans[k].phase_hi = h->chans[k].phase_lo = 0;   // m_Osc1 = (1,0)
        dc_chan_set_freq(h, k);
    }
    h->stream_pos = 0;
    return 0;
}
int pebble_downconvert_batch_path(const pebble_downconvert_t *h) { return h ? h->batch_path : PEBBLE_ERR_ARG; }
int pebble_downconvert_set_channel_frequency(pebble_downconvert_t *h, int ch, double f) {
    if (!h || ch < 0 || ch >= h->nch) return PEBBLE_ERR_ARG;
    h->chan_freq[ch] = -f + h->cw_offset;   // downconvert.cpp:105-108; the phase is NOT reset
    dc_chan_set_freq(h, ch);
    return 0;
}
int pebble_downconvert_process_channels_device(pebble_downconvert_t *h, int n, const void *d_in_v, double *d_out,
                                               long long out_stride, void *stream_v) {
    if (!h || n < 0 || (n && (!d_in_v || !d_out))) { pebble::set_error("downconvert: bad argument"); return PEBBLE_ERR_ARG; }
    if (h->nch < 1) { pebble::set_error("downconvert: set_channels not called"); return PEBBLE_ERR_STATE; }
    if (n == 0) return 0;
    cudaStream_t stream = (cudaStream_t)stream_v;
    const int S = h->chain.nstages, C = h->chunk;
    if (n & ((1 << S) - 1)) { pebble::set_error("downconvert: n=%d is not a multiple of 2^%d", n, S); return PEBBLE_ERR_ARG; }
    if (out_stride < (n >> S)) { pebble::set_error("downconvert: out_stride smaller than the output count"); return PEBBLE_ERR_ARG; }
    // ---- uniform channel grid: ONE mix + polyphase FIR + D-point DFT for all channels (chz_poly.cu) ----------------------
    if (h->in_order == 0 && h->batch_path != 0) {
        const int ok = chz_poly_prepare(h->poly, h->chain, h->chans);
        if (ok < 0) return ok;
        if (ok == 0 && h->batch_path == 1) {
            pebble::set_error("downconvert: a channel was retuned off the fs/2^S grid mid-stream; the polyphase front end cannot "
                              "follow (create the channelizer with PEBBLE_CHZ_DIRECT=1 for arbitrary retunes)");
            return PEBBLE_ERR_STATE;
        }
        h->batch_path = ok;
        if (ok == 1) {
            if (h->dirty) {
                PB_CUDA(cudaMemcpyAsync(h->d_chain, &h->chain, sizeof(DcChain), cudaMemcpyHostToDevice, stream), PEBBLE_ERR_CUDA);
                h->dirty = false;
            }
            const int r = chz_poly_process(h->poly, h->chain, h->d_chain, h->chans, h->stream_pos, h->in_fmt, h->in_scale, n, d_in_v,
                                           (cpx *)d_out, out_stride, stream);
            if (r < 0) return r;
            for (int k = 0; k < h->nch; ++k)
                pebble_host::phase_advance(h->chans[k].phase_hi, h->chans[k].phase_lo, h->chans[k].turn_hi, h->chans[k].turn_lo, (uint64_t)n);
            h->stream_pos += n;
            return n >> S;
        }
    }
    h->batch_path = 0;
    // ---- per-channel kernel (dc_cic_kernel, batched): raw formats / IQ orders are converted into a CPX scratch first --------
    const double *d_in = (const double *)d_in_v;
    if (h->in_fmt != PB_IQ_CPX || h->in_order != 0 || h->in_scale != 1.0) {
        if ((size_t)n > h->conv_cap) {
            PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
            cudaFree(h->d_conv); h->d_conv = nullptr; h->conv_cap = 0;
            PB_CUDA(cudaMalloc(&h->d_conv, sizeof(cpx) * (size_t)n), PEBBLE_ERR_CUDA);
            h->conv_cap = (size_t)n;
        }
        const int rc = iq_convert_launch(h->in_fmt, h->in_order, h->in_scale, n, d_in_v, h->d_conv, stream);
        if (rc < 0) return rc;
        d_in = (const double *)h->d_conv;
    }
    bool cic = false;
    const int front = h->front < S ? h->front : S;
    const void *fn = dc_pick_kernel(C, front, h->chain, &cic);
    if (!cic) {
        pebble::set_error("downconvert: the batched form needs a chain that starts with CIC3 (input rate >= ~667x max_bw)");
        return PEBBLE_ERR_STATE;
    }
    if (h->dirty) {
        PB_CUDA(cudaMemcpyAsync(h->d_chain, &h->chain, sizeof(DcChain), cudaMemcpyHostToDevice, stream), PEBBLE_ERR_CUDA);
        h->dirty = false;
    }
    PB_CUDA(cudaMemcpyAsync(h->d_chans, h->chans.data(), sizeof(DcChan) * (size_t)h->nch, cudaMemcpyHostToDevice, stream),
            PEBBLE_ERR_CUDA);
    DcCall call;
    call.in = (const cpx *)d_in;
    call.out = (cpx *)d_out;
    call.state_in = h->d_cstate[h->ccur];
    call.state_out = h->d_cstate[h->ccur ^ 1];
    call.n = n;
    call.phase_hi = call.phase_lo = 0;
    call.stream_pos = h->stream_pos;
    call.front = front;
    call.group = h->group;
    call.chan = h->d_chans;
    call.nch = h->nch;
    call.state_stride = PB_DC_MAX_STAGES * PB_DC_PREFIX;
    call.out_stride = out_stride;
    call.w1 = call.wchunk = cpx_make(0.0, 0.0); call.r_inf = 0.0; call.turn_hi = call.turn_lo = 0;   // per channel: DcChan
    const size_t smem = (size_t)(dc_smem_cpx(h->chain, C, front, call.group) - 2 * dc_stage_arr(h->chain.st[0], C)) * sizeof(cpx);
    if (smem > h->smem_set || fn != h->fn_set) {
        PB_CUDA(pebble::ensure_dynamic_smem(fn, smem), PEBBLE_ERR_CUDA);
        h->smem_set = smem;
        h->fn_set = fn;
    }
    int occ = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, PB_DC_THREADS, smem);
    if (occ < 1) occ = 1;
    // segments: at least 8x the pre-roll; enough (segment, channel) CTAs for ~4 waves when there are few channels
    long long min_seg = (long long)h->chain.preroll * 8;
    if (min_seg < 16LL * C) min_seg = 16LL * C;
    long long want = ((long long)pebble::num_sms() * occ * 4 + h->nch - 1) / h->nch;
    if (want < 1) want = 1;
    long long seg = (n + want - 1) / want;
    if (seg < min_seg) seg = min_seg;
    seg = ((seg + C - 1) / C) * C;
    call.seg_len = seg;
    const long long nseg = (n + seg - 1) / seg;
    void *args[2] = {(void *)&h->d_chain, (void *)&call};
    PB_CUDA(cudaLaunchKernel(fn, dim3((unsigned)(nseg * h->nch)), dim3(PB_DC_THREADS), args, smem, stream), PEBBLE_ERR_CUDA);
    pebble::count_launch();
    h->ccur ^= 1;
    for (int k = 0; k < h->nch; ++k)
        pebble_host::phase_advance(h->chans[k].phase_hi, h->chans[k].phase_lo, h->chans[k].turn_hi, h->chans[k].turn_lo, (uint64_t)n);
    h->stream_pos += n;
    return n >> S;
}

}  // extern "C"

// ---- Mixer (pebblelib/mixer.cpp:5-81): the same recursive oscillator as CDownConvert with no decimation stages ----
struct pebble_mixer {
    pebble_downconvert *dc;
    double frequency;      // m_frequency = -f (mixer.cpp:31)
    unsigned sample_rate, num_samples;
};

extern "C" {

pebble_mixer_t *pebble_mixer_create(unsigned sample_rate, unsigned num_samples) {
    if (sample_rate == 0) return nullptr;
    pebble_downconvert *dc = pebble_downconvert_create();
    if (!dc) return nullptr;
    pebble_mixer *m = new pebble_mixer();
    m->dc = dc;
    m->sample_rate = sample_rate;
    m->num_samples = num_samples;
    dc->in_rate = (double)sample_rate;
    dc->max_bw = 0.0;
    dc->lens.clear();
    pebble_host::dc_fill_chain(dc->chain, dc->lens);   // zero stages: process = dc_mix_kernel
    pebble_mixer_set_frequency(m, 0.0);                // mixer.cpp:13
    return m;
}

void pebble_mixer_destroy(pebble_mixer_t *m) {
    if (!m) return;
    pebble_downconvert_destroy(m->dc);
    delete m;
}

void pebble_mixer_set_frequency(pebble_mixer_t *m, double f) {
    if (!m) return;
    m->frequency = -f;
    m->dc->nco_freq = -f;
    dc_update_osc(m->dc);
    // m_lastOsc = (1, 0): unlike CDownConvert::SetFrequency, the phase AND the amplitude transient restart (mixer.cpp:36-37)
    m->dc->phase_hi = m->dc->phase_lo = 0;
    m->dc->stream_pos = 0;
}

int pebble_mixer_is_passthrough(const pebble_mixer_t *m) { return m ? (m->frequency == 0 ? 1 : 0) : PEBBLE_ERR_ARG; }

int pebble_mixer_process_device(pebble_mixer_t *m, int n, const double *d_in, double *d_out, void *stream) {
    if (!m || n < 0 || (n && (!d_in || !d_out))) { pebble::set_error("mixer: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    if (m->frequency == 0) {   // mixer.cpp:51-53 returns `in` itself; in the C ABI the samples are copied, the state is untouched
        if (d_in != d_out)
            PB_CUDA(cudaMemcpyAsync(d_out, d_in, sizeof(cpx) * (size_t)n, cudaMemcpyDeviceToDevice, (cudaStream_t)stream), PEBBLE_ERR_CUDA);
        return n;
    }
    return pebble_downconvert_process_device(m->dc, n, d_in, d_out, stream);
}

int pebble_mixer_process(pebble_mixer_t *m, int n, const double *in, double *out) {
    if (!m || n < 0 || (n && (!in || !out))) { pebble::set_error("mixer: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    if (m->frequency == 0) {
        if (in != out) memcpy(out, in, sizeof(cpx) * (size_t)n);
        return n;
    }
    return pebble_downconvert_process(m->dc, n, in, out);
}

}  // extern "C"

// ---- NCO (pebblelib/nco.h:20-60, nco.cpp:5-81): the oscillator primitive with its Nyquist safety net ----------------------
struct pebble_nco {
    pebble_downconvert *dc;
    double frequency;          // m_frequency after the clamp
    unsigned sample_rate, buffer_size;
    cpx *d_buf = nullptr; size_t cap = 0;
};

// out[i] = (mix ? in[i] : 0) + gain * Osc_i, Osc_i = NCO::nextSample (nco.h:29-38) = r_i exp(j (i+1) theta)
__global__ void __launch_bounds__(256) nco_gen_kernel(const DcChain *__restrict__ ch, const cpx *__restrict__ in, cpx *__restrict__ out,
                                                      long long n, uint64_t phase_hi, uint64_t phase_lo, long long stream_pos, double gain, int mix) {
    const long long i0 = (long long)blockIdx.x * 256 * 8 + threadIdx.x;
    if (i0 >= n) return;
    const cpx seed = dc_phase_to_osc(dc_phase_at(phase_hi, phase_lo, ch->turn_hi, ch->turn_lo, (uint64_t)(i0 + 1)));
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const long long i = i0 + (long long)k * 256;
        if (i < n) {
            const long long g = stream_pos + i;
            const double r = g >= PB_DC_RTAB ? ch->r_inf : ch->rtab[g];
            const cpx o = cpx_scale(cpx_mul(seed, ch->wtab[k]), r * gain);
            out[i] = mix ? cpx_add(in[i], o) : o;
        }
    }
}

extern "C" {
pebble_nco_t *pebble_nco_create(unsigned sample_rate, unsigned buffer_size) {
    if (sample_rate == 0) return nullptr;
    pebble_downconvert *dc = pebble_downconvert_create();
    if (!dc) return nullptr;
    pebble_nco *h = new pebble_nco();
    h->dc = dc; h->sample_rate = sample_rate; h->buffer_size = buffer_size;
    dc->in_rate = (double)sample_rate;
    dc->lens.clear();
    pebble_host::dc_fill_chain(dc->chain, dc->lens);
    pebble_nco_set_frequency(h, 0.0);
    return h;
}
void pebble_nco_destroy(pebble_nco_t *h) {
    if (!h) return;
    pebble_downconvert_destroy(h->dc);
    cudaFree(h->d_buf);
    delete h;
}
// NCO::setFrequency (nco.cpp:16-45): |f| is limited to (sampleRate / 2) - 1 (integer arithmetic) and the oscillator restarts
double pebble_nco_set_frequency(pebble_nco_t *h, double f) {
    if (!h) return 0.0;
    const int nyquist = (int)(h->sample_rate / 2) - 1;
    if (f < -nyquist) f = -nyquist;
    else if (f > nyquist) f = nyquist;
    h->frequency = f;
    h->dc->nco_freq = f;               // m_oscInc = TWOPI * m_frequency / m_sampleRate: not negated (contrast Mixer, mixer.cpp:31)
    dc_update_osc(h->dc);
    h->dc->phase_hi = h->dc->phase_lo = 0;
    h->dc->stream_pos = 0;
    return f;
}
double pebble_nco_frequency(const pebble_nco_t *h) { return h ? h->frequency : 0.0; }
// NCO::genSingle (nco.cpp:65-81): n consecutive nextSample() values times `gain`, written to (mix == 0) or added to (mix != 0) io
int pebble_nco_gen_single_device(pebble_nco_t *h, double *d_io, int n, double gain, int mix, void *stream_v) {
    if (!h || n < 0 || (n && !d_io)) { pebble::set_error("nco: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream_v;
    pebble_downconvert *dc = h->dc;
    if (dc->dirty) {
        PB_CUDA(cudaMemcpyAsync(dc->d_chain, &dc->chain, sizeof(DcChain), cudaMemcpyHostToDevice, st), PEBBLE_ERR_CUDA);
        dc->dirty = false;
    }
    nco_gen_kernel<<<(n + 2047) / 2048, 256, 0, st>>>(dc->d_chain, (const cpx *)d_io, (cpx *)d_io, n, dc->phase_hi, dc->phase_lo, dc->stream_pos,
                                                     gain, mix);
    pebble::count_launch();
    PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    pebble_host::phase_advance(dc->phase_hi, dc->phase_lo, dc->chain.turn_hi, dc->chain.turn_lo, (uint64_t)n);
    dc->stream_pos += n;
    return n;
}
int pebble_nco_gen_single(pebble_nco_t *h, double *io, int n, double gain, int mix) {
    if (!h || n < 0 || (n && !io)) { pebble::set_error("nco: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    if ((size_t)n > h->cap) {
        cudaFree(h->d_buf); h->d_buf = nullptr; h->cap = 0;
        PB_CUDA(cudaMalloc(&h->d_buf, sizeof(cpx) * (size_t)n), PEBBLE_ERR_CUDA);
        h->cap = (size_t)n;
    }
    if (mix) PB_CUDA(cudaMemcpyAsync(h->d_buf, io, sizeof(cpx) * (size_t)n, cudaMemcpyHostToDevice, h->dc->stream), PEBBLE_ERR_CUDA);
    const int r = pebble_nco_gen_single_device(h, (double *)h->d_buf, n, gain, mix, h->dc->stream);
    if (r < 0) return r;
    PB_CUDA(cudaMemcpyAsync(io, h->d_buf, sizeof(cpx) * (size_t)n, cudaMemcpyDeviceToHost, h->dc->stream), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamSynchronize(h->dc->stream), PEBBLE_ERR_CUDA);
    return r;
}
}  // extern "C"
