// channelizer.cu — the batched multi-channel receiver (BASELINE config 4): nch independent receiver chains
//   CDownConvert (own NCO each, shared stage plan) -> gain restore -> CFastFIR -> AGC_OFF -> Demod (AM / NBFM PLL / SSB)
// over ONE wideband capture, i.e. nch copies of the non-WFM branch of Receiver::processIQData
// (application/receiver.cpp:903-996) that differ only in mixer frequency and demod mode.
// Layout: channel-major. Decimated, filtered and demodulated streams are [channel][sample] rows, so a rank that owns
// channels [c0, c1) owns a contiguous slab (multi-GPU sharding by channel) and per-channel carried state is
// one row of each state array.
//   1. dc_cic_kernel, batched: blockIdx.x = segment*nch + channel (downconvert.cu). Every channel mixes the full-rate
//      stream (the reference algorithm, kept for parity): FP64-bound, the input is shared through L2.
//   2. fastfir_kernel, blockIdx.y = channel, the two scalar gains fused (fastfir.cu).
//   3. chan_demod_kernel: one thread per channel (the PLL and the DC-block IIR are serial recurrences, finding F10).
//   4. chan_cfir_kernel: the demod post filter, blockIdx.y = channel, per-mode tap set.
#include "lib_common.h"
#include "demod_kernels.cuh"
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

enum { CH_AM = 0, CH_FMN = 2 };   // DeviceInterface::DemodMode (device_interfaces.h:124-137); >= 5: pass-through

struct ChanState {      // per-channel demod state
    PllState pll;       // Demod_NFM (float members, demod_nfm.cpp:44-66)
    float pad;
    double am_dc_last;  // Demod_AM::m_amDcLast (demod_am.cpp:52-54)
};

// sqrt(x), x >= 0, without the library routine's branches: float rsqrt seed (2^-22), two Newton steps on 1/sqrt (2^-85), one
// correction of x * r: within 1 ulp of the correctly rounded root. Outside the float range (never for IQ envelopes) and for
// 0 it falls back to sqrt().
__device__ __forceinline__ double pb_sqrt_newton(double x) {
    if (!(x > 1e-36 && x < 1e36)) return sqrt(x);
    double r = (double)rsqrtf((float)x);
    const double hx = 0.5 * x;
    r = fma(r, fma(-hx, r * r, 0.5), r);
    r = fma(r, fma(-hx, r * r, 0.5), r);
    double s = x * r;
    s = fma(fma(-s, s, x), 0.5 * r, s);
    return s;
}

// one thread per channel: in/out rows of `stride` cpx
static __global__ void chan_demod_kernel(const cpx *__restrict__ in, cpx *__restrict__ out, int n, long long stride, int nch,
                                         const int *__restrict__ modes, const int *__restrict__ order, int slots,
                                         PllConst pc, double am_alpha, ChanState *__restrict__ st, int frame) {
    // order[]: channels grouped by demodulator and padded to whole warps (-1), so a warp runs ONE of the recurrences
    // below; with channels in natural order (NBFM and AM alternating in config 4) every warp ran both, one after the other.
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= slots) return;
    const int c = order[slot];
    if (c < 0 || c >= nch) return;
    const cpx *x = in + (long long)c * stride;
    cpx *y = out + (long long)c * stride;
    const int mode = modes[c];
    if (mode == CH_FMN) {                 // Demod_NFM::processBlockNCO, demod_nfm.cpp:225-257
        PllState v = st[c].pll;
        pll_run(x, y, n, pc, v, frame);
        st[c].pll = v;
    } else if (mode == CH_AM) {           // Demod_AM::processBlockFiltered, demod_am.cpp:40-59
        double d = st[c].am_dc_last;
        cpx cur[PB_SERIAL_BLK], nxt[PB_SERIAL_BLK];               // register double buffer, as in pll_run
#pragma unroll
        for (int j = 0; j < PB_SERIAL_BLK; ++j) cur[j] = j < n ? x[j] : cpx_make(0.0, 0.0);
        for (int i0 = 0; i0 < n; i0 += PB_SERIAL_BLK) {
#pragma unroll
            for (int j = 0; j < PB_SERIAL_BLK; ++j) nxt[j] = i0 + PB_SERIAL_BLK + j < n ? x[i0 + PB_SERIAL_BLK + j] : cpx_make(0.0, 0.0);
            // the envelopes of a block do not depend on the recurrence: formed first (independent, pipelined), so the serial
            // part is one multiply-add and one subtract per sample. sqrt: branch-free Newton form (the library routine's
            // special-case branches serialise the 16 evaluations); <= 1 ulp from the correctly rounded value.
            double mg[PB_SERIAL_BLK];
#pragma unroll
            for (int j = 0; j < PB_SERIAL_BLK; ++j) mg[j] = pb_sqrt_newton(cur[j].x * cur[j].x + cur[j].y * cur[j].y);   // DB::amplitude, db.h:28-30
#pragma unroll
            for (int j = 0; j < PB_SERIAL_BLK; ++j) {
                if (i0 + j < n) {
                    const double dn = (am_alpha * d) + mg[j];                             // demod_am.cpp:52
                    const double v = dn - d;                                              // demod_am.cpp:53
                    d = dn;
                    y[i0 + j] = cpx_make(v, v);
                }
            }
#pragma unroll
            for (int j = 0; j < PB_SERIAL_BLK; ++j) cur[j] = nxt[j];
        }
        st[c].am_dc_last = d;
    } else {                              // USB/LSB/...: Demod::processBlock returns `in` (demod.cpp:127-138)
        for (int i = 0; i < n; ++i) y[i] = x[i];
    }
}

// CFir::ProcessFilter over every channel (fir.cpp:106-132): virtual stream = hist ++ in, per-mode taps.
// taps: [2][PB_FIR_MAX] (row 0 AM low-pass, row 1 NFM low-pass); ntaps[2]. Pass-through channels are copied.
static __global__ void chan_cfir_kernel(const cpx *__restrict__ hist, const cpx *__restrict__ in, long long in_stride, int n,
                                        const double *__restrict__ taps, int ntaps_am, int ntaps_fm,
                                        const int *__restrict__ modes, cpx *__restrict__ out, long long out_stride,
                                        cpx *__restrict__ new_hist) {
    __shared__ double sh[PB_FIR_MAX];
    const int c = blockIdx.y;
    const int mode = modes[c];
    const cpx *x = in + c * in_stride;
    cpx *y = out + c * out_stride;
    if (mode != CH_AM && mode != CH_FMN) {
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) y[i] = x[i];
        return;
    }
    const int ntaps = mode == CH_AM ? ntaps_am : ntaps_fm;
    const double *h = taps + (mode == CH_AM ? 0 : PB_FIR_MAX);
    for (int k = threadIdx.x; k < ntaps; k += blockDim.x) sh[k] = h[k];
    __syncthreads();
    const int hl = PB_FIR_MAX;   // history rows are PB_FIR_MAX long; the last ntaps-1 entries are used
    const cpx *hc = hist + (long long)c * hl + (hl - (ntaps - 1));
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        double re = 0.0, im = 0.0;
        for (int k = 0; k < ntaps; ++k) {
            const int v = i + (ntaps - 1) - k;
            const cpx s = v < ntaps - 1 ? hc[v] : x[v - (ntaps - 1)];
            re += sh[k] * s.x;     // I and Q tap sets are equal for Init*Filter designs (fir.cpp:298-302)
            im += sh[k] * s.y;
        }
        y[i] = cpx_make(re, im);
    }
    // the last PB_FIR_MAX samples of hist ++ in become the next history (block 0 of each channel writes them)
    if (blockIdx.x == 0) {
        cpx *nh = new_hist + (long long)c * hl;
        const cpx *ho = hist + (long long)c * hl;
        for (int i = threadIdx.x; i < hl; i += blockDim.x) {
            const long long v = (long long)n + i;      // index into (hist ++ in)
            nh[i] = v < hl ? ho[v] : x[v - hl];
        }
    }
}

struct pebble_channelizer {
    int nch = 0, S = 0, fpb = 2048;
    double in_rate = 0, max_bw = 0, out_rate = 0;
    pebble_downconvert_t *dc = nullptr;
    pebble_fastfir_t *bp = nullptr;
    std::vector<int> modes;
    int *d_modes = nullptr, *d_order = nullptr;
    int slots = 0;                     // entries of d_order (multiple of 32)
    bool modes_dirty = true;
    PllConst pll;
    ChanState *d_state = nullptr;
    double taps[2 * PB_FIR_MAX];
    int ntaps_am = 0, ntaps_fm = 0;
    double *d_taps = nullptr;
    cpx *d_hist[2] = {nullptr, nullptr};
    int hcur = 0;
    int agc_threshold = 20;
    double am_bandwidth = 16000.0;     // Demod_AM ctor default (demod_am.cpp:9)
    cpx *d_dec = nullptr, *d_bp = nullptr, *d_tmp = nullptr;
    size_t row_cap = 0;                // samples per channel the work buffers hold
    // The per-channel demodulators are serial recurrences (PLL feedback): one thread per channel walks its row, and that
    // walk takes the same ~1 us per decimated sample whether 1024 or 128 channels are resident. A call is therefore cut
    // into `parts` runs of whole receiver frames; the demodulator + CFir of part k run on `tail` while the caller's stream
    // already down-converts and filters part k+1 (all operators carry their state, so split == unsplit).
    cudaStream_t tail = nullptr, tail2 = nullptr;   // tail: the serial demodulator walks; tail2: their post filters
    cudaEvent_t ev_front = nullptr, ev_tail = nullptr, ev_demod = nullptr, ev_cfir = nullptr;
    int parts = 4;
};

static int chz_upload_taps(pebble_channelizer *h) {
    h->ntaps_am = pebble_host::cfir_design_lp(0, 1.0, 50.0, h->am_bandwidth, h->am_bandwidth * 1.8, h->out_rate, h->taps);           // demod_am.cpp:21
    h->ntaps_fm = pebble_host::cfir_design_lp(0, 1.0, 50.0, 3000.0, 1.6 * 3000.0, h->out_rate, h->taps + PB_FIR_MAX);              // demod_nfm.cpp:60-63
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(h->d_taps, h->taps, sizeof(h->taps), cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    const size_t hb = sizeof(cpx) * PB_FIR_MAX * (size_t)h->nch;
    PB_CUDA(cudaMemset(h->d_hist[0], 0, hb), PEBBLE_ERR_CUDA);      // Init*Filter zeroes the delay line (fir.cpp:305-311)
    PB_CUDA(cudaMemset(h->d_hist[1], 0, hb), PEBBLE_ERR_CUDA);
    return 0;
}

static int chz_reserve(pebble_channelizer *h, size_t row) {
    if (row <= h->row_cap) return 0;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    cudaFree(h->d_dec); cudaFree(h->d_bp); cudaFree(h->d_tmp);
    h->d_dec = h->d_bp = h->d_tmp = nullptr; h->row_cap = 0;
    const size_t nb = sizeof(cpx) * row * (size_t)h->nch;
    PB_CUDA(cudaMalloc(&h->d_dec, nb), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMalloc(&h->d_bp, nb), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMalloc(&h->d_tmp, nb), PEBBLE_ERR_CUDA);
    h->row_cap = row;
    return 0;
}

extern "C" {

pebble_channelizer_t *pebble_channelizer_create(int nchannels, double in_rate, double max_bw) {
    if (nchannels < 1) { pebble::set_error("channelizer: nchannels < 1"); return nullptr; }
    pebble_channelizer *h = new pebble_channelizer();
    h->nch = nchannels; h->in_rate = in_rate; h->max_bw = max_bw;
    h->dc = pebble_downconvert_create();
    h->bp = pebble_fastfir_create();
    if (!h->dc || !h->bp) { pebble_channelizer_destroy(h); return nullptr; }
    if (const char *e = getenv("PEBBLE_CHZ_PARTS")) h->parts = atoi(e) > 0 ? atoi(e) : 1;
    // highest priority: the tail's few serial warps should get SM slots ahead of the next part's many front CTAs
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
    if (!pebble::cuda_ok(cudaStreamCreateWithPriority(&h->tail, cudaStreamNonBlocking, prio_hi), "cudaStreamCreate") ||
        !pebble::cuda_ok(cudaStreamCreateWithPriority(&h->tail2, cudaStreamNonBlocking, prio_hi), "cudaStreamCreate") ||
        !pebble::cuda_ok(cudaEventCreateWithFlags(&h->ev_demod, cudaEventDisableTiming), "cudaEventCreate") ||
        !pebble::cuda_ok(cudaEventCreateWithFlags(&h->ev_cfir, cudaEventDisableTiming), "cudaEventCreate") ||
        !pebble::cuda_ok(cudaEventCreateWithFlags(&h->ev_front, cudaEventDisableTiming), "cudaEventCreate") ||
        !pebble::cuda_ok(cudaEventCreateWithFlags(&h->ev_tail, cudaEventDisableTiming), "cudaEventCreate")) {
        pebble_channelizer_destroy(h);
        return nullptr;
    }
    h->out_rate = pebble_downconvert_set_data_rate(h->dc, in_rate, max_bw);
    h->S = pebble_downconvert_num_stages(h->dc);
    bool ok = pebble_downconvert_set_channels(h->dc, nchannels) == 0 && pebble_fastfir_set_streams(h->bp, nchannels) == 0;
    const size_t hb = sizeof(cpx) * PB_FIR_MAX * (size_t)nchannels;
    ok = ok && pebble::cuda_ok(cudaMalloc(&h->d_modes, sizeof(int) * nchannels), "cudaMalloc") &&
         pebble::cuda_ok(cudaMalloc(&h->d_order, sizeof(int) * ((size_t)nchannels + 3 * 32)), "cudaMalloc") &&
         pebble::cuda_ok(cudaMalloc(&h->d_state, sizeof(ChanState) * nchannels), "cudaMalloc") &&
         pebble::cuda_ok(cudaMemset(h->d_state, 0, sizeof(ChanState) * nchannels), "cudaMemset") &&
         pebble::cuda_ok(cudaMalloc(&h->d_taps, sizeof(h->taps)), "cudaMalloc") &&
         pebble::cuda_ok(cudaMalloc(&h->d_hist[0], hb), "cudaMalloc") && pebble::cuda_ok(cudaMalloc(&h->d_hist[1], hb), "cudaMalloc");
    if (!ok) { pebble_channelizer_destroy(h); return nullptr; }
    h->modes.assign(nchannels, CH_AM);   // Demod starts in dmAM
    // Demod_NFM::init (demod_nfm.cpp:44-66); members are float
    const double norm = 6.28318530717958647692528676656 / (double)(unsigned)h->out_rate;
    h->pll.lo = (float)(-15000.0 * norm);
    h->pll.hi = (float)(15000.0 * norm);
    h->pll.alpha = (float)(2.0 * .707 * 3000.0 * norm);
    h->pll.beta = (float)((h->pll.alpha * h->pll.alpha) / (4.0 * .707 * .707));
    h->pll.out_gain = 1.0f;
    h->pll.dc_alpha = (float)(1.0 - exp(-1.0 / ((double)(unsigned)h->out_rate * 0.001)));
    if (chz_upload_taps(h) < 0) { pebble_channelizer_destroy(h); return nullptr; }
    return h;
}
void pebble_channelizer_destroy(pebble_channelizer_t *h) {
    if (!h) return;
    if (h->dc) pebble_downconvert_destroy(h->dc);
    if (h->bp) pebble_fastfir_destroy(h->bp);
    cudaFree(h->d_modes); cudaFree(h->d_order); cudaFree(h->d_state); cudaFree(h->d_taps); cudaFree(h->d_hist[0]); cudaFree(h->d_hist[1]);
    cudaFree(h->d_dec); cudaFree(h->d_bp); cudaFree(h->d_tmp);
    if (h->tail) cudaStreamDestroy(h->tail);
    if (h->tail2) cudaStreamDestroy(h->tail2);
    if (h->ev_demod) cudaEventDestroy(h->ev_demod);
    if (h->ev_cfir) cudaEventDestroy(h->ev_cfir);
    if (h->ev_front) cudaEventDestroy(h->ev_front);
    if (h->ev_tail) cudaEventDestroy(h->ev_tail);
    delete h;
}
double pebble_channelizer_out_rate(const pebble_channelizer_t *h) { return h ? h->out_rate : 0.0; }
int pebble_channelizer_num_stages(const pebble_channelizer_t *h) { return h ? h->S : PEBBLE_ERR_ARG; }
int pebble_channelizer_num_channels(const pebble_channelizer_t *h) { return h ? h->nch : PEBBLE_ERR_ARG; }
int pebble_channelizer_set_frequency(pebble_channelizer_t *h, int ch, double f) {   // Receiver::mixerChanged per channel
    if (!h) return PEBBLE_ERR_ARG;
    return pebble_downconvert_set_channel_frequency(h->dc, ch, f);
}
int pebble_channelizer_set_filter(pebble_channelizer_t *h, double lo, double hi) {   // BandPassFilter::setBandPass, all channels
    if (!h) return PEBBLE_ERR_ARG;
    return pebble_fastfir_setup(h->bp, lo, hi, 0, h->out_rate);
}
int pebble_channelizer_set_demod_mode(pebble_channelizer_t *h, int ch, int mode) {
    if (!h || ch < 0 || ch >= h->nch || mode < 0 || mode > 12) return PEBBLE_ERR_ARG;
    if (mode == 1 || mode == 3 || mode == 4) { pebble::set_error("channelizer: SAM / WFM modes are out of scope"); return PEBBLE_ERR_STATE; }
    h->modes[ch] = mode;
    h->modes_dirty = true;
    return 0;
}
int pebble_channelizer_set_am_bandwidth(pebble_channelizer_t *h, double bw) {
    if (!h) return PEBBLE_ERR_ARG;
    h->am_bandwidth = bw;
    return chz_upload_taps(h);
}
int pebble_channelizer_set_agc_threshold(pebble_channelizer_t *h, int threshold) {   // AGC_OFF only (agc.cpp:86-94)
    if (!h) return PEBBLE_ERR_ARG;
    h->agc_threshold = threshold;
    return 0;
}
// Raw device samples into every channel (normalizeIQ fused into the front end's load; see "Device sample formats").
int pebble_channelizer_set_input_format(pebble_channelizer_t *h, int fmt, int order, double user_gain, double norm_gain) {
    if (!h) return PEBBLE_ERR_ARG;
    return pebble_downconvert_set_input_format(h->dc, fmt, order, user_gain, norm_gain);
}
// 1: polyphase-DFT front end (channels on the grid in_rate / 2^S), 0: one down-converter per channel, -1: no call yet
int pebble_channelizer_front_path(const pebble_channelizer_t *h) { return h ? pebble_downconvert_batch_path(h->dc) : PEBBLE_ERR_ARG; }
int pebble_channelizer_out_count(const pebble_channelizer_t *h, int n) {
    if (!h || n < 0) return PEBBLE_ERR_ARG;
    return n >> h->S;
}

// Decimated streams only (step 1), [channel][n >> S] rows `out_stride` cpx apart.
int pebble_channelizer_downconvert_device(pebble_channelizer_t *h, int n, const void *d_in, double *d_out,
                                          long long out_stride, void *stream) {
    if (!h) return PEBBLE_ERR_ARG;
    return pebble_downconvert_process_channels_device(h->dc, n, d_in, d_out, out_stride, stream);
}

// Whole chain. n must make (n >> S) a multiple of framesPerBuffer (2048), so that every call yields whole receiver
// frames (receiver.cpp:922-931) for every channel; returns the demodulated samples per channel (= n >> S).
int pebble_channelizer_process_device(pebble_channelizer_t *h, int n, const void *d_in, double *d_out, long long out_stride,
                                      void *stream_v) {
    if (!h || n < 0 || (n && (!d_in || !d_out))) { pebble::set_error("channelizer: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    const int nd = n >> h->S;
    if ((n & ((1 << h->S) - 1)) || nd % h->fpb) {
        pebble::set_error("channelizer: n=%d must be a multiple of 2^%d * %d", n, h->S, h->fpb);
        return PEBBLE_ERR_ARG;
    }
    if (out_stride < nd) { pebble::set_error("channelizer: out_stride smaller than the output count"); return PEBBLE_ERR_ARG; }
    cudaStream_t st = (cudaStream_t)stream_v;
    int r = chz_reserve(h, (size_t)nd);
    if (r < 0) return r;
    if (h->modes_dirty) {
        PB_CUDA(cudaMemcpyAsync(h->d_modes, h->modes.data(), sizeof(int) * h->nch, cudaMemcpyHostToDevice, st), PEBBLE_ERR_CUDA);
        // thread slots of chan_demod_kernel: NBFM channels, then AM, then pass-through, each group padded to a whole warp
        std::vector<int> order;
        for (int grp = 0; grp < 3; ++grp) {
            for (int c = 0; c < h->nch; ++c) {
                const int g = h->modes[c] == CH_FMN ? 0 : h->modes[c] == CH_AM ? 1 : 2;
                if (g == grp) order.push_back(c);
            }
            while (order.size() % 32) order.push_back(-1);
        }
        h->slots = (int)order.size();
        PB_CUDA(cudaMemcpy(h->d_order, order.data(), sizeof(int) * order.size(), cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
        h->modes_dirty = false;
    }
    const long long row = (long long)h->row_cap;
    const double gain = pow(10, (h->S * 2) / 20.0);                                                  // receiver.cpp:935-938
    const double agc_gain = pow(10, (h->agc_threshold / 5) / 20.0);                                  // agc.cpp:86-94 (integer division)
    pebble_fastfir_set_fused_scale(h->bp, gain, agc_gain);
    const int frames = nd / h->fpb;
    const int parts = h->parts < frames ? h->parts : frames;
    const size_t bps = (size_t)pebble_downconvert_input_bytes_per_sample(h->dc);
    // `tail` starts after everything already queued on the caller's stream (mode upload, the consumer of the last output)
    PB_CUDA(cudaEventRecord(h->ev_front, st), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamWaitEvent(h->tail, h->ev_front, 0), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamWaitEvent(h->tail2, h->ev_front, 0), PEBBLE_ERR_CUDA);
    int f0 = 0;
    for (int k = 0; k < parts; ++k) {
        const int f1 = (int)((long long)frames * (k + 1) / parts);
        const int pd = (f1 - f0) * h->fpb;             // decimated samples of this part
        const long long pos = (long long)f0 * h->fpb;  // its column in the work rows and in the output rows
        const long long pn = (long long)pd << h->S;
        r = pebble_downconvert_process_channels_device(h->dc, (int)pn, (const char *)d_in + bps * (size_t)(pos << h->S),
                                                       (double *)(h->d_dec + pos), row, st);   // receiver.cpp:908
        if (r < 0) return r;
        r = pebble_fastfir_process_batch_device(h->bp, pd, (const double *)(h->d_dec + pos), row, (double *)(h->d_bp + pos), row, st);   // :950
        if (r < 0) return r;
        if (r != pd) { pebble::set_error("channelizer: FastFIR returned %d of %d", r, pd); return PEBBLE_ERR_STATE; }
        PB_CUDA(cudaEventRecord(h->ev_front, st), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaStreamWaitEvent(h->tail, h->ev_front, 0), PEBBLE_ERR_CUDA);
        chan_demod_kernel<<<h->slots / 32, 32, 0, h->tail>>>(h->d_bp + pos, h->d_tmp + pos, pd, row, h->nch, h->d_modes, h->d_order,
                                                            h->slots, h->pll, (double)0.9999f, h->d_state, h->fpb);                         // :987
        // the walks of consecutive parts form the critical chain (32 warps, latency-bound); the post filter of part k runs
        // beside the walk of part k+1 on its own stream
        PB_CUDA(cudaEventRecord(h->ev_demod, h->tail), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaStreamWaitEvent(h->tail2, h->ev_demod, 0), PEBBLE_ERR_CUDA);
        int gx = (pd + 255) / 256;
        if (gx > 64) gx = 64;
        chan_cfir_kernel<<<dim3(gx, h->nch), 256, 0, h->tail2>>>(h->d_hist[h->hcur], h->d_tmp + pos, row, pd, h->d_taps, h->ntaps_am,
                                                                h->ntaps_fm, h->d_modes, (cpx *)d_out + pos, out_stride,
                                                                h->d_hist[h->hcur ^ 1]);
        pebble::count_launch(2);
        PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
        h->hcur ^= 1;
        f0 = f1;
    }
    // the call is complete on the caller's stream when both tails are
    PB_CUDA(cudaEventRecord(h->ev_tail, h->tail), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamWaitEvent(st, h->ev_tail, 0), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaEventRecord(h->ev_cfir, h->tail2), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamWaitEvent(st, h->ev_cfir, 0), PEBBLE_ERR_CUDA);
    return nd;
}

}  // extern "C"
