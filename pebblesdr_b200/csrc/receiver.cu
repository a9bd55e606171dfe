// receiver.cu — the non-WFM branch of Receiver::processIQData (application/receiver.cpp:903-996) as one
// device-resident chain: CDownConvert -> re-framing to framesPerBuffer -> gain restore -> BandPassFilter
// (CFastFIR) -> AGC -> Demod. Every step is the same GPU operator the stand-alone C ABI exposes; the
// chain only owns the intermediate device buffers and fuses the two scalar gains into the FastFIR kernel.
#include "lib_common.h"
#include "common.cuh"
#include "demod_internal.h"
#include <cmath>
#include <cstring>
#include <vector>

// resampler.cu
struct pebble_resampler;
int resampler_run_device(pebble_resampler *h, int n, int frame, double rate, const cpx *d_in, cpx *d_out, bool clip, cudaStream_t st);
long long resampler_peek_count(pebble_resampler *h, int n, int frame, double rate);

struct pebble_receiver {
    double sample_rate = 0, demod_rate = 0;
    int fpb = 2048;                // framesPerBuffer (settings.cpp:57)
    pebble_downconvert_t *dc = nullptr;
    // the WFM branch of processIQData (receiver.cpp:854-901): its own down-converter to < 400 kS/s (SetDataRateSimple, :217),
    // then straight to the demodulator — no band-pass, no gain restore, no AGC — and the resampler at wfmRate / audioRate
    pebble_downconvert_t *dc_wfm = nullptr;
    double wfm_rate = 0;
    int audio_rate = 0;
    bool was_wfm = false;
    pebble_fastfir_t *bp = nullptr;
    pebble_agc *agc = nullptr;
    pebble_demod *demod = nullptr;
    int gain_stages = -1;          // -1: use the CDownConvert stage count
    // audio tail (receiver.cpp:994-1007): fractional resampler demodRate -> audioRate once per frame, then the sink's clip
    pebble_resampler *rs = nullptr;
    double resamp_rate = 1.0;      // resampRate = demodSampleRate / audioOutRate (:994)
    bool audio_on = false, audio_clip = false;
    cpx *d_demod = nullptr; size_t demod_cap = 0;
    cpx *d_frames = nullptr; size_t frames_cap = 0;   // m_sampleBuf: decimated samples waiting for a full frame
    int fill = 0;                                      // m_sampleBufLen
    cpx *d_bp = nullptr, *d_agc = nullptr; size_t work_cap = 0;
    void *d_pin_in = nullptr, *d_pin_out = nullptr; size_t pin_in_cap = 0, pin_out_cap = 0;
    // pinned path: chunked, double-buffered upload on its own stream (the chain of chunk c runs under the copy of chunk c+1)
    void *d_pin_chunk[2] = {nullptr, nullptr}; size_t pin_chunk_cap = 0;
    cudaStream_t copy_st = nullptr;
    cudaEvent_t ev_up[2] = {nullptr, nullptr}, ev_free[2] = {nullptr, nullptr};
    bool pin_used[2] = {false, false};
    cudaStream_t stream = nullptr;
    pebble::Staging stg;
    // optional per-step device timing (bench.py's roofline leg): 5 events per process call, recorded on the
    // caller's stream: [before dc, after dc, after fastfir, after agc, after demod]
    bool profiling = false;
    std::vector<cudaEvent_t> ev;
    size_t ev_used = 0;
};

static void rx_mark(pebble_receiver *h, cudaStream_t st) {
    if (!h->profiling) return;
    if (h->ev_used == h->ev.size()) {
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return;
        h->ev.push_back(e);
    }
    cudaEventRecord(h->ev[h->ev_used++], st);
}

static int rx_reserve(pebble_receiver *h, size_t dec_total) {
    if (dec_total > h->frames_cap) {
        cpx *nb = nullptr;
        PB_CUDA(cudaMalloc(&nb, sizeof(cpx) * dec_total), PEBBLE_ERR_CUDA);
        if (h->d_frames) {
            PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
            PB_CUDA(cudaMemcpy(nb, h->d_frames, sizeof(cpx) * (size_t)h->fill, cudaMemcpyDeviceToDevice), PEBBLE_ERR_CUDA);
            cudaFree(h->d_frames);
        }
        h->d_frames = nb;
        h->frames_cap = dec_total;
    }
    if (dec_total > h->work_cap) {
        PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
        cudaFree(h->d_bp); cudaFree(h->d_agc);
        h->d_bp = h->d_agc = nullptr; h->work_cap = 0;
        PB_CUDA(cudaMalloc(&h->d_bp, sizeof(cpx) * dec_total), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMalloc(&h->d_agc, sizeof(cpx) * dec_total), PEBBLE_ERR_CUDA);
        h->work_cap = dec_total;
    }
    return 0;
}

extern "C" {

pebble_receiver_t *pebble_receiver_create(double sample_rate, int frames_per_buffer) {
    if (frames_per_buffer <= 0 || frames_per_buffer % 1024) {
        pebble::set_error("receiver: framesPerBuffer must be a positive multiple of 1024");
        return nullptr;
    }
    pebble_receiver *h = new pebble_receiver();
    h->sample_rate = sample_rate; h->fpb = frames_per_buffer;
    h->dc = pebble_downconvert_create();
    h->dc_wfm = pebble_downconvert_create();
    h->bp = pebble_fastfir_create();
    if (!h->dc || !h->dc_wfm || !h->bp || !pebble::cuda_ok(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "cudaStreamCreate")) {
        pebble_receiver_destroy(h);
        return nullptr;
    }
    h->demod_rate = sample_rate;
    return h;
}
void pebble_receiver_destroy(pebble_receiver_t *h) {
    if (!h) return;
    if (h->dc) pebble_downconvert_destroy(h->dc);
    if (h->dc_wfm) pebble_downconvert_destroy(h->dc_wfm);
    if (h->bp) pebble_fastfir_destroy(h->bp);
    if (h->agc) pebble_agc_destroy(h->agc);
    if (h->demod) pebble_demod_destroy(h->demod);
    cudaFree(h->d_frames); cudaFree(h->d_bp); cudaFree(h->d_agc); cudaFree(h->d_pin_in); cudaFree(h->d_pin_out); cudaFree(h->d_demod);
    cudaFree(h->d_pin_chunk[0]); cudaFree(h->d_pin_chunk[1]);
    if (h->copy_st) cudaStreamDestroy(h->copy_st);
    for (int b = 0; b < 2; ++b) { if (h->ev_up[b]) cudaEventDestroy(h->ev_up[b]); if (h->ev_free[b]) cudaEventDestroy(h->ev_free[b]); }
    if (h->rs) pebble_resampler_destroy(h->rs);
    if (h->stream) cudaStreamDestroy(h->stream);
    for (cudaEvent_t e : h->ev) cudaEventDestroy(e);
    h->stg.release();
    delete h;
}
// m_downConvert1.SetDataRate(sampleRate, maxBW) -> demod sample rate; (re)creates AGC and Demod at that rate
double pebble_receiver_set_data_rate(pebble_receiver_t *h, double max_bw) {
    if (!h) return 0.0;
    h->demod_rate = pebble_downconvert_set_data_rate(h->dc, h->sample_rate, max_bw);
    if (h->agc) pebble_agc_destroy(h->agc);
    if (h->demod) pebble_demod_destroy(h->demod);
    h->agc = pebble_agc_create((unsigned)h->demod_rate, (unsigned)h->fpb);
    h->demod = pebble_demod_create((unsigned)h->demod_rate, (unsigned)h->fpb);
    h->fill = 0;
    // receiver.cpp:217,228: m_demodWfmSampleRate = m_downConvertWfm1.SetDataRateSimple(sampleRate, 200000) (an int there);
    // Demod(demodSampleRate, demodWfmSampleRate, framesPerBuffer)
    h->wfm_rate = (double)(int)pebble_downconvert_set_data_rate_simple(h->dc_wfm, h->sample_rate, 200000.0);
    if (h->demod && h->wfm_rate > 0) pebble_demod_set_wfm_sample_rate(h->demod, h->wfm_rate);
    if (h->audio_rate > 0) pebble_receiver_set_audio(h, h->audio_rate, h->audio_clip);
    return (h->agc && h->demod) ? h->demod_rate : 0.0;
}
double pebble_receiver_get_wfm_rate(const pebble_receiver_t *h) { return h ? h->wfm_rate : 0.0; }
pebble_demod_t *pebble_receiver_demod(pebble_receiver_t *h) { return h ? h->demod : nullptr; }
void pebble_receiver_set_mixer(pebble_receiver_t *h, double freq) {   // receiver.cpp:716-719: both down-converters follow the mixer
    if (!h) return;
    pebble_downconvert_set_frequency(h->dc, freq);
    pebble_downconvert_set_frequency(h->dc_wfm, freq);
}
int pebble_receiver_set_filter(pebble_receiver_t *h, double lo, double hi) {  // BandPassFilter::setBandPass, bandpassfilter.cpp:35-45
    if (!h) return PEBBLE_ERR_ARG;
    return pebble_fastfir_setup(h->bp, lo, hi, 0, h->demod_rate);
}
int pebble_receiver_set_demod_mode(pebble_receiver_t *h, int mode) {
    if (!h || !h->demod) return PEBBLE_ERR_STATE;
    return pebble_demod_set_mode(h->demod, mode);
}
int pebble_receiver_set_demod_bandwidth(pebble_receiver_t *h, double bw) {
    if (!h || !h->demod) return PEBBLE_ERR_STATE;
    return pebble_demod_set_bandwidth(h->demod, bw);
}
int pebble_receiver_set_agc(pebble_receiver_t *h, int mode, int threshold) {
    if (!h || !h->agc) return PEBBLE_ERR_STATE;
    return pebble_agc_set_mode(h->agc, mode, threshold);
}
// Device sample format of the IQ stream (DeviceInterfaceBase::normalizeIQ, deviceinterfacebase.cpp:648-878): the
// conversion runs inside the down-converter's load stage, so `in` of the process calls holds raw device samples.
int pebble_receiver_set_input_format(pebble_receiver_t *h, int fmt, int order, double user_gain, double norm_gain) {
    if (!h) return PEBBLE_ERR_ARG;
    const int r = pebble_downconvert_set_input_format(h->dc, fmt, order, user_gain, norm_gain);
    return r < 0 ? r : pebble_downconvert_set_input_format(h->dc_wfm, fmt, order, user_gain, norm_gain);
}
int pebble_receiver_set_gain_stages(pebble_receiver_t *h, int stages) {
    if (!h) return PEBBLE_ERR_ARG;
    h->gain_stages = stages;
    return 0;
}
// Audio tail of Receiver::processIQData (receiver.cpp:994-1007): resampRate = demodSampleRate / audioOutRate; if it is not 1
// every demodulated frame goes through CFractResampler::Resample (else it is copied), then to the audio sink, whose only
// arithmetic is the +-0.9999 clip (AudioQT::SendToOutput, audioqt.cpp:190-201; clip != 0). audio_rate <= 0 switches the tail
// off again (the chain then ends at the demodulator, as before). The demod rate is truncated to an integer like
// m_demodSampleRate.
int pebble_receiver_set_audio(pebble_receiver_t *h, int audio_rate, int clip) {
    if (!h) return PEBBLE_ERR_ARG;
    if (audio_rate <= 0) { h->audio_on = false; h->audio_rate = 0; return 0; }
    if (!h->rs) h->rs = pebble_resampler_create(h->fpb);
    if (!h->rs) return PEBBLE_ERR_CUDA;
    h->audio_rate = audio_rate;
    h->resamp_rate = ((int)h->demod_rate * 1.0) / (audio_rate * 1.0);
    h->audio_on = true;
    h->audio_clip = clip != 0;
    return 0;
}
int pebble_receiver_set_profiling(pebble_receiver_t *h, int on) {
    if (!h) return PEBBLE_ERR_ARG;
    h->profiling = on != 0;
    h->ev_used = 0;
    return 0;
}
// Sums the device time of each step over the calls made since profiling was (re)enabled, then resets.
// ms[0] CDownConvert, ms[1] CFastFIR (+fused gains), ms[2] AGC (0 when fused), ms[3] Demod. Synchronises.
int pebble_receiver_get_profile(pebble_receiver_t *h, double *ms, int64_t *calls) {
    if (!h || !ms) return PEBBLE_ERR_ARG;
    ms[0] = ms[1] = ms[2] = ms[3] = 0.0;
    const size_t nc = h->ev_used / 5;
    if (nc) PB_CUDA(cudaEventSynchronize(h->ev[nc * 5 - 1]), PEBBLE_ERR_CUDA);
    for (size_t c = 0; c < nc; ++c)
        for (int k = 0; k < 4; ++k) {
            float t = 0.f;
            PB_CUDA(cudaEventElapsedTime(&t, h->ev[c * 5 + k], h->ev[c * 5 + k + 1]), PEBBLE_ERR_CUDA);
            ms[k] += t;
        }
    if (calls) *calls = (int64_t)nc;
    h->ev_used = 0;
    return 0;
}
static bool rx_wfm(const pebble_receiver *h) {   // receiver.cpp:854
    const int m = h->demod ? demod_mode(h->demod) : -1;
    return m == 3 || m == 4;
}
static double rx_resamp_rate(const pebble_receiver *h, bool wfm) {   // receiver.cpp:901 / :994
    return wfm ? ((int)h->wfm_rate * 1.0) / (h->audio_rate * 1.0) : h->resamp_rate;
}
int pebble_receiver_out_count(const pebble_receiver_t *h, int n) {
    if (!h || n < 0) return PEBBLE_ERR_ARG;
    const bool wfm = rx_wfm(h);
    const int S = pebble_downconvert_num_stages(wfm ? h->dc_wfm : h->dc);
    const int fill = wfm == h->was_wfm ? h->fill : 0;
    const int nproc = (int)((((long long)fill + (n >> S)) / h->fpb) * h->fpb);
    const double rr = rx_resamp_rate(h, wfm);
    if (h->audio_on && rr != 1.0 && nproc > 0) return (int)resampler_peek_count(h->rs, nproc, h->fpb, rr);
    return nproc;
}

// The WFM branch (receiver.cpp:854-901 + the common tail :994-1007): m_downConvertWfm1 -> full frames -> Demod (fmMono /
// fmStereo, one call per frame: the pilot lock decision and the loops' phase wrap are per call) -> resampler.
static int rx_process_wfm(pebble_receiver *h, int n, const void *d_in, double *d_out, cudaStream_t st) {
    const int S = pebble_downconvert_num_stages(h->dc_wfm);
    const int ndec = n >> S;
    int r = rx_reserve(h, (size_t)h->fill + (size_t)ndec);
    if (r < 0) return r;
    r = pebble_downconvert_process_device(h->dc_wfm, n, d_in, (double *)(h->d_frames + h->fill), st);
    if (r < 0) return r;
    const int total = h->fill + ndec, nframes = total / h->fpb, nproc = nframes * h->fpb;
    if (nframes == 0) { h->fill = total; return 0; }
    if (!d_out) { pebble::set_error("receiver: NULL output"); return PEBBLE_ERR_ARG; }
    const double rr = rx_resamp_rate(h, true);
    const bool resample = h->audio_on && rr != 1.0;
    cpx *demod_out = (cpx *)d_out;
    if (resample) {
        if ((size_t)nproc > h->demod_cap) {
            PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
            cudaFree(h->d_demod); h->d_demod = nullptr; h->demod_cap = 0;
            PB_CUDA(cudaMalloc(&h->d_demod, sizeof(cpx) * (size_t)nproc), PEBBLE_ERR_CUDA);
            h->demod_cap = (size_t)nproc;
        }
        demod_out = h->d_demod;
    }
    for (int f = 0; f < nframes; ++f) {
        r = demod_run_device(h->demod, h->fpb, h->d_frames + (size_t)f * h->fpb, demod_out + (size_t)f * h->fpb, st, nullptr, 0);
        if (r < 0) return r;
    }
    int produced = nproc;
    if (resample) {
        produced = resampler_run_device(h->rs, nproc, h->fpb, rr, h->d_demod, (cpx *)d_out, h->audio_clip, st);
        if (produced < 0) return produced;
    }
    const int rem = total - nproc;
    if (rem > 0)
        PB_CUDA(cudaMemcpyAsync(h->d_frames, h->d_frames + nproc, sizeof(cpx) * (size_t)rem, cudaMemcpyDeviceToDevice, st), PEBBLE_ERR_CUDA);
    h->fill = rem;
    return produced;
}

int pebble_receiver_process_device(pebble_receiver_t *h, int n, const void *d_in, double *d_out, void *stream_v) {
    if (!h || n < 0 || (n && !d_in)) { pebble::set_error("receiver: bad argument"); return PEBBLE_ERR_ARG; }
    if (!h->agc || !h->demod) { pebble::set_error("receiver: set_data_rate not called"); return PEBBLE_ERR_STATE; }
    if (n == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream_v;
    const bool wfm = rx_wfm(h);
    if (wfm != h->was_wfm) { h->fill = 0; h->was_wfm = wfm; }   // the two branches decimate to different rates: the partial frame is dropped
    if (wfm) return rx_process_wfm(h, n, d_in, d_out, st);
    const int S = pebble_downconvert_num_stages(h->dc);
    const int ndec = n >> S;
    int r = rx_reserve(h, (size_t)h->fill + (size_t)ndec);
    if (r < 0) return r;
    // receiver.cpp:908 + 922-924: decimated samples are appended to m_sampleBuf
    const size_t ev_entry = h->ev_used;   // roll-back point of the profiling events if a step fails (5 events per call)
    rx_mark(h, st);
    r = pebble_downconvert_process_device(h->dc, n, d_in, (double *)(h->d_frames + h->fill), st);
    if (r < 0) { h->ev_used = ev_entry; return r; }
    rx_mark(h, st);
    const int total = h->fill + ndec;
    const int nframes = total / h->fpb;
    const int nproc = nframes * h->fpb;
    if (nframes == 0) {  // receiver.cpp:927-928
        h->fill = total;
        rx_mark(h, st); rx_mark(h, st); rx_mark(h, st);
        return 0;
    }
    if (!d_out) { pebble::set_error("receiver: NULL output"); return PEBBLE_ERR_ARG; }
    // receiver.cpp:935-938: restore decimation gain, 2 dB per stage
    const int gs = h->gain_stages >= 0 ? h->gain_stages : S;
    const double gain = pow(10, (gs * 2) / 20.0);
    const bool fuse_agc = agc_is_off(h->agc);
    pebble_fastfir_set_fused_scale(h->bp, gain, fuse_agc ? agc_manual_gain(h->agc) : 1.0);
    if (fuse_agc) {
        // AM with the AGC off: the FastFIR output IS the detector's input, so its kernel also reduces each output block
        // to the detector's affine map (saves the demodulator's first pass over the data and one launch)
        double alpha = 0.0;
        if (void *tiles = demod_am_expect_tiles(h->demod, nproc, &alpha))
            if (fastfir_am_aggregate_next(h->bp, tiles, alpha, nproc) != 0) demod_am_cancel_tiles(h->demod);
    }
    r = pebble_fastfir_process_device(h->bp, nproc, (const double *)h->d_frames, (double *)h->d_bp, st);  // receiver.cpp:950
    if (r != nproc) {
        // the armed AM tile maps were never written: disarm them, and drop this call's profiling events
        demod_am_cancel_tiles(h->demod);
        if (h->profiling) h->ev_used = ev_entry;
        if (r < 0) return r;
        pebble::set_error("receiver: FastFIR returned %d of %d", r, nproc);
        return PEBBLE_ERR_STATE;
    }
    rx_mark(h, st);
    const cpx *next = h->d_bp;
    if (!fuse_agc) {  // receiver.cpp:983
        r = agc_run_device(h->agc, nproc, next, h->d_agc, st);
        if (r < 0) { if (h->profiling) h->ev_used = ev_entry; return r; }
        next = h->d_agc;
    }
    rx_mark(h, st);
    const bool resample = h->audio_on && h->resamp_rate != 1.0;
    cpx *demod_out = (cpx *)d_out;
    if (resample) {
        if ((size_t)nproc > h->demod_cap) {
            PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
            cudaFree(h->d_demod); h->d_demod = nullptr; h->demod_cap = 0;
            PB_CUDA(cudaMalloc(&h->d_demod, sizeof(cpx) * (size_t)nproc), PEBBLE_ERR_CUDA);
            h->demod_cap = (size_t)nproc;
        }
        demod_out = h->d_demod;
    }
    r = demod_run_device(h->demod, nproc, next, demod_out, st, nullptr, h->fpb);  // receiver.cpp:987, once per frame
    if (r < 0) { demod_am_cancel_tiles(h->demod); if (h->profiling) h->ev_used = ev_entry; return r; }
    int produced = nproc;
    if (resample) {   // receiver.cpp:1000-1002, one Resample call per frame
        produced = resampler_run_device(h->rs, nproc, h->fpb, h->resamp_rate, h->d_demod, (cpx *)d_out, h->audio_clip, st);
        if (produced < 0) { if (h->profiling) h->ev_used = ev_entry; return produced; }
    }
    rx_mark(h, st);
    // keep the incomplete frame for the next call
    const int rem = total - nproc;
    if (rem > 0)
        PB_CUDA(cudaMemcpyAsync(h->d_frames, h->d_frames + nproc, sizeof(cpx) * (size_t)rem, cudaMemcpyDeviceToDevice, st), PEBBLE_ERR_CUDA);
    h->fill = rem;
    return produced;
}

int pebble_receiver_process(pebble_receiver_t *h, int n, const void *in, double *out) {
    if (!h || n < 0 || (n && !in)) { pebble::set_error("receiver: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    const int k = pebble_receiver_out_count(h, n);
    const size_t bi = (size_t)pebble_downconvert_input_bytes_per_sample(h->dc) * (size_t)n, bo = sizeof(cpx) * (size_t)(k > 0 ? k : 1);
    if (!h->stg.reserve(bi, bo)) return PEBBLE_ERR_CUDA;
    memcpy(h->stg.h_in, in, bi);
    PB_CUDA(cudaMemcpyAsync(h->stg.d_in, h->stg.h_in, bi, cudaMemcpyHostToDevice, h->stream), PEBBLE_ERR_CUDA);
    int r = pebble_receiver_process_device(h, n, (const double *)h->stg.d_in, (double *)h->stg.d_out, h->stream);
    if (r < 0) return r;
    if (r > 0) PB_CUDA(cudaMemcpyAsync(h->stg.h_out, h->stg.d_out, sizeof(cpx) * (size_t)r, cudaMemcpyDeviceToHost, h->stream), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamSynchronize(h->stream), PEBBLE_ERR_CUDA);
    if (r > 0 && out) memcpy(out, h->stg.h_out, sizeof(cpx) * (size_t)r);
    return r;
}
// Pinned-buffer form for end-to-end timing: `in`/`out` must be page-locked host memory (no extra memcpy).
// The input is uploaded in chunks of PB_PIN_CHUNK samples on a copy stream into two alternating device buffers, and the chain of
// chunk c runs on `stream` while chunk c+1 is still crossing PCIe (the chain carries its state, so chunked == unchunked); across
// consecutive calls the first upload of call k+1 overlaps the last chain run and the download of call k. Everything is enqueued
// without synchronising; `out` is complete when `stream` is.
#define PB_PIN_CHUNK (1 << 24)
int pebble_receiver_process_pinned(pebble_receiver_t *h, int n, const void *in, double *out, void *stream_v) {
    if (!h || n <= 0 || !in) { pebble::set_error("receiver: bad argument"); return PEBBLE_ERR_ARG; }
    cudaStream_t st = (cudaStream_t)stream_v;
    const int k = pebble_receiver_out_count(h, n);
    const size_t bps = (size_t)pebble_downconvert_input_bytes_per_sample(h->dc);
    const size_t bo = sizeof(cpx) * (size_t)(k > 0 ? k : 1);
    if (!h->copy_st) {
        PB_CUDA(cudaStreamCreateWithFlags(&h->copy_st, cudaStreamNonBlocking), PEBBLE_ERR_CUDA);
        for (int b = 0; b < 2; ++b) {
            PB_CUDA(cudaEventCreateWithFlags(&h->ev_up[b], cudaEventDisableTiming), PEBBLE_ERR_CUDA);
            PB_CUDA(cudaEventCreateWithFlags(&h->ev_free[b], cudaEventDisableTiming), PEBBLE_ERR_CUDA);
        }
    }
    const size_t chunk = (size_t)n < (size_t)PB_PIN_CHUNK ? (size_t)n : (size_t)PB_PIN_CHUNK;
    if (chunk * bps > h->pin_chunk_cap) {
        PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
        for (int b = 0; b < 2; ++b) {
            cudaFree(h->d_pin_chunk[b]); h->d_pin_chunk[b] = nullptr;
            PB_CUDA(cudaMalloc(&h->d_pin_chunk[b], chunk * bps), PEBBLE_ERR_CUDA);
            h->pin_used[b] = false;
        }
        h->pin_chunk_cap = chunk * bps;
    }
    if (bo > h->pin_out_cap) {
        PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
        cudaFree(h->d_pin_out); h->d_pin_out = nullptr; h->pin_out_cap = 0;
        PB_CUDA(cudaMalloc(&h->d_pin_out, bo), PEBBLE_ERR_CUDA);
        h->pin_out_cap = bo;
    }
    int produced = 0, c = 0;
    for (size_t pos = 0; pos < (size_t)n; pos += chunk, ++c) {
        const int nc = (int)((size_t)n - pos < chunk ? (size_t)n - pos : chunk);
        const int b = c & 1;
        if (h->pin_used[b]) PB_CUDA(cudaStreamWaitEvent(h->copy_st, h->ev_free[b], 0), PEBBLE_ERR_CUDA);   // its last reader has finished
        PB_CUDA(cudaMemcpyAsync(h->d_pin_chunk[b], (const char *)in + pos * bps, (size_t)nc * bps, cudaMemcpyHostToDevice, h->copy_st), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaEventRecord(h->ev_up[b], h->copy_st), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaStreamWaitEvent(st, h->ev_up[b], 0), PEBBLE_ERR_CUDA);
        const int r = pebble_receiver_process_device(h, nc, (const double *)h->d_pin_chunk[b], (double *)h->d_pin_out + 2 * (size_t)produced, st);
        PB_CUDA(cudaEventRecord(h->ev_free[b], st), PEBBLE_ERR_CUDA);
        h->pin_used[b] = true;
        if (r < 0) return r;
        produced += r;
    }
    if (produced > 0 && out) PB_CUDA(cudaMemcpyAsync(out, h->d_pin_out, sizeof(cpx) * (size_t)produced, cudaMemcpyDeviceToHost, st), PEBBLE_ERR_CUDA);
    return produced;
}

}  // extern "C"
