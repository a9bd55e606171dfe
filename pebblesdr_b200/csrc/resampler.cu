// resampler.cu — CFractResampler on the GPU: the step after the demodulator, Receiver::processIQData's
// `m_fractResampler.Resample(numStepSamples, resampRate, nextStep, m_audioBuf)` (application/receiver.cpp:994-1002), and
// the audio clip of AudioQT::SendToOutput (pebblelib/audioqt.cpp:190-201).
//
// Reference algorithm (pebblelib/fractresampler.cpp:146-187): an output-time accumulator m_FloatTime advances by
// dt = inRate/outRate per output; each output is a 28-tap convolution of the input around (int)m_FloatTime with a
// Blackman-Harris windowed sinc read from a 10000-points-per-zero-crossing table at the truncated fractional position;
// after the call m_FloatTime -= InLength and the last 28 input samples move to the head of the work buffer.
//
// GPU form. The output times do not depend on the data, only on (dt, InLength, m_FloatTime): they are a chain of rounded
// double additions, which is serial — but inside one binade every addition after the second adds the SAME rounded step
// (t is a multiple of the binade's ulp u, so fl(t + dt) = t + RN_u(dt); in the round-half-even tie case the result is even
// from the second add on, which fixes the direction), so a run is t_k = fma(k, step, t_a) EXACTLY. rs_segments() walks one
// call in O(binades) real additions and emits (first output, t, step, count) runs; the host uses it to size the output and
// carry m_FloatTime from frame to frame, each CTA re-derives the runs of its own frame into shared memory, and every
// output sample is then independent: its time from fma, its 28 products with separate multiply and add roundings in the
// reference's order (bit-identical to the reference compiled with -ffp-contract=off), table reads served by L2.
// The receiver calls Resample once per frame (framesPerBuffer samples); `frame` keeps that call structure for a batch.
#include "lib_common.h"
#include "common.cuh"
#include <cmath>
#include <cstring>
#include <vector>

#define RS_SINC_PERIOD_PTS 10000          // fractresampler.cpp:52
#define RS_SINC_PERIODS 28                // fractresampler.cpp:55
#define RS_SINC_LENGTH (RS_SINC_PERIODS * RS_SINC_PERIOD_PTS + 1)
#define RS_MAX_SEGS 160

struct RsSeg { int k0, count; double t0, step; };   // outputs k0 .. k0+count-1 at t = fma(k - k0, step, t0)

// Output times of ONE Resample call (fractresampler.cpp:159-176): t starts at `t` (m_FloatTime), one output while
// (int)t < in_len, t += dt. Returns the number of runs; *count_out = outputs, *t_end = m_FloatTime after the call
// (already reduced by in_len, :177). Runs never span a binade boundary; the first three values of every binade are
// real additions (see the header), the rest of the binade is one arithmetic run.
PB_HD int rs_segments(double t, double dt, int in_len, RsSeg *segs, int max_segs, int *count_out, double *t_end) {
    int k = 0, ns = 0;
    const double lim = (double)in_len;
    while (t < lim) {               // (int)t < in_len  <=>  t < in_len for t >= 0
        // up to three real additions inside the current binade
        int e;
        frexp(t, &e);               // t in [2^(e-1), 2^e); t == 0 gives e = 0
        double bend = t == 0.0 ? 4.9e-324 : ldexp(1.0, e);
        if (bend > lim) bend = lim;
        double tp = t, tc = t;
        int singles = 0;
        while (singles < 3 && tc < bend) {
            if (ns < max_segs) { segs[ns].k0 = k; segs[ns].count = 1; segs[ns].t0 = tc; segs[ns].step = 0.0; }
            ++ns; ++k; ++singles;
            tp = tc;
            tc = tc + dt;
        }
        // tc = the next time (not yet emitted). After three in-binade values the step tc - tp repeats while tc stays inside.
        if (singles == 3 && tc < bend) {
            const double step = tc - tp;
            long long m = (long long)((bend - tc) / step);       // values tc, tc+step, ... below bend: m+1 of them (estimate)
            while (m > 0 && fma((double)m, step, tc) >= bend) --m;
            while (fma((double)(m + 1), step, tc) < bend) ++m;
            if (ns < max_segs) { segs[ns].k0 = k; segs[ns].count = (int)(m + 1); segs[ns].t0 = tc; segs[ns].step = step; }
            ++ns;
            k += (int)(m + 1);
            tc = fma((double)m, step, tc) + dt;                  // the addition that leaves the run is a real one again
        }
        t = tc;
    }
    *count_out = k;
    *t_end = t - lim;
    return ns;
}

struct RsFrame { double t0; long long out_off; int count, pad; };

// One CTA per (frame, stream). in: virtual stream = hist (28 samples) ++ in rows; frame f covers in[f*frame .. (f+1)*frame).
// MODE 0: CPX out; MODE 1: CPX out clipped to +-0.9999 (AudioQT::SendToOutput); MODE 2: stereo int16 with gain + clip
// (fractresampler.cpp:196-250).
template <int MODE>
__global__ void __launch_bounds__(256) rs_kernel(const cpx *__restrict__ in, long long in_stride, const cpx *__restrict__ hist,
                                                 int frame, int last_len, int nframes, const RsFrame *__restrict__ fr, double dt,
                                                 const double *__restrict__ sinc, void *__restrict__ out_v, long long out_stride,
                                                 double gain) {
    __shared__ RsSeg segs[RS_MAX_SEGS];
    __shared__ int s_ns;
    const int f = blockIdx.x, s = blockIdx.y;
    const RsFrame F = fr[f];
    const int len = f == nframes - 1 ? last_len : frame;
    if (threadIdx.x == 0) {
        int cnt; double te;
        s_ns = rs_segments(F.t0, dt, len, segs, RS_MAX_SEGS, &cnt, &te);
    }
    __syncthreads();
    const int ns = s_ns;
    const cpx *x = in + (long long)s * in_stride + (long long)f * frame;   // frame's first input sample
    const cpx *h = hist + (long long)s * RS_SINC_PERIODS;
    const long long first = (long long)f * frame;                          // stream index (within the call) of x[0]
    for (int k = threadIdx.x; k < F.count; k += blockDim.x) {
        int lo = 0, hi = ns - 1;                                            // run that holds output k
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (segs[mid].k0 <= k) lo = mid; else hi = mid - 1;
        }
        const RsSeg sg = segs[lo];
        const double t = fma((double)(k - sg.k0), sg.step, sg.t0);
        const int it = (int)t;
        double ar = 0.0, ai = 0.0;
#pragma unroll 4
        for (int i = 1; i <= RS_SINC_PERIODS; ++i) {
            const int j = it + i;                                           // position in the reference's work buffer
            const int sindx = (int)(((double)j - t) * (double)RS_SINC_PERIOD_PTS);
            const long long g = first + j - RS_SINC_PERIODS;                // sample index within the call (negative: history)
            const cpx v = g < 0 ? h[g + RS_SINC_PERIODS] : x[j - RS_SINC_PERIODS];
            const double c = __ldg(sinc + sindx);
            ar = __dadd_rn(ar, __dmul_rn(v.x, c));
            ai = __dadd_rn(ai, __dmul_rn(v.y, c));
        }
        const long long o = (long long)s * out_stride + F.out_off + k;
        if (MODE == 0) {
            ((cpx *)out_v)[o] = cpx_make(ar, ai);
        } else if (MODE == 1) {
            const double mx = (double)0.9999f;                              // `const float maxOutput = 0.9999`, audioqt.cpp:182
            ar = ar > mx ? mx : (ar < -mx ? -mx : ar);
            ai = ai > mx ? mx : (ai < -mx ? -mx : ai);
            ((cpx *)out_v)[o] = cpx_make(ar, ai);
        } else {
            double tr = __dmul_rn(ar, gain), ti = __dmul_rn(ai, gain);
            tr = tr > 32767.0 ? 32767.0 : (tr < -32767.0 ? -32767.0 : tr);
            ti = ti > 32767.0 ? 32767.0 : (ti < -32767.0 ? -32767.0 : ti);
            ((short2 *)out_v)[o] = make_short2((short)tr, (short)ti);       // (qint16) truncates toward zero
        }
    }
}

// the last 28 samples of (hist ++ in) of every stream become the next history
static __global__ void rs_hist_kernel(const cpx *__restrict__ in, long long in_stride, int n, const cpx *__restrict__ hist,
                                      cpx *__restrict__ new_hist) {
    const int s = blockIdx.x, i = threadIdx.x;
    if (i >= RS_SINC_PERIODS) return;
    const long long v = (long long)n + i;   // index into hist ++ in
    new_hist[(long long)s * RS_SINC_PERIODS + i] = v < RS_SINC_PERIODS ? hist[(long long)s * RS_SINC_PERIODS + v]
                                                                       : in[(long long)s * in_stride + v - RS_SINC_PERIODS];
}

struct pebble_resampler {
    int max_input = 0, streams = 1;
    double float_time = 0.0;           // m_FloatTime
    double *d_sinc = nullptr;
    cpx *d_hist[2] = {nullptr, nullptr};
    int hcur = 0;
    RsFrame *d_frames = nullptr; size_t frames_cap = 0;
    std::vector<RsFrame> frames;
    cudaStream_t stream = nullptr;
    pebble::Staging stg;
};

static int rs_alloc_hist(pebble_resampler *h) {
    for (int b = 0; b < 2; ++b) {
        cudaFree(h->d_hist[b]); h->d_hist[b] = nullptr;
        const size_t hb = sizeof(cpx) * RS_SINC_PERIODS * (size_t)h->streams;
        PB_CUDA(cudaMalloc(&h->d_hist[b], hb), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMemset(h->d_hist[b], 0, hb), PEBBLE_ERR_CUDA);   // Init zeroes the work buffer (fractresampler.cpp:98-102)
    }
    h->hcur = 0;
    return 0;
}

// Plans the frames of one call on the host: per-frame start time, output offset and count; returns the total output count
// and (commit) advances m_FloatTime.
static long long rs_plan(pebble_resampler *h, int n, int frame, double dt, bool commit) {
    if (frame <= 0 || frame > n) frame = n;
    const int nframes = (n + frame - 1) / frame;
    h->frames.resize(nframes);
    double t = h->float_time;
    long long off = 0;
    RsSeg scratch[1];
    for (int f = 0; f < nframes; ++f) {
        const int len = f == nframes - 1 ? n - f * frame : frame;
        int cnt; double te;
        rs_segments(t, dt, len, scratch, 0, &cnt, &te);
        h->frames[f].t0 = t; h->frames[f].out_off = off; h->frames[f].count = cnt; h->frames[f].pad = 0;
        off += cnt;
        t = te;
    }
    if (commit) h->float_time = t;
    return off;
}

template <int MODE>
static int rs_run(pebble_resampler *h, int n, int frame, double rate, const cpx *d_in, long long in_stride, void *d_out,
                  long long out_stride, double gain, cudaStream_t st) {
    if (frame <= 0 || frame > n) frame = n;
    const long long total = rs_plan(h, n, frame, rate, true);
    const int nframes = (int)h->frames.size();
    if ((size_t)nframes > h->frames_cap) {
        PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
        cudaFree(h->d_frames); h->d_frames = nullptr; h->frames_cap = 0;
        PB_CUDA(cudaMalloc(&h->d_frames, sizeof(RsFrame) * (size_t)nframes), PEBBLE_ERR_CUDA);
        h->frames_cap = (size_t)nframes;
    }
    PB_CUDA(cudaMemcpyAsync(h->d_frames, h->frames.data(), sizeof(RsFrame) * (size_t)nframes, cudaMemcpyHostToDevice, st), PEBBLE_ERR_CUDA);
    rs_kernel<MODE><<<dim3((unsigned)nframes, (unsigned)h->streams), 256, 0, st>>>(d_in, in_stride, h->d_hist[h->hcur], frame,
                                                                                 n - (nframes - 1) * frame, nframes, h->d_frames, rate,
                                                                                 h->d_sinc, d_out, out_stride, gain);
    rs_hist_kernel<<<h->streams, 32, 0, st>>>(d_in, in_stride, n, h->d_hist[h->hcur], h->d_hist[h->hcur ^ 1]);
    pebble::count_launch(2);
    PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    h->hcur ^= 1;
    return (int)total;
}

// entry points used by receiver.cu
int resampler_run_device(pebble_resampler *h, int n, int frame, double rate, const cpx *d_in, cpx *d_out, bool clip, cudaStream_t st) {
    return clip ? rs_run<1>(h, n, frame, rate, d_in, 0, d_out, 0, 0.0, st) : rs_run<0>(h, n, frame, rate, d_in, 0, d_out, 0, 0.0, st);
}
long long resampler_peek_count(pebble_resampler *h, int n, int frame, double rate) { return rs_plan(h, n, frame, rate, false); }

extern "C" {

pebble_resampler_t *pebble_resampler_create(int max_input_size) {
    pebble_resampler *h = new pebble_resampler();
    h->max_input = max_input_size;
    // CFractResampler::Init (fractresampler.cpp:87-121): Blackman-Harris windowed sinc, 10000 points per zero crossing
    std::vector<double> sinc(RS_SINC_LENGTH);
    const double TWOPI = 6.28318530717958647692528676656, PI = 3.14159265358979323846264338328;
    for (int i = 0; i < RS_SINC_LENGTH; i++) {
        const double window = (0.35875 - 0.48829 * cos((TWOPI * i) / (RS_SINC_LENGTH - 1)) + 0.14128 * cos((2.0 * TWOPI * i) / (RS_SINC_LENGTH - 1)) -
                               0.01168 * cos((3.0 * TWOPI * i) / (RS_SINC_LENGTH - 1)));
        const double fi = PI * (double)(i - RS_SINC_LENGTH / 2) / (double)RS_SINC_PERIOD_PTS;
        sinc[i] = i != RS_SINC_LENGTH / 2 ? window * (double)sin((double)fi) / (double)fi : 1.0;
    }
    if (!pebble::cuda_ok(cudaMalloc(&h->d_sinc, sizeof(double) * RS_SINC_LENGTH), "cudaMalloc(sinc)") ||
        !pebble::cuda_ok(cudaMemcpy(h->d_sinc, sinc.data(), sizeof(double) * RS_SINC_LENGTH, cudaMemcpyHostToDevice), "cudaMemcpy(sinc)") ||
        rs_alloc_hist(h) < 0 || !pebble::cuda_ok(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "cudaStreamCreate")) {
        pebble_resampler_destroy(h);
        return nullptr;
    }
    return h;
}
void pebble_resampler_destroy(pebble_resampler_t *h) {
    if (!h) return;
    cudaFree(h->d_sinc); cudaFree(h->d_hist[0]); cudaFree(h->d_hist[1]); cudaFree(h->d_frames);
    if (h->stream) cudaStreamDestroy(h->stream);
    h->stg.release();
    delete h;
}
int pebble_resampler_set_streams(pebble_resampler_t *h, int streams) {
    if (!h || streams < 1) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    h->streams = streams;
    h->float_time = 0.0;
    return rs_alloc_hist(h);
}
int pebble_resampler_out_count(pebble_resampler_t *h, int n, int frame, double rate) {
    if (!h || n < 0 || !(rate > 0.0)) return PEBBLE_ERR_ARG;
    if (n == 0) return 0;
    return (int)rs_plan(h, n, frame, rate, false);
}
int pebble_resampler_resample_device(pebble_resampler_t *h, int n, int frame, double rate, const double *d_in, long long in_stride,
                                     double *d_out, long long out_stride, int clip, void *stream) {
    if (!h || n < 0 || !(rate > 0.0) || (n && (!d_in || !d_out))) { pebble::set_error("resampler: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    return clip ? rs_run<1>(h, n, frame, rate, (const cpx *)d_in, in_stride, d_out, out_stride, 0.0, (cudaStream_t)stream)
                : rs_run<0>(h, n, frame, rate, (const cpx *)d_in, in_stride, d_out, out_stride, 0.0, (cudaStream_t)stream);
}
int pebble_resampler_resample_stereo16_device(pebble_resampler_t *h, int n, int frame, double rate, const double *d_in, long long in_stride,
                                              short *d_out, long long out_stride, double gain, void *stream) {
    if (!h || n < 0 || !(rate > 0.0) || (n && (!d_in || !d_out))) { pebble::set_error("resampler: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    return rs_run<2>(h, n, frame, rate, (const cpx *)d_in, in_stride, d_out, out_stride, gain, (cudaStream_t)stream);
}
// host pointers, one stream, one reference call: CFractResampler::Resample(InLength, Rate, pInBuf, pOutBuf)
int pebble_resampler_resample(pebble_resampler_t *h, int n, double rate, const double *in, double *out) {
    if (!h || n < 0 || !(rate > 0.0) || (n && (!in || !out))) { pebble::set_error("resampler: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    if (h->streams != 1) { pebble::set_error("resampler: host call needs streams == 1"); return PEBBLE_ERR_STATE; }
    const long long cnt = rs_plan(h, n, n, rate, false);
    const size_t bi = sizeof(cpx) * (size_t)n, bo = sizeof(cpx) * (size_t)(cnt > 0 ? cnt : 1);
    if (!h->stg.reserve(bi, bo)) return PEBBLE_ERR_CUDA;
    memcpy(h->stg.h_in, in, bi);
    PB_CUDA(cudaMemcpyAsync(h->stg.d_in, h->stg.h_in, bi, cudaMemcpyHostToDevice, h->stream), PEBBLE_ERR_CUDA);
    const int r = rs_run<0>(h, n, n, rate, (const cpx *)h->stg.d_in, 0, h->stg.d_out, 0, 0.0, h->stream);
    if (r < 0) return r;
    if (r > 0) PB_CUDA(cudaMemcpyAsync(h->stg.h_out, h->stg.d_out, sizeof(cpx) * (size_t)r, cudaMemcpyDeviceToHost, h->stream), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamSynchronize(h->stream), PEBBLE_ERR_CUDA);
    if (r > 0) memcpy(out, h->stg.h_out, sizeof(cpx) * (size_t)r);
    return r;
}
// carried state: double m_FloatTime, then 28 CPX per stream
int64_t pebble_resampler_state_size(const pebble_resampler_t *h) { return h ? 8 + (int64_t)sizeof(cpx) * RS_SINC_PERIODS * h->streams : PEBBLE_ERR_ARG; }
int pebble_resampler_get_state(pebble_resampler_t *h, void *buf, int64_t bytes) {
    if (!h || !buf || bytes < pebble_resampler_state_size(h)) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    memcpy(buf, &h->float_time, 8);
    PB_CUDA(cudaMemcpy((char *)buf + 8, h->d_hist[h->hcur], sizeof(cpx) * RS_SINC_PERIODS * h->streams, cudaMemcpyDeviceToHost), PEBBLE_ERR_CUDA);
    return 0;
}
int pebble_resampler_set_state(pebble_resampler_t *h, const void *buf, int64_t bytes) {
    if (!h || !buf || bytes < pebble_resampler_state_size(h)) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    memcpy(&h->float_time, buf, 8);
    PB_CUDA(cudaMemcpy(h->d_hist[h->hcur], (const char *)buf + 8, sizeof(cpx) * RS_SINC_PERIODS * h->streams, cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    return 0;
}
// Host-only (no device): output count and final m_FloatTime of one Resample call that starts at float_time (the planner the
// CPU tests pin against the reference's loop).
int pebble_resampler_plan(double float_time, double rate, int n, double *float_time_out) {
    if (n < 0 || !(rate > 0.0) || float_time < 0.0) return PEBBLE_ERR_ARG;
    RsSeg scratch[1];
    int cnt = 0; double te = float_time;
    if (n > 0) rs_segments(float_time, rate, n, scratch, 0, &cnt, &te);
    if (float_time_out) *float_time_out = te;
    return cnt;
}
// Host-only: the output times of that call, for the same purpose. Returns the count (writes at most max times).
int pebble_resampler_plan_times(double float_time, double rate, int n, double *times, int max) {
    if (n <= 0 || !(rate > 0.0) || float_time < 0.0) return PEBBLE_ERR_ARG;
    std::vector<RsSeg> segs(4096);
    int cnt = 0; double te;
    const int ns = rs_segments(float_time, rate, n, segs.data(), (int)segs.size(), &cnt, &te);
    if (ns > (int)segs.size()) return PEBBLE_ERR_STATE;
    int w = 0;
    for (int s = 0; s < ns; ++s)
        for (int k = 0; k < segs[s].count; ++k, ++w)
            if (times && w < max) times[w] = fma((double)k, segs[s].step, segs[s].t0);
    return cnt;
}

}  // extern "C"
