// noiseblanker.cu — NoiseBlanker::ProcessBlock / ProcessBlock2 (application/noiseblanker.cpp:47-98), the two blankers that sit between
// IQBalance and the spectrum in Receiver::processIQData (receiver.cpp:820-822). SURVEY 8f-3.
//
// Both keep a running average of the sample magnitude in a FLOAT member updated with double arithmetic,
//     avg = (float)(0.999 * avg + 0.001 * mag),                                   (noiseblanker.cpp:58, :90; members: noiseblanker.h:23-24)
// and compare every sample against threshold x avg. The per-step rounding to float makes the recurrence non-associative: a scan that
// re-associates it lands a few 1e-6 away from the reference's average, and every sample whose magnitude falls inside that sliver of the
// threshold would be blanked on one side and not on the other (a differing DECISION, i.e. an output error of the size of the signal,
// about once per million samples of noise). The walk is therefore serial and exact: one warp per stream, the lanes form the magnitudes
// of 32 samples (coalesced load, independent), then every lane runs the same 32 steps of the recurrence on shuffled magnitudes —
// 4 dependent operations per sample (~20 ns), 5x real time for one 10 MSPS stream. The blankers are off by default in the reference
// (noiseblanker.cpp:7-8); this is a correctness-first row, not a throughput kernel.
//   NB1 (:47-77): a sample above threshold x avg starts a run of nbSpike = 7 zeroed outputs (only when no run is active); otherwise
//        the output is the input delayed by two samples (DelayLine(8, 2): NextDelay(0) is the sample stored two NewSample calls ago,
//        delayline.cpp:16-29).
//   NB2 (:80-98): the output is the running complex average (0.75 / 0.25) where the sample exceeds threshold x avg, else the input.
#include "lib_common.h"
#include "common.cuh"
#include <cstring>

struct Nb1State { float avg; int spike; int pad0, pad1; cpx d1, d2; };   // d1 = the previous call's last sample, d2 = the one before it
struct Nb2State { float avg; float pad0; double pad1; cpx acpx; };

__device__ __forceinline__ float nb_mag(cpx x) {   // magCpx (cpx.h:228-238) into the reference's `float mag`
    return (float)sqrt(__dadd_rn(__dmul_rn(x.x, x.x), __dmul_rn(x.y, x.y)));
}
__device__ __forceinline__ float nb_avg(float avg, float mag) {
    return (float)__dadd_rn(__dmul_rn(0.999, (double)avg), __dmul_rn(0.001, (double)mag));
}

static __global__ void __launch_bounds__(32) nb1_kernel(const cpx *__restrict__ in, int n, double thr, int nspike, const Nb1State *__restrict__ st_in,
                                                        Nb1State *__restrict__ st_out, cpx *__restrict__ out) {
    const int lane = threadIdx.x;
    Nb1State v = *st_in;
    for (int i0 = 0; i0 < n; i0 += 32) {
        const int idx = i0 + lane;
        const cpx x = idx < n ? in[idx] : cpx_make(0.0, 0.0);
        const float mag = nb_mag(x);
        // the sample NextDelay(0) returns for this lane's step: in[idx - 2], reaching into the previous call for idx < 2
        const cpx xm2 = idx >= 2 ? (idx < n ? in[idx - 2] : cpx_make(0.0, 0.0)) : (idx == 1 ? v.d1 : v.d2);
        const int cnt = n - i0 < 32 ? n - i0 : 32;
        bool blank = false;
        for (int j = 0; j < cnt; ++j) {
            const float mj = __shfl_sync(0xffffffffu, mag, j);
            v.avg = nb_avg(v.avg, mj);                                                        // :58
            if (v.spike == 0 && (double)mj > __dmul_rn((double)v.avg, thr)) v.spike = nspike;   // :62-63
            const bool b = v.spike > 0;                                                         // :64-72
            if (b) --v.spike;
            blank = lane == j ? b : blank;
        }
        if (idx < n) out[idx] = blank ? cpx_make(0.0, 0.0) : xm2;
    }
    if (lane == 0) {
        Nb1State o = v;
        o.d1 = n >= 1 ? in[n - 1] : v.d1;
        o.d2 = n >= 2 ? in[n - 2] : (n == 1 ? v.d1 : v.d2);
        *st_out = o;
    }
}

static __global__ void __launch_bounds__(32) nb2_kernel(const cpx *__restrict__ in, int n, double thr, const Nb2State *__restrict__ st_in,
                                                        Nb2State *__restrict__ st_out, cpx *__restrict__ out) {
    const int lane = threadIdx.x;
    Nb2State v = *st_in;
    for (int i0 = 0; i0 < n; i0 += 32) {
        const int idx = i0 + lane;
        const cpx x = idx < n ? in[idx] : cpx_make(0.0, 0.0);
        const float mag = nb_mag(x);
        const int cnt = n - i0 < 32 ? n - i0 : 32;
        cpx mine = x;
        for (int j = 0; j < cnt; ++j) {
            const float mj = __shfl_sync(0xffffffffu, mag, j);
            const double xr = __shfl_sync(0xffffffffu, x.x, j), xi = __shfl_sync(0xffffffffu, x.y, j);
            // nb2AverageCPX = scaleCpx(nb2AverageCPX, 0.75) + scaleCpx(in[i], 0.25)                       :89
            v.acpx = cpx_make(__dadd_rn(__dmul_rn(v.acpx.x, 0.75), __dmul_rn(xr, 0.25)), __dadd_rn(__dmul_rn(v.acpx.y, 0.75), __dmul_rn(xi, 0.25)));
            v.avg = nb_avg(v.avg, mj);                                                                   // :90
            const bool over = (double)mj > __dmul_rn(thr, (double)v.avg);                                // :91
            if (lane == j && over) mine = v.acpx;
        }
        if (idx < n) out[idx] = mine;
    }
    if (lane == 0) *st_out = v;
}

struct pebble_noiseblanker {
    bool nb = false, nb2 = false;
    double thr = 3.3;      // nbThreshold
    int nspike = 7;        // nbSpike
    Nb1State *d_s1[2] = {nullptr, nullptr};
    Nb2State *d_s2[2] = {nullptr, nullptr};
    int c1 = 0, c2 = 0;
    cudaStream_t stream = nullptr;
    pebble::Staging stg;
};

static int nb_reset1(pebble_noiseblanker *h, float avg) {
    Nb1State s; memset(&s, 0, sizeof(s)); s.avg = avg;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(h->d_s1[h->c1], &s, sizeof(s), cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    return 0;
}
static int nb_reset2(pebble_noiseblanker *h, float avg) {
    Nb2State s; memset(&s, 0, sizeof(s)); s.avg = avg;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(h->d_s2[h->c2], &s, sizeof(s), cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    return 0;
}

extern "C" {

pebble_noiseblanker_t *pebble_noiseblanker_create(unsigned sample_rate, unsigned buffer_size) {
    (void)sample_rate; (void)buffer_size;
    pebble_noiseblanker *h = new pebble_noiseblanker();
    bool ok = pebble::cuda_ok(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "cudaStreamCreate");
    for (int b = 0; ok && b < 2; ++b)
        ok = pebble::cuda_ok(cudaMalloc(&h->d_s1[b], sizeof(Nb1State)), "cudaMalloc") && pebble::cuda_ok(cudaMalloc(&h->d_s2[b], sizeof(Nb2State)), "cudaMalloc");
    // NoiseBlanker::NoiseBlanker (noiseblanker.cpp:5-16): both averages start at 1, the delay line at zero
    if (!ok || nb_reset1(h, 1.0f) < 0 || nb_reset2(h, 1.0f) < 0) { pebble_noiseblanker_destroy(h); return nullptr; }
    return h;
}
void pebble_noiseblanker_destroy(pebble_noiseblanker_t *h) {
    if (!h) return;
    for (int b = 0; b < 2; ++b) { cudaFree(h->d_s1[b]); cudaFree(h->d_s2[b]); }
    if (h->stream) cudaStreamDestroy(h->stream);
    h->stg.release();
    delete h;
}
// setNbEnabled (:21-29): enabling restarts the spike count and the average at 0; the delay line keeps its contents
int pebble_noiseblanker_set_nb_enabled(pebble_noiseblanker_t *h, int on) {
    if (!h) return PEBBLE_ERR_ARG;
    if (on) {
        Nb1State s;
        PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMemcpy(&s, h->d_s1[h->c1], sizeof(s), cudaMemcpyDeviceToHost), PEBBLE_ERR_CUDA);
        s.avg = 0.0f; s.spike = 0;
        PB_CUDA(cudaMemcpy(h->d_s1[h->c1], &s, sizeof(s), cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    }
    h->nb = on != 0;
    return 0;
}
// setNb2Enabled (:30-37): enabling restarts the magnitude average and the complex average at 0
int pebble_noiseblanker_set_nb2_enabled(pebble_noiseblanker_t *h, int on) {
    if (!h) return PEBBLE_ERR_ARG;
    if (on && nb_reset2(h, 0.0f) < 0) return PEBBLE_ERR_CUDA;
    h->nb2 = on != 0;
    return 0;
}
int pebble_noiseblanker_enabled(const pebble_noiseblanker_t *h, int which) { return h ? (which == 2 ? h->nb2 : h->nb) : PEBBLE_ERR_ARG; }

// which = 1: ProcessBlock, 2: ProcessBlock2. Disabled: the reference returns `in` itself; here out is a copy (returns n either way).
int pebble_noiseblanker_process_device(pebble_noiseblanker_t *h, int which, int n, const double *d_in, double *d_out, void *stream) {
    if (!h || (which != 1 && which != 2) || n < 0 || (n && (!d_in || !d_out)) || d_in == d_out) {
        pebble::set_error("noiseblanker: bad argument (in-place calls are not supported)");
        return PEBBLE_ERR_ARG;
    }
    if (n == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (!(which == 1 ? h->nb : h->nb2)) {
        PB_CUDA(cudaMemcpyAsync(d_out, d_in, sizeof(cpx) * (size_t)n, cudaMemcpyDeviceToDevice, st), PEBBLE_ERR_CUDA);
        return n;
    }
    if (which == 1) {
        nb1_kernel<<<1, 32, 0, st>>>((const cpx *)d_in, n, h->thr, h->nspike, h->d_s1[h->c1], h->d_s1[h->c1 ^ 1], (cpx *)d_out);
        h->c1 ^= 1;
    } else {
        nb2_kernel<<<1, 32, 0, st>>>((const cpx *)d_in, n, h->thr, h->d_s2[h->c2], h->d_s2[h->c2 ^ 1], (cpx *)d_out);
        h->c2 ^= 1;
    }
    pebble::count_launch();
    PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    return n;
}
int pebble_noiseblanker_process(pebble_noiseblanker_t *h, int which, int n, const double *in, double *out) {
    if (!h || n < 0 || (n && (!in || !out))) { pebble::set_error("noiseblanker: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    const size_t b = sizeof(cpx) * (size_t)n;
    if (!h->stg.reserve(b, b)) return PEBBLE_ERR_CUDA;
    memcpy(h->stg.h_in, in, b);
    PB_CUDA(cudaMemcpyAsync(h->stg.d_in, h->stg.h_in, b, cudaMemcpyHostToDevice, h->stream), PEBBLE_ERR_CUDA);
    const int r = pebble_noiseblanker_process_device(h, which, n, (const double *)h->stg.d_in, (double *)h->stg.d_out, h->stream);
    if (r < 0) return r;
    PB_CUDA(cudaMemcpyAsync(h->stg.h_out, h->stg.d_out, b, cudaMemcpyDeviceToHost, h->stream), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamSynchronize(h->stream), PEBBLE_ERR_CUDA);
    memcpy(out, h->stg.h_out, b);
    return r;
}

}  // extern "C"
