// demod.cu — the demodulator tail of the receiver chain on the GPU + C ABI:
//   CFir (pebblelib/fir.cpp:106-132), AGC (application/agc.cpp:84-207), Demod_AM::processBlockFiltered
//   (demod/demod_am.cpp:40-64), Demod_NFM::processBlockNCO / processBlockFM2 (demod/demod_nfm.cpp:124-140,
//   225-257), Demod::processBlock dispatch (demod.cpp:100-141: USB/LSB/... return the input untouched).
// Parallel where the math allows it (FIR, envelope, phase difference; the AM DC-block IIR is a linear
// recurrence and runs as a 3-kernel affine scan); the PLL and the full AGC are true nonlinear
// recurrences and run one thread per stream (parallel only across streams/channels, finding F10).
#include "lib_common.h"
#include "common.cuh"
#include "host_design.h"
#include "wfm_coefs.h"
#include <cmath>
#include <cstring>
#include <vector>

#include "demod_kernels.cuh"

// =====================================================================================================
// host side
// =====================================================================================================
#include "demod_internal.h"

static int grid_for(int n, int threads = 256) {
    int g = (n + threads - 1) / threads;
    const int cap = pebble::num_sms() * 8;
    return g < 1 ? 1 : (g > cap ? cap : g);
}

int cfir_alloc(pebble_cfir *f) {
    PB_CUDA(cudaMalloc(&f->d_hI, sizeof(double) * PB_FIR_MAX), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMalloc(&f->d_hQ, sizeof(double) * PB_FIR_MAX), PEBBLE_ERR_CUDA);
    for (int b = 0; b < 2; ++b) {
        PB_CUDA(cudaMalloc(&f->d_hist[b], sizeof(cpx) * PB_FIR_MAX), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMemset(f->d_hist[b], 0, sizeof(cpx) * PB_FIR_MAX), PEBBLE_ERR_CUDA);
    }
    PB_CUDA(cudaStreamCreateWithFlags(&f->stream, cudaStreamNonBlocking), PEBBLE_ERR_CUDA);
    return 0;
}
void cfir_release(pebble_cfir *f) {
    cudaFree(f->d_hI); cudaFree(f->d_hQ); cudaFree(f->d_hist[0]); cudaFree(f->d_hist[1]);
    if (f->stream) cudaStreamDestroy(f->stream);
    f->stg.release();
}
int cfir_upload(pebble_cfir *f) {
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(f->d_hI, f->hI, sizeof(double) * f->ntaps, cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(f->d_hQ, f->hQ, sizeof(double) * f->ntaps, cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    // Init*Filter zeroes the delay line and m_State (fir.cpp:305-311)
    PB_CUDA(cudaMemset(f->d_hist[0], 0, sizeof(cpx) * PB_FIR_MAX), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemset(f->d_hist[1], 0, sizeof(cpx) * PB_FIR_MAX), PEBBLE_ERR_CUDA);
    return 0;
}
int cfir_run_device(pebble_cfir *f, int n, const cpx *d_in, cpx *d_out, cudaStream_t st, bool dup_input) {
    if (n <= 0) return 0;
    if (f->ntaps <= 0) { pebble::set_error("cfir: filter not initialised"); return PEBBLE_ERR_STATE; }
    if (d_in == d_out) { pebble::set_error("cfir: in-place device call not supported"); return PEBBLE_ERR_ARG; }
    FirTaps taps;
    memset(&taps, 0, sizeof(taps));
    memcpy(taps.hI, f->hI, sizeof(double) * f->ntaps);
    memcpy(taps.hQ, f->hQ, sizeof(double) * f->ntaps);
    if (dup_input && memcmp(f->hI, f->hQ, sizeof(double) * f->ntaps) == 0)
        cfir_dup_kernel<<<grid_for(n, PB_FIR_TILE), PB_FIR_THREADS, 0, st>>>(f->d_hist[f->cur], d_in, n, taps, f->ntaps, d_out,
                                                                            f->d_hist[f->cur ^ 1]);
    else
        cfir_kernel<<<grid_for(n, PB_FIR_TILE), PB_FIR_THREADS, 0, st>>>(f->d_hist[f->cur], d_in, n, taps, f->ntaps, d_out,
                                                                        f->d_hist[f->cur ^ 1]);
    pebble::count_launch();
    PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    f->cur ^= 1;
    return n;
}

// generic host-pointer wrapper: stage in, run fn(d_in, d_out, stream), stage out
template <class Fn>
static int host_roundtrip(pebble::Staging &stg, cudaStream_t st, int n, const double *in, double *out, Fn fn) {
    const size_t nb = sizeof(cpx) * (size_t)n;
    if (!stg.reserve(nb, nb)) return PEBBLE_ERR_CUDA;
    memcpy(stg.h_in, in, nb);
    PB_CUDA(cudaMemcpyAsync(stg.d_in, stg.h_in, nb, cudaMemcpyHostToDevice, st), PEBBLE_ERR_CUDA);
    int r = fn((const cpx *)stg.d_in, (cpx *)stg.d_out, st);
    if (r < 0) return r;
    PB_CUDA(cudaMemcpyAsync(stg.h_out, stg.d_out, nb, cudaMemcpyDeviceToHost, st), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamSynchronize(st), PEBBLE_ERR_CUDA);
    memcpy(out, stg.h_out, nb);
    return r;
}

extern "C" {
pebble_cfir_t *pebble_cfir_create(void) {
    pebble_cfir *f = new pebble_cfir();
    if (cfir_alloc(f) < 0) { cfir_release(f); delete f; return nullptr; }
    return f;
}
void pebble_cfir_destroy(pebble_cfir_t *f) { if (f) { cfir_release(f); delete f; } }
int pebble_cfir_init_lp(pebble_cfir_t *f, int num_taps, double scale, double astop, double fpass, double fstop, double fs) {
    if (!f || num_taps < 0 || num_taps > PB_FIR_MAX) return PEBBLE_ERR_ARG;
    f->fs = fs;
    f->ntaps = pebble_host::cfir_design_lp(num_taps, scale, astop, fpass, fstop, fs, f->hI);
    memcpy(f->hQ, f->hI, sizeof(double) * f->ntaps);  // fir.cpp:298-302
    int r = cfir_upload(f);
    return r < 0 ? r : f->ntaps;
}
int pebble_cfir_init_hp(pebble_cfir_t *f, int num_taps, double scale, double astop, double fpass, double fstop, double fs) {
    if (!f || num_taps < 0 || num_taps > PB_FIR_MAX) return PEBBLE_ERR_ARG;
    f->fs = fs;
    f->ntaps = pebble_host::cfir_design_hp(num_taps, scale, astop, fpass, fstop, fs, f->hI);
    memcpy(f->hQ, f->hI, sizeof(double) * f->ntaps);
    int r = cfir_upload(f);
    return r < 0 ? r : f->ntaps;
}
int pebble_cfir_init_const(pebble_cfir_t *f, int num_taps, const double *icoef, const double *qcoef, double fs) {
    if (!f || num_taps <= 0 || num_taps > PB_FIR_MAX || !icoef || !qcoef) return PEBBLE_ERR_ARG;  // fir.cpp:203-228
    f->fs = fs; f->ntaps = num_taps;
    memcpy(f->hI, icoef, sizeof(double) * num_taps);
    memcpy(f->hQ, qcoef, sizeof(double) * num_taps);
    int r = cfir_upload(f);
    return r < 0 ? r : f->ntaps;
}
int pebble_cfir_get_taps(pebble_cfir_t *f, double *out, int max) {
    if (!f || !out) return PEBBLE_ERR_ARG;
    for (int i = 0; i < f->ntaps && i < max; ++i) out[i] = f->hI[i];
    return f->ntaps;
}
int pebble_cfir_process_device(pebble_cfir_t *f, int n, const double *d_in, double *d_out, void *stream) {
    if (!f || n < 0 || (n && (!d_in || !d_out))) return PEBBLE_ERR_ARG;
    return cfir_run_device(f, n, (const cpx *)d_in, (cpx *)d_out, (cudaStream_t)stream);
}
int pebble_cfir_process(pebble_cfir_t *f, int n, const double *in, double *out) {
    if (!f || n < 0 || (n && (!in || !out))) return PEBBLE_ERR_ARG;
    if (n == 0) return 0;
    return host_roundtrip(f->stg, f->stream, n, in, out,
                          [&](const cpx *di, cpx *dout, cudaStream_t st) { return cfir_run_device(f, n, di, dout, st); });
}
}  // extern "C"

// ---- AGC handle (application/agc.cpp) ---------------------------------------------------------------------
struct pebble_agc {
    double sample_rate = 0, params_rate = 100.0;  // m_sampleRate = 100.0 triggers init on first setup (agc.cpp:26)
    int buffer_size = 0;
    int mode = 0;  // AGC_OFF
    int ui_threshold = 0;
    bool use_hang = false; int threshold = 0, slope = 0, decay = 0;  // agc.cpp:20-24
    double manual_gain = 1.0;
    pebble_host::AgcParams p;
    AgcState *d_state = nullptr; cpx *d_sig = nullptr; double *d_mag = nullptr;
    cudaStream_t stream = nullptr;
    pebble::Staging stg;
};
double agc_manual_gain(const pebble_agc *h) { return h->manual_gain; }
int agc_is_off(const pebble_agc *h) { return h->mode == 0; }

static int agc_reset_buffers(pebble_agc *h) {
    // agc.cpp:262-276
    std::vector<double> mag(PB_AGC_BUF, -16.0);
    AgcState s; s.sig_ptr = 0; s.mag_pos = 0; s.hang_timer = 0; s.pad = 0; s.peak = -16.0; s.attack_avg = -5.0; s.decay_avg = -5.0;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemset(h->d_sig, 0, sizeof(cpx) * PB_AGC_BUF), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(h->d_mag, mag.data(), sizeof(double) * PB_AGC_BUF, cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(h->d_state, &s, sizeof(s), cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    return 0;
}
// AGC::setParameters, agc.cpp:237-309
static int agc_set_parameters(pebble_agc *h, bool use_hang, int threshold, int slope, int decay) {
    if (h->mode == 0) {
        h->manual_gain = pow(10, (threshold / 5) / 20.0);  // DB::dBToAmplitude(threshold / 5): integer division
        return 0;
    }
    threshold = -threshold;
    h->manual_gain = 1;
    if (use_hang == h->use_hang && threshold == h->threshold && slope == h->slope && decay == h->decay) return 0;
    h->use_hang = use_hang; h->threshold = threshold; h->slope = slope; h->decay = decay;
    if (h->params_rate != h->sample_rate) {
        h->params_rate = h->sample_rate;
        int r = agc_reset_buffers(h);
        if (r < 0) return r;
    }
    pebble_host::agc_params(h->params_rate, use_hang, threshold, slope, decay, h->p);
    return 0;
}
int agc_run_device(pebble_agc *h, int n, const cpx *d_in, cpx *d_out, cudaStream_t st) {
    if (n <= 0) return 0;
    if (h->mode == 0) {
        scale_kernel<<<grid_for(n), 256, 0, st>>>(d_in, n, h->manual_gain, d_out);
    } else {
        agc_kernel<<<1, 32, 0, st>>>(d_in, n, 0, 1, h->p, h->d_state, h->d_sig, h->d_mag, d_out);
    }
    pebble::count_launch();
    PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    return n;
}

extern "C" {
pebble_agc_t *pebble_agc_create(unsigned sample_rate, unsigned buffer_size) {
    pebble_agc *h = new pebble_agc();
    h->sample_rate = sample_rate; h->buffer_size = (int)buffer_size;
    bool ok = pebble::cuda_ok(cudaMalloc(&h->d_state, sizeof(AgcState)), "cudaMalloc") &&
              pebble::cuda_ok(cudaMalloc(&h->d_sig, sizeof(cpx) * PB_AGC_BUF), "cudaMalloc") &&
              pebble::cuda_ok(cudaMalloc(&h->d_mag, sizeof(double) * PB_AGC_BUF), "cudaMalloc") &&
              pebble::cuda_ok(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "cudaStreamCreate");
    if (!ok) { pebble_agc_destroy(h); return nullptr; }
    pebble_agc_set_mode(h, 0, 1);  // setAgcMode(AGC_OFF, 1), agc.cpp:29
    return h;
}
void pebble_agc_destroy(pebble_agc_t *h) {
    if (!h) return;
    cudaFree(h->d_state); cudaFree(h->d_sig); cudaFree(h->d_mag);
    if (h->stream) cudaStreamDestroy(h->stream);
    h->stg.release();
    delete h;
}
// AGC::setAgcMode, agc.cpp:52-82. mode: 0 OFF, 1 FAST, 2 MED, 3 SLOW, 4 LONG (agc.h:11)
int pebble_agc_set_mode(pebble_agc_t *h, int mode, int threshold) {
    if (!h || mode < 0 || mode > 4) return PEBBLE_ERR_ARG;
    int decay = 200;
    h->mode = mode;
    if (mode == 1) decay = 100;
    else if (mode == 3) decay = 500;
    else if (mode == 2) decay = 250;
    else if (mode == 4) decay = 2000;
    return agc_set_parameters(h, false, threshold, 0, decay);
}
// AGC::setAgcThreshold, agc.cpp:43-50
int pebble_agc_set_threshold(pebble_agc_t *h, int g) {
    if (!h) return PEBBLE_ERR_ARG;
    h->ui_threshold = g;
    return agc_set_parameters(h, h->use_hang, g, h->slope, h->decay);
}
int pebble_agc_process_device(pebble_agc_t *h, int n, const double *d_in, double *d_out, void *stream) {
    if (!h || n < 0 || (n && (!d_in || !d_out))) return PEBBLE_ERR_ARG;
    return agc_run_device(h, n, (const cpx *)d_in, (cpx *)d_out, (cudaStream_t)stream);
}
int pebble_agc_process(pebble_agc_t *h, int n, const double *in, double *out) {
    if (!h || n < 0 || (n && (!in || !out))) return PEBBLE_ERR_ARG;
    if (n == 0) return 0;
    return host_roundtrip(h->stg, h->stream, n, in, out,
                          [&](const cpx *di, cpx *dout, cudaStream_t st) { return agc_run_device(h, n, di, dout, st); });
}
}  // extern "C"

// ---- Demod handle (application/demod.cpp:100-141 dispatch + demod_am.cpp + demod_nfm.cpp) --------------------
// WFM stereo pilot PLL (demod_wfm.cpp:369-430): constants of initPilotPll and the carried state (double, as there)
struct WfmPllConst { double alpha, beta, lo, hi, err_alpha, phase_adjust; };
struct WfmPllState { double phase, freq, err_ave; int locked; int pad; };
struct pebble_demod {
    double rate = 0;
    int mode = 0;                 // DeviceInterface::DemodMode numbering (device_interfaces.h:124-137); dmAM at create
    int nfm_variant = 0;          // 0 = processBlockNCO (live, demod.cpp:113), 1 = processBlockFM2
    // AM
    pebble_cfir am_lp;
    double *d_am_last[2] = {nullptr, nullptr};
    int am_cur = 0;
    bool am_pre_tiles = false;   // the next AM call finds its tile maps already in d_tiles (demod_am_expect_tiles)
    // NFM
    pebble_cfir nfm_lp;
    PllConst pll;
    PllState *d_pll = nullptr;
    cpx *d_fm_last[2] = {nullptr, nullptr};
    int fm_cur = 0;
    // SAM (demod_sam.cpp)
    pebble_cfir sam_bp;
    SamConst sam;
    SamState *d_sam = nullptr;
    // WFM mono (demod_wfm.cpp:207-231): input LP (>= 150 kS/s only), discriminator, audio LP, de-emphasis, pilot notch
    double wfm_rate = 0;
    pebble_iir_t *wfm_in_lp = nullptr, *wfm_deemph = nullptr, *wfm_notch = nullptr;
    pebble_cfir wfm_lp;
    cpx *d_wfm_last[2] = {nullptr, nullptr};
    int wfm_cur = 0;
    // WFM stereo (demod_wfm.cpp:258-366): Hilbert pair, pilot band-pass, pilot PLL, demux; the RDS branch lives in rds.cu
    RdsBranch *rds = nullptr;
    bool rds_on = true, rds_ok = false;   // rds_ok: the RDS chain could be designed for the current WFM rate
    pebble_cfir wfm_hilb;
    pebble_iir_t *wfm_pilot_bp = nullptr;
    WfmPllConst wfm_pll;
    WfmPllState *d_wfm_pll = nullptr;          // [2]: ping-pong; .locked of the current one = m_PilotLocked
    int wfm_pll_cur = 0;
    cpx *d_w1 = nullptr, *d_w2 = nullptr; double *d_wphase = nullptr; size_t w_cap = 0;
    // scratch (grow-only)
    cpx *d_tmp = nullptr; size_t tmp_cap = 0;
    Aff *d_tiles = nullptr; double *d_carry = nullptr; int tiles_cap = 0;
    cudaStream_t stream = nullptr;
    pebble::Staging stg;
};
// Demod_WFM::processDataMono's discriminator (demod_wfm.cpp:214-220): FMDEMOD_GAIN * atan2(D1.re D0.im - D0.re D1.im,
// D1.re D0.re + D1.im D0.im), D1 = the previous sample (carried over calls); products and sums separately rounded like the
// reference's build. Output (v, v).
static __global__ void wfm_disc_kernel(const cpx *__restrict__ in, int n, const cpx *__restrict__ last_in, cpx *__restrict__ out,
                                       cpx *__restrict__ last_out) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const cpx d0 = in[i];
        const cpx d1 = i ? in[i - 1] : *last_in;
        const double im = __dsub_rn(__dmul_rn(d1.x, d0.y), __dmul_rn(d0.x, d1.y));
        const double re = __dadd_rn(__dmul_rn(d1.x, d0.x), __dmul_rn(d1.y, d0.y));
        const double v = __dmul_rn(0.25, atan2(im, re));
        out[i] = cpx_make(v, v);
        if (i == n - 1) *last_out = d0;
    }
}
// ---- WFM stereo: pilot PLL (demod_wfm.cpp:388-430) and demux (:272-296) ---------------------------------------------------------
// One warp; the loop is a serial recurrence in double (sin/cos of the NCO phase, rotate, arctan2, loop filter), the lanes only
// fetch the samples of a 32-sample block. The reference takes sin/cos from x87 fsincos (:398); the double results agree to ~1 ulp
// and the loop is a contracting map, so the audio agrees to ~1e-12.
static __global__ void __launch_bounds__(32) wfm_pilot_pll_kernel(const cpx *__restrict__ in, int n, WfmPllConst c, const WfmPllState *__restrict__ st_in,
                                                                  WfmPllState *__restrict__ st_out, double *__restrict__ pilot_phase) {
    const int lane = threadIdx.x;
    WfmPllState v = *st_in;
    for (int i0 = 0; i0 < n; i0 += 32) {
        const int idx = i0 + lane;
        const cpx x = idx < n ? in[idx] : cpx_make(0.0, 0.0);
        const int cnt = n - i0 < 32 ? n - i0 : 32;
        double mine = 0.0;
        for (int j = 0; j < cnt; ++j) {
            const double xr = __shfl_sync(0xffffffffu, x.x, j), xi = __shfl_sync(0xffffffffu, x.y, j);
            double sn, cs;
            sincos(v.phase, &sn, &cs);
            const double tr = __dsub_rn(__dmul_rn(cs, xr), __dmul_rn(sn, xi));                    // :406-407
            const double ti = __dadd_rn(__dmul_rn(cs, xi), __dmul_rn(sn, xr));
            const double phzerror = -wfm_arctan2(ti, tr);                                         // :409
            v.freq = __dadd_rn(v.freq, __dmul_rn(c.beta, phzerror));                              // :412
            if (v.freq > c.hi) v.freq = c.hi; else if (v.freq < c.lo) v.freq = c.lo;              // :414-417
            v.phase = __dadd_rn(v.phase, __dadd_rn(v.freq, __dmul_rn(c.alpha, phzerror)));        // :419
            const double pp = __dadd_rn(v.phase, c.phase_adjust);                                 // :420
            v.err_ave = __dadd_rn(__dmul_rn(1.0 - c.err_alpha, v.err_ave), __dmul_rn(__dmul_rn(c.err_alpha, phzerror), phzerror));   // :422
            mine = lane == j ? pp : mine;
        }
        if (idx < n) pilot_phase[idx] = mine;
    }
    if (lane == 0) {
        v.phase = fmod(v.phase, 6.28318530717958647692528676656);                                 // :424
        v.locked = v.err_ave < 0.05 ? 1 : 0;                                                      // LOCK_MAG_THRESHOLD, :426-429
        *st_out = v;
    }
}
// raw = the discriminator output (v, v); locked: L-R = 2 v sin(2 pilotPhase), out = (v + L-R, v - L-R); else (v, v)   (:272-296)
static __global__ void wfm_demux_kernel(const cpx *__restrict__ raw, const double *__restrict__ pilot_phase, const WfmPllState *__restrict__ st, int n,
                                        cpx *__restrict__ out) {
    const bool locked = st->locked != 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const double v = raw[i].x;
        if (locked) {
            const double lmr = __dmul_rn(__dmul_rn(2.0, v), sin(__dmul_rn(pilot_phase[i], 2.0)));
            out[i] = cpx_make(__dadd_rn(v, lmr), __dsub_rn(v, lmr));
        } else {
            out[i] = cpx_make(v, v);
        }
    }
}

// Demod_WFM::setSampleRate (demod_wfm.cpp:155-196) for the mono path; out rate = sample rate (the constructor's init, :110-114)
static int wfm_setup(pebble_demod *h, double rate) {
    h->wfm_rate = rate;
    if (pebble_iir_init(h->wfm_in_lp, PEBBLE_IIR_LP, 75000.0, 1.0, rate) < 0) return PEBBLE_ERR_CUDA;      // m_MonoLPFilter.InitLP(75000, 1.0, rate)
    if (pebble_iir_init(h->wfm_notch, PEBBLE_IIR_BR, 19000.0, 5.0, rate) < 0) return PEBBLE_ERR_CUDA;     // m_NotchFilter.InitBR(19000, 5, rate)
    // de-emphasis, US version 75 us (:183, :452-470): Ave = (1 - a) Ave + a x, out = 2 Ave — a one-pole section written as the
    // biquad w0 = x + (1 - a) w1, y = 2 a w0 (same transfer function; the two orders differ by roundings, ~1e-16)
    const double a = 1.0 - exp(-1.0 / (rate * 75E-6));
    const double co[5] = {-(1.0 - a), 0.0, 2.0 * a, 0.0, 0.0};
    if (pebble_iir_set_coefficients(h->wfm_deemph, co) < 0) return PEBBLE_ERR_CUDA;
    h->wfm_lp.ntaps = pebble_host::cfir_design_lp(0, 1.0, 60.0, 15000.0, 1.4 * 15000.0, rate, h->wfm_lp.hI);   // m_LPFilter, :176
    memcpy(h->wfm_lp.hQ, h->wfm_lp.hI, sizeof(double) * h->wfm_lp.ntaps);
    if (cfir_upload(&h->wfm_lp) < 0) return PEBBLE_ERR_CUDA;
    for (int b = 0; b < 2; ++b) PB_CUDA(cudaMemset(h->d_wfm_last[b], 0, sizeof(cpx)), PEBBLE_ERR_CUDA);
    // stereo: m_HilbertFilter.InitConstFir(61, HILBLP_H, rate); GenerateHBFilter(42000) (:166-168, fir.cpp:454-468)
    {
        const double TWOPI = 6.28318530717958647692528676656;
        const int nt = pebble_wfm::kHilbLen;
        h->wfm_hilb.ntaps = nt;
        for (int k = 0; k < nt; ++k) {
            const double a = (TWOPI * 42000.0 / rate) * ((double)k - ((double)(nt - 1) / 2.0));
            h->wfm_hilb.hI[k] = 2.0 * pebble_wfm::kHilbLp[k] * cos(a);
            h->wfm_hilb.hQ[k] = 2.0 * pebble_wfm::kHilbLp[k] * sin(a);
        }
        if (cfir_upload(&h->wfm_hilb) < 0) return PEBBLE_ERR_CUDA;
        if (pebble_iir_init(h->wfm_pilot_bp, PEBBLE_IIR_BP, 19000.0, 500.0, rate) < 0) return PEBBLE_ERR_CUDA;   // m_PilotBPFilter.InitBP(19000, 500, rate)
        // initPilotPll (:369-384)
        const double norm = TWOPI / rate, f0 = -19000.0;
        h->wfm_pll.lo = (f0 - 20.0) * norm;
        h->wfm_pll.hi = (f0 + 20.0) * norm;
        h->wfm_pll.alpha = 2.0 * .707 * 10.0 * norm;
        h->wfm_pll.beta = (h->wfm_pll.alpha * h->wfm_pll.alpha) / (4.0 * .707 * .707);
        h->wfm_pll.err_alpha = (1.0 - exp(-1.0 / (rate * .5)));
        h->wfm_pll.phase_adjust = -7.267e-6 * rate + 3.677;                                                     // :163
        WfmPllState s0; memset(&s0, 0, sizeof(s0)); s0.freq = f0;                                               // m_PilotNcoFreq = -PILOTPLL_FREQ (Hz, as written)
        for (int b = 0; b < 2; ++b) PB_CUDA(cudaMemcpy(h->d_wfm_pll + b, &s0, sizeof(s0), cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
        h->wfm_pll_cur = 0;
    }
    // m_RdsDownConvert.SetDataRate(rate, 8000), SetFrequency(-57 kHz), initRds (:190-193). A rate the RDS chain cannot be
    // designed for (matched filter longer than CFir's 75 taps) leaves the audio path working and RDS off.
    h->rds_ok = rds_setup(h->rds, rate) == 0;
    return 0;
}

static int demod_reserve(pebble_demod *h, int n) {
    if ((size_t)n > h->tmp_cap) {
        cudaFree(h->d_tmp);
        h->d_tmp = nullptr; h->tmp_cap = 0;
        PB_CUDA(cudaMalloc(&h->d_tmp, sizeof(cpx) * (size_t)n), PEBBLE_ERR_CUDA);
        h->tmp_cap = (size_t)n;
    }
    const int tiles = (n + 1023) / 1024;   // enough for the 1024-sample tiles of the FastFIR-fused path as well
    if (tiles > h->tiles_cap) {
        cudaFree(h->d_tiles); cudaFree(h->d_carry);
        h->d_tiles = nullptr; h->d_carry = nullptr; h->tiles_cap = 0;
        PB_CUDA(cudaMalloc(&h->d_tiles, sizeof(Aff) * (size_t)tiles), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMalloc(&h->d_carry, sizeof(double) * (size_t)tiles), PEBBLE_ERR_CUDA);
        h->tiles_cap = tiles;
    }
    return 0;
}
// Receiver chain, AM with the AGC off: the FastFIR kernel that produces the demodulator's input also reduces every
// 1024-sample block to its affine map (fastfir.cu, ff_am_aggregate). Returns the device array it must fill for a call
// of n samples (n <= 1024 * PB_AM_DIRECT), or nullptr when the demodulator cannot take pre-computed maps.
void *demod_am_expect_tiles(pebble_demod *h, int n, double *alpha) {
    if (h->mode != 0 || n <= 0 || (n + 1023) / 1024 > PB_AM_DIRECT) return nullptr;
    if (demod_reserve(h, n) < 0) return nullptr;
    *alpha = (double)0.9999f;
    h->am_pre_tiles = true;
    return h->d_tiles;
}

void demod_am_cancel_tiles(pebble_demod *h) { h->am_pre_tiles = false; }
int demod_mode(const pebble_demod *h) { return h->mode; }

int demod_run_device(pebble_demod *h, int n, const cpx *d_in, cpx *d_out, cudaStream_t st, int *passthrough, int frame) {
    if (passthrough) *passthrough = 0;
    if (n <= 0) return 0;
    if (h->mode == 0) {  // dmAM -> processBlockFiltered (demod.cpp:105-108)
        int r = demod_reserve(h, n);
        if (r < 0) return r;
        const double alpha = (double)0.9999f;  // DC_ALPHA is a float literal, demod_am.cpp:36
        if (h->am_pre_tiles) {
            // the tile maps (one per 1024-sample FastFIR block) were produced by the FastFIR kernel that wrote d_in
            const int tiles = (n + 1023) / 1024;
            am_apply_kernel<128><<<tiles, 128, 0, st>>>(d_in, n, alpha, nullptr, h->d_tiles, h->d_am_last[h->am_cur], h->d_tmp,
                                                        h->d_am_last[h->am_cur ^ 1]);
            pebble::count_launch();
            h->am_pre_tiles = false;
        } else {
            const int tiles = (n + PB_SCAN_TILE - 1) / PB_SCAN_TILE;
            am_tile_kernel<<<tiles, PB_SCAN_THREADS, 0, st>>>(d_in, n, alpha, h->d_tiles);
            const bool direct = tiles <= PB_AM_DIRECT;
            if (!direct) am_carry_kernel<<<1, PB_SCAN_THREADS, 0, st>>>(h->d_tiles, tiles, h->d_am_last[h->am_cur], h->d_carry);
            am_apply_kernel<PB_SCAN_THREADS><<<tiles, PB_SCAN_THREADS, 0, st>>>(d_in, n, alpha, direct ? nullptr : h->d_carry, h->d_tiles,
                                                                              h->d_am_last[h->am_cur], h->d_tmp, h->d_am_last[h->am_cur ^ 1]);
            pebble::count_launch(direct ? 2 : 3);
        }
        PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
        h->am_cur ^= 1;
        return cfir_run_device(&h->am_lp, n, h->d_tmp, d_out, st, true);  // demod_am.cpp:62 (input is (y, y))
    }
    if (h->mode == 2) {  // dmFMN
        int r = demod_reserve(h, n);
        if (r < 0) return r;
        if (h->nfm_variant == 0) {
            pll_kernel<<<1, 32, 0, st>>>(d_in, n, 0, 1, h->pll, h->d_pll, h->d_tmp, frame);
            pebble::count_launch();
            PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
            return cfir_run_device(&h->nfm_lp, n, h->d_tmp, d_out, st);  // demod_nfm.cpp:255
        }
        fm2_kernel<<<grid_for(n), 256, 0, st>>>(d_in, n, h->d_fm_last[h->fm_cur], d_out, h->d_fm_last[h->fm_cur ^ 1]);
        pebble::count_launch();
        PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
        h->fm_cur ^= 1;
        return n;
    }
    if (h->mode == 1) {  // dmSAM -> Demod_SAM::processBlock (demod.cpp:109-111)
        int r = demod_reserve(h, n);
        if (r < 0) return r;
        sam_kernel<<<1, 32, 0, st>>>(d_in, n, h->sam, (double)0.9999f, h->d_sam, h->d_tmp);
        pebble::count_launch();
        PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
        r = cfir_run_device(&h->sam_bp, n, h->d_tmp, d_out, st);   // Hilbert pair: I taps on re, Q taps on im (demod_sam.cpp:103)
        if (r < 0) return r;
        sam_mix_kernel<<<grid_for(n), 256, 0, st>>>(d_out, n);
        pebble::count_launch();
        PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
        return n;
    }
    if (h->mode == 3) {  // dmFMM -> Demod::fmMono -> Demod_WFM::processDataMono (demod.cpp:168-175, demod_wfm.cpp:207-231)
        int r = demod_reserve(h, n);
        if (r < 0) return r;
        const cpx *x = d_in;
        if (h->wfm_rate >= 150000.0) {   // the 75 kHz IIR needs >= 150 kS/s to be stable (:209-211)
            r = pebble_iir_process_device(h->wfm_in_lp, n, (const double *)d_in, (double *)d_out, st);
            if (r < 0) return r;
            x = d_out;
        }
        wfm_disc_kernel<<<grid_for(n), 256, 0, st>>>(x, n, h->d_wfm_last[h->wfm_cur], h->d_tmp, h->d_wfm_last[h->wfm_cur ^ 1]);
        pebble::count_launch();
        PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
        h->wfm_cur ^= 1;
        r = cfir_run_device(&h->wfm_lp, n, h->d_tmp, d_out, st, true);                                   // roll off above 15 kHz; input is (v, v)
        if (r < 0) return r;
        r = pebble_iir_process_device(h->wfm_deemph, n, (const double *)d_out, (double *)h->d_tmp, st);  // 75 us de-emphasis
        if (r < 0) return r;
        return pebble_iir_process_device(h->wfm_notch, n, (const double *)h->d_tmp, (double *)d_out, st);   // 19 kHz pilot notch
    }
    if (h->mode == 4) {  // dmFMS -> Demod::fmStereo -> Demod_WFM::processDataStereo (demod_wfm.cpp:258-366)
        int r = demod_reserve(h, n);
        if (r < 0) return r;
        if ((size_t)n > h->w_cap) {
            cudaFree(h->d_w1); cudaFree(h->d_w2); cudaFree(h->d_wphase);
            h->d_w1 = h->d_w2 = nullptr; h->d_wphase = nullptr; h->w_cap = 0;
            PB_CUDA(cudaMalloc(&h->d_w1, sizeof(cpx) * (size_t)n), PEBBLE_ERR_CUDA);
            PB_CUDA(cudaMalloc(&h->d_w2, sizeof(cpx) * (size_t)n), PEBBLE_ERR_CUDA);
            PB_CUDA(cudaMalloc(&h->d_wphase, sizeof(double) * (size_t)n), PEBBLE_ERR_CUDA);
            h->w_cap = (size_t)n;
        }
        wfm_disc_kernel<<<grid_for(n), 256, 0, st>>>(d_in, n, h->d_wfm_last[h->wfm_cur], h->d_tmp, h->d_wfm_last[h->wfm_cur ^ 1]);   // m_RawFm as (v, v)
        pebble::count_launch();
        PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
        h->wfm_cur ^= 1;
        r = cfir_run_device(&h->wfm_hilb, n, h->d_tmp, h->d_w1, st);                                      // real -> complex: I taps on v, Q taps on v (:266)
        if (r < 0) return r;
        r = pebble_iir_process_device(h->wfm_pilot_bp, n, (const double *)h->d_w1, (double *)h->d_w2, st);   // :269
        if (r < 0) return r;
        wfm_pilot_pll_kernel<<<1, 32, 0, st>>>(h->d_w2, n, h->wfm_pll, h->d_wfm_pll + h->wfm_pll_cur, h->d_wfm_pll + (h->wfm_pll_cur ^ 1), h->d_wphase);
        h->wfm_pll_cur ^= 1;
        if (h->rds_on && h->rds_ok) {   // :298-357, from m_CpxRawFm (d_w1) before the demux below reuses the buffer
            r = rds_process(h->rds, n, h->d_w1, st);
            if (r < 0) return r;
        }
        wfm_demux_kernel<<<grid_for(n), 256, 0, st>>>(h->d_tmp, h->d_wphase, h->d_wfm_pll + h->wfm_pll_cur, n, h->d_w1);
        pebble::count_launch(2);
        PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
        r = cfir_run_device(&h->wfm_lp, n, h->d_w1, d_out, st);                                           // :361 (complex: left / right)
        if (r < 0) return r;
        r = pebble_iir_process_device(h->wfm_deemph, n, (const double *)d_out, (double *)h->d_w1, st);    // :362
        if (r < 0) return r;
        return pebble_iir_process_device(h->wfm_notch, n, (const double *)h->d_w1, (double *)d_out, st);  // :363
    }
    // USB/LSB/DSB/CW/DIG/NONE: Demod::processBlock returns `in` itself (demod.cpp:127-138)
    if (passthrough) *passthrough = 1;
    if (d_out != d_in) PB_CUDA(cudaMemcpyAsync(d_out, d_in, sizeof(cpx) * (size_t)n, cudaMemcpyDeviceToDevice, st), PEBBLE_ERR_CUDA);
    return n;
}

extern "C" {
pebble_demod_t *pebble_demod_create(unsigned input_rate, unsigned num_samples) {
    (void)num_samples;
    pebble_demod *h = new pebble_demod();
    h->rate = input_rate;
    h->wfm_in_lp = pebble_iir_create(); h->wfm_deemph = pebble_iir_create(); h->wfm_notch = pebble_iir_create(); h->wfm_pilot_bp = pebble_iir_create();
    h->rds = rds_create();
    bool ok = h->rds && h->wfm_in_lp && h->wfm_deemph && h->wfm_notch && h->wfm_pilot_bp && cfir_alloc(&h->wfm_lp) == 0 && cfir_alloc(&h->wfm_hilb) == 0 &&
              pebble::cuda_ok(cudaMalloc(&h->d_wfm_pll, 2 * sizeof(WfmPllState)), "cudaMalloc") &&
              cfir_alloc(&h->am_lp) == 0 && cfir_alloc(&h->nfm_lp) == 0 && cfir_alloc(&h->sam_bp) == 0 &&
              pebble::cuda_ok(cudaMalloc(&h->d_sam, sizeof(SamState)), "cudaMalloc") &&
              pebble::cuda_ok(cudaMemset(h->d_sam, 0, sizeof(SamState)), "cudaMemset") &&
              pebble::cuda_ok(cudaMalloc(&h->d_pll, sizeof(PllState)), "cudaMalloc") &&
              pebble::cuda_ok(cudaMemset(h->d_pll, 0, sizeof(PllState)), "cudaMemset") &&
              pebble::cuda_ok(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "cudaStreamCreate");
    for (int b = 0; ok && b < 2; ++b) {
        ok = pebble::cuda_ok(cudaMalloc(&h->d_am_last[b], sizeof(double)), "cudaMalloc") &&
             pebble::cuda_ok(cudaMemset(h->d_am_last[b], 0, sizeof(double)), "cudaMemset") &&
             pebble::cuda_ok(cudaMalloc(&h->d_fm_last[b], sizeof(cpx)), "cudaMalloc") &&
             pebble::cuda_ok(cudaMemset(h->d_fm_last[b], 0, sizeof(cpx)), "cudaMemset") &&
             pebble::cuda_ok(cudaMalloc(&h->d_wfm_last[b], sizeof(cpx)), "cudaMalloc");
    }
    if (!ok) { pebble_demod_destroy(h); return nullptr; }
    // Demod_AM ctor: setBandwidth(16000) (demod_am.cpp:9)
    pebble_demod_set_mode(h, 0);
    h->am_lp.ntaps = pebble_host::cfir_design_lp(0, 1.0, 50.0, 16000.0, 16000.0 * 1.8, h->rate, h->am_lp.hI);
    memcpy(h->am_lp.hQ, h->am_lp.hI, sizeof(double) * h->am_lp.ntaps);
    // Demod_NFM::init (demod_nfm.cpp:44-66); members are float
    const double norm = 6.28318530717958647692528676656 / h->rate;
    h->pll.lo = (float)(-15000.0 * norm);
    h->pll.hi = (float)(15000.0 * norm);
    h->pll.alpha = (float)(2.0 * .707 * 3000.0 * norm);
    h->pll.beta = (float)((h->pll.alpha * h->pll.alpha) / (4.0 * .707 * .707));
    h->pll.out_gain = 1.0f;
    h->pll.dc_alpha = (float)(1.0 - exp(-1.0 / (h->rate * 0.001)));
    h->nfm_lp.ntaps = pebble_host::cfir_design_lp(0, 1.0, 50.0, 3000.0, 1.6 * 3000.0, h->rate, h->nfm_lp.hI);
    memcpy(h->nfm_lp.hQ, h->nfm_lp.hI, sizeof(double) * h->nfm_lp.ntaps);
    {
        // Demod_SAM::Demod_SAM (demod_sam.cpp:5-34): members are float, pllBandwidth / pllLimit are ints, pllZeta a float
        const float pllZeta = 0.707f;
        const int pllBandwidth = 100, pllLimit = 1000;
        const double TWOPI = 6.28318530717958647692528676656;
        const unsigned sr = (unsigned)h->rate;
        h->sam.alpha = (float)(2.0 * pllZeta * pllBandwidth * TWOPI / sr);
        h->sam.beta = (float)(h->sam.alpha * h->sam.alpha / (4.0 * pllZeta * pllZeta));
        h->sam.lo = (float)(-pllLimit * TWOPI / sr);
        h->sam.hi = (float)(pllLimit * TWOPI / sr);
        // m_bpFilter.InitLPFilter(0, 1.0, 40.0, 4500, 5500, sampleRate); GenerateHBFilter(5000.0) (fir.cpp:454-468)
        double lp[75];
        const int nt = pebble_host::cfir_design_lp(0, 1.0, 40.0, 4500, 5500, (double)sr, lp);
        h->sam_bp.ntaps = nt;
        for (int k = 0; k < nt; ++k) {
            const double a = (TWOPI * 5000.0 / (double)sr) * ((double)k - ((double)(nt - 1) / 2.0));
            h->sam_bp.hI[k] = 2.0 * lp[k] * cos(a);
            h->sam_bp.hQ[k] = 2.0 * lp[k] * sin(a);
        }
    }
    if (cfir_upload(&h->am_lp) < 0 || cfir_upload(&h->nfm_lp) < 0 || cfir_upload(&h->sam_bp) < 0 ||
        wfm_setup(h, h->rate) < 0) {   // Demod_WFM(_inputRate, n): init(_inputRate, _inputRate) (demod_wfm.cpp:105-111)
        pebble_demod_destroy(h);
        return nullptr;
    }
    return h;
}
void pebble_demod_destroy(pebble_demod_t *h) {
    if (!h) return;
    cfir_release(&h->am_lp); cfir_release(&h->nfm_lp); cfir_release(&h->sam_bp); cfir_release(&h->wfm_lp);
    pebble_iir_destroy(h->wfm_in_lp); pebble_iir_destroy(h->wfm_deemph); pebble_iir_destroy(h->wfm_notch); pebble_iir_destroy(h->wfm_pilot_bp);
    rds_destroy(h->rds);
    cfir_release(&h->wfm_hilb); cudaFree(h->d_wfm_pll); cudaFree(h->d_w1); cudaFree(h->d_w2); cudaFree(h->d_wphase);
    cudaFree(h->d_wfm_last[0]); cudaFree(h->d_wfm_last[1]);
    cudaFree(h->d_sam);
    cudaFree(h->d_pll); cudaFree(h->d_tmp); cudaFree(h->d_tiles); cudaFree(h->d_carry);
    for (int b = 0; b < 2; ++b) { cudaFree(h->d_am_last[b]); cudaFree(h->d_fm_last[b]); }
    if (h->stream) cudaStreamDestroy(h->stream);
    h->stg.release();
    delete h;
}
int pebble_demod_set_mode(pebble_demod_t *h, int mode) {  // Demod::setDemodMode, demod.cpp:230-247
    if (!h || mode < 0 || mode > 12) return PEBBLE_ERR_ARG;
    h->mode = mode;
    return 0;
}
// Demod(sampleRate, wfmSampleRate, bufferSize) (demod.cpp:50-66): the WFM object runs at its own input rate
int pebble_demod_set_wfm_sample_rate(pebble_demod_t *h, double rate) {
    if (!h || !(rate > 0)) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    return wfm_setup(h, rate);
}
// Demod_WFM::getStereoLock (demod_wfm.cpp:436-448): the pilot lock decision of the last dmFMS call
int pebble_demod_get_stereo_lock(pebble_demod_t *h) {
    if (!h) return PEBBLE_ERR_ARG;
    WfmPllState s;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(&s, h->d_wfm_pll + h->wfm_pll_cur, sizeof(s), cudaMemcpyDeviceToHost), PEBBLE_ERR_CUDA);
    return s.locked;
}
// The RDS branch of dmFMS (demod_wfm.cpp:298-357) is on by default, as in the reference; off = audio path alone.
int pebble_demod_set_rds(pebble_demod_t *h, int on) {
    if (!h) return PEBBLE_ERR_ARG;
    h->rds_on = on != 0;
    return 0;
}
double pebble_demod_get_rds_rate(const pebble_demod_t *h) { return h && h->rds_ok ? rds_rate(h->rds) : 0.0; }
// Demod_WFM::getNextRdsGroupData (demod_wfm.cpp:764-788)
int pebble_demod_get_next_rds_group(pebble_demod_t *h, unsigned short *blocks4) {
    if (!h || !blocks4) return 0;
    return rds_next_group(h->rds, blocks4);
}
int pebble_demod_get_rds_tap(pebble_demod_t *h, double *out, int max_samples) {
    if (!h || !out || max_samples < 0) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    return rds_tap(h->rds, out, max_samples);
}
int pebble_demod_set_nfm_variant(pebble_demod_t *h, int variant) {
    if (!h || variant < 0 || variant > 1) return PEBBLE_ERR_ARG;
    h->nfm_variant = variant;
    return 0;
}
int pebble_demod_set_bandwidth(pebble_demod_t *h, double bw) {  // Demod::setBandwidth, demod.cpp:218-228: AM only
    if (!h) return PEBBLE_ERR_ARG;
    if (h->mode != 0) return 0;
    h->am_lp.ntaps = pebble_host::cfir_design_lp(0, 1.0, 50.0, bw, bw * 1.8, h->rate, h->am_lp.hI);  // demod_am.cpp:21
    memcpy(h->am_lp.hQ, h->am_lp.hI, sizeof(double) * h->am_lp.ntaps);
    return cfir_upload(&h->am_lp);
}
int pebble_demod_is_passthrough(const pebble_demod_t *h) {
    if (!h) return PEBBLE_ERR_ARG;
    return (h->mode >= 5) ? 1 : 0;
}
int pebble_demod_process_device(pebble_demod_t *h, int n, const double *d_in, double *d_out, void *stream) {
    if (!h || n < 0 || (n && (!d_in || !d_out))) return PEBBLE_ERR_ARG;
    return demod_run_device(h, n, (const cpx *)d_in, (cpx *)d_out, (cudaStream_t)stream, nullptr, 0);
}
int pebble_demod_process(pebble_demod_t *h, int n, const double *in, double *out) {
    if (!h || n < 0 || (n && (!in || !out))) return PEBBLE_ERR_ARG;
    if (n == 0) return 0;
    return host_roundtrip(h->stg, h->stream, n, in, out, [&](const cpx *di, cpx *dout, cudaStream_t st) {
        return demod_run_device(h, n, di, dout, st, nullptr, 0);
    });
}
}  // extern "C"
