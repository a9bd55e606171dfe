// oracle/ref_shim.cpp — TEST INFRASTRUCTURE ONLY (never linked into the product library).
//
// extern "C" wrappers around the UNMODIFIED reference classes compiled from /root/reference/pebblelib
// (see oracle/build_ref.sh). The resulting oracle/_ref/libpebbleref.so is the real reference, used
//   * by tests/ to pin the C restatement (oracle/pebble_oracle.c) and to generate tests/golden/*,
//   * by bench.py's cpu_baseline / --impl reference leg (kind "reference").
// Nothing here re-implements arithmetic: every function forwards to the reference object.
#include "cpx.h"
// Which wrapper groups a build exports (build_ref.sh): the full library exports all of them; the long-filter build
// (step 4) CFastFIR only; the MAX_DECSTAGES-12 build (step 6) CDownConvert only; the Ooura build (step 7) FFT only.
#if defined(REF_SHIM_FASTFIR_ONLY)
#define REF_HAVE_FASTFIR 1
#elif defined(REF_SHIM_DOWNCONVERT_ONLY)
#define REF_HAVE_DOWNCONVERT 1
#elif defined(REF_SHIM_FFT_ONLY)
#define REF_HAVE_FFT 1
#else
#define REF_HAVE_FASTFIR 1
#define REF_HAVE_DOWNCONVERT 1
#define REF_HAVE_FFT 1
#define REF_HAVE_REST 1
#endif
#if REF_HAVE_FASTFIR || REF_HAVE_FFT
#include "fft.h"
#endif
#if REF_HAVE_FASTFIR
#include "fastfir.h"
#endif
#if REF_HAVE_DOWNCONVERT
#include "downconvert.h"
#endif
#if REF_HAVE_FFT
#include "windowfunction.h"
#endif
#if REF_HAVE_REST
#include "fir.h"
#include "mixer.h"
#include "decimator.h"
#include "nco.h"
#endif
#include <cstring>

// pebblelib_global.h declares this extern; nothing on the hot path dereferences it.
PebbleLibGlobal *pebbleLibGlobal = nullptr;

extern "C" {

#if REF_HAVE_FASTFIR
// ---- CFastFIR (pebblelib/fastfir.h:50-77) -------------------------------------------------------
void *ref_fastfir_create() { return new CFastFIR(); }
void ref_fastfir_destroy(void *h) { delete (CFastFIR *)h; }
void ref_fastfir_setup(void *h, double lo, double hi, double offset, double fs) {
    ((CFastFIR *)h)->SetupParameters(lo, hi, offset, fs);
}
int ref_fastfir_process(void *h, int n, const double *in, double *out) {
    return ((CFastFIR *)h)->ProcessData(n, (CPX *)in, (CPX *)out);
}

#endif
#if REF_HAVE_DOWNCONVERT
// ---- CDownConvert (pebblelib/downconvert.h:28-37) -------------------------------------------------
void *ref_downconvert_create() { return new CDownConvert(); }
void ref_downconvert_destroy(void *h) { delete (CDownConvert *)h; }
double ref_downconvert_set_data_rate(void *h, double inRate, double maxBW) {
    return ((CDownConvert *)h)->SetDataRate(inRate, maxBW);
}
double ref_downconvert_set_data_rate_simple(void *h, double inRate, double maxBW) {
    return ((CDownConvert *)h)->SetDataRateSimple(inRate, maxBW);
}
void ref_downconvert_set_frequency(void *h, double f) { ((CDownConvert *)h)->SetFrequency(f); }
void ref_downconvert_set_cw_offset(void *h, double f) { ((CDownConvert *)h)->SetCwOffset(f); }
// NOTE: the reference clobbers `in` (downconvert.cpp:309-310,325); callers pass a scratch copy.
int ref_downconvert_process(void *h, int n, double *in, double *out) {
    return ((CDownConvert *)h)->ProcessData(n, (CPX *)in, (CPX *)out);
}

#endif
#if REF_HAVE_FFT
// ---- FFT (pebblelib/fft.h:29-62) via FFT::factory (USE_FFTW or USE_FFTCUTE at build time) ----------
void *ref_fft_create() { return FFT::factory("oracle"); }
void ref_fft_destroy(void *h) { delete (FFT *)h; }
void ref_fft_params(void *h, unsigned size, double dbComp, double fs, int samplesPerBuffer, int windowType) {
    ((FFT *)h)->fftParams(size, dbComp, fs, samplesPerBuffer, (WindowFunction::WINDOWTYPE)windowType);
}
int ref_fft_size(void *h) { return ((FFT *)h)->getFFTSize(); }
void ref_fft_forward(void *h, const double *in, double *out, int n) {
    ((FFT *)h)->fftForward((CPX *)in, (CPX *)out, n);
}
void ref_fft_inverse(void *h, const double *in, double *out, int n) {
    ((FFT *)h)->fftInverse((CPX *)in, (CPX *)out, n);
}
int ref_fft_spectrum(void *h, const double *in, double *out_db, int n) {
    return ((FFT *)h)->fftSpectrum((CPX *)in, out_db, n) ? 1 : 0;
}

// ---- WindowFunction (pebblelib/windowfunction.h) --------------------------------------------------
double ref_window(int type, int n, double *w_out) {
    WindowFunction w((WindowFunction::WINDOWTYPE)type, n);
    if (w_out) memcpy(w_out, w.window, sizeof(double) * n);
    return w.coherentGain;
}

#endif
#if REF_HAVE_REST
// ---- CFir (pebblelib/fir.h:31-56) -----------------------------------------------------------------
void *ref_cfir_create() { return new CFir(); }
void ref_cfir_destroy(void *h) { delete (CFir *)h; }
int ref_cfir_init_lp(void *h, int numTaps, double scale, double astop, double fpass, double fstop, double fs) {
    return ((CFir *)h)->InitLPFilter(numTaps, scale, astop, fpass, fstop, fs);
}
int ref_cfir_init_hp(void *h, int numTaps, double scale, double astop, double fpass, double fstop, double fs) {
    return ((CFir *)h)->InitHPFilter(numTaps, scale, astop, fpass, fstop, fs);
}
void ref_cfir_process_cpx(void *h, int n, const double *in, double *out) {
    ((CFir *)h)->ProcessFilter(n, (CPX *)in, (CPX *)out);
}
void ref_cfir_process_real(void *h, int n, const double *in, double *out) {
    ((CFir *)h)->ProcessFilter(n, (TYPEREAL *)in, (TYPEREAL *)out);
}

// ---- Mixer (pebblelib/mixer.h) --------------------------------------------------------------------
void *ref_mixer_create(unsigned fs, unsigned nsamples) { return new Mixer(fs, nsamples); }
void ref_mixer_destroy(void *h) { delete (Mixer *)h; }
void ref_mixer_set_frequency(void *h, double f) { ((Mixer *)h)->setFrequency(f); }
// Returns 1 if the reference returned `in` itself (f==0 pass-through), else 0; result copied to out.
int ref_mixer_process(void *h, int n, const double *in, double *out) {
    CPX *r = ((Mixer *)h)->processBlock((CPX *)in);
    memcpy(out, r, sizeof(CPX) * n);
    return r == (CPX *)in;
}

// ---- Decimator (pebblelib/decimator.h:228-284) ----------------------------------------------------
void *ref_decimator_create(unsigned fs, unsigned nsamples) { return new Decimator(fs, nsamples); }
void ref_decimator_destroy(void *h) { delete (Decimator *)h; }
float ref_decimator_build_chain(void *h, unsigned rateIn, unsigned protectBw, unsigned rateOut) {
    return ((Decimator *)h)->buildDecimationChain(rateIn, protectBw, rateOut);
}
int ref_decimator_dec_by2_stages(void *h) { return (int)((Decimator *)h)->decBy2Stages(); }
int ref_decimator_process(void *h, int n, const double *in, double *out) {
    return (int)((Decimator *)h)->process((CPX *)in, (CPX *)out, (quint32)n);
}

// ---- NCO (pebblelib/nco.h) -------------------------------------------------------------------------
void *ref_nco_create(unsigned fs, unsigned n) { return new NCO(fs, n); }
void ref_nco_destroy(void *h) { delete (NCO *)h; }
void ref_nco_set_frequency(void *h, double f) { ((NCO *)h)->setFrequency(f); }
void ref_nco_gen_single(void *h, double *io, unsigned n, double gain, int mix) { ((NCO *)h)->genSingle((CPX *)io, n, gain, mix != 0); }

// ---- CpxUtil (pebblelib/cpx.h:161-196, cpx.cpp:35-112) ---------------------------------------------
// op: 0 copy, 1 scale, 2 mult, 3 add, 4 mag, 5 sqrMag, 6 clear, 7 decimate(by = (int)a)
void ref_cpx_op(int op, double *out, const double *in, const double *in2, double a, int n) {
    CPX *o = (CPX *)out; CPX *i1 = (CPX *)in; CPX *i2 = (CPX *)in2;
    switch (op) {
        case 0: CpxUtil::copyCPX(o, i1, n); break;
        case 1: CpxUtil::scaleCPX(o, i1, a, n); break;
        case 2: CpxUtil::multCPX(o, i1, i2, n); break;
        case 3: CpxUtil::addCPX(o, i1, i2, n); break;
        case 4: CpxUtil::magCPX(o, i1, n); break;
        case 5: CpxUtil::sqrMagCPX(o, i1, n); break;
        case 6: CpxUtil::clearCPX(o, n); break;
        default: CpxUtil::decimateCPX(o, i1, (int)a, n); break;
    }
}
#endif  // REF_HAVE_REST
}  // extern "C"
