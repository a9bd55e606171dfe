// oracle/ref_app_shim.cpp — TEST INFRASTRUCTURE ONLY (never linked into the product library).
//
// extern "C" wrappers around the UNMODIFIED application-side steps of the reference, compiled where they lie under
// /root/reference/application (+ pebblelib/{delayline,iir,fractresampler}.cpp) by oracle/build_ref.sh step 5 against
// the empty Qt stand-ins of oracle/qt_shim and oracle/qt_shim_app:
//     AGC (agc.cpp), Demod_AM / Demod_NFM / Demod_SAM / Demod_WFM (demod/*.cpp), SignalSpectrum (signalspectrum.cpp),
//     SignalStrength (signalstrength.cpp), DCRemoval, IQBalance, NoiseBlanker, CFractResampler, CIir,
//     DeviceInterfaceBase::normalizeIQ (function bodies cut from deviceinterfacebase.cpp:532-908, see devbase_skel.h).
// The result, oracle/_ref/libpebbleref_app.so, pins oracle/pebble_oracle.c's restatement of those steps and is the
// checker of the -m gpu parity tests. Nothing here re-implements arithmetic: every function forwards to the reference
// object. demod.cpp itself is not compiled (it drags receiver.h and the whole GUI in); the two trivial Demod
// constructors it defines (demod.cpp:43-47, "ProcessStep(_sampleRate,_bufferSize) {}") are supplied below.
#include "agc.h"
#include "demod.h"
#include "demod/demod_am.h"
#include "demod/demod_nfm.h"
#include "demod/demod_sam.h"
#define private public   // ONLY for the RDS test tap below (reads m_RdsRaw / m_RdsOutputRate); layout and code are untouched
#include "demod/demod_wfm.h"
#undef private
#include "signalstrength.h"
#include "signalspectrum.h"
#include "dcremoval.h"
#include "iqbalance.h"
#include "noiseblanker.h"
#include "fractresampler.h"
#include "iir.h"
#include "devbase_skel.h"
#include <cstdlib>
#include <cstring>

extern "C" long long ref_fake_clock_ms = 0;   // read by the QElapsedTimer stand-in (qt_shim_app/QElapsedTimer)
PebbleLibGlobal *pebbleLibGlobal = nullptr;
Global *global = nullptr;
DeviceInterface::~DeviceInterface() {}

// demod.cpp:43-47 (the sub-class constructor) and :71-73. Members only; no arithmetic.
Demod::Demod(quint32 _sampleRate, quint32 _bufferSize) : ProcessStep(_sampleRate, _bufferSize) {}
Demod::~Demod() {}
// moc would generate these signal bodies; nothing on the path listens.
void Demod::bandData(QString, QString, QString, QString) {}
void Demod::outputBandData(QString, QString, QString, QString) {}
void SignalStrength::newSignalStrength(double, double, double, double, double) {}
void SignalSpectrum::newFftData() {}
void SignalSpectrum::newHiResFftData() {}

// `global->settings` is read by SignalSpectrum's constructor (signalspectrum.cpp:10-30). Global/Settings have
// GUI-laden constructors (global.cpp, settings.cpp) that are not compiled: raw zeroed storage with the four plain
// fields set is all the step looks at.
static void ensure_global(int bins, int hires_bins, int updates_per_sec) {
    if (!global) {
        global = (Global *)calloc(1, sizeof(Global));
        global->settings = (Settings *)calloc(1, sizeof(Settings));
    }
    global->settings->m_numSpectrumBins = bins;
    global->settings->m_numHiResSpectrumBins = hires_bins;
    global->settings->m_dbOffset = 0;
    global->settings->m_updatesPerSecond = updates_per_sec;
}

extern "C" {

void ref_set_clock_ms(long long t) { ref_fake_clock_ms = t; }

// ---- AGC (application/agc.h) ------------------------------------------------------------------------------------
void *ref_agc_create(unsigned fs, unsigned n) { return new AGC(fs, n); }
void ref_agc_destroy(void *h) { delete (AGC *)h; }
void ref_agc_set_mode(void *h, int mode, int threshold) { ((AGC *)h)->setAgcMode((AGC::AgcMode)mode, threshold); }
void ref_agc_set_threshold(void *h, int g) { ((AGC *)h)->setAgcThreshold(g); }
int ref_agc_get_threshold(void *h) { return ((AGC *)h)->getAgcThreshold(); }
// processes exactly the bufferSize the object was created with (agc.cpp:84-207 loops over numSamples)
void ref_agc_process(void *h, int n, const double *in, double *out) {
    CPX *r = ((AGC *)h)->processBlock((CPX *)in);
    memcpy(out, r, sizeof(CPX) * n);
}

// ---- Demod_AM (demod/demod_am.h) --------------------------------------------------------------------------------
void *ref_demod_am_create(int rate, int n) { return new Demod_AM(rate, n); }
void ref_demod_am_destroy(void *h) { delete (Demod_AM *)h; }
void ref_demod_am_set_bandwidth(void *h, double bw) { ((Demod_AM *)h)->setBandwidth(bw); }
void ref_demod_am_process_filtered(void *h, int n, const double *in, double *out) {
    ((Demod_AM *)h)->processBlockFiltered((CPX *)in, (CPX *)out, n);
}
void ref_demod_am_process(void *h, int n, const double *in, double *out) {
    ((Demod_AM *)h)->processBlock((CPX *)in, (CPX *)out, n);
}

// ---- Demod_NFM (demod/demod_nfm.h) ------------------------------------------------------------------------------
void *ref_demod_nfm_create(int rate, int n) { return new Demod_NFM(rate, n); }
void ref_demod_nfm_destroy(void *h) { delete (Demod_NFM *)h; }
void ref_demod_nfm_process_nco(void *h, int n, const double *in, double *out) {
    ((Demod_NFM *)h)->processBlockNCO((CPX *)in, (CPX *)out, n);
}
// NOTE: processBlockFM2 keeps its last sample in a FUNCTION-STATIC (demod_nfm.cpp:129): shared by all objects.
void ref_demod_nfm_process_fm2(void *h, int n, const double *in, double *out) {
    ((Demod_NFM *)h)->processBlockFM2((CPX *)in, (CPX *)out, n);
}

// ---- Demod_SAM (demod/demod_sam.h) ------------------------------------------------------------------------------
void *ref_demod_sam_create(int rate, int n) { return new Demod_SAM(rate, n); }
void ref_demod_sam_destroy(void *h) { delete (Demod_SAM *)h; }
void ref_demod_sam_process(void *h, int n, const double *in, double *out) {
    ((Demod_SAM *)h)->processBlock((CPX *)in, (CPX *)out, n);
}

// ---- Demod_WFM (demod/demod_wfm.h) ------------------------------------------------------------------------------
void *ref_demod_wfm_create(int rate, int n) { return new Demod_WFM(rate, n); }
void ref_demod_wfm_destroy(void *h) { delete (Demod_WFM *)h; }
double ref_demod_wfm_set_sample_rate(void *h, double rate, double out_rate, int us) {
    return ((Demod_WFM *)h)->setSampleRate(rate, out_rate, us != 0);
}
int ref_demod_wfm_process_mono(void *h, int n, const double *in, double *out) {
    return ((Demod_WFM *)h)->processDataMono(n, (CPX *)in, (CPX *)out);
}
int ref_demod_wfm_process_stereo(void *h, int n, const double *in, double *out) {
    return ((Demod_WFM *)h)->processDataStereo(n, (CPX *)in, (CPX *)out);
}
int ref_demod_wfm_stereo_lock(void *h, int *locked) { return ((Demod_WFM *)h)->getStereoLock(locked); }
// returns 1 and fills four 16-bit blocks when an RDS group is queued
int ref_demod_wfm_next_rds_group(void *h, unsigned short *blocks4) {
    tRDS_GROUPS g;
    if (!((Demod_WFM *)h)->getNextRdsGroupData(&g)) return 0;
    blocks4[0] = g.BlockA; blocks4[1] = g.BlockB; blocks4[2] = g.BlockC; blocks4[3] = g.BlockD;
    return 1;
}

// test tap: what processDataStereo itself leaves in m_RdsRaw (demod_wfm.cpp:336-352): per RDS-rate sample the pair
// (matched-filter value at a sampling instant or 0, matched-filter output). len = InLength / the RDS decimation.
int ref_demod_wfm_rds_tap(void *h, int len, double *out) {
    memcpy(out, ((Demod_WFM *)h)->m_RdsRaw, sizeof(double) * 2 * (size_t)len);
    return len;
}
double ref_demod_wfm_rds_rate(void *h) { return ((Demod_WFM *)h)->m_RdsOutputRate; }
// the block / group synchroniser alone: Demod_WFM::processNewRdsBit (demod_wfm.cpp:584) fed with data bits
void ref_demod_wfm_push_bits(void *h, const unsigned char *bits, int n) {
    for (int i = 0; i < n; ++i) ((Demod_WFM *)h)->processNewRdsBit(bits[i] & 1);
}

// ---- SignalSpectrum (application/signalspectrum.h) ---------------------------------------------------------------
void *ref_spectrum_create(unsigned fs, unsigned hires_fs, unsigned n, int bins, int hires_bins, int updates_per_sec) {
    ensure_global(bins, hires_bins, updates_per_sec);
    return new SignalSpectrum(fs, hires_fs, n);
}
void ref_spectrum_destroy(void *h) { /* the reference destructor free()s new[]-allocated arrays: leak instead */ (void)h; }
void ref_spectrum_set_hires(void *h, int on) { ((SignalSpectrum *)h)->setHiRes(on != 0); }
void ref_spectrum_set_updates_per_sec(void *h, int u) { ((SignalSpectrum *)h)->setUpdatesPerSec(u); }
void ref_spectrum_set_sample_rate(void *h, unsigned fs, unsigned hires_fs) { ((SignalSpectrum *)h)->setSampleRate(fs, hires_fs); }
void ref_spectrum_unprocessed(void *h, const double *in, int n) { ((SignalSpectrum *)h)->unprocessed((CPX *)in, n); }
void ref_spectrum_zoomed(void *h, const double *in, int n) { ((SignalSpectrum *)h)->zoomed((CPX *)in, n); }
int ref_spectrum_bin_count(void *h) { return ((SignalSpectrum *)h)->binCount(); }
int ref_spectrum_overload(void *h) { return ((SignalSpectrum *)h)->getOverload() ? 1 : 0; }
int ref_spectrum_update_pending(void *h) { return ((SignalSpectrum *)h)->m_displayUpdateComplete ? 0 : 1; }
void ref_spectrum_ack_update(void *h) { ((SignalSpectrum *)h)->m_displayUpdateComplete = true; }
int ref_spectrum_overrun(void *h) { return ((SignalSpectrum *)h)->m_displayUpdateOverrun; }
void ref_spectrum_get_unprocessed(void *h, double *out) {
    SignalSpectrum *s = (SignalSpectrum *)h;
    memcpy(out, s->getUnprocessed(), sizeof(double) * s->binCount());
}

// ---- SignalStrength (application/signalstrength.h) ---------------------------------------------------------------
void *ref_sigstrength_create(unsigned fs, unsigned n) { return new SignalStrength(fs, n); }
void ref_sigstrength_destroy(void *h) { delete (SignalStrength *)h; }
double ref_sigstrength_fd_estimate(void *h, const double *spectrum, int bins, unsigned fs, float lo, float hi, double mixer) {
    return ((SignalStrength *)h)->fdEstimate((double *)spectrum, bins, fs, lo, hi, mixer);
}
void ref_sigstrength_values(void *h, double *out5) {
    SignalStrength *s = (SignalStrength *)h;
    out5[0] = s->peakDb(); out5[1] = s->avgDb(); out5[2] = s->snrDb(); out5[3] = s->floorDb(); out5[4] = s->extValue();
}

// ---- DCRemoval / IQBalance / NoiseBlanker (application/) ---------------------------------------------------------
void *ref_dcremoval_create(unsigned fs, unsigned n) { return new DCRemoval(fs, n); }
void ref_dcremoval_destroy(void *h) { delete (DCRemoval *)h; }
void ref_dcremoval_enable(void *h, int on) { ((DCRemoval *)h)->enableStep(on != 0); }
void ref_dcremoval_process(void *h, int n, const double *in, double *out) {
    CPX *r = ((DCRemoval *)h)->process((CPX *)in, (quint32)n);
    memcpy(out, r, sizeof(CPX) * n);
}
void *ref_iqbalance_create(unsigned fs, unsigned n) { return new IQBalance(fs, n); }
void ref_iqbalance_destroy(void *h) { (void)h; /* reference dtor uses delete on new[] */ }
void ref_iqbalance_config(void *h, int enabled, double gain, double phase) {
    IQBalance *b = (IQBalance *)h;
    b->enableStep(enabled != 0); b->setGainFactor(gain); b->setPhaseFactor(phase);
}
int ref_iqbalance_process(void *h, int n, const double *in, double *out) {
    CPX *r = ((IQBalance *)h)->ProcessBlock((CPX *)in);
    memcpy(out, r, sizeof(CPX) * n);
    return r == (CPX *)in;
}
void *ref_noiseblanker_create(unsigned fs, unsigned n) { return new NoiseBlanker(fs, n); }
void ref_noiseblanker_destroy(void *h) { delete (NoiseBlanker *)h; }
void ref_noiseblanker_enable(void *h, int nb, int nb2) {
    ((NoiseBlanker *)h)->setNbEnabled(nb != 0); ((NoiseBlanker *)h)->setNb2Enabled(nb2 != 0);
}
int ref_noiseblanker_process(void *h, int n, const double *in, double *out) {
    CPX *r = ((NoiseBlanker *)h)->ProcessBlock((CPX *)in);
    memcpy(out, r, sizeof(CPX) * n);
    return r == (CPX *)in;
}
int ref_noiseblanker_process2(void *h, int n, const double *in, double *out) {
    CPX *r = ((NoiseBlanker *)h)->ProcessBlock2((CPX *)in);
    memcpy(out, r, sizeof(CPX) * n);
    return r == (CPX *)in;
}

// ---- CFractResampler (pebblelib/fractresampler.h) -----------------------------------------------------------------
void *ref_resampler_create(int max_input) { CFractResampler *r = new CFractResampler(); r->Init(max_input); return r; }
void ref_resampler_destroy(void *h) { (void)h; /* reference dtor uses delete on new[] */ }
int ref_resampler_resample_cpx(void *h, int n, double rate, const double *in, double *out) {
    return ((CFractResampler *)h)->Resample(n, rate, (TYPECPX *)in, (TYPECPX *)out);
}
int ref_resampler_resample_stereo16(void *h, int n, double rate, const double *in, short *out, double gain) {
    return ((CFractResampler *)h)->Resample(n, rate, (TYPECPX *)in, (TYPESTEREO16 *)out, gain);
}

// ---- CIir (pebblelib/iir.h) ----------------------------------------------------------------------------------------
void *ref_iir_create() { return new CIir(); }
void ref_iir_destroy(void *h) { delete (CIir *)h; }
void ref_iir_init(void *h, int kind, double f0, double q_or_bw, double fs) {
    CIir *f = (CIir *)h;
    if (kind == 0) f->InitLP(f0, q_or_bw, fs);
    else if (kind == 1) f->InitHP(f0, q_or_bw, fs);
    else if (kind == 2) f->InitBP(f0, q_or_bw, fs);
    else f->InitBR(f0, q_or_bw, fs);
}
void ref_iir_process_cpx(void *h, int n, const double *in, double *out) { ((CIir *)h)->ProcessFilter(n, (TYPECPX *)in, (TYPECPX *)out); }
void ref_iir_process_real(void *h, int n, const double *in, double *out) { ((CIir *)h)->ProcessFilter(n, (TYPEREAL *)in, (TYPEREAL *)out); }

// ---- DeviceInterfaceBase::normalizeIQ (pebblelib/deviceinterfacebase.cpp:648-878, block forms) --------------------
// fmt: 0 CPX8, 1 CPXU8, 2 CPX16, 3 CPXFLOAT, 4 CPX; order = DeviceInterface::IQOrder
void ref_normalize_iq(int fmt, int order, double user_gain, double norm_gain, int reverse, unsigned n, const void *in, double *out) {
    DeviceInterfaceBase d;
    d.m_userIQGain = user_gain; d.m_normalizeIQGain = norm_gain; d.m_iqOrder = (DeviceInterface::IQOrder)order;
    switch (fmt) {
        case 0: d.normalizeIQ((CPX *)out, (CPX8 *)in, n, reverse != 0); break;
        case 1: d.normalizeIQ((CPX *)out, (CPXU8 *)in, n, reverse != 0); break;
        case 2: d.normalizeIQ((CPX *)out, (CPX16 *)in, n, reverse != 0); break;
        case 3: d.normalizeIQ((CPX *)out, (CPXFLOAT *)in, n, reverse != 0); break;
        default: d.normalizeIQ((CPX *)out, (CPX *)in, n, reverse != 0); break;
    }
}

}  // extern "C"
