// pebble_dropin.hpp — the C++ drop-in: classes with the reference's own names, signatures and ownership rules that forward
// to the C ABI of libpebblecuda.so (pebble_cuda.h). A maintainer removes fastfir.cpp downconvert.cpp fft.cpp fftw.cpp fir.cpp
// mixer.cpp nco.cpp decimator.cpp fractresampler.cpp (pebblelib) and agc.cpp demod.cpp demod/*.cpp bandpassfilter.cpp
// signalspectrum.cpp (application) from the build, includes this header where those classes' headers were included, links
// -lpebblecuda, and Receiver::processIQData (application/receiver.cpp:758-1009) compiles and runs unchanged.
// tests/dropin/receiver_tail.cpp is that function's non-WFM branch (:903-1007) written against these classes; it is built
// by g++ and run by tests/test_gpu_dropin.py.
//
// Each class cites the reference interface it replaces. Conventions kept:
//   * ProcessStep shape (application/processstep.h:16-46): the step OWNS `out` (16-byte aligned, bufferSize CPX, allocated in
//     the constructor, freed in the destructor), returns a pointer the caller must not free, never modifies `in`, and a
//     pass-through step returns `in` itself;
//   * CuteSDR shape (fastfir.h:59, downconvert.h:35, fir.h:41-43): int ProcessData(int, CPX*, CPX*) returns the sample count;
//   * no exceptions and no error codes cross these classes (SURVEY 8b "Errors"): a failed GPU call returns 0 samples / leaves
//     `out` untouched and the text is available from pebble_last_error().
// Inside the Qt build define PEBBLE_DROPIN_QT (pebblelib's cpx.h and Qt's integer typedefs are then used instead of the
// stand-ins below, and ProcessStep keeps deriving from QObject in processstep.h).
#ifndef PEBBLE_DROPIN_HPP
#define PEBBLE_DROPIN_HPP

#include "pebble_cuda.h"
#include <chrono>
#include <complex>
#include <cstdlib>
#include <cstring>

#ifndef PEBBLE_DROPIN_QT
typedef std::complex<double> CPX;          // pebblelib/cpx.h:96
typedef double TYPEREAL;                   // pebblelib/cpx.h (CuteSDR names)
typedef std::complex<double> TYPECPX;
typedef unsigned int quint32;
typedef unsigned short quint16;
typedef int qint32;
typedef long long qint64;
#endif

namespace pebble_dropin {
// CpxUtil::memalign (pebblelib/cpx.cpp:24-33): 16-byte aligned CPX array
inline CPX *memalign(size_t n) {
    void *p = nullptr;
    if (posix_memalign(&p, 16, sizeof(CPX) * (n ? n : 1)) != 0) return nullptr;
    std::memset(p, 0, sizeof(CPX) * (n ? n : 1));
    return static_cast<CPX *>(p);
}
inline const double *dp(const CPX *p) { return reinterpret_cast<const double *>(p); }
inline double *dp(CPX *p) { return reinterpret_cast<double *>(p); }
inline long long now_ms() {
    using namespace std::chrono;
    return duration_cast<milliseconds>(steady_clock::now().time_since_epoch()).count();
}
}  // namespace pebble_dropin

// ---- CpxUtil helpers used by the receiver between steps (pebblelib/cpx.h:161-196, cpx.cpp:35-112): host loops, as in the
// reference (they touch framesPerBuffer samples; the per-sample work at the device rate lives in the operators) -----------
inline void copyCPX(CPX *out, const CPX *in, int n) { std::memcpy(out, in, sizeof(CPX) * (size_t)n); }
inline void clearCPX(CPX *out, int n) { std::memset(out, 0, sizeof(CPX) * (size_t)n); }
inline void scaleCPX(CPX *out, const CPX *in, double a, int n) { for (int i = 0; i < n; ++i) out[i] = CPX(in[i].real() * a, in[i].imag() * a); }
inline void addCPX(CPX *out, const CPX *a, const CPX *b, int n) { for (int i = 0; i < n; ++i) out[i] = a[i] + b[i]; }
inline void multCPX(CPX *out, const CPX *a, const CPX *b, int n) { for (int i = 0; i < n; ++i) out[i] = a[i] * b[i]; }
inline void decimateCPX(CPX *out, const CPX *in, int by, int n) { for (int i = 0, j = 0; i < n; i += by) out[j++] = in[i]; }

// ---- CFastFIR — pebblelib/fastfir.h:50-77, fastfir.cpp:77-334 ----------------------------------------------------------------
class CFastFIR {
public:
    CFastFIR() : h(pebble_fastfir_create()) {}
    virtual ~CFastFIR() { pebble_fastfir_destroy(h); }
    void SetupParameters(TYPEREAL FLoCut, TYPEREAL FHiCut, TYPEREAL Offset, TYPEREAL SampleRate) {   // fastfir.cpp:191-272
        pebble_fastfir_setup(h, FLoCut, FHiCut, Offset, SampleRate);
    }
    int ProcessData(int InLength, TYPECPX *InBuf, TYPECPX *OutBuf) {                                  // fastfir.cpp:281-319
        const int n = pebble_fastfir_process(h, InLength, pebble_dropin::dp(InBuf), pebble_dropin::dp(OutBuf));
        return n < 0 ? 0 : n;
    }
private:
    pebble_fastfir_t *h;
    CFastFIR(const CFastFIR &);
    CFastFIR &operator=(const CFastFIR &);
};

// ---- CDownConvert — pebblelib/downconvert.h:28-37, downconvert.cpp:64-535 ---------------------------------------------------
class CDownConvert {
public:
    CDownConvert() : h(pebble_downconvert_create()) {}
    virtual ~CDownConvert() { pebble_downconvert_destroy(h); }
    void SetFrequency(TYPEREAL NcoFreq) { pebble_downconvert_set_frequency(h, NcoFreq); }
    void SetCwOffset(TYPEREAL offset) { pebble_downconvert_set_cw_offset(h, offset); }
    TYPEREAL SetDataRate(TYPEREAL InRate, TYPEREAL MaxBW) { return pebble_downconvert_set_data_rate(h, InRate, MaxBW); }
    TYPEREAL SetDataRateSimple(TYPEREAL InRate, TYPEREAL MaxBW) { return pebble_downconvert_set_data_rate_simple(h, InRate, MaxBW); }
    int ProcessData(int InLength, TYPECPX *pInData, TYPECPX *pOutData) {   // downconvert.cpp:257-334; pInData is NOT clobbered (F6)
        const int n = pebble_downconvert_process(h, InLength, pebble_dropin::dp(pInData), pebble_dropin::dp(pOutData));
        return n < 0 ? 0 : n;
    }
    int NumStages() const { return pebble_downconvert_num_stages(h); }
private:
    pebble_downconvert_t *h;
    CDownConvert(const CDownConvert &);
    CDownConvert &operator=(const CDownConvert &);
};

// ---- ProcessStep — application/processstep.h:16-46, processstep.cpp:6-32 ----------------------------------------------------
#ifndef PEBBLE_DROPIN_QT
class ProcessStep {
public:
    ProcessStep(quint32 _sampleRate, quint32 _bufferSize)
        : numSamples(_bufferSize), bufferSize(_bufferSize), numSamplesX2(_bufferSize * 2), sampleRate(_sampleRate),
          out(pebble_dropin::memalign(_bufferSize)), enabled(false) {}
    virtual ~ProcessStep() { if (out) free(out); }
    virtual CPX *process(CPX *in, quint32) { return in; }
    quint32 getSampleRate() { return sampleRate; }
    quint32 getBufferSize() { return bufferSize; }
    void enableStep(bool _enableStep) { enabled = _enableStep; }
    bool isEnabled() { return enabled; }
protected:
    quint32 numSamples, bufferSize, numSamplesX2, sampleRate;
    CPX *out;          // every step returns its output in a buffer it owns (processstep.h:37-40)
    bool enabled;
private:
    ProcessStep(const ProcessStep &);
    ProcessStep &operator=(const ProcessStep &);
};
#endif

// ---- Mixer — pebblelib/mixer.h, mixer.cpp:5-81 ------------------------------------------------------------------------------------
class Mixer {
public:
    Mixer(quint32 _sampleRate, quint32 _bufferSize)
        : h(pebble_mixer_create(_sampleRate, _bufferSize)), m_out(pebble_dropin::memalign(_bufferSize)), m_numSamples(_bufferSize) {}
    ~Mixer() { pebble_mixer_destroy(h); free(m_out); }
    void setFrequency(double f) { pebble_mixer_set_frequency(h, f); }
    CPX *processBlock(CPX *in) {                                           // mixer.cpp:48-81: f == 0 returns `in` itself
        if (pebble_mixer_is_passthrough(h) == 1) return in;
        if (pebble_mixer_process(h, (int)m_numSamples, pebble_dropin::dp(in), pebble_dropin::dp(m_out)) < 0) return in;
        return m_out;
    }
private:
    pebble_mixer_t *h;
    CPX *m_out;
    quint32 m_numSamples;
};

// ---- Decimator — pebblelib/decimator.h:228-284, decimator.cpp:6-226 -----------------------------------------------------------
class Decimator {
public:
    Decimator(quint32 _sampleRate, quint32 _bufferSize) : h(pebble_decimator_create(_sampleRate, _bufferSize)) {}
    ~Decimator() { pebble_decimator_destroy(h); }
    float buildDecimationChain(quint32 _sampleRateIn, quint32 _protectBw, quint32 _sampleRateOut = 0) {
        return pebble_decimator_build_chain(h, _sampleRateIn, _protectBw, _sampleRateOut);
    }
    quint32 decBy2Stages() { const int s = pebble_decimator_dec_by2_stages(h); return s < 0 ? 0 : (quint32)s; }
    quint32 process(CPX *_in, CPX *_out, quint32 _numSamples) {
        const int n = pebble_decimator_process(h, (int)_numSamples, pebble_dropin::dp(_in), pebble_dropin::dp(_out));
        return n < 0 ? 0 : (quint32)n;
    }
private:
    pebble_decimator_t *h;
};

// ---- WindowFunction — pebblelib/windowfunction.h:8-33 ------------------------------------------------------------------------------
class WindowFunction {
public:
    enum WINDOWTYPE { RECTANGULAR = 1, HANNING, WELCH, PARZEN, BARTLETT, HAMMING, BLACKMAN2, BLACKMAN3, BLACKMAN4, EXPONENTIAL,
                      RIEMANN, BLACKMANHARRIS, BLACKMAN, NONE };
    WindowFunction(WINDOWTYPE _type, int _numSamples) : window(new double[_numSamples]), windowCpx(pebble_dropin::memalign(_numSamples)),
                                                        windowType(_type), numSamples(_numSamples) { regenerate(_type); }
    ~WindowFunction() { delete[] window; free(windowCpx); }
    void regenerate(WINDOWTYPE _type) {                                    // windowfunction.cpp:50-241
        windowType = _type;
        coherentGain = pebble_window((int)_type, numSamples, window);
        for (int i = 0; i < numSamples; ++i) windowCpx[i] = CPX(window[i], 0.0);
    }
    double *window;
    CPX *windowCpx;
    double coherentGain;
    WINDOWTYPE windowType;
    int numSamples;
};

// ---- FFT — pebblelib/fft.h:29-62 with the FFTfftw (fftw.cpp:17-81) or CFft (fftcute.cpp:58-277) contract ------------------------
class FFT {
public:
    FFT() : h(pebble_fft_create()), m_timeDomain(nullptr), m_freqDomain(nullptr), m_fftSize(0), m_isOverload(false) {}
    virtual ~FFT() { pebble_fft_destroy(h); free(m_timeDomain); free(m_freqDomain); }
    // FFT::factory (fft.cpp:45-65) picks the backend at compile time; here at run time: 0 FFTfftw, 1 CFft (Ooura)
    static FFT *factory(const char * /*label*/, int backend = PEBBLE_FFT_FFTW) {
        FFT *f = new FFT();
        pebble_fft_set_backend(f->h, backend);
        return f;
    }
    virtual void fftParams(quint32 _fftSize, double _dBCompensation, double _sampleRate, int _samplesPerBuffer,
                           WindowFunction::WINDOWTYPE _windowType) {      // fft.cpp:67-118
        if (pebble_fft_params(h, _fftSize, _dBCompensation, _sampleRate, _samplesPerBuffer, (int)_windowType) < 0) return;
        m_fftSize = pebble_fft_size(h);
        free(m_timeDomain); free(m_freqDomain);
        m_timeDomain = pebble_dropin::memalign(m_fftSize);
        m_freqDomain = pebble_dropin::memalign(m_fftSize);
    }
    // in == NULL: use the internal time-/frequency-domain buffer; out == NULL: leave the result there (fftw.cpp:31-67)
    virtual void fftForward(CPX *in, CPX *out, int numSamples) {
        if (m_fftSize <= 0) return;
        const CPX *src = in ? in : m_timeDomain;
        if (pebble_fft_forward(h, pebble_dropin::dp(src), pebble_dropin::dp(m_freqDomain), in ? numSamples : m_fftSize) < 0) return;
        if (out && out != m_freqDomain) copyCPX(out, m_freqDomain, m_fftSize);
    }
    virtual void fftInverse(CPX *in, CPX *out, int numSamples) {
        if (m_fftSize <= 0) return;
        const CPX *src = in ? in : m_freqDomain;
        if (pebble_fft_inverse(h, pebble_dropin::dp(src), pebble_dropin::dp(m_timeDomain), in ? numSamples : m_fftSize) < 0) return;
        if (out && out != m_timeDomain) copyCPX(out, m_timeDomain, m_fftSize);
    }
    virtual bool fftSpectrum(CPX *in, double *out, int numSamples) {      // fftw.cpp:70-81 + fft.cpp:324-399
        const int r = pebble_fft_spectrum(h, pebble_dropin::dp(in), out, numSamples);
        if (r >= 0) m_isOverload = r == 1;
        return m_isOverload;
    }
    CPX *getFreqDomain() { return m_freqDomain; }
    CPX *getTimeDomain() { return m_timeDomain; }
    qint32 getFFTSize() { return m_fftSize; }
    bool isOverload() { return m_isOverload; }
protected:
    pebble_fft_t *h;
    CPX *m_timeDomain, *m_freqDomain;
    int m_fftSize;
    bool m_isOverload;
};

// ---- SignalSpectrum — application/signalspectrum.h, signalspectrum.cpp:6-133 ------------------------------------------------------
// The reference reads the bin counts and the update rate from global->settings; they are constructor arguments here.
class SignalSpectrum : public ProcessStep {
public:
    SignalSpectrum(quint32 _sampleRate, quint32 _hiResSampleRate, quint32 _bufferSize, int _numSpectrumBins = 4096,
                   int _numHiResSpectrumBins = 2048, int _updatesPerSecond = 10)
        : ProcessStep(_sampleRate, _bufferSize),
          h(pebble_spectrum_create(_sampleRate, _hiResSampleRate, _bufferSize, _numSpectrumBins, _numHiResSpectrumBins, _updatesPerSecond)),
          m_hiResSampleRate(_hiResSampleRate), m_displayUpdateComplete(true) {}
    ~SignalSpectrum() { pebble_spectrum_destroy(h); }
    void setHiRes(bool _on) { pebble_spectrum_set_hires(h, _on ? 1 : 0); }
    void unprocessed(CPX *in, int _numSamples) {                           // signalspectrum.cpp:63-86
        if (m_displayUpdateComplete) pebble_spectrum_ack_update(h);
        if (pebble_spectrum_unprocessed(h, pebble_dropin::dp(in), _numSamples, pebble_dropin::now_ms()) == 1) m_displayUpdateComplete = false;
    }
    void zoomed(CPX *in, int _numSamples) { pebble_spectrum_zoomed(h, pebble_dropin::dp(in), _numSamples, pebble_dropin::now_ms()); }
    int binCount() { return pebble_spectrum_bin_count(h); }
    double *getUnprocessed() { return const_cast<double *>(pebble_spectrum_get_unprocessed(h)); }
    double *getHiRes() { return const_cast<double *>(pebble_spectrum_get_hires(h)); }
    void setUpdatesPerSec(int updatespersec) { pebble_spectrum_set_updates_per_sec(h, updatespersec); }
    void setSampleRate(quint32 _sampleRate, quint32 _hiResSampleRate) {
        sampleRate = _sampleRate; m_hiResSampleRate = _hiResSampleRate;
        pebble_spectrum_set_sample_rate(h, _sampleRate, _hiResSampleRate);
    }
    quint32 getHiResSampleRate() { return m_hiResSampleRate; }
    bool getOverload() { return pebble_spectrum_overload(h) == 1; }
    int displayUpdateOverrun() { return pebble_spectrum_overrun(h); }
    bool m_displayUpdateComplete;   // set true by the consumer (SpectrumWidget) when it has painted the spectrum
private:
    pebble_spectrum_t *h;
    quint32 m_hiResSampleRate;
};

// ---- CFir — pebblelib/fir.h:31-56, fir.cpp:106-132, 246-440 ----------------------------------------------------------------------
class CFir {
public:
    CFir() : h(pebble_cfir_create()) {}
    ~CFir() { pebble_cfir_destroy(h); }
    int InitLPFilter(int NumTaps, TYPEREAL Scale, TYPEREAL Astop, TYPEREAL Fpass, TYPEREAL Fstop, TYPEREAL Fsamprate) {
        return pebble_cfir_init_lp(h, NumTaps, Scale, Astop, Fpass, Fstop, Fsamprate);
    }
    int InitHPFilter(int NumTaps, TYPEREAL Scale, TYPEREAL Astop, TYPEREAL Fpass, TYPEREAL Fstop, TYPEREAL Fsamprate) {
        return pebble_cfir_init_hp(h, NumTaps, Scale, Astop, Fpass, Fstop, Fsamprate);
    }
    void InitConstFir(int NumTaps, const double *pICoef, const double *pQCoef, TYPEREAL Fsamprate) {
        pebble_cfir_init_const(h, NumTaps, pICoef, pQCoef, Fsamprate);
    }
    void ProcessFilter(int InLength, TYPECPX *InBuf, TYPECPX *OutBuf) {    // in == out allowed
        pebble_cfir_process(h, InLength, pebble_dropin::dp(InBuf), pebble_dropin::dp(OutBuf));
    }
private:
    pebble_cfir_t *h;
    CFir(const CFir &);
    CFir &operator=(const CFir &);
};

// ---- BandPassFilter — application/bandpassfilter.h, bandpassfilter.cpp:3-59 (m_useFastFIR = true) --------------------------------
class BandPassFilter : public ProcessStep {
public:
    BandPassFilter(quint32 _sampleRate, quint32 _bufferSize) : ProcessStep(_sampleRate, _bufferSize), m_lowFreq(0), m_highFreq(0) {}
    void setBandPass(float _low, float _high) {
        m_lowFreq = _low; m_highFreq = _high;
        m_bpFilter2.SetupParameters(m_lowFreq, m_highFreq, 0, sampleRate);
    }
    CPX *process(CPX *in, quint32 _numSamples) {                           // ignores the returned count, like the reference (:52-54)
        m_bpFilter2.ProcessData((int)_numSamples, in, out);
        return out;
    }
    float lowFreq() { return m_lowFreq; }
    float highFreq() { return m_highFreq; }
private:
    CFastFIR m_bpFilter2;
    float m_lowFreq, m_highFreq;
};

// ---- AGC — application/agc.h, agc.cpp:14-309 ---------------------------------------------------------------------------------------
class AGC : public ProcessStep {
public:
    enum AgcMode { AGC_OFF, ACG_FAST, AGC_MED, AGC_SLOW, AGC_LONG };
    AGC(quint32 _sampleRate, quint32 _bufferSize)
        : ProcessStep(_sampleRate, _bufferSize), h(pebble_agc_create(_sampleRate, _bufferSize)), uiSliderValueDb(0) {
        setAgcMode(AGC_OFF, 1);                                            // agc.cpp:30
    }
    ~AGC() { pebble_agc_destroy(h); }
    void setAgcMode(AgcMode m, int threshold) { pebble_agc_set_mode(h, (int)m, threshold); }
    void setAgcThreshold(int g) { uiSliderValueDb = g; pebble_agc_set_threshold(h, g); }
    int getAgcThreshold() { return uiSliderValueDb; }
    CPX *processBlock(CPX *in) {                                           // agc.cpp:84-207: numSamples per call, into its own `out`
        pebble_agc_process(h, (int)numSamples, pebble_dropin::dp(in), pebble_dropin::dp(out));
        return out;
    }
private:
    pebble_agc_t *h;
    int uiSliderValueDb;
};

// ---- Demod — application/demod.h, demod.cpp:100-141 -----------------------------------------------------------------------------
struct DeviceInterface {   // the one enum of pebblelib/device_interfaces.h:124-137 the demodulator needs
    enum DemodMode { dmAM, dmSAM, dmFMN, dmFMM, dmFMS, dmDSB, dmLSB, dmUSB, dmCWL, dmCWU, dmDIGL, dmDIGU, dmNONE };
};
struct tRDS_GROUPS { quint16 BlockA, BlockB, BlockC, BlockD; };   // application/demod/rbdsconstants.h:100-106
class Demod : public ProcessStep {
public:
    Demod(quint32 _sampleRate, quint32 _bufferSize)
        : ProcessStep(_sampleRate, _bufferSize), h(pebble_demod_create(_sampleRate, _bufferSize)), m_mode(DeviceInterface::dmAM) {}
    // demod.cpp:50-66: the WFM demodulator runs at its own (wider) input rate
    Demod(quint32 _sampleRate, quint32 _wfmSampleRate, quint32 _bufferSize)
        : ProcessStep(_sampleRate, _bufferSize), h(pebble_demod_create(_sampleRate, _bufferSize)), m_mode(DeviceInterface::dmAM) {
        pebble_demod_set_wfm_sample_rate(h, (double)_wfmSampleRate);
    }
    ~Demod() { pebble_demod_destroy(h); }
    void setDemodMode(DeviceInterface::DemodMode _mode, int /*_sourceSampleRate*/, int /*_audioSampleRate*/) {
        if (pebble_demod_set_mode(h, (int)_mode) == 0) m_mode = _mode;
    }
    DeviceInterface::DemodMode demodMode() const { return m_mode; }
    // what Demod::fmStereo reads from its Demod_WFM after every block (demod.cpp:196-219, demod_wfm.cpp:436-448, 764-788)
    int getStereoLock(int *pPilotLock) {
        const int now = pebble_demod_get_stereo_lock(h);
        const int changed = now != m_lastLock;
        m_lastLock = now;
        if (pPilotLock) *pPilotLock = now > 0;
        return changed;
    }
    int getNextRdsGroupData(tRDS_GROUPS *pGroupData) {
        unsigned short b[4];
        if (!pGroupData || pebble_demod_get_next_rds_group(h, b) != 1) return 0;
        pGroupData->BlockA = b[0]; pGroupData->BlockB = b[1]; pGroupData->BlockC = b[2]; pGroupData->BlockD = b[3];
        return 1;
    }
    void setBandwidth(double bandwidth) { if (m_mode == DeviceInterface::dmAM) pebble_demod_set_bandwidth(h, bandwidth); }   // demod.cpp:230-240
    CPX *processBlock(CPX *in, int bufSize) {
        if (pebble_demod_is_passthrough(h) == 1) return in;                // USB/LSB/DSB/CW/DIG: `return in` (demod.cpp:127-138)
        if (pebble_demod_process(h, bufSize, pebble_dropin::dp(in), pebble_dropin::dp(out)) < 0) return in;
        return out;
    }
private:
    pebble_demod_t *h;
    DeviceInterface::DemodMode m_mode;
    int m_lastLock = -1;
};

// ---- CFractResampler — pebblelib/fractresampler.h:20-40 ------------------------------------------------------------------------------
class CFractResampler {
public:
    CFractResampler() : h(nullptr) {}
    virtual ~CFractResampler() { pebble_resampler_destroy(h); }
    void Init(int MaxInputSize) { pebble_resampler_destroy(h); h = pebble_resampler_create(MaxInputSize); }
    int Resample(int InLength, TYPEREAL Rate, TYPECPX *pInBuf, TYPECPX *pOutBuf) {
        const int n = pebble_resampler_resample(h, InLength, Rate, pebble_dropin::dp(pInBuf), pebble_dropin::dp(pOutBuf));
        return n < 0 ? 0 : n;
    }
private:
    pebble_resampler_t *h;
};

#endif  // PEBBLE_DROPIN_HPP
