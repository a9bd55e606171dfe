/*
 * pebble_cuda.h — C ABI of libpebblecuda.so: PebbleSDR's receiver IQ hot path on B200 (sm_100a).
 *
 * This is the drop-in boundary. Every entry point replaces one method of a pebblelib /
 * application operator; the reference interface it replaces is cited as file:line under the
 * reference tree. Plain pointers and sizes only; no C++/torch types cross this boundary.
 *
 * Conventions
 *   - CPX is std::complex<double> (pebblelib/cpx.h:96): buffers are interleaved re,im doubles, so a
 *     `double*` of n complex samples holds 2n doubles.
 *   - `*_process(h, n, in, out)` takes HOST pointers (the reference call shape); the library stages
 *     through pinned memory on the handle's stream and returns when `out` is valid.
 *   - `*_process_device(h, n, d_in, d_out, stream)` takes DEVICE pointers and enqueues on `stream`
 *     (a cudaStream_t passed as void*; NULL = the legacy default stream) without synchronising.
 *     The returned count is computed on the host (block arithmetic is deterministic).
 *   - Return values: sample counts are >= 0. Errors are negative: PEBBLE_ERR_*; the text of the last
 *     error on the calling thread is pebble_last_error(). There is NO CPU fallback: without a
 *     usable CUDA device every create/process call fails with PEBBLE_ERR_CUDA.
 *   - One handle = one stream of carried state, not thread-safe (the reference runs one DSP thread
 *     per receiver). Setters take effect at the next process call.
 */
#ifndef PEBBLE_CUDA_H
#define PEBBLE_CUDA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PEBBLE_ERR_ARG (-1)   /* bad argument (NULL handle, n not a multiple the chain requires, ...) */
#define PEBBLE_ERR_CUDA (-2)  /* CUDA runtime/driver error, or no device */
#define PEBBLE_ERR_STATE (-3) /* object not configured (e.g. process before SetDataRate) */

/* ---- library / device ------------------------------------------------------------------------- */
const char *pebble_last_error(void);
int pebble_abi_version(void);
int pebble_device_count(void);          /* >=0, or PEBBLE_ERR_CUDA */
int pebble_set_device(int ordinal);     /* 0 or error; handles bind to the device current at create */
int pebble_device_synchronize(void);
/* Number of kernel launches this library has issued in this process (bench.py's gpu_launches). */
uint64_t pebble_launch_count(void);

/* ---- Device sample formats: DeviceInterfaceBase::normalizeIQ ------------------------------------------------------------
 * Replaces pebblelib/deviceinterfacebase.cpp:648-878 (block forms of normalizeIQ) for the device sample structs of
 * pebblelib/cpx.h:43-96: CPX16 (qint16 re,im), CPXU8 (quint8, offset binary: 128 = 0), CPX8 (qint8), CPXFLOAT, CPX.
 *   out = code * scale,  scale = 1/32768.0 (CS16) | 1/128.0 (CU8: (code-128), CS8) | 1 (CF32, CPX), times user_gain * norm_gain
 *   (m_userIQGain, m_normalizeIQGain, deviceinterfacebase.h:73-74); order = DeviceInterface::IQOrder (device_interfaces.h:140-145);
 *   reverse swaps IQ <-> QI as the `_reverse` argument does (:653-657).
 * pebble_normalize_iq[_device]() is the conversion on its own: bit-identical to the reference for every code.
 * pebble_{downconvert,receiver,channelizer}_set_input_format() make the process calls of that object take RAW device
 * samples (`in` then points at n samples of that format, 16-byte aligned for the device-pointer forms): the conversion is
 * fused into the mixer's load stage — 2..8 bytes per sample cross PCIe and HBM instead of 16 — with `scale` folded into the
 * oscillator amplitude (one rounding apart from convert-then-mix unless scale is a power of two, as it is for unit gains).
 * Chains/orders without a fused kernel convert into an internal CPX buffer first (same results as the stand-alone call). */
enum { PEBBLE_IQ_CPX = 0, PEBBLE_IQ_CS16 = 1, PEBBLE_IQ_CU8 = 2, PEBBLE_IQ_CS8 = 3, PEBBLE_IQ_CF32 = 4 };
enum { PEBBLE_IQO_IQ = 0, PEBBLE_IQO_QI = 1, PEBBLE_IQO_IONLY = 2, PEBBLE_IQO_QONLY = 3 };
int pebble_iq_bytes_per_sample(int fmt);
int pebble_normalize_iq(int fmt, int order, double user_gain, double norm_gain, int reverse, long long n, const void *in,
                        double *out);
int pebble_normalize_iq_device(int fmt, int order, double user_gain, double norm_gain, int reverse, long long n,
                               const void *d_in, double *d_out, void *stream);

/* ---- CFastFIR : overlap-save FFT band-pass ------------------------------------------------------
 * Replaces pebblelib/fastfir.h:50-77 (class CFastFIR), implementation fastfir.cpp:77-334.
 * pebble_fastfir_create()         <-> CFastFIR::CFastFIR()              fastfir.cpp:77-145 (2048/1025)
 * pebble_fastfir_create_sized()   <-> same algorithm with CONV_FFT_SIZE/CONV_FIR_SIZE as parameters
 *                                     (fastfir.cpp:65-66 are compile-time in the reference; BASELINE
 *                                     config 5 needs 131072/65537). fir_size must be fft_size/2+1.
 *                                     fft_size in {2048, 4096, 8192}: one block per CTA in shared memory;
 *                                     fft_size in {32768, 131072}: four-step transform (fastfir_long.cu), all
 *                                     entry points below (incl. set_streams / process_batch_device) unchanged.
 * pebble_fastfir_setup()          <-> CFastFIR::SetupParameters         fastfir.cpp:191-272
 *      returns 0 = new filter installed, 1 = unchanged (fastfir.cpp:195-199),
 *      2 = "Filter Parameter error": old filter kept, new parameters remembered (fastfir.cpp:208-216)
 * pebble_fastfir_process()        <-> CFastFIR::ProcessData             fastfir.cpp:281-319
 */
typedef struct pebble_fastfir pebble_fastfir_t;
pebble_fastfir_t *pebble_fastfir_create(void);
pebble_fastfir_t *pebble_fastfir_create_sized(int fft_size, int fir_size);
void pebble_fastfir_destroy(pebble_fastfir_t *h);
int pebble_fastfir_setup(pebble_fastfir_t *h, double f_lo, double f_hi, double offset, double sample_rate);
int pebble_fastfir_process(pebble_fastfir_t *h, int n, const double *in, double *out);
/* d_in and d_out must not overlap (PEBBLE_ERR_ARG): blocks are filtered concurrently. The host-pointer call may be in place. */
int pebble_fastfir_process_device(pebble_fastfir_t *h, int n, const double *d_in, double *d_out, void *stream);
/* Output count a process call with n inputs WOULD return, without running it. */
int pebble_fastfir_out_count(const pebble_fastfir_t *h, int n);
/* Carried state = the (fir_size-1)-sample overlap + the partially filled block (m_InBufInPos).
 * Layout: int64 fill, then (fir_size-1 + fill) CPX. *_state_size returns bytes. */
int64_t pebble_fastfir_state_size(const pebble_fastfir_t *h);
int pebble_fastfir_get_state(pebble_fastfir_t *h, void *buf, int64_t bytes);
int pebble_fastfir_set_state(pebble_fastfir_t *h, const void *buf, int64_t bytes);
/* Filter spectrum as the reference holds it (natural FFTW bin order), fft_size CPX, host buffer. */
int pebble_fastfir_get_filter(pebble_fastfir_t *h, double *coef_out);

/* ---- CDownConvert : NCO mix + decimate-by-2 cascade ------------------------------------------------
 * Replaces pebblelib/downconvert.h:28-37, implementation downconvert.cpp:64-535.
 * pebble_downconvert_set_data_rate()        <-> CDownConvert::SetDataRate        downconvert.cpp:139-213
 * pebble_downconvert_set_data_rate_simple() <-> CDownConvert::SetDataRateSimple  downconvert.cpp:220-244
 * pebble_downconvert_set_frequency()        <-> CDownConvert::SetFrequency       downconvert.cpp:102-113
 * pebble_downconvert_set_cw_offset()        <-> CDownConvert::SetCwOffset        downconvert.h:34
 * pebble_downconvert_process()              <-> CDownConvert::ProcessData        downconvert.cpp:257-334
 *      n must be a multiple of 2^stages. The reference mutates `in` (mix and decimate in place,
 *      downconvert.cpp:309-310,325); this library leaves `in` untouched (documented deviation F6:
 *      nothing downstream reads the clobbered buffer, receiver.cpp:908).
 * The reference's MAX_DECSTAGES=10 leaves no NULL sentinel for a 10-stage chain (downconvert.h:23);
 * this library plans up to 16 stages and runs up to 12 (2^12 = /4096; longer chains: PEBBLE_ERR_STATE).
 * Short calls (documented deviation): a reference half-band stage that receives fewer samples than it has taps skips
 * the filter and returns n/2 ("safety net", downconvert.cpp:362-363), and one that receives fewer than 2*(taps-1)
 * re-reads its carried history from the caller's buffer after its in-place outputs overwrote it (:387-390), so the
 * reference's output depends on how the stream is cut into calls (framesPerBuffer 2048 at 61.44 MSPS leaves 4 samples
 * for the 35-tap stage; 32768-sample calls leave 64 < 68). This library filters every stage for every call length:
 * any split of a stream gives what the reference gives for calls long enough (>= 2*(taps-1) samples into every
 * stage). pebble_downconvert_set_strict_lengths(h, 1) instead refuses shorter calls with PEBBLE_ERR_ARG.
 */
typedef struct pebble_downconvert pebble_downconvert_t;
pebble_downconvert_t *pebble_downconvert_create(void);
void pebble_downconvert_destroy(pebble_downconvert_t *h);
double pebble_downconvert_set_data_rate(pebble_downconvert_t *h, double in_rate, double max_bw);
double pebble_downconvert_set_data_rate_simple(pebble_downconvert_t *h, double in_rate, double max_bw);
void pebble_downconvert_set_frequency(pebble_downconvert_t *h, double nco_freq);
void pebble_downconvert_set_cw_offset(pebble_downconvert_t *h, double offset);
int pebble_downconvert_set_strict_lengths(pebble_downconvert_t *h, int on);
int pebble_downconvert_num_stages(const pebble_downconvert_t *h);
/* Host-only planner, no device needed: the stage list CDownConvert::SetDataRate (simple = 0, downconvert.cpp:139-213)
 * or SetDataRateSimple (simple = 1, :220-244) builds for (in_rate, max_bw). Writes up to `max` FIR lengths
 * (3 = CIC3), stores the output rate in *out_rate (may be NULL) and returns the stage count. */
int pebble_downconvert_plan(double in_rate, double max_bw, int simple, int *lens, int max, double *out_rate);
/* Host-only: the planned cascade as ONE decimate-by-2^S FIR, y[n] = sum_T h[T] x[2^S n + c - T] (every stage is an FIR:
 * CIC3 downconvert.cpp:519-535, HB11 :421-496, generic half-band :357-393 incl. its doubled oldest tap). This is the
 * prototype filter of the polyphase channelizer front end. Writes min(len, max) taps to h (may be NULL), *c = alignment,
 * returns len. */
int pebble_downconvert_composite_response(double in_rate, double max_bw, int simple, double *h, int max, int *c);
/* Stage description for tests: writes up to max entries of the FIR length per stage
 * (3 = CIC3, 11 = unrolled 11-tap half-band, 15..51 = generic half-band). Returns stage count. */
int pebble_downconvert_stage_lengths(const pebble_downconvert_t *h, int *lens, int max);
/* `in` holds n samples of the format set by pebble_downconvert_set_input_format (default PEBBLE_IQ_CPX = CPX*). */
int pebble_downconvert_set_input_format(pebble_downconvert_t *h, int fmt, int order, double user_gain, double norm_gain);
int pebble_downconvert_input_bytes_per_sample(const pebble_downconvert_t *h);
int pebble_downconvert_process(pebble_downconvert_t *h, int n, const void *in, double *out);
/* d_in and d_out must not overlap (PEBBLE_ERR_ARG; the same holds for pebble_decimator_process_device). */
int pebble_downconvert_process_device(pebble_downconvert_t *h, int n, const void *d_in, double *d_out,
                                      void *stream);
int64_t pebble_downconvert_state_size(const pebble_downconvert_t *h);
int pebble_downconvert_get_state(pebble_downconvert_t *h, void *buf, int64_t bytes);
int pebble_downconvert_set_state(pebble_downconvert_t *h, const void *buf, int64_t bytes);

/* ---- Decimator : the reference's alternative decimation chain (Matlab-rounded half-bands, merged stages) -----------
 * Replaces pebblelib/decimator.h:228-284, implementation decimator.cpp:6-1339 (live when Receiver::m_useDemodDecimator,
 * receiver.cpp:194-195,910-911: Mixer::processBlock then Decimator::process).
 * pebble_decimator_create()          <-> Decimator::Decimator(sampleRate, bufferSize)   decimator.cpp:6-45
 * pebble_decimator_build_chain()     <-> Decimator::buildDecimationChain                decimator.cpp:64-149
 *      returns the decimated rate (float, as the reference); consecutive steps that pick the same design are MERGED into
 *      one stage that filters once and keeps every 2^k-th sample (m_combineStages, decimator.cpp:130-143).
 * pebble_decimator_dec_by2_stages()  <-> Decimator::decBy2Stages()                      decimator.h:241
 * pebble_decimator_process()         <-> Decimator::process                             decimator.cpp:152-226
 *      n must be a multiple of the total decimation and large enough that every half-band stage sees >= taps samples;
 *      otherwise PEBBLE_ERR_ARG. (The reference degrades to plain decimation for such calls and refills its delay
 *      line from beyond the valid input, decimator.cpp:605-626; that path is not reproduced.) No chain -> copy.
 *      in and out must not overlap. Results are bit-identical to the reference's scalar arithmetic.
 * pebble_decimator_plan(): host-only planner (no device): per merged stage taps (0 = CIC3) and decimation. */
typedef struct pebble_decimator pebble_decimator_t;
pebble_decimator_t *pebble_decimator_create(unsigned sample_rate, unsigned buffer_size);
void pebble_decimator_destroy(pebble_decimator_t *h);
float pebble_decimator_build_chain(pebble_decimator_t *h, unsigned rate_in, unsigned protect_bw, unsigned rate_out);
int pebble_decimator_dec_by2_stages(const pebble_decimator_t *h);
int pebble_decimator_total_decimation(const pebble_decimator_t *h);
int pebble_decimator_chain(const pebble_decimator_t *h, int *taps, int *decimate, int max);
int pebble_decimator_plan(unsigned rate_in, unsigned protect_bw, unsigned rate_out, int *taps, int *decimate, int max,
                          float *out_rate, int *dec_by2_stages);
int pebble_decimator_out_count(const pebble_decimator_t *h, int n);
int pebble_decimator_process(pebble_decimator_t *h, int n, const double *in, double *out);
int pebble_decimator_process_device(pebble_decimator_t *h, int n, const double *d_in, double *d_out, void *stream);
/* Carried state = the last `hist` RAW input samples (oldest first): every stage is a pure FIR, so all delay lines
 * (m_lastXVDsp, m_xOdd/m_xEven) are re-derived from them. *_state_size returns bytes. */
int64_t pebble_decimator_state_size(const pebble_decimator_t *h);
int pebble_decimator_get_state(pebble_decimator_t *h, void *buf, int64_t bytes);
int pebble_decimator_set_state(pebble_decimator_t *h, const void *buf, int64_t bytes);

/* ---- Mixer : frequency shift by the recursive oscillator ------------------------------------------------------------
 * Replaces pebblelib/mixer.h, implementation mixer.cpp:5-81 (and the NCO primitive nco.cpp:16-45 it inlines).
 * pebble_mixer_set_frequency() <-> Mixer::setFrequency  mixer.cpp:25-39  (m_frequency = -f; oscillator restarts at (1,0))
 * pebble_mixer_process()       <-> Mixer::processBlock  mixer.cpp:48-81  (f == 0: the reference returns `in` itself;
 *                                  here the samples are copied to out and pebble_mixer_is_passthrough() reports 1) */
typedef struct pebble_mixer pebble_mixer_t;
pebble_mixer_t *pebble_mixer_create(unsigned sample_rate, unsigned num_samples);
void pebble_mixer_destroy(pebble_mixer_t *h);
void pebble_mixer_set_frequency(pebble_mixer_t *h, double f);
int pebble_mixer_is_passthrough(const pebble_mixer_t *h);
int pebble_mixer_process(pebble_mixer_t *h, int n, const double *in, double *out);
int pebble_mixer_process_device(pebble_mixer_t *h, int n, const double *d_in, double *d_out, void *stream);

/* ---- NCO : the quadrature oscillator primitive -------------------------------------------------------------------------------
 * Replaces pebblelib/nco.h:20-60, nco.cpp:5-81.
 * pebble_nco_set_frequency() <-> NCO::setFrequency (nco.cpp:16-45): the "safety net" limits |f| to (sampleRate / 2) - 1
 *      (integer arithmetic) and the oscillator restarts at (1, 0); returns the frequency actually set. (Mixer::setFrequency,
 *      mixer.cpp:25-39, and CDownConvert::SetFrequency, downconvert.cpp:102-113, do NOT clamp in the reference, and do not here.)
 * pebble_nco_gen_single() <-> NCO::genSingle (nco.cpp:65-81): n x nextSample() (nco.h:29-38) scaled by `gain`, written over
 *      (mix == 0) or added to (mix != 0) the buffer. */
typedef struct pebble_nco pebble_nco_t;
pebble_nco_t *pebble_nco_create(unsigned sample_rate, unsigned buffer_size);
void pebble_nco_destroy(pebble_nco_t *h);
double pebble_nco_set_frequency(pebble_nco_t *h, double f);
double pebble_nco_frequency(const pebble_nco_t *h);
int pebble_nco_gen_single(pebble_nco_t *h, double *io, int n, double gain, int mix);
int pebble_nco_gen_single_device(pebble_nco_t *h, double *d_io, int n, double gain, int mix, void *stream);

/* ---- CpxUtil : element-wise CPX vector helpers on DEVICE buffers ---------------------------------------------------------------
 * Replace pebblelib/cpx.h:161-196, cpx.cpp:35-112 for callers that keep CPX buffers on the device between operator calls
 * (the fused chain does not need them). copyCPX, scaleCPX, multCPX, addCPX, magCPX / sqrMagCPX (which keep in.re in out.im,
 * cpx.cpp:82-96), clearCPX, decimateCPX (out[j] = in[j*by]; returns the output count). Results are bit-identical to the
 * host functions. n counts CPX samples. */
int pebble_cpx_copy_device(double *d_out, const double *d_in, long long n, void *stream);
int pebble_cpx_scale_device(double *d_out, const double *d_in, double a, long long n, void *stream);
int pebble_cpx_mult_device(double *d_out, const double *d_in, const double *d_in2, long long n, void *stream);
int pebble_cpx_add_device(double *d_out, const double *d_in, const double *d_in2, long long n, void *stream);
int pebble_cpx_mag_device(double *d_out, const double *d_in, long long n, void *stream);
int pebble_cpx_sqrmag_device(double *d_out, const double *d_in, long long n, void *stream);
int pebble_cpx_clear_device(double *d_out, long long n, void *stream);
int pebble_cpx_decimate_device(double *d_out, const double *d_in, int by, long long n, void *stream);

/* ---- FFT / SignalSpectrum : windowed FFT, unfold, 2-tap averaged amplitude dB -------------------------
 * Replaces pebblelib/fft.h:29-62 (class FFT + FFTfftw, fftw.cpp:17-81) and the FFT use of
 * application/signalspectrum.cpp:54-113.
 * pebble_fft_params()    <-> FFT::fftParams       fft.cpp:67-118  (the reference clamps the size to [2048, 65535] and accepts
 *                            any value in between; here sizes below 2048 are raised to 2048 like there, and the supported
 *                            sizes are 2048, 4096 and 8192 — the transforms that fit one CTA's shared memory; 16384 and
 *                            32768 and non-powers of two return PEBBLE_ERR_ARG)
 * pebble_fft_forward()   <-> FFTfftw::fftForward  fftw.cpp:31-46  (window applied iff n==samples_per_buffer
 *                            and window!=NONE, fft.cpp:129-157; zero-padded; natural bin order)
 * pebble_fft_inverse()   <-> FFTfftw::fftInverse  fftw.cpp:49-67  (unnormalised)
 * pebble_fft_spectrum()  <-> FFTfftw::fftSpectrum fftw.cpp:70-81 + FFT::calcPowerAverages fft.cpp:324-399
 *                            returns the overload flag (1/0), writes fft_size doubles (amplitude dB).
 * pebble_fft_spectrum_batch_device(): `frames` consecutive frames of n samples each; the 2-tap average
 *                            state is carried frame to frame exactly as successive calls would.
 * window_type uses WindowFunction::WINDOWTYPE numbering (windowfunction.h): see PEBBLE_WINDOW_*.
 */
enum {
    PEBBLE_WINDOW_RECTANGULAR = 1, PEBBLE_WINDOW_HANNING = 2, PEBBLE_WINDOW_WELCH = 3,
    PEBBLE_WINDOW_PARZEN = 4, PEBBLE_WINDOW_BARTLETT = 5, PEBBLE_WINDOW_HAMMING = 6,
    PEBBLE_WINDOW_BLACKMAN2 = 7, PEBBLE_WINDOW_BLACKMAN3 = 8, PEBBLE_WINDOW_BLACKMAN4 = 9,
    PEBBLE_WINDOW_EXPONENTIAL = 10, PEBBLE_WINDOW_RIEMANN = 11, PEBBLE_WINDOW_BLACKMANHARRIS = 12,
    PEBBLE_WINDOW_BLACKMAN = 13, PEBBLE_WINDOW_NONE = 14
};
typedef struct pebble_fft pebble_fft_t;
pebble_fft_t *pebble_fft_create(void);
void pebble_fft_destroy(pebble_fft_t *h);
int pebble_fft_params(pebble_fft_t *h, unsigned fft_size, double db_compensation, double sample_rate,
                      int samples_per_buffer, int window_type);
int pebble_fft_size(const pebble_fft_t *h);
int pebble_fft_forward(pebble_fft_t *h, const double *in, double *out, int n);
int pebble_fft_inverse(pebble_fft_t *h, const double *in, double *out, int n);
int pebble_fft_spectrum(pebble_fft_t *h, const double *in, double *out_db, int n);
int pebble_fft_spectrum_batch_device(pebble_fft_t *h, int frames, int n, const double *d_in, double *d_out_db,
                                     int *d_overload_flags, void *stream);
/* Window table + coherent gain exactly as WindowFunction::regenerate builds them (float-rounded
 * constants, windowfunction.cpp:51-235). w_out may be NULL. */
double pebble_window(int window_type, int n, double *w_out);
/* Which reference FFT class's contract the forward / inverse calls follow (fft.cpp:45-65 selects one at compile time):
 * PEBBLE_FFT_FFTW  = FFTfftw (fftw.cpp): forward e^{-j}, inverse e^{+j}, both unnormalised;
 * PEBBLE_FFT_OOURA = CFft (fftcute.cpp:183-277): fftForward swaps I and Q before Ooura's transform (:194-201), which makes
 *                    its output the FFTW_FORWARD result with re and im exchanged; fftInverse (bitrv2conj + cftbsub) is again the
 *                    e^{-j} transform, so inverse(forward(x)) = N * (x with I and Q swapped). fftSpectrum is the same for both. */
enum { PEBBLE_FFT_FFTW = 0, PEBBLE_FFT_OOURA = 1 };
int pebble_fft_set_backend(pebble_fft_t *h, int backend);
/* averaging state (m_fftPower, m_fftAmplitude: 2*fft_size doubles) */
int pebble_fft_get_state(pebble_fft_t *h, double *buf, int64_t count);
int pebble_fft_set_state(pebble_fft_t *h, const double *buf, int64_t count);

/* ---- SignalSpectrum : the display-rate spectrum operator ------------------------------------------------------------------
 * Replaces application/signalspectrum.h (class SignalSpectrum : ProcessStep), signalspectrum.cpp:6-133: two FFT objects
 * (full-rate `unprocessed`, default 4096 bins, and the post-decimation `zoomed` / hi-res one, default 2048 bins), both
 * BLACKMANHARRIS over bufferSize samples, each behind an update gate of 1000 / updatesPerSec milliseconds.
 * create <-> SignalSpectrum(sampleRate, hiResSampleRate, bufferSize) with global->settings {numSpectrumBins,
 * numHiResSpectrumBins, updatesPerSecond} (:8-33); unprocessed / zoomed <-> :63-106 (the first call of each only starts
 * its timer; return 1 = a new spectrum was made, where the reference emits newFftData()); now_ms replaces QElapsedTimer.
 * get_unprocessed / get_hires return the dB arrays (amplitude dB, -f..0..+f, bin_count doubles), valid until the next call. */
typedef struct pebble_spectrum pebble_spectrum_t;
pebble_spectrum_t *pebble_spectrum_create(unsigned sample_rate, unsigned hires_sample_rate, unsigned buffer_size, int num_bins,
                                          int num_hires_bins, int updates_per_sec);
void pebble_spectrum_destroy(pebble_spectrum_t *h);
int pebble_spectrum_set_sample_rate(pebble_spectrum_t *h, unsigned sample_rate, unsigned hires_sample_rate);
void pebble_spectrum_set_hires(pebble_spectrum_t *h, int on);
void pebble_spectrum_set_updates_per_sec(pebble_spectrum_t *h, int updates_per_sec);
int pebble_spectrum_unprocessed(pebble_spectrum_t *h, const double *in, int n, long long now_ms);
int pebble_spectrum_zoomed(pebble_spectrum_t *h, const double *in, int n, long long now_ms);
int pebble_spectrum_bin_count(const pebble_spectrum_t *h);
int pebble_spectrum_hires_bin_count(const pebble_spectrum_t *h);
int pebble_spectrum_overload(const pebble_spectrum_t *h);
int pebble_spectrum_overrun(const pebble_spectrum_t *h);
void pebble_spectrum_ack_update(pebble_spectrum_t *h);
const double *pebble_spectrum_get_unprocessed(const pebble_spectrum_t *h);
const double *pebble_spectrum_get_hires(const pebble_spectrum_t *h);

/* Library extension (no reference counterpart): scalar gains fused into the FastFIR kernel's load and
 * store, used by the receiver chain for receiver.cpp:935-938 (gain restore) and AGC_OFF (agc.cpp:86-94). */
int pebble_fastfir_set_fused_scale(pebble_fastfir_t *h, double in_scale, double out_scale);

/* ---- CFir : Kaiser-windowed time-domain FIR (demod post filters) -------------------------------------
 * Replaces pebblelib/fir.h:31-56. init_lp/init_hp <-> CFir::InitLPFilter / InitHPFilter (fir.cpp:246-337,
 * 352-440; return the tap count), init_const <-> CFir::InitConstFir (fir.cpp:203-228),
 * process <-> CFir::ProcessFilter(int, TYPECPX*, TYPECPX*) (fir.cpp:106-132; I taps on re, Q taps on im). */
typedef struct pebble_cfir pebble_cfir_t;
pebble_cfir_t *pebble_cfir_create(void);
void pebble_cfir_destroy(pebble_cfir_t *h);
int pebble_cfir_init_lp(pebble_cfir_t *h, int num_taps, double scale, double astop, double fpass, double fstop, double fs);
int pebble_cfir_init_hp(pebble_cfir_t *h, int num_taps, double scale, double astop, double fpass, double fstop, double fs);
int pebble_cfir_init_const(pebble_cfir_t *h, int num_taps, const double *icoef, const double *qcoef, double fs);
int pebble_cfir_get_taps(pebble_cfir_t *h, double *out, int max);
int pebble_cfir_process(pebble_cfir_t *h, int n, const double *in, double *out);
int pebble_cfir_process_device(pebble_cfir_t *h, int n, const double *d_in, double *d_out, void *stream);

/* ---- AGC ----------------------------------------------------------------------------------------------
 * Replaces application/agc.h (class AGC : ProcessStep). mode: 0 AGC_OFF, 1 ACG_FAST, 2 AGC_MED, 3 AGC_SLOW,
 * 4 AGC_LONG (agc.h:11). set_mode <-> AGC::setAgcMode (agc.cpp:52-82), set_threshold <-> AGC::setAgcThreshold
 * (agc.cpp:43-50), process <-> AGC::processBlock (agc.cpp:84-207; n replaces the fixed numSamples). */
typedef struct pebble_agc pebble_agc_t;
pebble_agc_t *pebble_agc_create(unsigned sample_rate, unsigned buffer_size);
void pebble_agc_destroy(pebble_agc_t *h);
int pebble_agc_set_mode(pebble_agc_t *h, int mode, int threshold);
int pebble_agc_set_threshold(pebble_agc_t *h, int g);
int pebble_agc_process(pebble_agc_t *h, int n, const double *in, double *out);
int pebble_agc_process_device(pebble_agc_t *h, int n, const double *d_in, double *d_out, void *stream);

/* ---- Demod -------------------------------------------------------------------------------------------
 * Replaces application/demod.h Demod::processBlock (demod.cpp:100-141). mode = DeviceInterface::DemodMode
 * (device_interfaces.h:124-137): 0 dmAM (Demod_AM::processBlockFiltered, demod_am.cpp:40-64), 2 dmFMN
 * (Demod_NFM::processBlockNCO, demod_nfm.cpp:225-257; variant 1 = processBlockFM2, :124-140),
 * 1 dmSAM (Demod_SAM::processBlock, demod/demod_sam.cpp:41-112), 3 dmFMM (Demod::fmMono -> Demod_WFM::processDataMono,
 * demod.cpp:168-175, demod/demod_wfm.cpp:207-231: 75 kHz IIR low-pass at >= 150 kS/s, discriminator, 15 kHz CFir, 75 us
 * de-emphasis, 19 kHz notch; pebble_demod_set_wfm_sample_rate() = the second rate of Demod(sampleRate, wfmSampleRate, n),
 * demod.cpp:50-66, default = input_rate),
 * 5..12 DSB/LSB/USB/CW/DIG/NONE = pass-through (the reference returns `in`; here out is a copy and
 * pebble_demod_is_passthrough() reports it). 4 dmFMS (Demod::fmStereo -> Demod_WFM::processDataStereo, demod_wfm.cpp:258-366):
 * the AUDIO path — discriminator, Hilbert pair, 19 kHz pilot band-pass, pilot PLL (with the reference's own arctan2), stereo
 * demux when locked, 15 kHz CFir, de-emphasis, notch; pebble_demod_get_stereo_lock() = getStereoLock — and the RDS branch
 * (:298-357: CDownConvert by -57 kHz to ~8-20 kS/s, 2400 Hz complex low-pass, RDS PLL, bi-phase matched filter, bit-rate
 * resonator, bit slicer on the GPU; differential decode, (26,16) syndrome check with burst correction and the block / group
 * state machine of processNewRdsBit / checkBlock :584-758 on the host, fed through a page-locked ring so that the audio
 * path never waits for it). pebble_demod_get_next_rds_group() = Demod_WFM::getNextRdsGroupData (:764-788): pops one group,
 * returns 1 when it differs from the last one returned. With RDS on (default, as in the reference) a dmFMS call must be a
 * multiple of the RDS down-converter's decimation (rate / pebble_demod_get_rds_rate()); pebble_demod_set_rds(h, 0) runs the
 * audio path alone. pebble_demod_get_rds_tap() (tests): the last call's m_RdsRaw as the reference leaves it (:336-352),
 * pairs (value at the sampling instant or 0, matched-filter output). rdsdecode.cpp (groups -> text) is not part of the path. */
typedef struct pebble_demod pebble_demod_t;
pebble_demod_t *pebble_demod_create(unsigned input_rate, unsigned num_samples);
void pebble_demod_destroy(pebble_demod_t *h);
int pebble_demod_set_mode(pebble_demod_t *h, int mode);
int pebble_demod_set_nfm_variant(pebble_demod_t *h, int variant);
int pebble_demod_set_wfm_sample_rate(pebble_demod_t *h, double rate);
int pebble_demod_get_stereo_lock(pebble_demod_t *h);
int pebble_demod_set_rds(pebble_demod_t *h, int on);
double pebble_demod_get_rds_rate(const pebble_demod_t *h);
int pebble_demod_get_next_rds_group(pebble_demod_t *h, unsigned short *blocks4);
int pebble_demod_get_rds_tap(pebble_demod_t *h, double *out, int max_samples);
/* The block / group synchroniser of the RDS branch by itself (Demod_WFM::processNewRdsBit + checkBlock, demod_wfm.cpp:584-758):
 * pure host logic, fed with DATA bits (after the differential decode); pebble_rds_sync_next_group as
 * pebble_demod_get_next_rds_group. The demodulator owns one of these; this handle exists for callers that get their bits
 * elsewhere and for the CPU tests. */
typedef struct pebble_rds_sync pebble_rds_sync_t;
pebble_rds_sync_t *pebble_rds_sync_create(void);
void pebble_rds_sync_destroy(pebble_rds_sync_t *h);
int pebble_rds_sync_push_bits(pebble_rds_sync_t *h, const unsigned char *bits, int n);
int pebble_rds_sync_next_group(pebble_rds_sync_t *h, unsigned short *blocks4);
int pebble_demod_set_bandwidth(pebble_demod_t *h, double bandwidth);
int pebble_demod_is_passthrough(const pebble_demod_t *h);
int pebble_demod_process(pebble_demod_t *h, int n, const double *in, double *out);
int pebble_demod_process_device(pebble_demod_t *h, int n, const double *d_in, double *d_out, void *stream);

/* ---- CFractResampler : fractional resampler to the audio rate -----------------------------------------------
 * Replaces pebblelib/fractresampler.h:20-40, implementation fractresampler.cpp:64-187 (complex form) and :196-250 (stereo
 * int16 form), called by Receiver::processIQData after the demodulator (application/receiver.cpp:994-1002) with
 * Rate = demodRate / audioRate.
 * pebble_resampler_create()     <-> CFractResampler() + Init(MaxInputSize)            fractresampler.cpp:64-121
 * pebble_resampler_resample()   <-> Resample(InLength, Rate, TYPECPX*, TYPECPX*)      fractresampler.cpp:146-187 (one call)
 * pebble_resampler_resample_device(): `frame` > 0 reproduces the call structure of a batch: the n samples are resampled as
 *      n/frame consecutive Resample calls of `frame` samples (the receiver calls it once per framesPerBuffer frame; the
 *      time accumulator m_FloatTime is rounded differently depending on where the calls end). streams > 1: row s of
 *      d_in/d_out (in_stride/out_stride CPX apart), all rows with the same Rate. clip != 0 also applies the +-0.9999 clip
 *      of AudioQT::SendToOutput (pebblelib/audioqt.cpp:190-201). Returns the output samples per stream.
 * Output counts and every output sample are bit-identical to the reference (separate multiply and add roundings in its
 * order). pebble_resampler_out_count() = what the next call WOULD return (the times depend only on Rate and lengths).
 * pebble_resampler_plan*(): host-only planner (no device). */
typedef struct pebble_resampler pebble_resampler_t;
pebble_resampler_t *pebble_resampler_create(int max_input_size);
void pebble_resampler_destroy(pebble_resampler_t *h);
int pebble_resampler_set_streams(pebble_resampler_t *h, int streams);
int pebble_resampler_out_count(pebble_resampler_t *h, int n, int frame, double rate);
int pebble_resampler_resample(pebble_resampler_t *h, int n, double rate, const double *in, double *out);
int pebble_resampler_resample_device(pebble_resampler_t *h, int n, int frame, double rate, const double *d_in, long long in_stride,
                                     double *d_out, long long out_stride, int clip, void *stream);
int pebble_resampler_resample_stereo16_device(pebble_resampler_t *h, int n, int frame, double rate, const double *d_in,
                                              long long in_stride, short *d_out, long long out_stride, double gain, void *stream);
int64_t pebble_resampler_state_size(const pebble_resampler_t *h);   /* double m_FloatTime + 28 CPX per stream */
int pebble_resampler_get_state(pebble_resampler_t *h, void *buf, int64_t bytes);
int pebble_resampler_set_state(pebble_resampler_t *h, const void *buf, int64_t bytes);
int pebble_resampler_plan(double float_time, double rate, int n, double *float_time_out);
int pebble_resampler_plan_times(double float_time, double rate, int n, double *times, int max);

/* ---- CIir, DCRemoval, IQBalance, squelch (the cheap steps either side of the path) --------------------------------
 * pebble_iir_*  <-> pebblelib/iir.h:17-45 (class CIir): init <-> InitLP / InitHP / InitBP / InitBR (iir.cpp:50-170), process <->
 *      ProcessFilter(int, TYPECPX*, TYPECPX*) (iir.cpp:191-206; re and im through the same biquad with separate delay values).
 *      DCRemoval (application/dcremoval.cpp:3-18) is pebble_iir_init(h, PEBBLE_IIR_HP, 10, 0.7071, sampleRate) + process; a
 *      disabled DCRemoval copies (dcremoval.cpp:11-14). The biquad is a linear recurrence and is evaluated as a three-level
 *      scan (prefilter.cu); each sample is formed with the reference's own operations, the carried-in state is a re-associated
 *      sum. For filters with poles ~1e-6 from the unit circle (the DC blocker at MHz rates) the reference's direct form 2 itself
 *      keeps ~8 digits, and so does the agreement between the two.
 * pebble_iqbalance_process_device() <-> IQBalance::ProcessBlock (application/iqbalance.cpp:67-88) for n / block_len consecutive
 *      blocks of numSamples (the correction recurrence restarts in every block there). Bit-identical to the reference.
 * pebble_sigstrength_fd_estimate_device() <-> SignalStrength::fdEstimate (application/signalstrength.cpp:289-380) for nch
 *      channels over one spectrum (Receiver: the unprocessed spectrum, receiver.cpp:959-961), without its 100 ms display
 *      timer; out5[k] = {peakDb, avgDb, snrDb, floorDb, open}, open = avgDb >= squelch_db (receiver.cpp:963-966). */
enum { PEBBLE_IIR_LP = 0, PEBBLE_IIR_HP = 1, PEBBLE_IIR_BP = 2, PEBBLE_IIR_BR = 3 };
typedef struct pebble_iir pebble_iir_t;
pebble_iir_t *pebble_iir_create(void);
void pebble_iir_destroy(pebble_iir_t *h);
int pebble_iir_init(pebble_iir_t *h, int kind, double f0, double q, double sample_rate);
int pebble_iir_get_coefficients(const pebble_iir_t *h, double *a1a2b0b1b2);
int pebble_iir_set_coefficients(pebble_iir_t *h, const double *a1a2b0b1b2);   /* raw biquad (m_A1.. are public members of CIir, iir.h) */
int pebble_iir_process(pebble_iir_t *h, int n, const double *in, double *out);
int pebble_iir_process_device(pebble_iir_t *h, int n, const double *d_in, double *d_out, void *stream);
int pebble_iir_get_state(pebble_iir_t *h, double *w4);
int pebble_iir_set_state(pebble_iir_t *h, const double *w4);
int pebble_iqbalance_process_device(int n, int block_len, double gain_factor, double phase_factor, const double *d_in, double *d_out,
                                    void *stream);
int pebble_sigstrength_fd_estimate_device(const double *d_spectrum, int bins, unsigned spectrum_rate, float bp_lo, float bp_hi,
                                          const double *mixer_freqs, int nch, double squelch_db, double *out5, void *stream);
/* pebble_sigstrength_fd_estimate_noalloc_device(): NO counterpart in the reference — added for
 *      callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). Same arguments,
 *      same kernel and same results as pebble_sigstrength_fd_estimate_device(), but it calls no cudaMalloc / cudaFree / allocating
 *      routine: the scratch the other entry point allocates per call (per-channel bin ranges in, per-channel sums out) lives in
 *      one file-scope device array of PEBBLE_FD_NOALLOC_MAX_CH entries (48 KiB per GPU that loads the library, held until exit),
 *      host temporaries are on the stack (52 KiB). nch > PEBBLE_FD_NOALLOC_MAX_CH returns PEBBLE_ERR_ARG with a message naming
 *      the limit. ALL calls share that array: calls whose GPU work overlaps (other streams / threads) race on it; the call
 *      synchronises `stream` before returning, so consecutive calls from one thread are safe. Nothing in the array needs
 *      clearing between calls (every entry read was written by the same call). */
#define PEBBLE_FD_NOALLOC_MAX_CH 1024
int pebble_sigstrength_fd_estimate_noalloc_max_channels(void);
int pebble_sigstrength_fd_estimate_noalloc_device(const double *d_spectrum, int bins, unsigned spectrum_rate, float bp_lo, float bp_hi,
                                                  const double *mixer_freqs, int nch, double squelch_db, double *out5, void *stream);

/* ---- NoiseBlanker ----------------------------------------------------------------------------------------------
 * Replaces application/noiseblanker.h: ProcessBlock (NB1: a sample above 3.3 x the running magnitude average starts 7 zeroed
 * outputs, otherwise the input delayed by two samples; noiseblanker.cpp:47-77) and ProcessBlock2 (NB2: the running complex
 * average where the sample exceeds 3.3 x the average, else the input; :80-98), called back to back at the device rate
 * (receiver.cpp:820-822). which = 1 / 2 selects the block. The running average is a float updated in double arithmetic, so the
 * walk is serial and bit-faithful (decisions identical to the reference's). Disabled (the reference's default): the reference
 * returns `in` itself; here out is a copy. In-place calls are refused. */
typedef struct pebble_noiseblanker pebble_noiseblanker_t;
pebble_noiseblanker_t *pebble_noiseblanker_create(unsigned sample_rate, unsigned buffer_size);
void pebble_noiseblanker_destroy(pebble_noiseblanker_t *h);
int pebble_noiseblanker_set_nb_enabled(pebble_noiseblanker_t *h, int on);    /* setNbEnabled, noiseblanker.cpp:21-29 */
int pebble_noiseblanker_set_nb2_enabled(pebble_noiseblanker_t *h, int on);   /* setNb2Enabled, :30-37 */
int pebble_noiseblanker_enabled(const pebble_noiseblanker_t *h, int which);
int pebble_noiseblanker_process(pebble_noiseblanker_t *h, int which, int n, const double *in, double *out);
int pebble_noiseblanker_process_device(pebble_noiseblanker_t *h, int which, int n, const double *d_in, double *d_out, void *stream);

/* ---- Receiver chain : Receiver::processIQData, both branches ---------------------------------------------
 * Non-WFM (application/receiver.cpp:903-996): m_downConvert1.ProcessData -> m_sampleBuf re-framing to
 * framesPerBuffer (:922-931) -> scaleCPX gain restore (:935-938) -> m_bpFilter->process (:950) ->
 * m_agc->processBlock (:983) -> m_demod->processBlock (:987). Returns the number of demodulated samples
 * written (a multiple of framesPerBuffer; 0 while a frame is still filling). SignalSpectrum, squelch,
 * noise filter and modem plugins are separate entry points.
 * WFM (demod mode dmFMM / dmFMS, receiver.cpp:854-901): m_downConvertWfm1 (SetDataRateSimple(sampleRate, 200000), :217:
 * 51-tap stages until the rate is below 400 kS/s; pebble_receiver_get_wfm_rate) -> re-framing -> m_demod->processBlock
 * (Demod::fmMono / fmStereo at the WFM rate, one call per frame) — no band-pass, gain restore or AGC on this branch —
 * and, with pebble_receiver_set_audio, the resampler at wfmRate / audioRate (:901). Switching between the branches drops the
 * partial frame (they decimate to different rates). pebble_receiver_demod() is the receiver's own demodulator, for
 * pebble_demod_get_stereo_lock / pebble_demod_get_next_rds_group after a call (demod.cpp:196-219); it stays owned by the receiver
 * and is replaced by pebble_receiver_set_data_rate. */
typedef struct pebble_receiver pebble_receiver_t;
pebble_receiver_t *pebble_receiver_create(double sample_rate, int frames_per_buffer);
void pebble_receiver_destroy(pebble_receiver_t *h);
double pebble_receiver_set_data_rate(pebble_receiver_t *h, double max_bw);
void pebble_receiver_set_mixer(pebble_receiver_t *h, double freq);
int pebble_receiver_set_filter(pebble_receiver_t *h, double lo, double hi);
int pebble_receiver_set_demod_mode(pebble_receiver_t *h, int mode);
int pebble_receiver_set_demod_bandwidth(pebble_receiver_t *h, double bandwidth);
int pebble_receiver_set_agc(pebble_receiver_t *h, int mode, int threshold);
int pebble_receiver_set_gain_stages(pebble_receiver_t *h, int stages);
/* Raw device samples into the chain (normalizeIQ fused into the mixer load, see "Device sample formats" above). */
int pebble_receiver_set_input_format(pebble_receiver_t *h, int fmt, int order, double user_gain, double norm_gain);
/* Audio tail (receiver.cpp:994-1007): CFractResampler::Resample once per frame with Rate = (int)demodRate / audio_rate, and,
 * with clip != 0, the audio sink's +-0.9999 clip (AudioQT::SendToOutput, pebblelib/audioqt.cpp:190-201). The process calls
 * then return AUDIO samples (a data-independent count, pebble_receiver_out_count). audio_rate <= 0 switches it off. */
int pebble_receiver_set_audio(pebble_receiver_t *h, int audio_rate, int clip);
double pebble_receiver_get_wfm_rate(const pebble_receiver_t *h);
pebble_demod_t *pebble_receiver_demod(pebble_receiver_t *h);
int pebble_receiver_out_count(const pebble_receiver_t *h, int n);
/* Library extension (the reference's Perform timer call sites, receiver.cpp:788,805,827, are commented out):
 * per-step DEVICE time via CUDA events recorded on the caller's stream. get_profile sums the calls made since
 * profiling was enabled (ms[0] CDownConvert, ms[1] CFastFIR, ms[2] AGC, ms[3] Demod), then resets. */
int pebble_receiver_set_profiling(pebble_receiver_t *h, int on);
int pebble_receiver_get_profile(pebble_receiver_t *h, double *ms4, int64_t *calls);
int pebble_receiver_process(pebble_receiver_t *h, int n, const void *in, double *out);
int pebble_receiver_process_device(pebble_receiver_t *h, int n, const void *d_in, double *d_out, void *stream);
/* in/out are page-locked HOST buffers; copies and kernels are enqueued without synchronising and `out` is complete when `stream`
 * is. The upload runs in 2^24-sample chunks on an internal copy stream into two alternating device buffers; the chain of chunk c
 * runs on `stream` under the copy of chunk c+1 (the chain carries its state: chunked == one call up to the roundings of the
 * down-converter's per-call oscillator seeds, 1e-16). `in` must stay untouched until `stream` has passed this call. */
int pebble_receiver_process_pinned(pebble_receiver_t *h, int n, const void *in, double *out, void *stream);

/* ---- Batched forms (library extensions for BASELINE config 4; no single reference counterpart) -------------------------
 * nch CDownConvert objects with the same (InRate, MaxBW) and one NCO each, run in one launch over ONE input stream:
 * set_channels <-> nch x CDownConvert::CDownConvert + SetDataRate, set_channel_frequency <-> SetFrequency of object
 * `ch`, process_channels_device <-> ProcessData of every object on the same input; outputs are [channel][n >> S] rows
 * `out_stride` CPX apart. Needs a chain that starts with CIC3 (in_rate >= ~667 x max_bw), else PEBBLE_ERR_STATE. */
int pebble_downconvert_set_channels(pebble_downconvert_t *h, int nch);
int pebble_downconvert_set_channel_frequency(pebble_downconvert_t *h, int ch, double nco_freq);
int pebble_downconvert_process_channels_device(pebble_downconvert_t *h, int n, const void *d_in, double *d_out,
                                               long long out_stride, void *stream);
/* 1 if the last batched call ran the polyphase-DFT front end (channels on the uniform grid fs / 2^S: one mix + polyphase FIR
 * + 2^S-point DFT for ALL channels, chz_poly.cu), 0 if it ran one down-converter per channel, -1 before the first call. */
int pebble_downconvert_batch_path(const pebble_downconvert_t *h);
/* `streams` independent CFastFIR objects sharing one filter: stream s reads row s of d_in (rows in_stride CPX apart). */
int pebble_fastfir_set_streams(pebble_fastfir_t *h, int streams);
int pebble_fastfir_process_batch_device(pebble_fastfir_t *h, int n, const double *d_in, long long in_stride, double *d_out,
                                        long long out_stride, void *stream);

/* ---- Channelizer: nch receiver chains (receiver.cpp:903-996) over one wideband capture, BASELINE config 4 ----------------
 * Every channel = CDownConvert(own mixer frequency) -> gain restore -> CFastFIR(lo,hi) -> AGC_OFF(threshold) -> Demod(mode),
 * framesPerBuffer 2048. Channel-major device output: row `ch` holds that channel's demodulated samples. Channels shard
 * across GPUs by giving each rank its own channelizer over its channel range (no data-path collective). */
typedef struct pebble_channelizer pebble_channelizer_t;
pebble_channelizer_t *pebble_channelizer_create(int nchannels, double in_rate, double max_bw);
void pebble_channelizer_destroy(pebble_channelizer_t *h);
double pebble_channelizer_out_rate(const pebble_channelizer_t *h);
int pebble_channelizer_num_stages(const pebble_channelizer_t *h);
int pebble_channelizer_num_channels(const pebble_channelizer_t *h);
int pebble_channelizer_set_frequency(pebble_channelizer_t *h, int ch, double freq);
int pebble_channelizer_set_filter(pebble_channelizer_t *h, double lo, double hi);
int pebble_channelizer_set_demod_mode(pebble_channelizer_t *h, int ch, int mode);   /* 0 dmAM, 2 dmFMN (PLL), >= 5 pass-through */
int pebble_channelizer_set_am_bandwidth(pebble_channelizer_t *h, double bandwidth);
int pebble_channelizer_set_agc_threshold(pebble_channelizer_t *h, int threshold);
int pebble_channelizer_out_count(const pebble_channelizer_t *h, int n);
/* Squelch per channel (receiver.cpp:955-965: a frame whose band average is below the squelch level returns before AGC and
 * demodulator). open[ch] == 0: channel ch yields zeros and its demodulator / post-filter state does not advance; NULL: all open.
 * _from_spectrum: SignalStrength::fdEstimate (signalstrength.cpp:289-380) for every channel over ONE wideband dB spectrum on the
 * device, with the channelizer's band-pass edges and channel centres; open = avgDb >= squelch_db; returns the open count. */
int pebble_channelizer_set_squelch(pebble_channelizer_t *h, const unsigned char *open, int nch);
int pebble_channelizer_squelch_from_spectrum(pebble_channelizer_t *h, const double *d_spectrum, int bins, unsigned spectrum_rate,
                                             double squelch_db, void *stream);
int pebble_channelizer_set_input_format(pebble_channelizer_t *h, int fmt, int order, double user_gain, double norm_gain);
/* Front end of the last call: 1 = polyphase-DFT (all channel centres on the grid in_rate / 2^S: ONE mix, a 2^S-branch
 * polyphase FIR with the cascade's composite response and a 2^S-point DFT per output sample serve every channel; the
 * per-channel rounding of the oscillator angle, downconvert.cpp:108-110, is re-applied per channel at the output rate),
 * 0 = one CDownConvert per channel (any frequencies), -1 = no call yet. PEBBLE_CHZ_DIRECT=1 forces 0. */
int pebble_channelizer_front_path(const pebble_channelizer_t *h);
int pebble_channelizer_downconvert_device(pebble_channelizer_t *h, int n, const void *d_in, double *d_out,
                                          long long out_stride, void *stream);
int pebble_channelizer_process_device(pebble_channelizer_t *h, int n, const void *d_in, double *d_out, long long out_stride,
                                      void *stream);

#ifdef __cplusplus
}
#endif
#endif /* PEBBLE_CUDA_H */
