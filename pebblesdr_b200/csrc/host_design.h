// host_design.h — control-path (setup-time) computations done on the host in FP64, mirroring the
// reference's setters: chain planning, oscillator constants, filter design, window tables.
// These run once per parameter change (the reference runs them on the GUI thread under a mutex);
// all per-sample work is on the GPU.
#pragma once
#include "dc_types.h"
#include <vector>

namespace pebble_host {

// CDownConvert::SetDataRate / SetDataRateSimple (downconvert.cpp:139-244). Returns the output rate
// and fills `lens` with the FIR length of every stage (3 = CIC3).
double dc_plan(double in_rate, double max_bw, bool simple, std::vector<int> &lens);
// Fill the stage table, coefficients and pre-roll of a DcChain from stage lengths.
void dc_fill_chain(DcChain &ch, const std::vector<int> &lens);
// Oscillator constants for cos/sin of the NCO increment as the reference rounds them
// (downconvert.cpp:108-110). `threads` is the CTA size and `chunk` the chunk length the step tables are built for.
void dc_set_osc(DcChain &ch, double osc_cos, double osc_sin, int threads, int chunk);
// Composite impulse response h and alignment c of the whole cascade:  y[n] = sum_T h[T] x[2^S n + c - T]  (host_design.cpp).
int dc_composite(const DcChain &ch, std::vector<double> &h, int &c);
// phase (0.128 turns) += k * turn
void phase_advance(uint64_t &hi, uint64_t &lo, uint64_t thi, uint64_t tlo, uint64_t k);

// CFastFIR::SetupParameters (fastfir.cpp:191-264) up to, not including, the forward FFT:
// taps[0..fft_size) = zero-padded, frequency-shifted, 1/fft_size-scaled windowed-sinc.
// Returns false for the reference's "Filter Parameter error" (fastfir.cpp:208-216).
bool fastfir_design(int fft_size, int fir_size, double lo, double hi, double offset, double fs,
                    std::vector<double> &taps_interleaved);

// exp(-j 2 pi m / n), m < n, interleaved (computed in long double, rounded once).
void twiddle_table(int n, std::vector<double> &tw_interleaved);

// WindowFunction::regenerate (windowfunction.cpp:50-241). Returns coherentGain.
double window_table(int type, int n, std::vector<double> &w);

}  // namespace pebble_host

namespace pebble_host {
// CFir::InitLPFilter / InitHPFilter (fir.cpp:246-337, 352-440): Kaiser-windowed sinc, taps clamped to
// [3,75] (LP) / [3,74]|1 (HP). Returns the tap count and fills coef[0..ntaps).
int cfir_design_lp(int num_taps, double scale, double astop, double fpass, double fstop, double fs, double *coef);
int cfir_design_hp(int num_taps, double scale, double astop, double fpass, double fstop, double fs, double *coef);

// AGC::setParameters (application/agc.cpp:237-309) for the non-OFF modes.
struct AgcParams {
    int use_hang;
    double knee, gain_slope, fixed_gain;
    double attack_rise, attack_fall, decay_rise, decay_fall;
    int hang_time, delay_samples, window_samples;
};
void agc_params(double sample_rate, bool use_hang, int threshold, int slope_factor, int decay_ms, AgcParams &p);
}  // namespace pebble_host
