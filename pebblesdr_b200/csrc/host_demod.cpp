// host_demod.cpp — setup-time designs for the demodulator tail: CFir Kaiser filters and AGC constants.
#include "host_design.h"
#include <cmath>

namespace pebble_host {
static const double kTwoPi = 6.28318530717958647692528676656;
static const double kPi = 3.14159265358979323846264338328;
static const int kMaxNumCoef = 75;  // MAX_NUMCOEF, fir.h:25

// CFir::Izero, fir.cpp:494-512: I0(x) series until the term drops below 1e-9 of the sum
static double izero(double x) {
    double x2 = x / 2.0, sum = 1.0, ds = 1.0, di = 1.0, tmp;
    do {
        tmp = x2 / di;
        tmp *= tmp;
        ds *= tmp;
        sum += ds;
        di += 1.0;
    } while (ds >= 1e-9 * sum);
    return sum;
}
static double kaiser_beta(double astop) {  // fir.cpp:259-264
    if (astop < 20.96) return 0;
    if (astop >= 50.0) return .1102 * (astop - 8.71);
    return .5842 * pow((astop - 20.96), 0.4) + .07886 * (astop - 20.96);
}

int cfir_design_lp(int NumTaps, double Scale, double Astop, double Fpass, double Fstop, double Fs, double *coef) {
    double normFpass = Fpass / Fs, normFstop = Fstop / Fs;
    double normFcut = (normFstop + normFpass) / 2.0;
    double Beta = kaiser_beta(Astop);
    int n_taps = (Astop - 8.0) / (2.285 * kTwoPi * (normFstop - normFpass)) + 1;  // fir.cpp:267 (double -> int)
    if (n_taps > kMaxNumCoef) n_taps = kMaxNumCoef;
    if (n_taps < 3) n_taps = 3;
    if (NumTaps) n_taps = NumTaps;
    double fCenter = .5 * (double)(n_taps - 1);
    double izb = izero(Beta);
    for (int n = 0; n < n_taps; n++) {
        double x = (double)n - fCenter;
        double c;
        if ((double)n == fCenter)
            c = 2.0 * normFcut;
        else
            c = (double)sin(kTwoPi * x * normFcut) / (kPi * x);
        x = ((double)n - ((double)n_taps - 1.0) / 2.0) / (((double)n_taps - 1.0) / 2.0);
        coef[n] = Scale * c * izero(Beta * sqrt(1 - (x * x))) / izb;
    }
    return n_taps;
}

int cfir_design_hp(int NumTaps, double Scale, double Astop, double Fpass, double Fstop, double Fs, double *coef) {
    double normFpass = Fpass / Fs, normFstop = Fstop / Fs;
    double normFcut = (normFstop + normFpass) / 2.0;
    double Beta = kaiser_beta(Astop);
    int n_taps = (Astop - 8.0) / (2.285 * kTwoPi * (normFpass - normFstop)) + 1;  // fir.cpp:373
    if (n_taps > (kMaxNumCoef - 1)) n_taps = kMaxNumCoef - 1;
    if (n_taps < 3) n_taps = 3;
    n_taps |= 1;
    if (NumTaps) n_taps = NumTaps;
    double izb = izero(Beta);
    double fCenter = .5 * (double)(n_taps - 1);
    for (int n = 0; n < n_taps; n++) {
        double x = (double)n - (double)(n_taps - 1) / 2.0;
        double c;
        if ((double)n == fCenter)
            c = 1.0 - 2.0 * normFcut;
        else
            c = (double)(sin(kPi * x) / (kPi * x) - sin(kTwoPi * x * normFcut) / (kPi * x));
        x = ((double)n - ((double)n_taps - 1.0) / 2.0) / (((double)n_taps - 1.0) / 2.0);
        coef[n] = Scale * c * izero(Beta * sqrt(1 - (x * x))) / izb;
    }
    return n_taps;
}

void agc_params(double fs, bool use_hang, int threshold, int slope_factor, int decay, AgcParams &p) {
    // agc.h:31-58 constants are `static constexpr float`: their float roundings enter the formulas
    const float DELAY_TIMECONST = .015f, WINDOW_TIMECONST = .018f, ATTACK_RISE_TIMECONST = .002f,
                ATTACK_FALL_TIMECONST = .005f, DECAY_RISEFALL_RATIO = .3f, RELEASE_TIMECONST = .05f,
                AGC_OUTSCALE = 0.7f;
    const int MAX_DELAY_BUF = 2048;
    p.use_hang = use_hang ? 1 : 0;
    p.knee = (double)threshold / 20.0;
    p.gain_slope = slope_factor / (100.0);
    p.fixed_gain = AGC_OUTSCALE * pow(10.0, p.knee * (p.gain_slope - 1.0));
    p.attack_rise = (1.0 - exp(-1.0 / (fs * ATTACK_RISE_TIMECONST)));
    p.attack_fall = (1.0 - exp(-1.0 / (fs * ATTACK_FALL_TIMECONST)));
    p.decay_rise = (1.0 - exp(-1.0 / (fs * (double)decay * .001 * DECAY_RISEFALL_RATIO)));
    p.hang_time = (int)(fs * (double)decay * .001);
    if (use_hang)
        p.decay_fall = (1.0 - exp(-1.0 / (fs * RELEASE_TIMECONST)));
    else
        p.decay_fall = (1.0 - exp(-1.0 / (fs * (double)decay * .001)));
    p.delay_samples = (int)(fs * DELAY_TIMECONST);
    p.window_samples = (int)(fs * WINDOW_TIMECONST);
    if (p.delay_samples >= MAX_DELAY_BUF - 1) p.delay_samples = MAX_DELAY_BUF - 1;
}
}  // namespace pebble_host
