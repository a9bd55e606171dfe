// host_window.cpp — window tables for the FFT operator, built on the host once per fftParams call.
// Restates WindowFunction::regenerate (pebblelib/windowfunction.cpp:50-241) INCLUDING its mixed
// float/double arithmetic (finding F13): the loop variables freq/rate/angle/expn/expsum/cx/two_pi are
// `float` there, and because <cmath> is in scope cos()/sin() of a float resolve to the float overloads.
// The same declarations are used here so the tables are bit-identical to the reference's.
#include "host_design.h"
#include <math.h>
#include <cmath>
#include <complex>

namespace pebble_host {

double window_table(int type, int numSamples, std::vector<double> &window) {
    window.assign((size_t)(numSamples > 0 ? numSamples : 0), 0.0);
    if (numSamples <= 0) return 0.0;
    int i, j, midn, midp1, midm1;
    float freq, rate, sr1, angle, expn, expsum, cx, two_pi;
    midn = numSamples / 2;
    midp1 = (numSamples + 1) / 2;
    midm1 = (numSamples - 1) / 2;
    two_pi = 6.28318530717958647692528676656;
    freq = two_pi / numSamples;
    rate = 1.0 / midn;
    angle = 0.0;
    expn = log(2.0) / midn + 1.0;
    expsum = 1.0;
    double sumCoeff = 0;
    double coherentGain = 0.0;
    double *w = window.data();
    switch (type) {
    case 1:  // RECTANGULAR
        for (i = 0; i < numSamples; i++) { w[i] = 1.0; sumCoeff += w[i]; }
        break;
    case 2:  // HANNING
        for (i = 0, j = numSamples - 1, angle = 0.0; i <= midn; i++, j--, angle += freq) {
            w[j] = (w[i] = 0.5 - 0.5 * cos(angle));
            sumCoeff += w[i];
        }
        break;
    case 3:  // WELCH (integer division inside the sqrt, as in the reference)
        for (i = 0, j = numSamples - 1; i <= midn; i++, j--) {
            w[j] = (w[i] = 1.0 - (float)sqrt((float)((i - midm1) / midp1)));
            sumCoeff += w[i];
        }
        break;
    case 4:  // PARZEN
        for (i = 0, j = numSamples - 1; i <= midn; i++, j--) {
            w[j] = (w[i] = 1.0 - ((float)fabs((float)(i - midm1) / midp1)));
            sumCoeff += w[i];
        }
        break;
    case 5:  // BARTLETT
        for (i = 0, j = numSamples - 1, angle = 0.0; i <= midn; i++, j--, angle += rate) {
            w[j] = (w[i] = angle);
            sumCoeff += w[i];
        }
        break;
    case 6:  // HAMMING
        for (i = 0, j = numSamples - 1, angle = 0.0; i <= midn; i++, j--, angle += freq) {
            w[j] = (w[i] = 0.54 - 0.46 * cos(angle));
            sumCoeff += w[i];
        }
        break;
    case 7:  // BLACKMAN2
        for (i = 0, j = numSamples - 1, angle = 0.0; i <= midn; i++, j--, angle += freq) {
            cx = cos(angle);
            w[j] = (w[i] = (.34401 + (cx * (-.49755 + (cx * .15844)))));
            sumCoeff += w[i];
        }
        break;
    case 8:  // BLACKMAN3
        for (i = 0, j = numSamples - 1, angle = 0.0; i <= midn; i++, j--, angle += freq) {
            cx = cos(angle);
            w[j] = (w[i] = (.21747 + (cx * (-.45325 + (cx * (.28256 - (cx * .04672)))))));
            sumCoeff += w[i];
        }
        break;
    case 9:  // BLACKMAN4
        for (i = 0, j = numSamples - 1, angle = 0.0; i <= midn; i++, j--, angle += freq) {
            cx = cos(angle);
            w[j] = (w[i] = (.084037 + (cx * (-.29145 + (cx * (.375696 + (cx * (-.20762 + (cx * .041194)))))))));
            sumCoeff += w[i];
        }
        break;
    case 10:  // EXPONENTIAL
        for (i = 0, j = numSamples - 1; i <= midn; i++, j--) {
            w[j] = (w[i] = expsum - 1.0);
            sumCoeff += w[i];
            expsum *= expn;
        }
        break;
    case 11:  // RIEMANN
        sr1 = two_pi / numSamples;
        for (i = 0, j = numSamples - 1; i <= midn; i++, j--) {
            if (i == midn) {
                w[j] = (w[i] = 1.0);
            } else {
                cx = sr1 * (midn - i);
                w[i] = sin(cx) / cx;
                w[j] = w[i];
            }
            sumCoeff += w[i];
        }
        break;
    case 12: {  // BLACKMANHARRIS (the only window SignalSpectrum uses, signalspectrum.cpp:58-59)
        float a0 = 0.35875F, a1 = 0.48829F, a2 = 0.14128F, a3 = 0.01168F;
        for (i = 0; i < numSamples; i++) {
            w[i] = a0 - a1 * cos(two_pi * (i + 0.5) / numSamples) + a2 * cos(2.0 * two_pi * (i + 0.5) / numSamples) -
                   a3 * cos(3.0 * two_pi * (i + 0.5) / numSamples);
            sumCoeff += w[i];
        }
        break;
    }
    default:  // NONE, BLACKMAN (reads an uninitialised coherentGain in the reference: refused), unknown
        return 0.0;
    }
    coherentGain = sumCoeff / numSamples;
    return coherentGain;
}

}  // namespace pebble_host
