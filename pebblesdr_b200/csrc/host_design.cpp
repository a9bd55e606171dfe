// host_design.cpp — see host_design.h. Plain C++ (g++), no CUDA.
#include "host_design.h"
#include "hb_coefs.h"
#include <quadmath.h>
#include <cmath>
#include <cstring>

namespace pebble_host {

static const double kTwoPi = 6.28318530717958647692528676656;  // TWOPI, cpx.h:15
static const double kPi = 3.14159265358979323846264338328;     // PI, cpx.h:13

// downconvert.cpp:56-57
static const double kMinOutputRate = 7900.0 * 2.0;

double dc_plan(double in_rate, double max_bw, bool simple, std::vector<int> &lens) {
    using namespace pebble_hb;
    lens.clear();
    double f = in_rate;
    if (simple) {
        // downconvert.cpp:231-236: 51-tap half-bands until f <= 400 kHz
        while (f > 400000.0 && (int)lens.size() < PB_DC_MAX_STAGES) {
            lens.push_back(51);
            f /= 2.0;
        }
        return f;
    }
    const double last_max = kHalfBandMax[kNumHalfBands - 1];
    // downconvert.cpp:152-206: cheapest stage whose alias-free bandwidth still covers max_bw
    while ((f > (max_bw / last_max)) && (f > kMinOutputRate) && (int)lens.size() < PB_DC_MAX_STAGES) {
        if (f >= (max_bw / kCic3Max)) {
            lens.push_back(3);
        } else {
            for (int k = 0; k < kNumHalfBands; ++k) {
                if (f >= (max_bw / kHalfBandMax[k])) {
                    lens.push_back(kHalfBandLen[k]);
                    break;
                }
            }
            // (the loop condition guarantees the 51-tap test succeeds, so a stage is always added)
        }
        f /= 2.0;
    }
    return f;
}

void dc_fill_chain(DcChain &ch, const std::vector<int> &lens) {
    using namespace pebble_hb;
    ch.nstages = (int)lens.size();
    int off = 0;
    long long mem = 0;
    for (int s = 0; s < ch.nstages; ++s) {
        DcStage &st = ch.st[s];
        st.len = lens[s];
        st.coef_off = off;
        if (lens[s] == 3) {
            st.type = PB_DC_CIC3;
            st.hist = 2;
        } else {
            st.type = lens[s] == 11 ? PB_DC_HB11 : PB_DC_HBGEN;
            st.hist = lens[s] - 1;
            for (int k = 0; k < kNumHalfBands; ++k)
                if (kHalfBandLen[k] == lens[s])
                    memcpy(ch.coef + off, kHalfBandTaps + kHalfBandOff[k], sizeof(double) * lens[s]);
            off += 52;
        }
        mem += (long long)st.hist << s;
    }
    const long long g = 1LL << ch.nstages;
    ch.preroll = (int)(((mem + g - 1) / g) * g);
}

void dc_set_osc(DcChain &ch, double osc_cos, double osc_sin, int threads, int chunk) {
    // effective rotation per sample of the recursive oscillator = angle of the ROUNDED (cos, sin)
    __float128 th = atan2q((__float128)osc_sin, (__float128)osc_cos);
    const __float128 two_pi = 8 * atanq((__float128)1.0);
    __float128 turn = th / two_pi;
    turn -= floorq(turn);  // [0,1)
    __float128 t64 = ldexpq(turn, 64);
    __float128 hi = floorq(t64);
    __float128 lo = floorq(ldexpq(t64 - hi, 64));
    ch.turn_hi = (uint64_t)hi;
    ch.turn_lo = (uint64_t)lo;
    ch.wtab_stride = threads;
    for (int k = 0; k < PB_DC_MAX_WTAB; ++k) {
        __float128 a = th * (__float128)((long long)k * threads);
        ch.wtab[k].x = (double)cosq(a);
        ch.wtab[k].y = (double)sinq(a);
    }
    for (int k = 0; k < PB_DC_MAX_PAIRS; ++k) {
        __float128 a = th * (__float128)((long long)k * 2 * threads);
        ch.wpair[k].x = (double)cosq(a);
        ch.wpair[k].y = (double)sinq(a);
    }
    ch.w1.x = (double)cosq(th);
    ch.w1.y = (double)sinq(th);
    ch.chunk = chunk;
    ch.wchunk.x = (double)cosq(th * (__float128)chunk);
    ch.wchunk.y = (double)sinq(th * (__float128)chunk);
    // amplitude recurrence r_{i+1} = r_i * (1.95 - r_i^2), r_0 = 1 (downconvert.cpp:281-287)
    double r = 1.0;
    for (int i = 0; i < PB_DC_RTAB; ++i) {
        ch.rtab[i] = r;
        r = r * (1.95 - r * r);
    }
    ch.r_inf = r;
    ch.inv_r_inf = 1.0 / r;
}

// The cascade as ONE decimate-by-2^S FIR. Stage s computes  y_s[n] = sum_t g_s[t] * x_s[2n + a_s - t]  with
//   CIC3   (downconvert.cpp:519-535):  g = {.125, .375, .375, .125}, a = 1  (the newest sample used is x[2n+1])
//   HB11   (downconvert.cpp:421-496):  g = H0..H10,                 a = 0
//   HBgen  (downconvert.cpp:357-393):  g = coef[0..L-1] with the term of the OLDEST sample counted twice (finding F7:
//                                      the accumulator is seeded with buf[i]*coef[0] and the loop starts at j = 0), a = 0
// so  y[n] = sum_T h[T] * x[2^S n + c - T]  with  h = g_0 * (g_1 up 2) * (g_2 up 4) * ...  and  c = sum_s 2^s a_s.
// Products are accumulated in binary128 and rounded once. Returns the length of h.
int dc_composite(const DcChain &ch, std::vector<double> &h_out, int &c_out) {
    std::vector<__float128> h(1, (__float128)1.0);
    int c = 0;
    for (int s = 0; s < ch.nstages; ++s) {
        const DcStage &st = ch.st[s];
        std::vector<__float128> g;
        int a = 0;
        if (st.type == PB_DC_CIC3) {
            g = {(__float128)0.125, (__float128)0.375, (__float128)0.375, (__float128)0.125};
            a = 1;
        } else {
            g.resize(st.len);
            for (int t = 0; t < st.len; ++t) g[t] = (__float128)ch.coef[st.coef_off + (st.len - 1 - t)];
            if (st.type == PB_DC_HBGEN) g[st.len - 1] += (__float128)ch.coef[st.coef_off];   // F7
        }
        const long long up = 1LL << s;
        std::vector<__float128> nh(h.size() + (g.size() - 1) * (size_t)up, (__float128)0.0);
        for (size_t t = 0; t < g.size(); ++t)
            for (size_t T = 0; T < h.size(); ++T) nh[t * (size_t)up + T] += g[t] * h[T];
        h.swap(nh);
        c += (int)(up * a);
    }
    h_out.resize(h.size());
    for (size_t i = 0; i < h.size(); ++i) h_out[i] = (double)h[i];
    c_out = c;
    return (int)h.size();
}

void phase_advance(uint64_t &hi, uint64_t &lo, uint64_t thi, uint64_t tlo, uint64_t k) {
    unsigned __int128 p = ((unsigned __int128)hi << 64) | lo;
    unsigned __int128 t = ((unsigned __int128)thi << 64) | tlo;
    p += t * k;
    hi = (uint64_t)(p >> 64);
    lo = (uint64_t)p;
}

bool fastfir_design(int fft_size, int fir_size, double lo, double hi, double offset, double fs,
                    std::vector<double> &taps) {
    // fastfir.cpp:204-216
    double FLoCut = lo + offset, FHiCut = hi + offset;
    if ((FLoCut >= FHiCut) || (FLoCut >= fs / 2.0) || (FLoCut <= -fs / 2.0) || (FHiCut >= fs / 2.0) ||
        (FHiCut <= -fs / 2.0))
        return false;
    // fastfir.cpp:219-244
    double nFL = FLoCut / fs;
    double nFH = FHiCut / fs;
    double nFc = (nFH - nFL) / 2.0;
    double nFs = kTwoPi * (nFH + nFL) / 2.0;
    double fCenter = 0.5 * (double)(fir_size - 1);
    taps.assign((size_t)fft_size * 2, 0.0);
    for (int i = 0; i < fir_size; ++i) {
        double x = (double)i - fCenter;
        double z;
        if ((double)i == fCenter) {
            z = 2.0 * nFc;
        } else {
            // Blackman-Nuttall window, fastfir.cpp:103-111
            double w = (0.3635819 - 0.4891775 * cos((kTwoPi * i) / (fir_size - 1)) +
                        0.1365995 * cos((2.0 * kTwoPi * i) / (fir_size - 1)) -
                        0.0106411 * cos((3.0 * kTwoPi * i) / (fir_size - 1)));
            z = (double)sin(kTwoPi * x * nFc) / (kPi * x) * w;
        }
        taps[2 * i] = z * cos(nFs * x) / (double)fft_size;
        taps[2 * i + 1] = z * sin(nFs * x) / (double)fft_size;
    }
    return true;
}

void twiddle_table(int n, std::vector<double> &tw) {
    tw.resize((size_t)n * 2);
    for (int m = 0; m < n; ++m) {
        long double ang = -2.0L * 3.14159265358979323846264338327950288L * (long double)m / (long double)n;
        tw[2 * m] = (double)cosl(ang);
        tw[2 * m + 1] = (double)sinl(ang);
    }
}

}  // namespace pebble_host
