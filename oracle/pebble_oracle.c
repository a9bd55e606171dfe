/*
 * oracle/pebble_oracle.c — TEST INFRASTRUCTURE ONLY. Not linked into, imported by or shipped with the product.
 *
 * Plain-C, single-threaded CPU restatement of the reference algorithms on the hot path, each function citing
 * the reference file:line it follows. Used by tests/ as the checker, by __graft_entry__.smoke() and by bench.py's
 * cpu_baseline leg (kind "port") when the real reference build (oracle/_ref, kind "reference") is not present.
 *
 * Parity pinning: the reference ships no golden vectors. This restatement is pinned against the REAL
 * reference compiled from /root/reference (oracle/_ref/libpebbleref.so: unmodified pebblelib + FFTW 3.3.4) in
 * tests/test_oracle.py, and against tests/golden/ fixtures generated from that library (tests/golden/make_golden.py).
 * The application-side steps (AGC, Demod_AM, Demod_NFM, receiver ordering) inherit QObject/global and cannot be
 * compiled without Qt; they are restated here line by line and are "pinned by inspection" only (their CFir/FFT
 * building blocks ARE pinned).
 *
 * The FFT is a textbook iterative radix-2 (the reference delegates to FFTW 3.3.4, pebblelib/fftw.cpp:24-25,41,62:
 * forward = exp(-j..), backward = exp(+j..), both unnormalised).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "hb_coefs_c.h"

#define ORC_PI 3.14159265358979323846264338328     /* PI, cpx.h:13 */
#define ORC_TWOPI 6.28318530717958647692528676656  /* TWOPI, cpx.h:15 */

typedef struct { double re, im; } ocpx;

/* ------------------------------------------------------------------------------------------------------------
 * FFT (power of two), in place. sign = -1 forward, +1 backward; unnormalised (FFTW convention).
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct { int n; ocpx *tw; int *rev; } orc_fft;

static orc_fft *orc_fft_new(int n) {
    orc_fft *f = (orc_fft *)malloc(sizeof(orc_fft));
    f->n = n;
    f->tw = (ocpx *)malloc(sizeof(ocpx) * n);
    f->rev = (int *)malloc(sizeof(int) * n);
    int bits = 0;
    while ((1 << bits) < n) bits++;
    for (int i = 0; i < n; i++) {
        long double a = -2.0L * 3.14159265358979323846264338327950288L * i / n;
        f->tw[i].re = (double)cosl(a);
        f->tw[i].im = (double)sinl(a);
        int r = 0;
        for (int b = 0; b < bits; b++) if (i & (1 << b)) r |= 1 << (bits - 1 - b);
        f->rev[i] = r;
    }
    return f;
}
static void orc_fft_free(orc_fft *f) { if (f) { free(f->tw); free(f->rev); free(f); } }
static void orc_fft_run(const orc_fft *f, ocpx *a, int sign) {
    const int n = f->n;
    for (int i = 0; i < n; i++) {
        int r = f->rev[i];
        if (r > i) { ocpx t = a[i]; a[i] = a[r]; a[r] = t; }
    }
    for (int len = 2; len <= n; len <<= 1) {
        const int half = len >> 1, step = n / len;
        for (int i = 0; i < n; i += len) {
            for (int k = 0; k < half; k++) {
                const double wr = f->tw[k * step].re, wi = sign < 0 ? f->tw[k * step].im : -f->tw[k * step].im;
                ocpx *u = &a[i + k], *v = &a[i + k + half];
                const double tr = v->re * wr - v->im * wi, ti = v->re * wi + v->im * wr;
                v->re = u->re - tr; v->im = u->im - ti;
                u->re += tr; u->im += ti;
            }
        }
    }
}

/* ------------------------------------------------------------------------------------------------------------
 * CFastFIR (pebblelib/fastfir.cpp:77-334), sizes as parameters (fastfir.cpp:65-66 fixes them at 2048/1025).
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct {
    int N, FIR;
    double *window;            /* m_pWindowTbl */
    ocpx *coef, *buf, *ovl;    /* m_pFilterCoef, m_pFFTBuf, m_pFFTOverlapBuf */
    int inpos;                 /* m_InBufInPos */
    double lo, hi, off, fs;
    orc_fft *fft;
} orc_fastfir;

void *orc_fastfir_create(int fft_size, int fir_size) {
    orc_fastfir *h = (orc_fastfir *)calloc(1, sizeof(orc_fastfir));
    h->N = fft_size; h->FIR = fir_size;
    h->window = (double *)malloc(sizeof(double) * fir_size);
    h->coef = (ocpx *)calloc(fft_size, sizeof(ocpx));
    h->buf = (ocpx *)calloc(fft_size, sizeof(ocpx));       /* fastfir.cpp:96-100 zeroes m_pFFTBuf */
    h->ovl = (ocpx *)calloc(fir_size, sizeof(ocpx));
    h->inpos = fir_size - 1;                               /* fastfir.cpp:95 */
    for (int i = 0; i < fir_size; i++)                     /* Blackman-Nuttall, fastfir.cpp:103-111 */
        h->window[i] = (0.3635819 - 0.4891775 * cos((ORC_TWOPI * i) / (fir_size - 1)) +
                        0.1365995 * cos((2.0 * ORC_TWOPI * i) / (fir_size - 1)) -
                        0.0106411 * cos((3.0 * ORC_TWOPI * i) / (fir_size - 1)));
    h->lo = -1.0; h->hi = 1.0; h->off = 1.0; h->fs = 1.0;  /* fastfir.cpp:140-143 */
    h->fft = orc_fft_new(fft_size);
    return h;
}
void orc_fastfir_destroy(void *p) {
    orc_fastfir *h = (orc_fastfir *)p;
    if (!h) return;
    free(h->window); free(h->coef); free(h->buf); free(h->ovl); orc_fft_free(h->fft); free(h);
}
/* fastfir.cpp:191-272. Returns 0 installed, 1 unchanged, 2 parameter error (old filter kept). */
int orc_fastfir_setup(void *p, double FLoCut, double FHiCut, double Offset, double SampleRate) {
    orc_fastfir *h = (orc_fastfir *)p;
    if (FLoCut == h->lo && FHiCut == h->hi && Offset == h->off && SampleRate == h->fs) return 1;
    h->lo = FLoCut; h->hi = FHiCut; h->off = Offset; h->fs = SampleRate;
    FLoCut += Offset; FHiCut += Offset;
    if ((FLoCut >= FHiCut) || (FLoCut >= SampleRate / 2.0) || (FLoCut <= -SampleRate / 2.0) ||
        (FHiCut >= SampleRate / 2.0) || (FHiCut <= -SampleRate / 2.0))
        return 2;
    double nFL = FLoCut / SampleRate, nFH = FHiCut / SampleRate;
    double nFc = (nFH - nFL) / 2.0;
    double nFs = ORC_TWOPI * (nFH + nFL) / 2.0;
    double fCenter = 0.5 * (double)(h->FIR - 1);
    for (int i = 0; i < h->N; i++) h->coef[i].re = h->coef[i].im = 0.0;
    for (int i = 0; i < h->FIR; i++) {
        double x = (double)i - fCenter, z;
        if ((double)i == fCenter) z = 2.0 * nFc;
        else z = (double)sin(ORC_TWOPI * x * nFc) / (ORC_PI * x) * h->window[i];
        h->coef[i].re = z * cos(nFs * x) / (double)h->N;
        h->coef[i].im = z * sin(nFs * x) / (double)h->N;
    }
    orc_fft_run(h->fft, h->coef, -1);  /* fastfir.cpp:267 */
    return 0;
}
/* fastfir.cpp:281-319 */
int orc_fastfir_process(void *p, int InLength, const double *in_, double *out_) {
    orc_fastfir *h = (orc_fastfir *)p;
    const ocpx *InBuf = (const ocpx *)in_;
    ocpx *OutBuf = (ocpx *)out_;
    int i = 0, j, len = InLength, outpos = 0;
    if (!InLength) return 0;
    while (len--) {
        j = h->inpos - (h->N - h->FIR + 1);
        if (j >= 0) h->ovl[j] = InBuf[i];
        h->buf[h->inpos++] = InBuf[i++];
        if (h->inpos >= h->N) {
            orc_fft_run(h->fft, h->buf, -1);
            for (int k = 0; k < h->N; k++) {  /* CpxMpy, fastfir.cpp:324-334 */
                double sr = h->buf[k].re, si = h->buf[k].im;
                h->buf[k].re = h->coef[k].re * sr - h->coef[k].im * si;
                h->buf[k].im = h->coef[k].re * si + h->coef[k].im * sr;
            }
            orc_fft_run(h->fft, h->buf, +1);
            for (j = (h->FIR - 1); j < h->N; j++) OutBuf[outpos++] = h->buf[j];
            for (j = 0; j < (h->FIR - 1); j++) h->buf[j] = h->ovl[j];
            h->inpos = h->FIR - 1;
        }
    }
    return outpos;
}

/* ------------------------------------------------------------------------------------------------------------
 * CDownConvert (pebblelib/downconvert.cpp:64-535): recursive quadrature oscillator + in-place decimate-by-2 chain.
 * The in-place semantics (stages read and write the same buffer) are kept, because they are visible: a generic
 * half-band stage copies its delay line from the buffer AFTER the outputs were written over its head
 * (downconvert.cpp:388-390), and returns InLength/2 unfiltered when InLength < taps (downconvert.cpp:362-363).
 * ---------------------------------------------------------------------------------------------------------- */
#define ORC_MAX_STAGES 16          /* reference: MAX_DECSTAGES 10 incl. sentinel (downconvert.h:23), finding F8 */
#define ORC_HB_BUFSIZE 32768       /* MAX_HALF_BAND_BUFSIZE, downconvert.cpp:59 */
#define ORC_MIN_OUTPUT_RATE (7900.0 * 2.0)

typedef struct {
    int len;                 /* 3 = CIC3, 11 = unrolled 11-tap, else generic half-band */
    const double *coef;
    ocpx *buf;               /* generic: m_pHBFirBuf */
    ocpx d[10];              /* 11-tap: d0..d9 */
    ocpx xodd, xeven;        /* CIC3 */
} orc_stage;

typedef struct {
    double nco_inc, nco_freq, cw_offset, in_rate, max_bw, out_rate;
    ocpx osc1;
    double osc_cos, osc_sin;
    int nstages;
    orc_stage st[ORC_MAX_STAGES];
} orc_downconvert;

static void orc_dc_delete_filters(orc_downconvert *h) {
    for (int i = 0; i < h->nstages; i++) free(h->st[i].buf);
    h->nstages = 0;
}
static void orc_dc_add_stage(orc_downconvert *h, int len) {
    orc_stage *s = &h->st[h->nstages++];
    memset(s, 0, sizeof(*s));
    s->len = len;
    if (len > 3) {
        for (int k = 0; k < ORC_NUM_HB; k++) if (orc_hb_len[k] == len) s->coef = orc_hb_taps + orc_hb_off[k];
        if (len != 11) s->buf = (ocpx *)calloc(ORC_HB_BUFSIZE, sizeof(ocpx));
    }
}
void *orc_downconvert_create(void) {
    orc_downconvert *h = (orc_downconvert *)calloc(1, sizeof(orc_downconvert));
    h->in_rate = 100000.0; h->max_bw = 10000.0;   /* downconvert.cpp:71-72 */
    h->osc1.re = 1.0; h->osc1.im = 0.0;           /* downconvert.cpp:75-76 */
    return h;
}
void orc_downconvert_destroy(void *p) { if (p) { orc_dc_delete_filters((orc_downconvert *)p); free(p); } }
/* downconvert.cpp:102-113 */
void orc_downconvert_set_frequency(void *p, double NcoFreq) {
    orc_downconvert *h = (orc_downconvert *)p;
    NcoFreq = -NcoFreq;
    h->nco_freq = NcoFreq + h->cw_offset;
    h->nco_inc = ORC_TWOPI * h->nco_freq / h->in_rate;
    h->osc_cos = cos(h->nco_inc);
    h->osc_sin = sin(h->nco_inc);
}
void orc_downconvert_set_cw_offset(void *p, double off) { ((orc_downconvert *)p)->cw_offset = off; }
/* downconvert.cpp:139-213 (simple=0) and :220-244 (simple=1) */
double orc_downconvert_set_data_rate(void *p, double InRate, double MaxBW, int simple) {
    orc_downconvert *h = (orc_downconvert *)p;
    double f = InRate;
    if (h->in_rate != InRate || h->max_bw != MaxBW) {
        h->in_rate = InRate; h->max_bw = MaxBW;
        orc_dc_delete_filters(h);
        if (simple) {
            while (f > 400000.0 && h->nstages < ORC_MAX_STAGES) { orc_dc_add_stage(h, 51); f /= 2.0; }
        } else {
            while ((f > (MaxBW / orc_hb_max[ORC_NUM_HB - 1])) && (f > ORC_MIN_OUTPUT_RATE) && h->nstages < ORC_MAX_STAGES) {
                if (f >= (MaxBW / orc_cic3_max)) orc_dc_add_stage(h, 3);
                else for (int k = 0; k < ORC_NUM_HB; k++) if (f >= (MaxBW / orc_hb_max[k])) { orc_dc_add_stage(h, orc_hb_len[k]); break; }
                f /= 2.0;
            }
        }
        h->out_rate = f;
        orc_downconvert_set_frequency(h, h->nco_freq);  /* downconvert.cpp:210: re-negates the stored value */
    }
    return h->out_rate;
}
int orc_downconvert_num_stages(void *p) { return ((orc_downconvert *)p)->nstages; }
int orc_downconvert_stage_len(void *p, int i) { return ((orc_downconvert *)p)->st[i].len; }

/* CCicN3DecimateBy2::DecBy2, downconvert.cpp:519-535 */
static int orc_cic3(orc_stage *s, int n, ocpx *in, ocpx *out) {
    int i, j;
    for (i = 0, j = 0; i < n; i += 2, j++) {
        ocpx even = in[i], odd = in[i + 1];
        out[j].re = .125 * (odd.re + s->xeven.re + 3.0 * (s->xodd.re + even.re));
        out[j].im = .125 * (odd.im + s->xeven.im + 3.0 * (s->xodd.im + even.im));
        s->xodd = odd; s->xeven = even;
    }
    return j;
}
/* CHalfBandDecimateBy2::DecBy2, downconvert.cpp:357-393 (tap 0 counted twice, finding F7) */
static int orc_hbgen(orc_stage *s, int n, ocpx *in, ocpx *out) {
    const int L = s->len;
    const double *c = s->coef;
    int i, j, nout = 0;
    if (n < L) return n / 2;
    for (i = 0, j = L - 1; i < n; i++) s->buf[j++] = in[i];
    for (i = 0; i < n; i += 2) {
        ocpx acc;
        acc.re = s->buf[i].re * c[0];
        acc.im = s->buf[i].im * c[0];
        for (j = 0; j < L; j += 2) {
            acc.re += s->buf[i + j].re * c[j];
            acc.im += s->buf[i + j].im * c[j];
        }
        acc.re += s->buf[i + (L - 1) / 2].re * c[(L - 1) / 2];
        acc.im += s->buf[i + (L - 1) / 2].im * c[(L - 1) / 2];
        out[nout++] = acc;
    }
    for (i = 0, j = (n - L) + 1; i < L - 1; i++) s->buf[i] = in[j++];  /* reads `in` after `out` (== in) was written */
    return nout;
}
/* CHalfBand11TapDecimateBy2::DecBy2, downconvert.cpp:421-496 */
static int orc_hb11(orc_stage *s, int n, ocpx *in, ocpx *out) {
    const double *H = s->coef;
    ocpx tmp[9];
    /* w(k): sample k of the virtual sequence d0..d9 followed by the input */
#define W(k) ((k) < 10 ? s->d[(k)] : in[(k) - 10])
    for (int o = 0; o < 9; o++) {
        const int b = 2 * o;
        tmp[o].re = H[0] * W(b).re + H[2] * W(b + 2).re + H[4] * W(b + 4).re + H[5] * W(b + 5).re + H[6] * W(b + 6).re +
                    H[8] * W(b + 8).re + H[10] * W(b + 10).re;
        tmp[o].im = H[0] * W(b).im + H[2] * W(b + 2).im + H[4] * W(b + 4).im + H[5] * W(b + 5).im + H[6] * W(b + 6).im +
                    H[8] * W(b + 8).im + H[10] * W(b + 10).im;
    }
#undef W
    ocpx *pIn = &in[8], *pOut = &out[9];
    for (int i = 0; i < (n - 11 - 6) / 2; i++) {
        ocpx y;
        y.re = H[0] * pIn[0].re + H[2] * pIn[2].re + H[4] * pIn[4].re + H[5] * pIn[5].re + H[6] * pIn[6].re +
               H[8] * pIn[8].re + H[10] * pIn[10].re;
        y.im = H[0] * pIn[0].im + H[2] * pIn[2].im + H[4] * pIn[4].im + H[5] * pIn[5].im + H[6] * pIn[6].im +
               H[8] * pIn[8].im + H[10] * pIn[10].im;
        *pOut++ = y;
        pIn += 2;
    }
    for (int o = 0; o < 9; o++) out[o] = tmp[o];
    for (int k = 0; k < 10; k++) s->d[k] = in[n - 10 + k];  /* downconvert.cpp:489-493 */
    return n / 2;
}
/* downconvert.cpp:257-334. `in` is clobbered exactly as in the reference. */
int orc_downconvert_process(void *p, int InLength, double *in_, double *out_) {
    orc_downconvert *h = (orc_downconvert *)p;
    ocpx *in = (ocpx *)in_, *out = (ocpx *)out_;
    for (int i = 0; i < InLength; i++) {
        ocpx dtmp = in[i], Osc;
        Osc.re = h->osc1.re * h->osc_cos - h->osc1.im * h->osc_sin;
        Osc.im = h->osc1.im * h->osc_cos + h->osc1.re * h->osc_sin;
        double OscGn = 1.95 - (h->osc1.re * h->osc1.re + h->osc1.im * h->osc1.im);
        h->osc1.re = OscGn * Osc.re;
        h->osc1.im = OscGn * Osc.im;
        in[i].re = ((dtmp.re * Osc.re) - (dtmp.im * Osc.im));
        in[i].im = ((dtmp.re * Osc.im) + (dtmp.im * Osc.re));
    }
    int n = InLength;
    for (int j = 0; j < h->nstages; j++) {
        orc_stage *s = &h->st[j];
        if (s->len == 3) n = orc_cic3(s, n, in, in);
        else if (s->len == 11) n = orc_hb11(s, n, in, in);
        else n = orc_hbgen(s, n, in, in);
    }
    for (int i = 0; i < n; i++) out[i] = in[i];
    return n;
}

/* ------------------------------------------------------------------------------------------------------------
 * WindowFunction::regenerate (pebblelib/windowfunction.cpp:50-241). The reference's loop variables are `float` and,
 * with <cmath> in scope, cos()/sin() of a float call the float overloads: spelled cosf/sinf here (finding F13).
 * Returns coherentGain; type numbering = WindowFunction::WINDOWTYPE (RECTANGULAR=1 .. NONE=14).
 * ---------------------------------------------------------------------------------------------------------- */
double orc_window(int type, int numSamples, double *w) {
    int i, j, midn = numSamples / 2, midp1 = (numSamples + 1) / 2, midm1 = (numSamples - 1) / 2;
    float freq, rate, sr1, angle, expn, expsum, cx, two_pi;
    two_pi = ORC_TWOPI;
    freq = two_pi / numSamples;
    rate = 1.0 / midn;
    angle = 0.0;
    expn = log(2.0) / midn + 1.0;
    expsum = 1.0;
    double sum = 0;
    switch (type) {
    case 1: for (i = 0; i < numSamples; i++) { w[i] = 1.0; sum += w[i]; } break;
    case 2: for (i = 0, j = numSamples - 1, angle = 0.0; i <= midn; i++, j--, angle += freq) { w[j] = (w[i] = 0.5 - 0.5 * cosf(angle)); sum += w[i]; } break;
    case 3: for (i = 0, j = numSamples - 1; i <= midn; i++, j--) { w[j] = (w[i] = 1.0 - (float)sqrtf((float)((i - midm1) / midp1))); sum += w[i]; } break;
    case 4: for (i = 0, j = numSamples - 1; i <= midn; i++, j--) { w[j] = (w[i] = 1.0 - ((float)fabsf((float)(i - midm1) / midp1))); sum += w[i]; } break;
    case 5: for (i = 0, j = numSamples - 1, angle = 0.0; i <= midn; i++, j--, angle += rate) { w[j] = (w[i] = angle); sum += w[i]; } break;
    case 6: for (i = 0, j = numSamples - 1, angle = 0.0; i <= midn; i++, j--, angle += freq) { w[j] = (w[i] = 0.54 - 0.46 * cosf(angle)); sum += w[i]; } break;
    case 7: for (i = 0, j = numSamples - 1, angle = 0.0; i <= midn; i++, j--, angle += freq) { cx = cosf(angle); w[j] = (w[i] = (.34401 + (cx * (-.49755 + (cx * .15844))))); sum += w[i]; } break;
    case 8: for (i = 0, j = numSamples - 1, angle = 0.0; i <= midn; i++, j--, angle += freq) { cx = cosf(angle); w[j] = (w[i] = (.21747 + (cx * (-.45325 + (cx * (.28256 - (cx * .04672))))))); sum += w[i]; } break;
    case 9: for (i = 0, j = numSamples - 1, angle = 0.0; i <= midn; i++, j--, angle += freq) { cx = cosf(angle); w[j] = (w[i] = (.084037 + (cx * (-.29145 + (cx * (.375696 + (cx * (-.20762 + (cx * .041194))))))))); sum += w[i]; } break;
    case 10: for (i = 0, j = numSamples - 1; i <= midn; i++, j--) { w[j] = (w[i] = expsum - 1.0); sum += w[i]; expsum *= expn; } break;
    case 11:
        sr1 = two_pi / numSamples;
        for (i = 0, j = numSamples - 1; i <= midn; i++, j--) {
            if (i == midn) w[j] = (w[i] = 1.0);
            else { cx = sr1 * (midn - i); w[i] = sinf(cx) / cx; w[j] = w[i]; }
            sum += w[i];
        }
        break;
    case 12: {
        float a0 = 0.35875F, a1 = 0.48829F, a2 = 0.14128F, a3 = 0.01168F;
        for (i = 0; i < numSamples; i++) {
            w[i] = a0 - a1 * cos(two_pi * (i + 0.5) / numSamples) + a2 * cos(2.0 * two_pi * (i + 0.5) / numSamples) -
                   a3 * cos(3.0 * two_pi * (i + 0.5) / numSamples);
            sum += w[i];
        }
        break;
    }
    default: return 0.0;
    }
    return sum / numSamples;
}

/* ------------------------------------------------------------------------------------------------------------
 * FFT operator: fftParams / fftForward / fftInverse / fftSpectrum (pebblelib/fft.cpp:67-157,183-225,324-399,
 * fftw.cpp:17-81). Power-of-two sizes only (FFTW accepts any size; the reference clamps to [2048,65535]).
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct {
    int N, spb, wtype, overload;
    double cg, maxbin;
    double *win, *power, *amp;
    ocpx *time, *freq;
    orc_fft *fft;
} orc_fftop;

void *orc_fftop_create(int fft_size, int samples_per_buffer, int window_type) {
    orc_fftop *h = (orc_fftop *)calloc(1, sizeof(orc_fftop));
    h->N = fft_size < 2048 ? 2048 : fft_size;           /* fft.cpp:74-79 (upper clamp not restated) */
    h->spb = samples_per_buffer; h->wtype = window_type;
    h->maxbin = 1.0 * samples_per_buffer;               /* fft.cpp:84 */
    h->win = (double *)calloc(samples_per_buffer, sizeof(double));
    h->cg = window_type == 14 ? 1.0 : orc_window(window_type, samples_per_buffer, h->win);
    h->power = (double *)calloc(h->N, sizeof(double));  /* reference: new double[] (uninitialised) */
    h->amp = (double *)calloc(h->N, sizeof(double));
    h->time = (ocpx *)calloc(h->N, sizeof(ocpx));
    h->freq = (ocpx *)calloc(h->N, sizeof(ocpx));
    h->fft = orc_fft_new(h->N);
    return h;
}
void orc_fftop_destroy(void *p) {
    orc_fftop *h = (orc_fftop *)p;
    if (!h) return;
    free(h->win); free(h->power); free(h->amp); free(h->time); free(h->freq); orc_fft_free(h->fft); free(h);
}
/* m_applyWindow, fft.cpp:129-157 */
static void orc_apply_window(orc_fftop *h, const ocpx *in, int n) {
    if (h->wtype != 14 && n == h->spb) {
        h->overload = 0;
        for (int i = 0; i < h->spb; i++) {
            if (fabs(in[i].re) > 0.9 || fabs(in[i].im) > 0.9) h->overload = 1;
            /* in[i] * windowCpx[i] with windowCpx = (w, 0): full complex product of the reference */
            h->time[i].re = in[i].re * h->win[i] - in[i].im * 0.0;
            h->time[i].im = in[i].re * 0.0 + in[i].im * h->win[i];
        }
        for (int i = h->spb; i < h->N; i++) h->time[i].re = h->time[i].im = 0.0;
    } else {
        memset(h->time, 0, sizeof(ocpx) * h->N);
        memcpy(h->time, in, sizeof(ocpx) * n);
    }
}
void orc_fftop_forward(void *p, const double *in, double *out, int n) {  /* fftw.cpp:31-46 */
    orc_fftop *h = (orc_fftop *)p;
    orc_apply_window(h, (const ocpx *)in, n);
    memcpy(h->freq, h->time, sizeof(ocpx) * h->N);
    orc_fft_run(h->fft, h->freq, -1);
    memcpy(out, h->freq, sizeof(ocpx) * h->N);
}
void orc_fftop_inverse(void *p, const double *in, double *out, int n) {  /* fftw.cpp:49-67 */
    orc_fftop *h = (orc_fftop *)p;
    memset(h->freq, 0, sizeof(ocpx) * h->N);
    memcpy(h->freq, in, sizeof(ocpx) * n);
    orc_fft_run(h->fft, h->freq, +1);
    memcpy(out, h->freq, sizeof(ocpx) * h->N);
}
int orc_fftop_spectrum(void *p, const double *in, double *out, int n) {  /* fftw.cpp:70-81 */
    orc_fftop *h = (orc_fftop *)p;
    orc_apply_window(h, (const ocpx *)in, n);
    memcpy(h->freq, h->time, sizeof(ocpx) * h->N);
    orc_fft_run(h->fft, h->freq, -1);
    const int mid = h->N / 2;
    for (int i = 0; i < h->N; i++) {
        const ocpx X = h->freq[(i + mid) % h->N];            /* m_unfoldInOrder, fft.cpp:183-225 */
        double asd = sqrt(X.re * X.re + X.im * X.im) / h->cg;  /* fft.cpp:350 */
        double psd = asd * asd;                              /* fft.cpp:353 */
        psd /= h->maxbin;
        asd /= h->maxbin;
        double binAmp = (asd + h->amp[i]) / 2;               /* fft.cpp:380 */
        h->power[i] = psd;
        h->amp[i] = asd;
        double db = binAmp == 0 ? -120.0 : 20 * log10(binAmp);  /* db.h:44-48 */
        if (db < -120.0) db = -120.0;                        /* DB::clip */
        if (db > 0.0) db = 0.0;
        out[i] = db;
    }
    return h->overload;
}
void orc_fftop_set_amp_state(void *p, const double *amp) { memcpy(((orc_fftop *)p)->amp, amp, sizeof(double) * ((orc_fftop *)p)->N); }

/* ------------------------------------------------------------------------------------------------------------
 * CFir (pebblelib/fir.cpp): Kaiser LP/HP design (:246-337, :352-440) and the ring-buffer ProcessFilter (:106-132).
 * ---------------------------------------------------------------------------------------------------------- */
#define ORC_MAX_NUMCOEF 75
typedef struct {
    int ntaps, state;
    double icoef[ORC_MAX_NUMCOEF * 2], qcoef[ORC_MAX_NUMCOEF * 2];
    ocpx z[ORC_MAX_NUMCOEF];
} orc_cfir;

static double orc_izero(double x) {  /* fir.cpp:494-512 */
    double x2 = x / 2.0, sum = 1.0, ds = 1.0, di = 1.0, tmp;
    do { tmp = x2 / di; tmp *= tmp; ds *= tmp; sum += ds; di += 1.0; } while (ds >= 1e-9 * sum);
    return sum;
}
static double orc_beta(double Astop) {
    if (Astop < 20.96) return 0;
    if (Astop >= 50.0) return .1102 * (Astop - 8.71);
    return .5842 * pow((Astop - 20.96), 0.4) + .07886 * (Astop - 20.96);
}
static void orc_cfir_finish(orc_cfir *f, const double *coef) {
    for (int n = 0; n < f->ntaps; n++) {
        f->icoef[n] = f->icoef[n + f->ntaps] = coef[n];
        f->qcoef[n] = f->qcoef[n + f->ntaps] = coef[n];
    }
    memset(f->z, 0, sizeof(f->z));
    f->state = 0;
}
void *orc_cfir_create(void) { return calloc(1, sizeof(orc_cfir)); }
void orc_cfir_destroy(void *p) { free(p); }
int orc_cfir_init_lp(void *p, int NumTaps, double Scale, double Astop, double Fpass, double Fstop, double Fs) {
    orc_cfir *f = (orc_cfir *)p;
    double coef[ORC_MAX_NUMCOEF];
    double normFpass = Fpass / Fs, normFstop = Fstop / Fs, normFcut = (normFstop + normFpass) / 2.0;
    double Beta = orc_beta(Astop);
    f->ntaps = (Astop - 8.0) / (2.285 * ORC_TWOPI * (normFstop - normFpass)) + 1;
    if (f->ntaps > ORC_MAX_NUMCOEF) f->ntaps = ORC_MAX_NUMCOEF;
    if (f->ntaps < 3) f->ntaps = 3;
    if (NumTaps) f->ntaps = NumTaps;
    double fCenter = .5 * (double)(f->ntaps - 1), izb = orc_izero(Beta);
    for (int n = 0; n < f->ntaps; n++) {
        double x = (double)n - fCenter, c;
        if ((double)n == fCenter) c = 2.0 * normFcut;
        else c = (double)sin(ORC_TWOPI * x * normFcut) / (ORC_PI * x);
        x = ((double)n - ((double)f->ntaps - 1.0) / 2.0) / (((double)f->ntaps - 1.0) / 2.0);
        coef[n] = Scale * c * orc_izero(Beta * sqrt(1 - (x * x))) / izb;
    }
    orc_cfir_finish(f, coef);
    return f->ntaps;
}
int orc_cfir_init_hp(void *p, int NumTaps, double Scale, double Astop, double Fpass, double Fstop, double Fs) {
    orc_cfir *f = (orc_cfir *)p;
    double coef[ORC_MAX_NUMCOEF];
    double normFpass = Fpass / Fs, normFstop = Fstop / Fs, normFcut = (normFstop + normFpass) / 2.0;
    double Beta = orc_beta(Astop);
    f->ntaps = (Astop - 8.0) / (2.285 * ORC_TWOPI * (normFpass - normFstop)) + 1;
    if (f->ntaps > (ORC_MAX_NUMCOEF - 1)) f->ntaps = ORC_MAX_NUMCOEF - 1;
    if (f->ntaps < 3) f->ntaps = 3;
    f->ntaps |= 1;
    if (NumTaps) f->ntaps = NumTaps;
    double izb = orc_izero(Beta), fCenter = .5 * (double)(f->ntaps - 1);
    for (int n = 0; n < f->ntaps; n++) {
        double x = (double)n - (double)(f->ntaps - 1) / 2.0, c;
        if ((double)n == fCenter) c = 1.0 - 2.0 * normFcut;
        else c = (double)(sin(ORC_PI * x) / (ORC_PI * x) - sin(ORC_TWOPI * x * normFcut) / (ORC_PI * x));
        x = ((double)n - ((double)f->ntaps - 1.0) / 2.0) / (((double)f->ntaps - 1.0) / 2.0);
        coef[n] = Scale * c * orc_izero(Beta * sqrt(1 - (x * x))) / izb;
    }
    orc_cfir_finish(f, coef);
    return f->ntaps;
}
int orc_cfir_taps(void *p, double *out) {
    orc_cfir *f = (orc_cfir *)p;
    for (int i = 0; i < f->ntaps; i++) out[i] = f->icoef[i];
    return f->ntaps;
}
void orc_cfir_process(void *p, int n, const double *in_, double *out_) {  /* fir.cpp:106-132; in == out allowed */
    orc_cfir *f = (orc_cfir *)p;
    const ocpx *in = (const ocpx *)in_;
    ocpx *out = (ocpx *)out_;
    for (int i = 0; i < n; i++) {
        f->z[f->state] = in[i];
        const double *HI = f->icoef + f->ntaps - f->state, *HQ = f->qcoef + f->ntaps - f->state;
        const ocpx *Z = f->z;
        ocpx acc;
        acc.re = (*HI++ * Z->re);
        acc.im = (*HQ++ * (Z++)->im);
        for (int j = 1; j < f->ntaps; j++) {
            acc.re += (*HI++ * Z->re);
            acc.im += (*HQ++ * (Z++)->im);
        }
        if (--f->state < 0) f->state += f->ntaps;
        out[i] = acc;
    }
}

/* ------------------------------------------------------------------------------------------------------------
 * Demod_AM::processBlockFiltered (application/demod/demod_am.cpp:40-64), Demod_NFM (demod_nfm.cpp)
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct {
    double rate;
    double am_dc, am_dc_last; orc_cfir am_lp;
    float pll_phase, nco_freq, freq_err_dc, lo, hi, alpha, beta, dc_alpha, out_gain; orc_cfir nfm_lp;
    ocpx fm2_last;  /* a function-static in the reference (demod_nfm.cpp:128) */
} orc_demod;

void *orc_demod_create(double rate) {
    orc_demod *h = (orc_demod *)calloc(1, sizeof(orc_demod));
    h->rate = rate;
    orc_cfir_init_lp(&h->am_lp, 0, 1.0, 50.0, 16000.0, 16000.0 * 1.8, rate);  /* Demod_AM ctor: setBandwidth(16000) */
    double norm = ORC_TWOPI / rate;                                           /* Demod_NFM::init, demod_nfm.cpp:44-66 */
    h->lo = -15000.0 * norm; h->hi = 15000.0 * norm;
    h->alpha = 2.0 * .707 * 3000.0 * norm;
    h->beta = (h->alpha * h->alpha) / (4.0 * .707 * .707);
    h->out_gain = 1.0;
    h->dc_alpha = (1.0 - exp(-1.0 / (rate * 0.001)));
    orc_cfir_init_lp(&h->nfm_lp, 0, 1.0, 50.0, 3000.0, 1.6 * 3000.0, rate);
    return h;
}
void orc_demod_destroy(void *p) { free(p); }
void orc_demod_set_am_bandwidth(void *p, double bw) {  /* demod_am.cpp:18-22 */
    orc_demod *h = (orc_demod *)p;
    orc_cfir_init_lp(&h->am_lp, 0, 1.0, 50.0, bw, bw * 1.8, h->rate);
}
void orc_demod_am(void *p, int n, const double *in_, double *out_) {
    orc_demod *h = (orc_demod *)p;
    const ocpx *in = (const ocpx *)in_;
    ocpx *out = (ocpx *)out_;
    for (int i = 0; i < n; i++) {
        double mag = sqrt(in[i].re * in[i].re + in[i].im * in[i].im);
        h->am_dc = (0.9999f * h->am_dc_last) + mag;   /* DC_ALPHA is a float literal */
        double amOut = h->am_dc - h->am_dc_last;
        h->am_dc_last = h->am_dc;
        out[i].re = amOut; out[i].im = amOut;
    }
    orc_cfir_process(&h->am_lp, n, out_, out_);
}
static double orc_phase_cpx(ocpx c) {  /* CpxUtil::phaseCpx, cpx.cpp:5-21 */
    double tmp = atan(c.im / ((c.re == 0) ? 1e-200 : c.re));
    if (c.re < 0 && c.im < 0) tmp -= M_PI;
    else if (c.re < 0 && c.im >= 0) tmp += M_PI;
    return tmp;
}
void orc_demod_fm2(void *p, int n, const double *in_, double *out_) {  /* demod_nfm.cpp:124-140 */
    orc_demod *h = (orc_demod *)p;
    const ocpx *in = (const ocpx *)in_;
    ocpx *out = (ocpx *)out_;
    for (int i = 0; i < n; i++) {
        ocpx prod;
        prod.re = in[i].re * h->fm2_last.re + in[i].im * h->fm2_last.im;
        prod.im = in[i].im * h->fm2_last.re - in[i].re * h->fm2_last.im;
        out[i].re = orc_phase_cpx(prod) * .0005;
        out[i].im = orc_phase_cpx(prod) * .0005;
        h->fm2_last = in[i];
    }
}
void orc_demod_nfm_pll(void *p, int n, const double *in_, double *out_) {  /* demod_nfm.cpp:225-257 */
    orc_demod *h = (orc_demod *)p;
    const ocpx *in = (const ocpx *)in_;
    ocpx *out = (ocpx *)out_;
    for (int i = 0; i < n; i++) {
        double ncoSin = sinf(h->pll_phase);   /* sin(float) -> float overload in the reference build */
        double ncoCos = cosf(h->pll_phase);
        double tr = ncoCos * in[i].re - ncoSin * in[i].im;
        double ti = ncoCos * in[i].im + ncoSin * in[i].re;
        double phzerror = -atan2(ti, tr);
        h->nco_freq += (h->beta * phzerror);
        if (h->nco_freq > h->hi) h->nco_freq = h->hi;
        else if (h->nco_freq < h->lo) h->nco_freq = h->lo;
        h->pll_phase += (h->nco_freq + h->alpha * phzerror);
        h->freq_err_dc = (1.0 - h->dc_alpha) * h->freq_err_dc + h->dc_alpha * h->nco_freq;
        out[i].re = (h->nco_freq - h->freq_err_dc) * h->out_gain;
        out[i].im = 0.0;
    }
    h->pll_phase = fmod(h->pll_phase, ORC_TWOPI);
    orc_cfir_process(&h->nfm_lp, n, out_, out_);
}

/* ------------------------------------------------------------------------------------------------------------
 * AGC (application/agc.cpp:52-82 setAgcMode, :84-207 processBlock, :237-309 setParameters; constants agc.h:25-58)
 * ---------------------------------------------------------------------------------------------------------- */
#define ORC_AGC_BUF 2048
typedef struct {
    double sample_rate, params_rate;
    int mode, use_hang, threshold, slope, decay;
    double manual_gain, knee, gain_slope, fixed_gain, attack_rise, attack_fall, decay_rise, decay_fall;
    int hang_time, delay_samples, window_samples, sig_ptr, mag_pos, hang_timer;
    double peak, attack_avg, decay_avg;
    ocpx sig[ORC_AGC_BUF];
    double mag[ORC_AGC_BUF];
} orc_agc;

static void orc_agc_set_parameters(orc_agc *h, int useHang, int threshold, int slopeFactor, int decay) {
    const float DELAY_TIMECONST = .015, WINDOW_TIMECONST = .018, ATTACK_RISE_TIMECONST = .002, ATTACK_FALL_TIMECONST = .005,
                DECAY_RISEFALL_RATIO = .3, RELEASE_TIMECONST = .05, AGC_OUTSCALE = 0.7;
    if (h->mode == 0) { h->manual_gain = pow(10, (threshold / 5) / 20.0); return; }
    threshold = -threshold;
    h->manual_gain = 1;
    if (useHang == h->use_hang && threshold == h->threshold && slopeFactor == h->slope && decay == h->decay) return;
    h->use_hang = useHang; h->threshold = threshold; h->slope = slopeFactor; h->decay = decay;
    if (h->params_rate != h->sample_rate) {
        h->params_rate = h->sample_rate;
        for (int i = 0; i < ORC_AGC_BUF; i++) { h->sig[i].re = h->sig[i].im = 0.0; h->mag[i] = -16.0; }
        h->sig_ptr = 0; h->hang_timer = 0; h->peak = -16.0; h->decay_avg = -5.0; h->attack_avg = -5.0; h->mag_pos = 0;
    }
    const double fs = h->params_rate;
    h->knee = (double)h->threshold / 20.0;
    h->gain_slope = h->slope / (100.0);
    h->fixed_gain = AGC_OUTSCALE * pow(10.0, h->knee * (h->gain_slope - 1.0));
    h->attack_rise = (1.0 - exp(-1.0 / (fs * ATTACK_RISE_TIMECONST)));
    h->attack_fall = (1.0 - exp(-1.0 / (fs * ATTACK_FALL_TIMECONST)));
    h->decay_rise = (1.0 - exp(-1.0 / (fs * (double)h->decay * .001 * DECAY_RISEFALL_RATIO)));
    h->hang_time = (int)(fs * (double)h->decay * .001);
    if (h->use_hang) h->decay_fall = (1.0 - exp(-1.0 / (fs * RELEASE_TIMECONST)));
    else h->decay_fall = (1.0 - exp(-1.0 / (fs * (double)h->decay * .001)));
    h->delay_samples = (int)(fs * DELAY_TIMECONST);
    h->window_samples = (int)(fs * WINDOW_TIMECONST);
    if (h->delay_samples >= ORC_AGC_BUF - 1) h->delay_samples = ORC_AGC_BUF - 1;
}
void orc_agc_set_mode(void *p, int mode, int threshold) {
    orc_agc *h = (orc_agc *)p;
    int decay = 200;
    h->mode = mode;
    if (mode == 1) decay = 100; else if (mode == 3) decay = 500; else if (mode == 2) decay = 250; else if (mode == 4) decay = 2000;
    orc_agc_set_parameters(h, 0, threshold, 0, decay);
}
void orc_agc_set_threshold(void *p, int g) {  /* AGC::setAgcThreshold, agc.cpp:43-50: the delay line and averages are kept */
    orc_agc *h = (orc_agc *)p;
    orc_agc_set_parameters(h, h->use_hang, g, h->slope, h->decay);
}
void *orc_agc_create(double sample_rate) {
    orc_agc *h = (orc_agc *)calloc(1, sizeof(orc_agc));
    h->sample_rate = sample_rate; h->params_rate = 100.0;
    orc_agc_set_mode(h, 0, 1);
    return h;
}
void orc_agc_destroy(void *p) { free(p); }
double orc_agc_manual_gain(void *p) { return ((orc_agc *)p)->manual_gain; }
void orc_agc_process(void *p, int n, const double *in_, double *out_) {
    orc_agc *h = (orc_agc *)p;
    const ocpx *in = (const ocpx *)in_;
    ocpx *out = (ocpx *)out_;
    const float MIN_CONSTANT = 1e-8, AGC_OUTSCALE = 0.7;
    if (h->mode == 0) {
        for (int i = 0; i < n; i++) { out[i].re = h->manual_gain * in[i].re; out[i].im = h->manual_gain * in[i].im; }
        return;
    }
    for (int i = 0; i < n; i++) {
        ocpx x = in[i];
        ocpx delayed = h->sig[h->sig_ptr];
        h->sig[h->sig_ptr++] = x;
        if (h->sig_ptr >= h->delay_samples) h->sig_ptr = 0;
        double mag = fabs(x.re), mim = fabs(x.im);
        if (mim > mag) mag = mim;
        mag = log10(mag + MIN_CONSTANT) - log10f(1.0f);
        double tmp = h->mag[h->mag_pos];
        h->mag[h->mag_pos++] = mag;
        if (h->mag_pos >= h->window_samples) h->mag_pos = 0;
        if (mag > h->peak) h->peak = mag;
        else if (tmp == h->peak) {
            h->peak = -8.0;
            for (int k = 0; k < h->window_samples; k++) { tmp = h->mag[k]; if (tmp > h->peak) h->peak = tmp; }
        }
        if (h->peak > h->attack_avg) h->attack_avg = (1.0 - h->attack_rise) * h->attack_avg + h->attack_rise * h->peak;
        else h->attack_avg = (1.0 - h->attack_fall) * h->attack_avg + h->attack_fall * h->peak;
        if (h->use_hang) {
            if (h->peak > h->decay_avg) { h->decay_avg = (1.0 - h->decay_rise) * h->decay_avg + h->decay_rise * h->peak; h->hang_timer = 0; }
            else if (h->hang_timer < h->hang_time) h->hang_timer++;
            else h->decay_avg = (1.0 - h->decay_fall) * h->decay_avg + h->decay_fall * h->peak;
        } else {
            if (h->peak > h->decay_avg) h->decay_avg = (1.0 - h->decay_rise) * h->decay_avg + h->decay_rise * (h->peak);
            else h->decay_avg = (1.0 - h->decay_fall) * h->decay_avg + h->decay_fall * (h->peak);
        }
        mag = h->attack_avg > h->decay_avg ? h->attack_avg : h->decay_avg;
        double gain = mag <= h->knee ? h->fixed_gain : AGC_OUTSCALE * pow(10.0, mag * (h->gain_slope - 1.0));
        out[i].re = delayed.re * gain; out[i].im = delayed.im * gain;
    }
}

/* ------------------------------------------------------------------------------------------------------------
 * Receiver::processIQData, non-WFM branch (application/receiver.cpp:903-996): downconvert -> frame accumulation ->
 * gain restore -> FastFIR -> AGC -> demod. demod_mode: DeviceInterface::DemodMode (0 AM, 2 FMN, >=5 pass-through).
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct {
    int fpb, fill, demod_mode, gain_stages;
    double rate, demod_rate;
    void *dc, *bp, *agc, *demod;
    ocpx *sample_buf, *work, *tmp;
} orc_receiver;
void *orc_receiver_create(double sample_rate, int fpb, double max_bw) {
    orc_receiver *h = (orc_receiver *)calloc(1, sizeof(orc_receiver));
    h->fpb = fpb; h->rate = sample_rate; h->gain_stages = -1;
    h->dc = orc_downconvert_create();
    h->demod_rate = orc_downconvert_set_data_rate(h->dc, sample_rate, max_bw, 0);
    h->bp = orc_fastfir_create(2048, 1025);
    h->agc = orc_agc_create((unsigned)h->demod_rate);
    h->demod = orc_demod_create((unsigned)h->demod_rate);
    h->sample_buf = (ocpx *)calloc(2 * fpb + 65536, sizeof(ocpx));
    h->work = (ocpx *)calloc(65536, sizeof(ocpx));
    h->tmp = (ocpx *)calloc(fpb + 2048, sizeof(ocpx));
    return h;
}
void orc_receiver_destroy(void *p) {
    orc_receiver *h = (orc_receiver *)p;
    if (!h) return;
    orc_downconvert_destroy(h->dc); orc_fastfir_destroy(h->bp); orc_agc_destroy(h->agc); orc_demod_destroy(h->demod);
    free(h->sample_buf); free(h->work); free(h->tmp); free(h);
}
double orc_receiver_demod_rate(void *p) { return ((orc_receiver *)p)->demod_rate; }
void orc_receiver_config(void *p, double mixer, double lo, double hi, int demod_mode, int agc_mode, int agc_threshold) {
    orc_receiver *h = (orc_receiver *)p;
    orc_downconvert_set_frequency(h->dc, mixer);
    orc_fastfir_setup(h->bp, lo, hi, 0, h->demod_rate);
    h->demod_mode = demod_mode;
    orc_agc_set_mode(h->agc, agc_mode, agc_threshold);
}
void orc_receiver_set_am_bandwidth(void *p, double bw) { orc_demod_set_am_bandwidth(((orc_receiver *)p)->demod, bw); }
/* One processIQData call with n <= 65536 input samples (clobbers `in` like the reference). Frames are processed as
 * soon as framesPerBuffer decimated samples are available; returns the number of output samples written. */
int orc_receiver_process(void *p, int n, double *in, double *out_) {
    orc_receiver *h = (orc_receiver *)p;
    ocpx *out = (ocpx *)out_;
    int nout = 0;
    int k = orc_downconvert_process(h->dc, n, in, (double *)h->work);                 /* receiver.cpp:908 */
    for (int i = 0; i < k; i++) h->sample_buf[h->fill++] = h->work[i];                 /* :922-924 */
    while (h->fill >= h->fpb) {                                                        /* :927-931 */
        const int S = h->gain_stages >= 0 ? h->gain_stages : orc_downconvert_num_stages(h->dc);
        const double g = pow(10, (S * 2) / 20.0);                                      /* :935-938 */
        for (int i = 0; i < h->fpb; i++) { h->sample_buf[i].re *= g; h->sample_buf[i].im *= g; }
        orc_fastfir_process(h->bp, h->fpb, (double *)h->sample_buf, (double *)h->tmp); /* :950 */
        orc_agc_process(h->agc, h->fpb, (double *)h->tmp, (double *)h->tmp);           /* :983 (AGC_OFF and delay-line modes are in-place safe) */
        ocpx *dst = out + nout;
        if (h->demod_mode == 0) orc_demod_am(h->demod, h->fpb, (double *)h->tmp, (double *)dst);
        else if (h->demod_mode == 2) orc_demod_nfm_pll(h->demod, h->fpb, (double *)h->tmp, (double *)dst);
        else memcpy(dst, h->tmp, sizeof(ocpx) * h->fpb);                               /* demod.cpp:127-138 */
        nout += h->fpb;
        memmove(h->sample_buf, h->sample_buf + h->fpb, sizeof(ocpx) * (h->fill - h->fpb));
        h->fill -= h->fpb;
    }
    return nout;
}

/* ------------------------------------------------------------------------------------------------------------
 * Decimator — the alternative decimation chain (pebblelib/decimator.cpp). Restates the path the reference runs:
 * m_useVdsp = m_combineStages = true (decimator.cpp:9-10), USE_MATLAB taps (:4), convolveVDsp2 (:593-659) with
 * vDSP_zrdesampD as the scalar loop Apple documents (quoted at :637-647), processCIC3 (:719-740).
 * ---------------------------------------------------------------------------------------------------------- */
#include "dec_coefs_c.h"
#define ORC_DEC_MAXLEN 32768  /* maxResultLen, decimator.h:193 */
typedef struct {
    int taps, D;              /* taps == 0: CIC3 (m_useCIC3) */
    const double *h;
    ocpx x_odd, x_even;       /* m_xOdd, m_xEven */
    ocpx *last_x;             /* m_lastXVDsp: taps-1 carried samples, then the call's input */
} orc_dec_stage;
typedef struct {
    orc_dec_stage st[16];
    int nst, dec_by2;
    float rate;
    ocpx *buf[2];
    int buf_len;
} orc_decimator;

void *orc_decimator_create(unsigned sample_rate, unsigned buffer_size) {  /* decimator.cpp:6-45 */
    orc_decimator *d = (orc_decimator *)calloc(1, sizeof(orc_decimator));
    d->rate = (float)sample_rate;
    d->buf_len = ORC_DEC_MAXLEN;   /* the reference sizes its split buffers 2*bufferSize; the stage buffers cap a call at maxResultLen */
    (void)buffer_size;
    d->buf[0] = (ocpx *)calloc(d->buf_len, sizeof(ocpx));
    d->buf[1] = (ocpx *)calloc(d->buf_len, sizeof(ocpx));
    return d;
}
static void orc_dec_clear(orc_decimator *d) {
    for (int i = 0; i < d->nst; i++) free(d->st[i].last_x);
    d->nst = 0;
}
void orc_decimator_destroy(void *p) {
    orc_decimator *d = (orc_decimator *)p;
    if (!d) return;
    orc_dec_clear(d);
    free(d->buf[0]); free(d->buf[1]); free(d);
}
/* decimator.cpp:64-149 */
float orc_decimator_build_chain(void *p, unsigned rate_in, unsigned protect_bw, unsigned rate_out) {
    orc_decimator *d = (orc_decimator *)p;
    orc_dec_clear(d);
    float rate = (float)rate_in;                        /* m_decimatedSampleRate is a float (decimator.h:250) */
    const double bw = (double)protect_bw;               /* m_protectBandWidth is a double (:251) */
    const unsigned min_out = rate_out > 0 ? rate_out : 15000u;
    d->dec_by2 = 0;
    while (rate > min_out) {
        int taps = -1;
        const double *h = NULL;
        if (rate >= (bw / orc_DecCicWpass)) taps = 0;
        else
            for (int k = 1; k < ORC_DEC_NUM_HB; k++)    /* hb7 (k = 0) is compiled out of the chain (:81-86) */
                if (rate >= (bw / orc_DecHbWpass[k])) { taps = orc_DecHbLen[k]; h = orc_DecHbTaps + orc_DecHbOff[k]; break; }
        if (taps < 0) break;                            /* :124-126 */
        d->dec_by2++;
        if (d->nst == 0 || d->st[d->nst - 1].taps != taps) {   /* :130-136 */
            orc_dec_stage *s = &d->st[d->nst++];
            memset(s, 0, sizeof(*s));
            s->taps = taps; s->D = 2; s->h = h;
            s->last_x = (ocpx *)calloc(ORC_DEC_MAXLEN, sizeof(ocpx));
        } else {
            d->st[d->nst - 1].D *= 2;                   /* :141 */
        }
        rate /= 2;
    }
    d->rate = rate;
    return rate;
}
int orc_decimator_dec_by2_stages(void *p) { return ((orc_decimator *)p)->dec_by2; }
int orc_decimator_num_stages(void *p) { return ((orc_decimator *)p)->nst; }
int orc_decimator_stage(void *p, int i, int *taps, int *D) {
    orc_decimator *d = (orc_decimator *)p;
    if (i < 0 || i >= d->nst) return -1;
    *taps = d->st[i].taps; *D = d->st[i].D;
    return 0;
}
static int orc_dec_cic3(orc_dec_stage *s, const ocpx *in, ocpx *out, int n) {   /* decimator.cpp:719-740 */
    int m = 0;
    for (int i = 0; i < n; i += s->D) {
        const ocpx even = in[i], odd = in[i + 1];
        out[m].re = .125 * (odd.re + s->x_even.re + 3.0 * (s->x_odd.re + even.re));
        out[m].im = .125 * (odd.im + s->x_even.im + 3.0 * (s->x_odd.im + even.im));
        s->x_odd = odd; s->x_even = even;
        m++;
    }
    return m;
}
static int orc_dec_hb(orc_dec_stage *s, const ocpx *x, ocpx *y, int xLen) {     /* convolveVDsp2, decimator.cpp:593-659 */
    const int hLen = s->taps, dLen = hLen - 1, D = s->D;
    if (xLen < hLen) {   /* :604-626 — plain decimation; the refill below indexes x past xLen exactly as the reference does */
        int yCnt = 0;
        for (int i = 0; i < xLen; i += D) y[yCnt++] = x[i];
        const int save = dLen - xLen;
        for (int i = 0; i < hLen; i++) {
            if (i < save) { s->last_x[i].re = 0; s->last_x[i].im = 0; }
            else s->last_x[i] = x[i];
        }
        return yCnt;
    }
    memcpy(s->last_x + dLen, x, sizeof(ocpx) * xLen);                            /* :635 */
    const int yLen = xLen / D;
    for (int n = 0; n < yLen; n++) {                                              /* vDSP_zrdesampD, :637-648 */
        double sr = 0, si = 0;
        for (int k = 0; k < hLen; k++) {
            sr += s->last_x[n * D + k].re * s->h[k];
            si += s->last_x[n * D + k].im * s->h[k];
        }
        y[n].re = sr; y[n].im = si;
    }
    memcpy(s->last_x, x + (xLen - dLen), sizeof(ocpx) * dLen);                    /* :652-654 */
    return yLen;
}
/* decimator.cpp:152-226; n <= 32768 - 58 */
int orc_decimator_process(void *p, int n, const double *in, double *out) {
    orc_decimator *d = (orc_decimator *)p;
    if (d->nst == 0) { memcpy(out, in, sizeof(ocpx) * n); return n; }             /* :155-160 */
    memcpy(d->buf[0], in, sizeof(ocpx) * n);
    int cur = 0, rem = n;
    for (int i = 0; i < d->nst; i++) {
        orc_dec_stage *s = &d->st[i];
        rem = s->taps ? orc_dec_hb(s, d->buf[cur], d->buf[cur ^ 1], rem) : orc_dec_cic3(s, d->buf[cur], d->buf[cur ^ 1], rem);
        cur ^= 1;
    }
    memcpy(out, d->buf[cur], sizeof(ocpx) * rem);
    return rem;
}

/* ------------------------------------------------------------------------------------------------------------
 * DeviceInterfaceBase::normalizeIQ, block forms (pebblelib/deviceinterfacebase.cpp:648-878; sample structs
 * pebblelib/cpx.h:43-92). fmt: 0 CPX8 (qint8 re,im), 1 CPXU8 (quint8, offset binary), 2 CPX16 (qint16),
 * 3 CPXFLOAT (float), 4 CPX (double). order: DeviceInterface::IQOrder 0 IQ, 1 QI, 2 I only, 3 Q only
 * (device_interfaces.h:140-145); `reverse` swaps IQ <-> QI (:653-657). scale is formed left to right as the
 * reference writes it: (1/128.0 * userGain) * normalizeGain, (1/32768.0 * ...) or userGain * normalizeGain.
 * ---------------------------------------------------------------------------------------------------------- */
void orc_normalize_iq(int fmt, int order, double user_gain, double norm_gain, int reverse, unsigned n, const void *in, double *out_) {
    ocpx *out = (ocpx *)out_;
    double scale;
    if (fmt == 0 || fmt == 1) scale = 1 / 128.0 * user_gain * norm_gain;
    else if (fmt == 2) scale = 1 / 32768.0 * user_gain * norm_gain;
    else scale = user_gain * norm_gain;
    if (order == 0 && reverse) order = 1;
    else if (order == 1 && reverse) order = 0;
    for (unsigned i = 0; i < n; i++) {
        double I, Q;
        switch (fmt) {
            case 0: I = ((const signed char *)in)[2 * i]; Q = ((const signed char *)in)[2 * i + 1]; break;
            case 1: I = ((const unsigned char *)in)[2 * i] - 128.0; Q = ((const unsigned char *)in)[2 * i + 1] - 128.0; break;  /* :700-701 */
            case 2: I = ((const short *)in)[2 * i]; Q = ((const short *)in)[2 * i + 1]; break;
            case 3: I = ((const float *)in)[2 * i]; Q = ((const float *)in)[2 * i + 1]; break;
            default: I = ((const double *)in)[2 * i]; Q = ((const double *)in)[2 * i + 1]; break;
        }
        switch (order) {
            case 0: out[i].re = I * scale; out[i].im = Q * scale; break;
            case 1: out[i].re = Q * scale; out[i].im = I * scale; break;
            case 2: out[i].re = I * scale; out[i].im = out[i].re; break;
            default: out[i].im = Q * scale; out[i].re = out[i].im; break;
        }
    }
}
