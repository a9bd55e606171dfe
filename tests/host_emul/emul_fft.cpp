// Host emulation of the shared-memory FFT passes (pebblesdr_b200/csrc/fft_core.cuh): runs every
// "thread" of the CTA sequentially, one pass at a time (a pass boundary == __syncthreads), and checks
// forward against a long-double naive DFT and inverse(forward(x)) == N*x. Built with plain g++.
#include "../../pebblesdr_b200/csrc/fft_core.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cmath>

template <int N, int P, bool INV> struct RunPasses {
    static void fwd(cpx *a, const cpx *tw) {
        RunPasses<N, P - 1, INV>::fwd(a, tw);
        for (int t = 0; t < FftPlan<N>::THREADS; ++t) FftPlan<N>::template pass<P, INV>(a, tw, t, FftPlan<N>::THREADS);
    }
    static void inv(cpx *a, const cpx *tw) {
        for (int t = 0; t < FftPlan<N>::THREADS; ++t) FftPlan<N>::template pass<P, INV>(a, tw, t, FftPlan<N>::THREADS);
        RunPasses<N, P - 1, INV>::inv(a, tw);
    }
};
template <int N, bool INV> struct RunPasses<N, -1, INV> {
    static void fwd(cpx *, const cpx *) {}
    static void inv(cpx *, const cpx *) {}
};

template <int N> int check() {
    typedef FftPlan<N> PL;
    std::vector<cpx> x(N), a(N), tw(N);
    for (int m = 0; m < N; ++m) {
        long double ang = -2.0L * M_PIl * m / N;
        tw[m] = cpx_make((double)cosl(ang), (double)sinl(ang));
    }
    srand(N);
    for (int i = 0; i < N; ++i) x[i] = cpx_make(rand() / (double)RAND_MAX - 0.5, rand() / (double)RAND_MAX - 0.5);
    for (int i = 0; i < N; ++i) a[fft_swz<PL::SH>(i)] = x[i];
    RunPasses<N, PL::NP - 1, false>::fwd(a.data(), tw.data());
    // naive DFT on a subset of bins (all bins for N<=2048)
    double maxerr = 0, maxmag = 0;
    int step = N <= 2048 ? 1 : 7;
    for (int f = 0; f < N; f += step) {
        long double sr = 0, si = 0;
        for (int n = 0; n < N; ++n) {
            long double ang = -2.0L * M_PIl * ((long long)f * n % N) / N;
            long double c = cosl(ang), s = sinl(ang);
            sr += x[n].x * c - x[n].y * s;
            si += x[n].x * s + x[n].y * c;
        }
        cpx got = a[fft_swz<PL::SH>(PL::freq2pos(f))];
        if (PL::pos2freq(PL::freq2pos(f)) != f) { printf("N=%d map mismatch f=%d\n", N, f); return 1; }
        double e = hypot(got.x - (double)sr, got.y - (double)si);
        if (e > maxerr) maxerr = e;
        double m = hypot((double)sr, (double)si);
        if (m > maxmag) maxmag = m;
    }
    RunPasses<N, PL::NP - 1, true>::inv(a.data(), tw.data());
    double maxinv = 0;
    for (int i = 0; i < N; ++i) {
        cpx g = a[fft_swz<PL::SH>(i)];
        double e = hypot(g.x / N - x[i].x, g.y / N - x[i].y);
        if (e > maxinv) maxinv = e;
    }
    printf("N=%d fwd max err %.3e (max |X| %.3e) inverse roundtrip err %.3e\n", N, maxerr, maxmag, maxinv);
    return (maxerr / maxmag < 1e-14 && maxinv < 1e-14) ? 0 : 1;
}

int main() {
    int rc = 0;
    rc |= check<2048>();
    rc |= check<4096>();
    rc |= check<8192>();
    printf(rc ? "FAIL\n" : "OK\n");
    return rc;
}
