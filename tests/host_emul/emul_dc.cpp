// Host emulation of the fused downconvert CTA (pebblesdr_b200/csrc/dc_core.cuh): a pthread team with
// a barrier plays the CTA's threads, memcpy plays the TMA chunk load. Exposed as a C function so the
// CPU test-suite (tests/test_emul.py) can compare the exact kernel data-flow with the reference
// library without a GPU. Test infrastructure only.
#include "../../pebblesdr_b200/csrc/host_design.h"
#include <pthread.h>
#include <cstring>
#include <vector>
#include <cmath>

struct TeamShared {
    pthread_barrier_t bar;
    int T;
};
struct HostCtx {
    TeamShared *sh;
    int t;
    int tid() const { return t; }
    int nthreads() const { return sh->T; }
    void sync() { pthread_barrier_wait(&sh->bar); }
    void load_chunk(int, cpx *dst, const cpx *src, int len) {
        if (t == 0) memcpy(dst, src, sizeof(cpx) * (size_t)len);
    }
    void wait_chunk(int) { pthread_barrier_wait(&sh->bar); }  // copy by thread 0 precedes this barrier
};
struct ThreadArg {
    TeamShared *sh; int t; const DcChain *ch; const DcCall *call; cpx *smem; long long seg;
};
static void *thread_main(void *p) {
    ThreadArg *a = (ThreadArg *)p;
    HostCtx ctx{a->sh, a->t};
    dc_run_segment(ctx, a->ch, a->ch->st, a->ch->coef, *a->call, a->smem, a->seg);
    return nullptr;
}

extern "C" {
// Emulated object: one CDownConvert with carried state across calls.
struct EmulDc {
    DcChain ch;
    std::vector<cpx> state[2];
    int cur;
    uint64_t phase_hi, phase_lo;
    long long stream_pos;
    double in_rate, nco_freq, cw, out_rate;
    int threads;
};
void *emul_dc_create(int threads) {
    EmulDc *d = new EmulDc();
    memset(&d->ch, 0, sizeof(d->ch));
    d->state[0].assign(PB_DC_MAX_STAGES * PB_DC_PREFIX, cpx_make(0, 0));
    d->state[1] = d->state[0];
    d->cur = 0; d->phase_hi = d->phase_lo = 0; d->stream_pos = 0;
    d->in_rate = 100000.0; d->nco_freq = 0; d->cw = 0; d->threads = threads;
    return d;
}
void emul_dc_destroy(void *h) { delete (EmulDc *)h; }
void emul_dc_set_frequency(void *h, double f) {
    EmulDc *d = (EmulDc *)h;
    d->nco_freq = -f + d->cw;  // downconvert.cpp:105-108
    double inc = 6.28318530717958647692528676656 * d->nco_freq / d->in_rate;
    pebble_host::dc_set_osc(d->ch, cos(inc), sin(inc), d->threads);
}
double emul_dc_set_data_rate(void *h, double in_rate, double max_bw) {
    EmulDc *d = (EmulDc *)h;
    std::vector<int> lens;
    d->out_rate = pebble_host::dc_plan(in_rate, max_bw, false, lens);
    pebble_host::dc_fill_chain(d->ch, lens);
    d->in_rate = in_rate;
    double inc = 6.28318530717958647692528676656 * d->nco_freq / d->in_rate;
    pebble_host::dc_set_osc(d->ch, cos(inc), sin(inc), d->threads);
    return d->out_rate;
}
int emul_dc_num_stages(void *h) { return ((EmulDc *)h)->ch.nstages; }
int emul_dc_preroll(void *h) { return ((EmulDc *)h)->ch.preroll; }
int emul_dc_process(void *h, long long n, const double *in, double *out, long long seg_len, int chunk,
                    int front, int group) {
    EmulDc *d = (EmulDc *)h;
    const int S = d->ch.nstages;
    if (n % (1LL << S)) return -1;
    DcCall call;
    call.in = (const cpx *)in; call.out = (cpx *)out;
    call.state_in = d->state[d->cur].data(); call.state_out = d->state[d->cur ^ 1].data();
    call.n = n; call.seg_len = seg_len; call.phase_hi = d->phase_hi; call.phase_lo = d->phase_lo;
    call.stream_pos = d->stream_pos; call.chunk = chunk; call.front = front; call.group = group;
    DcSmem L = dc_smem_layout(call, S);
    long long nseg = (n + seg_len - 1) / seg_len;
    // run segments in REVERSE order to prove they are independent of each other
    for (long long seg = nseg - 1; seg >= 0; --seg) {
        std::vector<cpx> smem((size_t)L.total, cpx_make(NAN, NAN));
        TeamShared sh; sh.T = d->threads;
        pthread_barrier_init(&sh.bar, nullptr, d->threads);
        std::vector<pthread_t> th(d->threads);
        std::vector<ThreadArg> args(d->threads);
        for (int t = 0; t < d->threads; ++t) {
            args[t] = ThreadArg{&sh, t, &d->ch, &call, smem.data(), seg};
            pthread_create(&th[t], nullptr, thread_main, &args[t]);
        }
        for (int t = 0; t < d->threads; ++t) pthread_join(th[t], nullptr);
        pthread_barrier_destroy(&sh.bar);
    }
    d->cur ^= 1;
    pebble_host::phase_advance(d->phase_hi, d->phase_lo, d->ch.turn_hi, d->ch.turn_lo, (uint64_t)n);
    d->stream_pos += n;
    return (int)(n >> S);
}
}
