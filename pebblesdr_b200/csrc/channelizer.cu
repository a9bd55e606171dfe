// channelizer.cu — the batched multi-channel receiver (BASELINE config 4): nch independent receiver chains
//   CDownConvert (own NCO each, shared stage plan) -> gain restore -> CFastFIR -> AGC_OFF -> Demod (AM / NBFM PLL / SSB)
// over ONE wideband capture, i.e. nch copies of the non-WFM branch of Receiver::processIQData
// (application/receiver.cpp:903-996) that differ only in mixer frequency and demod mode.
// Layout: channel-major. Decimated, filtered and demodulated streams are [channel][sample] rows, so a rank that owns
// channels [c0, c1) owns a contiguous slab (multi-GPU sharding by channel) and per-channel carried state is
// one row of each state array.
//   1. dc_cic_kernel, batched: blockIdx.x = segment*nch + channel (downconvert.cu). Every channel mixes the full-rate
//      stream (the reference algorithm, kept for parity): FP64-bound, the input is shared through L2.
//   2. fastfir_kernel, blockIdx.y = channel, the two scalar gains fused (fastfir.cu).
//   3. chan_demod_kernel: four channels per warp, eight lanes each (the PLL and the DC-block IIR are serial recurrences, F10).
//   4. chan_cfir_kernel: the demod post filter, blockIdx.y = channel, per-mode tap set.
#include "lib_common.h"
#include "demod_kernels.cuh"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

enum { CH_AM = 0, CH_FMN = 2 };   // DeviceInterface::DemodMode (device_interfaces.h:124-137); >= 5: pass-through

struct ChanState {      // per-channel demod state
    PllState pll;       // Demod_NFM (float members, demod_nfm.cpp:44-66)
    float pad;
    double am_dc_last;  // Demod_AM::m_amDcLast (demod_am.cpp:52-54)
};

// chan_demod_kernel: in/out rows of `stride` cpx. The demodulators are serial recurrences (PLL feedback, DC-block IIR), so a
// channel's speed is the latency of ONE sample's dependent chain and nothing else; what the kernel can choose is how many
// issue slots it spends around that chain, because the walks run BESIDE the next part's front end (other streams):
//   * one thread per channel (first version, 32 warps): that thread also issued everything that is NOT on the chain — the
//     uncoalesced row loads, the double->float conversions, the 20-instruction arg(x) polynomial, the stores: 265 cycles per
//     sample for a chain of ~14 FP32 operations;
//   * one warp per channel (1024 warps, every lane repeating the chain on shuffled inputs): 99 cycles per sample alone
//     (207 us per 4096-sample part) but 1024 x 30 instructions per sample step — a third of the GPU's issue slots while it
//     runs; beside the front end both slowed down 1.5-2.8x (event timeline, tools/trace_chz.py);
//   * this form: a warp carries PB_CHZ_C = 8 channels, PB_CHZ_L = 4 lanes each. A block is 32 consecutive samples of each
//     channel: lane s of a channel's group loads samples s, s + 4, ... (64-byte runs), forms the loop-independent part of ITS
//     samples (arg(x) / |x|), then all lanes run the same 32 chain steps on the group's state, picking step j's input out of
//     its owner lane with a width-4 shuffle (issued ahead of the chain: it does not depend on the state). One warp
//     instruction advances 8 recurrences, so the walk costs 1/8 of the issue slots of the warp-per-channel form at the same
//     latency; one warp per CTA, 128 CTAs for 1024 channels.
#define PB_CHZ_L 4                     // lanes per channel
#define PB_CHZ_C (32 / PB_CHZ_L)       // channels per warp
#define PB_CHZ_R (32 / PB_CHZ_L)       // samples per lane in a 32-sample block
// One PLL step (demod_nfm.cpp:225-257) on the chain. th = arg(x) (0 for an exactly-zero sample), th_turns = th / 2 pi,
// nb / na = -beta / -alpha (0 for an exactly-zero sample: phzerror = 0 there, atan2(0, 0) = 0 in the reference).
// arg(x e^{j phase}) = th + phase reduced to [-pi, pi]: the reduction (Cody-Waite, k by the 1.5 * 2^23 trick) is applied to the
// SUM, so no separate wrap sits on the chain, and the zero-sample case rides on the two gains instead of a select: the chain
// is 11 plain FP32 operations, none of which produces or consumes a predicate.
__device__ __forceinline__ float pll_chain(const float th, const float th_turns, const float nb, const float na, const PllConst &c, PllState &v) {
    const float q = fmaf(v.phase, 0.15915494309189535f, th_turns);
    const float k = (q + 12582912.0f) - 12582912.0f;
    float r = fmaf(k, -PB_TWO_PI_HI, v.phase);
    r = fmaf(k, -PB_TWO_PI_LO, r);
    const float e = th + r;                                   // phzerror = -e
    float f = fmaf(nb, e, v.freq);
    f = fminf(fmaxf(f, c.lo), c.hi);
    v.freq = f;
    v.phase = v.phase + fmaf(na, e, f);
    v.dc = fmaf(c.dc_alpha, f, (1.0f - c.dc_alpha) * v.dc);   // its own, shorter chain
    return (f - v.dc) * c.out_gain;
}
static __global__ void __launch_bounds__(32) chan_demod_kernel(const cpx *__restrict__ in, cpx *__restrict__ out_rows, int n, long long stride,
                                         int nch, const int *__restrict__ modes, const int *__restrict__ order, int slots,
                                         PllConst pc, double am_alpha, ChanState *__restrict__ st, int frame,
                                         const unsigned char *__restrict__ open) {
    // order[]: channels grouped by demodulator, each group padded (-1) to whole warps: a warp runs ONE kind of recurrence
    const int lane = threadIdx.x;
    const int s = lane % PB_CHZ_L;                           // this lane's position inside its channel's group
    const int slot0 = blockIdx.x * PB_CHZ_C;
    if (slot0 >= slots) return;
    const int slot = slot0 + lane / PB_CHZ_L;
    const int c = slot < slots ? order[slot] : -1;
    // a squelched channel (receiver.cpp:955-965 returns before AGC and demodulator: their state does not advance) rides along
    // in its warp like a pad slot: nothing is stored for it
    const bool valid = c >= 0 && c < nch && (open == nullptr || open[c] != 0);
    const int c0 = order[slot0];                             // never a pad: groups are padded at their END
    const int mode = modes[c0];
    const cpx *x = in + (long long)(valid ? c : c0) * stride;
    // Both detectors produce a REAL stream (NBFM: (o, 0), AM: (v, v)); the row written here holds those reals, and the post
    // filter, which duplicates / zero-extends them, reads 8 instead of 16 bytes per sample. Same row pitch as the input.
    double *y = (double *)(out_rows + (long long)(valid ? c : c0) * stride);
    // a block = 32 consecutive samples of every channel of the warp; lane s of a group holds samples s, s + L, s + 2L, ...
    // The next block's loads are issued before the current block's 32 chain steps (~2000 cycles: well past the L2 latency).
    if (mode == CH_FMN) {                 // Demod_NFM::processBlockNCO, demod_nfm.cpp:225-257
        PllState v = st[valid ? c : c0].pll;
        if (frame <= 0) frame = n;
        // Three blocks are in flight per channel: block b is walked (th / nbv), block b+1's samples (raw) are turned into
        // arg(x) INSIDE block b's 32 chain steps — one sample every fourth step, straight-line code the scheduler drops into
        // the chain's 4-cycle dependent-issue gaps — and block b+2's loads (nxt) are outstanding. Doing arg(x), the
        // conversions and the stores between two blocks instead cost ~1000 cycles per block on top of the chain's ~1700
        // (tools/lat_probe.cu: the chain with its shuffles, DC tracker and selects alone takes 53 cycles per step; the kernel
        // took 85).
        auto sample = [&](int i) -> cpx { return i < n ? x[i] : cpx_make(1.0, 0.0); };
        cpx raw[PB_CHZ_R], nxt[PB_CHZ_R];
        float th[PB_CHZ_R], nbv[PB_CHZ_R];
#pragma unroll
        for (int m = 0; m < PB_CHZ_R; ++m) {
            const float t = pll_theta(sample(m * PB_CHZ_L + s));            // NaN marks an exactly-zero sample
            th[m] = t != t ? 0.0f : t;
            nbv[m] = t != t ? 0.0f : -pc.beta;
            raw[m] = sample(32 + m * PB_CHZ_L + s);
        }
        for (int i = 0; i < n; i += 32) {
#pragma unroll
            for (int m = 0; m < PB_CHZ_R; ++m) nxt[m] = sample(i + 64 + m * PB_CHZ_L + s);
            const int cnt = n - i < 32 ? n - i : 32;
            float thn[PB_CHZ_R], nbn[PB_CHZ_R], mine[PB_CHZ_R];
#pragma unroll
            for (int m = 0; m < PB_CHZ_R; ++m) mine[m] = 0.0f;
            if (cnt == 32) {
                // no per-step branch here: a branch between steps pins every shuffle right in front of its use and puts
                // its ~25-cycle latency on the chain (measured: 164 instead of ~100 cycles per step)
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const float tj = __shfl_sync(0xffffffffu, th[j / PB_CHZ_L], j % PB_CHZ_L, PB_CHZ_L);
                    const float nbj = __shfl_sync(0xffffffffu, nbv[j / PB_CHZ_L], j % PB_CHZ_L, PB_CHZ_L);
                    const float o = pll_chain(tj, tj * 0.15915494309189535f, nbj, nbj == 0.0f ? 0.0f : -pc.alpha, pc, v);
                    mine[j / PB_CHZ_L] = s == j % PB_CHZ_L ? o : mine[j / PB_CHZ_L];
                    if (j % PB_CHZ_L == 1) {                                // next block's sample j / L: off the chain
                        const float t = pll_theta(raw[j / PB_CHZ_L]);
                        thn[j / PB_CHZ_L] = t != t ? 0.0f : t;
                        nbn[j / PB_CHZ_L] = t != t ? 0.0f : -pc.beta;
                    }
                }
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    if (j < cnt) {                           // uniform
                        const float tj = __shfl_sync(0xffffffffu, th[j / PB_CHZ_L], j % PB_CHZ_L, PB_CHZ_L);
                        const float nbj = __shfl_sync(0xffffffffu, nbv[j / PB_CHZ_L], j % PB_CHZ_L, PB_CHZ_L);
                        const float o = pll_chain(tj, tj * 0.15915494309189535f, nbj, nbj == 0.0f ? 0.0f : -pc.alpha, pc, v);
                        mine[j / PB_CHZ_L] = s == j % PB_CHZ_L ? o : mine[j / PB_CHZ_L];
                    }
                }
#pragma unroll
                for (int m = 0; m < PB_CHZ_R; ++m) { thn[m] = 0.0f; nbn[m] = 0.0f; }
            }
#pragma unroll
            for (int m = 0; m < PB_CHZ_R; ++m)
                if (valid && m * PB_CHZ_L + s < cnt) y[i + m * PB_CHZ_L + s] = (double)mine[m];   // CPX = float there; imaginary part 0
#pragma unroll
            for (int m = 0; m < PB_CHZ_R; ++m) { th[m] = thn[m]; nbv[m] = nbn[m]; raw[m] = nxt[m]; }
            // the reference wraps the phase once per receiver frame (:252) and at the end of the call
            if ((i + 32) % frame == 0 || i + 32 >= n) v.phase = (float)fmod((double)v.phase, 6.28318530717958647692528676656);
        }
        if (valid && s == 0) st[c].pll = v;
    } else if (mode == CH_AM) {           // Demod_AM::processBlockFiltered, demod_am.cpp:40-59
        double d = st[valid ? c : c0].am_dc_last;
        cpx cur[PB_CHZ_R], nxt[PB_CHZ_R];
#pragma unroll
        for (int m = 0; m < PB_CHZ_R; ++m) nxt[m] = m * PB_CHZ_L + s < n ? x[m * PB_CHZ_L + s] : cpx_make(0.0, 0.0);
        for (int i = 0; i < n; i += 32) {
#pragma unroll
            for (int m = 0; m < PB_CHZ_R; ++m) cur[m] = nxt[m];
            if (i + 32 < n) {
#pragma unroll
                for (int m = 0; m < PB_CHZ_R; ++m) nxt[m] = i + 32 + m * PB_CHZ_L + s < n ? x[i + 32 + m * PB_CHZ_L + s] : cpx_make(0.0, 0.0);
            }
            // sqrt: branch-free Newton form, <= 1 ulp from the correctly rounded value (DB::amplitude, db.h:28-30)
            double mg[PB_CHZ_R], mine[PB_CHZ_R];
#pragma unroll
            for (int m = 0; m < PB_CHZ_R; ++m) { mg[m] = pb_sqrt_newton(cur[m].x * cur[m].x + cur[m].y * cur[m].y); mine[m] = 0.0; }
            const int cnt = n - i < 32 ? n - i : 32;
            // all 32 magnitudes of the block into registers first (64 shuffles issued back to back, pipelined), then a recurrence
            // that touches registers only: with the shuffle inside the step the 64-bit shuffle's latency sat on every step
            // (tools/lat_probe.cu: 46.6 cycles per step for a chain of one 8-cycle DFMA)
            double mall[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) mall[j] = __shfl_sync(0xffffffffu, mg[j / PB_CHZ_L], j % PB_CHZ_L, PB_CHZ_L);
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                if (j < cnt) {                                                        // uniform; cnt == 32 except in a ragged tail
                    const double dn = (am_alpha * d) + mall[j];                       // demod_am.cpp:52
                    const double w = dn - d;                                          // demod_am.cpp:53
                    d = dn;
                    mine[j / PB_CHZ_L] = s == j % PB_CHZ_L ? w : mine[j / PB_CHZ_L];
                }
            }
#pragma unroll
            for (int m = 0; m < PB_CHZ_R; ++m)
                if (valid && m * PB_CHZ_L + s < cnt) y[i + m * PB_CHZ_L + s] = mine[m];             // (v, v) in the reference
        }
        if (valid && s == 0) st[c].am_dc_last = d;
    }
    // USB/LSB/...: Demod::processBlock returns `in` (demod.cpp:127-138): nothing to walk, chan_cfir_kernel copies the row
}

// CFir::ProcessFilter over every channel (fir.cpp:106-132): virtual stream = hist ++ in, per-mode taps, blockIdx.y = channel,
// blockIdx.x = tile of PB_FIR_TILE outputs. The detectors' outputs are real streams whose I and Q tap sets are equal
// (Init*Filter designs, fir.cpp:298-302), so ONE real convolution per channel gives both components: AM writes (r, r), NBFM
// (r, +0) — what the complex filter returns for (v, v) and (o, 0) inputs, bit for bit (same tap order, same FMAs; the
// imaginary sum of +-0 products is +0). Register-blocked like cfir_dup_kernel (demod_kernels.cuh): each thread makes
// PB_FIR_R consecutive outputs from a sliding register window, one shared load per tap.
// taps: [2][PB_FIR_MAX] (row 0 AM low-pass, row 1 NFM low-pass). hist rows: PB_FIR_MAX reals, the last ntaps-1 are used.
// Pass-through channels are copied from the band-pass output `bp`.
static __global__ void __launch_bounds__(PB_FIR_THREADS)
chan_cfir_kernel(const double *__restrict__ hist, const cpx *__restrict__ in_rows, long long in_stride, const cpx *__restrict__ bp,
                 int n, const double *__restrict__ taps, int ntaps_am, int ntaps_fm, const int *__restrict__ modes,
                 cpx *__restrict__ out, long long out_stride, double *__restrict__ new_hist, const unsigned char *__restrict__ open) {
    __shared__ double win[PB_FIR_TILE + PB_FIR_MAX + (PB_FIR_TILE + PB_FIR_MAX) / 8 + 1];
    __shared__ double sh_t[PB_FIR_MAX];
    const int c = blockIdx.y;
    const int mode = modes[c];
    cpx *y = out + c * out_stride;
    const int t0 = blockIdx.x * PB_FIR_TILE;
    const int valid = min(PB_FIR_TILE, n - t0);
    if (open != nullptr && open[c] == 0) {   // squelched: no audio for this call (zeros), the filter's delay line is kept
        for (int i = threadIdx.x; i < valid; i += PB_FIR_THREADS) y[t0 + i] = cpx_make(0.0, 0.0);
        if (blockIdx.x == 0)
            for (int i = threadIdx.x; i < PB_FIR_MAX; i += PB_FIR_THREADS) new_hist[(long long)c * PB_FIR_MAX + i] = hist[(long long)c * PB_FIR_MAX + i];
        return;
    }
    if (mode != CH_AM && mode != CH_FMN) {
        const cpx *xb = bp + c * in_stride;
        for (int i = threadIdx.x; i < valid; i += PB_FIR_THREADS) y[t0 + i] = xb[t0 + i];
        return;
    }
    const double *x = (const double *)(in_rows + c * in_stride);
    const int ntaps = mode == CH_AM ? ntaps_am : ntaps_fm;
    const double *tp = taps + (mode == CH_AM ? 0 : PB_FIR_MAX);
    const int hl = PB_FIR_MAX, h = ntaps - 1;
    const double *ho = hist + (long long)c * hl;
    for (int k = threadIdx.x; k < ntaps; k += PB_FIR_THREADS) sh_t[k] = tp[k];
    for (int p = threadIdx.x; p < valid + h; p += PB_FIR_THREADS) {
        const int idx = t0 + p - h;                        // index into `in`; negative: history
        win[fir_slot(p)] = idx < 0 ? ho[hl + idx] : x[idx];
    }
    // the last PB_FIR_MAX samples of hist ++ in become the next history (tile 0 of each channel writes them)
    if (blockIdx.x == 0) {
        double *nh = new_hist + (long long)c * hl;
        for (int i = threadIdx.x; i < hl; i += PB_FIR_THREADS) {
            const long long v = (long long)n + i;          // index into (hist ++ in)
            nh[i] = v < hl ? ho[v] : x[v - hl];
        }
    }
    __syncthreads();
    const int i0 = threadIdx.x * PB_FIR_R;
    if (i0 >= valid) return;
    double re[PB_FIR_R], w[PB_FIR_R];
#pragma unroll
    for (int j = 0; j < PB_FIR_R; ++j) { re[j] = 0.0; w[j] = win[fir_slot(i0 + h + j)]; }
    for (int k0 = 0; k0 < ntaps; k0 += PB_FIR_R) {
#pragma unroll
        for (int kk = 0; kk < PB_FIR_R; ++kk) {
            const int k = k0 + kk;
            if (k < ntaps) {
                const double cI = sh_t[k];
#pragma unroll
                for (int j = 0; j < PB_FIR_R; ++j) re[j] += cI * w[j];
#pragma unroll
                for (int j = PB_FIR_R - 1; j > 0; --j) w[j] = w[j - 1];
                if (k + 1 < ntaps) w[0] = win[fir_slot(i0 + h - k - 1)];
            }
        }
    }
    const bool dup = mode == CH_AM;
#pragma unroll
    for (int j = 0; j < PB_FIR_R; ++j)
        if (i0 + j < valid) y[t0 + i0 + j] = cpx_make(re[j], dup ? re[j] : 0.0);
}

struct pebble_channelizer {
    int nch = 0, S = 0, fpb = 2048;
    double in_rate = 0, max_bw = 0, out_rate = 0;
    pebble_downconvert_t *dc = nullptr;
    pebble_fastfir_t *bp = nullptr;
    std::vector<int> modes;
    int *d_modes = nullptr, *d_order = nullptr;
    std::vector<double> freqs;         // channel centres (for the squelch estimate)
    double bp_lo = 0, bp_hi = 0;       // band-pass edges
    std::vector<unsigned char> open;   // per-channel squelch decision; empty = every channel open
    unsigned char *d_open = nullptr;
    bool open_dirty = false;
    int slots = 0;                     // entries of d_order
    bool modes_dirty = true;
    PllConst pll;
    ChanState *d_state = nullptr;
    double taps[2 * PB_FIR_MAX];
    int ntaps_am = 0, ntaps_fm = 0;
    double *d_taps = nullptr;
    cpx *d_hist[2] = {nullptr, nullptr};
    int hcur = 0;
    int agc_threshold = 20;
    double am_bandwidth = 16000.0;     // Demod_AM ctor default (demod_am.cpp:9)
    cpx *d_dec = nullptr, *d_bp = nullptr, *d_tmp = nullptr;
    size_t row_cap = 0;                // samples per channel the work buffers hold
    // The per-channel demodulators are serial recurrences (PLL feedback): one thread per channel walks its row, and that
    // walk takes the same ~1 us per decimated sample whether 1024 or 128 channels are resident. A call is therefore cut
    // into `parts` runs of whole receiver frames; the demodulator + CFir of part k run on `tail` while the caller's stream
    // already down-converts and filters part k+1 (all operators carry their state, so split == unsplit).
    cudaStream_t tail = nullptr, tail2 = nullptr;   // tail: the serial demodulator walks; tail2: their post filters
    cudaEvent_t ev_front = nullptr, ev_tail = nullptr, ev_demod = nullptr, ev_cfir = nullptr;
    int parts = 4;
};

static int chz_upload_taps(pebble_channelizer *h) {
    h->ntaps_am = pebble_host::cfir_design_lp(0, 1.0, 50.0, h->am_bandwidth, h->am_bandwidth * 1.8, h->out_rate, h->taps);           // demod_am.cpp:21
    h->ntaps_fm = pebble_host::cfir_design_lp(0, 1.0, 50.0, 3000.0, 1.6 * 3000.0, h->out_rate, h->taps + PB_FIR_MAX);              // demod_nfm.cpp:60-63
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(h->d_taps, h->taps, sizeof(h->taps), cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    const size_t hb = sizeof(cpx) * PB_FIR_MAX * (size_t)h->nch;
    PB_CUDA(cudaMemset(h->d_hist[0], 0, hb), PEBBLE_ERR_CUDA);      // Init*Filter zeroes the delay line (fir.cpp:305-311)
    PB_CUDA(cudaMemset(h->d_hist[1], 0, hb), PEBBLE_ERR_CUDA);
    return 0;
}

static int chz_reserve(pebble_channelizer *h, size_t row) {
    if (row <= h->row_cap) return 0;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    cudaFree(h->d_dec); cudaFree(h->d_bp); cudaFree(h->d_tmp);
    h->d_dec = h->d_bp = h->d_tmp = nullptr; h->row_cap = 0;
    const size_t nb = sizeof(cpx) * row * (size_t)h->nch;
    PB_CUDA(cudaMalloc(&h->d_dec, nb), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMalloc(&h->d_bp, nb), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMalloc(&h->d_tmp, nb), PEBBLE_ERR_CUDA);
    h->row_cap = row;
    return 0;
}

// PEBBLE_CHZ_TRACE=1: time-stamp every stage of every part with events on its own stream and print the timeline (us from the
// call's first event) to stderr after the call — the only way to see how far the three streams really overlap without nsys.
namespace {
struct ChzTrace {
    struct Mark { const char *what; int part; cudaEvent_t a, b; };
    std::vector<Mark> marks;
    bool on = false;
    void begin(const char *what, int part, cudaStream_t s) {
        if (!on) return;
        Mark m{what, part, nullptr, nullptr};
        cudaEventCreate(&m.a); cudaEventCreate(&m.b);
        cudaEventRecord(m.a, s);
        marks.push_back(m);
    }
    void end(cudaStream_t s) { if (on) cudaEventRecord(marks.back().b, s); }
    void report() {
        if (!on || marks.empty()) return;
        cudaDeviceSynchronize();
        for (const Mark &m : marks) {
            float t0 = 0, t1 = 0;
            const cudaError_t e0 = cudaEventElapsedTime(&t0, marks[0].a, m.a), e1 = cudaEventElapsedTime(&t1, marks[0].a, m.b);
            if (e0 != cudaSuccess || e1 != cudaSuccess) fprintf(stderr, "chz-trace: %s / %s\n", cudaGetErrorString(e0), cudaGetErrorString(e1));
            fprintf(stderr, "chz-trace part %d %-8s %9.1f -> %9.1f us (%7.1f)\n", m.part, m.what, t0 * 1e3, t1 * 1e3, (t1 - t0) * 1e3);
        }
        for (const Mark &m : marks) { cudaEventDestroy(m.a); cudaEventDestroy(m.b); }
        marks.clear();
    }
};
}  // namespace

extern "C" {

pebble_channelizer_t *pebble_channelizer_create(int nchannels, double in_rate, double max_bw) {
    if (nchannels < 1) { pebble::set_error("channelizer: nchannels < 1"); return nullptr; }
    pebble_channelizer *h = new pebble_channelizer();
    h->nch = nchannels; h->in_rate = in_rate; h->max_bw = max_bw;
    h->dc = pebble_downconvert_create();
    h->bp = pebble_fastfir_create();
    if (!h->dc || !h->bp) { pebble_channelizer_destroy(h); return nullptr; }
    if (const char *e = getenv("PEBBLE_CHZ_PARTS")) h->parts = atoi(e) > 0 ? atoi(e) : 1;
    // highest priority: the tail's few serial warps should get SM slots ahead of the next part's many front CTAs
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
    if (!pebble::cuda_ok(cudaStreamCreateWithPriority(&h->tail, cudaStreamNonBlocking, prio_hi), "cudaStreamCreate") ||
        !pebble::cuda_ok(cudaStreamCreateWithPriority(&h->tail2, cudaStreamNonBlocking, prio_hi), "cudaStreamCreate") ||
        !pebble::cuda_ok(cudaEventCreateWithFlags(&h->ev_demod, cudaEventDisableTiming), "cudaEventCreate") ||
        !pebble::cuda_ok(cudaEventCreateWithFlags(&h->ev_cfir, cudaEventDisableTiming), "cudaEventCreate") ||
        !pebble::cuda_ok(cudaEventCreateWithFlags(&h->ev_front, cudaEventDisableTiming), "cudaEventCreate") ||
        !pebble::cuda_ok(cudaEventCreateWithFlags(&h->ev_tail, cudaEventDisableTiming), "cudaEventCreate")) {
        pebble_channelizer_destroy(h);
        return nullptr;
    }
    h->out_rate = pebble_downconvert_set_data_rate(h->dc, in_rate, max_bw);
    h->S = pebble_downconvert_num_stages(h->dc);
    bool ok = pebble_downconvert_set_channels(h->dc, nchannels) == 0 && pebble_fastfir_set_streams(h->bp, nchannels) == 0;
    const size_t hb = sizeof(cpx) * PB_FIR_MAX * (size_t)nchannels;
    ok = ok && pebble::cuda_ok(cudaMalloc(&h->d_modes, sizeof(int) * nchannels), "cudaMalloc") &&
         pebble::cuda_ok(cudaMalloc(&h->d_order, sizeof(int) * ((size_t)nchannels + 3 * 32)), "cudaMalloc") &&
         pebble::cuda_ok(cudaMalloc(&h->d_open, (size_t)nchannels), "cudaMalloc") &&
         pebble::cuda_ok(cudaMalloc(&h->d_state, sizeof(ChanState) * nchannels), "cudaMalloc") &&
         pebble::cuda_ok(cudaMemset(h->d_state, 0, sizeof(ChanState) * nchannels), "cudaMemset") &&
         pebble::cuda_ok(cudaMalloc(&h->d_taps, sizeof(h->taps)), "cudaMalloc") &&
         pebble::cuda_ok(cudaMalloc(&h->d_hist[0], hb), "cudaMalloc") && pebble::cuda_ok(cudaMalloc(&h->d_hist[1], hb), "cudaMalloc");
    if (!ok) { pebble_channelizer_destroy(h); return nullptr; }
    h->modes.assign(nchannels, CH_AM);   // Demod starts in dmAM
    h->freqs.assign(nchannels, 0.0);
    // Demod_NFM::init (demod_nfm.cpp:44-66); members are float
    const double norm = 6.28318530717958647692528676656 / (double)(unsigned)h->out_rate;
    h->pll.lo = (float)(-15000.0 * norm);
    h->pll.hi = (float)(15000.0 * norm);
    h->pll.alpha = (float)(2.0 * .707 * 3000.0 * norm);
    h->pll.beta = (float)((h->pll.alpha * h->pll.alpha) / (4.0 * .707 * .707));
    h->pll.out_gain = 1.0f;
    h->pll.dc_alpha = (float)(1.0 - exp(-1.0 / ((double)(unsigned)h->out_rate * 0.001)));
    if (chz_upload_taps(h) < 0) { pebble_channelizer_destroy(h); return nullptr; }
    return h;
}
void pebble_channelizer_destroy(pebble_channelizer_t *h) {
    if (!h) return;
    if (h->dc) pebble_downconvert_destroy(h->dc);
    if (h->bp) pebble_fastfir_destroy(h->bp);
    cudaFree(h->d_modes); cudaFree(h->d_order); cudaFree(h->d_open); cudaFree(h->d_state); cudaFree(h->d_taps); cudaFree(h->d_hist[0]); cudaFree(h->d_hist[1]);
    cudaFree(h->d_dec); cudaFree(h->d_bp); cudaFree(h->d_tmp);
    if (h->tail) cudaStreamDestroy(h->tail);
    if (h->tail2) cudaStreamDestroy(h->tail2);
    if (h->ev_demod) cudaEventDestroy(h->ev_demod);
    if (h->ev_cfir) cudaEventDestroy(h->ev_cfir);
    if (h->ev_front) cudaEventDestroy(h->ev_front);
    if (h->ev_tail) cudaEventDestroy(h->ev_tail);
    delete h;
}
double pebble_channelizer_out_rate(const pebble_channelizer_t *h) { return h ? h->out_rate : 0.0; }
int pebble_channelizer_num_stages(const pebble_channelizer_t *h) { return h ? h->S : PEBBLE_ERR_ARG; }
int pebble_channelizer_num_channels(const pebble_channelizer_t *h) { return h ? h->nch : PEBBLE_ERR_ARG; }
int pebble_channelizer_set_frequency(pebble_channelizer_t *h, int ch, double f) {   // Receiver::mixerChanged per channel
    if (!h) return PEBBLE_ERR_ARG;
    if (ch >= 0 && ch < h->nch) h->freqs[ch] = f;
    return pebble_downconvert_set_channel_frequency(h->dc, ch, f);
}
int pebble_channelizer_set_filter(pebble_channelizer_t *h, double lo, double hi) {   // BandPassFilter::setBandPass, all channels
    if (!h) return PEBBLE_ERR_ARG;
    h->bp_lo = lo; h->bp_hi = hi;
    return pebble_fastfir_setup(h->bp, lo, hi, 0, h->out_rate);
}
// Squelch (receiver.cpp:955-965): a frame whose band average is below the squelch level returns before AGC and demodulator.
// open[ch] != 0: channel ch is demodulated; 0: its calls produce zeros and its demodulator / post-filter state does not
// advance (the down-converter and the band-pass keep running, as in the reference). open == NULL: every channel open.
int pebble_channelizer_set_squelch(pebble_channelizer_t *h, const unsigned char *open, int nch) {
    if (!h || (open && nch != h->nch)) return PEBBLE_ERR_ARG;
    if (open) h->open.assign(open, open + nch); else h->open.clear();
    h->open_dirty = true;
    return 0;
}
// The decision itself, for every channel at once, from ONE wideband dB spectrum on the device (SignalSpectrum's unprocessed
// output of the capture): SignalStrength::fdEstimate per channel with the channelizer's own band-pass edges and channel
// centres, open = avgDb >= squelch_db. Returns the number of open channels.
int pebble_channelizer_squelch_from_spectrum(pebble_channelizer_t *h, const double *d_spectrum, int bins, unsigned spectrum_rate,
                                             double squelch_db, void *stream) {
    if (!h || !d_spectrum || bins <= 0) return PEBBLE_ERR_ARG;
    std::vector<double> out5((size_t)h->nch * 5);
    const int r = pebble_sigstrength_fd_estimate_device(d_spectrum, bins, spectrum_rate, (float)h->bp_lo, (float)h->bp_hi, h->freqs.data(),
                                                        h->nch, squelch_db, out5.data(), stream);
    if (r < 0) return r;
    h->open.resize(h->nch);
    int nopen = 0;
    for (int c = 0; c < h->nch; ++c) { h->open[c] = out5[(size_t)c * 5 + 4] != 0.0; nopen += h->open[c]; }
    h->open_dirty = true;
    return nopen;
}
int pebble_channelizer_set_demod_mode(pebble_channelizer_t *h, int ch, int mode) {
    if (!h || ch < 0 || ch >= h->nch || mode < 0 || mode > 12) return PEBBLE_ERR_ARG;
    if (mode == 1 || mode == 3 || mode == 4) { pebble::set_error("channelizer: SAM / WFM modes are out of scope"); return PEBBLE_ERR_STATE; }
    h->modes[ch] = mode;
    h->modes_dirty = true;
    return 0;
}
int pebble_channelizer_set_am_bandwidth(pebble_channelizer_t *h, double bw) {
    if (!h) return PEBBLE_ERR_ARG;
    h->am_bandwidth = bw;
    return chz_upload_taps(h);
}
int pebble_channelizer_set_agc_threshold(pebble_channelizer_t *h, int threshold) {   // AGC_OFF only (agc.cpp:86-94)
    if (!h) return PEBBLE_ERR_ARG;
    h->agc_threshold = threshold;
    return 0;
}
// Raw device samples into every channel (normalizeIQ fused into the front end's load; see "Device sample formats").
int pebble_channelizer_set_input_format(pebble_channelizer_t *h, int fmt, int order, double user_gain, double norm_gain) {
    if (!h) return PEBBLE_ERR_ARG;
    return pebble_downconvert_set_input_format(h->dc, fmt, order, user_gain, norm_gain);
}
// 1: polyphase-DFT front end (channels on the grid in_rate / 2^S), 0: one down-converter per channel, -1: no call yet
int pebble_channelizer_front_path(const pebble_channelizer_t *h) { return h ? pebble_downconvert_batch_path(h->dc) : PEBBLE_ERR_ARG; }
int pebble_channelizer_out_count(const pebble_channelizer_t *h, int n) {
    if (!h || n < 0) return PEBBLE_ERR_ARG;
    return n >> h->S;
}

// Decimated streams only (step 1), [channel][n >> S] rows `out_stride` cpx apart.
int pebble_channelizer_downconvert_device(pebble_channelizer_t *h, int n, const void *d_in, double *d_out,
                                          long long out_stride, void *stream) {
    if (!h) return PEBBLE_ERR_ARG;
    return pebble_downconvert_process_channels_device(h->dc, n, d_in, d_out, out_stride, stream);
}

// Whole chain. n must make (n >> S) a multiple of framesPerBuffer (2048), so that every call yields whole receiver
// frames (receiver.cpp:922-931) for every channel; returns the demodulated samples per channel (= n >> S).
int pebble_channelizer_process_device(pebble_channelizer_t *h, int n, const void *d_in, double *d_out, long long out_stride,
                                      void *stream_v) {
    if (!h || n < 0 || (n && (!d_in || !d_out))) { pebble::set_error("channelizer: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    const int nd = n >> h->S;
    if ((n & ((1 << h->S) - 1)) || nd % h->fpb) {
        pebble::set_error("channelizer: n=%d must be a multiple of 2^%d * %d", n, h->S, h->fpb);
        return PEBBLE_ERR_ARG;
    }
    if (out_stride < nd) { pebble::set_error("channelizer: out_stride smaller than the output count"); return PEBBLE_ERR_ARG; }
    cudaStream_t st = (cudaStream_t)stream_v;
    int r = chz_reserve(h, (size_t)nd);
    if (r < 0) return r;
    if (h->modes_dirty) {
        PB_CUDA(cudaMemcpyAsync(h->d_modes, h->modes.data(), sizeof(int) * h->nch, cudaMemcpyHostToDevice, st), PEBBLE_ERR_CUDA);
        // channel slots of chan_demod_kernel: NBFM channels (the longest walks) first, then AM, then pass-through; a warp takes
        // PB_CHZ_C consecutive slots, so every group is padded to a multiple of that
        std::vector<int> order;
        for (int grp = 0; grp < 3; ++grp) {
            for (int c = 0; c < h->nch; ++c) {
                const int g = h->modes[c] == CH_FMN ? 0 : h->modes[c] == CH_AM ? 1 : 2;
                if (g == grp) order.push_back(c);
            }
            while (order.size() % PB_CHZ_C) order.push_back(-1);
        }
        h->slots = (int)order.size();
        PB_CUDA(cudaMemcpy(h->d_order, order.data(), sizeof(int) * order.size(), cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
        h->modes_dirty = false;
    }
    if (h->open_dirty) {
        if (!h->open.empty()) PB_CUDA(cudaMemcpyAsync(h->d_open, h->open.data(), (size_t)h->nch, cudaMemcpyHostToDevice, st), PEBBLE_ERR_CUDA);
        h->open_dirty = false;
    }
    const unsigned char *d_open = h->open.empty() ? nullptr : h->d_open;
    const long long row = (long long)h->row_cap;
    const double gain = pow(10, (h->S * 2) / 20.0);                                                  // receiver.cpp:935-938
    const double agc_gain = pow(10, (h->agc_threshold / 5) / 20.0);                                  // agc.cpp:86-94 (integer division)
    pebble_fastfir_set_fused_scale(h->bp, gain, agc_gain);
    const int frames = nd / h->fpb;
    const int parts = h->parts < frames ? h->parts : frames;
    const size_t bps = (size_t)pebble_downconvert_input_bytes_per_sample(h->dc);
    // `tail` starts after everything already queued on the caller's stream (mode upload, the consumer of the last output)
    PB_CUDA(cudaEventRecord(h->ev_front, st), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamWaitEvent(h->tail, h->ev_front, 0), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamWaitEvent(h->tail2, h->ev_front, 0), PEBBLE_ERR_CUDA);
    ChzTrace tr;
    tr.on = getenv("PEBBLE_CHZ_TRACE") != nullptr;
    int f0 = 0;
    for (int k = 0; k < parts; ++k) {
        const int f1 = (int)((long long)frames * (k + 1) / parts);
        const int pd = (f1 - f0) * h->fpb;             // decimated samples of this part
        const long long pos = (long long)f0 * h->fpb;  // its column in the work rows and in the output rows
        const long long pn = (long long)pd << h->S;
        tr.begin("front", k, st);
        r = pebble_downconvert_process_channels_device(h->dc, (int)pn, (const char *)d_in + bps * (size_t)(pos << h->S),
                                                       (double *)(h->d_dec + pos), row, st);   // receiver.cpp:908
        tr.end(st);
        if (r < 0) return r;
        tr.begin("fastfir", k, st);
        r = pebble_fastfir_process_batch_device(h->bp, pd, (const double *)(h->d_dec + pos), row, (double *)(h->d_bp + pos), row, st);   // :950
        tr.end(st);
        if (r < 0) return r;
        if (r != pd) { pebble::set_error("channelizer: FastFIR returned %d of %d", r, pd); return PEBBLE_ERR_STATE; }
        PB_CUDA(cudaEventRecord(h->ev_front, st), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaStreamWaitEvent(h->tail, h->ev_front, 0), PEBBLE_ERR_CUDA);
        tr.begin("demod", k, h->tail);
        chan_demod_kernel<<<(h->slots + PB_CHZ_C - 1) / PB_CHZ_C, 32, 0, h->tail>>>(h->d_bp + pos, h->d_tmp + pos, pd, row, h->nch, h->d_modes, h->d_order,
                                                            h->slots, h->pll, (double)0.9999f, h->d_state, h->fpb, d_open);                         // :987
        // the walks of consecutive parts form the critical chain (32 warps, latency-bound); the post filter of part k runs
        // beside the walk of part k+1 on its own stream
        tr.end(h->tail);
        PB_CUDA(cudaEventRecord(h->ev_demod, h->tail), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaStreamWaitEvent(h->tail2, h->ev_demod, 0), PEBBLE_ERR_CUDA);
        tr.begin("cfir", k, h->tail2);
        chan_cfir_kernel<<<dim3((pd + PB_FIR_TILE - 1) / PB_FIR_TILE, h->nch), PB_FIR_THREADS, 0, h->tail2>>>(
            (const double *)h->d_hist[h->hcur], h->d_tmp + pos, row, h->d_bp + pos, pd, h->d_taps, h->ntaps_am, h->ntaps_fm, h->d_modes,
            (cpx *)d_out + pos, out_stride, (double *)h->d_hist[h->hcur ^ 1], d_open);
        tr.end(h->tail2);
        pebble::count_launch(2);
        PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
        h->hcur ^= 1;
        f0 = f1;
    }
    // the call is complete on the caller's stream when both tails are
    PB_CUDA(cudaEventRecord(h->ev_tail, h->tail), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamWaitEvent(st, h->ev_tail, 0), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaEventRecord(h->ev_cfir, h->tail2), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamWaitEvent(st, h->ev_cfir, 0), PEBBLE_ERR_CUDA);
    tr.report();
    return nd;
}

}  // extern "C"
