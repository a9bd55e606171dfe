// dc_core.cuh — fused NCO mix + decimate-by-2 cascade (the CDownConvert::ProcessData hot loops,
// pebblelib/downconvert.cpp:274-328) as one streaming pass over the input.
//
// The body is written against a small execution-context policy (Ctx) so the identical source runs
//   * on the GPU  : Ctx = CTA (threadIdx, __syncthreads, cp.async.bulk + mbarrier chunk loads), and
//   * on the host : Ctx = pthread team with a barrier (tests/host_emul/emul_dc.cpp) for CPU-only CI.
//
// Work decomposition: the call's n input samples are cut into SEGMENTS (one per CTA). A CTA walks
// its segment in CHUNKS of <= C samples: chunk -> mix in place -> stage 0 .. F-1 ("front" stages, run
// every chunk) accumulating into stage F's buffer; every G chunks the "back" stages F..S-1 run on the
// accumulated samples and the last stage writes the decimated output. Per-stage FIR history lives
// in shared memory in front of each stage buffer, exactly like the delay lines the reference
// carries between calls (downconvert.cpp:366-390, 489-493, 528-531). Segment 0 loads that history from
// the handle's carried state; every other segment rebuilds it by PRE-ROLLING over the `preroll`
// input samples before its start (total cascade memory, rounded up to 2^S) and discarding the
// outputs; the last segment stores the history back. No inter-CTA communication.
//
// Oscillator (finding F5): the reference's recursive quadrature oscillator is
//   Osc_i = r_i * exp(j*(phi0 + (i+1)*theta_eff)),  r_{i+1} = r_i*(1.95 - r_i^2),
// theta_eff = atan2(m_OscSin, m_OscCos) of the ROUNDED doubles. The host passes theta_eff as a
// 128-bit fixed-point turn fraction; a thread seeds with one exact sincospi per chunk and steps with a
// small table of exp(j*k*T*theta) (T = threads), so every sample costs one complex multiply.
#pragma once
#include "common.cuh"

#define PB_DC_MAX_STAGES 16
#define PB_DC_PREFIX 64      // history slots in front of every stage buffer (max hist = 50)
#define PB_DC_MAX_WTAB 32    // chunk/threads steps of the oscillator table
#define PB_DC_RTAB 2048      // amplitude-transient table length (r_i for the first samples)

enum { PB_DC_CIC3 = 0, PB_DC_HB11 = 1, PB_DC_HBGEN = 2 };

struct DcStage {
    int type;      // PB_DC_*
    int len;       // FIR length (3 for CIC3)
    int hist;      // carried input samples: 2 for CIC3, len-1 for half-bands
    int coef_off;  // offset into DcChain::coef
};

// Device-resident description of one CDownConvert object (chain + oscillator constants).
struct DcChain {
    int nstages;
    int preroll;                 // total cascade memory in input samples, multiple of 2^nstages
    DcStage st[PB_DC_MAX_STAGES];
    double coef[PB_DC_MAX_STAGES * 52];
    // oscillator
    uint64_t turn_hi, turn_lo;   // theta_eff / 2pi as a 0.128 fixed-point fraction
    double r_inf;                // converged oscillator amplitude (~sqrt(0.95))
    int wtab_stride;             // thread count the table below was built for
    cpx wtab[PB_DC_MAX_WTAB];    // exp(j*k*wtab_stride*theta_eff), k = 0..
    double rtab[PB_DC_RTAB];     // r_i for stream sample i < PB_DC_RTAB (exact recurrence)
};

// Per-call arguments.
struct DcCall {
    const cpx *in;       // n input samples
    cpx *out;            // n >> S output samples
    const cpx *state_in; // carried history read by segment 0: stage s at + s*PB_DC_PREFIX, hist_s used
    cpx *state_out;      // history stored by the last segment (ping-pong: must differ from state_in)
    long long n;         // input samples in this call (multiple of 2^S)
    long long seg_len;   // samples per segment (multiple of 2^S, >= preroll)
    uint64_t phase_hi, phase_lo;  // oscillator phase before sample 0 of this call (0.128 turns)
    long long stream_pos;         // samples mixed since the oscillator was created (for rtab)
    int chunk;           // C, multiple of 2^S and of nthreads
    int front;           // F
    int group;           // G
};

// shared-memory carve-up, in cpx units
struct DcSmem {
    int in_off[2];                       // two input chunk buffers [PREFIX + C]
    int st_off[PB_DC_MAX_STAGES + 1];    // stage s input buffer [PREFIX + cap_s], s >= 1
    int total;
};

PB_HD int dc_stage_cap(const DcCall &c, int S, int s) {
    // new samples stage s can receive between two runs of its consumer
    int F = c.front < S ? c.front : S;
    long long per_chunk = c.chunk >> s;
    return (int)(s < F ? per_chunk : per_chunk * c.group);
}

PB_HD DcSmem dc_smem_layout(const DcCall &c, int S) {
    DcSmem L;
    int off = 0;
    L.in_off[0] = off; off += PB_DC_PREFIX + c.chunk;
    L.in_off[1] = off; off += PB_DC_PREFIX + c.chunk;
    L.st_off[0] = -1;
    for (int s = 1; s < S; ++s) {
        L.st_off[s] = off;
        off += PB_DC_PREFIX + dc_stage_cap(c, S, s);
    }
    L.total = off;
    return L;
}

// 128-bit phase: (hi,lo) + k*(thi,tlo), returns the top 64 bits.
PB_HD uint64_t dc_phase_at(uint64_t phi, uint64_t plo, uint64_t thi, uint64_t tlo, uint64_t k) {
#if defined(__CUDA_ARCH__)
    uint64_t lo_prod = tlo * k;
    uint64_t carry_mul = __umul64hi(tlo, k);
#else
    unsigned __int128 pr = (unsigned __int128)tlo * k;
    uint64_t lo_prod = (uint64_t)pr;
    uint64_t carry_mul = (uint64_t)(pr >> 64);
#endif
    uint64_t lo = plo + lo_prod;
    uint64_t c = lo < plo ? 1u : 0u;
    return phi + thi * k + carry_mul + c;
}

// exp(j*2pi*p/2^64)
PB_HD cpx dc_phase_to_osc(uint64_t p) {
    double x = (double)(int64_t)p * (1.0 / 9223372036854775808.0);  // in [-1,1): angle / pi
    double s, c;
#if defined(__CUDA_ARCH__)
    sincospi(x, &s, &c);
#else
    s = sin(x * 3.14159265358979323846264338328);
    c = cos(x * 3.14159265358979323846264338328);
#endif
    return cpx_make(c, s);
}

// One decimate-by-2 output. w[k] = stage input sample (2j - hist + k).
PB_HD cpx dc_stage_eval(const DcStage &sd, const double *coef, const cpx *w) {
    if (sd.type == PB_DC_CIC3) {
        // downconvert.cpp:526-527: .125*( odd + Xeven_prev + 3.0*(Xodd_prev + even) )
        return cpx_make(.125 * (w[3].x + w[0].x + 3.0 * (w[1].x + w[2].x)),
                        .125 * (w[3].y + w[0].y + 3.0 * (w[1].y + w[2].y)));
    }
    const double *h = coef + sd.coef_off;
    if (sd.type == PB_DC_HB11) {
        // downconvert.cpp:478-483: H0,H2,H4,H5,H6,H8,H10 in that order
        double re = h[0] * w[0].x + h[2] * w[2].x + h[4] * w[4].x + h[5] * w[5].x + h[6] * w[6].x +
                    h[8] * w[8].x + h[10] * w[10].x;
        double im = h[0] * w[0].y + h[2] * w[2].y + h[4] * w[4].y + h[5] * w[5].y + h[6] * w[6].y +
                    h[8] * w[8].y + h[10] * w[10].y;
        return cpx_make(re, im);
    }
    // generic half-band, downconvert.cpp:372-384: the accumulator is seeded with tap 0 and the loop
    // starts at j=0 again (tap 0 counted twice, finding F7), then the centre tap.
    double re = w[0].x * h[0], im = w[0].y * h[0];
    for (int j = 0; j < sd.len; j += 2) {
        re += w[j].x * h[j];
        im += w[j].y * h[j];
    }
    const int m = (sd.len - 1) / 2;
    re += w[m].x * h[m];
    im += w[m].y * h[m];
    return cpx_make(re, im);
}

// ---- phases (each is followed by ctx.sync() in dc_run_segment) -------------------------------------

// In-place history carry of a single-buffered stage: the hist samples that follow the m samples
// just consumed move to the front. b points at slot 0 of the history (= buffer + PREFIX - hist).
template <class Ctx> PB_HD void dc_carry(Ctx &ctx, cpx *b, int hist, int m) {
    if (m <= 0) return;
    if (m >= hist) {
        for (int i = ctx.tid(); i < hist; i += ctx.nthreads()) b[i] = b[i + m];
    } else if (ctx.tid() == 0) {
        for (int i = 0; i < hist; ++i) b[i] = b[i + m];  // overlapping: ascending order is safe
    }
}

// Mix `len` raw samples in place: x[i] *= r * exp(j*phase(a0+i)).
template <class Ctx>
PB_HD void dc_mix(Ctx &ctx, const DcChain *ch, const DcCall &call, cpx *x, long long a0, int len) {
    const int T = ctx.nthreads(), tid = ctx.tid();
    if (tid >= len) return;
    const uint64_t p = dc_phase_at(call.phase_hi, call.phase_lo, ch->turn_hi, ch->turn_lo,
                                   (uint64_t)(a0 + tid + 1));
    const cpx seed = dc_phase_to_osc(p);
    const long long g0 = call.stream_pos + a0 + tid;
    const bool steady = g0 >= PB_DC_RTAB;
    const double rinf = ch->r_inf;
    int k = 0;
    for (int i = tid; i < len; i += T, ++k) {
        cpx o = cpx_mul(seed, ch->wtab[k]);
        const long long g = g0 + (long long)k * T;
        const double r = (steady || g >= PB_DC_RTAB) ? rinf : ch->rtab[g];
        o = cpx_scale(o, r);
        x[i] = cpx_mul(x[i], o);
    }
}

// Run one decimate-by-2 stage over nout outputs. src = history slot 0 of the stage's input buffer.
// Output goes to dst[j] (shared) or, when gout != nullptr, to gout[gidx0 + j] for indices >= first_out.
template <class Ctx>
PB_HD void dc_consume(Ctx &ctx, const DcStage &sd, const double *coef, const cpx *src, int nout, cpx *dst,
                      cpx *gout, long long gidx0, long long first_out) {
    for (int j = ctx.tid(); j < nout; j += ctx.nthreads()) {
        cpx y = dc_stage_eval(sd, coef, src + 2 * j);
        if (gout) {
            if (gidx0 + j >= first_out) gout[gidx0 + j] = y;
        } else {
            dst[j] = y;
        }
    }
}

// ---- one segment == one CTA ------------------------------------------------------------------------
// sh_chain: a shared-memory copy of the stage table + coefficients (ctx-visible), ch: global struct.
template <class Ctx>
PB_HD void dc_run_segment(Ctx &ctx, const DcChain *ch, const DcStage *st, const double *coef,
                          const DcCall &call, cpx *smem, long long seg) {
    const int S = ch->nstages;
    const int T = ctx.nthreads(), tid = ctx.tid();
    const int F = call.front < S ? call.front : S;
    const int C = call.chunk;
    const DcSmem L = dc_smem_layout(call, S);
    const long long seg_begin = seg * call.seg_len;
    long long seg_end = seg_begin + call.seg_len;
    if (seg_end > call.n) seg_end = call.n;
    const long long start = seg == 0 ? 0 : seg_begin - ch->preroll;
    const long long first_out = seg_begin >> S;
    const bool last_seg = seg_end == call.n;
    const int h0 = st[0].hist;

    // ---- initial history: carried state for segment 0, zeros (then pre-roll) otherwise
    {
        cpx *b0 = smem + L.in_off[0] + PB_DC_PREFIX - h0;
        for (int i = tid; i < h0; i += T) b0[i] = seg == 0 ? call.state_in[i] : cpx_make(0.0, 0.0);
        for (int s = 1; s < S; ++s) {
            cpx *b = smem + L.st_off[s] + PB_DC_PREFIX - st[s].hist;
            for (int i = tid; i < st[s].hist; i += T)
                b[i] = seg == 0 ? call.state_in[s * PB_DC_PREFIX + i] : cpx_make(0.0, 0.0);
        }
    }
    const long long total = seg_end - start;
    const long long nchunks = (total + C - 1) / C;
    // prologue: chunks 0 and 1 in flight
    for (int c = 0; c < 2 && c < nchunks; ++c) {
        long long pos = start + (long long)c * C;
        int len = (int)((seg_end - pos) < C ? (seg_end - pos) : C);
        ctx.load_chunk(c, smem + L.in_off[c] + PB_DC_PREFIX, call.in + pos, len);
    }
    ctx.sync();

    int acc = 0;             // stage-F input samples accumulated since the last back run
    int ingroup = 0;         // chunks since the last back run
    long long group_pos = start;  // input position where the current group started
    int prev_len = 0;        // chunk length whose front-stage carries are still pending
    int prev_back = 0;       // stage-F sample count whose back-stage carries are still pending

    for (long long ci = 0; ci < nchunks; ++ci) {
        const int buf = (int)(ci & 1);
        const long long pos = start + ci * C;
        const int len = (int)((seg_end - pos) < C ? (seg_end - pos) : C);
        cpx *inb = smem + L.in_off[buf];
        ctx.wait_chunk(buf);

        // phase MIX (+ pending in-place carries of the single-buffered stages)
        dc_mix(ctx, ch, call, inb + PB_DC_PREFIX, pos, len);
        for (int s = 1; s < S; ++s) {
            int m = s < F ? (prev_len >> s) : (prev_back >> (s - F));
            dc_carry(ctx, smem + L.st_off[s] + PB_DC_PREFIX - st[s].hist, st[s].hist, m);
        }
        prev_back = 0;
        ctx.sync();

        // phase STAGE 0 (+ its history carry into the other input buffer's prefix)
        {
            const bool to_g = (S == 1);
            cpx *dst = to_g ? nullptr : smem + L.st_off[1] + PB_DC_PREFIX + (F == 1 ? acc : 0);
            dc_consume(ctx, st[0], coef, inb + PB_DC_PREFIX - h0, len >> 1, dst, to_g ? call.out : nullptr,
                       pos >> 1, first_out);
            cpx *nb = smem + L.in_off[buf ^ 1] + PB_DC_PREFIX - h0;
            const cpx *cb = inb + PB_DC_PREFIX - h0 + len;
            for (int i = tid; i < h0; i += T) nb[i] = cb[i];
        }
        ctx.sync();
        // the input buffer is free again: prefetch chunk ci+2 into it
        if (ci + 2 < nchunks) {
            long long p2 = start + (ci + 2) * C;
            int l2 = (int)((seg_end - p2) < C ? (seg_end - p2) : C);
            ctx.load_chunk(buf, inb + PB_DC_PREFIX, call.in + p2, l2);
        }
        // phases STAGE 1 .. F-1
        for (int s = 1; s < F; ++s) {
            const bool to_g = (s == S - 1);
            cpx *dst = to_g ? nullptr : smem + L.st_off[s + 1] + PB_DC_PREFIX + (s + 1 == F ? acc : 0);
            dc_consume(ctx, st[s], coef, smem + L.st_off[s] + PB_DC_PREFIX - st[s].hist, len >> (s + 1), dst,
                       to_g ? call.out : nullptr, pos >> (s + 1), first_out);
            ctx.sync();
        }
        prev_len = len;
        if (F < S) {
            acc += len >> F;
            ++ingroup;
            if (ingroup == call.group || ci == nchunks - 1) {
                // back stages F .. S-1 on the accumulated samples
                for (int s = F; s < S; ++s) {
                    const int nin = acc >> (s - F);
                    const bool to_g = (s == S - 1);
                    cpx *dst = to_g ? nullptr : smem + L.st_off[s + 1] + PB_DC_PREFIX;
                    dc_consume(ctx, st[s], coef, smem + L.st_off[s] + PB_DC_PREFIX - st[s].hist, nin >> 1, dst,
                               to_g ? call.out : nullptr, group_pos >> (s + 1), first_out);
                    ctx.sync();
                }
                prev_back = acc;
                acc = 0;
                ingroup = 0;
                group_pos = pos + len;
            }
        }
    }

    if (last_seg) {
        // final carries, then store every stage's history as the carried state
        for (int s = 1; s < S; ++s) {
            int m = s < F ? (prev_len >> s) : (prev_back >> (s - F));
            dc_carry(ctx, smem + L.st_off[s] + PB_DC_PREFIX - st[s].hist, st[s].hist, m);
        }
        ctx.sync();
        const int nb = (int)(nchunks & 1);  // stage 0 history sits in the prefix of the next buffer
        const cpx *b0 = smem + L.in_off[nb] + PB_DC_PREFIX - h0;
        for (int i = tid; i < h0; i += T) call.state_out[i] = b0[i];
        for (int s = 1; s < S; ++s) {
            const cpx *b = smem + L.st_off[s] + PB_DC_PREFIX - st[s].hist;
            for (int i = tid; i < st[s].hist; i += T) call.state_out[s * PB_DC_PREFIX + i] = b[i];
        }
    }
}
