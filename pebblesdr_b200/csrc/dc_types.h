// dc_types.h — host+device description of one CDownConvert object (pebblelib/downconvert.cpp:64-535):
// the decimate-by-2 stage chain, the oscillator constants and the per-call arguments of the fused kernel
// (dc_fused.cuh). Plain structs, shared by the setup-time host code (host_design.cpp) and the kernel.
//
// Oscillator (finding F5): the reference's recursive quadrature oscillator (downconvert.cpp:281-287) is
//   Osc_i = r_i * exp(j*(phi0 + (i+1)*theta_eff)),  r_{i+1} = r_i*(1.95 - r_i^2),  r_0 = 1,
// theta_eff = atan2(m_OscSin, m_OscCos) of the ROUNDED doubles. theta_eff is kept as a 128-bit fixed-point
// turn fraction so that the phase of any sample of an arbitrarily long stream is exact to 2^-64 turns.
#pragma once
#include "common.cuh"

#define PB_DC_MAX_STAGES 16
#define PB_DC_PREFIX 64      // state slots per stage (max history = 50 samples)
#define PB_DC_MAX_WTAB 32    // oscillator step table of the mix-only kernel
#define PB_DC_MAX_PAIRS 8    // sample pairs per thread per chunk in the fused kernel
#define PB_DC_RTAB 2048      // amplitude-transient table length (r_i for the first samples of a stream)
#define PB_DC_THREADS 256    // CTA size of both kernels

enum { PB_DC_CIC3 = 0, PB_DC_HB11 = 1, PB_DC_HBGEN = 2 };

struct DcStage {
    int type;      // PB_DC_*
    int len;       // FIR length (3 for CIC3)
    int hist;      // carried input samples: 2 for CIC3, len-1 for half-bands (always even)
    int coef_off;  // offset into DcChain::coef (host-side copy of the taps; the kernel uses __constant__ tables)
};

struct DcChain {
    int nstages;
    int preroll;                 // total cascade memory in input samples, multiple of 2^nstages
    DcStage st[PB_DC_MAX_STAGES];
    double coef[PB_DC_MAX_STAGES * 52];
    // oscillator
    uint64_t turn_hi, turn_lo;   // theta_eff / 2pi as a 0.128 fixed-point fraction
    double r_inf, inv_r_inf;     // converged oscillator amplitude (~sqrt(0.95)) and its reciprocal
    int wtab_stride;             // sample stride of wtab (PB_DC_THREADS)
    int chunk;                   // chunk length wchunk was built for
    cpx wtab[PB_DC_MAX_WTAB];    // exp(j*k*wtab_stride*theta_eff)            (mix-only kernel)
    cpx wpair[PB_DC_MAX_PAIRS];  // exp(j*k*2*PB_DC_THREADS*theta_eff)        (fused kernel: pair k of a thread)
    cpx w1;                      // exp(j*theta_eff)
    cpx wchunk;                  // exp(j*chunk*theta_eff)
    double rtab[PB_DC_RTAB];     // r_i for stream sample i < PB_DC_RTAB (exact recurrence)
};

// Per-channel oscillator of the batched (multi-channel) form: all channels share the stage chain of one DcChain,
// each has its own NCO. phase = oscillator phase before sample 0 of the call.
struct DcChan {
    uint64_t turn_hi, turn_lo, phase_hi, phase_lo;
    cpx w1, wchunk;
};

// Per-call arguments of the fused kernel.
struct DcCall {
    const cpx *in;       // n input samples
    cpx *out;            // n >> S output samples
    const cpx *state_in; // carried history read by segment 0: stage s at s*PB_DC_PREFIX, hist_s samples, oldest first
    cpx *state_out;      // history stored by the last segment (ping-pong: differs from state_in)
    long long n;         // input samples in this call (multiple of 2^S)
    long long seg_len;   // samples per segment == per CTA (multiple of the chunk length, >= preroll)
    uint64_t phase_hi, phase_lo;  // oscillator phase before sample 0 of this call (0.128 turns)
    long long stream_pos;         // samples mixed since the oscillator was created (for rtab)
    int front;           // F: stages 0..F-1 run every chunk
    int group;           // G: stages F..S-1 run every G chunks on the accumulated samples
    // batched form (dc_cic_kernel only): blockIdx.x = segment * nch + channel, so that the CTAs resident at the
    // same time read the SAME input segments (served by L2) for different channels
    const DcChan *chan;       // nullptr: single channel, oscillator from DcChain + phase_hi/lo above
    int nch;                  // channels (>= 1)
    int state_stride;         // cpx between two channels' carried state
    long long out_stride;     // cpx between two channels' outputs
    // single-channel oscillator constants as kernel parameters (constant bank, no registers held across the chunk
    // loop): used by dc_pipe_kernel; copies of DcChain::w1 / wchunk / r_inf / turn
    cpx w1, wchunk;
    double r_inf;
    uint64_t turn_hi, turn_lo;
};

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

// ---- shared-memory plan of the fused kernel (cpx units), computed identically on host and device -----------
// Stage s reads its input from two arrays, EVEN and ODD input samples (history first), stored with one pad slot
// per four (slot(i) = i + i/4) so that a thread computing four consecutive outputs reads conflict-free.
PB_HD int dc_slot(int i) { return i + (i >> 2); }

PB_HD int dc_stage_cap(int chunk, int front, int group, int S, int s) {
    const int F = front < S ? front : S;
    const int per_chunk = chunk >> s;
    return s < F ? per_chunk : per_chunk * group;   // new samples buffered between two runs of stage s
}
// slots of ONE of the two arrays of stage s (== 4 mod 8, so EVEN and ODD start 64 B apart modulo 128 B)
PB_HD int dc_stage_arr(const DcStage &st, int cap) {
    int n = dc_slot(st.hist / 2 + cap / 2 + 8) + 1;
    return ((n + 7) & ~7) + 4;
}
PB_HD int dc_smem_cpx(const DcChain &ch, int chunk, int front, int group) {
    int tot = 0;
    for (int s = 0; s < ch.nstages; ++s) tot += 2 * dc_stage_arr(ch.st[s], dc_stage_cap(chunk, front, group, ch.nstages, s));
    return tot;
}
