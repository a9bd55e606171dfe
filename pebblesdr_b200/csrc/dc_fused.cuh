// dc_fused.cuh — fused NCO mix + decimate-by-2 cascade: the two hot loops of CDownConvert::ProcessData
// (pebblelib/downconvert.cpp:274-311 oscillator+mix, :323-328 stage loop) as ONE streaming pass over the input.
//
// Decomposition. The call's n input samples are cut into SEGMENTS, one per CTA (a single wave over the SMs).
// A CTA walks its segment in CHUNKS of C samples:
//   P1   every thread owns K = C/(2T) sample PAIRS (2p, 2p+1), p = tid + k*T. The pair is one coalesced 256-bit
//        global load (a warp reads 1 KB contiguous), issued one chunk AHEAD into registers so that HBM latency is
//        covered by a whole chunk of arithmetic; the stream is touched exactly once and never staged through
//        shared memory (it is consumed by exactly one thread; staging would only add shared-memory traffic, which
//        for 16-byte FP64 samples is the second-scarcest resource after HBM). Mix = two complex multiplies per
//        sample: Osc = seed * wpair[k] (* w1 for the odd sample), x *= Osc.
//   S_s  stage s (CIC3 / half-band of compile-time length) reads its input from shared memory, split into EVEN and
//        ODD sample arrays (the polyphase form of a decimate-by-2 FIR: all even taps hit one array, the centre
//        tap the other), each thread producing FOUR consecutive outputs from a sliding register window
//        (3.25 loads per output for the 11-tap stage instead of 7); taps come from the constant bank.
//   Stages 0..F-1 run every chunk; stages F..S-1 run every G chunks on the accumulated samples.
// Per-stage FIR history sits in front of each stage array exactly like the delay lines the reference carries
// between calls (downconvert.cpp:366-390, 489-493, 528-531). Segment 0 loads it from the handle's carried state;
// every other segment rebuilds it by PRE-ROLLING over the `preroll` input samples before its start (total cascade
// memory rounded up to 2^S) and discarding those outputs; the last segment stores the history back.
// There is no inter-CTA communication.
//
// Arithmetic order differs from the reference's scalar loops (sliding-window accumulation, closed-form
// oscillator), so parity is relative-RMS (<= 1e-10, tests/), not bit-exact; sample counts and block boundaries
// are exact.
#pragma once
#include "dc_types.h"

#if defined(__CUDACC__)

__constant__ double c_hb[341];   // pebble_hb::kHalfBandTaps (filtercoef.h:36-426), uploaded once per device

__host__ __device__ constexpr int hb_off(int L) {
    // pebble_hb::kHalfBandOff[(L-11)/4]
    return L == 11 ? 0 : L == 15 ? 11 : L == 19 ? 26 : L == 23 ? 45 : L == 27 ? 68 : L == 31 ? 95 : L == 35 ? 126
         : L == 39 ? 161 : L == 43 ? 200 : L == 47 ? 243 : 290;
}

struct DcBuf {            // one stage's input arrays; pointers address logical index 0 (oldest history sample)
    cpx *ev, *od;
    int half_hist;        // hist/2
};

// Where a stage writes its outputs: the next stage's arrays at logical new-sample offset `acc`, or global memory.
struct DcSink {
    cpx *ev, *od;         // next stage arrays (nullptr => global)
    int base;             // next stage: half_hist*2 + acc (sample index of output 0 in history-first numbering)
    cpx *gout;            // global output
    long long gidx0, first_out;
};

__device__ __forceinline__ void dc_emit(const DcSink &k, int j, cpx y) {
    if (k.gout) {
        if (k.gidx0 + j >= k.first_out) k.gout[k.gidx0 + j] = y;
    } else {
        const int q = k.base + j;
        cpx *a = (q & 1) ? k.od : k.ev;
        a[dc_slot(q >> 1)] = y;
    }
}

// Half-band of length L, decimate by 2 (downconvert.cpp:341-393 generic, :401-496 11-tap), outputs 4u..4u+3.
// With x' = history-first sample numbering, y[j] = sum_m h[2m] * ev[j+m] + h[c] * od[j + (c-1)/2], c = (L-1)/2.
// The generic stage counts tap 0 twice (finding F7, downconvert.cpp:372-379); the 11-tap stage does not.
template <int L>
__device__ __forceinline__ void dc_tile_hb(const DcBuf &b, int u, int nout, const DcSink &snk) {
    constexpr int E = (L + 1) / 2, CO = (L - 1) / 2, OC = (CO - 1) / 2, OFF = hb_off(L);
    const cpx *e = b.ev + 5 * u, *o = b.od + 5 * u;
    double ar[4] = {0, 0, 0, 0}, ai[4] = {0, 0, 0, 0};
#pragma unroll
    for (int k = 0; k < E + 3; ++k) {
        const cpx v = e[k + (k >> 2)];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int m = k - r;
            if (m >= 0 && m < E) {
                const double h = (m == 0 && L != 11) ? 2.0 * c_hb[OFF] : c_hb[OFF + 2 * m];
                ar[r] = fma(h, v.x, ar[r]);
                ai[r] = fma(h, v.y, ai[r]);
            }
        }
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const cpx v = o[(OC + r) + ((OC + r) >> 2)];
        ar[r] = fma(c_hb[OFF + CO], v.x, ar[r]);
        ai[r] = fma(c_hb[OFF + CO], v.y, ai[r]);
    }
#pragma unroll
    for (int r = 0; r < 4; ++r)
        if (4 * u + r < nout) dc_emit(snk, 4 * u + r, cpx_make(ar[r], ai[r]));
}

// CIC3 (downconvert.cpp:519-535): y[j] = .125*( x[2j+1] + x[2j-2] + 3*(x[2j-1] + x[2j]) ); history = 2 samples,
// so ev[j] = x[2j-2], od[j] = x[2j-1], ev[j+1] = x[2j], od[j+1] = x[2j+1].
__device__ __forceinline__ void dc_tile_cic3(const DcBuf &b, int u, int nout, const DcSink &snk) {
    const cpx *e = b.ev + 5 * u, *o = b.od + 5 * u;
    cpx ve[5], vo[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        ve[k] = e[k + (k >> 2)];
        vo[k] = o[k + (k >> 2)];
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const double yr = .125 * (vo[r + 1].x + ve[r].x + 3.0 * (vo[r].x + ve[r + 1].x));
        const double yi = .125 * (vo[r + 1].y + ve[r].y + 3.0 * (vo[r].y + ve[r + 1].y));
        if (4 * u + r < nout) dc_emit(snk, 4 * u + r, cpx_make(yr, yi));
    }
}

// Run stage `sd` over nout outputs with the whole CTA.
__device__ __forceinline__ void dc_run_stage(const DcStage &sd, const DcBuf &b, int nout, const DcSink &snk) {
    const int tiles = (nout + 3) >> 2;
    for (int u = threadIdx.x; u < tiles; u += PB_DC_THREADS) {
        switch (sd.len) {
            case 3: dc_tile_cic3(b, u, nout, snk); break;
            case 11: dc_tile_hb<11>(b, u, nout, snk); break;
            case 15: dc_tile_hb<15>(b, u, nout, snk); break;
            case 19: dc_tile_hb<19>(b, u, nout, snk); break;
            case 23: dc_tile_hb<23>(b, u, nout, snk); break;
            case 27: dc_tile_hb<27>(b, u, nout, snk); break;
            case 31: dc_tile_hb<31>(b, u, nout, snk); break;
            case 35: dc_tile_hb<35>(b, u, nout, snk); break;
            case 39: dc_tile_hb<39>(b, u, nout, snk); break;
            case 43: dc_tile_hb<43>(b, u, nout, snk); break;
            case 47: dc_tile_hb<47>(b, u, nout, snk); break;
            default: dc_tile_hb<51>(b, u, nout, snk); break;
        }
    }
}

// History carry after a run consumed n (even) input samples: the hist samples that follow move to the front.
// Done by the LAST threads of the CTA (the stage tiles occupy the first ones).
__device__ __forceinline__ void dc_carry(const DcBuf &b, int n) {
    const int hh = b.half_hist, m = n >> 1;
    if (m <= 0) return;
    const int i = PB_DC_THREADS - 1 - (int)threadIdx.x;
    if (m >= hh) {
        if (i < 2 * hh) {
            cpx *a = (i & 1) ? b.od : b.ev;
            const int idx = i >> 1;
            a[dc_slot(idx)] = a[dc_slot(idx + m)];
        }
    } else if (i < 2) {
        cpx *a = i ? b.od : b.ev;
        for (int idx = 0; idx < hh; ++idx) a[dc_slot(idx)] = a[dc_slot(idx + m)];  // overlapping: ascending is safe
    }
}

__device__ __forceinline__ void dc_ld256(const cpx *p, cpx &a, cpx &b) {
    asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];"
                 : "=d"(a.x), "=d"(a.y), "=d"(b.x), "=d"(b.y)
                 : "l"(p));
}

// Generic fallback (any chain, any F): stage descriptors and array pointers live in shared memory and every
// stage goes through the runtime switch. Correct for every chain the planner can produce, but its bookkeeping
// costs several times the arithmetic; dc_fast_kernel below is the path the headline configurations take.
template <int C>
__global__ void __launch_bounds__(PB_DC_THREADS) dc_fused_kernel(const DcChain *__restrict__ ch, const DcCall call) {
    constexpr int T = PB_DC_THREADS, K = C / (2 * T);
    static_assert(K >= 1 && K <= PB_DC_MAX_PAIRS, "chunk must be 2*T .. 16*T samples");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ DcStage st[PB_DC_MAX_STAGES];
    __shared__ DcBuf buf[PB_DC_MAX_STAGES];
    cpx *smem = reinterpret_cast<cpx *>(smem_raw);
    const int tid = threadIdx.x;
    const int S = ch->nstages;
    const int F = call.front < S ? call.front : S;
    const int G = call.group;
    if (tid < S) st[tid] = ch->st[tid];
    __syncthreads();
    if (tid == 0) {
        int off = 0;
        for (int s = 0; s < S; ++s) {
            const int arr = dc_stage_arr(st[s], dc_stage_cap(C, call.front, G, S, s));
            buf[s].ev = smem + off;
            buf[s].od = smem + off + arr;
            buf[s].half_hist = st[s].hist >> 1;
            off += 2 * arr;
        }
    }
    __syncthreads();

    const long long seg = blockIdx.x;
    const long long seg_begin = seg * call.seg_len;
    long long seg_end = seg_begin + call.seg_len;
    if (seg_end > call.n) seg_end = call.n;
    const long long start = seg == 0 ? 0 : seg_begin - ch->preroll;
    const long long first_out = seg_begin >> S;
    const bool last_seg = seg_end == call.n;

    // ---- initial history: carried state for segment 0, zeros (then pre-roll) otherwise
    for (int s = 0; s < S; ++s)
        for (int i = tid; i < st[s].hist; i += T) {
            cpx *a = (i & 1) ? buf[s].od : buf[s].ev;
            a[dc_slot(i >> 1)] = seg == 0 ? call.state_in[s * PB_DC_PREFIX + i] : cpx_make(0.0, 0.0);
        }

    const long long total = seg_end - start;
    const long long nchunks = (total + C - 1) / C;
    const cpx w1 = ch->w1, wchunk = ch->wchunk;
    const double rinf = ch->r_inf;
    cpx wp[K];
#pragma unroll
    for (int k = 0; k < K; ++k) wp[k] = ch->wpair[k];

    cpx nx0[K], nx1[K];   // prefetched pairs of the next chunk
    {
        const int len0 = (int)(total < C ? total : C);
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int p = tid + k * T;
            nx0[k] = nx1[k] = cpx_make(0.0, 0.0);
            if (2 * p < len0) dc_ld256(call.in + start + 2 * p, nx0[k], nx1[k]);
        }
    }
    __syncthreads();

    cpx seed = cpx_make(0.0, 0.0);   // r_inf * exp(j*phase(pos + 2*tid + 1)) of the current chunk
    int accF = 0, ingroup = 0;
    long long group_pos = start;
    int pend_s = -1, pend_n = 0;     // history carry still owed (the last stage that ran)

    for (long long ci = 0; ci < nchunks; ++ci) {
        const long long pos = start + ci * C;
        const int len = (int)((seg_end - pos) < C ? (seg_end - pos) : C);
        // ---------------- P1: oscillator + mix into stage 0's arrays -------------------------------------------
        cpx c0[K], c1[K];
#pragma unroll
        for (int k = 0; k < K; ++k) { c0[k] = nx0[k]; c1[k] = nx1[k]; }
        if (ci + 1 < nchunks) {
            const long long p2 = pos + C;
            const int l2 = (int)((seg_end - p2) < C ? (seg_end - p2) : C);
#pragma unroll
            for (int k = 0; k < K; ++k) {
                const int p = tid + k * T;
                if (2 * p < l2) dc_ld256(call.in + p2 + 2 * p, nx0[k], nx1[k]);
            }
        }
        if (pend_s == 0) {            // single-stage chain: stage 0's carry must precede the new mix
            dc_carry(buf[0], pend_n);
            pend_s = -1;
            __syncthreads();
        }
        if ((ci & 31) == 0) {
            seed = cpx_scale(dc_phase_to_osc(dc_phase_at(call.phase_hi, call.phase_lo, ch->turn_hi, ch->turn_lo,
                                                         (uint64_t)(pos + 2 * tid + 1))), rinf);
        } else {
            seed = cpx_mul(seed, wchunk);
        }
        const bool transient = call.stream_pos + pos < PB_DC_RTAB;
        {
            cpx *ev0 = buf[0].ev, *od0 = buf[0].od;
            const int hh0 = buf[0].half_hist;
#pragma unroll
            for (int k = 0; k < K; ++k) {
                const int p = tid + k * T;
                if (2 * p < len) {
                    cpx o0 = cpx_mul(seed, wp[k]);
                    cpx o1 = cpx_mul(o0, w1);
                    if (transient) {   // amplitude still converging (first PB_DC_RTAB samples of a stream)
                        const long long g = call.stream_pos + pos + 2 * p;
                        const double q0 = (g < PB_DC_RTAB ? ch->rtab[g] : rinf) * ch->inv_r_inf;
                        const double q1 = (g + 1 < PB_DC_RTAB ? ch->rtab[g + 1] : rinf) * ch->inv_r_inf;
                        o0 = cpx_scale(o0, q0);
                        o1 = cpx_scale(o1, q1);
                    }
                    const int sl = dc_slot(hh0 + p);
                    ev0[sl] = cpx_mul(c0[k], o0);
                    od0[sl] = cpx_mul(c1[k], o1);
                }
            }
        }
        if (pend_s > 0) { dc_carry(buf[pend_s], pend_n); pend_s = -1; }
        __syncthreads();

        // ---------------- front stages 0..F-1 on this chunk ----------------------------------------------------
        for (int s = 0; s < F; ++s) {
            const int nin = len >> s;
            DcSink snk;
            if (s == S - 1) {
                snk.ev = snk.od = nullptr; snk.base = 0;
                snk.gout = call.out; snk.gidx0 = pos >> (s + 1); snk.first_out = first_out;
            } else {
                snk.ev = buf[s + 1].ev; snk.od = buf[s + 1].od;
                snk.base = 2 * buf[s + 1].half_hist + (s + 1 == F ? accF : 0);
                snk.gout = nullptr; snk.gidx0 = 0; snk.first_out = 0;
            }
            dc_run_stage(st[s], buf[s], nin >> 1, snk);
            if (s > 0) dc_carry(buf[s - 1], len >> (s - 1));
            __syncthreads();
        }
        pend_s = F - 1;
        pend_n = len >> (F - 1);
        // ---------------- back stages F..S-1 every G chunks ----------------------------------------------------
        if (F < S) {
            accF += len >> F;
            ++ingroup;
            if (ingroup == G || ci == nchunks - 1) {
                for (int s = F; s < S; ++s) {
                    const int nin = accF >> (s - F);
                    DcSink snk;
                    if (s == S - 1) {
                        snk.ev = snk.od = nullptr; snk.base = 0;
                        snk.gout = call.out; snk.gidx0 = group_pos >> (s + 1); snk.first_out = first_out;
                    } else {
                        snk.ev = buf[s + 1].ev; snk.od = buf[s + 1].od;
                        snk.base = 2 * buf[s + 1].half_hist;
                        snk.gout = nullptr; snk.gidx0 = 0; snk.first_out = 0;
                    }
                    dc_run_stage(st[s], buf[s], nin >> 1, snk);
                    dc_carry(buf[s - 1], s == F ? (len >> (F - 1)) : (accF >> (s - 1 - F)));
                    __syncthreads();
                }
                pend_s = S - 1;
                pend_n = accF >> (S - 1 - F);
                accF = 0;
                ingroup = 0;
                group_pos = pos + len;
            }
        }
    }

    if (last_seg) {
        if (pend_s >= 0) dc_carry(buf[pend_s], pend_n);
        __syncthreads();
        for (int s = 0; s < S; ++s)
            for (int i = tid; i < st[s].hist; i += T) {
                const cpx *a = (i & 1) ? buf[s].od : buf[s].ev;
                call.state_out[s * PB_DC_PREFIX + i] = a[dc_slot(i >> 1)];
            }
    }
}

// ================================================================================================================
// dc_fast_kernel: the same algorithm with the three FRONT stages fixed at compile time (lengths L0,L1,L2, F = 3).
// Array offsets are immediates, every stage phase is entered only by the warps that own a tile, and the back
// stages (3..S-1, run every G chunks) use the runtime switch. Requires S >= 3 and C <= 8*T.
// ================================================================================================================
__host__ __device__ constexpr int dc_hh(int L) { return L == 3 ? 1 : (L - 1) / 2; }
__host__ __device__ constexpr int dc_arr_slots(int hh, int cap) {
    return ((((hh + cap / 2 + 8) + ((hh + cap / 2 + 8) >> 2) + 1) + 7) & ~7) + 4;   // == dc_stage_arr
}

#ifndef PB_DC_TILE_LOADS
#define PB_DC_TILE_LOADS 5
#endif
struct SmemSink {          // next stage's arrays; `half` = its hist/2 + (samples already accumulated)/2
    cpx *ev, *od;
    int half;
    __device__ __forceinline__ void emit4(int u, const cpx (&y)[4], int nout) const {
        const int i = half + 2 * u, s0 = i + (i >> 2), s1 = (i + 1) + ((i + 1) >> 2);
        if (4 * u + 0 < nout) ev[s0] = y[0];
        if (4 * u + 1 < nout) od[s0] = y[1];
        if (4 * u + 2 < nout) ev[s1] = y[2];
        if (4 * u + 3 < nout) od[s1] = y[3];
    }
};
struct GlobalSink {
    cpx *out;
    long long gidx0, first_out;
    __device__ __forceinline__ void emit4(int u, const cpx (&y)[4], int nout) const {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const long long g = gidx0 + 4 * u + r;
            if (4 * u + r < nout && g >= first_out) out[g] = y[r];
        }
    }
};

template <int L, class Sink>
__device__ __forceinline__ void dc_tile4(const cpx *e, const cpx *o, int u, int nout, const Sink &snk) {
    cpx y[4];
    if (L == 3) {   // CIC3, downconvert.cpp:519-535
        cpx ve[5], vo[5];
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            ve[k] = e[k + (k >> 2)];
            vo[k] = o[k + (k >> 2)];
        }
#pragma unroll
        for (int r = 0; r < 4; ++r)
            y[r] = cpx_make(.125 * (vo[r + 1].x + ve[r].x + 3.0 * (vo[r].x + ve[r + 1].x)),
                            .125 * (vo[r + 1].y + ve[r].y + 3.0 * (vo[r].y + ve[r + 1].y)));
    } else {        // half-band, downconvert.cpp:341-393 / 401-496 (tap 0 doubled except for L == 11, finding F7)
        constexpr int E = (L + 1) / 2, CO = (L - 1) / 2, OC = (CO - 1) / 2, OFF = hb_off(L);
        double ar[4] = {0, 0, 0, 0}, ai[4] = {0, 0, 0, 0};
#pragma unroll
        for (int k = 0; k < E + 3; ++k) {
            const cpx v = e[k + (k >> 2)];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int m = k - r;
                if (m >= 0 && m < E) {
                    const double h = (m == 0 && L != 11) ? 2.0 * c_hb[OFF] : c_hb[OFF + 2 * m];
                    ar[r] = fma(h, v.x, ar[r]);
                    ai[r] = fma(h, v.y, ai[r]);
                }
            }
            // keep at most PB_DC_TILE_LOADS shared loads in flight: without the fence ptxas hoists all E+3 loads (up to
            // 80 registers for L = 27) and spills the caller's in-flight sample prefetch around the tile
            if (PB_DC_TILE_LOADS > 0 && (k % PB_DC_TILE_LOADS) == PB_DC_TILE_LOADS - 1) asm volatile("" ::: "memory");
        }
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const cpx v = o[(OC + r) + ((OC + r) >> 2)];
            y[r] = cpx_make(fma(c_hb[OFF + CO], v.x, ar[r]), fma(c_hb[OFF + CO], v.y, ai[r]));
        }
    }
    snk.emit4(u, y, nout);
}

template <class Sink>
__device__ __forceinline__ void dc_tile4_rt(int len, const cpx *e, const cpx *o, int u, int nout, const Sink &snk) {
    switch (len) {
        case 3: dc_tile4<3>(e, o, u, nout, snk); break;
        case 11: dc_tile4<11>(e, o, u, nout, snk); break;
        case 15: dc_tile4<15>(e, o, u, nout, snk); break;
        case 19: dc_tile4<19>(e, o, u, nout, snk); break;
        case 23: dc_tile4<23>(e, o, u, nout, snk); break;
        case 27: dc_tile4<27>(e, o, u, nout, snk); break;
        case 31: dc_tile4<31>(e, o, u, nout, snk); break;
        case 35: dc_tile4<35>(e, o, u, nout, snk); break;
        case 39: dc_tile4<39>(e, o, u, nout, snk); break;
        case 43: dc_tile4<43>(e, o, u, nout, snk); break;
        case 47: dc_tile4<47>(e, o, u, nout, snk); break;
        default: dc_tile4<51>(e, o, u, nout, snk); break;
    }
}

// history carry by the LAST warp: after n (even) samples were consumed, the hh-per-array samples that follow move
// to the front of both arrays.
__device__ __forceinline__ void dc_carry2(cpx *ev, cpx *od, int hh, int n) {
    const int i = PB_DC_THREADS - 1 - (int)threadIdx.x, m = n >> 1;
    if (i >= 32 || m <= 0) return;
    if (m >= hh) {
        for (int q = i; q < 2 * hh; q += 32) {
            cpx *a = (q & 1) ? od : ev;
            const int idx = q >> 1;
            a[dc_slot(idx)] = a[dc_slot(idx + m)];
        }
    } else if (i < 2) {
        cpx *a = i ? od : ev;
        for (int idx = 0; idx < hh; ++idx) a[dc_slot(idx)] = a[dc_slot(idx + m)];   // overlapping: ascending is safe
    }
}

template <int C, int L0, int L1, int L2>
__global__ void __launch_bounds__(PB_DC_THREADS, 3) dc_fast_kernel(const DcChain *__restrict__ ch, const DcCall call) {
    constexpr int T = PB_DC_THREADS, K = C / (2 * T);
    static_assert(K >= 1 && K <= 4, "chunk must be 2*T .. 8*T samples");
    constexpr int HH0 = dc_hh(L0), HH1 = dc_hh(L1), HH2 = dc_hh(L2);
    constexpr int A0 = dc_arr_slots(HH0, C), A1 = dc_arr_slots(HH1, C / 2), A2 = dc_arr_slots(HH2, C / 4);
    constexpr int EV0 = 0, OD0 = A0, EV1 = 2 * A0, OD1 = EV1 + A1, EV2 = EV1 + 2 * A1, OD2 = EV2 + A2, B3 = EV2 + 2 * A2;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ int blen[PB_DC_MAX_STAGES], bhh[PB_DC_MAX_STAGES], boff[PB_DC_MAX_STAGES + 1];
    cpx *smem = reinterpret_cast<cpx *>(smem_raw);
    const int tid = threadIdx.x;
    const int S = ch->nstages;
    const int G = call.group;
    if (tid == 0) {
        int off = B3;
        for (int s = 3; s < S; ++s) {
            blen[s] = ch->st[s].len;
            bhh[s] = ch->st[s].hist >> 1;
            boff[s] = off;
            off += 2 * dc_stage_arr(ch->st[s], dc_stage_cap(C, 3, G, S, s));
        }
        boff[S] = off;
    }
    __syncthreads();

    const long long seg = blockIdx.x;
    const long long seg_begin = seg * call.seg_len;
    long long seg_end = seg_begin + call.seg_len;
    if (seg_end > call.n) seg_end = call.n;
    const long long start = seg == 0 ? 0 : seg_begin - ch->preroll;
    const long long first_out = seg_begin >> S;
    const bool last_seg = seg_end == call.n;

    // ---- initial history: carried state for segment 0, zeros (then pre-roll) otherwise
    for (int s = 0; s < S; ++s) {
        const int hist = ch->st[s].hist;
        const int ev = s == 0 ? EV0 : s == 1 ? EV1 : s == 2 ? EV2 : boff[s];
        const int od = s == 0 ? OD0 : s == 1 ? OD1 : s == 2 ? OD2
                     : boff[s] + dc_stage_arr(ch->st[s], dc_stage_cap(C, 3, G, S, s));
        for (int i = tid; i < hist; i += T)
            smem[((i & 1) ? od : ev) + dc_slot(i >> 1)] = seg == 0 ? call.state_in[s * PB_DC_PREFIX + i] : cpx_make(0.0, 0.0);
    }

    const long long total = seg_end - start;
    const int nchunks = (int)((total + C - 1) / C);
    const cpx w1 = ch->w1, wchunk = ch->wchunk;
    const double rinf = ch->r_inf;
    cpx wp[K];
    int sl[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
        wp[k] = ch->wpair[k];
        sl[k] = dc_slot(HH0 + tid + k * T);
    }
    cpx nx0[K], nx1[K];   // prefetched pairs of the next chunk
    {
        const int len0 = (int)(total < C ? total : C);
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int p = tid + k * T;
            nx0[k] = nx1[k] = cpx_make(0.0, 0.0);
            if (2 * p < len0) dc_ld256(call.in + start + 2 * p, nx0[k], nx1[k]);
        }
    }
    __syncthreads();

    cpx seed = cpx_make(0.0, 0.0);   // r_inf * exp(j*phase(pos + 2*tid + 1)) of the current chunk
    int accF = 0, ingroup = 0;
    long long group_pos = start;
    int pend_s = -1, pend_n = 0;     // history carry still owed (2 = front stage 2, >= 3 = last back stage)
    const cpx *e_t = smem + 5 * tid; // tile base of this thread (array offset added per stage)

    for (int ci = 0; ci < nchunks; ++ci) {
        const long long pos = start + (long long)ci * C;
        const int len = (int)((seg_end - pos) < C ? (seg_end - pos) : C);
        // ---------------- P1: oscillator + mix into stage 0's arrays -------------------------------------------
        cpx c0[K], c1[K];
#pragma unroll
        for (int k = 0; k < K; ++k) { c0[k] = nx0[k]; c1[k] = nx1[k]; }
        if (ci + 1 < nchunks) {
            const long long p2 = pos + C;
            const int l2 = (int)((seg_end - p2) < C ? (seg_end - p2) : C);
#pragma unroll
            for (int k = 0; k < K; ++k)
                if (2 * (tid + k * T) < l2) dc_ld256(call.in + p2 + 2 * (tid + k * T), nx0[k], nx1[k]);
        }
        if ((ci & 31) == 0) {
            seed = cpx_scale(dc_phase_to_osc(dc_phase_at(call.phase_hi, call.phase_lo, ch->turn_hi, ch->turn_lo,
                                                         (uint64_t)(pos + 2 * tid + 1))), rinf);
        } else {
            seed = cpx_mul(seed, wchunk);
        }
        const bool transient = call.stream_pos + pos < PB_DC_RTAB;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int p = tid + k * T;
            if (2 * p < len) {
                cpx o0 = cpx_mul(seed, wp[k]);
                cpx o1 = cpx_mul(o0, w1);
                if (transient) {   // amplitude still converging (first PB_DC_RTAB samples of a stream)
                    const long long g = call.stream_pos + pos + 2 * p;
                    o0 = cpx_scale(o0, (g < PB_DC_RTAB ? ch->rtab[g] : rinf) * ch->inv_r_inf);
                    o1 = cpx_scale(o1, (g + 1 < PB_DC_RTAB ? ch->rtab[g + 1] : rinf) * ch->inv_r_inf);
                }
                smem[EV0 + sl[k]] = cpx_mul(c0[k], o0);
                smem[OD0 + sl[k]] = cpx_mul(c1[k], o1);
            }
        }
        if (pend_s == 2) dc_carry2(smem + EV2, smem + OD2, HH2, pend_n);
        else if (pend_s > 2)
            dc_carry2(smem + boff[pend_s], smem + boff[pend_s] + ((boff[pend_s + 1] - boff[pend_s]) >> 1), bhh[pend_s], pend_n);
        pend_s = -1;
        __syncthreads();

        // ---------------- front stages ---------------------------------------------------------------------------
        {   // stage 0: buf0 -> buf1
            const int nout = len >> 1;
            if (4 * tid < nout) dc_tile4<L0>(e_t + EV0, e_t + OD0, tid, nout, SmemSink{smem + EV1, smem + OD1, HH1});
        }
        __syncthreads();
        {   // stage 1: buf1 -> buf2 ; carry buf0
            const int nout = len >> 2;
            if (4 * tid < nout) dc_tile4<L1>(e_t + EV1, e_t + OD1, tid, nout, SmemSink{smem + EV2, smem + OD2, HH2});
            dc_carry2(smem + EV0, smem + OD0, HH0, len);
        }
        __syncthreads();
        {   // stage 2: buf2 -> buf3 (accumulating) or global ; carry buf1
            const int nout = len >> 3;
            if (4 * tid < nout) {
                if (S > 3) dc_tile4<L2>(e_t + EV2, e_t + OD2, tid, nout,
                                        SmemSink{smem + B3, smem + B3 + ((boff[4 > S ? S : 4] - B3) >> 1), bhh[3] + (accF >> 1)});
                else dc_tile4<L2>(e_t + EV2, e_t + OD2, tid, nout, GlobalSink{call.out, pos >> 3, first_out});
            }
            dc_carry2(smem + EV1, smem + OD1, HH1, len >> 1);
        }
        __syncthreads();
        pend_s = 2;
        pend_n = len >> 2;
        // ---------------- back stages 3..S-1 every G chunks ------------------------------------------------------
        if (S > 3) {
            accF += len >> 3;
            ++ingroup;
            if (ingroup == G || ci == nchunks - 1) {
                for (int s = 3; s < S; ++s) {
                    const int nout = accF >> (s - 2);            // inputs accF >> (s-3), halved
                    const int ev = boff[s], half = (boff[s + 1] - boff[s]) >> 1;
                    if (4 * tid < nout) {
                        if (s == S - 1)
                            dc_tile4_rt(blen[s], e_t + ev, e_t + ev + half, tid, nout,
                                        GlobalSink{call.out, group_pos >> (s + 1), first_out});
                        else
                            dc_tile4_rt(blen[s], e_t + ev, e_t + ev + half, tid, nout,
                                        SmemSink{smem + boff[s + 1], smem + boff[s + 1] + ((boff[s + 2] - boff[s + 1]) >> 1), bhh[s + 1]});
                    }
                    if (s == 3) dc_carry2(smem + EV2, smem + OD2, HH2, len >> 2);
                    else dc_carry2(smem + boff[s - 1], smem + boff[s - 1] + ((boff[s] - boff[s - 1]) >> 1), bhh[s - 1], accF >> (s - 4));
                    __syncthreads();
                }
                pend_s = S - 1;
                pend_n = accF >> (S - 4);
                accF = 0;
                ingroup = 0;
                group_pos = pos + len;
            }
        }
    }

    if (last_seg) {
        if (pend_s == 2) dc_carry2(smem + EV2, smem + OD2, HH2, pend_n);
        else if (pend_s > 2)
            dc_carry2(smem + boff[pend_s], smem + boff[pend_s] + ((boff[pend_s + 1] - boff[pend_s]) >> 1), bhh[pend_s], pend_n);
        __syncthreads();
        for (int s = 0; s < S; ++s) {
            const int hist = ch->st[s].hist;
            const int ev = s == 0 ? EV0 : s == 1 ? EV1 : s == 2 ? EV2 : boff[s];
            const int od = s == 0 ? OD0 : s == 1 ? OD1 : s == 2 ? OD2 : boff[s] + ((boff[s + 1] - boff[s]) >> 1);
            for (int i = tid; i < hist; i += T)
                call.state_out[s * PB_DC_PREFIX + i] = smem[((i & 1) ? od : ev) + dc_slot(i >> 1)];
        }
    }
}
#endif
