// dc_pipe.cuh — dc_pipe_kernel: the CIC3-first down-converter (dc_cic.cuh) with ALL six half-band stages fixed at
// compile time and SOFTWARE-PIPELINED ACROSS CHUNKS, one CTA barrier per chunk.
//
// dc_cic_kernel runs mix+CIC3, stage 1, stage 2 and (every G chunks) the back stages as separate phases of one chunk,
// each ended by a CTA barrier and each keeping fewer warps busy than the last: ncu puts ~42 % of its warp-stall samples
// at barriers (profiles/r1_dc_history.md). Here phase p does, between two barriers,
//     all warps     : oscillator + mix + CIC3 of chunk p                     -> stage 1 input, half p&1
//     warps of s = 1: stage 1 on chunk p-1                                   -> stage 2 input
//     warps of s = 2: stage 2 on chunk p-2                                   -> stage 3 input
//     ...            stage s, every P_s phases, on the P_s chunks it has accumulated (a full warp of 4-output tiles)
// Every stage input is double-buffered ("halves"): job j of stage s reads half j&1 while the stage before it fills the
// other half, and the hist_s samples a job leaves for the next one are copied to the head of the other half in the
// same phase. Schedule (all compile-time):
//     tiles per chunk  TPC_s = C / 2^(s+3)          period  P_s = max(1, TJ / TPC_s)       warps  NW_s = max(1, TPC_s / 32)
//     job j of stage s runs in phase  D_s + j P_s,   D_1 = 1,  D_s = D_(s-1) + P_(s-1) (P_s / P_(s-1) - 1) + 1
// i.e. one phase after the last job of stage s-1 that feeds it. The half it reads is written again no earlier than two
// phases later (checked in DESIGN.md section 4), so one barrier per phase orders everything.
// Reference: downconvert.cpp:281-287 (oscillator), 519-535 (CIC3), 341-496 (half-bands); same arithmetic as dc_cic.cuh.
#pragma once
#include "dc_cic.cuh"
#include "dc_input.cuh"

#if defined(__CUDACC__)
// TJ = tiles a job of a late stage waits for (32 = a full warp; 16 halves the late stages' buffers: 3 CTAs per SM)
template <int C, int TJ, int T, int L1, int L2, int L3, int L4, int L5, int L6> struct DcPipe {
    static constexpr int NST = 6;   // half-band stages 1..6 after the CIC3
    __host__ __device__ static constexpr int len(int s) { return s == 1 ? L1 : s == 2 ? L2 : s == 3 ? L3 : s == 4 ? L4 : s == 5 ? L5 : L6; }
    __host__ __device__ static constexpr int hh(int s) { return dc_hh(len(s)); }
    __host__ __device__ static constexpr int tpc(int s) { return C >> (s + 3); }
    __host__ __device__ static constexpr int per(int s) { return tpc(s) >= TJ ? 1 : TJ / tpc(s); }
    __host__ __device__ static constexpr int nw(int s) { return tpc(s) >= 32 ? tpc(s) / 32 : 1; }
    __host__ __device__ static constexpr int cap(int s) { return per(s) * (C >> s); }
    __host__ __device__ static constexpr int arr(int s) { return dc_arr_slots(hh(s), cap(s)); }
    __host__ __device__ static constexpr int delay(int s) { return s == 1 ? 1 : delay(s - 1) + per(s - 1) * (per(s) / per(s - 1) - 1) + 1; }
    __host__ __device__ static constexpr int slots(int s) { return s == 1 ? 3 : 2; }   // stage 1 input is triple-buffered
    __host__ __device__ static constexpr int base(int s) { return s == 1 ? 0 : base(s - 1) + 2 * slots(s - 1) * arr(s - 1); }   // [slot][parity]
    __host__ __device__ static constexpr int warp0(int s) {
        int w = 0;
        for (int q = 1; q < s; ++q) w += nw(q);
        return w < T / 32 - 1 ? w : T / 32 - 1;   // stages that do not fit share the last warp
    }
    __host__ __device__ static constexpr int total() { return base(NST + 1); }
};

// Per-segment scalars. They live in shared memory and are re-read (volatile) where needed: held in registers across the
// phase loop they were what ptxas spilled to local memory (28 KB of L1 left beside 2 x 111 KB of shared memory, so every
// reload went to L2: long-scoreboard 2.0 stall cycles per issue in the first version of this kernel).
struct DcSeg {
    int N, len_last, n_transient, phases, last_seg, pad;
    long long start, first_out;
    cpx b_state;
};

// Stage s of the pipeline in phase p (see the header). N = chunks of this segment, len_last = length of the last one.
template <class G, int C, int s>
__device__ __forceinline__ void dc_pipe_stage(cpx *smem, int p, int N, int len_last, const volatile DcSeg *si, cpx *gout) {
    constexpr int P = G::per(s), D = G::delay(s), A = G::arr(s), HH = G::hh(s);
    const int u = (int)threadIdx.x - 32 * G::warp0(s);
    if (u < 0 || u >= 32 * G::nw(s)) return;
    const int q = p - D;
    if (q < 0 || (q & (P - 1))) return;
    const int j = q / P;
    const int c0 = j * P;
    if (c0 >= N) return;
    const int c1 = c0 + P < N ? c0 + P : N;
    const int nin = (c1 - c0) * (C >> s) - (c1 == N ? ((C - len_last) >> s) : 0);
    const int nout = nin >> 1;
    constexpr int NSL = G::slots(s);
    cpx *ev = smem + G::base(s) + 2 * (j % NSL) * A, *od = ev + A;
    if (4 * u < nout) {
        if (s == G::NST) {
            dc_tile4<G::len(s)>(ev + 5 * u, od + 5 * u, u, nout, GlobalSink{gout, (si->start + (long long)c0 * C) >> (s + 1), si->first_out});
        } else {
            constexpr int sn = s < G::NST ? s + 1 : s;     // (keeps the constexpr calls in range for the last stage)
            constexpr int PN = G::per(sn), AN = G::arr(sn);
            const int jn = c0 / PN, off = (c0 - jn * PN) * (C >> sn);
            cpx *nev = smem + G::base(sn) + 2 * (jn % G::slots(sn)) * AN;
            dc_tile4<G::len(s)>(ev + 5 * u, od + 5 * u, u, nout, SmemSink{nev, nev + AN, G::hh(sn) + (off >> 1)});
        }
    }
    // the hist_s samples after the consumed ones become the head of the other half (the next job's history; after the
    // last job, the state a following call needs)
    if (u < 2 * HH) {
        const int par = u & 1, idx = u >> 1;
        const cpx *src = (par ? od : ev);
        cpx *dst = smem + G::base(s) + (2 * ((j + 1) % NSL) + par) * A;
        dst[dc_slot(idx)] = src[dc_slot(idx + (nin >> 1))];
    }
}

// Warp-role dispatch: a warp tests only the stages from s on until it finds its own (stages that share the last warp
// are all run by it), instead of every thread testing all six stages every phase.
template <class G, int C, int s>
__device__ __forceinline__ void dc_pipe_stages(cpx *smem, int warp, int p, int N, int len_last, const volatile DcSeg *si, cpx *gout) {
    if constexpr (s <= G::NST) {
        constexpr bool shared_next = s < G::NST && G::warp0(s < G::NST ? s + 1 : s) == G::warp0(s);
        if (warp < G::warp0(s) + G::nw(s)) {
            dc_pipe_stage<G, C, s>(smem, p, N, len_last, si, gout);
            if constexpr (shared_next) dc_pipe_stages<G, C, s + 1>(smem, warp, p, N, len_last, si, gout);
        } else {
            dc_pipe_stages<G, C, s + 1>(smem, warp, p, N, len_last, si, gout);
        }
    }
}

// FMT = sample format of call.in (dc_input.cuh): the raw device samples are converted in registers where the mix consumes
// them; the normalizeIQ scale rides on call.r_inf (the oscillator seed's amplitude).
template <int C, int TJ, int T, int MINB, int L1, int L2, int L3, int L4, int L5, int L6, int FMT = PB_IQ_CPX>
__global__ void __launch_bounds__(T, MINB) dc_pipe_kernel(const DcChain *__restrict__ ch, const DcCall call) {
    using G = DcPipe<C, TJ, T, L1, L2, L3, L4, L5, L6>;
    constexpr int K = C / (2 * T), NS = 2 * K, S = 7;
    static_assert(K == 2 || K == 4, "chunk must be 4*T or 8*T samples");
    static_assert(G::nw(1) <= T / 32, "stage 1 tiles must fit the CTA");
    constexpr int HH1 = G::hh(1), A1 = G::arr(1);
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cpx *smem = reinterpret_cast<cpx *>(smem_raw);
    const int tid = threadIdx.x;
    // Which NS input samples a thread mixes. With K == 2 a thread stores ONE slot per array, and lanes storing consecutive
    // slots span 10 padded slots per quarter-warp (two wavefronts per STS.128). Dealing the warp's 32 groups out as
    // 4*(lane & 7) + (lane >> 3) gives every quarter-warp the stride-5 pattern the tile reads already use: one wavefront.
    constexpr bool PERM = K == 2;
    const int mtid = PERM ? (tid & ~31) | (4 * (tid & 7) + ((tid & 31) >> 3)) : tid;

    // single channel only (the batched form of the 10-stage channelizer chain stays on dc_cic_kernel)
    __shared__ DcSeg seg_info;
    __shared__ __align__(8) unsigned long long phase_bar;   // mbarrier: completes once per phase, when all T threads arrived
    const volatile DcSeg *si = &seg_info;
    const unsigned bar_addr = (unsigned)__cvta_generic_to_shared(&phase_bar);
    if (tid == 0) asm volatile("mbarrier.init.shared.b64 [%0], %1;" ::"r"(bar_addr), "r"(T));
    constexpr int SB = IqRaw<FMT, K>::BYTES;   // bytes per input sample
    const char *in_t;
    IqRaw<FMT, K> raw;
    IqHalo<FMT> rawh;
    rawh.zero();
    const bool halo_lane = (tid & 31) == 0;
    {
        const long long seg = blockIdx.x;
        const cpx *state_in = call.state_in;
        const long long seg_begin = seg * call.seg_len;
        long long seg_end = seg_begin + call.seg_len;
        if (seg_end > call.n) seg_end = call.n;
        const long long start = seg == 0 ? 0 : seg_begin - ch->preroll;
        const long long total = seg_end - start;
        const int N = (int)((total + C - 1) / C);
        // ---- history of the first job of every stage (head of half 0): carried state for segment 0, else zeros + pre-roll
#pragma unroll
        for (int s = 1; s <= G::NST; ++s) {
            const int hist = 2 * G::hh(s);
            if (tid < hist)
                smem[G::base(s) + (tid & 1) * G::arr(s) + dc_slot(tid >> 1)] = seg == 0 ? state_in[s * PB_DC_PREFIX + tid] : cpx_make(0.0, 0.0);
        }
        if (tid == 0) {
            seg_info.N = N;
            seg_info.len_last = (int)(total - (long long)(N - 1) * C);
            const long long tr = PB_DC_RTAB - (call.stream_pos + start);
            seg_info.n_transient = tr > 0 ? (int)((tr + C - 1) / C) : 0;
            int phases = N;   // last phase in which any stage still has a job
#pragma unroll
            for (int s = 1; s <= G::NST; ++s) {
                const int J = (N + G::per(s) - 1) / G::per(s);
                const int e = G::delay(s) + (J - 1) * G::per(s) + 1;
                phases = e > phases ? e : phases;
            }
            seg_info.phases = phases;
            seg_info.last_seg = seg_end == call.n;
            seg_info.start = start;
            seg_info.first_out = seg_begin >> S;
            cpx b_state = cpx_make(0.0, 0.0);   // CIC3 state = B of the last mixed pair (downconvert.cpp:528-531)
            if (seg == 0) {
                const cpx e = state_in[0], o = state_in[1];
                b_state = cpx_make(fma(.375, o.x, .125 * e.x), fma(.375, o.y, .125 * e.y));
            }
            seg_info.b_state = b_state;
        }
        in_t = reinterpret_cast<const char *>(call.in) + (start + NS * mtid) * SB;
        const int len0 = (int)(total < C ? total : C);
        raw.zero();
        if (NS * mtid < len0) raw.load(in_t);
        if (halo_lane && tid > 0 && NS * mtid < len0) rawh.load(in_t - 2 * SB);
    }
    int sl[K];   // slot of this thread's q-th CIC3 output inside half 0 of stage 1's arrays (even q -> even array)
#pragma unroll
    for (int q = 0; q < K; ++q) sl[q] = G::base(1) + (q & 1) * A1 + dc_slot(HH1 + ((K * mtid + q) >> 1));
    __syncthreads();

    cpx seed = cpx_make(0.0, 0.0);
    for (int p = 0; p < si->phases; ++p) {
        const int N = si->N, len_last = si->len_last;
        if (p < N) {
            // ---------------- oscillator + mix + CIC3 of chunk p in registers -> stage 1, half p&1 -------------------
            const int ci = p;
            const bool full = ci > 0 && ci >= si->n_transient && ci < N - 1;
            const int len = ci < N - 1 ? C : len_last;
            const int hoff = (ci % 3) * 2 * A1;
            cpx y[K], ml0, ml1;
            if ((ci & 31) == 0) {
                seed = cpx_scale(dc_phase_to_osc(dc_phase_at(call.phase_hi, call.phase_lo, call.turn_hi, call.turn_lo,
                                                             (uint64_t)(si->start + (long long)ci * C + NS * mtid + 1))), call.r_inf);
            } else {
                seed = cpx_mul(seed, call.wchunk);
            }
            const cpx w1 = call.w1, w2 = cpx_mul(w1, w1);
            cpx nx0[K], nx1[K], nxh0, nxh1;
#pragma unroll
            for (int k = 0; k < K; ++k) raw.get(k, nx0[k], nx1[k]);
            rawh.get(nxh0, nxh1);
            if (full) {
                dc_mix_cic<K, true, PERM>(nx0, nx1, nxh0, nxh1, seed, w1, w2, ch, 0, false, cpx_make(0.0, 0.0), y, ml0, ml1);
            } else {
                const cpx b_state = cpx_make(si->b_state.x, si->b_state.y);
                dc_mix_cic<K, false, PERM>(nx0, nx1, nxh0, nxh1, seed, w1, w2, ch, call.stream_pos + si->start + (long long)ci * C + NS * mtid,
                                     ci == 0 && tid == 0, b_state, y, ml0, ml1);
            }
            // the samples of chunk p+1 are loaded INTO THE SAME REGISTERS as soon as the mix has consumed chunk p's: they
            // are in flight during the stage tiles and the barrier below (no second register set, no copies)
            in_t += C * SB;
            if (ci + 2 < N || (ci + 1 < N && NS * mtid < len_last)) {
                raw.load(in_t);
                if (halo_lane) rawh.load(in_t - 2 * SB);
            }
            if (full) {
#pragma unroll
                for (int q = 0; q < K; ++q) smem[sl[q] + hoff] = y[q];
            } else if (NS * mtid < len) {
#pragma unroll
                for (int q = 0; q < K; ++q) smem[sl[q] + hoff] = y[q];
                if (si->last_seg && ci == N - 1 && NS * mtid + NS == len) {   // CIC3 delay line for the next call
                    call.state_out[0] = ml0;
                    call.state_out[1] = ml1;
                }
            }
        }
        // Split barrier: the tiles of phase p read what phase p-1 produced, so a warp waits for "everyone finished
        // phase p-1" only HERE, after its own mix of chunk p (which depends on nothing from phase p-1: stage 1's input
        // has three slots). Warps with less tile work run ahead by up to one mix and absorb the imbalance.
        if (p > 0) {
            const unsigned parity = (unsigned)(p - 1) & 1u;
            unsigned done;
            do {
                // suspend-time hint: a waiting warp sleeps in hardware instead of re-issuing the probe (the plain form
                // spun: 16 % of all executed warp instructions in the first profile of this kernel)
                asm volatile("{\n\t.reg .pred q;\n\tmbarrier.try_wait.parity.shared.b64 q, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, q;\n\t}"
                             : "=r"(done) : "r"(bar_addr), "r"(parity), "r"(4000u) : "memory");
            } while (!done);
        }
        dc_pipe_stages<G, C, 1>(smem, tid >> 5, p, N, len_last, si, call.out);
        asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared.b64 st, [%0];\n\t}" ::"r"(bar_addr) : "memory");
    }
    __syncthreads();

    if (si->last_seg) {
        // after its last job every stage's carried history sits at the head of half (jobs & 1)
        const int N = si->N;
#pragma unroll
        for (int s = 1; s <= G::NST; ++s) {
            const int J = (N + G::per(s) - 1) / G::per(s);
            const int hist = 2 * G::hh(s);
            if (tid < hist) call.state_out[s * PB_DC_PREFIX + tid] = smem[G::base(s) + (2 * (J % G::slots(s)) + (tid & 1)) * G::arr(s) + dc_slot(tid >> 1)];
        }
    }
}
#endif
