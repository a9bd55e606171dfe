// dc_cic.cuh — dc_cic_kernel: CDownConvert chains whose first stage is the CIC3 (every chain SetDataRate plans for
// an input rate above ~667x the output bandwidth, downconvert.cpp:152-165 — BASELINE configs 2 and 4).
//
// Same decomposition as dc_fast_kernel (dc_fused.cuh) with the oscillator, the mix AND the first decimate-by-2 fused
// in registers, so the full-rate stream never touches shared memory at all:
//   * a thread owns NS = 2K CONSECUTIVE samples of the chunk (K coalesced 256-bit loads, prefetched one chunk ahead);
//   * the CIC3 (downconvert.cpp:519-535)  y[P] = .125*x[2P+1] + .375*x[2P] + .375*x[2P-1] + .125*x[2P-2]  is split
//     by input pair:  y[P] = A[P] + B[P-1],  A[P] = .125*m[2P+1] + .375*m[2P],  B[P] = .125*m[2P] + .375*m[2P+1]
//     (m = mixed samples). A and B of a pair need only that pair; B[P-1] comes from the thread's own previous pair,
//     from the neighbouring lane by one warp shuffle, or — for lane 0 — from a redundantly mixed HALO pair;
//   * the thread stores its K CIC3 outputs straight into stage 1's even/odd arrays.
// Stages 1 and 2 (compile-time lengths) then run once per chunk, stages 3..S-1 every G chunks, as in dc_fast_kernel.
// The FULL template flag strips every bounds/transient test from the steady-state chunk; the first chunk of a
// segment, the last one, partial chunks and the first PB_DC_RTAB samples of a stream take the guarded variant.
#pragma once
#include "dc_fused.cuh"

#if defined(__CUDACC__)

__device__ __forceinline__ cpx dc_shfl_up1(cpx v) {
    return cpx_make(__shfl_up_sync(0xffffffffu, v.x, 1), __shfl_up_sync(0xffffffffu, v.y, 1));
}

// Oscillator + mix + CIC3 for one thread's NS samples. Returns the K outputs in y[]; m_last0/1 = the thread's last
// mixed pair (the stage-0 state a following call needs).
// PERM: the warp's 32 sample groups are dealt to its lanes as group = 4*(lane & 7) + (lane >> 3) (dc_pipe_kernel, K == 2),
// so the previous group lives 8 lanes down, or 23 lanes up for the first quarter.
template <int K, bool FULL, bool PERM = false>
__device__ __forceinline__ void dc_mix_cic(const cpx (&x0)[K], const cpx (&x1)[K], cpx xh0, cpx xh1, cpx seed, cpx w1, cpx w2,
                                           const DcChain *__restrict__ ch, long long g0 /* stream index of x0[0] */,
                                           bool halo_from_state, cpx b_state, cpx (&y)[K], cpx &m_last0, cpx &m_last1) {
    // oscillator of the even / odd sample of the current pair, stepped by w^2 per pair (two short dependent chains
    // instead of 2K values held at once: the register budget is what bounds the chunk length)
    cpx oe = seed, oo = cpx_mul(seed, w1);
    cpx oh0 = cpx_mulc(seed, w2), oh1 = cpx_mulc(seed, w1);   // oscillator two and one samples BEFORE x0[0]
    const bool transient = !FULL && g0 < PB_DC_RTAB;   // amplitude still converging (first PB_DC_RTAB samples, finding F5)
    const double ir = transient ? ch->inv_r_inf : 1.0, rinf = transient ? ch->r_inf : 1.0;
    if (transient) {
        oh0 = cpx_scale(oh0, (g0 >= 2 ? ch->rtab[g0 - 2] : rinf) * ir);
        oh1 = cpx_scale(oh1, (g0 >= 1 ? ch->rtab[g0 - 1] : rinf) * ir);
    }
    cpx bprev;
    {
        const cpx h0 = cpx_mul(xh0, oh0), h1 = cpx_mul(xh1, oh1);
        bprev = cpx_make(fma(.375, h1.x, .125 * h0.x), fma(.375, h1.y, .125 * h0.y));
        if (!FULL && halo_from_state) bprev = b_state;
    }
    cpx b = cpx_make(0.0, 0.0);
#pragma unroll
    for (int q = 0; q < K; ++q) {
        cpx ce = oe, co = oo;
        if (transient) {
            ce = cpx_scale(oe, (g0 + 2 * q < PB_DC_RTAB ? ch->rtab[g0 + 2 * q] : rinf) * ir);
            co = cpx_scale(oo, (g0 + 2 * q + 1 < PB_DC_RTAB ? ch->rtab[g0 + 2 * q + 1] : rinf) * ir);
        }
        const cpx m0 = cpx_mul(x0[q], ce), m1 = cpx_mul(x1[q], co);
        if (q + 1 < K) { oe = cpx_mul(oe, w2); oo = cpx_mul(oo, w2); }
        const cpx a = cpx_make(fma(.375, m0.x, .125 * m1.x), fma(.375, m0.y, .125 * m1.y));
        const cpx bn = cpx_make(fma(.375, m1.x, .125 * m0.x), fma(.375, m1.y, .125 * m0.y));
        if (q == 0) y[0] = a;                       // completed below once the neighbour's B is known
        else y[q] = cpx_add(a, b);
        b = bn;
        if (q == K - 1) { m_last0 = m0; m_last1 = m1; }
    }
    cpx bup;                                        // the previous group's last pair
    if (PERM) {
        const int lane = threadIdx.x & 31, src = lane >= 8 ? lane - 8 : lane + 23;
        bup = cpx_make(__shfl_sync(0xffffffffu, b.x, src), __shfl_sync(0xffffffffu, b.y, src));
    } else {
        bup = dc_shfl_up1(b);
    }
    y[0] = cpx_add(y[0], (threadIdx.x & 31) ? bup : bprev);
}

template <int C, int L1, int L2, int MINB>
__global__ void __launch_bounds__(PB_DC_THREADS, MINB) dc_cic_kernel(const DcChain *__restrict__ ch, const DcCall call) {
    constexpr int T = PB_DC_THREADS, K = C / (2 * T), NS = 2 * K;
    static_assert(K == 2 || K == 4, "chunk must be 4*T or 8*T samples");
    constexpr int HH1 = dc_hh(L1), HH2 = dc_hh(L2);
    constexpr int A1 = dc_arr_slots(HH1, C / 2), A2 = dc_arr_slots(HH2, C / 4);
    constexpr int EV1 = 0, OD1 = A1, EV2 = 2 * A1, OD2 = EV2 + A2, B3 = EV2 + 2 * A2;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ int blen[PB_DC_MAX_STAGES + 2], bhh[PB_DC_MAX_STAGES + 2], bev[PB_DC_MAX_STAGES + 2], bod[PB_DC_MAX_STAGES + 2];
    cpx *smem = reinterpret_cast<cpx *>(smem_raw);
    const int tid = threadIdx.x;
    const int S = ch->nstages;
    const int G = call.group;
    if (tid == 0) {
        int off = B3;
        for (int s = 3; s < S; ++s) {
            const int arr = dc_stage_arr(ch->st[s], dc_stage_cap(C, 3, G, S, s));
            blen[s] = ch->st[s].len;
            bhh[s] = ch->st[s].hist >> 1;
            bev[s] = off;
            bod[s] = off + arr;
            off += 2 * arr;
        }
        bev[1] = EV1; bod[1] = OD1; bev[2] = EV2; bod[2] = OD2;
        bhh[1] = HH1; bhh[2] = HH2;
        bev[S] = bod[S] = 0; bhh[S] = 0;
    }
    __syncthreads();

    const int chn = call.nch > 1 ? (int)(blockIdx.x % (unsigned)call.nch) : 0;
    const long long seg = call.nch > 1 ? blockIdx.x / (unsigned)call.nch : blockIdx.x;
    const cpx *state_in = call.state_in + (long long)chn * call.state_stride;
    cpx *state_out = call.state_out + (long long)chn * call.state_stride;
    cpx *gout = call.out + chn * call.out_stride;
    const long long seg_begin = seg * call.seg_len;
    long long seg_end = seg_begin + call.seg_len;
    if (seg_end > call.n) seg_end = call.n;
    const long long start = seg == 0 ? 0 : seg_begin - ch->preroll;
    const long long first_out = seg_begin >> S;
    const bool last_seg = seg_end == call.n;

    // ---- initial history of stages 1..S-1: carried state for segment 0, zeros (then pre-roll) otherwise
    for (int s = 1; s < S; ++s) {
        const int hist = 2 * bhh[s];
        for (int i = tid; i < hist; i += T)
            smem[((i & 1) ? bod[s] : bev[s]) + dc_slot(i >> 1)] = seg == 0 ? state_in[s * PB_DC_PREFIX + i] : cpx_make(0.0, 0.0);
    }
    // stage 0 (CIC3) state = the last mixed pair (downconvert.cpp:528-531): B of that pair, used by thread 0 of chunk 0
    cpx b_state = cpx_make(0.0, 0.0);
    if (seg == 0) {
        const cpx e = state_in[0], o = state_in[1];
        b_state = cpx_make(fma(.375, o.x, .125 * e.x), fma(.375, o.y, .125 * e.y));
    }

    const long long total = seg_end - start;
    const int nchunks = (int)((total + C - 1) / C);
    // chunks that must take the guarded path because the oscillator amplitude is still converging
    const long long tr = PB_DC_RTAB - (call.stream_pos + start);
    const int n_transient = tr > 0 ? (int)((tr + C - 1) / C) : 0;
    const cpx w1 = call.chan ? call.chan[chn].w1 : ch->w1, w2 = cpx_mul(w1, w1);
    const cpx wchunk = call.chan ? call.chan[chn].wchunk : ch->wchunk;
    const uint64_t turn_hi = call.chan ? call.chan[chn].turn_hi : ch->turn_hi, turn_lo = call.chan ? call.chan[chn].turn_lo : ch->turn_lo;
    const uint64_t phase_hi = call.chan ? call.chan[chn].phase_hi : call.phase_hi;
    const uint64_t phase_lo = call.chan ? call.chan[chn].phase_lo : call.phase_lo;
    const double rinf = ch->r_inf;
    int sl[K];   // slot of this thread's q-th CIC3 output inside stage 1's arrays (even q -> EV1, odd q -> OD1)
#pragma unroll
    for (int q = 0; q < K; ++q) sl[q] = ((q & 1) ? OD1 : EV1) + dc_slot(HH1 + ((K * tid + q) >> 1));

    const cpx *in_t = call.in + start + NS * tid;    // this thread's samples of the current chunk
    const bool halo_lane = (tid & 31) == 0;
    cpx nx0[K], nx1[K], nxh0 = cpx_make(0.0, 0.0), nxh1 = cpx_make(0.0, 0.0);
    {
        const int len0 = (int)(total < C ? total : C);
#pragma unroll
        for (int k = 0; k < K; ++k) {
            nx0[k] = nx1[k] = cpx_make(0.0, 0.0);
            if (NS * tid < len0) dc_ld256(in_t + 2 * k, nx0[k], nx1[k]);
        }
        if (halo_lane && tid > 0 && NS * tid < len0) dc_ld256(in_t - 2, nxh0, nxh1);
    }
    __syncthreads();

    cpx seed = cpx_make(0.0, 0.0);   // r_inf * exp(j*phase(pos + NS*tid + 1)) of the current chunk
    int accF = 0, ingroup = 0;
    long long group_pos = start;
    int pend_s = -1, pend_n = 0;     // history carry still owed (2 = front stage 2, >= 3 = last back stage)
    const cpx *e_t = smem + 5 * tid; // tile base of this thread (array offset added per stage)

    for (int ci = 0; ci < nchunks; ++ci) {
        const bool full = ci > 0 && ci >= n_transient && ci < nchunks - 1;   // => len == C
        const long long pos = start + (long long)ci * C;
        const int len = full ? C : (int)((seg_end - pos) < C ? (seg_end - pos) : C);
        // ---------------- P1: oscillator + mix + CIC3 in registers -> stage 1's arrays ---------------------------
        cpx c0[K], c1[K], y[K], ml0, ml1;
        const cpx ch0 = nxh0, ch1 = nxh1;
#pragma unroll
        for (int k = 0; k < K; ++k) { c0[k] = nx0[k]; c1[k] = nx1[k]; }
        in_t += C;
        if (ci + 1 < nchunks) {
            if (ci + 2 < nchunks) {          // next chunk is complete
#pragma unroll
                for (int k = 0; k < K; ++k) dc_ld256(in_t + 2 * k, nx0[k], nx1[k]);
                if (halo_lane) dc_ld256(in_t - 2, nxh0, nxh1);
            } else {
                const int l2 = (int)(seg_end - (pos + C));
                if (NS * tid < l2) {
#pragma unroll
                    for (int k = 0; k < K; ++k) dc_ld256(in_t + 2 * k, nx0[k], nx1[k]);
                    if (halo_lane) dc_ld256(in_t - 2, nxh0, nxh1);
                }
            }
        }
        if ((ci & 31) == 0) {
            seed = cpx_scale(dc_phase_to_osc(dc_phase_at(phase_hi, phase_lo, turn_hi, turn_lo,
                                                         (uint64_t)(pos + NS * tid + 1))), rinf);
        } else {
            seed = cpx_mul(seed, wchunk);
        }
        if (full) {
            dc_mix_cic<K, true>(c0, c1, ch0, ch1, seed, w1, w2, ch, 0, false, b_state, y, ml0, ml1);
#pragma unroll
            for (int q = 0; q < K; ++q) smem[sl[q]] = y[q];
        } else {
            dc_mix_cic<K, false>(c0, c1, ch0, ch1, seed, w1, w2, ch, call.stream_pos + pos + NS * tid,
                                 ci == 0 && tid == 0, b_state, y, ml0, ml1);
            if (NS * tid < len) {
#pragma unroll
                for (int q = 0; q < K; ++q) smem[sl[q]] = y[q];
                if (last_seg && ci == nchunks - 1 && NS * tid + NS == len) {   // CIC3 delay line for the next call
                    state_out[0] = ml0;
                    state_out[1] = ml1;
                }
            }
        }
        if (pend_s == 2) dc_carry2(smem + EV2, smem + OD2, HH2, pend_n);
        else if (pend_s > 2) dc_carry2(smem + bev[pend_s], smem + bod[pend_s], bhh[pend_s], pend_n);
        pend_s = -1;
        __syncthreads();

        // ---------------- stage 1: buf1 -> buf2 ------------------------------------------------------------------
        {
            const int nout = len >> 2;
            if (4 * tid < nout) dc_tile4<L1>(e_t + EV1, e_t + OD1, tid, nout, SmemSink{smem + EV2, smem + OD2, HH2});
        }
        __syncthreads();
        // ---------------- stage 2: buf2 -> buf3 (accumulating) or global ; carry buf1 -----------------------------
        {
            const int nout = len >> 3;
            if (4 * tid < nout) {
                if (S > 3) dc_tile4<L2>(e_t + EV2, e_t + OD2, tid, nout, SmemSink{smem + bev[3], smem + bod[3], bhh[3] + (accF >> 1)});
                else dc_tile4<L2>(e_t + EV2, e_t + OD2, tid, nout, GlobalSink{gout, pos >> 3, first_out});
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
                    if (4 * tid < nout) {
                        if (s == S - 1)
                            dc_tile4_rt(blen[s], e_t + bev[s], e_t + bod[s], tid, nout,
                                        GlobalSink{gout, group_pos >> (s + 1), first_out});
                        else
                            dc_tile4_rt(blen[s], e_t + bev[s], e_t + bod[s], tid, nout,
                                        SmemSink{smem + bev[s + 1], smem + bod[s + 1], bhh[s + 1]});
                    }
                    dc_carry2(smem + bev[s - 1], smem + bod[s - 1], bhh[s - 1], s == 3 ? (len >> 2) : (accF >> (s - 4)));
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
        if (pend_s >= 2) dc_carry2(smem + bev[pend_s], smem + bod[pend_s], bhh[pend_s], pend_n);
        __syncthreads();
        for (int s = 1; s < S; ++s) {
            const int hist = 2 * bhh[s];
            for (int i = tid; i < hist; i += T)
                state_out[s * PB_DC_PREFIX + i] = smem[((i & 1) ? bod[s] : bev[s]) + dc_slot(i >> 1)];
        }
    }
}
#endif
