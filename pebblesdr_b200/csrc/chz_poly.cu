// chz_poly.cu — polyphase-DFT front end of the batched channelizer (BASELINE config 4).
//
// The reference runs one CDownConvert per channel (receiver.cpp:908): every channel mixes the full-rate capture with its
// own oscillator and pushes it through the same decimate-by-2 cascade — ~22 kflop per capture sample for 1024 channels.
// Every stage of that cascade is an FIR (CIC3 downconvert.cpp:519-535, HB11 :421-496, generic half-band :357-393), so the
// cascade is ONE decimate-by-D FIR (D = 2^S):  y_k[n] = sum_T h[T] x[Dn + c - T] Osc_k[Dn + c - T]  (h, c: host_design.cpp
// dc_composite). When the channel centres sit on the grid  theta_k = theta_b - 2 pi g_k / D  (config 4: 61.44 MHz / 1024 =
// 60 kHz exactly) the oscillators factor,
//     Osc_k[i] = r_i e^{j phi_b(i)} * e^{-j 2 pi g_k (i+1) / D} * e^{j (i+1) delta_k},
// (r_i: the reference's amplitude transient, downconvert.cpp:281-287, common to all channels; delta_k ~ 1e-16 rad: what
// rounding cos/sin to doubles does to channel k's angle, downconvert.cpp:108-110) and with T = D q + p the grid term loses q:
//     y_k[n] = e^{j E_k(n)} * sum_p e^{+j 2 pi g_k p / D} v_p[n],     v_p[n] = sum_q h[Dq + p] u[D(n - q) + c - p],
//     u[i] = x[i] r_i e^{j phi_b(i)},   E_k(n) = (phi_k - phi_b)(call start) - 2 pi g_k (c+1)/D + (Dn + c - t_c + 1) delta_k.
// That is: one mix of the capture (poly_fir_kernel: the mix fused into a D-branch polyphase FIR, ~25 taps per branch), one
// D-point DFT per output sample (poly_dft_kernel, the shared-memory FFT of fft_core.cuh) and one rotation per channel output
// — ~150 flop per capture sample instead of ~22 500, for ANY number of channels on the grid. The residual delta_k is applied
// at the centre t_c of the filter support; across the support it moves the phase by < L/2 * |delta_k| <= 1.7e-11 rad (the
// path is used only while |delta_k| < 2^-52 turns), inside the 1e-10 parity tolerance. Channels off the grid, D != 1024 or
// more than 32 taps per branch keep the per-channel kernel (dc_cic_kernel, batched).
//
// Carried state = the last Q*D RAW input samples (in the device's sample format): every stage is an FIR, so all delay
// lines are re-derived from them. NOTE (documented deviation): after a retune those samples are re-mixed with the NEW
// oscillator, whereas the reference's delay lines hold samples mixed with the old one; the two differ for the L samples
// (0.4 ms at 61.44 MSPS) after a SetFrequency, which the reference spends in a filter transient of its own.
#include "chz_poly.h"
#include "fft_core.cuh"
#include "dc_input.cuh"
#include "host_design.h"
#include <cstdlib>
#include <cstring>
#include <vector>

struct PolyChan {            // per channel, rebuilt every call (32 B + 8)
    uint64_t a_hi, a_lo;     // E_k at n = 0 of this call, 0.128 turns
    uint64_t d_hi, d_lo;     // D * delta_k, 0.128 turns (per output sample)
    int bin, pad;            // DFT bin that holds channel k: (D - g_k) mod D  (forward transform of v gives sum e^{-j...})
};

struct PolyArgs {
    const void *in, *hist;   // n new samples; Q*D samples preceding them (both in the raw sample format)
    long long n, hist_len;   // hist_len = Q*D
    int D, c, nout, nt;      // decimation, alignment, outputs this call, outputs per CTA tile
    const double *taps;      // [Q][D]: taps[q*D + p] = h[D*q + p]
    cpx *v;                  // branch outputs [nout][D]
    const DcChain *chain;    // rtab / inv_r_inf of the amplitude transient
    uint64_t ph_hi, ph_lo;   // base oscillator phase BEFORE virtual sample 0 (= first history sample)
    uint64_t turn_hi, turn_lo;
    cpx wD;                  // exp(j D theta_b): one row step
    double seed_amp;         // r_inf * normalizeIQ scale
    long long stream_v0;     // stream index of virtual sample 0 (negative while the history predates the stream)
};

template <int FMT> __device__ __forceinline__ cpx poly_load(const void *hist, const void *in, long long j, long long hl) {
    const void *base = j < hl ? hist : in;
    const long long i = j < hl ? j : j - hl;
    if (FMT == PB_IQ_CPX) return __ldg((const cpx *)base + i);
    if (FMT == PB_IQ_CS16) {
        const unsigned w = __ldg((const unsigned *)base + i) ^ 0x80008000u;
        return cpx_make(iq_code_to_double(w & 0xffffu, PB_IQ_MAGIC16), iq_code_to_double(w >> 16, PB_IQ_MAGIC16));
    }
    if (FMT == PB_IQ_CF32) { const float2 f = __ldg((const float2 *)base + i); return cpx_make((double)f.x, (double)f.y); }
    unsigned w = __ldg((const unsigned short *)base + i);
    if (FMT == PB_IQ_CS8) w ^= 0x8080u;
    return cpx_make(iq_code_to_double(w & 0xffu, PB_IQ_MAGIC8), iq_code_to_double(w >> 8, PB_IQ_MAGIC8));
}

// Mix + polyphase FIR. Thread = branch p; it walks the rows m = n0-(Q-1) .. of its branch (samples D apart) with a register
// window of the last Q mixed samples and its own Q taps in registers: 2Q DFMA per row against one 16-byte (or 2..8-byte)
// coalesced load. blockIdx.x = group of 128 branches, blockIdx.y = tile of `nt` outputs.
template <int FMT, int Q>
__global__ void __launch_bounds__(128, 2) poly_fir_kernel(const PolyArgs a) {
    const int D = a.D;
    const int p = blockIdx.x * 128 + threadIdx.x;
    const int n0 = blockIdx.y * a.nt;
    const int nt = min(a.nt, a.nout - n0);
    double h[Q];
#pragma unroll
    for (int q = 0; q < Q; ++q) h[q] = __ldg(a.taps + (long long)q * D + p);
    cpx w[Q];
#pragma unroll
    for (int q = 0; q < Q; ++q) w[q] = cpx_make(0.0, 0.0);
    const long long jmax = a.hist_len + a.n - 1;
    // row k of this tile: m = n0 - (Q-1) + k, virtual index j = D (m + Q) + c - p
    long long j = (long long)D * (n0 + 1) + a.c - p;
    const int total = nt + Q - 1;
    cpx osc = cpx_make(0.0, 0.0);
    const double inv_r = a.chain->inv_r_inf;
    // The rows of a branch are D samples apart, so every row is a fresh DRAM line and a load issued where it is used costs the
    // warp the full memory latency per row (ncu of that form: 1070 cycles per row, FP64 pipe 19.5 % busy with 8 warps per SM;
    // the compiler may not hoist the load over the store to `v`). PF rows are kept in flight in a register ring instead: row
    // k + PF is requested before row k is mixed and filtered (~2 x 110 pipe cycles per row and warp: PF = 5 rows cover ~1100).
    constexpr int PF = (Q % 5 == 0) ? 5 : 4;
    static_assert(Q % PF == 0, "the ring index must be static under the Q-fold unroll");
    cpx ring[PF];
#pragma unroll
    for (int f = 0; f < PF; ++f) {
        const long long jf = j + (long long)f * D;
        ring[f] = poly_load<FMT>(a.hist, a.in, jf < jmax ? jf : jmax, a.hist_len);
    }
    for (int kb = 0; kb < total; kb += Q) {
#pragma unroll
        for (int r = 0; r < Q; ++r) {
            const int k = kb + r;
            const long long jj = j < jmax ? j : jmax;          // rows past the tile end are computed but never stored
            const cpx x = ring[r % PF];
            {
                const long long jn = j + (long long)PF * D;
                ring[r % PF] = poly_load<FMT>(a.hist, a.in, jn < jmax ? jn : jmax, a.hist_len);
            }
            if ((k & 31) == 0)
                osc = cpx_scale(dc_phase_to_osc(dc_phase_at(a.ph_hi, a.ph_lo, a.turn_hi, a.turn_lo, (uint64_t)(jj + 1))), a.seed_amp);
            else
                osc = cpx_mul(osc, a.wD);
            cpx o = osc;
            const long long g = a.stream_v0 + jj;               // stream index: the first PB_DC_RTAB samples carry the transient r_i
            if (g < PB_DC_RTAB) o = cpx_scale(osc, g >= 0 ? a.chain->rtab[g] * inv_r : 0.0);
            w[r] = cpx_mul(x, o);
            double ar0 = 0.0, ai0 = 0.0, ar1 = 0.0, ai1 = 0.0;   // two partial sums per component: shorter DFMA chains
#pragma unroll
            for (int q = 0; q < Q; ++q) {
                const cpx s = w[(r - q + Q) % Q];
                if (q & 1) { ar1 = fma(h[q], s.x, ar1); ai1 = fma(h[q], s.y, ai1); }
                else { ar0 = fma(h[q], s.x, ar0); ai0 = fma(h[q], s.y, ai0); }
            }
            if (k >= Q - 1 && k < total) a.v[(long long)(n0 + k - (Q - 1)) * D + p] = cpx_make(ar0 + ar1, ai0 + ai1);
            j += D;
        }
    }
}

// D-point DFT over the branches of NB consecutive output samples + per-channel rotation, written channel-major.
// Forward transform X[f] = sum_p v_p e^{-j 2 pi f p / D}; channel k needs sum_p v_p e^{+j 2 pi g_k p / D} = X[(D - g_k) mod D].
template <int NB>
__global__ void __launch_bounds__(256) poly_dft_kernel(const cpx *__restrict__ v, int nout, const cpx *__restrict__ tw,
                                                       const PolyChan *__restrict__ chan, int nch, cpx *__restrict__ out,
                                                       long long out_stride) {
    constexpr int N = 1024, SH = FftPlan<N>::SH;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cpx *a = reinterpret_cast<cpx *>(smem_raw);
    const int tid = threadIdx.x;
    const int n0 = blockIdx.x * NB;
    const int nb = min(NB, nout - n0);
    for (int i = tid; i < nb * N; i += 256) {
        const int t = i >> 10, p = i & (N - 1);
        a[t * N + fft_swz<SH>(p)] = __ldg(v + (long long)(n0 + t) * N + p);
    }
    __syncthreads();
    for (int b = tid; b < nb * 64; b += 256) fft_butterfly<N, 16, 1024, false, SH>(a + (b >> 6) * N, tw, b & 63);
    __syncthreads();
    for (int b = tid; b < nb * 128; b += 256) fft_butterfly<N, 8, 64, false, SH>(a + (b >> 7) * N, tw, b & 127);
    __syncthreads();
    for (int b = tid; b < nb * 128; b += 256) fft_butterfly<N, 8, 8, false, SH>(a + (b >> 7) * N, tw, b & 127);
    __syncthreads();
    for (int ch = tid; ch < nch; ch += 256) {
        const PolyChan pc = chan[ch];
        // the residual moves by D*delta ~ 1e-13 rad per output: one rotation at the middle of the NB outputs serves them all
        const cpx rot = dc_phase_to_osc(dc_phase_at(pc.a_hi, pc.a_lo, pc.d_hi, pc.d_lo, (uint64_t)(n0 + nb / 2)));
        const int slot = fft_swz<SH>(FftPlan<N>::freq2pos(pc.bin));
        cpx *o = out + (long long)ch * out_stride + n0;
#pragma unroll
        for (int t = 0; t < NB; ++t)
            if (t < nb) o[t] = cpx_mul(a[t * N + slot], rot);
    }
}

struct ChzPoly {
    int D = 0, Q = 0, c = 0, L = 0, S = 0;
    bool planned = false, eligible = false;
    double *d_taps = nullptr;
    cpx *d_tw = nullptr;
    void *d_hist[2] = {nullptr, nullptr};
    int hcur = 0, hist_fmt = -1;
    size_t hist_bytes = 0;
    cpx *d_v = nullptr; size_t v_cap = 0;
    PolyChan *d_chan = nullptr; int chan_cap = 0;
    std::vector<PolyChan> chan;
    std::vector<int> grid;                 // g_k
    std::vector<unsigned __int128> delta;  // delta_k, 0.128 turns (two's complement)
    bool smem_set = false;
};

typedef unsigned __int128 u128;
static inline u128 mk128(uint64_t hi, uint64_t lo) { return ((u128)hi << 64) | lo; }

ChzPoly *chz_poly_create() { return new ChzPoly(); }
void chz_poly_destroy(ChzPoly *p) {
    if (!p) return;
    cudaFree(p->d_taps); cudaFree(p->d_tw); cudaFree(p->d_hist[0]); cudaFree(p->d_hist[1]); cudaFree(p->d_v); cudaFree(p->d_chan);
    delete p;
}
void chz_poly_invalidate(ChzPoly *p) { if (p) { p->planned = false; p->hist_fmt = -1; } }
bool chz_poly_active(const ChzPoly *p) { return p && p->planned && p->eligible; }

// Builds the prototype filter for `chain` and decides whether the channels sit on a uniform grid. 1 = polyphase path
// applies, 0 = it does not (caller keeps the per-channel kernel), < 0 = CUDA error.
int chz_poly_prepare(ChzPoly *p, const DcChain &chain, const std::vector<DcChan> &chans) {
    if (getenv("PEBBLE_CHZ_DIRECT")) { p->planned = true; p->eligible = false; return 0; }
    if (!p->planned) {
        p->planned = true;
        p->eligible = false;
        p->S = chain.nstages;
        p->D = 1 << chain.nstages;
        if (p->D != 1024) return 0;            // the branch DFT is instantiated for 1024 points (config 4's /1024 chain)
        std::vector<double> h;
        p->L = pebble_host::dc_composite(chain, h, p->c);
        const int q = (p->L + p->D - 1) / p->D;
        p->Q = q <= 25 ? 25 : q <= 32 ? 32 : 0;
        if (!p->Q || p->c >= p->D) return 0;
        std::vector<double> taps((size_t)p->Q * p->D, 0.0);
        for (int t = 0; t < p->L; ++t) taps[t] = h[t];                 // taps[q*D + p] = h[D*q + p]
        std::vector<double> tw;
        pebble_host::twiddle_table(p->D, tw);
        cudaFree(p->d_taps); cudaFree(p->d_tw);
        p->d_taps = nullptr; p->d_tw = nullptr;
        PB_CUDA(cudaMalloc(&p->d_taps, sizeof(double) * taps.size()), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMalloc(&p->d_tw, sizeof(double) * tw.size()), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMemcpy(p->d_taps, taps.data(), sizeof(double) * taps.size(), cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMemcpy(p->d_tw, tw.data(), sizeof(double) * tw.size(), cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
        p->eligible = true;
    }
    if (!p->eligible) return 0;
    // grid test: turn_k - turn_0 = -g_k / D + delta_k with |delta_k| < 2^-52 turns
    const int nch = (int)chans.size();
    p->grid.assign(nch, 0);
    p->delta.assign(nch, 0);
    const u128 t0 = mk128(chans[0].turn_hi, chans[0].turn_lo);
    for (int k = 0; k < nch; ++k) {
        const u128 d = mk128(chans[k].turn_hi, chans[k].turn_lo) - t0;
        const u128 nd = (u128)0 - d;
        const int g = (int)((((nd >> (128 - p->S - 1)) + 1) >> 1) & (u128)(p->D - 1));
        const u128 delta = d + ((u128)g << (128 - p->S));
        const u128 mag = (delta >> 127) ? (u128)0 - delta : delta;
        if (mag >> (128 - 52)) return 0;       // off the grid: per-channel kernel
        p->grid[k] = g;
        p->delta[k] = delta;
    }
    return 1;
}

template <int FMT> static void poly_launch_fir(int Q, dim3 grid, cudaStream_t st, const PolyArgs &a) {
    if (Q == 25) poly_fir_kernel<FMT, 25><<<grid, 128, 0, st>>>(a);
    else poly_fir_kernel<FMT, 32><<<grid, 128, 0, st>>>(a);
}

int chz_poly_process(ChzPoly *p, const DcChain &chain, const DcChain *d_chain, const std::vector<DcChan> &chans, long long stream_pos,
                     int fmt, double in_scale, long long n, const void *d_in, cpx *d_out, long long out_stride, cudaStream_t st) {
    const int D = p->D, Q = p->Q, nch = (int)chans.size();
    const long long hl = (long long)Q * D;
    const int bps = iq_bytes_per_sample(fmt);
    const size_t hbytes = (size_t)hl * bps;
    if (p->hist_fmt != fmt || p->hist_bytes != hbytes) {
        // (re)start: the history before the stream is silence; 0x80 is the code of zero in offset-binary uint8
        PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
        for (int b = 0; b < 2; ++b) {
            cudaFree(p->d_hist[b]); p->d_hist[b] = nullptr;
            PB_CUDA(cudaMalloc(&p->d_hist[b], hbytes), PEBBLE_ERR_CUDA);
            PB_CUDA(cudaMemset(p->d_hist[b], fmt == PB_IQ_CU8 ? 0x80 : 0, hbytes), PEBBLE_ERR_CUDA);
        }
        p->hist_fmt = fmt; p->hist_bytes = hbytes; p->hcur = 0;
    }
    const int nout = (int)(n / D);
    if ((size_t)nout * D > p->v_cap) {
        PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
        cudaFree(p->d_v); p->d_v = nullptr; p->v_cap = 0;
        PB_CUDA(cudaMalloc(&p->d_v, sizeof(cpx) * (size_t)nout * D), PEBBLE_ERR_CUDA);
        p->v_cap = (size_t)nout * D;
    }
    if (nch > p->chan_cap) {
        PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
        cudaFree(p->d_chan); p->d_chan = nullptr; p->chan_cap = 0;
        PB_CUDA(cudaMalloc(&p->d_chan, sizeof(PolyChan) * (size_t)nch), PEBBLE_ERR_CUDA);
        p->chan_cap = nch;
    }
    // per-channel rotation  E_k(n) = (phi_k - phi_b)(call start) - g_k (c+1)/D + (c - t_c + 1) delta_k + n D delta_k
    p->chan.resize(nch);
    const u128 ph_b = mk128(chans[0].phase_hi, chans[0].phase_lo);
    const long long tc = (p->L - 1) / 2;
    for (int k = 0; k < nch; ++k) {
        const u128 delta = p->delta[k];
        u128 e = mk128(chans[k].phase_hi, chans[k].phase_lo) - ph_b;
        e -= (u128)(((long long)p->grid[k] * (p->c + 1)) & (D - 1)) << (128 - p->S);
        e += delta * (u128)(unsigned long long)(long long)(p->c - tc + 1) + ((p->c - tc + 1) < 0 ? ((u128)0 - (delta << 64)) : (u128)0);
        const u128 dd = delta * (u128)D;
        PolyChan &pc = p->chan[k];
        pc.a_hi = (uint64_t)(e >> 64); pc.a_lo = (uint64_t)e;
        pc.d_hi = (uint64_t)(dd >> 64); pc.d_lo = (uint64_t)dd;
        pc.bin = (D - p->grid[k]) & (D - 1);
        pc.pad = 0;
    }
    PB_CUDA(cudaMemcpyAsync(p->d_chan, p->chan.data(), sizeof(PolyChan) * (size_t)nch, cudaMemcpyHostToDevice, st), PEBBLE_ERR_CUDA);

    PolyArgs a;
    a.in = d_in; a.hist = p->d_hist[p->hcur];
    a.n = n; a.hist_len = hl;
    a.D = D; a.c = p->c; a.nout = nout;
    int nt = 126;                                   // Q*B - (Q-1) rows per tile keeps the unrolled blocks full (Q = 25, B = 6)
    if (const char *e = getenv("PEBBLE_CHZ_TILE")) nt = atoi(e) > 0 ? atoi(e) : nt;
    a.nt = nt;
    a.taps = p->d_taps; a.v = p->d_v; a.chain = d_chain;
    // base oscillator = channel 0's; its phase before the first HISTORY sample = phase0 - hl * turn
    {
        const u128 tb = mk128(chans[0].turn_hi, chans[0].turn_lo);
        const u128 pv = ph_b - tb * (u128)(unsigned long long)hl;
        a.ph_hi = (uint64_t)(pv >> 64); a.ph_lo = (uint64_t)pv;
        a.turn_hi = chans[0].turn_hi; a.turn_lo = chans[0].turn_lo;
        const u128 td = tb * (u128)D;               // D * theta_b, exact mod 1
        const double x = (double)(int64_t)(uint64_t)(td >> 64) * (1.0 / 9223372036854775808.0);
        a.wD = cpx_make(cos(x * 3.14159265358979323846264338328), sin(x * 3.14159265358979323846264338328));
    }
    a.seed_amp = chain.r_inf * in_scale;
    a.stream_v0 = stream_pos - hl;
    const dim3 grid((unsigned)(D / 128), (unsigned)((nout + nt - 1) / nt));
    switch (fmt) {
        case PB_IQ_CPX: poly_launch_fir<PB_IQ_CPX>(Q, grid, st, a); break;
        case PB_IQ_CS16: poly_launch_fir<PB_IQ_CS16>(Q, grid, st, a); break;
        case PB_IQ_CU8: poly_launch_fir<PB_IQ_CU8>(Q, grid, st, a); break;
        case PB_IQ_CS8: poly_launch_fir<PB_IQ_CS8>(Q, grid, st, a); break;
        default: poly_launch_fir<PB_IQ_CF32>(Q, grid, st, a); break;
    }
    constexpr int NB = 4;
    const size_t smem = sizeof(cpx) * 1024 * NB;
    if (!p->smem_set) {
        PB_CUDA(cudaFuncSetAttribute(poly_dft_kernel<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), PEBBLE_ERR_CUDA);
        p->smem_set = true;
    }
    poly_dft_kernel<NB><<<(nout + NB - 1) / NB, 256, smem, st>>>(p->d_v, nout, p->d_tw, p->d_chan, nch, d_out, out_stride);
    pebble::count_launch(2);
    PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    // next history = the last hl samples of (history ++ input)
    void *nh = p->d_hist[p->hcur ^ 1];
    if (n >= hl) {
        PB_CUDA(cudaMemcpyAsync(nh, (const char *)d_in + (size_t)(n - hl) * bps, hbytes, cudaMemcpyDeviceToDevice, st), PEBBLE_ERR_CUDA);
    } else {
        const size_t keep = (size_t)(hl - n) * bps;
        PB_CUDA(cudaMemcpyAsync(nh, (const char *)p->d_hist[p->hcur] + (size_t)n * bps, keep, cudaMemcpyDeviceToDevice, st), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMemcpyAsync((char *)nh + keep, d_in, (size_t)n * bps, cudaMemcpyDeviceToDevice, st), PEBBLE_ERR_CUDA);
    }
    p->hcur ^= 1;
    return nout;
}
