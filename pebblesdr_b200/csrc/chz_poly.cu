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

// One raw sample as it sits in memory (16 / 4 / 8 / 2 bytes) and its conversion. The load is a volatile asm statement: a plain
// __ldg may be SUNK by the compiler to just before the first use of its result (fewer live registers), which put the DRAM
// latency of every row back on the critical path (ncu: the three DMULs that first touch a loaded sample held 20 % of all stall
// samples although the loads were written PF rows earlier). The conversion to double happens where the sample is consumed.
template <int FMT> struct PolyRaw;
template <> struct PolyRaw<PB_IQ_CPX> {
    double x, y;
    __device__ __forceinline__ void load(const void *base, long long i) {
        asm volatile("ld.global.nc.v2.f64 {%0,%1}, [%2];" : "=d"(x), "=d"(y) : "l"((const cpx *)base + i));
    }
    __device__ __forceinline__ cpx get() const { return cpx_make(x, y); }
};
template <> struct PolyRaw<PB_IQ_CS16> {
    unsigned w;
    __device__ __forceinline__ void load(const void *base, long long i) { asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(w) : "l"((const unsigned *)base + i)); }
    __device__ __forceinline__ cpx get() const {
        const unsigned u = w ^ 0x80008000u;
        return cpx_make(iq_code_to_double(u & 0xffffu, PB_IQ_MAGIC16), iq_code_to_double(u >> 16, PB_IQ_MAGIC16));
    }
};
template <> struct PolyRaw<PB_IQ_CF32> {
    float x, y;
    __device__ __forceinline__ void load(const void *base, long long i) {
        asm volatile("ld.global.nc.v2.f32 {%0,%1}, [%2];" : "=f"(x), "=f"(y) : "l"((const float2 *)base + i));
    }
    __device__ __forceinline__ cpx get() const { return cpx_make((double)x, (double)y); }
};
template <int FMT> struct PolyRaw8 {
    unsigned short w;
    __device__ __forceinline__ void load(const void *base, long long i) { asm volatile("ld.global.nc.u16 %0, [%1];" : "=h"(w) : "l"((const unsigned short *)base + i)); }
    __device__ __forceinline__ cpx get() const {
        unsigned u = w;
        if (FMT == PB_IQ_CS8) u ^= 0x8080u;
        return cpx_make(iq_code_to_double(u & 0xffu, PB_IQ_MAGIC8), iq_code_to_double(u >> 8, PB_IQ_MAGIC8));
    }
};
template <> struct PolyRaw<PB_IQ_CU8> : PolyRaw8<PB_IQ_CU8> {};
template <> struct PolyRaw<PB_IQ_CS8> : PolyRaw8<PB_IQ_CS8> {};
template <int FMT> __device__ __forceinline__ void poly_fetch(PolyRaw<FMT> &r, const void *hist, const void *in, long long j, long long hl) {
    r.load(j < hl ? hist : in, j < hl ? j : j - hl);
}

// Mix + polyphase FIR. Thread = branch p; it walks the rows m = n0-(Q-1) .. of its branch (samples D apart) with a register
// window of the last Q mixed samples and its own Q taps in registers: 2Q DFMA per row against one 16-byte (or 2..8-byte)
// coalesced load. blockIdx.x = group of 128 branches, blockIdx.y = tile of `nt` outputs.
//   * PF rows are kept in flight in a register ring of RAW samples: row k + PF is requested before row k is mixed and filtered
//     (rows are D samples apart: every row is a fresh DRAM line; the first version, load at use, spent 1070 cycles per row with the
//     FP64 pipe 19.5 % busy).
//   * The window has STATIC register indices: PF rows are unrolled over Q + PF - 1 entries which are then shifted down by PF
//     (~10 moves per row). Unrolling Q-fold instead (indices (r - q) mod Q, no moves) made 64 KB of code: a quarter of the stall
//     samples were instruction fetches.
//   * No divergent branch in the row loop: the oscillator is re-seeded exactly once per RS rows at the top of an iteration
//     (uniform), and the amplitude transient r_i of the first PB_DC_RTAB stream samples — a per-thread test with a table load
//     behind it — is compiled into a separate instance of the loop that only tiles at the very start of a stream enter.
template <int FMT, int Q, int PF, bool TRANSIENT>
__device__ __forceinline__ void poly_rows(const PolyArgs &a, const int p, const int n0, const int total, const double (&h)[Q], cpx (&w)[Q + PF - 1],
                                          PolyRaw<FMT> (&ring)[PF], cpx &osc, long long &j, int kb0, int kb1) {
    constexpr int RS = 6 * PF;               // rows between exact re-seeds of the oscillator recurrence
    const int D = a.D;
    const long long jmax = a.hist_len + a.n - 1;
    const double inv_r = a.chain->inv_r_inf;
    for (int kb = kb0; kb < kb1; kb += PF) {
        const bool reseed = (kb % RS) == 0;
#pragma unroll
        for (int r = 0; r < PF; ++r) {
            const int k = kb + r;
            const long long jj = j < jmax ? j : jmax;          // rows past the tile end are computed but never stored
            const cpx x = ring[r].get();
            {
                const long long jn = j + (long long)PF * D;
                poly_fetch<FMT>(ring[r], a.hist, a.in, jn < jmax ? jn : jmax, a.hist_len);
            }
            if (r == 0 && reseed)
                osc = cpx_scale(dc_phase_to_osc(dc_phase_at(a.ph_hi, a.ph_lo, a.turn_hi, a.turn_lo, (uint64_t)(jj + 1))), a.seed_amp);
            else
                osc = cpx_mul(osc, a.wD);
            cpx o = osc;
            if (TRANSIENT) {
                const long long g = a.stream_v0 + jj;           // stream index: the first PB_DC_RTAB samples carry the transient r_i
                if (g < PB_DC_RTAB) o = cpx_scale(osc, g >= 0 ? a.chain->rtab[g] * inv_r : 0.0);
            }
            w[Q - 1 + r] = cpx_mul(x, o);                       // newest sample of row k; tap q multiplies the sample q rows back
            double ar0 = 0.0, ai0 = 0.0, ar1 = 0.0, ai1 = 0.0;   // two partial sums per component: shorter DFMA chains
#pragma unroll
            for (int q = 0; q < Q; ++q) {
                const cpx s = w[Q - 1 + r - q];
                if (q & 1) { ar1 = fma(h[q], s.x, ar1); ai1 = fma(h[q], s.y, ai1); }
                else { ar0 = fma(h[q], s.x, ar0); ai0 = fma(h[q], s.y, ai0); }
            }
            if (k >= Q - 1 && k < total) a.v[(long long)(n0 + k - (Q - 1)) * D + p] = cpx_make(ar0 + ar1, ai0 + ai1);
            j += D;
        }
#pragma unroll
        for (int q = 0; q < Q - 1; ++q) w[q] = w[q + PF];
    }
}

template <int FMT, int Q>
__global__ void __launch_bounds__(128, 2) poly_fir_kernel(const PolyArgs a) {
    const int D = a.D;
    const int p = blockIdx.x * 128 + threadIdx.x;
    const int n0 = blockIdx.y * a.nt;
    const int nt = min(a.nt, a.nout - n0);
    double h[Q];
#pragma unroll
    for (int q = 0; q < Q; ++q) h[q] = __ldg(a.taps + (long long)q * D + p);
    constexpr int PF = (Q % 5 == 0) ? 5 : 4;
    cpx w[Q + PF - 1];
#pragma unroll
    for (int q = 0; q < Q + PF - 1; ++q) w[q] = cpx_make(0.0, 0.0);
    const long long jmax = a.hist_len + a.n - 1;
    // row k of this tile: m = n0 - (Q-1) + k, virtual index j = D (m + Q) + c - p
    long long j = (long long)D * (n0 + 1) + a.c - p;
    const int total = nt + Q - 1;
    cpx osc = cpx_make(0.0, 0.0);
    PolyRaw<FMT> ring[PF];
#pragma unroll
    for (int f = 0; f < PF; ++f) {
        const long long jf = j + (long long)f * D;
        poly_fetch<FMT>(ring[f], a.hist, a.in, jf < jmax ? jf : jmax, a.hist_len);
    }
    // rows whose smallest stream index in this CTA (largest p) is still inside the transient table take the TRANSIENT instance
    const long long j_lo = (long long)D * (n0 + 1) + a.c - (blockIdx.x * 128 + 127);
    long long kt = (PB_DC_RTAB - a.stream_v0 - j_lo + D - 1) / D;    // first row with stream index >= PB_DC_RTAB for every thread
    if (kt < 0) kt = 0;
    const int total_pf = (total + PF - 1) / PF * PF;
    constexpr int RS = 6 * PF;
    int ksplit = kt >= total_pf ? total_pf : (int)((kt + RS - 1) / RS * RS);   // on a re-seed boundary, so both instances re-seed alike
    if (ksplit > total_pf) ksplit = total_pf;
    poly_rows<FMT, Q, PF, true>(a, p, n0, total, h, w, ring, osc, j, 0, ksplit);
    poly_rows<FMT, Q, PF, false>(a, p, n0, total, h, w, ring, osc, j, ksplit, total_pf);
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
    // Pass 0 (radix 16, points b + 64 k) takes its inputs straight from global memory: consecutive b are consecutive branches,
    // so the 16 loads of a thread are coalesced across the warp and all in flight at once; staging the rows in shared memory
    // first cost one more shared-memory store and load per point in a kernel whose busiest unit is the shared-memory pipe
    // (ncu: LSU data pipe 69 %, FP64 18 %).
    static_assert(NB * 64 == 256, "one pass-0 butterfly per thread");
    {
        const int t = tid >> 6, b = tid & 63;
        if (t < nb) {
            const cpx *src = v + (long long)(n0 + t) * N + b;
            cpx x[16], w[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) x[k] = __ldg(src + k * 64);
            fft_twiddle_powers<16>(tw, b, N - 1, w);
            DftR<16, false>::run(x);
#pragma unroll
            for (int q = 1; q < 16; ++q) x[q] = cpx_mul(x[q], w[q]);
#pragma unroll
            for (int k = 0; k < 16; ++k) a[t * N + fft_swz<SH>(b + k * 64)] = x[k];
        }
    }
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
