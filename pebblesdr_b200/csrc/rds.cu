// rds.cu — the RDS branch of Demod_WFM::processDataStereo (application/demod/demod_wfm.cpp:298-357) on the GPU, and the
// block / group synchroniser behind it (processNewRdsBit :584-700, checkBlock :707-758, getNextRdsGroupData :764-788).
//
//   m_CpxRawFm (the Hilbert-filtered discriminator output, already on the device for the stereo demux)
//     -> CDownConvert: mix by -57 kHz, decimate to ~8-20 kS/s      (m_RdsDownConvert, :190-191, :298)  pebble_downconvert
//     -> complex low-pass 2400 Hz                                  (m_RdsBPFilter, :497, :303)         cfir
//     -> PLL that de-rotates the DSB signal, output = Im           (processRdsPll :541-573)            rds_pll_kernel (serial)
//     -> matched filter: one bi-phase symbol                       (m_RdsMatchedFilter :505-522, :311) cfir (real taps)
//     -> squared, Q = 500 resonator at the bit rate                (:314-319)                          rds_square_kernel + iir
//     -> peak of the resonator = sampling instant, sign of the previous matched-filter sample = raw bit (:321-350)
//                                                                                                     rds_slice_kernel (pointwise)
// The sampling decisions leave the GPU as one byte per RDS-rate sample through a page-locked ring; the differential decode
// (:343) and the 26-bit block / 104-bit group state machine run on the host when the caller asks for groups
// (pebble_demod_get_next_rds_group) or when the ring is half full — never a synchronisation in the audio path otherwise.
// The (26,16) code's parity-check rows and offset syndromes are generated from the generator polynomial
// x^10 + x^8 + x^7 + x^5 + x^4 + x^3 + 1 and the five offset words of the RDS standard (IEC 62106 annex B), not tabulated.
// rdsdecode.cpp (group -> text for the GUI) is out of scope: the groups are the interface.
#include "demod_internal.h"
#include "demod_kernels.cuh"
#include <cmath>
#include <cstring>
#include <vector>

namespace {
constexpr double RDS_FREQUENCY = 57000.0, RDS_BITRATE = RDS_FREQUENCY / 48.0;   // 1187.5 bit/s
constexpr double RDSPLL_RANGE = 12.0, RDSPLL_BW = 1.0, RDSPLL_ZETA = .707;       // demod_wfm.cpp:70-72
constexpr double TWOPI = 6.28318530717958647692528676656;

struct RdsPllConst { double alpha, beta, lo, hi; };
struct RdsPllState { double phase, freq; };
struct RdsSliceState { double last_sync, last_slope, last_data; int pad[2]; };

// processRdsPll (:541-573). One thread: a nonlinear recurrence at ~16 kS/s. The reference takes sin/cos from x87 fsincos;
// sincos() agrees with it to ~1 ulp. Output (v, v) with v = Im of the de-rotated sample: the next step is a real FIR.
__global__ void __launch_bounds__(32) rds_pll_kernel(const cpx *__restrict__ in, int n, RdsPllConst c, const RdsPllState *__restrict__ st_in,
                                                     RdsPllState *__restrict__ st_out, cpx *__restrict__ out) {
    if (threadIdx.x) return;
    RdsPllState v = *st_in;
    for (int i = 0; i < n; ++i) {
        const cpx x = in[i];
        double sn, cs;
        sincos(v.phase, &sn, &cs);
        const double tr = __dsub_rn(__dmul_rn(cs, x.x), __dmul_rn(sn, x.y));               // :553-554
        const double ti = __dadd_rn(__dmul_rn(cs, x.y), __dmul_rn(sn, x.x));
        const double phzerror = -wfm_arctan2(ti, tr);                                      // :556
        v.freq = __dadd_rn(v.freq, __dmul_rn(c.beta, phzerror));                           // :558
        if (v.freq > c.hi) v.freq = c.hi; else if (v.freq < c.lo) v.freq = c.lo;           // :560-563
        v.phase = __dadd_rn(v.phase, __dadd_rn(v.freq, __dmul_rn(c.alpha, phzerror)));     // :565
        out[i] = cpx_make(ti, ti);                                                         // :566
    }
    v.phase = fmod(v.phase, TWOPI);                                                        // :568
    *st_out = v;
}

// m_RdsMag[i] = m_RdsData[i]^2 (:314-315), as (v, v) for the resonator
__global__ void rds_square_kernel(const cpx *__restrict__ data, int n, cpx *__restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { const double d = data[i].x, q = __dmul_rn(d, d); out[i] = cpx_make(q, q); }
}

// The bit slicer (:321-350) without its loop: slope_i = sync_i - sync_(i-1); a sampling instant is a sample whose slope is
// negative after a positive one; the raw bit is the sign of the matched filter ONE sample earlier. code[i]: 0 = no bit,
// 1 = raw bit 0, 2 = raw bit 1. tap[i] = what the reference leaves in m_RdsRaw (value sampled or 0, matched filter output).
__global__ void rds_slice_kernel(const cpx *__restrict__ data, const cpx *__restrict__ sync, int n, const RdsSliceState *__restrict__ st_in,
                                 RdsSliceState *__restrict__ st_out, unsigned char *__restrict__ code, cpx *__restrict__ tap) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const RdsSliceState s = *st_in;
    const double s0 = sync[i].x;
    const double s1 = i >= 1 ? sync[i - 1].x : s.last_sync;
    const double slope = __dsub_rn(s0, s1);
    double prev_slope;
    if (i >= 1) prev_slope = __dsub_rn(s1, i >= 2 ? sync[i - 2].x : s.last_sync);
    else prev_slope = s.last_slope;
    const double prev_data = i >= 1 ? data[i - 1].x : s.last_data;
    const bool hit = slope < 0.0 && __dmul_rn(prev_slope, slope) < 0.0;
    code[i] = hit ? (prev_data >= 0.0 ? 2 : 1) : 0;
    tap[i] = cpx_make(hit ? prev_data : 0.0, data[i].x);
    if (i == n - 1) {
        RdsSliceState o;
        o.last_sync = s0; o.last_slope = slope; o.last_data = data[i].x; o.pad[0] = o.pad[1] = 0;
        *st_out = o;
    }
}

// ---- the (26,16) shortened cyclic code of the RDS standard, from its generator polynomial -----------------------------------
constexpr unsigned RDS_POLY = 0x5B9;   // x^10 + x^8 + x^7 + x^5 + x^4 + x^3 + 1
constexpr unsigned RDS_OFFSET_WORD[5] = {0x0FC, 0x198, 0x168, 0x350, 0x1B4};   // A, B, C, C', D
struct RdsCode {
    unsigned row[16];      // syndrome of message bit 15 - i: x^(340 - i) mod g (the code is the (341,331) cyclic code shortened)
    unsigned offset[5];    // syndrome of each offset word: word(x) x^325 mod g
    RdsCode() {
        auto mulx = [](unsigned v) { v <<= 1; if (v & 0x400) v ^= RDS_POLY; return v & 0x3FFu; };
        unsigned pw[341];
        pw[0] = 1;
        for (int k = 1; k <= 340; ++k) pw[k] = mulx(pw[k - 1]);
        for (int i = 0; i < 16; ++i) row[i] = pw[340 - i];
        for (int w = 0; w < 5; ++w) {
            unsigned s = 0;
            for (int b = 0; b < 10; ++b) if ((RDS_OFFSET_WORD[w] >> b) & 1) s ^= pw[325 + b];
            offset[w] = s;
        }
    }
};
const RdsCode &rds_code() { static const RdsCode c; return c; }

// Block and group synchroniser: the state machine of processNewRdsBit, restated. States as there: 0 bit sync (slide until a
// block passes as A without correction), 1 block sync (B, C/C', D must follow at 26-bit spacing), 2 group decode (burst
// correction allowed; more than 5 bad blocks -> queue cleared, an all-zero group posted, back to 0), 3 skip to next group.
struct RdsSync {
    unsigned shift = 0;
    int state = 0, block = 0, bitpos = 0, version_b = 0, bad_blocks = 0, last_raw_bit = 0;
    unsigned short data[4] = {0, 0, 0, 0};
    static constexpr int QSIZE = 100;   // RDS_Q_SIZE (demod_wfm.h:21)
    unsigned short queue[QSIZE][4];
    int head = 0, tail = 0;
    unsigned short last_group[4] = {0, 0, 0, 0};

    // syndrome of the low 26 bits of `shift` against the offset expected for (blk, version); with `fix`, Meggitt burst
    // correction of up to 5 bits applied to `shift` itself (checkBlock :707-758). 0 = block accepted.
    unsigned check(int blk, bool fix) {
        const RdsCode &c = rds_code();
        const unsigned word = shift & 0x3FFFFFFu;
        unsigned syn = word >> 16;                       // the check matrix starts with a 10 x 10 identity
        for (int i = 0; i < 16; ++i) if (word & (0x8000u >> i)) syn ^= c.row[i];
        const int w = blk == 2 && version_b ? 3 : blk == 3 ? 4 : blk;
        syn ^= c.offset[w];
        if (syn && fix) {
            unsigned mask = 1u << 25;
            for (int i = 0; i < 16; ++i, mask >>= 1) {
                if (syn & 0x200) {
                    if ((syn & 0x1F) == 0) { shift ^= mask; syn <<= 1; }   // a burst that fits: flip this bit
                    else syn = (syn << 1) ^ RDS_POLY;
                } else {
                    syn <<= 1;
                }
            }
            syn &= 0x3FF;
        }
        return syn;
    }
    void post(unsigned short a, unsigned short b, unsigned short c, unsigned short d) {
        queue[head][0] = a; queue[head][1] = b; queue[head][2] = c; queue[head][3] = d;
        if (++head >= QSIZE) head = 0;
    }
    void accept_block() {
        data[block] = (unsigned short)(shift >> 10);
        version_b = block == 1 && (data[1] & 0x0800) ? 1 : 0;   // GROUPB_BIT; reset by every other block, as in the reference
    }
    void bit(int b) {
        shift = (shift << 1) | (unsigned)b;
        switch (state) {
        case 0:
            version_b = 0;   // the reference tests BLK_OFFSET_TBL-independent OFFSET_SYNDROME_BLOCK_A here
            block = 0;
            if (check(0, false) == 0) {
                bitpos = 0;
                data[0] = (unsigned short)(shift >> 10);
                block = 1;
                state = 1;
            }
            break;
        case 1:
            if (++bitpos < 26) break;
            bitpos = 0;
            if (check(block, false)) { state = 0; break; }
            accept_block();
            if (block >= 3) { post(data[0], data[1], data[2], data[3]); block = 0; bad_blocks = 0; state = 2; }
            else ++block;
            break;
        case 2:
            if (++bitpos < 26) break;
            bitpos = 0;
            if (check(block, true)) {
                if (++bad_blocks > 5) {           // BLOCK_ERROR_LIMIT: signal lost
                    head = tail = 0;
                    post(0, 0, 0, 0);
                    state = 0;
                } else {
                    if (++block > 3) block = 0;
                    if (block != 0) state = 3;
                }
            } else {
                accept_block();
                if (++block > 3) { post(data[0], data[1], data[2], data[3]); block = 0; bad_blocks = 0; }
            }
            break;
        default:
            if (++bitpos < 26) break;
            bitpos = 0;
            if (++block > 3) { block = 0; state = 2; }
            break;
        }
    }
    void raw_bit(int raw) { bit(raw ^ last_raw_bit); last_raw_bit = raw; }   // differential decode (:343-344)
    // getNextRdsGroupData: pops one group; 1 only when it differs from the last one handed out
    int next(unsigned short *out) {
        if (head == tail) return 0;
        memcpy(out, queue[tail], sizeof(unsigned short) * 4);
        if (++tail >= QSIZE) tail = 0;
        if (memcmp(out, last_group, sizeof(last_group)) != 0) { memcpy(last_group, out, sizeof(last_group)); return 1; }
        return 0;
    }
};
}  // namespace

struct RdsBranch {
    pebble_downconvert_t *dc = nullptr;
    int decim = 1;
    double rate = 0.0;                 // m_RdsOutputRate
    pebble_cfir lp, match;
    pebble_iir_t *sync_bp = nullptr;
    RdsPllConst pllc;
    RdsPllState *d_pll = nullptr; int pll_cur = 0;
    RdsSliceState *d_slice = nullptr; int slice_cur = 0;
    cpx *d_a = nullptr, *d_b = nullptr, *d_c = nullptr, *d_tap = nullptr; size_t cap = 0;
    unsigned char *d_code = nullptr;
    // page-locked ring of sampling decisions, consumed in order on the host
    unsigned char *h_ring = nullptr; size_t ring = 0, wr = 0, rd = 0;
    cudaEvent_t ev = nullptr; bool pending = false;
    int last_len = 0;
    RdsSync sync;
};

static void rds_drain(RdsBranch *r) {
    if (r->pending) { cudaEventSynchronize(r->ev); r->pending = false; }
    for (; r->rd != r->wr; r->rd = (r->rd + 1) % r->ring) {
        const unsigned char c = r->h_ring[r->rd];
        if (c) r->sync.raw_bit(c == 2);
    }
}

RdsBranch *rds_create() {
    RdsBranch *r = new RdsBranch();
    r->dc = pebble_downconvert_create();
    r->sync_bp = pebble_iir_create();
    r->ring = 1 << 16;
    bool ok = r->dc && r->sync_bp && cfir_alloc(&r->lp) == 0 && cfir_alloc(&r->match) == 0 &&
              pebble::cuda_ok(cudaMalloc(&r->d_pll, 2 * sizeof(RdsPllState)), "cudaMalloc") &&
              pebble::cuda_ok(cudaMalloc(&r->d_slice, 2 * sizeof(RdsSliceState)), "cudaMalloc") &&
              pebble::cuda_ok(cudaHostAlloc(&r->h_ring, r->ring, cudaHostAllocDefault), "cudaHostAlloc") &&
              pebble::cuda_ok(cudaEventCreateWithFlags(&r->ev, cudaEventDisableTiming), "cudaEventCreate");
    if (!ok) { rds_destroy(r); return nullptr; }
    return r;
}

void rds_destroy(RdsBranch *r) {
    if (!r) return;
    if (r->dc) pebble_downconvert_destroy(r->dc);
    if (r->sync_bp) pebble_iir_destroy(r->sync_bp);
    cfir_release(&r->lp); cfir_release(&r->match);
    cudaFree(r->d_pll); cudaFree(r->d_slice); cudaFree(r->d_a); cudaFree(r->d_b); cudaFree(r->d_c); cudaFree(r->d_tap); cudaFree(r->d_code);
    if (r->h_ring) cudaFreeHost(r->h_ring);
    if (r->ev) cudaEventDestroy(r->ev);
    delete r;
}

// Demod_WFM::setSampleRate's RDS part (:190-193) + initRds (:489-537)
int rds_setup(RdsBranch *r, double wfm_rate) {
    cudaDeviceSynchronize();   // state rewrites below must not overtake a call in flight
    r->rate = pebble_downconvert_set_data_rate(r->dc, wfm_rate, 8000.0);
    if (!(r->rate > 0.0)) return PEBBLE_ERR_ARG;
    pebble_downconvert_set_frequency(r->dc, -RDS_FREQUENCY);
    r->decim = 1 << pebble_downconvert_num_stages(r->dc);
    r->lp.ntaps = pebble_host::cfir_design_lp(0, 1.0, 40.0, 2400.0, 1.3 * 2400.0, r->rate, r->lp.hI);   // m_RdsBPFilter (:497)
    if (r->lp.ntaps <= 0) return PEBBLE_ERR_ARG;
    memcpy(r->lp.hQ, r->lp.hI, sizeof(double) * r->lp.ntaps);
    if (cfir_upload(&r->lp) < 0) return PEBBLE_ERR_CUDA;
    const double norm = TWOPI / r->rate;
    r->pllc.lo = (0.0 - RDSPLL_RANGE) * norm;                                                            // :501-504
    r->pllc.hi = (0.0 + RDSPLL_RANGE) * norm;
    r->pllc.alpha = 2.0 * RDSPLL_ZETA * RDSPLL_BW * norm;
    r->pllc.beta = (r->pllc.alpha * r->pllc.alpha) / (4.0 * RDSPLL_ZETA * RDSPLL_ZETA);
    // matched filter: one bi-phase symbol, anti-symmetric about its centre (:508-519); 1/x at x = 0 is +inf and the term 0
    const int half = (int)(r->rate / RDS_BITRATE);
    if (2 * half > 75 || half < 1) { pebble::set_error("rds: matched filter of %d taps does not fit CFir (75)", 2 * half); return PEBBLE_ERR_ARG; }
    std::vector<double> co(2 * half + 1);
    for (int i = 0; i <= half; ++i) {
        const double t = (double)i / r->rate, x = t * RDS_BITRATE, x64 = 64.0 * x;
        const double v = .75 * cos(2.0 * TWOPI * x) * ((1.0 / (1.0 / x - x64)) - (1.0 / (9.0 / x - x64)));
        co[i + half] = v;
        co[half - i] = -v;
    }
    r->match.ntaps = 2 * half;
    for (int k = 0; k < 2 * half; ++k) r->match.hI[k] = r->match.hQ[k] = co[k];
    if (cfir_upload(&r->match) < 0) return PEBBLE_ERR_CUDA;
    if (pebble_iir_init(r->sync_bp, PEBBLE_IIR_BP, RDS_BITRATE, 500.0, r->rate) < 0) return PEBBLE_ERR_CUDA;   // :524
    RdsPllState p0 = {0.0, 0.0};
    RdsSliceState s0; memset(&s0, 0, sizeof(s0));
    for (int b = 0; b < 2; ++b) {
        PB_CUDA(cudaMemcpy(r->d_pll + b, &p0, sizeof(p0), cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMemcpy(r->d_slice + b, &s0, sizeof(s0), cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    }
    r->pll_cur = r->slice_cur = 0;
    r->wr = r->rd = 0; r->pending = false; r->last_len = 0;
    r->sync = RdsSync();
    return 0;
}

double rds_rate(const RdsBranch *r) { return r->rate; }

// demod_wfm.cpp:298-357 for one call of n samples of m_CpxRawFm
int rds_process(RdsBranch *r, int n, const cpx *d_cpx_raw_fm, cudaStream_t st) {
    if (n % r->decim) {
        pebble::set_error("demod (dmFMS, RDS on): %d samples per call is not a multiple of the RDS down-converter's decimation %d "
                          "(pebble_demod_set_rds(h, 0) runs the audio path alone)", n, r->decim);
        return PEBBLE_ERR_ARG;
    }
    const size_t need = (size_t)(n / r->decim) + 8;
    if (need > r->cap) {
        cudaFree(r->d_a); cudaFree(r->d_b); cudaFree(r->d_c); cudaFree(r->d_tap); cudaFree(r->d_code);
        r->d_a = r->d_b = r->d_c = r->d_tap = nullptr; r->d_code = nullptr; r->cap = 0;
        PB_CUDA(cudaMalloc(&r->d_a, sizeof(cpx) * need), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMalloc(&r->d_b, sizeof(cpx) * need), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMalloc(&r->d_c, sizeof(cpx) * need), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMalloc(&r->d_tap, sizeof(cpx) * need), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMalloc(&r->d_code, need), PEBBLE_ERR_CUDA);
        r->cap = need;
    }
    const int len = pebble_downconvert_process_device(r->dc, n, d_cpx_raw_fm, (double *)r->d_a, st);     // :298
    if (len < 0) return len;
    r->last_len = len;
    if (len == 0) return 0;
    int e = cfir_run_device(&r->lp, len, r->d_a, r->d_b, st);                                             // :303
    if (e < 0) return e;
    rds_pll_kernel<<<1, 32, 0, st>>>(r->d_b, len, r->pllc, r->d_pll + r->pll_cur, r->d_pll + (r->pll_cur ^ 1), r->d_a);   // :307 -> m_RdsMag
    r->pll_cur ^= 1;
    e = cfir_run_device(&r->match, len, r->d_a, r->d_b, st, true);                                        // :311 -> m_RdsData as (d, d)
    if (e < 0) return e;
    const int blocks = (len + 127) / 128;
    rds_square_kernel<<<blocks, 128, 0, st>>>(r->d_b, len, r->d_a);                                       // :314-315
    pebble::count_launch(2);
    e = pebble_iir_process_device(r->sync_bp, len, (const double *)r->d_a, (double *)r->d_c, st);         // :319
    if (e < 0) return e;
    rds_slice_kernel<<<blocks, 128, 0, st>>>(r->d_b, r->d_c, len, r->d_slice + r->slice_cur, r->d_slice + (r->slice_cur ^ 1), r->d_code, r->d_tap);
    r->slice_cur ^= 1;
    pebble::count_launch();
    PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    // decisions -> host ring (two pieces when it wraps); drain first if this call would overrun what the host has not read
    const size_t used = (r->wr + r->ring - r->rd) % r->ring;
    if (used + (size_t)len >= r->ring / 2) rds_drain(r);
    if ((size_t)len >= r->ring / 2) { pebble::set_error("rds: %d decisions in one call exceed the host ring", len); return PEBBLE_ERR_ARG; }
    const size_t first = (size_t)len < r->ring - r->wr ? (size_t)len : r->ring - r->wr;
    PB_CUDA(cudaMemcpyAsync(r->h_ring + r->wr, r->d_code, first, cudaMemcpyDeviceToHost, st), PEBBLE_ERR_CUDA);
    if (first < (size_t)len) PB_CUDA(cudaMemcpyAsync(r->h_ring, r->d_code + first, (size_t)len - first, cudaMemcpyDeviceToHost, st), PEBBLE_ERR_CUDA);
    r->wr = (r->wr + (size_t)len) % r->ring;
    PB_CUDA(cudaEventRecord(r->ev, st), PEBBLE_ERR_CUDA);
    r->pending = true;
    return len;
}

int rds_next_group(RdsBranch *r, unsigned short *blocks4) {
    rds_drain(r);
    return r->sync.next(blocks4);
}

// test tap: the last call's m_RdsRaw as the reference leaves it (:336-352): (sampled value or 0, matched filter output)
int rds_tap(RdsBranch *r, double *out, int max_samples) {
    const int k = r->last_len < max_samples ? r->last_len : max_samples;
    if (k > 0) PB_CUDA(cudaMemcpy(out, r->d_tap, sizeof(cpx) * (size_t)k, cudaMemcpyDeviceToHost), PEBBLE_ERR_CUDA);
    return k;
}

// ---- the synchroniser alone (host logic; no device involved): tests feed it bit streams directly ------------------------------
struct pebble_rds_sync { RdsSync s; };
extern "C" {
pebble_rds_sync_t *pebble_rds_sync_create(void) { return new pebble_rds_sync(); }
void pebble_rds_sync_destroy(pebble_rds_sync_t *h) { delete h; }
int pebble_rds_sync_push_bits(pebble_rds_sync_t *h, const unsigned char *bits, int n) {
    if (!h || (!bits && n > 0) || n < 0) return PEBBLE_ERR_ARG;
    for (int i = 0; i < n; ++i) h->s.bit(bits[i] & 1);
    return n;
}
int pebble_rds_sync_next_group(pebble_rds_sync_t *h, unsigned short *blocks4) { return h && blocks4 ? h->s.next(blocks4) : 0; }
}
