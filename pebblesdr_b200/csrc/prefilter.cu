// prefilter.cu — the cheap steps either side of the hot path: CIir biquads (pebblelib/iir.cpp) and with them
// DCRemoval (application/dcremoval.cpp), IQBalance (application/iqbalance.cpp:67-88) and the squelch estimate
// SignalStrength::fdEstimate (application/signalstrength.cpp:289-380).
//
// CIir::ProcessFilter (iir.cpp:173-206) is a direct-form-2 biquad, w0 = x - A1 w1 - A2 w2, y = B0 w0 + B1 w1 + B2 w2, a
// LINEAR recurrence in the state s = (w1, w2): s' = M s + (x, 0), M = [[-A1, -A2], [1, 0]]. It runs at the full device rate
// in front of the chain (receiver.cpp:814-816), so it is evaluated as a three-level scan instead of one serial walk:
//   iir_tile_kernel   every thread walks its SEG samples from a zero state; a Hillis-Steele scan over the CTA (with the host's
//                     M^(SEG 2^d) tables) gives each tile's zero-state end state;
//   iir_carry_kernel  one CTA turns the tile end states into each tile's true entry state (same scan, one level up);
//   iir_apply_kernel  every thread re-derives its own entry state and walks its SEG samples again, now writing y.
// Inside a walk the operations and their order are the reference's (separate multiplies and adds), so a segment that enters
// with the reference's state reproduces it bit for bit; the entry states themselves are re-associated sums. For the DC
// blocker (10 Hz at 10 MHz: poles 4e-6 inside the unit circle) the reference itself keeps only ~8 digits — w is ~1e8 x the
// signal and y is its second difference — so two evaluation orders agree to ~1e-8, not 1e-16; the tests say so.
#include "lib_common.h"
#include "common.cuh"
#include <quadmath.h>
#include <cmath>
#include <cstring>
#include <vector>
#include <mutex>

#define IIR_SEG 16
#define IIR_T 256
#define IIR_TILE (IIR_SEG * IIR_T)

struct IirCoef { double a1, a2, b0, b1, b2; };
struct Mat2 { double m00, m01, m10, m11; };
struct IirTables {
    Mat2 seg_pow2[8];     // M^(SEG * 2^d), d = 0..7: scan over the 256 threads of a tile
    Mat2 tile_pow2[10];   // M^(TILE * G * 2^d): scan over the 1024 threads of the carry kernel (G tiles per thread)
    Mat2 tile1;           // M^TILE
};

__device__ __forceinline__ void mat_apply(const Mat2 &m, double &w1, double &w2) {
    const double a = m.m00 * w1 + m.m01 * w2, b = m.m10 * w1 + m.m11 * w2;
    w1 = a; w2 = b;
}
// one sample of iir.cpp:195-198 with the reference's operation order
__device__ __forceinline__ double iir_step(const IirCoef &c, double x, double &w1, double &w2) {
    const double w0 = __dsub_rn(__dsub_rn(x, __dmul_rn(c.a1, w1)), __dmul_rn(c.a2, w2));
    const double y = __dadd_rn(__dadd_rn(__dmul_rn(c.b0, w0), __dmul_rn(c.b1, w1)), __dmul_rn(c.b2, w2));
    w2 = w1; w1 = w0;
    return y;
}

// Zero-state walk of every segment + CTA scan. st4[t] = (w1a, w2a, w1b, w2b) at the END of segment t, from a zero state at the
// tile start. Shared by the tile and the apply kernels.
__device__ __forceinline__ void iir_tile_scan(const IirCoef &c, const IirTables *tb, const cpx (&x)[IIR_SEG], int nvalid, double4 *sh) {
    double w1a = 0, w2a = 0, w1b = 0, w2b = 0;
#pragma unroll
    for (int j = 0; j < IIR_SEG; ++j) {
        if (j < nvalid) { iir_step(c, x[j].x, w1a, w2a); iir_step(c, x[j].y, w1b, w2b); }
        else {   // past the end of the stream: the state just propagates with zero input, keeping the segment lengths uniform
            iir_step(c, 0.0, w1a, w2a); iir_step(c, 0.0, w1b, w2b);
        }
    }
    const int t = threadIdx.x;
    sh[t] = make_double4(w1a, w2a, w1b, w2b);
    __syncthreads();
#pragma unroll
    for (int d = 0; d < 8; ++d) {
        const int off = 1 << d;
        double4 prev = make_double4(0, 0, 0, 0);
        if (t >= off) prev = sh[t - off];
        __syncthreads();
        if (t >= off) {
            mat_apply(tb->seg_pow2[d], prev.x, prev.y);
            mat_apply(tb->seg_pow2[d], prev.z, prev.w);
            double4 cur = sh[t];
            sh[t] = make_double4(cur.x + prev.x, cur.y + prev.y, cur.z + prev.z, cur.w + prev.w);
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(IIR_T) iir_tile_kernel(const cpx *__restrict__ in, long long n, IirCoef c, const IirTables *__restrict__ tb,
                                                         double4 *__restrict__ tile_end) {
    __shared__ double4 sh[IIR_T];
    const long long base = (long long)blockIdx.x * IIR_TILE + (long long)threadIdx.x * IIR_SEG;
    cpx x[IIR_SEG];
    const int nvalid = (int)max(0LL, min((long long)IIR_SEG, n - base));
#pragma unroll
    for (int j = 0; j < IIR_SEG; ++j) x[j] = j < nvalid ? in[base + j] : cpx_make(0.0, 0.0);
    iir_tile_scan(c, tb, x, nvalid, sh);
    if (threadIdx.x == IIR_T - 1) tile_end[blockIdx.x] = sh[IIR_T - 1];
}

// entry[i] = state at the start of tile i; entry[ntiles] = state after the call. One CTA of 1024 threads, G tiles per thread.
__global__ void __launch_bounds__(1024) iir_carry_kernel(const double4 *__restrict__ tile_end, int ntiles, int G, const IirTables *__restrict__ tb,
                                                         const double4 *__restrict__ state_in, double4 *__restrict__ entry) {
    __shared__ double4 sh[1024];
    const int t = threadIdx.x;
    const int i0 = t * G;
    double4 z = make_double4(0, 0, 0, 0);
    for (int g = 0; g < G; ++g) {            // zero-state composition of this thread's G tiles (missing tiles: zero input)
        mat_apply(tb->tile1, z.x, z.y);
        mat_apply(tb->tile1, z.z, z.w);
        if (i0 + g < ntiles) { const double4 e = tile_end[i0 + g]; z.x += e.x; z.y += e.y; z.z += e.z; z.w += e.w; }
    }
    sh[t] = z;
    __syncthreads();
    for (int d = 0; d < 10; ++d) {
        const int off = 1 << d;
        double4 prev = make_double4(0, 0, 0, 0);
        if (t >= off) prev = sh[t - off];
        __syncthreads();
        if (t >= off) {
            mat_apply(tb->tile_pow2[d], prev.x, prev.y);
            mat_apply(tb->tile_pow2[d], prev.z, prev.w);
            const double4 cur = sh[t];
            sh[t] = make_double4(cur.x + prev.x, cur.y + prev.y, cur.z + prev.z, cur.w + prev.w);
        }
        __syncthreads();
    }
    // state at the start of tile i0 = (zero-state prefix of the threads before) + M^(TILE * i0) * state_in: walk it serially
    // from thread to thread instead of forming M^(TILE*i0): s = prefix(t-1) + homogeneous part, the latter by one more scan
    // would need more tables; G-step serial replay per thread is cheap, so carry the homogeneous part the same way:
    double4 s = *state_in;
    // homogeneous part M^(TILE * G * t) * state_in by binary decomposition of t over tile_pow2
    for (int d = 0; d < 10; ++d)
        if (t & (1 << d)) { mat_apply(tb->tile_pow2[d], s.x, s.y); mat_apply(tb->tile_pow2[d], s.z, s.w); }
    if (t > 0) { const double4 p = sh[t - 1]; s.x += p.x; s.y += p.y; s.z += p.z; s.w += p.w; }
    for (int g = 0; g < G; ++g) {
        if (i0 + g <= ntiles) entry[i0 + g] = s;
        if (i0 + g < ntiles) {
            mat_apply(tb->tile1, s.x, s.y);
            mat_apply(tb->tile1, s.z, s.w);
            const double4 e = tile_end[i0 + g];
            s.x += e.x; s.y += e.y; s.z += e.z; s.w += e.w;
        }
    }
}

__global__ void __launch_bounds__(IIR_T) iir_apply_kernel(const cpx *__restrict__ in, cpx *__restrict__ out, long long n, IirCoef c,
                                                          const IirTables *__restrict__ tb, const Mat2 *__restrict__ seg_pow /* M^(SEG t), t < 256 */,
                                                          const double4 *__restrict__ entry, double4 *__restrict__ state_out) {
    __shared__ double4 sh[IIR_T];
    const int t = threadIdx.x;
    const long long base = (long long)blockIdx.x * IIR_TILE + (long long)t * IIR_SEG;
    cpx x[IIR_SEG];
    const int nvalid = (int)max(0LL, min((long long)IIR_SEG, n - base));
#pragma unroll
    for (int j = 0; j < IIR_SEG; ++j) x[j] = j < nvalid ? in[base + j] : cpx_make(0.0, 0.0);
    iir_tile_scan(c, tb, x, nvalid, sh);
    // entry state of this segment = M^(SEG t) * (tile entry state) + zero-state prefix of the segments before it
    double4 s = entry[blockIdx.x];
    const Mat2 p = seg_pow[t];
    mat_apply(p, s.x, s.y);
    mat_apply(p, s.z, s.w);
    if (t > 0) { const double4 z = sh[t - 1]; s.x += z.x; s.y += z.y; s.z += z.z; s.w += z.w; }
#pragma unroll
    for (int j = 0; j < IIR_SEG; ++j) {
        if (j < nvalid) {
            const double yr = iir_step(c, x[j].x, s.x, s.y);
            const double yi = iir_step(c, x[j].y, s.z, s.w);
            out[base + j] = cpx_make(yr, yi);
        }
    }
    if (nvalid > 0 && base + nvalid == n) *state_out = s;   // m_w1a, m_w2a, m_w1b, m_w2b after the call
}

// ---- IQBalance::ProcessBlock (iqbalance.cpp:67-88): t1/t2 restart at 0 in EVERY block of numSamples, so blocks are independent
// serial recurrences: one thread per block (register double buffer as in the demodulator walks).
__global__ void __launch_bounds__(128) iqbalance_kernel(const cpx *__restrict__ in, cpx *__restrict__ out, long long nblocks, int block_len,
                                                        double gain, double phase) {
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nblocks) return;
    const cpx *x = in + b * block_len;
    cpx *y = out + b * block_len;
    const double mu = (double)0.0025f;                     // `float mu = 0.0025`
    const double keep = 1.0 - mu * 0.000001;
    double t2r = 0.0, t2i = 0.0;
    for (int i = 0; i < block_len; ++i) {
        const cpx v = x[i];
        const double orr = __dmul_rn(v.x, gain);                                   // :76
        const double oi = __dadd_rn(v.y, __dmul_rn(v.x, phase));                   // :78
        // t1 = out + t2 * conj(out)  (std::complex product: ac - bd, ad + bc with b -> -b)
        const double pr = __dsub_rn(__dmul_rn(t2r, orr), __dmul_rn(t2i, -oi));
        const double pi = __dadd_rn(__dmul_rn(t2r, -oi), __dmul_rn(t2i, orr));
        const double t1r = __dadd_rn(orr, pr), t1i = __dadd_rn(oi, pi);
        // t2 = scaleCpx(t2, 1 - mu*1e-6) - scaleCpx(t1*t1, mu)
        const double sr = __dsub_rn(__dmul_rn(t1r, t1r), __dmul_rn(t1i, t1i));
        const double si = __dadd_rn(__dmul_rn(t1r, t1i), __dmul_rn(t1i, t1r));
        t2r = __dsub_rn(__dmul_rn(t2r, keep), __dmul_rn(sr, mu));
        t2i = __dsub_rn(__dmul_rn(t2i, keep), __dmul_rn(si, mu));
        y[i] = cpx_make(t1r, t1i);
    }
}

// ---- SignalStrength::fdEstimate (signalstrength.cpp:289-380): sums of dBToPower over the band-pass bins and over one band-pass
// width either side. One CTA per channel; fixed-order tree reduction (deterministic).
struct FdBins { int bp_lo, bp_hi, noise_lo, noise_hi; };
__global__ void __launch_bounds__(256) fd_estimate_kernel(const double *__restrict__ spectrum, int bins, const FdBins *__restrict__ fb,
                                                          double4 *__restrict__ res /* bpTotal, bpPeak, noiseTotal, noiseBins */) {
    __shared__ double s_tot[256], s_pk[256], s_nt[256], s_nb[256];
    const FdBins f = fb[blockIdx.x];
    double tot = 0, pk = 0, nt = 0, nb = 0;
    for (int i = f.noise_lo + threadIdx.x; i < bins && i <= f.noise_hi; i += 256) {    // the loop leaves at i >= noiseHighBin (:349-350)
        const double pwr = pow(10.0, spectrum[i] / 10.0);                                // DB::dBToPower
        if (i >= f.bp_lo && i <= f.bp_hi) { tot += pwr; pk = pwr > pk ? pwr : pk; }
        else { nt += pwr; nb += 1.0; }
    }
    s_tot[threadIdx.x] = tot; s_pk[threadIdx.x] = pk; s_nt[threadIdx.x] = nt; s_nb[threadIdx.x] = nb;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) {
            s_tot[threadIdx.x] += s_tot[threadIdx.x + o];
            s_pk[threadIdx.x] = fmax(s_pk[threadIdx.x], s_pk[threadIdx.x + o]);
            s_nt[threadIdx.x] += s_nt[threadIdx.x + o];
            s_nb[threadIdx.x] += s_nb[threadIdx.x + o];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) res[blockIdx.x] = make_double4(s_tot[0], s_pk[0], s_nt[0], s_nb[0]);
}

// ---------------------------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------------------------
struct QMat { __float128 a, b, c, d; };
static QMat qmul(const QMat &x, const QMat &y) { return {x.a * y.a + x.b * y.c, x.a * y.b + x.b * y.d, x.c * y.a + x.d * y.c, x.c * y.b + x.d * y.d}; }
static QMat qpow(QMat m, unsigned long long k) {
    QMat r = {1, 0, 0, 1};
    while (k) { if (k & 1) r = qmul(r, m); m = qmul(m, m); k >>= 1; }
    return r;
}
static Mat2 qround(const QMat &m) { return {(double)m.a, (double)m.b, (double)m.c, (double)m.d}; }

struct pebble_iir {
    IirCoef c;
    double4 state = {0, 0, 0, 0};              // host mirror is NOT kept in sync; device state is authoritative
    double4 *d_state[2] = {nullptr, nullptr};
    int cur = 0;
    IirTables tb; int tb_G = -1;
    IirTables *d_tb = nullptr;
    Mat2 *d_segpow = nullptr;
    double4 *d_tile_end = nullptr, *d_entry = nullptr; size_t tiles_cap = 0;
    cudaStream_t stream = nullptr;
    pebble::Staging stg;
};

static int iir_upload_tables(pebble_iir *h, int G, cudaStream_t st) {
    const QMat M = {-(__float128)h->c.a1, -(__float128)h->c.a2, 1, 0};
    if (h->tb_G < 0) {
        for (int d = 0; d < 8; ++d) h->tb.seg_pow2[d] = qround(qpow(M, (unsigned long long)IIR_SEG << d));
        h->tb.tile1 = qround(qpow(M, IIR_TILE));
        std::vector<Mat2> sp(IIR_T);
        QMat step = qpow(M, IIR_SEG), acc = {1, 0, 0, 1};
        for (int t = 0; t < IIR_T; ++t) { sp[t] = qround(acc); acc = qmul(acc, step); }
        PB_CUDA(cudaMemcpyAsync(h->d_segpow, sp.data(), sizeof(Mat2) * IIR_T, cudaMemcpyHostToDevice, st), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaStreamSynchronize(st), PEBBLE_ERR_CUDA);   // sp is a local
    }
    if (h->tb_G != G) {
        QMat p = qpow(M, (unsigned long long)IIR_TILE * (unsigned long long)G);
        for (int d = 0; d < 10; ++d) { h->tb.tile_pow2[d] = qround(p); p = qmul(p, p); }
        h->tb_G = G;
        PB_CUDA(cudaMemcpyAsync(h->d_tb, &h->tb, sizeof(IirTables), cudaMemcpyHostToDevice, st), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaStreamSynchronize(st), PEBBLE_ERR_CUDA);
    }
    return 0;
}

static int iir_run_device(pebble_iir *h, long long n, const cpx *d_in, cpx *d_out, cudaStream_t st) {
    if (n <= 0) return 0;
    const int ntiles = (int)((n + IIR_TILE - 1) / IIR_TILE);
    const int G = (ntiles + 1023) / 1024;
    if ((size_t)ntiles + 1 > h->tiles_cap) {
        PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
        cudaFree(h->d_tile_end); cudaFree(h->d_entry); h->d_tile_end = h->d_entry = nullptr; h->tiles_cap = 0;
        PB_CUDA(cudaMalloc(&h->d_tile_end, sizeof(double4) * ((size_t)ntiles + 1)), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMalloc(&h->d_entry, sizeof(double4) * ((size_t)ntiles + 1)), PEBBLE_ERR_CUDA);
        h->tiles_cap = (size_t)ntiles + 1;
    }
    int r = iir_upload_tables(h, G, st);
    if (r < 0) return r;
    iir_tile_kernel<<<ntiles, IIR_T, 0, st>>>(d_in, n, h->c, h->d_tb, h->d_tile_end);
    iir_carry_kernel<<<1, 1024, 0, st>>>(h->d_tile_end, ntiles, G, h->d_tb, h->d_state[h->cur], h->d_entry);
    iir_apply_kernel<<<ntiles, IIR_T, 0, st>>>(d_in, d_out, n, h->c, h->d_tb, h->d_segpow, h->d_entry, h->d_state[h->cur ^ 1]);
    pebble::count_launch(3);
    PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    h->cur ^= 1;
    return (int)n;
}

extern "C" {

// ---- CIir (pebblelib/iir.h:17-45) -----------------------------------------------------------------------------------------
pebble_iir_t *pebble_iir_create(void) {
    pebble_iir *h = new pebble_iir();
    bool ok = pebble::cuda_ok(cudaMalloc(&h->d_tb, sizeof(IirTables)), "cudaMalloc") &&
              pebble::cuda_ok(cudaMalloc(&h->d_segpow, sizeof(Mat2) * IIR_T), "cudaMalloc") &&
              pebble::cuda_ok(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "cudaStreamCreate");
    for (int b = 0; ok && b < 2; ++b)
        ok = pebble::cuda_ok(cudaMalloc(&h->d_state[b], sizeof(double4)), "cudaMalloc") &&
             pebble::cuda_ok(cudaMemset(h->d_state[b], 0, sizeof(double4)), "cudaMemset");
    if (!ok) { pebble_iir_destroy(h); return nullptr; }
    pebble_iir_init(h, PEBBLE_IIR_BR, 25000, 1000.0, 100000);   // CIir::CIir (iir.cpp:45-48)
    return h;
}
void pebble_iir_destroy(pebble_iir_t *h) {
    if (!h) return;
    cudaFree(h->d_tb); cudaFree(h->d_segpow); cudaFree(h->d_state[0]); cudaFree(h->d_state[1]); cudaFree(h->d_tile_end); cudaFree(h->d_entry);
    if (h->stream) cudaStreamDestroy(h->stream);
    h->stg.release();
    delete h;
}
// CIir::InitLP / InitHP / InitBP / InitBR (iir.cpp:50-170): coefficients scaled by 1/A0, the four delay values cleared
int pebble_iir_init(pebble_iir_t *h, int kind, double f0, double q, double fs) {
    if (!h || kind < 0 || kind > 3) return PEBBLE_ERR_ARG;
    const double TWOPI = 6.28318530717958647692528676656;
    const double w0 = TWOPI * f0 / fs;
    const double alpha = sin(w0) / (2.0 * q);
    const double A = 1.0 / (1.0 + alpha);
    IirCoef &c = h->c;
    if (kind == PEBBLE_IIR_LP) { c.b0 = A * ((1.0 - cos(w0)) / 2.0); c.b1 = A * (1.0 - cos(w0)); c.b2 = A * ((1.0 - cos(w0)) / 2.0); }
    else if (kind == PEBBLE_IIR_HP) { c.b0 = A * ((1.0 + cos(w0)) / 2.0); c.b1 = -A * (1.0 + cos(w0)); c.b2 = A * ((1.0 + cos(w0)) / 2.0); }
    else if (kind == PEBBLE_IIR_BP) { c.b0 = A * alpha; c.b1 = 0.0; c.b2 = A * -alpha; }
    else { c.b0 = A * 1.0; c.b1 = A * (-2.0 * cos(w0)); c.b2 = A * 1.0; }
    c.a1 = A * (-2.0 * cos(w0));
    c.a2 = A * (1.0 - alpha);
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    for (int b = 0; b < 2; ++b) PB_CUDA(cudaMemset(h->d_state[b], 0, sizeof(double4)), PEBBLE_ERR_CUDA);
    h->tb_G = -1;
    return 0;
}
// Raw biquad (iir.h keeps m_A1, m_A2, m_B0, m_B1, m_B2 public): w0 = x - A1 w1 - A2 w2, y = B0 w0 + B1 w1 + B2 w2. State restarts at 0.
int pebble_iir_set_coefficients(pebble_iir_t *h, const double *a1a2b0b1b2) {
    if (!h || !a1a2b0b1b2) return PEBBLE_ERR_ARG;
    h->c.a1 = a1a2b0b1b2[0]; h->c.a2 = a1a2b0b1b2[1]; h->c.b0 = a1a2b0b1b2[2]; h->c.b1 = a1a2b0b1b2[3]; h->c.b2 = a1a2b0b1b2[4];
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    for (int b = 0; b < 2; ++b) PB_CUDA(cudaMemset(h->d_state[b], 0, sizeof(double4)), PEBBLE_ERR_CUDA);
    h->tb_G = -1;
    return 0;
}
int pebble_iir_get_coefficients(const pebble_iir_t *h, double *a1a2b0b1b2) {
    if (!h || !a1a2b0b1b2) return PEBBLE_ERR_ARG;
    a1a2b0b1b2[0] = h->c.a1; a1a2b0b1b2[1] = h->c.a2; a1a2b0b1b2[2] = h->c.b0; a1a2b0b1b2[3] = h->c.b1; a1a2b0b1b2[4] = h->c.b2;
    return 0;
}
int pebble_iir_process_device(pebble_iir_t *h, int n, const double *d_in, double *d_out, void *stream) {
    if (!h || n < 0 || (n && (!d_in || !d_out))) { pebble::set_error("iir: bad argument"); return PEBBLE_ERR_ARG; }
    return iir_run_device(h, n, (const cpx *)d_in, (cpx *)d_out, (cudaStream_t)stream);
}
int pebble_iir_process(pebble_iir_t *h, int n, const double *in, double *out) {
    if (!h || n < 0 || (n && (!in || !out))) { pebble::set_error("iir: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    const size_t b = sizeof(cpx) * (size_t)n;
    if (!h->stg.reserve(b, b)) return PEBBLE_ERR_CUDA;
    memcpy(h->stg.h_in, in, b);
    PB_CUDA(cudaMemcpyAsync(h->stg.d_in, h->stg.h_in, b, cudaMemcpyHostToDevice, h->stream), PEBBLE_ERR_CUDA);
    const int r = iir_run_device(h, n, (const cpx *)h->stg.d_in, (cpx *)h->stg.d_out, h->stream);
    if (r < 0) return r;
    PB_CUDA(cudaMemcpyAsync(h->stg.h_out, h->stg.d_out, b, cudaMemcpyDeviceToHost, h->stream), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamSynchronize(h->stream), PEBBLE_ERR_CUDA);
    memcpy(out, h->stg.h_out, b);
    return r;
}
// m_w1a, m_w2a, m_w1b, m_w2b
int pebble_iir_get_state(pebble_iir_t *h, double *w4) {
    if (!h || !w4) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(w4, h->d_state[h->cur], sizeof(double4), cudaMemcpyDeviceToHost), PEBBLE_ERR_CUDA);
    return 0;
}
int pebble_iir_set_state(pebble_iir_t *h, const double *w4) {
    if (!h || !w4) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(h->d_state[h->cur], w4, sizeof(double4), cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    return 0;
}

// ---- IQBalance (application/iqbalance.h) ----------------------------------------------------------------------------------------
// n must be a multiple of block_len (= the step's numSamples): every block restarts the correction recurrence (iqbalance.cpp:72-73)
int pebble_iqbalance_process_device(int n, int block_len, double gain_factor, double phase_factor, const double *d_in, double *d_out,
                                    void *stream) {
    if (n < 0 || block_len <= 0 || n % block_len || (n && (!d_in || !d_out))) { pebble::set_error("iqbalance: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    const long long nb = n / block_len;
    iqbalance_kernel<<<(unsigned)((nb + 127) / 128), 128, 0, (cudaStream_t)stream>>>((const cpx *)d_in, (cpx *)d_out, nb, block_len, gain_factor,
                                                                                    phase_factor);
    pebble::count_launch();
    PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    return n;
}

// ---- SignalStrength::fdEstimate (application/signalstrength.cpp:289-380), one or many channels over ONE spectrum ---------------------
// d_spectrum: `bins` amplitude-dB values (-f..0..+f) on the device. Channel k: band-pass (lo, hi) around mixer[k]. out5[k] =
// {peakDb, avgDb, snrDb, floorDb, open} where open = avgDb >= squelch_db (receiver.cpp:963-966 drops the frame otherwise).
int pebble_sigstrength_fd_estimate_device(const double *d_spectrum, int bins, unsigned spectrum_rate, float bp_lo, float bp_hi,
                                          const double *mixer_freqs, int nch, double squelch_db, double *out5, void *stream_v) {
    if (!d_spectrum || bins <= 0 || !mixer_freqs || nch <= 0 || !out5) { pebble::set_error("sigstrength: bad argument"); return PEBBLE_ERR_ARG; }
    cudaStream_t st = (cudaStream_t)stream_v;
    std::vector<FdBins> fb(nch);
    std::vector<int> bpbins(nch);
    const double binWidth = (double)(spectrum_rate / (unsigned)bins);    // integer division, as the reference writes it (:311)
    const int centerBin = bins / 2;
    auto bound = [bins](int v) { return v < 0 ? 0 : (v > bins ? bins : v); };
    for (int k = 0; k < nch; ++k) {
        int mixerBin = (int)(centerBin + (mixer_freqs[k] / binWidth));   // :324
        mixerBin = bound(mixerBin);
        int lo = (int)(mixerBin + ((double)bp_lo / binWidth)); lo = bound(lo);
        int hi = (int)(mixerBin + ((double)bp_hi / binWidth)); hi = bound(hi);
        const int bpBins = hi - lo;
        fb[k].bp_lo = lo; fb[k].bp_hi = hi;
        fb[k].noise_lo = bound(lo - bpBins); fb[k].noise_hi = bound(hi + bpBins);
        bpbins[k] = bpBins;
    }
    FdBins *d_fb = nullptr; double4 *d_res = nullptr;
    PB_CUDA(cudaMalloc(&d_fb, sizeof(FdBins) * nch), PEBBLE_ERR_CUDA);
    if (!pebble::cuda_ok(cudaMalloc(&d_res, sizeof(double4) * nch), "cudaMalloc")) { cudaFree(d_fb); return PEBBLE_ERR_CUDA; }
    std::vector<double4> res(nch);
    bool ok = pebble::cuda_ok(cudaMemcpyAsync(d_fb, fb.data(), sizeof(FdBins) * nch, cudaMemcpyHostToDevice, st), "cudaMemcpy");
    if (ok) {
        fd_estimate_kernel<<<nch, 256, 0, st>>>(d_spectrum, bins, d_fb, d_res);
        pebble::count_launch();
        ok = pebble::cuda_ok(cudaMemcpyAsync(res.data(), d_res, sizeof(double4) * nch, cudaMemcpyDeviceToHost, st), "cudaMemcpy") &&
             pebble::cuda_ok(cudaStreamSynchronize(st), "cudaStreamSynchronize");
    }
    cudaFree(d_fb); cudaFree(d_res);
    if (!ok) return PEBBLE_ERR_CUDA;
    auto p2db = [](double p) { return p == 0 ? -120.0 : 10 * log10(p); };
    auto clip = [](double d) { return d < -120.0 ? -120.0 : (d > 0.0 ? 0.0 : d); };
    for (int k = 0; k < nch; ++k) {
        const double bpAvg = res[k].x / bpbins[k], noiseAvg = res[k].z / res[k].w;     // :352-353 (0/0 -> NaN there too)
        double *o = out5 + 5 * k;
        o[0] = clip(p2db(res[k].y));
        o[1] = clip(p2db(bpAvg));
        o[3] = clip(p2db(noiseAvg));
        double snr = (noiseAvg == 0 || res[k].y == 0) ? -120.0 : 10.0 * log10(res[k].y / noiseAvg);   // DB::powerRatioToDb
        o[2] = snr < 0.0 ? 0.0 : (snr > 120.0 ? 120.0 : snr);
        o[4] = o[1] < squelch_db ? 0.0 : 1.0;
    }
    return nch;
}


// ---- the same estimate for callers that may not allocate -------------------------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc (code
// under CUDA graph capture, tight serving loops, harnesses that forbid it). pebble_sigstrength_fd_estimate_device() above takes its
// two scratch arrays (the per-channel bin ranges going in, the per-channel sums coming out) from cudaMalloc on every call; this
// variant keeps them in ONE file-scope device array of PEBBLE_FD_NOALLOC_MAX_CH entries (48 KiB, held by every GPU that loads
// the library until the process exits) and its host-side temporaries on the stack. It runs the same kernel on the same numbers.
//   * g_fd_scratch as written in HOST code is a host-side stand-in for the symbol, not a device address; the device address is
//     asked for once per device, with that device current (cudaGetSymbolAddress), and that pointer is what the kernel is given.
//   * Every call uses the same array: two calls whose GPU work overlaps (different streams, different host threads) race on
//     it. Calls on one stream, or separated by a synchronisation, are safe; this entry point synchronises `stream` before
//     it returns (it has to read the sums back), so back-to-back calls from one thread are.
//   * The array keeps the previous call's contents. Nothing in it has to start at zero: the nch FdBins entries are copied in
//     and each of the nch result entries is written in full by its own CTA before any of them is read back; entries past nch
//     are never read.
static __device__ __align__(256) unsigned char g_fd_scratch[PEBBLE_FD_NOALLOC_MAX_CH * (sizeof(double4) + sizeof(FdBins))];
#define PB_FD_MAX_DEVICES 64
static void *g_fd_scratch_dev[PB_FD_MAX_DEVICES];     // device address of g_fd_scratch per device, filled on first use
static std::mutex g_fd_mu;

static void *fd_scratch_on_current_device() {
    int dev = -1;
    if (!pebble::cuda_ok(cudaGetDevice(&dev), "cudaGetDevice")) return nullptr;
    if (dev < 0 || dev >= PB_FD_MAX_DEVICES) { pebble::set_error("sigstrength(noalloc): device ordinal %d out of range", dev); return nullptr; }
    std::lock_guard<std::mutex> lk(g_fd_mu);
    if (!g_fd_scratch_dev[dev]) {
        void *p = nullptr;
        if (!pebble::cuda_ok(cudaGetSymbolAddress(&p, g_fd_scratch), "cudaGetSymbolAddress(g_fd_scratch)") || !p) return nullptr;
        g_fd_scratch_dev[dev] = p;
    }
    return g_fd_scratch_dev[dev];
}

int pebble_sigstrength_fd_estimate_noalloc_max_channels(void) { return PEBBLE_FD_NOALLOC_MAX_CH; }

int pebble_sigstrength_fd_estimate_noalloc_device(const double *d_spectrum, int bins, unsigned spectrum_rate, float bp_lo, float bp_hi,
                                                  const double *mixer_freqs, int nch, double squelch_db, double *out5, void *stream_v) {
    if (!d_spectrum || bins <= 0 || !mixer_freqs || nch <= 0 || !out5) { pebble::set_error("sigstrength(noalloc): bad argument"); return PEBBLE_ERR_ARG; }
    if (nch > PEBBLE_FD_NOALLOC_MAX_CH) {
        pebble::set_error("sigstrength(noalloc): %d channels exceed the static scratch array's %d (PEBBLE_FD_NOALLOC_MAX_CH); split the call or "
                          "use pebble_sigstrength_fd_estimate_device", nch, PEBBLE_FD_NOALLOC_MAX_CH);
        return PEBBLE_ERR_ARG;
    }
    unsigned char *scratch = (unsigned char *)fd_scratch_on_current_device();
    if (!scratch) return PEBBLE_ERR_CUDA;
    double4 *d_res = (double4 *)scratch;
    FdBins *d_fb = (FdBins *)(scratch + sizeof(double4) * PEBBLE_FD_NOALLOC_MAX_CH);
    cudaStream_t st = (cudaStream_t)stream_v;
    FdBins fb[PEBBLE_FD_NOALLOC_MAX_CH];          // 16 KiB + 4 KiB + 32 KiB of stack: no heap either
    int bpbins[PEBBLE_FD_NOALLOC_MAX_CH];
    double4 res[PEBBLE_FD_NOALLOC_MAX_CH];
    const double binWidth = (double)(spectrum_rate / (unsigned)bins);    // signalstrength.cpp:311, integer division
    const int centerBin = bins / 2;
    for (int k = 0; k < nch; ++k) {                                       // :324-345, as in the allocating entry point
        int mixerBin = (int)(centerBin + (mixer_freqs[k] / binWidth));
        mixerBin = mixerBin < 0 ? 0 : (mixerBin > bins ? bins : mixerBin);
        int lo = (int)(mixerBin + ((double)bp_lo / binWidth)); lo = lo < 0 ? 0 : (lo > bins ? bins : lo);
        int hi = (int)(mixerBin + ((double)bp_hi / binWidth)); hi = hi < 0 ? 0 : (hi > bins ? bins : hi);
        const int bpBins = hi - lo;
        int nlo = lo - bpBins, nhi = hi + bpBins;
        nlo = nlo < 0 ? 0 : (nlo > bins ? bins : nlo);
        nhi = nhi < 0 ? 0 : (nhi > bins ? bins : nhi);
        fb[k].bp_lo = lo; fb[k].bp_hi = hi; fb[k].noise_lo = nlo; fb[k].noise_hi = nhi;
        bpbins[k] = bpBins;
    }
    PB_CUDA(cudaMemcpyAsync(d_fb, fb, sizeof(FdBins) * nch, cudaMemcpyHostToDevice, st), PEBBLE_ERR_CUDA);
    fd_estimate_kernel<<<nch, 256, 0, st>>>(d_spectrum, bins, d_fb, d_res);
    pebble::count_launch();
    PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpyAsync(res, d_res, sizeof(double4) * nch, cudaMemcpyDeviceToHost, st), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamSynchronize(st), PEBBLE_ERR_CUDA);
    for (int k = 0; k < nch; ++k) {
        const double bpAvg = res[k].x / bpbins[k], noiseAvg = res[k].z / res[k].w;     // :352-353
        double *o = out5 + 5 * k;
        const double pk = res[k].y == 0 ? -120.0 : 10 * log10(res[k].y);
        const double av = bpAvg == 0 ? -120.0 : 10 * log10(bpAvg);
        const double fl = noiseAvg == 0 ? -120.0 : 10 * log10(noiseAvg);
        o[0] = pk < -120.0 ? -120.0 : (pk > 0.0 ? 0.0 : pk);
        o[1] = av < -120.0 ? -120.0 : (av > 0.0 ? 0.0 : av);
        o[3] = fl < -120.0 ? -120.0 : (fl > 0.0 ? 0.0 : fl);
        const double snr = (noiseAvg == 0 || res[k].y == 0) ? -120.0 : 10.0 * log10(res[k].y / noiseAvg);
        o[2] = snr < 0.0 ? 0.0 : (snr > 120.0 ? 120.0 : snr);
        o[4] = o[1] < squelch_db ? 0.0 : 1.0;
    }
    return nch;
}

}  // extern "C"
