// demod.cu — the demodulator tail of the receiver chain on the GPU + C ABI:
//   CFir (pebblelib/fir.cpp:106-132), AGC (application/agc.cpp:84-207), Demod_AM::processBlockFiltered
//   (demod/demod_am.cpp:40-64), Demod_NFM::processBlockNCO / processBlockFM2 (demod/demod_nfm.cpp:124-140,
//   225-257), Demod::processBlock dispatch (demod.cpp:100-141: USB/LSB/... return the input untouched).
// Parallel where the math allows it (FIR, envelope, phase difference; the AM DC-block IIR is a linear
// recurrence and runs as a 3-kernel affine scan); the PLL and the full AGC are true nonlinear
// recurrences and run one thread per stream (parallel only across streams/channels, finding F10).
#include "lib_common.h"
#include "common.cuh"
#include "host_design.h"
#include <cmath>
#include <cstring>
#include <vector>

#define PB_FIR_MAX 75
#define PB_SCAN_THREADS 256
#define PB_SCAN_PER 8
#define PB_SCAN_TILE (PB_SCAN_THREADS * PB_SCAN_PER)

// ---- CFir ---------------------------------------------------------------------------------------
// virtual stream = hist (ntaps-1 samples) ++ in; out[i] = sum_k h[k] * xv[i + ntaps-1 - k]
__global__ void cfir_kernel(const cpx *__restrict__ hist, const cpx *__restrict__ in, int n, const double *__restrict__ hI,
                            const double *__restrict__ hQ, int ntaps, cpx *__restrict__ out) {
    __shared__ double sI[PB_FIR_MAX], sQ[PB_FIR_MAX];
    for (int k = threadIdx.x; k < ntaps; k += blockDim.x) { sI[k] = hI[k]; sQ[k] = hQ[k]; }
    __syncthreads();
    const int h = ntaps - 1;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        double re = 0.0, im = 0.0;
        for (int k = 0; k < ntaps; ++k) {
            const int v = i + h - k;  // index into the virtual stream
            const cpx x = v < h ? hist[v] : in[v - h];
            re += sI[k] * x.x;
            im += sQ[k] * x.y;
        }
        out[i] = cpx_make(re, im);
    }
}
__global__ void cfir_state_kernel(const cpx *__restrict__ hist, const cpx *__restrict__ in, int n, int h,
                                  cpx *__restrict__ new_hist) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < h) {
        const int v = n + i;  // last h samples of hist ++ in
        new_hist[i] = v < h ? hist[v] : in[v - h];
    }
}

// ---- affine scan helpers (d <- A*d + B composition) ---------------------------------------------------
struct Aff { double a, b; };
__device__ __forceinline__ Aff aff_then(Aff first, Aff second) {  // apply `first`, then `second`
    Aff r;
    r.a = first.a * second.a;
    r.b = second.a * first.b + second.b;
    return r;
}
// Exclusive scan of per-thread maps over the CTA (in thread order). Returns the composition of all
// lower-numbered threads' maps; *total (valid in every thread) = composition of all threads' maps.
__device__ __forceinline__ Aff block_aff_exclusive(Aff mine, Aff *total) {
    __shared__ Aff warp_tot[PB_SCAN_THREADS / 32];
    __shared__ Aff warp_pre[PB_SCAN_THREADS / 32 + 1];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    Aff inc = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        Aff o;
        o.a = __shfl_up_sync(0xffffffffu, inc.a, d);
        o.b = __shfl_up_sync(0xffffffffu, inc.b, d);
        if (lane >= d) inc = aff_then(o, inc);
    }
    if (lane == 31) warp_tot[w] = inc;
    __syncthreads();
    if (threadIdx.x == 0) {
        Aff run; run.a = 1.0; run.b = 0.0;
        for (int i = 0; i < PB_SCAN_THREADS / 32; ++i) { warp_pre[i] = run; run = aff_then(run, warp_tot[i]); }
        warp_pre[PB_SCAN_THREADS / 32] = run;
    }
    __syncthreads();
    Aff ex;
    ex.a = __shfl_up_sync(0xffffffffu, inc.a, 1);
    ex.b = __shfl_up_sync(0xffffffffu, inc.b, 1);
    if (lane == 0) { ex.a = 1.0; ex.b = 0.0; }
    *total = warp_pre[PB_SCAN_THREADS / 32];
    return aff_then(warp_pre[w], ex);
}

// ---- Demod_AM::processBlockFiltered, demod_am.cpp:40-59 ------------------------------------------------
// d_n = alpha*d_{n-1} + |x_n| ; y_n = d_n - d_{n-1} ; out = (y, y)
// pass 1: per-tile aggregate map; pass 2: serial carry over tiles; pass 3: apply.
__global__ void __launch_bounds__(PB_SCAN_THREADS) am_tile_kernel(const cpx *__restrict__ in, int n, double alpha,
                                                                  Aff *__restrict__ tile_map) {
    const int base = blockIdx.x * PB_SCAN_TILE + threadIdx.x * PB_SCAN_PER;
    Aff m; m.a = 1.0; m.b = 0.0;
#pragma unroll
    for (int k = 0; k < PB_SCAN_PER; ++k) {
        if (base + k < n) {
            const cpx x = in[base + k];
            m.b = alpha * m.b + sqrt(x.x * x.x + x.y * x.y);  // DB::amplitude, db.h:28-30
            m.a *= alpha;
        }
    }
    Aff tot;
    block_aff_exclusive(m, &tot);
    if (threadIdx.x == 0) tile_map[blockIdx.x] = tot;
}
__global__ void am_carry_kernel(const Aff *__restrict__ tile_map, int ntiles, const double *__restrict__ d_last_in,
                                double *__restrict__ carry) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        double d = *d_last_in;
        for (int t = 0; t < ntiles; ++t) {
            carry[t] = d;
            d = tile_map[t].a * d + tile_map[t].b;
        }
    }
}
__global__ void __launch_bounds__(PB_SCAN_THREADS) am_apply_kernel(const cpx *__restrict__ in, int n, double alpha,
                                                                   const double *__restrict__ carry,
                                                                   cpx *__restrict__ out, double *__restrict__ d_last_out) {
    const int base = blockIdx.x * PB_SCAN_TILE + threadIdx.x * PB_SCAN_PER;
    double mag[PB_SCAN_PER];
    Aff m; m.a = 1.0; m.b = 0.0;
#pragma unroll
    for (int k = 0; k < PB_SCAN_PER; ++k) {
        mag[k] = 0.0;
        if (base + k < n) {
            const cpx x = in[base + k];
            mag[k] = sqrt(x.x * x.x + x.y * x.y);
            m.b = alpha * m.b + mag[k];
            m.a *= alpha;
        }
    }
    Aff tot;
    const Aff ex = block_aff_exclusive(m, &tot);
    double d = ex.a * carry[blockIdx.x] + ex.b;  // m_amDcLast entering this thread's first sample
#pragma unroll
    for (int k = 0; k < PB_SCAN_PER; ++k) {
        if (base + k < n) {
            const double dn = (alpha * d) + mag[k];  // demod_am.cpp:52
            const double y = dn - d;                 // demod_am.cpp:53
            d = dn;
            out[base + k] = cpx_make(y, y);
            if (base + k == n - 1) *d_last_out = d;
        }
    }
}

// ---- AGC_OFF / gain: out = g * in (agc.cpp:86-94, receiver.cpp:935-938) ---------------------------------
__global__ void scale_kernel(const cpx *__restrict__ in, int n, double g, cpx *__restrict__ out) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[i] = cpx_scale(in[i], g);
}

// ---- Demod_NFM::processBlockFM2, demod_nfm.cpp:124-140 (phase difference) --------------------------------
__device__ __forceinline__ double phase_cpx(cpx c) {  // CpxUtil::phaseCpx, cpx.cpp:5-21
    double tmp = atan(c.y / ((c.x == 0) ? 1e-200 : c.x));
    if (c.x < 0 && c.y < 0) tmp -= 3.14159265358979323846;
    else if (c.x < 0 && c.y >= 0) tmp += 3.14159265358979323846;
    return tmp;
}
__global__ void fm2_kernel(const cpx *__restrict__ in, int n, const cpx *__restrict__ last_in, cpx *__restrict__ out,
                           cpx *__restrict__ last_out) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const cpx prev = i ? in[i - 1] : *last_in;
        const cpx prod = cpx_mulc(in[i], prev);
        const double v = phase_cpx(prod) * .0005;
        out[i] = cpx_make(v, v);
        if (i == n - 1) *last_out = in[i];
    }
}

// ---- Demod_NFM::processBlockNCO, demod_nfm.cpp:225-257 (PLL; state and constants are `float`) ----------
struct PllConst { float alpha, beta, dc_alpha, lo, hi, out_gain; };
struct PllState { float phase, freq, dc; };
// one thread per stream; stream s occupies in[s*stride .. s*stride + n)
__global__ void pll_kernel(const cpx *__restrict__ in, int n, long long stride, int nstreams, PllConst c,
                           PllState *__restrict__ st, cpx *__restrict__ out) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= nstreams) return;
    PllState v = st[s];
    const cpx *x = in + (long long)s * stride;
    cpx *y = out + (long long)s * stride;
    for (int i = 0; i < n; ++i) {
        // sin/cos of a float resolve to the float overloads in the reference build
        const double ncoSin = (double)sinf(v.phase);
        const double ncoCos = (double)cosf(v.phase);
        const cpx xi = x[i];
        const double tr = ncoCos * xi.x - ncoSin * xi.y;
        const double ti = ncoCos * xi.y + ncoSin * xi.x;
        const double phzerror = -atan2(ti, tr);
        v.freq = (float)((double)v.freq + ((double)c.beta * phzerror));
        if (v.freq > c.hi) v.freq = c.hi;
        else if (v.freq < c.lo) v.freq = c.lo;
        v.phase = (float)((double)v.phase + ((double)v.freq + (double)c.alpha * phzerror));
        v.dc = (float)((1.0 - (double)c.dc_alpha) * (double)v.dc + (double)c.dc_alpha * (double)v.freq);
        const float o = (v.freq - v.dc) * c.out_gain;
        y[i] = cpx_make((double)o, 0.0);  // CPX = float: imaginary part is zero
    }
    v.phase = (float)fmod((double)v.phase, 6.28318530717958647692528676656);
    st[s] = v;
}

// ---- AGC::processBlock, agc.cpp:84-207 (modes other than OFF) --------------------------------------------
#define PB_AGC_BUF 2048  // MAX_DELAY_BUF, agc.h:27
struct AgcState {
    int sig_ptr, mag_pos, hang_timer, pad;
    double peak, attack_avg, decay_avg;
};
__global__ void agc_kernel(const cpx *__restrict__ in, int n, long long stride, int nstreams, pebble_host::AgcParams p,
                           AgcState *__restrict__ st, cpx *__restrict__ sig_buf, double *__restrict__ mag_buf,
                           cpx *__restrict__ out) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= nstreams) return;
    AgcState v = st[s];
    cpx *sig = sig_buf + (long long)s * PB_AGC_BUF;
    double *mb = mag_buf + (long long)s * PB_AGC_BUF;
    const cpx *x = in + (long long)s * stride;
    cpx *y = out + (long long)s * stride;
    const double min_const = (double)1e-8f;   // MIN_CONSTANT is a float constant, agc.h:58
    const double outscale = (double)0.7f;     // AGC_OUTSCALE, agc.h:49
    for (int i = 0; i < n; ++i) {
        const cpx xin = x[i];
        const cpx delayed = sig[v.sig_ptr];
        sig[v.sig_ptr++] = xin;
        if (v.sig_ptr >= p.delay_samples) v.sig_ptr = 0;
        double mag = fabs(xin.x);
        const double mim = fabs(xin.y);
        if (mim > mag) mag = mim;
        mag = log10(mag + min_const);  // - log10(MAX_AMPLITUDE) == 0
        double tmp = mb[v.mag_pos];
        mb[v.mag_pos++] = mag;
        if (v.mag_pos >= p.window_samples) v.mag_pos = 0;
        if (mag > v.peak) {
            v.peak = mag;
        } else if (tmp == v.peak) {
            v.peak = -8.0;
            for (int k = 0; k < p.window_samples; ++k) {
                tmp = mb[k];
                if (tmp > v.peak) v.peak = tmp;
            }
        }
        if (v.peak > v.attack_avg) v.attack_avg = (1.0 - p.attack_rise) * v.attack_avg + p.attack_rise * v.peak;
        else v.attack_avg = (1.0 - p.attack_fall) * v.attack_avg + p.attack_fall * v.peak;
        if (p.use_hang) {
            if (v.peak > v.decay_avg) {
                v.decay_avg = (1.0 - p.decay_rise) * v.decay_avg + p.decay_rise * v.peak;
                v.hang_timer = 0;
            } else if (v.hang_timer < p.hang_time) {
                v.hang_timer++;
            } else {
                v.decay_avg = (1.0 - p.decay_fall) * v.decay_avg + p.decay_fall * v.peak;
            }
        } else {
            if (v.peak > v.decay_avg) v.decay_avg = (1.0 - p.decay_rise) * v.decay_avg + p.decay_rise * v.peak;
            else v.decay_avg = (1.0 - p.decay_fall) * v.decay_avg + p.decay_fall * v.peak;
        }
        mag = v.attack_avg > v.decay_avg ? v.attack_avg : v.decay_avg;
        double gain;
        if (mag <= p.knee) gain = p.fixed_gain;
        else gain = outscale * pow(10.0, mag * (p.gain_slope - 1.0));
        y[i] = cpx_make(delayed.x * gain, delayed.y * gain);
    }
    st[s] = v;
}

// =====================================================================================================
// host side
// =====================================================================================================
#include "demod_internal.h"

static int grid_for(int n, int threads = 256) {
    int g = (n + threads - 1) / threads;
    const int cap = pebble::num_sms() * 8;
    return g < 1 ? 1 : (g > cap ? cap : g);
}

int cfir_alloc(pebble_cfir *f) {
    PB_CUDA(cudaMalloc(&f->d_hI, sizeof(double) * PB_FIR_MAX), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMalloc(&f->d_hQ, sizeof(double) * PB_FIR_MAX), PEBBLE_ERR_CUDA);
    for (int b = 0; b < 2; ++b) {
        PB_CUDA(cudaMalloc(&f->d_hist[b], sizeof(cpx) * PB_FIR_MAX), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMemset(f->d_hist[b], 0, sizeof(cpx) * PB_FIR_MAX), PEBBLE_ERR_CUDA);
    }
    PB_CUDA(cudaStreamCreateWithFlags(&f->stream, cudaStreamNonBlocking), PEBBLE_ERR_CUDA);
    return 0;
}
void cfir_release(pebble_cfir *f) {
    cudaFree(f->d_hI); cudaFree(f->d_hQ); cudaFree(f->d_hist[0]); cudaFree(f->d_hist[1]);
    if (f->stream) cudaStreamDestroy(f->stream);
    f->stg.release();
}
int cfir_upload(pebble_cfir *f) {
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(f->d_hI, f->hI, sizeof(double) * f->ntaps, cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(f->d_hQ, f->hQ, sizeof(double) * f->ntaps, cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    // Init*Filter zeroes the delay line and m_State (fir.cpp:305-311)
    PB_CUDA(cudaMemset(f->d_hist[0], 0, sizeof(cpx) * PB_FIR_MAX), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemset(f->d_hist[1], 0, sizeof(cpx) * PB_FIR_MAX), PEBBLE_ERR_CUDA);
    return 0;
}
int cfir_run_device(pebble_cfir *f, int n, const cpx *d_in, cpx *d_out, cudaStream_t st) {
    if (n <= 0) return 0;
    if (f->ntaps <= 0) { pebble::set_error("cfir: filter not initialised"); return PEBBLE_ERR_STATE; }
    if (d_in == d_out) { pebble::set_error("cfir: in-place device call not supported"); return PEBBLE_ERR_ARG; }
    cfir_kernel<<<grid_for(n), 256, 0, st>>>(f->d_hist[f->cur], d_in, n, f->d_hI, f->d_hQ, f->ntaps, d_out);
    cfir_state_kernel<<<1, 128, 0, st>>>(f->d_hist[f->cur], d_in, n, f->ntaps - 1, f->d_hist[f->cur ^ 1]);
    pebble::count_launch(2);
    PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    f->cur ^= 1;
    return n;
}

// generic host-pointer wrapper: stage in, run fn(d_in, d_out, stream), stage out
template <class Fn>
static int host_roundtrip(pebble::Staging &stg, cudaStream_t st, int n, const double *in, double *out, Fn fn) {
    const size_t nb = sizeof(cpx) * (size_t)n;
    if (!stg.reserve(nb, nb)) return PEBBLE_ERR_CUDA;
    memcpy(stg.h_in, in, nb);
    PB_CUDA(cudaMemcpyAsync(stg.d_in, stg.h_in, nb, cudaMemcpyHostToDevice, st), PEBBLE_ERR_CUDA);
    int r = fn((const cpx *)stg.d_in, (cpx *)stg.d_out, st);
    if (r < 0) return r;
    PB_CUDA(cudaMemcpyAsync(stg.h_out, stg.d_out, nb, cudaMemcpyDeviceToHost, st), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamSynchronize(st), PEBBLE_ERR_CUDA);
    memcpy(out, stg.h_out, nb);
    return r;
}

extern "C" {
pebble_cfir_t *pebble_cfir_create(void) {
    pebble_cfir *f = new pebble_cfir();
    if (cfir_alloc(f) < 0) { cfir_release(f); delete f; return nullptr; }
    return f;
}
void pebble_cfir_destroy(pebble_cfir_t *f) { if (f) { cfir_release(f); delete f; } }
int pebble_cfir_init_lp(pebble_cfir_t *f, int num_taps, double scale, double astop, double fpass, double fstop, double fs) {
    if (!f || num_taps < 0 || num_taps > PB_FIR_MAX) return PEBBLE_ERR_ARG;
    f->fs = fs;
    f->ntaps = pebble_host::cfir_design_lp(num_taps, scale, astop, fpass, fstop, fs, f->hI);
    memcpy(f->hQ, f->hI, sizeof(double) * f->ntaps);  // fir.cpp:298-302
    int r = cfir_upload(f);
    return r < 0 ? r : f->ntaps;
}
int pebble_cfir_init_hp(pebble_cfir_t *f, int num_taps, double scale, double astop, double fpass, double fstop, double fs) {
    if (!f || num_taps < 0 || num_taps > PB_FIR_MAX) return PEBBLE_ERR_ARG;
    f->fs = fs;
    f->ntaps = pebble_host::cfir_design_hp(num_taps, scale, astop, fpass, fstop, fs, f->hI);
    memcpy(f->hQ, f->hI, sizeof(double) * f->ntaps);
    int r = cfir_upload(f);
    return r < 0 ? r : f->ntaps;
}
int pebble_cfir_init_const(pebble_cfir_t *f, int num_taps, const double *icoef, const double *qcoef, double fs) {
    if (!f || num_taps <= 0 || num_taps > PB_FIR_MAX || !icoef || !qcoef) return PEBBLE_ERR_ARG;  // fir.cpp:203-228
    f->fs = fs; f->ntaps = num_taps;
    memcpy(f->hI, icoef, sizeof(double) * num_taps);
    memcpy(f->hQ, qcoef, sizeof(double) * num_taps);
    int r = cfir_upload(f);
    return r < 0 ? r : f->ntaps;
}
int pebble_cfir_get_taps(pebble_cfir_t *f, double *out, int max) {
    if (!f || !out) return PEBBLE_ERR_ARG;
    for (int i = 0; i < f->ntaps && i < max; ++i) out[i] = f->hI[i];
    return f->ntaps;
}
int pebble_cfir_process_device(pebble_cfir_t *f, int n, const double *d_in, double *d_out, void *stream) {
    if (!f || n < 0 || (n && (!d_in || !d_out))) return PEBBLE_ERR_ARG;
    return cfir_run_device(f, n, (const cpx *)d_in, (cpx *)d_out, (cudaStream_t)stream);
}
int pebble_cfir_process(pebble_cfir_t *f, int n, const double *in, double *out) {
    if (!f || n < 0 || (n && (!in || !out))) return PEBBLE_ERR_ARG;
    if (n == 0) return 0;
    return host_roundtrip(f->stg, f->stream, n, in, out,
                          [&](const cpx *di, cpx *dout, cudaStream_t st) { return cfir_run_device(f, n, di, dout, st); });
}
}  // extern "C"

// ---- AGC handle (application/agc.cpp) ---------------------------------------------------------------------
struct pebble_agc {
    double sample_rate = 0, params_rate = 100.0;  // m_sampleRate = 100.0 triggers init on first setup (agc.cpp:26)
    int buffer_size = 0;
    int mode = 0;  // AGC_OFF
    int ui_threshold = 0;
    bool use_hang = false; int threshold = 0, slope = 0, decay = 0;  // agc.cpp:20-24
    double manual_gain = 1.0;
    pebble_host::AgcParams p;
    AgcState *d_state = nullptr; cpx *d_sig = nullptr; double *d_mag = nullptr;
    cudaStream_t stream = nullptr;
    pebble::Staging stg;
};
double agc_manual_gain(const pebble_agc *h) { return h->manual_gain; }
int agc_is_off(const pebble_agc *h) { return h->mode == 0; }

static int agc_reset_buffers(pebble_agc *h) {
    // agc.cpp:262-276
    std::vector<double> mag(PB_AGC_BUF, -16.0);
    AgcState s; s.sig_ptr = 0; s.mag_pos = 0; s.hang_timer = 0; s.pad = 0; s.peak = -16.0; s.attack_avg = -5.0; s.decay_avg = -5.0;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemset(h->d_sig, 0, sizeof(cpx) * PB_AGC_BUF), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(h->d_mag, mag.data(), sizeof(double) * PB_AGC_BUF, cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(h->d_state, &s, sizeof(s), cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    return 0;
}
// AGC::setParameters, agc.cpp:237-309
static int agc_set_parameters(pebble_agc *h, bool use_hang, int threshold, int slope, int decay) {
    if (h->mode == 0) {
        h->manual_gain = pow(10, (threshold / 5) / 20.0);  // DB::dBToAmplitude(threshold / 5): integer division
        return 0;
    }
    threshold = -threshold;
    h->manual_gain = 1;
    if (use_hang == h->use_hang && threshold == h->threshold && slope == h->slope && decay == h->decay) return 0;
    h->use_hang = use_hang; h->threshold = threshold; h->slope = slope; h->decay = decay;
    if (h->params_rate != h->sample_rate) {
        h->params_rate = h->sample_rate;
        int r = agc_reset_buffers(h);
        if (r < 0) return r;
    }
    pebble_host::agc_params(h->params_rate, use_hang, threshold, slope, decay, h->p);
    return 0;
}
int agc_run_device(pebble_agc *h, int n, const cpx *d_in, cpx *d_out, cudaStream_t st) {
    if (n <= 0) return 0;
    if (h->mode == 0) {
        scale_kernel<<<grid_for(n), 256, 0, st>>>(d_in, n, h->manual_gain, d_out);
    } else {
        agc_kernel<<<1, 32, 0, st>>>(d_in, n, 0, 1, h->p, h->d_state, h->d_sig, h->d_mag, d_out);
    }
    pebble::count_launch();
    PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    return n;
}

extern "C" {
pebble_agc_t *pebble_agc_create(unsigned sample_rate, unsigned buffer_size) {
    pebble_agc *h = new pebble_agc();
    h->sample_rate = sample_rate; h->buffer_size = (int)buffer_size;
    bool ok = pebble::cuda_ok(cudaMalloc(&h->d_state, sizeof(AgcState)), "cudaMalloc") &&
              pebble::cuda_ok(cudaMalloc(&h->d_sig, sizeof(cpx) * PB_AGC_BUF), "cudaMalloc") &&
              pebble::cuda_ok(cudaMalloc(&h->d_mag, sizeof(double) * PB_AGC_BUF), "cudaMalloc") &&
              pebble::cuda_ok(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "cudaStreamCreate");
    if (!ok) { pebble_agc_destroy(h); return nullptr; }
    pebble_agc_set_mode(h, 0, 1);  // setAgcMode(AGC_OFF, 1), agc.cpp:29
    return h;
}
void pebble_agc_destroy(pebble_agc_t *h) {
    if (!h) return;
    cudaFree(h->d_state); cudaFree(h->d_sig); cudaFree(h->d_mag);
    if (h->stream) cudaStreamDestroy(h->stream);
    h->stg.release();
    delete h;
}
// AGC::setAgcMode, agc.cpp:52-82. mode: 0 OFF, 1 FAST, 2 MED, 3 SLOW, 4 LONG (agc.h:11)
int pebble_agc_set_mode(pebble_agc_t *h, int mode, int threshold) {
    if (!h || mode < 0 || mode > 4) return PEBBLE_ERR_ARG;
    int decay = 200;
    h->mode = mode;
    if (mode == 1) decay = 100;
    else if (mode == 3) decay = 500;
    else if (mode == 2) decay = 250;
    else if (mode == 4) decay = 2000;
    return agc_set_parameters(h, false, threshold, 0, decay);
}
// AGC::setAgcThreshold, agc.cpp:43-50
int pebble_agc_set_threshold(pebble_agc_t *h, int g) {
    if (!h) return PEBBLE_ERR_ARG;
    h->ui_threshold = g;
    return agc_set_parameters(h, h->use_hang, g, h->slope, h->decay);
}
int pebble_agc_process_device(pebble_agc_t *h, int n, const double *d_in, double *d_out, void *stream) {
    if (!h || n < 0 || (n && (!d_in || !d_out))) return PEBBLE_ERR_ARG;
    return agc_run_device(h, n, (const cpx *)d_in, (cpx *)d_out, (cudaStream_t)stream);
}
int pebble_agc_process(pebble_agc_t *h, int n, const double *in, double *out) {
    if (!h || n < 0 || (n && (!in || !out))) return PEBBLE_ERR_ARG;
    if (n == 0) return 0;
    return host_roundtrip(h->stg, h->stream, n, in, out,
                          [&](const cpx *di, cpx *dout, cudaStream_t st) { return agc_run_device(h, n, di, dout, st); });
}
}  // extern "C"

// ---- Demod handle (application/demod.cpp:100-141 dispatch + demod_am.cpp + demod_nfm.cpp) --------------------
struct pebble_demod {
    double rate = 0;
    int mode = 0;                 // DeviceInterface::DemodMode numbering (device_interfaces.h:124-137); dmAM at create
    int nfm_variant = 0;          // 0 = processBlockNCO (live, demod.cpp:113), 1 = processBlockFM2
    // AM
    pebble_cfir am_lp;
    double *d_am_last[2] = {nullptr, nullptr};
    int am_cur = 0;
    // NFM
    pebble_cfir nfm_lp;
    PllConst pll;
    PllState *d_pll = nullptr;
    cpx *d_fm_last[2] = {nullptr, nullptr};
    int fm_cur = 0;
    // scratch (grow-only)
    cpx *d_tmp = nullptr; size_t tmp_cap = 0;
    Aff *d_tiles = nullptr; double *d_carry = nullptr; int tiles_cap = 0;
    cudaStream_t stream = nullptr;
    pebble::Staging stg;
};
static int demod_reserve(pebble_demod *h, int n) {
    if ((size_t)n > h->tmp_cap) {
        cudaFree(h->d_tmp);
        h->d_tmp = nullptr; h->tmp_cap = 0;
        PB_CUDA(cudaMalloc(&h->d_tmp, sizeof(cpx) * (size_t)n), PEBBLE_ERR_CUDA);
        h->tmp_cap = (size_t)n;
    }
    const int tiles = (n + PB_SCAN_TILE - 1) / PB_SCAN_TILE;
    if (tiles > h->tiles_cap) {
        cudaFree(h->d_tiles); cudaFree(h->d_carry);
        h->d_tiles = nullptr; h->d_carry = nullptr; h->tiles_cap = 0;
        PB_CUDA(cudaMalloc(&h->d_tiles, sizeof(Aff) * (size_t)tiles), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMalloc(&h->d_carry, sizeof(double) * (size_t)tiles), PEBBLE_ERR_CUDA);
        h->tiles_cap = tiles;
    }
    return 0;
}
int demod_run_device(pebble_demod *h, int n, const cpx *d_in, cpx *d_out, cudaStream_t st, int *passthrough) {
    if (passthrough) *passthrough = 0;
    if (n <= 0) return 0;
    if (h->mode == 0) {  // dmAM -> processBlockFiltered (demod.cpp:105-108)
        int r = demod_reserve(h, n);
        if (r < 0) return r;
        const double alpha = (double)0.9999f;  // DC_ALPHA is a float literal, demod_am.cpp:36
        const int tiles = (n + PB_SCAN_TILE - 1) / PB_SCAN_TILE;
        am_tile_kernel<<<tiles, PB_SCAN_THREADS, 0, st>>>(d_in, n, alpha, h->d_tiles);
        am_carry_kernel<<<1, 32, 0, st>>>(h->d_tiles, tiles, h->d_am_last[h->am_cur], h->d_carry);
        am_apply_kernel<<<tiles, PB_SCAN_THREADS, 0, st>>>(d_in, n, alpha, h->d_carry, h->d_tmp, h->d_am_last[h->am_cur ^ 1]);
        pebble::count_launch(3);
        PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
        h->am_cur ^= 1;
        return cfir_run_device(&h->am_lp, n, h->d_tmp, d_out, st);  // demod_am.cpp:62
    }
    if (h->mode == 2) {  // dmFMN
        int r = demod_reserve(h, n);
        if (r < 0) return r;
        if (h->nfm_variant == 0) {
            pll_kernel<<<1, 32, 0, st>>>(d_in, n, 0, 1, h->pll, h->d_pll, h->d_tmp);
            pebble::count_launch();
            PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
            return cfir_run_device(&h->nfm_lp, n, h->d_tmp, d_out, st);  // demod_nfm.cpp:255
        }
        fm2_kernel<<<grid_for(n), 256, 0, st>>>(d_in, n, h->d_fm_last[h->fm_cur], d_out, h->d_fm_last[h->fm_cur ^ 1]);
        pebble::count_launch();
        PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
        h->fm_cur ^= 1;
        return n;
    }
    if (h->mode == 1 || h->mode == 3 || h->mode == 4) {
        pebble::set_error("demod: SAM / WFM modes are outside this build's hot-path scope (SURVEY 8f rank 4)");
        return PEBBLE_ERR_STATE;
    }
    // USB/LSB/DSB/CW/DIG/NONE: Demod::processBlock returns `in` itself (demod.cpp:127-138)
    if (passthrough) *passthrough = 1;
    if (d_out != d_in) PB_CUDA(cudaMemcpyAsync(d_out, d_in, sizeof(cpx) * (size_t)n, cudaMemcpyDeviceToDevice, st), PEBBLE_ERR_CUDA);
    return n;
}

extern "C" {
pebble_demod_t *pebble_demod_create(unsigned input_rate, unsigned num_samples) {
    (void)num_samples;
    pebble_demod *h = new pebble_demod();
    h->rate = input_rate;
    bool ok = cfir_alloc(&h->am_lp) == 0 && cfir_alloc(&h->nfm_lp) == 0 &&
              pebble::cuda_ok(cudaMalloc(&h->d_pll, sizeof(PllState)), "cudaMalloc") &&
              pebble::cuda_ok(cudaMemset(h->d_pll, 0, sizeof(PllState)), "cudaMemset") &&
              pebble::cuda_ok(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "cudaStreamCreate");
    for (int b = 0; ok && b < 2; ++b) {
        ok = pebble::cuda_ok(cudaMalloc(&h->d_am_last[b], sizeof(double)), "cudaMalloc") &&
             pebble::cuda_ok(cudaMemset(h->d_am_last[b], 0, sizeof(double)), "cudaMemset") &&
             pebble::cuda_ok(cudaMalloc(&h->d_fm_last[b], sizeof(cpx)), "cudaMalloc") &&
             pebble::cuda_ok(cudaMemset(h->d_fm_last[b], 0, sizeof(cpx)), "cudaMemset");
    }
    if (!ok) { pebble_demod_destroy(h); return nullptr; }
    // Demod_AM ctor: setBandwidth(16000) (demod_am.cpp:9)
    pebble_demod_set_mode(h, 0);
    h->am_lp.ntaps = pebble_host::cfir_design_lp(0, 1.0, 50.0, 16000.0, 16000.0 * 1.8, h->rate, h->am_lp.hI);
    memcpy(h->am_lp.hQ, h->am_lp.hI, sizeof(double) * h->am_lp.ntaps);
    // Demod_NFM::init (demod_nfm.cpp:44-66); members are float
    const double norm = 6.28318530717958647692528676656 / h->rate;
    h->pll.lo = (float)(-15000.0 * norm);
    h->pll.hi = (float)(15000.0 * norm);
    h->pll.alpha = (float)(2.0 * .707 * 3000.0 * norm);
    h->pll.beta = (float)((h->pll.alpha * h->pll.alpha) / (4.0 * .707 * .707));
    h->pll.out_gain = 1.0f;
    h->pll.dc_alpha = (float)(1.0 - exp(-1.0 / (h->rate * 0.001)));
    h->nfm_lp.ntaps = pebble_host::cfir_design_lp(0, 1.0, 50.0, 3000.0, 1.6 * 3000.0, h->rate, h->nfm_lp.hI);
    memcpy(h->nfm_lp.hQ, h->nfm_lp.hI, sizeof(double) * h->nfm_lp.ntaps);
    if (cfir_upload(&h->am_lp) < 0 || cfir_upload(&h->nfm_lp) < 0) { pebble_demod_destroy(h); return nullptr; }
    return h;
}
void pebble_demod_destroy(pebble_demod_t *h) {
    if (!h) return;
    cfir_release(&h->am_lp); cfir_release(&h->nfm_lp);
    cudaFree(h->d_pll); cudaFree(h->d_tmp); cudaFree(h->d_tiles); cudaFree(h->d_carry);
    for (int b = 0; b < 2; ++b) { cudaFree(h->d_am_last[b]); cudaFree(h->d_fm_last[b]); }
    if (h->stream) cudaStreamDestroy(h->stream);
    h->stg.release();
    delete h;
}
int pebble_demod_set_mode(pebble_demod_t *h, int mode) {  // Demod::setDemodMode, demod.cpp:230-247
    if (!h || mode < 0 || mode > 12) return PEBBLE_ERR_ARG;
    h->mode = mode;
    return 0;
}
int pebble_demod_set_nfm_variant(pebble_demod_t *h, int variant) {
    if (!h || variant < 0 || variant > 1) return PEBBLE_ERR_ARG;
    h->nfm_variant = variant;
    return 0;
}
int pebble_demod_set_bandwidth(pebble_demod_t *h, double bw) {  // Demod::setBandwidth, demod.cpp:218-228: AM only
    if (!h) return PEBBLE_ERR_ARG;
    if (h->mode != 0) return 0;
    h->am_lp.ntaps = pebble_host::cfir_design_lp(0, 1.0, 50.0, bw, bw * 1.8, h->rate, h->am_lp.hI);  // demod_am.cpp:21
    memcpy(h->am_lp.hQ, h->am_lp.hI, sizeof(double) * h->am_lp.ntaps);
    return cfir_upload(&h->am_lp);
}
int pebble_demod_is_passthrough(const pebble_demod_t *h) {
    if (!h) return PEBBLE_ERR_ARG;
    return (h->mode >= 5) ? 1 : 0;
}
int pebble_demod_process_device(pebble_demod_t *h, int n, const double *d_in, double *d_out, void *stream) {
    if (!h || n < 0 || (n && (!d_in || !d_out))) return PEBBLE_ERR_ARG;
    return demod_run_device(h, n, (const cpx *)d_in, (cpx *)d_out, (cudaStream_t)stream, nullptr);
}
int pebble_demod_process(pebble_demod_t *h, int n, const double *in, double *out) {
    if (!h || n < 0 || (n && (!in || !out))) return PEBBLE_ERR_ARG;
    if (n == 0) return 0;
    return host_roundtrip(h->stg, h->stream, n, in, out, [&](const cpx *di, cpx *dout, cudaStream_t st) {
        return demod_run_device(h, n, di, dout, st, nullptr);
    });
}
}  // extern "C"
