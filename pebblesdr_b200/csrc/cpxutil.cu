// cpxutil.cu — CpxUtil vector helpers on device buffers: pebblelib/cpx.h:161-196, cpx.cpp:35-112. The fused chain
// never calls them (its scalings ride on filter spectra and oscillator seeds); they exist for callers that keep their own
// CPX buffers on the device between operator calls. Element-wise, memory-bound; each forms its result with the reference's
// own operations and roundings (std::complex<double> product = ac - bd, ad + bc with separate roundings), so results are
// bit-identical to the host functions.
#include "lib_common.h"
#include "common.cuh"

enum { CU_COPY, CU_SCALE, CU_MULT, CU_ADD, CU_MAG, CU_SQRMAG, CU_CLEAR };

template <int OP>
__global__ void __launch_bounds__(256) cpxutil_kernel(cpx *__restrict__ out, const cpx *__restrict__ a, const cpx *__restrict__ b, double s,
                                                      long long n) {
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
        cpx r;
        if (OP == CU_CLEAR) r = cpx_make(0.0, 0.0);
        else {
            const cpx x = a[i];
            if (OP == CU_COPY) r = x;
            else if (OP == CU_SCALE) r = cpx_make(__dmul_rn(x.x, s), __dmul_rn(x.y, s));                        // scaleCpx
            else if (OP == CU_ADD) { const cpx y = b[i]; r = cpx_make(__dadd_rn(x.x, y.x), __dadd_rn(x.y, y.y)); }
            else if (OP == CU_MULT) {
                const cpx y = b[i];
                r = cpx_make(__dsub_rn(__dmul_rn(x.x, y.x), __dmul_rn(x.y, y.y)), __dadd_rn(__dmul_rn(x.x, y.y), __dmul_rn(x.y, y.x)));
            } else if (OP == CU_MAG) r = cpx_make(__dsqrt_rn(__dadd_rn(__dmul_rn(x.x, x.x), __dmul_rn(x.y, x.y))), x.x);   // magCPX keeps in.re in .im (cpx.cpp:82-88)
            else r = cpx_make(__dadd_rn(__dmul_rn(x.x, x.x), __dmul_rn(x.y, x.y)), x.x);                          // sqrMagCPX (:90-96)
        }
        out[i] = r;
    }
}
// decimateCPX (cpx.cpp:105-111): out[j] = in[j * by]
__global__ void __launch_bounds__(256) cpx_decimate_kernel(cpx *__restrict__ out, const cpx *__restrict__ in, int by, long long nout) {
    for (long long j = (long long)blockIdx.x * 256 + threadIdx.x; j < nout; j += (long long)gridDim.x * 256) out[j] = in[j * by];
}

template <int OP> static int cu_launch(double *out, const double *a, const double *b, double s, long long n, void *stream) {
    if (n < 0 || (n && (!out || (OP != CU_CLEAR && !a) || ((OP == CU_MULT || OP == CU_ADD) && !b)))) {
        pebble::set_error("cpx util: bad argument");
        return PEBBLE_ERR_ARG;
    }
    if (n == 0) return 0;
    long long blocks = (n + 255) / 256;
    const long long cap = (long long)pebble::num_sms() * 16;
    if (blocks > cap) blocks = cap;
    cpxutil_kernel<OP><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((cpx *)out, (const cpx *)a, (const cpx *)b, s, n);
    pebble::count_launch();
    PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    return 0;
}

extern "C" {
int pebble_cpx_copy_device(double *d_out, const double *d_in, long long n, void *stream) { return cu_launch<CU_COPY>(d_out, d_in, nullptr, 0, n, stream); }
int pebble_cpx_scale_device(double *d_out, const double *d_in, double a, long long n, void *stream) { return cu_launch<CU_SCALE>(d_out, d_in, nullptr, a, n, stream); }
int pebble_cpx_mult_device(double *d_out, const double *d_in, const double *d_in2, long long n, void *stream) { return cu_launch<CU_MULT>(d_out, d_in, d_in2, 0, n, stream); }
int pebble_cpx_add_device(double *d_out, const double *d_in, const double *d_in2, long long n, void *stream) { return cu_launch<CU_ADD>(d_out, d_in, d_in2, 0, n, stream); }
int pebble_cpx_mag_device(double *d_out, const double *d_in, long long n, void *stream) { return cu_launch<CU_MAG>(d_out, d_in, nullptr, 0, n, stream); }
int pebble_cpx_sqrmag_device(double *d_out, const double *d_in, long long n, void *stream) { return cu_launch<CU_SQRMAG>(d_out, d_in, nullptr, 0, n, stream); }
int pebble_cpx_clear_device(double *d_out, long long n, void *stream) { return cu_launch<CU_CLEAR>(d_out, nullptr, nullptr, 0, n, stream); }
int pebble_cpx_decimate_device(double *d_out, const double *d_in, int by, long long n, void *stream) {
    if (by < 1 || n < 0 || (n && (!d_out || !d_in))) { pebble::set_error("cpx util: bad argument"); return PEBBLE_ERR_ARG; }
    const long long nout = (n + by - 1) / by;
    if (nout == 0) return 0;
    long long blocks = (nout + 255) / 256;
    const long long cap = (long long)pebble::num_sms() * 16;
    if (blocks > cap) blocks = cap;
    cpx_decimate_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((cpx *)d_out, (const cpx *)d_in, by, nout);
    pebble::count_launch();
    PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    return (int)(nout > 0x7fffffff ? 0x7fffffff : nout);
}
}  // extern "C"
