// fastfir.cu — CFastFIR (overlap-save FFT band-pass) on the GPU + its C ABI.
// Reference: pebblelib/fastfir.cpp:77-334. One CTA per overlap-save block: 2L = N samples (L old + L new)
// are staged into shared memory, forward FFT, multiplied by the filter spectrum H (kept in the same
// digit-reversed slot order the forward passes produce), inverse FFT, and the upper L samples written.
// The 1/N of the unnormalised FFTW inverse is folded into H by the reference itself (fastfir.cpp:243).
#include "lib_common.h"
#include "fft_kernels.cuh"
#include "host_design.h"
#include <cstring>
#include <vector>

// The "virtual stream" of a call is state ++ input, state = overlap (L) + not yet processed samples.
__device__ __forceinline__ cpx vs_load(const cpx *__restrict__ state, int s0, const cpx *__restrict__ in, long long v) {
    return v < s0 ? state[v] : in[v - s0];
}

template <int N>
__global__ void __launch_bounds__(FftPlan<N>::THREADS)
fastfir_kernel(const cpx *__restrict__ state, int s0, const cpx *__restrict__ in, cpx *__restrict__ out,
               const cpx *__restrict__ hpos, const cpx *__restrict__ tw, int nblocks, double in_scale,
               double out_scale, long long in_stride, long long out_stride, int state_stride) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cpx *a = reinterpret_cast<cpx *>(smem_raw);
    constexpr int L = N / 2;
    constexpr int SH = FftPlan<N>::SH;
    // batched form: blockIdx.y = independent stream (own overlap state, own rows of in/out), same filter
    state += (long long)blockIdx.y * state_stride;
    in += blockIdx.y * in_stride;
    out += blockIdx.y * out_stride;
    for (int blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
        const long long v0 = (long long)blk * L;
        for (int i = threadIdx.x; i < N; i += blockDim.x)
            a[fft_swz<SH>(i)] = cpx_scale(vs_load(state, s0, in, v0 + i), in_scale);
        __syncthreads();
        fft_forward_smem<N>(a, tw);
        // CpxMpy(N, m_pFilterCoef, m_pFFTBuf, m_pFFTBuf), fastfir.cpp:306,324-334
        for (int p = threadIdx.x; p < N; p += blockDim.x) a[fft_swz<SH>(p)] = cpx_mul(hpos[p], a[fft_swz<SH>(p)]);
        __syncthreads();
        fft_inverse_smem<N>(a, tw);
        // fastfir.cpp:308-311: output = samples FIR-1 .. N-1 of the inverse transform
        for (int i = threadIdx.x; i < L; i += blockDim.x) out[v0 + i] = cpx_scale(a[fft_swz<SH>(L + i)], out_scale);
        __syncthreads();
    }
}

// Forward FFT of the zero-padded taps, stored in slot order (what fastfir_kernel multiplies by).
template <int N>
__global__ void __launch_bounds__(FftPlan<N>::THREADS)
fastfir_filter_kernel(const cpx *__restrict__ taps, cpx *__restrict__ hpos, const cpx *__restrict__ tw) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    cpx *a = reinterpret_cast<cpx *>(smem_raw);
    constexpr int SH = FftPlan<N>::SH;
    for (int i = threadIdx.x; i < N; i += blockDim.x) a[fft_swz<SH>(i)] = taps[i];
    __syncthreads();
    fft_forward_smem<N>(a, tw);
    for (int p = threadIdx.x; p < N; p += blockDim.x) hpos[p] = a[fft_swz<SH>(p)];
}

// new_state[i] = V[start + i], V = state ++ in
__global__ void fastfir_state_kernel(const cpx *__restrict__ state, int s0, const cpx *__restrict__ in,
                                     cpx *__restrict__ new_state, long long start, int count, long long in_stride,
                                     int state_stride) {
    state += (long long)blockIdx.y * state_stride;
    new_state += (long long)blockIdx.y * state_stride;
    in += blockIdx.y * in_stride;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) new_state[i] = vs_load(state, s0, in, start + i);
}

struct pebble_fastfir {
    int N, fir, L;
    // reference members (fastfir.cpp:140-143): "unset" sentinels so the first setup always runs
    double lo, hi, offset, fs;
    cpx *d_tw, *d_hpos, *d_taps, *d_state[2];
    int cur, fill;
    int streams;          // independent streams sharing this filter (batched form); state is streams x N cpx
    bool have_filter;
    double in_scale, out_scale;
    cudaStream_t stream;
    pebble::Staging stg;
};

template <int N> static cudaError_t ff_launch_filter(pebble_fastfir *h, cudaStream_t st) {
    const size_t smem = sizeof(cpx) * N;
    cudaError_t e = cudaFuncSetAttribute(fastfir_filter_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    fastfir_filter_kernel<N><<<1, FftPlan<N>::THREADS, smem, st>>>(h->d_taps, h->d_hpos, h->d_tw);
    pebble::count_launch();
    return cudaGetLastError();
}
template <int N>
static cudaError_t ff_launch_process(pebble_fastfir *h, const cpx *in, cpx *out, int nblocks, cudaStream_t st,
                                     int streams = 1, long long in_stride = 0, long long out_stride = 0) {
    const size_t smem = sizeof(cpx) * N;
    cudaError_t e = cudaFuncSetAttribute(fastfir_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int occ = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fastfir_kernel<N>, FftPlan<N>::THREADS, smem);
    if (occ < 1) occ = 1;
    int grid = pebble::num_sms() * occ * 4;
    if (streams > 1) grid = (grid + streams - 1) / streams;
    if (grid > nblocks) grid = nblocks;
    if (grid < 1) grid = 1;
    fastfir_kernel<N><<<dim3(grid, streams), FftPlan<N>::THREADS, smem, st>>>(
        h->d_state[h->cur], h->L + h->fill, in, out, h->d_hpos, h->d_tw, nblocks, h->in_scale, h->out_scale, in_stride,
        out_stride, streams > 1 ? h->N : 0);
    pebble::count_launch();
    return cudaGetLastError();
}

#define FF_DISPATCH(h, expr2048, expr4096, expr8192)                 \
    ((h)->N == 2048 ? (expr2048) : (h)->N == 4096 ? (expr4096) : (expr8192))

extern "C" {

pebble_fastfir_t *pebble_fastfir_create_sized(int fft_size, int fir_size) {
    if ((fft_size != 2048 && fft_size != 4096 && fft_size != 8192) || fir_size != fft_size / 2 + 1) {
        pebble::set_error("fastfir: unsupported size %d/%d (fft in {2048,4096,8192}, fir = fft/2+1)", fft_size, fir_size);
        return nullptr;
    }
    pebble_fastfir *h = new pebble_fastfir();
    h->N = fft_size; h->fir = fir_size; h->L = fft_size / 2;
    h->lo = -1.0; h->hi = 1.0; h->offset = 1.0; h->fs = 1.0;   // fastfir.cpp:140-143
    h->d_tw = h->d_hpos = h->d_taps = nullptr; h->d_state[0] = h->d_state[1] = nullptr;
    h->cur = 0; h->fill = 0; h->streams = 1; h->have_filter = false; h->in_scale = h->out_scale = 1.0; h->stream = nullptr;
    const size_t nb = sizeof(cpx) * (size_t)fft_size;
    std::vector<double> tw;
    pebble_host::twiddle_table(fft_size, tw);
    bool ok = pebble::cuda_ok(cudaMalloc(&h->d_tw, nb), "cudaMalloc") && pebble::cuda_ok(cudaMalloc(&h->d_hpos, nb), "cudaMalloc") &&
              pebble::cuda_ok(cudaMalloc(&h->d_taps, nb), "cudaMalloc") && pebble::cuda_ok(cudaMalloc(&h->d_state[0], nb), "cudaMalloc") &&
              pebble::cuda_ok(cudaMalloc(&h->d_state[1], nb), "cudaMalloc") &&
              pebble::cuda_ok(cudaMemcpy(h->d_tw, tw.data(), nb, cudaMemcpyHostToDevice), "cudaMemcpy(tw)") &&
              pebble::cuda_ok(cudaMemset(h->d_state[0], 0, nb), "cudaMemset") && pebble::cuda_ok(cudaMemset(h->d_state[1], 0, nb), "cudaMemset") &&
              // until SetupParameters runs the reference multiplies by an uninitialised buffer; we use zeros
              pebble::cuda_ok(cudaMemset(h->d_hpos, 0, nb), "cudaMemset") &&
              pebble::cuda_ok(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "cudaStreamCreate");
    if (!ok) { pebble_fastfir_destroy(h); return nullptr; }
    return h;
}
pebble_fastfir_t *pebble_fastfir_create(void) { return pebble_fastfir_create_sized(2048, 1025); }  // fastfir.cpp:65-66

void pebble_fastfir_destroy(pebble_fastfir_t *h) {
    if (!h) return;
    cudaFree(h->d_tw); cudaFree(h->d_hpos); cudaFree(h->d_taps); cudaFree(h->d_state[0]); cudaFree(h->d_state[1]);
    if (h->stream) cudaStreamDestroy(h->stream);
    h->stg.release();
    delete h;
}

int pebble_fastfir_setup(pebble_fastfir_t *h, double lo, double hi, double offset, double fs) {
    if (!h) return PEBBLE_ERR_ARG;
    if (lo == h->lo && hi == h->hi && offset == h->offset && fs == h->fs) return 1;  // fastfir.cpp:195-199
    h->lo = lo; h->hi = hi; h->offset = offset; h->fs = fs;
    std::vector<double> taps;
    if (!pebble_host::fastfir_design(h->N, h->fir, lo, hi, offset, fs, taps)) return 2;  // fastfir.cpp:208-216
    // in-order with any process call already queued on the legacy stream
    PB_CUDA(cudaMemcpy(h->d_taps, taps.data(), sizeof(cpx) * (size_t)h->N, cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    cudaError_t e = FF_DISPATCH(h, ff_launch_filter<2048>(h, 0), ff_launch_filter<4096>(h, 0), ff_launch_filter<8192>(h, 0));
    PB_CUDA(e, PEBBLE_ERR_CUDA);
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    h->have_filter = true;
    return 0;
}

int pebble_fastfir_set_fused_scale(pebble_fastfir_t *h, double in_scale, double out_scale) {
    if (!h) return PEBBLE_ERR_ARG;
    h->in_scale = in_scale; h->out_scale = out_scale;
    return 0;
}

int pebble_fastfir_out_count(const pebble_fastfir_t *h, int n) {
    if (!h || n < 0) return PEBBLE_ERR_ARG;
    return (int)((((long long)h->fill + n) / h->L) * h->L);
}

int pebble_fastfir_set_streams(pebble_fastfir_t *h, int streams) {
    if (!h || streams < 1) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    const size_t nb = sizeof(cpx) * (size_t)h->N * (size_t)streams;
    for (int b = 0; b < 2; ++b) {
        cudaFree(h->d_state[b]); h->d_state[b] = nullptr;
        PB_CUDA(cudaMalloc(&h->d_state[b], nb), PEBBLE_ERR_CUDA);
        PB_CUDA(cudaMemset(h->d_state[b], 0, nb), PEBBLE_ERR_CUDA);
    }
    h->streams = streams; h->cur = 0; h->fill = 0;
    return 0;
}

int pebble_fastfir_process_batch_device(pebble_fastfir_t *h, int n, const double *d_in, long long in_stride,
                                        double *d_out, long long out_stride, void *stream_v) {
    if (!h || n < 0 || (n && !d_in)) { pebble::set_error("fastfir: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream_v;
    const long long total = (long long)h->fill + n;
    const int nblocks = (int)(total / h->L);
    if (nblocks > 0) {
        if (!d_out) { pebble::set_error("fastfir: NULL output"); return PEBBLE_ERR_ARG; }
        cudaError_t e = FF_DISPATCH(h,
            ff_launch_process<2048>(h, (const cpx *)d_in, (cpx *)d_out, nblocks, st, h->streams, in_stride, out_stride),
            ff_launch_process<4096>(h, (const cpx *)d_in, (cpx *)d_out, nblocks, st, h->streams, in_stride, out_stride),
            ff_launch_process<8192>(h, (const cpx *)d_in, (cpx *)d_out, nblocks, st, h->streams, in_stride, out_stride));
        PB_CUDA(e, PEBBLE_ERR_CUDA);
    }
    const int new_fill = (int)(total % h->L);
    const int count = h->L + new_fill;
    const long long vlen = (long long)h->L + total;
    fastfir_state_kernel<<<dim3((count + 255) / 256, h->streams), 256, 0, st>>>(
        h->d_state[h->cur], h->L + h->fill, (const cpx *)d_in, h->d_state[h->cur ^ 1], vlen - count, count, in_stride,
        h->streams > 1 ? h->N : 0);
    pebble::count_launch();
    PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    h->cur ^= 1;
    h->fill = new_fill;
    return nblocks * h->L;
}

int pebble_fastfir_process_device(pebble_fastfir_t *h, int n, const double *d_in, double *d_out, void *stream_v) {
    if (!h || n < 0 || (n && !d_in)) { pebble::set_error("fastfir: bad argument"); return PEBBLE_ERR_ARG; }
    if (h->streams != 1) { pebble::set_error("fastfir: handle is batched, use process_batch_device"); return PEBBLE_ERR_STATE; }
    if (n == 0) return 0;  // fastfir.cpp:287-288
    cudaStream_t st = (cudaStream_t)stream_v;
    const long long total = (long long)h->fill + n;
    const int nblocks = (int)(total / h->L);
    if (nblocks > 0) {
        if (!d_out) { pebble::set_error("fastfir: NULL output"); return PEBBLE_ERR_ARG; }
        cudaError_t e = FF_DISPATCH(h, ff_launch_process<2048>(h, (const cpx *)d_in, (cpx *)d_out, nblocks, st),
                                    ff_launch_process<4096>(h, (const cpx *)d_in, (cpx *)d_out, nblocks, st),
                                    ff_launch_process<8192>(h, (const cpx *)d_in, (cpx *)d_out, nblocks, st));
        PB_CUDA(e, PEBBLE_ERR_CUDA);
    }
    // carry: the last L samples before the first unprocessed one, plus the unprocessed remainder
    const int new_fill = (int)(total % h->L);
    const int count = h->L + new_fill;
    const long long vlen = (long long)h->L + total;  // length of state ++ in
    fastfir_state_kernel<<<(count + 255) / 256, 256, 0, st>>>(h->d_state[h->cur], h->L + h->fill, (const cpx *)d_in,
                                                             h->d_state[h->cur ^ 1], vlen - count, count, 0, 0);
    pebble::count_launch();
    PB_CUDA(cudaGetLastError(), PEBBLE_ERR_CUDA);
    h->cur ^= 1;
    h->fill = new_fill;
    return nblocks * h->L;
}

int pebble_fastfir_process(pebble_fastfir_t *h, int n, const double *in, double *out) {
    if (!h || n < 0 || (n && (!in || !out))) { pebble::set_error("fastfir: bad argument"); return PEBBLE_ERR_ARG; }
    if (n == 0) return 0;
    const int k = pebble_fastfir_out_count(h, n);
    const size_t bi = (size_t)n * sizeof(cpx), bo = (size_t)k * sizeof(cpx);
    if (!h->stg.reserve(bi, bo ? bo : 16)) return PEBBLE_ERR_CUDA;
    memcpy(h->stg.h_in, in, bi);
    PB_CUDA(cudaMemcpyAsync(h->stg.d_in, h->stg.h_in, bi, cudaMemcpyHostToDevice, h->stream), PEBBLE_ERR_CUDA);
    int r = pebble_fastfir_process_device(h, n, (const double *)h->stg.d_in, (double *)h->stg.d_out, h->stream);
    if (r < 0) return r;
    if (bo) PB_CUDA(cudaMemcpyAsync(h->stg.h_out, h->stg.d_out, bo, cudaMemcpyDeviceToHost, h->stream), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaStreamSynchronize(h->stream), PEBBLE_ERR_CUDA);
    if (bo) memcpy(out, h->stg.h_out, bo);
    return r;
}

int64_t pebble_fastfir_state_size(const pebble_fastfir_t *h) {
    if (!h) return PEBBLE_ERR_ARG;
    return 8 + (int64_t)sizeof(cpx) * (h->L + h->fill);
}
int pebble_fastfir_get_state(pebble_fastfir_t *h, void *buf, int64_t bytes) {
    if (!h || !buf || bytes < pebble_fastfir_state_size(h)) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    int64_t f = h->fill;
    memcpy(buf, &f, 8);
    PB_CUDA(cudaMemcpy((char *)buf + 8, h->d_state[h->cur], sizeof(cpx) * (h->L + h->fill), cudaMemcpyDeviceToHost), PEBBLE_ERR_CUDA);
    return 0;
}
int pebble_fastfir_set_state(pebble_fastfir_t *h, const void *buf, int64_t bytes) {
    if (!h || !buf || bytes < 8) return PEBBLE_ERR_ARG;
    int64_t f;
    memcpy(&f, buf, 8);
    if (f < 0 || f >= h->L || bytes < 8 + (int64_t)sizeof(cpx) * (h->L + f)) return PEBBLE_ERR_ARG;
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(h->d_state[h->cur], (const char *)buf + 8, sizeof(cpx) * (h->L + f), cudaMemcpyHostToDevice), PEBBLE_ERR_CUDA);
    h->fill = (int)f;
    return 0;
}
int pebble_fastfir_get_filter(pebble_fastfir_t *h, double *coef_out) {
    if (!h || !coef_out) return PEBBLE_ERR_ARG;
    std::vector<cpx> hp(h->N);
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    PB_CUDA(cudaMemcpy(hp.data(), h->d_hpos, sizeof(cpx) * h->N, cudaMemcpyDeviceToHost), PEBBLE_ERR_CUDA);
    for (int f = 0; f < h->N; ++f) {
        int p = FF_DISPATCH(h, FftPlan<2048>::freq2pos(f), FftPlan<4096>::freq2pos(f), FftPlan<8192>::freq2pos(f));
        coef_out[2 * f] = hp[p].x;
        coef_out[2 * f + 1] = hp[p].y;
    }
    return 0;
}

}  // extern "C"
