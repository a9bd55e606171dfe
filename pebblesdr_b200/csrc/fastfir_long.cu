// fastfir_long.cu — CFastFIR for long filters: N-point overlap-save with N = N1 x 2048 too large for shared memory
// (BASELINE config 5: CONV_FFT_SIZE 131072 / CONV_FIR_SIZE 65537, fastfir.cpp:65-66 with the sizes lifted; algorithm
// fastfir.cpp:281-334 unchanged: FFT of L old + L new samples, x H, unnormalised inverse, keep samples L..N-1).
//
// Four-step transform, index split n = 2048 n1 + n2, k = k1 + N1 k2 (N1 = R x R, R = 8 for 131072, 4 for 32768):
//   cols fwd : for every column n2 an N1-point DFT over n1 (stride 2048)                 -> scratch[k1][n2]
//   rows     : x W_N^(n2 k1), 2048-point DFT over n2 (shared memory, fft_core.cuh), x H (CpxMpy, fastfir.cpp:324-334),
//              inverse 2048-point DFT, x conj W_N^(n2 k1)                                -> scratch[k1][n2] in place
//   cols inv : inverse N1-point DFT over k1; only n1 >= N1/2 (the L valid samples) is formed and written.
// H is kept in the rows kernel's own [k1][slot] order (slot = digit-reversed k2), so nothing is ever reordered.
//
// Super-blocks: overlap-save at N = 2 (FIR - 1) transforms every input sample twice. Seven consecutive reference blocks
// are the same linear convolution as ONE block of a 4N-point transform that keeps its last 4N - (FIR - 1) = 7L samples
// (8/7 transforms per sample instead of 2: 36 % fewer flops and 39 % less scratch traffic per output sample), so a call
// with >= 7 blocks per stream runs floor(nblocks / 7) super-blocks on a second plan (4N = 16 x 16 x 2048 for 131072) and
// only the remaining < 7 blocks on the N-point plan. Same samples as the block-by-block form up to FP64 rounding.
//
// HBM: algorithmic 32 B per input sample (16 read + 16 written). The scratch round trips (3 x 32 B per sample) are kept
// out of DRAM by processing `chunk` blocks at a time with chunk x N x 16 B well inside the 126 MB L2: the three launches
// of a chunk run back to back on one stream, so the rows and the inverse columns find their input in L2.
#include "fastfir_long.cuh"
#include "fft_kernels.cuh"
#include "ff_fused.cuh"
#include "host_design.h"
#include <cstdlib>
#include <vector>

namespace {
constexpr int LF_N2 = 2048;
// columns per CTA of the column kernels: 16 x 16 B = 256 B contiguous per row (8 for the radix-16 plan: 32 KB of shared memory)
#ifndef LF_DISCARD
#define LF_DISCARD 1
#endif
template <int R> struct LfTc { static constexpr int v = R == 16 ? 8 : 16; };

__device__ __forceinline__ cpx lf_ld_stream(const cpx *p) {
    cpx r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
    return r;
}
__device__ __forceinline__ void lf_st_stream(cpx *p, cpx v) {
    asm volatile("st.global.cs.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(v.x), "d"(v.y) : "memory");
}

struct LfGeom {
    int nblocks;            // blocks per stream in this call (of this plan)
    int step;               // samples a block advances = samples it writes (N - (FIR - 1))
    long long off;          // sample of the virtual stream / of `out` where this plan's block 0 starts
    int s0;                 // samples of the virtual stream that come from `state`
    int state_stride;       // cpx between the states of consecutive streams
    long long in_stride, out_stride;
    long long job0;         // first job (stream * nblocks + block) of this launch
};

// Column transforms of LF_TC adjacent columns of one block. Thread (c, i): column c, butterfly i of each radix-R pass.
// Forward: pass 0 on rows i + R k (twiddle W_N1^(i q) after the DFT), pass 1 on rows R i + k; slot R i + q = bin i + R q.
// Inverse: the mirror image (pass 1, conj twiddle, pass 0); SKIPK = (FIR - 1) / (2048 R): samples n1 < R SKIPK are the
// overlap the inverse discards.
// TW: the inter-step twiddle W_N^(col k1) is applied HERE (on the way out of the forward transform, on the way into the
// inverse) instead of in the row kernel: these kernels wait on L2 with the FP64 pipe idle, the row kernel does not. The
// R factors of a thread are W^(col i) (W^(col R))^q: two exact table entries and a chain of at most R - 1 products
// (<= 2e-15 relative). TW = false leaves it to lf_rows_kernel (filter build, exact table entries throughout).
// lf_pipe_kernel only: the previous item's completion signal (release-add on *sig), sent by thread 0 once THIS item's
// global loads are in flight, so that the fence inside it (which waits for the previous item's stores to reach L2) costs
// no time of its own.
__device__ __forceinline__ void lf_signal(int *&sig) {
    if (sig) {
        asm volatile("red.release.gpu.global.add.s32 [%0], 1;" ::"l"(sig) : "memory");
        sig = nullptr;
    }
}

template <int R, bool INV, int SKIPK, bool TW>
__device__ __forceinline__ void lf_cols_tile(cpx *a, const cpx *w, const cpx *__restrict__ state, const cpx *__restrict__ in,
                                             cpx *__restrict__ out, cpx *__restrict__ scratch_block,
                                             const cpx *__restrict__ twN, const LfGeom &g, long long job, int tile, double scale,
                                             int *&sig) {
    constexpr int LF_TC = LfTc<R>::v;
    const int t = threadIdx.x, c = t % LF_TC, i = t / LF_TC;
    const int s = (int)(job / g.nblocks), b = (int)(job % g.nblocks);
    const int col = tile * LF_TC + c;
    cpx *sc = scratch_block + col;
    cpx v[R];
    cpx tw0 = cpx_make(1.0, 0.0), tws = tw0;
    if (TW) { tw0 = twN[col * i]; tws = twN[col * R]; }
    if (!INV) {
        const cpx *st = state + (long long)s * g.state_stride;
        const cpx *ip = in + s * g.in_stride;
        const long long v0 = g.off + (long long)b * g.step + col;
#pragma unroll
        for (int k = 0; k < R; ++k) {
            const long long u = v0 + (long long)LF_N2 * (i + R * k);
            v[k] = lf_ld_stream(u < g.s0 ? st + u : ip + (u - g.s0));
        }
        lf_signal(sig);
        DftR<R, false>::run(v);
#pragma unroll
        for (int q = 0; q < R; ++q) a[(i + R * q) * LF_TC + c] = q ? cpx_mul(v[q], w[i * q]) : v[q];
        __syncthreads();
#pragma unroll
        for (int k = 0; k < R; ++k) v[k] = a[(R * i + k) * LF_TC + c];
        DftR<R, false>::run(v);
        if (TW) {
#pragma unroll
            for (int q = 0; q < R; ++q) { v[q] = cpx_mul(v[q], tw0); if (q + 1 < R) tw0 = cpx_mul(tw0, tws); }
        }
#pragma unroll
        for (int q = 0; q < R; ++q) __stcg(sc + (size_t)(i + R * q) * LF_N2, v[q]);
    } else {
#pragma unroll
        for (int k = 0; k < R; ++k) v[k] = __ldcg(sc + (size_t)(i + R * k) * LF_N2);
        lf_signal(sig);
        if (TW) {
#pragma unroll
            for (int k = 0; k < R; ++k) { v[k] = cpx_mulc(v[k], tw0); if (k + 1 < R) tw0 = cpx_mul(tw0, tws); }
        }
        DftR<R, true>::run(v);
#pragma unroll
        for (int q = 0; q < R; ++q) a[(R * i + q) * LF_TC + c] = v[q];
        __syncthreads();
        // The scratch of this tile has been read for the last time (every thread's loads were consumed above): tell L2 the
        // lines are dead, so that they are dropped instead of written back to DRAM when the output stream pushes them out.
        // One 128-byte line = 8 columns of one row; the next forward pass rewrites whole lines before anything reads them.
        if (LF_DISCARD && (c & 7) == 0) {
#pragma unroll
            for (int k = 0; k < R; ++k) asm volatile("discard.global.L2 [%0], 128;" ::"l"(sc + (size_t)(i + R * k) * LF_N2) : "memory");
        }
#pragma unroll
        for (int q = 0; q < R; ++q) {
            v[q] = a[(i + R * q) * LF_TC + c];
            if (q) v[q] = cpx_mulc(v[q], w[i * q]);
        }
        DftR<R, true>::run(v);
        // v[k] = sample n1 = i + R k of column col; the valid part is n1 >= (FIR - 1) / 2048 (fastfir.cpp:308-311)
        cpx *op = out + s * g.out_stride + g.off + (long long)b * g.step + col;
#pragma unroll
        for (int k = SKIPK; k < R; ++k) lf_st_stream(op + (long long)LF_N2 * (i + R * k - R * SKIPK), cpx_scale(v[k], scale));
    }
}

template <int R, bool INV, int SKIPK, bool TW>
__global__ void __launch_bounds__(LfTc<R>::v *R, R == 16 ? 4 : 1)
lf_cols_kernel(const cpx *__restrict__ state, const cpx *__restrict__ in, cpx *__restrict__ out, cpx *__restrict__ scratch,
               const cpx *__restrict__ twN, LfGeom g, double scale) {
    constexpr int N1 = R * R, LF_TC = LfTc<R>::v;
    __shared__ __align__(16) cpx a[N1 * LF_TC];
    __shared__ __align__(16) cpx w[N1];
    for (int m = threadIdx.x; m < N1; m += LF_TC * R) w[m] = twN[(size_t)m * LF_N2];   // W_N1^m = W_N^(2048 m)
    __syncthreads();
    int *sig = nullptr;
    lf_cols_tile<R, INV, SKIPK, TW>(a, w, state, in, out, scratch + (size_t)blockIdx.y * ((size_t)N1 * LF_N2), twN, g,
                                    g.job0 + blockIdx.y, blockIdx.x, scale, sig);
}

// One row k1 of one block: twiddle, 2048-point forward, x H, inverse, conj twiddle, in place. FILTER: forward only -> H.
template <bool FILTER>
__global__ void __launch_bounds__(FftPlan<LF_N2>::THREADS)
lf_rows_kernel(cpx *__restrict__ scratch, const cpx *__restrict__ hpos, cpx *__restrict__ hout, const cpx *__restrict__ tw2,
               const cpx *__restrict__ twN, int N1) {
    __shared__ __align__(128) cpx a[LF_N2];
    constexpr int SH = FftPlan<LF_N2>::SH;
    const int k1 = blockIdx.x;
    cpx *row = scratch + ((size_t)blockIdx.y * N1 + k1) * LF_N2;
    for (int n2 = threadIdx.x; n2 < LF_N2; n2 += blockDim.x) a[fft_swz<SH>(n2)] = cpx_mul(row[n2], twN[n2 * k1]);
    __syncthreads();
    fft_forward_smem<LF_N2>(a, tw2);
    if (FILTER) {
        for (int p = threadIdx.x; p < LF_N2; p += blockDim.x) hout[(size_t)k1 * LF_N2 + p] = a[fft_swz<SH>(p)];
        return;
    }
    for (int p = threadIdx.x; p < LF_N2; p += blockDim.x)
        a[fft_swz<SH>(p)] = cpx_mul(hpos[(size_t)k1 * LF_N2 + p], a[fft_swz<SH>(p)]);
    __syncthreads();
    fft_inverse_smem<LF_N2>(a, tw2);
    for (int n2 = threadIdx.x; n2 < LF_N2; n2 += blockDim.x) row[n2] = cpx_mulc(a[fft_swz<SH>(n2)], twN[n2 * k1]);
}

// H of every row from the generic kernel's slot order to the fused schedule's [k][thread] order (ff_fused.cuh).
__global__ void lf_permute_kernel(const cpx *__restrict__ hpos, cpx *__restrict__ ht) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;   // idx = k * 128 + t  <->  slot 16 t + k
    const size_t r = (size_t)blockIdx.y * LF_N2;
    ht[r + idx] = hpos[r + FftPlan<LF_N2>::freq2pos(ff2048_pos2freq(16 * (idx & 127) + (idx >> 7)))];
}

// The process-path row kernel on the register-fused 8 x 16 x 16 schedule of fastfir2048_kernel (8 shared-memory accesses
// per point instead of 16): F0 reads the row from scratch (L2), I0 writes it back; the inter-step twiddle W_N^(n2 k1) and
// its conjugate are applied by the column kernels (lf_cols_kernel<.., TW = true>).
__device__ __forceinline__ void lf_rows_fused_row(cpx *a, const cpx *t1, cpx *__restrict__ row, const cpx *__restrict__ ht,
                                                  const cpx *__restrict__ tw2, int *&sig) {
    const int t = threadIdx.x;
    const cpx wa1 = tw2[t], wa2 = tw2[2 * t], wa4 = tw2[4 * t];
    const cpx wb1 = tw2[t + 128], wb2 = tw2[2 * t + 256], wb4 = tw2[4 * t + 512];
#pragma unroll
    for (int g = 0; g < 2; ++g) {
        const int i = t + 128 * g;
        cpx u[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) u[k] = __ldcg(row + i + 256 * k);
        lf_signal(sig);
        dft8<false>(u);
        ff_tw8<false>(u, g ? wb1 : wa1, g ? wb2 : wa2, g ? wb4 : wa4);
#pragma unroll
        for (int k = 0; k < 8; ++k) a[ff_swz(i + 256 * k)] = u[k];
    }
    __syncthreads();
    ff_mid_phases<128, false>(a, t1, ht, t, 0);
#pragma unroll
    for (int g = 0; g < 2; ++g) {
        const int i = t + 128 * g;
        cpx u[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) u[k] = a[ff_swz(i + 256 * k)];
        ff_tw8<true>(u, g ? wb1 : wa1, g ? wb2 : wa2, g ? wb4 : wa4);
        dft8<true>(u);
#pragma unroll
        for (int k = 0; k < 8; ++k) __stcg(row + i + 256 * k, u[k]);
    }
}

template <int MINB>
__global__ void __launch_bounds__(128, MINB)
lf_rows_fused_kernel(cpx *__restrict__ scratch, const cpx *__restrict__ ht_all, const cpx *__restrict__ tw2, int N1) {
    __shared__ __align__(128) cpx a[LF_N2 + 64];
    cpx *t1 = a + LF_N2;
    const int t = threadIdx.x, k1 = blockIdx.x;
    if (t < 64) t1[t] = tw2[8 * ((t & 15) << (t >> 4))];   // t = 16 b + i -> exp(-j 2 pi i 2^b / 256)
    __syncthreads();
    int *sig = nullptr;
    lf_rows_fused_row(a, t1, scratch + ((size_t)blockIdx.y * N1 + k1) * LF_N2, ht_all + (size_t)k1 * LF_N2, tw2, sig);
}

// ---- the three phases as ONE persistent kernel (opt-in: PEBBLE_LF_PIPE_MIN) ------------------------------------------
// The launch triples above are single waves of CTAs that start together and therefore wait on L2 together, compute together
// and drain together; three lanes of them overlap only coarsely. lf_pipe_kernel runs the same tile functions as a dataflow
// pipeline: 4 CTAs per SM take work items from one ordered queue (an atomic counter),
//     group j = [forward column tiles of block j] ++ [rows of block j - LAG] ++ [inverse column tiles of block j - 2 LAG],
// and an item waits (thread 0 spins on a counter, acquire) until the phase it consumes has finished for its block:
//     rows(b)  <- all CT forward tiles of b        inv(b) <- all N1 rows of b        fwd(b) <- all CT inverse tiles of b - RING
// (the last one because block b reuses scratch slot b % RING). Every dependency is an EARLIER item of the queue, taken by a
// CTA that is running and never waits on a later one, so the queue cannot deadlock whatever the number of resident CTAs;
// LAG (1 for the 768-item groups of R = 16, 2 for the 320-item groups of R = 8) puts the dependencies more than the 592
// resident CTAs back in the queue, so in steady state nothing waits; RING >= 2 LAG + 2. The CTAs
// of an SM are at different phases at any time: FP64-bound rows overlap L2-bound column tiles on every SM, not per lane.
// Counters are per ring slot and only ever grow (block b = generation b / RING of its slot), zeroed by the host per launch.
// Scratch is read with ld.cg / written with st.cg (other SMs' data; L1 must not serve a previous generation of the slot).
template <int R> struct LfLag { static constexpr int v = R == 16 ? 1 : 2; };
struct LfPipe {
    int *work;              // next item
    int *done;              // [3][RING]: forward tiles, rows, inverse tiles finished per slot (cumulative)
    volatile int *error;    // host-mapped; set when a wait gave up (PB_LF_SPIN_LIMIT): the NEXT call reports it, nothing hangs
    int ring;
    long long jobs;
};
#define PB_LF_SPIN_LIMIT (1ll << 31)   // ~1 s of clock64 at 2 GHz

__device__ __forceinline__ int lf_ld_relaxed(const int *p) {
    int v;
    asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

template <int R, int SKIPK>
__global__ void __launch_bounds__(128, 4)
lf_pipe_kernel(const cpx *__restrict__ state, const cpx *__restrict__ in, cpx *__restrict__ out, cpx *__restrict__ scratch,
               const cpx *__restrict__ twN, const cpx *__restrict__ tw2, const cpx *__restrict__ ht_all, LfGeom g, double scale,
               LfPipe q) {
    constexpr int N1 = R * R, TC = LfTc<R>::v, CT = LF_N2 / TC, GROUP = 2 * CT + N1, LAG = LfLag<R>::v;
    static_assert(TC * R == 128, "one CTA shape for all three phases");
    __shared__ __align__(128) cpx a[LF_N2 + N1 + 64];
    __shared__ int s_item;
    cpx *w = a + LF_N2, *t1 = a + LF_N2 + N1;
    const int t = threadIdx.x;
    for (int m = t; m < N1; m += 128) w[m] = twN[(size_t)m * LF_N2];
    if (t < 64) t1[t] = tw2[8 * ((t & 15) << (t >> 4))];
    const long long items = (q.jobs + 2 * LAG) * GROUP;
    if (t == 0) s_item = atomicAdd(q.work, 1);
    int *sig = nullptr;   // thread 0: the finished item whose completion has not been published yet
    for (;;) {
        __syncthreads();   // the previous item's stores are issued, its shared memory is free; s_item visible
        const long long item = s_item;
        if (item >= items) break;
        __syncthreads();   // everyone has read s_item
        if (t == 0) s_item = atomicAdd(q.work, 1);   // the next item's number is fetched under this item's work
        const long long grp = item / GROUP;
        const int r = (int)(item - grp * GROUP);
        const int phase = r < CT ? 0 : r < CT + N1 ? 1 : 2;
        const long long job = grp - phase * LAG;
        if (job < 0 || job >= q.jobs) continue;
        const int slot = (int)(job % q.ring);
        const int gen = (int)(job / q.ring);
        cpx *blk = scratch + (size_t)slot * ((size_t)N1 * LF_N2);
        // what this item consumes. The counter is read relaxed (no L1 invalidate): everything a consumer reads from another
        // CTA comes through ld.cg, and the barrier below orders those loads behind this one.
        const int *dep = phase == 0 ? q.done + 2 * q.ring + slot : phase == 1 ? q.done + slot : q.done + q.ring + slot;
        const int need = phase == 0 ? gen * CT : phase == 1 ? (gen + 1) * CT : (gen + 1) * N1;
        if (t == 0 && need > 0 && lf_ld_relaxed(dep) < need) {
            lf_signal(sig);   // never wait while holding back a signal somebody else may be waiting for
            const long long t0 = clock64();
            while (lf_ld_relaxed(dep) < need) {
                __nanosleep(200);
                if (clock64() - t0 > PB_LF_SPIN_LIMIT) { *q.error = 1; __threadfence_system(); break; }
            }
        }
        __syncthreads();
        if (phase == 0) lf_cols_tile<R, false, SKIPK, true>(a, w, state, in, out, blk, twN, g, job, r, 1.0, sig);
        else if (phase == 1) lf_rows_fused_row(a, t1, blk + (size_t)(r - CT) * LF_N2, ht_all + (size_t)(r - CT) * LF_N2, tw2, sig);
        else lf_cols_tile<R, true, SKIPK, true>(a, w, state, in, out, blk, twN, g, job, r - CT - N1, scale, sig);
        // published after the next item's loads are in flight (lf_signal inside the tile functions), or below at the end;
        // the loop-top barrier puts every thread's stores of this item before thread 0's release (cumulativity)
        if (t == 0) sig = q.done + phase * q.ring + slot;
    }
    if (t == 0) lf_signal(sig);
}
}  // namespace

// One transform size: N = N1 x 2048 with N1 = R x R.
struct LfPlan {
    int N, N1, R;
    int skipk;         // (FIR - 1) / (2048 R)
    int step;          // N - (FIR - 1): samples per block
    int chunk;         // blocks per launch triple
    const cpx *d_twN;  // exp(-2 pi j m / N)
    cpx *d_ht;         // H in the fused row kernel's order (lf_permute_kernel)
};

struct LongFir {
    LfPlan base, sup;  // sup.N == 0: no super-block plan
    cpx *d_twN_sup, *d_taps_sup, *d_hpos_sup;
    cpx *d_tw2;        // exp(-2 pi j m / 2048)
    cpx *d_scratch;
    size_t lane_cpx;   // scratch of one lane
    bool fused;        // PEBBLE_LF_GENERIC=1 keeps the generic in-place row kernel (cross-check)
    // Consecutive chunks rotate over `lanes` (default 3) internal streams (own scratch each), forked from and joined to the
    // caller's stream by events: the column kernels are L2-bandwidth-bound and the row kernel FP64/shared-memory-bound,
    // so chunk c's rows overlap chunk c+1's forward columns.
    int lanes;
    int *d_pipe;       // lf_pipe_kernel's queue: [0] next item, [1 .. 3 ring] done counters
    int *h_pipe_err;   // page-locked, device-visible
    int pipe_ring, pipe_min;   // scratch slots of the persistent kernel; fewest blocks of a plan worth launching it for
    cudaStream_t lane[4];
    cudaEvent_t fork, join[4];
};

bool lf_size_supported(int fft_size) { return fft_size == 32768 || fft_size == 131072; }

static void lf_plan_init(LfPlan &pl, int n, int fir_m1, const cpx *d_twN, size_t chunk_bytes) {
    pl.N = n; pl.N1 = n / LF_N2; pl.R = pl.N1 == 256 ? 16 : pl.N1 == 64 ? 8 : 4;
    pl.skipk = fir_m1 / (LF_N2 * pl.R); pl.step = n - fir_m1; pl.d_twN = d_twN; pl.d_ht = nullptr;
    pl.chunk = (int)(chunk_bytes / (sizeof(cpx) * (size_t)n));
    if (pl.chunk < 1) pl.chunk = 1;
}

LongFir *lf_create(int fft_size, const cpx *d_twN) {
    if (!lf_size_supported(fft_size)) return nullptr;
    LongFir *p = new LongFir();
    p->d_twN_sup = p->d_taps_sup = p->d_hpos_sup = nullptr;
    p->d_tw2 = nullptr; p->d_scratch = nullptr; p->fused = !getenv("PEBBLE_LF_GENERIC");
    p->d_pipe = nullptr; p->h_pipe_err = nullptr;
    // default (N = 131072): 9 blocks per chunk and three lanes = 54 MB of scratch in flight, inside the 126 MB L2 together
    // with the streaming input/output. 9 x 64 = 576 row CTAs = one wave of 148 SMs x 4; 1152 column CTAs = 0.86 wave of
    // 148 x 9: every launch is a single wave and three of them (one per lane) are in flight. Sweep in
    // profiles/r1_longfir_history.md: lanes x chunk = 2x18 9.51 ms, 3x12 9.48, 3x9 9.12-9.15, 3x6 9.82, 3x18 12.3 (L2 spill).
    // The super-block plan takes the same bytes: two 8 MB blocks = 512 row CTAs per launch.
    const char *e = getenv("PEBBLE_LF_CHUNK_MB");
    const size_t mb = e && atoi(e) > 0 ? atoi(e) : 18;
    lf_plan_init(p->base, fft_size, fft_size / 2, d_twN, mb << 20);
    if ((e = getenv("PEBBLE_LF_CHUNK_BLOCKS")) && atoi(e) > 0) p->base.chunk = atoi(e);
    p->sup.N = 0; p->sup.d_ht = nullptr;
    const bool want_sup = !getenv("PEBBLE_LF_NOSUPER") && p->fused;
    e = getenv("PEBBLE_LF_LANES");
    p->lanes = e && atoi(e) >= 1 && atoi(e) <= 4 ? atoi(e) : 3;
    for (int l = 0; l < 4; ++l) { p->lane[l] = nullptr; p->join[l] = nullptr; }
    p->fork = nullptr;
    const size_t nb = sizeof(cpx) * (size_t)fft_size, nbs = 4 * nb;
    bool ok = pebble::cuda_ok(cudaMalloc(&p->d_tw2, sizeof(cpx) * LF_N2), "cudaMalloc") &&
              pebble::cuda_ok(cudaMalloc(&p->base.d_ht, nb), "cudaMalloc") &&
              pebble::cuda_ok(cudaMemset(p->base.d_ht, 0, nb), "cudaMemset") &&
              pebble::cuda_ok(cudaEventCreateWithFlags(&p->fork, cudaEventDisableTiming), "cudaEventCreate") &&
              // W_2048^m = W_N^(N1 m): strided copy out of the caller's table
              pebble::cuda_ok(cudaMemcpy2D(p->d_tw2, sizeof(cpx), d_twN, sizeof(cpx) * p->base.N1, sizeof(cpx), LF_N2,
                                           cudaMemcpyDeviceToDevice), "cudaMemcpy2D(tw2)");
    if (ok && want_sup) {
        std::vector<double> tw;
        pebble_host::twiddle_table(4 * fft_size, tw);
        ok = pebble::cuda_ok(cudaMalloc(&p->d_twN_sup, nbs), "cudaMalloc") &&
             pebble::cuda_ok(cudaMalloc(&p->d_taps_sup, nbs), "cudaMalloc") &&
             pebble::cuda_ok(cudaMalloc(&p->d_hpos_sup, nbs), "cudaMalloc") &&
             pebble::cuda_ok(cudaMalloc(&p->sup.d_ht, nbs), "cudaMalloc") &&
             pebble::cuda_ok(cudaMemset(p->sup.d_ht, 0, nbs), "cudaMemset") &&
             pebble::cuda_ok(cudaMemcpy(p->d_twN_sup, tw.data(), nbs, cudaMemcpyHostToDevice), "cudaMemcpy(twN 4N)");
        cpx *ht = p->sup.d_ht;
        if (ok) { lf_plan_init(p->sup, 4 * fft_size, fft_size / 2, p->d_twN_sup, mb << 20); p->sup.d_ht = ht; }
        if (ok && (e = getenv("PEBBLE_LF_SUPER_CHUNK_BLOCKS")) && atoi(e) > 0) p->sup.chunk = atoi(e);
    }
    p->lane_cpx = (size_t)p->base.N * p->base.chunk;
    if (p->sup.N && (size_t)p->sup.N * p->sup.chunk > p->lane_cpx) p->lane_cpx = (size_t)p->sup.N * p->sup.chunk;
    ok = ok && pebble::cuda_ok(cudaMalloc(&p->d_scratch, sizeof(cpx) * p->lane_cpx * p->lanes), "cudaMalloc") &&
         pebble::cuda_ok(cudaMalloc(&p->d_pipe, sizeof(int) * 64), "cudaMalloc") &&
         pebble::cuda_ok(cudaHostAlloc(&p->h_pipe_err, sizeof(int), cudaHostAllocMapped), "cudaHostAlloc");
    if (ok) *p->h_pipe_err = 0;
    // the persistent kernel's ring: as many blocks of the larger plan as the scratch holds (>= 4 for LAG 1, 6 for LAG 2)
    e = getenv("PEBBLE_LF_RING");
    p->pipe_ring = e && atoi(e) >= 3 && atoi(e) <= 16 ? atoi(e) : 0;
    // Opt-in (PEBBLE_LF_PIPE_MIN=<fewest blocks of a plan to launch it for>): measured 6.57 ms per config-5 step against
    // 6.17 ms for the launch triples over three lanes (profiles/r2_longfir_history.md), so the triples stay the default.
    e = getenv("PEBBLE_LF_PIPE_MIN");
    p->pipe_min = !p->fused || !e || atoi(e) <= 0 ? 0 : atoi(e);
    for (int l = 0; ok && l < p->lanes && p->lanes > 1; ++l)
        ok = pebble::cuda_ok(cudaStreamCreateWithFlags(&p->lane[l], cudaStreamNonBlocking), "cudaStreamCreate") &&
             pebble::cuda_ok(cudaEventCreateWithFlags(&p->join[l], cudaEventDisableTiming), "cudaEventCreate");
    if (!ok) { lf_destroy(p); return nullptr; }
    // Keep the scratch in L2: a persisting-access window over each lane's scratch on that lane's stream (the set-aside is
    // a device-wide limit, sized to the scratch and capped by the device maximum). Everything else the kernels touch
    // (input, output, H, twiddles) lies outside the windows and keeps the normal policy. PEBBLE_LF_NOPERSIST=1 disables.
    if (p->lanes > 1 && !getenv("PEBBLE_LF_NOPERSIST")) {
        int dev = 0, max_persist = 0, max_win = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, dev);
        cudaDeviceGetAttribute(&max_win, cudaDevAttrMaxAccessPolicyWindowSize, dev);
        const size_t lane_bytes = sizeof(cpx) * p->lane_cpx;
        size_t want = lane_bytes * p->lanes;
        if (want > (size_t)max_persist) want = (size_t)max_persist;
        if (want > 0 && max_win > 0 && cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want) == cudaSuccess) {
            for (int l = 0; l < p->lanes; ++l) {
                cudaStreamAttrValue v = {};
                v.accessPolicyWindow.base_ptr = p->d_scratch + (size_t)l * p->lane_cpx;
                v.accessPolicyWindow.num_bytes = lane_bytes < (size_t)max_win ? lane_bytes : (size_t)max_win;
                const double share = (double)(want / p->lanes) / (double)v.accessPolicyWindow.num_bytes;
                v.accessPolicyWindow.hitRatio = share < 1.0 ? (float)share : 1.0f;
                v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
                v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
                cudaStreamSetAttribute(p->lane[l], cudaStreamAttributeAccessPolicyWindow, &v);
            }
        }
        cudaGetLastError();   // the policy is an optimisation: a refusal is not an error
    }
    return p;
}

void lf_destroy(LongFir *p) {
    if (!p) return;
    cudaFree(p->d_tw2); cudaFree(p->d_scratch); cudaFree(p->base.d_ht); cudaFree(p->sup.d_ht);
    cudaFree(p->d_twN_sup); cudaFree(p->d_taps_sup); cudaFree(p->d_hpos_sup); cudaFree(p->d_pipe);
    if (p->h_pipe_err) cudaFreeHost(p->h_pipe_err);
    for (int l = 0; l < 4; ++l) {
        if (p->lane[l]) cudaStreamDestroy(p->lane[l]);
        if (p->join[l]) cudaEventDestroy(p->join[l]);
    }
    if (p->fork) cudaEventDestroy(p->fork);
    delete p;
}

int lf_freq2index(int fft_size, int f) {
    const int N1 = fft_size / LF_N2;
    return (f % N1) * LF_N2 + FftPlan<LF_N2>::freq2pos(f / N1);
}

// The launch triples of `jobs` blocks of one plan, chunk by chunk; chunk number `c` (continued across plans) picks the lane.
template <int R, int SKIPK>
static void lf_enqueue(LongFir *p, const LfPlan &pl, const cpx *hpos, cpx *hout, const cpx *state, const cpx *in, cpx *out,
                       LfGeom g, long long jobs, double scale, cudaStream_t st, int lanes, int &c) {
    constexpr int TC = LfTc<R>::v;
    const dim3 cgrid(LF_N2 / TC, 1), cblock(TC * R);
    for (long long j0 = 0; j0 < jobs; j0 += pl.chunk, ++c) {
        const unsigned nj = (unsigned)(jobs - j0 < pl.chunk ? jobs - j0 : pl.chunk);
        cudaStream_t cs = lanes > 1 ? p->lane[c % lanes] : st;
        cpx *scr = p->d_scratch + (size_t)(c % lanes) * p->lane_cpx;
        g.job0 = j0;
        const bool tw = p->fused && !hout;
        if (tw) lf_cols_kernel<R, false, SKIPK, true><<<dim3(cgrid.x, nj), cblock, 0, cs>>>(state, in, nullptr, scr, pl.d_twN, g, 1.0);
        else lf_cols_kernel<R, false, SKIPK, false><<<dim3(cgrid.x, nj), cblock, 0, cs>>>(state, in, nullptr, scr, pl.d_twN, g, 1.0);
        if (hout) {
            lf_rows_kernel<true><<<dim3(pl.N1, nj), FftPlan<LF_N2>::THREADS, 0, cs>>>(scr, nullptr, hout, p->d_tw2, pl.d_twN, pl.N1);
            pebble::count_launch(2);
            continue;
        }
        if (p->fused)
            lf_rows_fused_kernel<4><<<dim3(pl.N1, nj), 128, 0, cs>>>(scr, pl.d_ht, p->d_tw2, pl.N1);
        else
            lf_rows_kernel<false><<<dim3(pl.N1, nj), FftPlan<LF_N2>::THREADS, 0, cs>>>(scr, hpos, nullptr, p->d_tw2, pl.d_twN, pl.N1);
        if (tw) lf_cols_kernel<R, true, SKIPK, true><<<dim3(cgrid.x, nj), cblock, 0, cs>>>(nullptr, nullptr, out, scr, pl.d_twN, g, scale);
        else lf_cols_kernel<R, true, SKIPK, false><<<dim3(cgrid.x, nj), cblock, 0, cs>>>(nullptr, nullptr, out, scr, pl.d_twN, g, scale);
        pebble::count_launch(3);
    }
}

// The same blocks through the persistent kernel, on the caller's stream.
template <int R, int SKIPK>
static bool lf_enqueue_pipe(LongFir *p, const LfPlan &pl, const cpx *state, const cpx *in, cpx *out, LfGeom g, long long jobs,
                            double scale, cudaStream_t st) {
    constexpr int LAG = LfLag<R>::v;
    const size_t blk = (size_t)pl.N, total = p->lane_cpx * p->lanes;
    int ring = (int)(total / blk < 16 ? total / blk : 16);
    if (p->pipe_ring && p->pipe_ring < ring) ring = p->pipe_ring;
    if (ring > 2 * LAG + 2 && !p->pipe_ring) ring = 2 * LAG + 2;
    if (ring < 2 * LAG + 2) return false;
    LfPipe q = {p->d_pipe, p->d_pipe + 1, p->h_pipe_err, ring, jobs};
    cudaMemsetAsync(p->d_pipe, 0, sizeof(int) * (1 + 3 * ring), st);
    g.job0 = 0;
    const long long items = (jobs + 2 * LAG) * (2 * (LF_N2 / LfTc<R>::v) + R * R);
    long long grid = (long long)pebble::num_sms() * 4;
    if (grid > items) grid = items;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3(128); cfg.stream = st;
    cudaLaunchAttribute at[1];
    int nat = 0;
    if (!getenv("PEBBLE_LF_NOPERSIST")) {
        int dev = 0, max_win = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&max_win, cudaDevAttrMaxAccessPolicyWindowSize, dev);
        const size_t bytes = sizeof(cpx) * blk * ring;
        if (max_win > 0) {
            at[0].id = cudaLaunchAttributeAccessPolicyWindow;
            at[0].val.accessPolicyWindow.base_ptr = p->d_scratch;
            at[0].val.accessPolicyWindow.num_bytes = bytes < (size_t)max_win ? bytes : (size_t)max_win;
            at[0].val.accessPolicyWindow.hitRatio = 1.0f;
            at[0].val.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
            at[0].val.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
            nat = 1;
        }
    }
    cfg.attrs = at; cfg.numAttrs = nat;
    const cpx *twN = pl.d_twN, *tw2 = p->d_tw2, *ht = pl.d_ht;
    cpx *scr = p->d_scratch;
    cudaError_t e = cudaLaunchKernelEx(&cfg, lf_pipe_kernel<R, SKIPK>, state, in, out, scr, twN, tw2, ht, g, scale, q);
    pebble::count_launch();
    return e == cudaSuccess;
}

static bool lf_pipe_plan(LongFir *p, const LfPlan &pl, const cpx *state, const cpx *in, cpx *out, const LfGeom &g, long long jobs,
                         double scale, cudaStream_t st) {
    if (!p->pipe_min || jobs < p->pipe_min) return false;
    if (pl.R == 16) return lf_enqueue_pipe<16, 2>(p, pl, state, in, out, g, jobs, scale, st);
    if (pl.R == 8 && pl.skipk == 4) return lf_enqueue_pipe<8, 4>(p, pl, state, in, out, g, jobs, scale, st);
    if (pl.R == 8) return lf_enqueue_pipe<8, 1>(p, pl, state, in, out, g, jobs, scale, st);
    return false;
}

static void lf_enqueue_plan(LongFir *p, const LfPlan &pl, const cpx *hpos, cpx *hout, const cpx *state, const cpx *in, cpx *out,
                            const LfGeom &g, long long jobs, double scale, cudaStream_t st, int lanes, int &c) {
    if (pl.R == 16) lf_enqueue<16, 2>(p, pl, hpos, hout, state, in, out, g, jobs, scale, st, lanes, c);          // 524288 / 65537
    else if (pl.R == 8 && pl.skipk == 4) lf_enqueue<8, 4>(p, pl, hpos, hout, state, in, out, g, jobs, scale, st, lanes, c);   // 131072 / 65537
    else if (pl.R == 8) lf_enqueue<8, 1>(p, pl, hpos, hout, state, in, out, g, jobs, scale, st, lanes, c);       // 131072 / 16385
    else lf_enqueue<4, 2>(p, pl, hpos, hout, state, in, out, g, jobs, scale, st, lanes, c);                      // 32768 / 16385
}

static void lf_fork(LongFir *p, cudaStream_t st, int lanes) {
    if (lanes <= 1) return;
    cudaEventRecord(p->fork, st);
    for (int l = 0; l < lanes; ++l) cudaStreamWaitEvent(p->lane[l], p->fork, 0);
}
static void lf_join(LongFir *p, cudaStream_t st, int lanes) {
    if (lanes <= 1) return;
    for (int l = 0; l < lanes; ++l) {
        cudaEventRecord(p->join[l], p->lane[l]);
        cudaStreamWaitEvent(st, p->join[l], 0);
    }
}

cudaError_t lf_build_filter(LongFir *p, const cpx *d_taps, cpx *d_hpos, cudaStream_t st) {
    // the taps as a one-block stream with no carried state: V = taps
    int c = 0;
    LfGeom g = {1, p->base.step, 0, 0, 0, 0, 0, 0};
    lf_enqueue_plan(p, p->base, nullptr, d_hpos, d_taps, d_taps, nullptr, g, 1, 1.0, st, 1, c);
    lf_permute_kernel<<<dim3(LF_N2 / 256, p->base.N1), 256, 0, st>>>(d_hpos, p->base.d_ht);
    pebble::count_launch();
    if (p->sup.N) {
        // the same taps zero-padded to 4N (d_taps holds N values, zero beyond the FIR length: fastfir.cpp:246-252)
        cudaMemsetAsync(p->d_taps_sup, 0, sizeof(cpx) * (size_t)p->sup.N, st);
        cudaMemcpyAsync(p->d_taps_sup, d_taps, sizeof(cpx) * (size_t)p->base.N, cudaMemcpyDeviceToDevice, st);
        LfGeom gs = {1, p->sup.step, 0, 0, 0, 0, 0, 0};
        lf_enqueue_plan(p, p->sup, nullptr, p->d_hpos_sup, p->d_taps_sup, p->d_taps_sup, nullptr, gs, 1, 1.0, st, 1, c);
        lf_permute_kernel<<<dim3(LF_N2 / 256, p->sup.N1), 256, 0, st>>>(p->d_hpos_sup, p->sup.d_ht);
        pebble::count_launch();
    }
    return cudaGetLastError();
}

cudaError_t lf_process(LongFir *p, const cpx *d_hpos, const cpx *state, int s0, int state_stride, const cpx *in,
                       long long in_stride, cpx *out, long long out_stride, int nblocks, int streams, double scale,
                       cudaStream_t st) {
    if (*p->h_pipe_err) {
        pebble::set_error("fastfir (long): a wait inside lf_pipe_kernel gave up in an earlier call; its output is incomplete");
        *p->h_pipe_err = 0;
        return cudaErrorLaunchFailure;
    }
    // floor(nblocks / 7) super-blocks per stream, then the remaining blocks on the N-point plan behind them
    const int per = p->sup.N ? p->sup.step / p->base.step : 0;   // 7
    const int nsup = per ? nblocks / per : 0, rem = nblocks - nsup * per;
    long long jobs_sup = (long long)nsup * streams, jobs_base = (long long)rem * streams;
    // the unnormalised inverse of the 4N-point transform carries 4N where the reference's carries N: exact factor 1/4
    const double scale_sup = p->sup.N ? scale * ((double)p->base.N / p->sup.N) : scale;
    const LfGeom gs = {nsup, p->sup.step, 0, s0, state_stride, in_stride, out_stride, 0};
    const LfGeom gb = {rem, p->base.step, (long long)nsup * p->sup.step, s0, state_stride, in_stride, out_stride, 0};
    // the persistent kernel where a plan has enough blocks to fill its pipeline (one after the other on the caller's stream:
    // both use the whole scratch as their ring) ...
    if (jobs_sup && lf_pipe_plan(p, p->sup, state, in, out, gs, jobs_sup, scale_sup, st)) jobs_sup = 0;
    if (jobs_base && lf_pipe_plan(p, p->base, state, in, out, gb, jobs_base, scale, st)) jobs_base = 0;
    // ... and launch triples over the lanes for what is left
    const long long chunks = (jobs_sup ? (jobs_sup + p->sup.chunk - 1) / p->sup.chunk : 0) +
                             (jobs_base + p->base.chunk - 1) / p->base.chunk;
    if (!chunks) return cudaGetLastError();
    const int lanes = chunks <= 1 ? 1 : p->lanes;
    int c = 0;
    lf_fork(p, st, lanes);
    if (jobs_sup) lf_enqueue_plan(p, p->sup, nullptr, nullptr, state, in, out, gs, jobs_sup, scale_sup, st, lanes, c);
    if (jobs_base) lf_enqueue_plan(p, p->base, d_hpos, nullptr, state, in, out, gb, jobs_base, scale, st, lanes, c);
    lf_join(p, st, lanes);
    return cudaGetLastError();
}
