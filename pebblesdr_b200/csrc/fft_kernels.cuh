// fft_kernels.cuh — device-side drivers for the in-place shared-memory FFT (fft_core.cuh):
// run all passes of a plan with CTA barriers in between.
#pragma once
#include "fft_core.cuh"

#if defined(__CUDACC__)
template <int N, int P> struct FftRun {
    static PB_D void fwd(cpx *a, const cpx *tw) {
        FftRun<N, P - 1>::fwd(a, tw);
        FftPlan<N>::template pass<P, false>(a, tw, threadIdx.x, blockDim.x);
        __syncthreads();
    }
    static PB_D void inv(cpx *a, const cpx *tw) {
        FftPlan<N>::template pass<P, true>(a, tw, threadIdx.x, blockDim.x);
        __syncthreads();
        FftRun<N, P - 1>::inv(a, tw);
    }
};
template <int N> struct FftRun<N, -1> {
    static PB_D void fwd(cpx *, const cpx *) {}
    static PB_D void inv(cpx *, const cpx *) {}
};
// a[] must be complete (barrier) before the call; complete again (barrier) on return.
template <int N> PB_D void fft_forward_smem(cpx *a, const cpx *tw) { FftRun<N, FftPlan<N>::NP - 1>::fwd(a, tw); }
template <int N> PB_D void fft_inverse_smem(cpx *a, const cpx *tw) { FftRun<N, FftPlan<N>::NP - 1>::inv(a, tw); }
#endif
