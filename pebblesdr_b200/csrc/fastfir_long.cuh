// fastfir_long.cuh — interface of the long-filter CFastFIR path (fastfir_long.cu) used by fastfir.cu for
// CONV_FFT_SIZE > 8192 (BASELINE config 5: 131072-point overlap-save, 65537 taps, 64 concurrent streams).
#pragma once
#include "lib_common.h"
#include "common.cuh"

struct LongFir;
bool lf_size_supported(int fft_size);
LongFir *lf_create(int fft_size, const cpx *d_twN);   // d_twN: exp(-2 pi j m / N), m < N (owned by the caller)
void lf_destroy(LongFir *p);
// taps: N cpx, zero-padded (fastfir.cpp:246-252); hpos receives the filter spectrum in the kernels' [k1][slot] layout
cudaError_t lf_build_filter(LongFir *p, const cpx *d_taps, cpx *d_hpos, cudaStream_t st);
// nblocks overlap-save blocks of every stream: block b of stream s transforms V_s[b L .. b L + N), V_s = state_s ++ in_s,
// and writes out_s[b L .. b L + L) (samples L..N-1 of the inverse transform, fastfir.cpp:308-311) times `scale`.
cudaError_t lf_process(LongFir *p, const cpx *d_hpos, const cpx *state, int s0, int state_stride, const cpx *in,
                       long long in_stride, cpx *out, long long out_stride, int nblocks, int streams, double scale,
                       cudaStream_t st);
// bin f of the spectrum lives at hpos[lf_freq2index(N, f)]
int lf_freq2index(int fft_size, int f);
