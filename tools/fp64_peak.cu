// tools/fp64_peak.cu — measures the FP64 FMA peak of the GPU (the denominator of every "FP64 roof" this repo quotes).
// Dependent-free DFMA loop: 8 independent accumulators per thread (FP64 FMA latency on sm_100 is ~8 cycles at one
// warp-instruction per 2 cycles per sub-partition, so 4 accumulators x 1+ warps already saturate; 8 and 8 warps per
// sub-partition leave margin), every SM filled with 2 CTAs x 512 threads, launches repeated for >= `ms` milliseconds,
// timed with CUDA events. Prints one JSON object; tools/measure_fp64_peak.py adds the clocks record.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

template <int ACC>
__global__ void __launch_bounds__(512) dfma_kernel(double *out, double a, double b, int iters) {
    double acc[ACC];
#pragma unroll
    for (int k = 0; k < ACC; ++k) acc[k] = (double)(threadIdx.x + k);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int k = 0; k < ACC; ++k) acc[k] = fma(acc[k], a, b);
        }
    }
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < ACC; ++k) s += acc[k];
    if (s == 123.456) out[blockIdx.x * blockDim.x + threadIdx.x] = s;   // never true: keeps the loop alive
}

int main(int argc, char **argv) {
    const double want_ms = argc > 1 ? atof(argv[1]) : 400.0;
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, 0) != cudaSuccess) { fprintf(stderr, "no CUDA device\n"); return 2; }
    const int sms = p.multiProcessorCount, ctas = sms * 2, threads = 512, iters = 4096;
    constexpr int ACC = 8;
    double *d_out = nullptr;
    cudaMalloc(&d_out, sizeof(double) * ctas * threads);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int w = 0; w < 3; ++w) dfma_kernel<ACC><<<ctas, threads>>>(d_out, 0.999999, 1e-9, iters);
    cudaDeviceSynchronize();
    // one launch timed alone (burst), then back-to-back launches for >= want_ms (sustained)
    cudaEventRecord(e0);
    dfma_kernel<ACC><<<ctas, threads>>>(d_out, 0.999999, 1e-9, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms1 = 0.f;
    cudaEventElapsedTime(&ms1, e0, e1);
    const double flop_per_launch = 2.0 * ACC * 8.0 * iters * (double)ctas * threads;
    int reps = (int)(want_ms / ms1) + 1;
    cudaEventRecord(e0);
    for (int r = 0; r < reps; ++r) dfma_kernel<ACC><<<ctas, threads>>>(d_out, 0.999999, 1e-9, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float msN = 0.f;
    cudaEventElapsedTime(&msN, e0, e1);
    if (cudaGetLastError() != cudaSuccess) { fprintf(stderr, "kernel failed\n"); return 3; }
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"ctas\": %d, \"threads\": %d, \"accumulators\": %d, \"burst_ms\": %.4f, "
           "\"fp64_tflops_burst\": %.3f, \"sustained_ms\": %.2f, \"launches\": %d, \"fp64_tflops_sustained\": %.3f, "
           "\"clock_khz_prop\": %d}\n",
           p.name, sms, ctas, threads, ACC, ms1, flop_per_launch / (ms1 * 1e-3) / 1e12, msN, reps,
           flop_per_launch * reps / (msN * 1e-3) / 1e12, p.clockRate);
    return 0;
}
