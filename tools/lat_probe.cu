// Development aid: dependent-issue latency of the FP32 operations on the PLL chain (one warp, clock64 around N dependent ops).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/lat_probe tools/lat_probe.cu ; run: tools/lat_probe
#include <cstdio>
#include <cuda_runtime.h>
#define N 4096
template <int KIND> __global__ void k(float *out, long long *cyc, float a, float b, float lo, float hi) {
    float x = a, y = b;
    long long t0 = clock64();
#pragma unroll 64
    for (int i = 0; i < N; ++i) {
        if (KIND == 0) x = fmaf(x, a, b);                                   // FFMA chain
        if (KIND == 1) x = x + a;                                           // FADD chain
        if (KIND == 2) x = fminf(fmaxf(x, lo), hi) + a;                     // FMNMX, FMNMX, FADD
        if (KIND == 3) { x = fmaf(x, a, b); x = fmaxf(x, lo); }             // FFMA -> FMNMX (cross pipe)
        if (KIND == 4) { float q = fmaf(x, 0.15915494f, y); float kk = (q + 12582912.0f) - 12582912.0f; float r = fmaf(kk, -6.28125f, x);
                         r = fmaf(kk, -1.9353072e-3f, r); float e = y + r; float f = fmaf(a, e, b); f = fminf(fmaxf(f, lo), hi);
                         x = x + fmaf(a, e, f); }                           // the PLL chain (11 ops), constant inputs
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) { *out = x; *cyc = t1 - t0; }
}
// the walk's inner block in stages: what, beyond the bare chain, costs cycles per step?
//   STAGE 1: chain fed by width-4 shuffles of lane-held inputs (2 SHFL + FMUL + FSETP/FSEL per step)
//   STAGE 2: + the DC tracker and the output scaling        STAGE 3: + "keep my own step's output" selects
template <int STAGE> __global__ void walk(float *out, long long *cyc, float beta, float alpha, float lo, float hi, float dca, float gain, int blocks) {
    const int lane = threadIdx.x, s = lane % 4;
    float phase = 0.1f, freq = 0.01f, dc = 0.0f, acc = 0.0f;
    float th[8], nb[8], mine[8];
    for (int m = 0; m < 8; ++m) { th[m] = 0.3f * (lane + m) - 2.0f; nb[m] = -beta; mine[m] = 0.f; }
    long long t0 = clock64();
    for (int b = 0; b < blocks; ++b) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const float tj = __shfl_sync(0xffffffffu, th[j / 4], j % 4, 4);
            const float nbj = __shfl_sync(0xffffffffu, nb[j / 4], j % 4, 4);
            const float na = nbj == 0.0f ? 0.0f : -alpha;
            const float q = fmaf(phase, 0.15915494f, tj * 0.15915494f);
            const float kk = (q + 12582912.0f) - 12582912.0f;
            float r = fmaf(kk, -6.28125f, phase);
            r = fmaf(kk, -1.9353072e-3f, r);
            const float e = tj + r;
            float f = fmaf(nbj, e, freq);
            f = fminf(fmaxf(f, lo), hi);
            freq = f;
            phase = phase + fmaf(na, e, f);
            if (STAGE >= 2) {
                dc = fmaf(dca, f, (1.0f - dca) * dc);
                const float o = (f - dc) * gain;
                if (STAGE >= 3) mine[j / 4] = s == j % 4 ? o : mine[j / 4];
                else acc += o;
            }
        }
        for (int m = 0; m < 8; ++m) { th[m] += 0.01f; acc += mine[m]; }
    }
    long long t1 = clock64();
    if (lane == 0) { *out = phase + freq + dc + acc; *cyc = t1 - t0; }
}
__global__ void dfma_chain(double *out, long long *cyc, double a, double b) {
    double x = a;
    long long t0 = clock64();
#pragma unroll 64
    for (int i = 0; i < N; ++i) x = fma(x, a, b);
    long long t1 = clock64();
    if (threadIdx.x == 0) { *out = x; *cyc = t1 - t0; }
}
// the AM walk's inner block: d = alpha d + m_j with the magnitudes shuffled in, w = dn - d kept by the owner lane
__global__ void am_walk(double *out, long long *cyc, double alpha, int blocks) {
    const int lane = threadIdx.x, s = lane % 4;
    double d = 0.1, acc = 0.0, mg[8], mine[8];
    for (int m = 0; m < 8; ++m) { mg[m] = 0.3 * (lane + m); mine[m] = 0.0; }
    long long t0 = clock64();
    for (int b = 0; b < blocks; ++b) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const double mj = __shfl_sync(0xffffffffu, mg[j / 4], j % 4, 4);
            const double dn = (alpha * d) + mj;
            const double w = dn - d;
            d = dn;
            mine[j / 4] = s == j % 4 ? w : mine[j / 4];
        }
        for (int m = 0; m < 8; ++m) { mg[m] += 0.01; acc += mine[m]; }
    }
    long long t1 = clock64();
    if (lane == 0) { *out = d + acc; *cyc = t1 - t0; }
}
int main() {
    {
        double *o; long long *c, h;
        cudaMalloc(&o, 8); cudaMalloc(&c, 8);
        for (int rep = 0; rep < 2; ++rep) { dfma_chain<<<1, 32>>>(o, c, 0.999, 1e-3); cudaDeviceSynchronize(); }
        cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
        printf("DFMA chain: %.2f cycles per dependent op\n", (double)h / N);
        for (int rep = 0; rep < 2; ++rep) { am_walk<<<1, 32>>>(o, c, 0.9999, 128); cudaDeviceSynchronize(); }
        cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
        printf("AM walk inner block: %.1f cycles per step\n", (double)h / (128 * 32));
    }
    {
        float *o; long long *c, h;
        cudaMalloc(&o, 4); cudaMalloc(&c, 8);
        for (int st = 1; st <= 3; ++st) {
            for (int rep = 0; rep < 2; ++rep) {
                if (st == 1) walk<1><<<1, 32>>>(o, c, 0.05f, 0.3f, -1.5f, 1.5f, 0.01f, 1.0f, 128);
                if (st == 2) walk<2><<<1, 32>>>(o, c, 0.05f, 0.3f, -1.5f, 1.5f, 0.01f, 1.0f, 128);
                if (st == 3) walk<3><<<1, 32>>>(o, c, 0.05f, 0.3f, -1.5f, 1.5f, 0.01f, 1.0f, 128);
                cudaDeviceSynchronize();
            }
            cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
            printf("walk stage %d: %.1f cycles per step\n", st, (double)h / (128 * 32));
        }
    }

    float *o; long long *c, h;
    cudaMalloc(&o, 4); cudaMalloc(&c, 8);
    const char *names[5] = {"FFMA chain", "FADD chain", "FMNMX+FMNMX+FADD", "FFMA->FMNMX", "PLL chain (11 ops)"};
    const int ops[5] = {1, 1, 3, 2, 11};
    for (int kind = 0; kind < 5; ++kind) {
        for (int rep = 0; rep < 2; ++rep) {
            if (kind == 0) k<0><<<1, 32>>>(o, c, 0.999f, 1e-3f, -1.f, 1.f);
            if (kind == 1) k<1><<<1, 32>>>(o, c, 1e-3f, 0.f, -1.f, 1.f);
            if (kind == 2) k<2><<<1, 32>>>(o, c, 1e-3f, 0.f, -1.f, 1.f);
            if (kind == 3) k<3><<<1, 32>>>(o, c, 0.999f, 1e-3f, -1.f, 1.f);
            if (kind == 4) k<4><<<1, 32>>>(o, c, -0.01f, 0.3f, -1.5f, 1.5f);
            cudaDeviceSynchronize();
        }
        cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
        printf("%-22s %.2f cycles per iteration, %.2f per dependent op\n", names[kind], (double)h / N, (double)h / N / ops[kind]);
    }
    return 0;
}
