// cta_ctx.cuh — the GPU execution context for the Ctx-templated kernel bodies (dc_core.cuh):
// threads = the CTA, sync = __syncthreads, chunk loads = one cp.async.bulk (TMA bulk copy, SASS UBLKCP)
// per chunk completing on an mbarrier, double-buffered.
#pragma once
#include "common.cuh"

#if defined(__CUDACC__)
struct CtaCtx {
    uint64_t *mbar;       // two mbarriers in shared memory
    uint32_t parity;      // bit b = phase parity to wait for on buffer b

    PB_HD int tid() const {
#if defined(__CUDA_ARCH__)
        return threadIdx.x;
#else
        return 0;
#endif
    }
    PB_HD int nthreads() const {
#if defined(__CUDA_ARCH__)
        return blockDim.x;
#else
        return 1;
#endif
    }
    PB_HD void sync() {
#if defined(__CUDA_ARCH__)
        __syncthreads();
#endif
    }
    // Called by every thread (after a sync that retired all generic-proxy accesses to dst).
    PB_HD void load_chunk(int buf, cpx *dst, const cpx *src, int len) {
#if defined(__CUDA_ARCH__)
        if (threadIdx.x == 0) {
            const uint32_t bytes = (uint32_t)len * (uint32_t)sizeof(cpx);
            const uint32_t bar = (uint32_t)__cvta_generic_to_shared(mbar + buf);
            const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
            asm volatile(
                "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(d),
                "l"(src), "r"(bytes), "r"(bar)
                : "memory");
        }
#endif
    }
    PB_HD void wait_chunk(int buf) {
#if defined(__CUDA_ARCH__)
        const uint32_t bar = (uint32_t)__cvta_generic_to_shared(mbar + buf);
        const uint32_t ph = (parity >> buf) & 1u;
        uint32_t done = 0;
        while (!done) {
            asm volatile(
                "{\n\t.reg .pred p;\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                "selp.u32 %0, 1, 0, p;\n\t}"
                : "=r"(done)
                : "r"(bar), "r"(ph)
                : "memory");
        }
        parity ^= (1u << buf);
#endif
    }
    // one thread initialises both barriers (arrival count 1: the expect_tx arrive of the issuing thread)
    PB_D void init() {
        if (threadIdx.x == 0) {
            for (int b = 0; b < 2; ++b) {
                const uint32_t bar = (uint32_t)__cvta_generic_to_shared(mbar + b);
                asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
            }
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        parity = 0;
        __syncthreads();
    }
};
#endif
