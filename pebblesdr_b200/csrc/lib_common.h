// lib_common.h — host-side plumbing shared by the C-ABI translation units: error reporting, launch
// accounting, pinned staging. No CPU compute fallback exists anywhere in this library: if CUDA is not
// usable every entry point fails with PEBBLE_ERR_CUDA.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/pebble_cuda.h"

namespace pebble {
void set_error(const char *fmt, ...);
bool cuda_ok(cudaError_t e, const char *what);   // records the error text when e != cudaSuccess
void count_launch(int n = 1);
int num_sms();
// cudaFuncAttributeMaxDynamicSharedMemorySize belongs to the FUNCTION (per device), not to a handle: two handles that launch
// the same kernel with different sizes must never lower each other's limit. Raises it to at least `bytes`, never lowers.
cudaError_t ensure_dynamic_smem(const void *fn, size_t bytes);
// Device entry points whose grids read the input while other CTAs already write the output refuse overlapping buffers.
inline bool ranges_overlap(const void *a, size_t abytes, const void *b, size_t bbytes) {
    const uintptr_t a0 = (uintptr_t)a, b0 = (uintptr_t)b;
    return a0 < b0 + bbytes && b0 < a0 + abytes;
}

// Grow-only pinned host + device scratch used by the host-pointer (compat) entry points.
struct Staging {
    void *h_in = nullptr, *h_out = nullptr, *d_in = nullptr, *d_out = nullptr;
    size_t cap_in = 0, cap_out = 0;
    bool reserve(size_t bytes_in, size_t bytes_out);
    void release();
};
}  // namespace pebble

#define PB_CUDA(call, ret)                                   \
    do {                                                     \
        if (!pebble::cuda_ok((call), #call)) return (ret);   \
    } while (0)
