#include "lib_common.h"
#include <atomic>
#include <map>
#include <mutex>
#include <utility>
#include <cstdarg>
#include <cstdio>

namespace pebble {
static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
bool cuda_ok(cudaError_t e, const char *what) {
    if (e == cudaSuccess) return true;
    set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
    return false;
}
void count_launch(int n) { g_launches += (uint64_t)n; }
int num_sms() {
    // per device ordinal: a process may drive several devices (pebble_set_device)
    static std::atomic<int> sms[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    int v = sms[dev].load(std::memory_order_relaxed);
    if (!v) {
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
        sms[dev].store(v, std::memory_order_relaxed);
    }
    return v;
}
cudaError_t ensure_dynamic_smem(const void *fn, size_t bytes) {
    static std::mutex mu;
    static std::map<std::pair<int, const void *>, size_t> limit;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lk(mu);
    size_t &cur = limit[std::make_pair(dev, fn)];
    if (bytes <= cur) return cudaSuccess;
    e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e == cudaSuccess) cur = bytes;
    return e;
}
bool Staging::reserve(size_t bi, size_t bo) {
    if (bi > cap_in) {
        if (h_in) cudaFreeHost(h_in);
        if (d_in) cudaFree(d_in);
        h_in = d_in = nullptr; cap_in = 0;
        if (!cuda_ok(cudaMallocHost(&h_in, bi), "cudaMallocHost(in)")) return false;
        if (!cuda_ok(cudaMalloc(&d_in, bi), "cudaMalloc(in)")) return false;
        cap_in = bi;
    }
    if (bo > cap_out) {
        if (h_out) cudaFreeHost(h_out);
        if (d_out) cudaFree(d_out);
        h_out = d_out = nullptr; cap_out = 0;
        if (!cuda_ok(cudaMallocHost(&h_out, bo), "cudaMallocHost(out)")) return false;
        if (!cuda_ok(cudaMalloc(&d_out, bo), "cudaMalloc(out)")) return false;
        cap_out = bo;
    }
    return true;
}
void Staging::release() {
    if (h_in) cudaFreeHost(h_in);
    if (h_out) cudaFreeHost(h_out);
    if (d_in) cudaFree(d_in);
    if (d_out) cudaFree(d_out);
    h_in = h_out = d_in = d_out = nullptr;
    cap_in = cap_out = 0;
}
}  // namespace pebble

extern "C" {
const char *pebble_last_error(void) { return pebble::g_err; }
int pebble_abi_version(void) { return 1; }
int pebble_device_count(void) {
    int n = 0;
    PB_CUDA(cudaGetDeviceCount(&n), PEBBLE_ERR_CUDA);
    return n;
}
int pebble_set_device(int ordinal) {
    PB_CUDA(cudaSetDevice(ordinal), PEBBLE_ERR_CUDA);
    return 0;
}
int pebble_device_synchronize(void) {
    PB_CUDA(cudaDeviceSynchronize(), PEBBLE_ERR_CUDA);
    return 0;
}
uint64_t pebble_launch_count(void) { return pebble::g_launches.load(); }
}
