// Runtime glue: thread-local error string, launch counter, device query.
#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "common.cuh"

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void b200_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

void b200_count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

extern "C" const char* b200_last_error(void) { return g_err; }
extern "C" uint64_t b200_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

extern "C" int b200_device_info(int* out4) {
    int dev = 0;
    B200_CUDA_OK(cudaGetDevice(&dev));
    B200_CUDA_OK(cudaDeviceGetAttribute(&out4[0], cudaDevAttrMultiProcessorCount, dev));
    B200_CUDA_OK(cudaDeviceGetAttribute(&out4[1], cudaDevAttrComputeCapabilityMajor, dev));
    B200_CUDA_OK(cudaDeviceGetAttribute(&out4[2], cudaDevAttrComputeCapabilityMinor, dev));
    B200_CUDA_OK(cudaDeviceGetAttribute(&out4[3], cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    return 0;
}

int b200_gemm_init();

// Forces module load of every kernel (CUDA loads lazily) and applies per-kernel
// attributes.  Called once from the Python binding before anything is captured.
extern "C" int b200_init(void) {
    B200_CUDA_OK(cudaFree(0));
    return b200_gemm_init();
}
