// Error reporting and device queries shared by all entry points.
#include <stdarg.h>
#include <stdlib.h>
#include <time.h>

#include "common.cuh"

namespace hc {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}

double wall_seconds() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (!cached[dev]) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

}  // namespace hc

extern "C" int hc_version(void) { return HC_VERSION; }

extern "C" int hc_last_error(char *h_buf, size_t n) {
    if (!h_buf || !n) return HC_EINVAL;
    strncpy(h_buf, hc::g_err, n - 1);
    h_buf[n - 1] = 0;
    return HC_OK;
}

extern "C" int hc_device_info(int *h_sm_count, size_t *h_total_mem, int *h_cc_major, int *h_cc_minor) {
    int dev = 0;
    HC_CUDA_OK(cudaGetDevice(&dev));
    cudaDeviceProp p;
    HC_CUDA_OK(cudaGetDeviceProperties(&p, dev));
    if (h_sm_count) *h_sm_count = p.multiProcessorCount;
    if (h_total_mem) *h_total_mem = p.totalGlobalMem;
    if (h_cc_major) *h_cc_major = p.major;
    if (h_cc_minor) *h_cc_minor = p.minor;
    return HC_OK;
}

extern "C" int hc_set_device(int device) {
    HC_CUDA_OK(cudaSetDevice(device));
    return HC_OK;
}
