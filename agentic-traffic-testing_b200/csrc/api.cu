// Error plumbing and device queries of the C ABI.
#include "common.cuh"

#include <cstdarg>
#include <cstdio>

#include "../../include/atx_b200.h"

namespace atx {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
    set_error("CUDA error %d (%s) at %s", static_cast<int>(e), cudaGetErrorString(e), what);
    return ATX_ERR_CUDA;
}

int num_sms() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

}  // namespace atx

extern "C" {

const char* atx_last_error(void) { return atx::g_err; }

int atx_version(void) { return ATX_VERSION; }

int atx_device_info(int* sm_count, int* cc_major, int* cc_minor) {
    int dev = 0;
    ATX_CUDA(cudaGetDevice(&dev));
    int v = 0;
    if (sm_count) {
        ATX_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev));
        *sm_count = v;
    }
    if (cc_major) {
        ATX_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMajor, dev));
        *cc_major = v;
    }
    if (cc_minor) {
        ATX_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMinor, dev));
        *cc_minor = v;
    }
    return ATX_OK;
}

}  // extern "C"
