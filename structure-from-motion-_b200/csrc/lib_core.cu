// lib_core.cu -- error plumbing, launch accounting and device queries of libsfm_b200.
#include "sfm_common.cuh"
#include <atomic>
#include <stdio.h>
#include <string.h>

namespace sfm {

static char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void record_cuda_error(cudaError_t e, const char *where) {
    snprintf(g_err, sizeof(g_err), "%s: %s (%s)", where, cudaGetErrorString(e), cudaGetErrorName(e));
    // clear the sticky "last error" so that an unrelated later call does not trip on it
    (void)cudaGetLastError();
}

// a validation failure that deserves words (read back through sfm_last_cuda_error like the CUDA ones)
void record_error_text(const char *msg) { snprintf(g_err, sizeof(g_err), "%s", msg ? msg : ""); }

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int sm_count() {
    static int cached = 0;
    if (cached) return cached;
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) {
        (void)cudaGetLastError();
        return 148;  // B200; only reached without a device (grid sizing hint, nothing runs)
    }
    cached = n;
    return n;
}

}  // namespace sfm

extern "C" int sfm_abi_version(void) { return 1; }
extern "C" const char *sfm_last_cuda_error(void) { return sfm::g_err; }
extern "C" int64_t sfm_launch_count(void) { return (int64_t)sfm::g_launches.load(); }
