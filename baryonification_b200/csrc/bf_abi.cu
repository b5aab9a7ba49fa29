// bf_abi.cu -- error string, version and device queries of the C ABI.
#include <stdarg.h>
#include <string.h>

#include "bf_common.cuh"

static thread_local char g_err[512] = "";

void bf_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char *bf_last_error(void) { return g_err; }
extern "C" int bf_version(void) { return 100; }

extern "C" int bf_device_sm_count(int *sm_count_host) {
    BF_REQUIRE(sm_count_host != nullptr, "sm_count_host is NULL");
    int dev = 0, n = 0;
    BF_CUDA_OK(cudaGetDevice(&dev));
    BF_CUDA_OK(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
    *sm_count_host = n;
    return BF_OK;
}
