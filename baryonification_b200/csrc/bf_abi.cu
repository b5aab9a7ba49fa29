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

// Number of kernel launches of THIS library since load (every launch site passes through BF_LAUNCH_OK): what
// bench.py reports as `gpu_launches` -- counted, not assumed.  CUB's own kernels (radix sort, scan) are not included.
static unsigned long long g_launches = 0;
void bf_count_launch() { __atomic_add_fetch(&g_launches, 1ull, __ATOMIC_RELAXED); }
extern "C" unsigned long long bf_launch_count(void) { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }

// fp64 throughput probe: 8 independent DFMA chains per thread, one CTA of 256 threads per SM slot.  bench.py measures
// the device's fp64 instruction rate with it, live, as the denominator of the pair kernels' fp64 roofline.
__global__ void __launch_bounds__(256) probe_fp64_kernel(int iters, double *out) {
    double a[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = 1.0 + 1e-3 * (double)(threadIdx.x + k);
    const double b = 1.0 - 1e-9, c = 1e-9;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = fma(a[k], b, c);
    }
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += a[k];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

extern "C" int bf_probe_fp64(int blocks, int iters, double *out, unsigned long long *dfma_per_thread_host, void *stream) {
    BF_REQUIRE(blocks > 0 && iters > 0 && out, "blocks / iters / out");
    probe_fp64_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(iters, out);
    BF_LAUNCH_OK();
    if (dfma_per_thread_host) *dfma_per_thread_host = 8ull * (unsigned long long)iters;
    return BF_OK;
}

extern "C" int bf_device_sm_count(int *sm_count_host) {
    BF_REQUIRE(sm_count_host != nullptr, "sm_count_host is NULL");
    int dev = 0, n = 0;
    BF_CUDA_OK(cudaGetDevice(&dev));
    BF_CUDA_OK(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
    *sm_count_host = n;
    return BF_OK;
}

// Page-locking a caller's host array where it lies (the Python side registers large inputs that are processed
// repeatedly, _device.pin_in_place).  The CUDA runtime keeps a failed call's error code until somebody asks for it, and
// the next unrelated check (torch's, ours) would then report it -- so these two wrappers always leave the error state
// clean.  bf_host_pin: 0 = registered now, 1 = already page-locked (cudaHostAlloc / registered before: nothing to do),
// BF_E_INVALID = could not be registered.  bf_host_unpin undoes a 0.
extern "C" int bf_host_pin(void *ptr, size_t bytes) {
    BF_REQUIRE(ptr != nullptr && bytes > 0, "ptr / bytes");
    cudaPointerAttributes at;
    cudaError_t e = cudaPointerGetAttributes(&at, ptr);
    if (e == cudaSuccess && at.type == cudaMemoryTypeHost) {
        cudaPointerAttributes last;
        e = cudaPointerGetAttributes(&last, (const char *)ptr + (bytes - 1));
        if (e == cudaSuccess && last.type == cudaMemoryTypeHost) return 1;
    }
    (void)cudaGetLastError();
    e = cudaHostRegister(ptr, bytes, cudaHostRegisterDefault);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        bf_set_error("bf_host_pin: cudaHostRegister(%p, %zu) -> %s", ptr, bytes, cudaGetErrorString(e));
        return BF_E_INVALID;
    }
    return BF_OK;
}

extern "C" int bf_host_unpin(void *ptr) {
    BF_REQUIRE(ptr != nullptr, "ptr");
    cudaError_t e = cudaHostUnregister(ptr);
    (void)cudaGetLastError();
    if (e != cudaSuccess) {
        bf_set_error("bf_host_unpin: cudaHostUnregister(%p) -> %s", ptr, cudaGetErrorString(e));
        return BF_E_INVALID;
    }
    return BF_OK;
}
