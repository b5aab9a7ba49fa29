// bf_common.cuh -- shared host/device helpers of the sm_100a kernels.
//
// The whole library is compiled with -fmad=false: every a*b+c in index-determining arithmetic
// (ring ranges, interpolation-neighbour indices, table cells) then rounds twice, exactly like the
// reference's numpy / healpix_cxx arithmetic, so pixel lists are bit-identical.  Value arithmetic
// that wants a fused multiply-add calls fma() explicitly.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/baryonification_b200.h"

typedef long long i64;
typedef unsigned long long u64;

void bf_set_error(const char *fmt, ...);

#define BF_CUDA_OK(expr)                                                              \
    do {                                                                              \
        cudaError_t _e = (expr);                                                      \
        if (_e != cudaSuccess) {                                                      \
            bf_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return (int)_e;                                                           \
        }                                                                             \
    } while (0)

#define BF_REQUIRE(cond, msg)                                        \
    do {                                                             \
        if (!(cond)) {                                               \
            bf_set_error("%s:%d invalid argument: %s", __FILE__, __LINE__, msg); \
            return BF_E_INVALID;                                     \
        }                                                            \
    } while (0)

void bf_count_launch();  // bf_abi.cu: every kernel launch of the library is counted (bf_launch_count)
#define BF_LAUNCH_OK()                     \
    do {                                   \
        bf_count_launch();                 \
        BF_CUDA_OK(cudaGetLastError());    \
    } while (0)

static inline int bf_sm_count() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}

// ------------------------------------------------------------------------------------------
// r-axis descriptor of a table.  scipy's interval search (find_interval_ascending: grid[i] <= x <
// grid[i+1], last interval closed, out of range -> NaN fill) is reproduced exactly: a uniform-grid
// guess floor((x - g0) * inv_dx) is corrected against the stored grid, so the cell equals bisection.
// ------------------------------------------------------------------------------------------
struct bf_raxis {
    const double *g; // [nr] ln r grid (global or shared)
    int nr;
    double g0, glast, inv_dx;
};

__device__ __forceinline__ int bf_find_cell(const bf_raxis &ax, double x) {
    // precondition: g0 <= x <= glast
    int k = (int)((x - ax.g0) * ax.inv_dx);
    k = max(0, min(k, ax.nr - 2));
    while (k > 0 && x < ax.g[k]) --k;
    while (k < ax.nr - 2 && x >= ax.g[k + 1]) ++k;
    return k;
}

// Same cell with one corrective step each way (enough whenever the uniform-grid guess is within +-1 of the
// bisection cell, which holds for np.geomspace axes); falls back to the search loops otherwise.
__device__ __forceinline__ int bf_find_cell_fast(const bf_raxis &ax, double x) {
    int k = (int)((x - ax.g0) * ax.inv_dx);
    k = max(0, min(k, ax.nr - 2));
    k -= (k > 0 && x < ax.g[k]) ? 1 : 0;
    k += (k < ax.nr - 2 && x >= ax.g[k + 1]) ? 1 : 0;
    if (x < ax.g[k] || (k < ax.nr - 2 && x >= ax.g[k + 1])) {  // non-uniform axis: search
        while (k > 0 && x < ax.g[k]) --k;
        while (k < ax.nr - 2 && x >= ax.g[k + 1]) ++k;
    }
    return k;
}

// ------------------------------------------------------------------------------------------
// ln(x) for finite x > 0 from a 128-entry shared-memory table: x = 2^e m, m in [1, 2); the top 7 mantissa
// bits pick c with |m / c - 1| <= 2^-8, and ln m = ln c' + log1p(r), r = m * (1/c)' - 1 (one FMA, exact for the
// ROUNDED reciprocal, whose own logarithm the table stores), log1p by its degree-6 series (truncation 2e-18).
// Absolute error ~2e-16 * (1 + |e|): the table coordinate ln r only needs ~1e-13.  ~14 fp64 instructions
// instead of the ~80 of the correctly rounded library log.  x = 0 or subnormal gives a large negative number
// (below any table), which is all the callers need.
// ------------------------------------------------------------------------------------------
#define BF_LOGTAB 128
__device__ __forceinline__ void bf_logtab_init(double2 *tab, int t, int nthreads) {
    for (int i = t; i < BF_LOGTAB; i += nthreads) {
        double ic = 1.0 / (1.0 + ((double)i + 0.5) / (double)BF_LOGTAB);
        tab[i] = make_double2(ic, -log(ic));
    }
}

// Polynomial coefficients of the pair loops live in constant memory: an fp64 instruction takes a c[bank][offset]
// operand directly, whereas a 64-bit literal costs two MOVs per use once the register budget is spent (the hot
// loops run at 64 registers per thread, so ptxas rematerialises them every iteration).
static __constant__ double bf_kc[8] = {-1.0 / 6.0, 0.2, 1.0 / 3.0, 0.693147180559945309417232121458, 0.375, 0.2734375,
                                       -0.3125, 1e-3};

__device__ __forceinline__ double bf_fastlog(const double2 *__restrict__ tab, double x) {
    const int hi = __double2hiint(x);
    const int e = (hi >> 20) - 1023;
    const double m = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(x));
    const double2 t = tab[(hi >> 13) & (BF_LOGTAB - 1)];
    const double r = fma(m, t.x, -1.0);
    double p = fma(r, bf_kc[0], bf_kc[1]);
    p = fma(r, p, -0.25);
    p = fma(r, p, bf_kc[2]);
    p = fma(r, p, -0.5);
    p = fma(r, p, 1.0);
    return fma((double)e, bf_kc[3], fma(r, p, t.y));
}

// 1 / sqrt(x) for normal x > 0: the hardware seed (MUFU.RSQ64H, 2^-22) refined by one third-order step -- the
// instruction sequence of the library rsqrt() without its special-case branch.  x = 0 gives NaN (inf seed), which
// every caller turns into 0 like the reference's `offset[~isfinite(offset)] = 0`.
__device__ __forceinline__ double bf_rsqrt(double x) {
    double y0;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y0) : "d"(x));
    const double e = fma(x, -(y0 * y0), 1.0);
    const double p = fma(e, bf_kc[4], 0.5);
    return fma(p, y0 * e, y0);
}

// ------------------------------------------------------------------------------------------
// TMA bulk copy (cp.async.bulk, global -> shared, completion on an mbarrier): the per-call constant tables of a pair
// kernel (ln table, radial axis) are prepared ONCE in global memory and every CTA stages them with a single
// asynchronous bulk copy issued by one thread, instead of 256 threads recomputing them with divisions and logarithms.
// Sizes and addresses must be multiples of 16 bytes.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned bf_smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void bf_mbar_init(u64 *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bf_smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");  // visible to the async proxy
}

__device__ __forceinline__ void bf_tma_load_1d(void *dst_smem, const void *src_gmem, unsigned bytes, u64 *bar) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bf_smem_u32(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     bf_smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(bf_smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ void bf_mbar_wait(u64 *bar, unsigned parity) {
    unsigned done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(bf_smem_u32(bar)), "r"(parity)
                     : "memory");
    } while (!done);
}

// value of one collapsed row at ln r = x; NaN outside the table (fill_value).  The two-product form
// keeps scipy's inf/NaN propagation: a zero weight still multiplies (0 * -inf = NaN).
__device__ __forceinline__ double bf_row_lerp(const bf_raxis &ax, const double *__restrict__ row, double x) {
    if (!(x >= ax.g0 && x <= ax.glast)) return nan("");
    int k = bf_find_cell(ax, x);
    double gk = ax.g[k];
    double t = (x - gk) / (ax.g[k + 1] - gk);
    return row[k] * (1.0 - t) + row[k + 1] * t;
}

static inline bf_raxis bf_make_raxis(const bf_table *t) {
    bf_raxis ax;
    ax.g = t->gr;
    ax.nr = (int)t->nr;
    ax.g0 = t->lnr_first;
    ax.glast = t->lnr_last;
    ax.inv_dx = (double)(t->nr - 1) / (ax.glast - ax.g0);
    return ax;
}

static inline int bf_check_table(const bf_table *t) {
    if (!t || !t->gz || !t->gM || !t->gr || !t->values) return 0;
    if (t->nz < 2 || t->nM < 2 || t->nr < 2 || t->nr > (1 << 30)) return 0;
    if (!(t->lnr_last > t->lnr_first)) return 0;
    return 1;
}

__device__ __forceinline__ bool bf_finite(double v) { return isfinite(v); }
