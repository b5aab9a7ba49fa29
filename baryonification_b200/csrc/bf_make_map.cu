// bf_make_map.cu -- ParticleSnapshot.make_map / HaloNDCatalog-style mass maps on the device.
//
// Replaces utils/io.py:657-670: `np.histogramdd(coords, bins=(linspace(0, L, N + 1),) * D, weights=M)`.  numpy's
// histogramdd gives every sample the bin `searchsorted(edges, x, 'right') - 1`, moves samples that sit exactly on the
// last edge into the last bin, drops everything outside [edges[0], edges[-1]], and then adds the weights with
// np.bincount -- a serial loop in SAMPLE order.  Here:
//   1. key    : cell index of every particle against the SAME edges array (the host's np.linspace values are uploaded;
//               a uniform-grid guess is corrected against them, so the bins equal searchsorted's bit for bit);
//               out-of-range particles get the key `ncell`;
//   2. sort   : stable CUB radix sort of (cell, mass) pairs -- particles of one cell stay in input order;
//   3. reduce : the first particle of every run of equal keys adds its run left to right and writes the cell once.
// That is numpy's summation order, so the map is bit-identical to the reference's whatever the masses are, and
// bit-reproducible (no atomics).  24 B of workspace per particle plus the sort's temporary storage.
#include <cub/cub.cuh>

#include "bf_common.cuh"

typedef unsigned int u32;

struct mm_ws {
    u32 *keys, *keys_alt;
    double *vals, *vals_alt;
    void *cub_tmp;
    size_t cub_bytes, total;
};

static size_t mm_al(size_t v) { return (v + 255) & ~(size_t)255; }

static mm_ws mm_layout(void *base, i64 np) {
    mm_ws w;
    char *p = (char *)base;
    size_t o = 0;
    const size_t n = (size_t)(np > 0 ? np : 1);
    w.keys = (u32 *)(p + o);        o += mm_al(sizeof(u32) * n);
    w.keys_alt = (u32 *)(p + o);    o += mm_al(sizeof(u32) * n);
    w.vals = (double *)(p + o);     o += mm_al(sizeof(double) * n);
    w.vals_alt = (double *)(p + o); o += mm_al(sizeof(double) * n);
    size_t b = 0;
    cub::DoubleBuffer<u32> dk(nullptr, nullptr);
    cub::DoubleBuffer<double> dv(nullptr, nullptr);
    cub::DeviceRadixSort::SortPairs(nullptr, b, dk, dv, (int)n, 0, 32);
    w.cub_bytes = b + 256;
    w.cub_tmp = (void *)(p + o);    o += mm_al(w.cub_bytes);
    w.total = o;
    return w;
}

// searchsorted(edges, x, 'right') - 1 with numpy's last-edge rule; -1 = outside [edges[0], edges[n]] (or NaN)
__device__ __forceinline__ int mm_bin(const double *__restrict__ edges, int n, double inv_step, double x) {
    const double e0 = edges[0], e1 = edges[n];
    if (!(x >= e0 && x <= e1)) return -1;
    if (x == e1) return n - 1;
    int k = (int)((x - e0) * inv_step);
    k = max(0, min(k, n - 1));
    while (k > 0 && x < edges[k]) --k;
    while (k < n - 1 && x >= edges[k + 1]) ++k;
    return k;
}

template <int DIM, bool REC>
__global__ void __launch_bounds__(256) mm_key_kernel(i64 np, const double *__restrict__ x, const double *__restrict__ y,
                                                     const double *__restrict__ z, const double *__restrict__ m,
                                                     const double4 *__restrict__ rec, const double *__restrict__ edges,
                                                     int n, double inv_step, u32 ncell, u32 *__restrict__ keys,
                                                     double *__restrict__ vals) {
    const i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= np) return;
    double px, py, pz = 0.0, pm;
    if (REC) {
        const double2 *r = reinterpret_cast<const double2 *>(rec + p);
        const double2 a = __ldcs(r), b = __ldcs(r + 1);
        pm = a.x; px = a.y; py = b.x; pz = b.y;
    } else {
        px = __ldcs(x + p); py = __ldcs(y + p);
        if (DIM == 3) pz = __ldcs(z + p);
        pm = __ldcs(m + p);
    }
    const int ix = mm_bin(edges, n, inv_step, px), iy = mm_bin(edges, n, inv_step, py);
    const int iz = (DIM == 3) ? mm_bin(edges, n, inv_step, pz) : 0;
    u32 c = ncell;
    if (ix >= 0 && iy >= 0 && iz >= 0) c = (DIM == 3) ? ((u32)ix * (u32)n + (u32)iy) * (u32)n + (u32)iz : (u32)ix * (u32)n + (u32)iy;
    keys[p] = c;
    vals[p] = pm;
}

__global__ void __launch_bounds__(256) mm_reduce_kernel(i64 np, const u32 *__restrict__ keys, const double *__restrict__ vals,
                                                        u32 ncell, double *__restrict__ grid) {
    const i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= np) return;
    const u32 c = keys[i];
    if (c >= ncell || (i > 0 && keys[i - 1] == c)) return;  // dropped, or not the head of its run
    double s = 0.0 + vals[i];  // (np.bincount starts from 0.0)
    for (i64 q = i + 1; q < np && keys[q] == c; ++q) s += vals[q];
    grid[c] = s;
}

extern "C" size_t bf_make_map_ws_bytes(int64_t np) {
    if (np < 0 || np >= (1ll << 31)) return 0;
    return mm_layout(nullptr, np).total;
}

extern "C" int bf_make_map(int dim, int64_t np, const double *x, const double *y, const double *z, const double *m,
                           const void *records, const double *edges, int64_t ngrid, double *grid, void *ws,
                           size_t ws_bytes, void *stream) {
    BF_REQUIRE(dim == 2 || dim == 3, "dim must be 2 or 3");
    BF_REQUIRE(np >= 0 && np < (1ll << 31), "np");
    BF_REQUIRE(ngrid >= 1 && edges && grid, "ngrid / edges / grid");
    const double ncell_d = (dim == 3) ? (double)ngrid * ngrid * ngrid : (double)ngrid * ngrid;
    BF_REQUIRE(ncell_d < 4294967295.0, "ngrid^dim must stay below 2^32");
    BF_REQUIRE(np == 0 || records || (x && y && m && (dim == 2 || z)), "particle arrays");
    BF_REQUIRE(!records || (((uintptr_t)records) & 15) == 0, "particle records must be 16-byte aligned");
    const u32 ncell = (u32)ncell_d;
    cudaStream_t s = (cudaStream_t)stream;
    BF_CUDA_OK(cudaMemsetAsync(grid, 0, sizeof(double) * (size_t)ncell, s));  // empty cells: histogramdd's zeros
    if (np == 0) return BF_OK;
    mm_ws w = mm_layout(ws, np);
    if (!ws || ws_bytes < w.total) {
        bf_set_error("bf_make_map: workspace %zu < %zu bytes", ws_bytes, w.total);
        return BF_E_WORKSPACE;
    }
    double e[2] = {0.0, 0.0};  // the uniform-grid guess needs the step of the edges: read the two ends back
    BF_CUDA_OK(cudaMemcpyAsync(&e[0], edges, sizeof(double), cudaMemcpyDeviceToHost, s));
    BF_CUDA_OK(cudaMemcpyAsync(&e[1], edges + ngrid, sizeof(double), cudaMemcpyDeviceToHost, s));
    BF_CUDA_OK(cudaStreamSynchronize(s));
    BF_REQUIRE(e[1] > e[0], "edges must increase");
    const double inv_step = (double)ngrid / (e[1] - e[0]);
    const unsigned blocks = (unsigned)((np + 255) / 256);
    const double4 *rec = (const double4 *)records;
    if (dim == 2) {
        if (rec) mm_key_kernel<2, true><<<blocks, 256, 0, s>>>(np, x, y, z, m, rec, edges, (int)ngrid, inv_step, ncell, w.keys, w.vals);
        else mm_key_kernel<2, false><<<blocks, 256, 0, s>>>(np, x, y, z, m, rec, edges, (int)ngrid, inv_step, ncell, w.keys, w.vals);
    } else {
        if (rec) mm_key_kernel<3, true><<<blocks, 256, 0, s>>>(np, x, y, z, m, rec, edges, (int)ngrid, inv_step, ncell, w.keys, w.vals);
        else mm_key_kernel<3, false><<<blocks, 256, 0, s>>>(np, x, y, z, m, rec, edges, (int)ngrid, inv_step, ncell, w.keys, w.vals);
    }
    BF_LAUNCH_OK();
    int bits = 1;
    while (bits < 32 && (1ull << bits) <= (unsigned long long)ncell) ++bits;  // keys run up to ncell inclusive
    cub::DoubleBuffer<u32> dk(w.keys, w.keys_alt);
    cub::DoubleBuffer<double> dv(w.vals, w.vals_alt);
    size_t cb = w.cub_bytes;
    BF_CUDA_OK(cub::DeviceRadixSort::SortPairs(w.cub_tmp, cb, dk, dv, (int)np, 0, bits, s));  // stable: input order inside a cell
    bf_count_launch();
    mm_reduce_kernel<<<blocks, 256, 0, s>>>(np, dk.Current(), dv.Current(), ncell, grid);
    BF_LAUNCH_OK();
    return BF_OK;
}
