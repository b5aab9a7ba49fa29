// bf_regrid_sorted.cu -- deterministic re-binning: sort-by-destination + in-order segmented reduction.
//
// The atomic re-binning kernels (bf_shell.cu: shell_regrid_kernel, bf_grid.cu: grid_regrid_kernel) add the
// 4 (HEALPix) / 2^D (grid) deposits of every displaced pixel with fp64 RED.ADD, so the summation order -- and
// the last bits of the map -- change from run to run.  This variant reproduces the SERIAL order of the
// reference's numba loops (regrid_pixels_hpix, Runners/HealpixRunner.py:13-67: for i in parents, for j in 0..3:
// hmap[child[i,j]] += w[i,j]*val[i];  regrid_pixels_2D/3D, Runners/Map2DRunner.py:14-163):
//   1. emit    : entry e = i*K + j  ->  key[e] = destination pixel, val[e] = w*val  (masked pixels: key = sentinel)
//   2. sort    : a counting sort by destination.  rank[e] = atomicAdd(count[key[e]], 1) is the entry's arrival
//                order inside its destination (any order); an exclusive scan of the counts gives every destination
//                a contiguous bucket; entry e goes to bucket slot start[key] + rank, together with its id e.
//                Deposits land next to their sources, so the buckets of neighbouring entries share cache lines
//                and this costs ~3 streaming passes instead of the 4 x (read + write) of a 28-bit radix sort.
//   3. reduce  : one thread per destination adds its bucket in ascending entry id -- the order in which the
//                reference's loop would have added them -- and does ONE non-atomic += on the map.
// The integer atomics only decide WHERE inside a bucket an entry is parked; the sum never depends on them, so the
// result is bit-reproducible.  HBM-bound: ~60 B per entry over all passes.
#include <cub/cub.cuh>

#include "bf_healpix.cuh"

typedef unsigned int u32;
typedef unsigned char u8;

struct sorted_ws {
    u32 *key;      // [nent] destination of every entry (emit output)
    double *val;   // [nent]
    u8 *rank;      // [nent] arrival order inside the destination bucket (saturating: see overflow)
    u32 *bid;      // [nent] bucketed entry ids
    double *bval;  // [nent] bucketed values
    int *count;    // [nout + 1]
    int *start;    // [nout + 2]
    int *overflow; // [1] set when a destination receives > 255 entries (buckets are then filled with a cursor)
    void *cub_tmp;
    size_t cub_bytes, total;
};

static size_t al256(size_t v) { return (v + 255) & ~(size_t)255; }

static sorted_ws sorted_layout(void *base, i64 nent, i64 nout) {
    sorted_ws w;
    char *p = (char *)base;
    size_t o = 0;
    size_t n = (size_t)(nent > 0 ? nent : 1);
    w.key = (u32 *)(p + o);     o += al256(sizeof(u32) * n);
    w.val = (double *)(p + o);  o += al256(sizeof(double) * n);
    w.rank = (u8 *)(p + o);     o += al256(n);
    w.bid = (u32 *)(p + o);     o += al256(sizeof(u32) * n);
    w.bval = (double *)(p + o); o += al256(sizeof(double) * n);
    w.count = (int *)(p + o);   o += al256(sizeof(int) * (size_t)(nout + 2));
    w.start = (int *)(p + o);   o += al256(sizeof(int) * (size_t)(nout + 2));
    w.overflow = (int *)(p + o); o += 256;
    size_t cb = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, cb, (int *)nullptr, (int *)nullptr, (int)(nout + 1));
    w.cub_bytes = cb + 256;
    w.cub_tmp = (void *)(p + o); o += al256(w.cub_bytes);
    w.total = o;
    return w;
}

// arrival order of one entry inside its destination (saturating u8; the flag sends the bucket pass to its cursor mode)
__device__ __forceinline__ u8 take_rank(u32 k, u32 nout, int *__restrict__ count, int *__restrict__ overflow) {
    if (k >= nout) return 0;  // sentinel: masked source
    int r = atomicAdd(count + k, 1);
    if (r > 255) { r = 255; *overflow = 1; }
    return (u8)r;
}

// only when a bucket overflowed: the counters become the cursors of the bucket pass (decided on the device, so the
// call never waits for the host)
__global__ void __launch_bounds__(256) reset_if_overflow_kernel(const int *__restrict__ overflow, int *__restrict__ count,
                                                                i64 n) {
    if (*overflow == 0) return;
    for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x) count[i] = 0;
}

__global__ void __launch_bounds__(256) rank_kernel(i64 n, const u32 *__restrict__ key, u32 nout, int *__restrict__ count,
                                                   u8 *__restrict__ rank, int *__restrict__ overflow) {
    i64 e = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    u32 k = key[e];
    if (k >= nout) return;  // sentinel: masked source
    int r = atomicAdd(count + k, 1);
    if (r > 255) { r = 255; *overflow = 1; }
    rank[e] = (u8)r;
}

__global__ void __launch_bounds__(256) bucket_kernel(i64 n, const u32 *__restrict__ key, const double *__restrict__ val,
                                                     const u8 *__restrict__ rank, u32 nout,
                                                     const int *__restrict__ start, int *__restrict__ cursor,
                                                     const int *__restrict__ overflow, u32 *__restrict__ bid,
                                                     double *__restrict__ bval) {
    i64 e = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    u32 k = key[e];
    if (k >= nout) return;
    // normally the recorded rank is the slot; if a bucket overflowed u8, every entry takes a fresh one from the cursors
    int slot = start[k] + (*overflow ? atomicAdd(cursor + k, 1) : (int)rank[e]);
    bid[slot] = (u32)e;
    bval[slot] = val[e];
}

// one thread per destination: add the bucket in ascending entry id (selection by "smallest id above the last one";
// buckets hold ~K entries, so this is a handful of cached loads)
__global__ void __launch_bounds__(256) bucket_reduce_kernel(u32 nout, const int *__restrict__ start,
                                                            const u32 *__restrict__ bid, const double *__restrict__ bval,
                                                            double *__restrict__ out) {
    u32 d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= nout) return;
    const int b = start[d], e = start[d + 1];
    if (e <= b) return;
    double s = 0.0;
    i64 last = -1;
    for (int t = b; t < e; ++t) {
        i64 best = 0x7fffffffffffffffLL;
        double bv = 0.0;
        for (int q = b; q < e; ++q) {
            i64 id = (i64)bid[q];
            if (id > last && id < best) { best = id; bv = bval[q]; }
        }
        s = (t == b) ? bv : s + bv;
        last = best;
    }
    out[d] += s;
}

// zero the counters and the flag; the emit kernels take the ranks themselves (`ranked`), saving a pass over the keys
static int sort_prepare(sorted_ws &w, u32 nout, cudaStream_t s) {
    BF_CUDA_OK(cudaMemsetAsync(w.count, 0, sizeof(int) * ((size_t)nout + 2), s));
    BF_CUDA_OK(cudaMemsetAsync(w.overflow, 0, sizeof(int), s));
    return BF_OK;
}

static int sort_and_reduce(sorted_ws &w, i64 nent, u32 nout, double *out, bool ranked, cudaStream_t s) {
    unsigned eb = (unsigned)((nent + 255) / 256);
    if (!ranked) {
        int rc = sort_prepare(w, nout, s);
        if (rc) return rc;
        rank_kernel<<<eb, 256, 0, s>>>(nent, w.key, nout, w.count, w.rank, w.overflow);
        BF_LAUNCH_OK();
    }
    size_t cb = w.cub_bytes;
    BF_CUDA_OK(cub::DeviceScan::ExclusiveSum(w.cub_tmp, cb, w.count, w.start, (int)(nout + 1), s));
    reset_if_overflow_kernel<<<bf_sm_count() * 8, 256, 0, s>>>(w.overflow, w.count, (i64)nout + 2);
    BF_LAUNCH_OK();
    bucket_kernel<<<eb, 256, 0, s>>>(nent, w.key, w.val, w.rank, nout, w.start, w.count, w.overflow, w.bid, w.bval);
    BF_LAUNCH_OK();
    bucket_reduce_kernel<<<(nout + 255) / 256, 256, 0, s>>>(nout, w.start, w.bid, w.bval, out);
    BF_LAUNCH_OK();
    return BF_OK;
}

// ---------------------------------------------------------------------------------------------------
// HEALPix shell (HealpixRunner.py:333-341)
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) shell_emit_kernel(bf_hpx h, i64 pix_first, i64 pix_last,
                                                         const double *__restrict__ orig,
                                                         const double *__restrict__ off, u32 *__restrict__ key,
                                                         double *__restrict__ val, int *__restrict__ count,
                                                         u8 *__restrict__ rank, int *__restrict__ overflow) {
    i64 q = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    i64 p = pix_first + q;
    if (p >= pix_last) return;
    double m = orig[q];
    uint4 kk = make_uint4((u32)h.npix, (u32)h.npix, (u32)h.npix, (u32)h.npix);
    double vv[4] = {0.0, 0.0, 0.0, 0.0};
    if (m > 0) {  // :335
        const double o3[3] = {off[3 * q], off[3 * q + 1], off[3 * q + 2]};
        i64 cp[4]; double w[4];
        bfh_displaced_deposits(h, p, o3, cp, w);
        kk = make_uint4((u32)cp[0], (u32)cp[1], (u32)cp[2], (u32)cp[3]);
#pragma unroll
        for (int k = 0; k < 4; ++k) vv[k] = w[k] * m;
    }
    const u32 nout = (u32)h.npix;
    reinterpret_cast<uchar4 *>(rank)[q] = make_uchar4(take_rank(kk.x, nout, count, overflow), take_rank(kk.y, nout, count, overflow),
                                                      take_rank(kk.z, nout, count, overflow), take_rank(kk.w, nout, count, overflow));
    reinterpret_cast<uint4 *>(key)[q] = kk;
    reinterpret_cast<double2 *>(val)[2 * q] = make_double2(vv[0], vv[1]);
    reinterpret_cast<double2 *>(val)[2 * q + 1] = make_double2(vv[2], vv[3]);
}

extern "C" size_t bf_shell_regrid_sorted_ws_bytes(int64_t nside, int64_t npix_local) {
    if (nside <= 0 || npix_local < 0) return 0;
    return sorted_layout(nullptr, 4 * npix_local, 12 * nside * nside).total;
}

extern "C" int bf_shell_regrid_sorted(int64_t nside, int64_t pix_first, int64_t pix_last, const double *orig_map,
                                      const double *pix_offsets, double *new_map, void *ws, size_t ws_bytes,
                                      void *stream) {
    BF_REQUIRE(nside > 0 && orig_map && pix_offsets && new_map, "null / nside");
    bf_hpx h = bf_hpx_make(nside);
    BF_REQUIRE(h.npix < 0xffffffffll, "nside too large for 32-bit destination keys");
    BF_REQUIRE(pix_first >= 0 && pix_first <= pix_last && pix_last <= h.npix, "pixel range");
    i64 n = pix_last - pix_first;
    if (n == 0) return BF_OK;
    BF_REQUIRE(4 * n < (1ll << 31), "pixel range too large for one call (4 entries per pixel must fit int32): shard it");
    sorted_ws w = sorted_layout(ws, 4 * n, h.npix);
    if (!ws || ws_bytes < w.total) {
        bf_set_error("bf_shell_regrid_sorted: workspace %zu < %zu bytes", ws_bytes, w.total);
        return BF_E_WORKSPACE;
    }
    cudaStream_t s = (cudaStream_t)stream;
    int rc = sort_prepare(w, (u32)h.npix, s);
    if (rc) return rc;
    shell_emit_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(h, pix_first, pix_last, orig_map, pix_offsets, w.key,
                                                                  w.val, w.count, w.rank, w.overflow);
    BF_LAUNCH_OK();
    return sort_and_reduce(w, 4 * n, (u32)h.npix, new_map, true, s);
}

// ---------------------------------------------------------------------------------------------------
// periodic grids (Map2DRunner.py:583-607): same CIC decomposition as grid_regrid_kernel (bf_grid.cu)
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ double pymod_s(double a, double n) {
    double r = fmod(a, n);
    if (r < 0) r += n;
    return r;
}

template <int DIM>
__global__ void __launch_bounds__(256) grid_emit_kernel(i64 N, const double *__restrict__ off,
                                                        const double *__restrict__ orig, u32 *__restrict__ key,
                                                        double *__restrict__ val) {
    const i64 ncell = (DIM == 2) ? N * N : N * N * N;
    const int K = 1 << DIM;
    i64 cell = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= ncell) return;
    i64 a, b, c = 0;
    if (DIM == 2) { a = cell / N; b = cell - a * N; }
    else { a = cell / (N * N); i64 rem = cell - a * N * N; b = rem / N; c = rem - b * N; }
    double o[3] = {0, 0, 0};
#pragma unroll
    for (int m = 0; m < DIM; ++m) {
        double v = off[DIM * cell + m];
        o[m] = bf_finite(v) ? v : 0.0;
    }
    double sft[3] = {pymod_s(o[1] + (double)a, (double)N), pymod_s(o[0] + (double)b, (double)N),
                     (DIM == 3) ? pymod_s(o[2] + (double)c, (double)N) : 0.0};
    i64 i0[3]; double w0[3], w1[3];
#pragma unroll
    for (int m = 0; m < DIM; ++m) {
        i0[m] = (i64)sft[m];
        w0[m] = (double)(i0[m] + 1) - sft[m];
        w1[m] = (sft[m] + 1.0) - (double)(i0[m] + 1);
    }
    double m0 = orig[cell];
    // the reference scans its neighbourhood with the first axis outermost: (ca, cb[, cc]) ascending in the
    // UNWRAPPED index is the order in which it adds to the map; a destination cell receives at most one deposit
    // per source cell (N >= 5), so only the source order matters for the sum and any fixed slot order will do
#pragma unroll
    for (int e = 0; e < K; ++e) {
        int ca = (DIM == 2) ? (e >> 1) : (e >> 2), cb = (DIM == 2) ? (e & 1) : ((e >> 1) & 1), cc = e & 1;
        double wa = ca ? w1[0] : w0[0], wb = cb ? w1[1] : w0[1], wc = (DIM == 3) ? (cc ? w1[2] : w0[2]) : 1.0;
        bool okw = (wa > 0) && (wb > 0) && (DIM == 2 || wc > 0);
        i64 ia = (i0[0] + ca) % N, ib = (i0[1] + cb) % N, ic = (DIM == 3) ? (i0[2] + cc) % N : 0;
        i64 dst = (DIM == 2) ? ia * N + ib : (ia * N + ib) * N + ic;
        double wv = (DIM == 2) ? (wb * wa) * m0 : ((wb * wa) * wc) * m0;
        key[cell * K + e] = okw ? (u32)dst : (u32)ncell;
        val[cell * K + e] = okw ? wv : 0.0;
    }
}

extern "C" size_t bf_grid_regrid_sorted_ws_bytes(int dim, int64_t npix) {
    if ((dim != 2 && dim != 3) || npix < 1) return 0;
    i64 ncell = (dim == 2) ? npix * npix : npix * npix * npix;
    return sorted_layout(nullptr, ncell << dim, ncell).total;
}

extern "C" int bf_grid_regrid_sorted(int dim, int64_t npix, const double *offsets, const double *orig, double *new_map,
                                     void *ws, size_t ws_bytes, void *stream) {
    BF_REQUIRE(dim == 2 || dim == 3, "dim must be 2 or 3");
    BF_REQUIRE(npix >= 5 && offsets && orig && new_map, "npix >= 5 and non-null arrays");
    i64 ncell = (dim == 2) ? npix * npix : npix * npix * npix;
    BF_REQUIRE(ncell < 0xffffffffll, "grid too large for 32-bit destination keys");
    i64 nent = ncell << dim;
    BF_REQUIRE(nent < (1ll << 31), "grid too large for the sorted variant (2^D entries per cell must fit int32)");
    sorted_ws w = sorted_layout(ws, nent, ncell);
    if (!ws || ws_bytes < w.total) {
        bf_set_error("bf_grid_regrid_sorted: workspace %zu < %zu bytes", ws_bytes, w.total);
        return BF_E_WORKSPACE;
    }
    cudaStream_t s = (cudaStream_t)stream;
    unsigned blocks = (unsigned)((ncell + 255) / 256);
    if (dim == 2) grid_emit_kernel<2><<<blocks, 256, 0, s>>>(npix, offsets, orig, w.key, w.val);
    else grid_emit_kernel<3><<<blocks, 256, 0, s>>>(npix, offsets, orig, w.key, w.val);
    BF_LAUNCH_OK();
    return sort_and_reduce(w, nent, (u32)ncell, new_map, false, s);
}
