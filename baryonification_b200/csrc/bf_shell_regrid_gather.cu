// bf_shell_regrid_gather.cu -- re-binning on the sphere as a DESTINATION GATHER (HealpixRunner.py:333-341 +
// regrid_pixels_hpix :13-67): deterministic, no sort, no atomics on the map, 48 B of workspace per pixel.
//
// The reference's serial loop adds, for source pixels i ascending and j = 0..3, hmap[child[i, j]] += w[i, j] * val[i].
// The counting-sort variant (bf_regrid_sorted.cu) reproduces that order by bucketing the 4 * Npix deposits by
// destination.  Displaced mass lands next to its source, though, so a destination can simply LOOK at the sources
// that can reach it:
//   1. emit    : every source pixel writes its 4 (destination, w * m) deposits (the same bits as the sorted variant:
//                bfh_displaced_deposits) and the kernel measures the WINDOW of the field: the range of
//                ring(destination) - ring(source), and the largest in-ring distance between a source and the
//                projection of its destination onto the source's ring;
//   2. gather  : one thread per destination walks the source rings of the window in ascending order and, inside each
//                ring, the in-ring window in ascending pixel index; a source whose key names the destination adds its
//                deposit.  That is ascending (i, j): the reference's order, and the sorted variant's bits.
// Rings within BF_SHELL_GATHER_POLAR rings of a pole are tiny and their interpolation stencils reach across the pole
// ((pix + 2) & 3), so their destinations scan whole source rings and are left out of the in-ring measurement.
// A field whose window would cost more than BF_SHELL_GATHER_MAX_WINDOW sources per destination returns BF_E_WINDOW:
// the caller falls back to bf_shell_regrid_sorted.
#include "bf_healpix.cuh"

typedef unsigned int u32;

#define RG_POLAR BF_SHELL_GATHER_POLAR

struct rg_window {  // device-resident, filled by the emit kernel
    int dr_min, dr_max;  // ring(destination) - ring(source)
    int dj_max;          // |in-ring index of the source - projection of the destination onto the source's ring|
    int pad;
};

// 0-based in-ring index on a ring of `nr_s` pixels closest in longitude to 0-based index j_d of a ring of `nr_d`
// (equal rings -- the whole equatorial belt -- need no division; 32-bit arithmetic whenever (2 j + 1) nr fits)
__device__ __forceinline__ int rg_project(const bf_hpx &h, int j_d, int nr_d, int nr_s) {
    if (nr_s == nr_d) return j_d;
    if (h.nside <= 8192) return (int)(((unsigned)(2 * j_d + 1) * (unsigned)nr_s) / (unsigned)(2 * nr_d));
    return (int)(((u64)(2 * j_d + 1) * (u64)nr_s) / (u64)(2 * (i64)nr_d));
}

__device__ __forceinline__ bool rg_polar(const bf_hpx &h, int ring) {
    return ring <= RG_POLAR || ring >= 4 * (int)h.nside - RG_POLAR;
}

// first pixel and length of a ring, 32-bit (npix < 2^32 is required by the entry point)
__device__ __forceinline__ void rg_ring(const bf_hpx &h, int ring, u32 &start, int &len) {
    const int ns = (int)h.nside;
    if (ring < ns) { len = 4 * ring; start = 2u * (u32)ring * (u32)(ring - 1); }
    else if (ring < 3 * ns) { len = 4 * ns; start = (u32)h.ncap + (u32)(ring - ns) * (u32)len; }
    else { const int nr = 4 * ns - ring; len = 4 * nr; start = (u32)h.npix - 2u * (u32)nr * (u32)(nr + 1); }
}

__global__ void __launch_bounds__(256) rg_init_kernel(rg_window *w) {
    w->dr_min = 0x7fffffff; w->dr_max = -0x7fffffff; w->dj_max = 0; w->pad = 0;
}

// Sources p_first <= p < p_last (orig / off are that slice); pixels of the margin e_first <= p < e_last around it only
// get the "no deposit" key, so that a ring-band shard can gather its partial sums without reading stale keys.
// co-latitudes of rings 1 .. 4 nside - 1 exactly as the interpolation routine computes them (it is the routine that
// fills the table), so that the emit pass reads two doubles per pixel instead of evaluating two acos / atan2
__global__ void __launch_bounds__(256) rg_theta_kernel(bf_hpx h, double *__restrict__ theta_tab) {
    const i64 ring = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (ring > 4 * h.nside) return;
    double theta = 0.0;
    if (ring >= 1 && ring < 4 * h.nside) {
        i64 sp, nr; bool sh;
        bfh_ring_info2(h, ring, sp, nr, theta, sh);
    }
    theta_tab[ring] = theta;
}

__global__ void __launch_bounds__(256) rg_emit_kernel(bf_hpx h, i64 p_first, i64 p_last, i64 e_first, i64 e_last,
                                                      const double *__restrict__ orig_s,
                                                      const double *__restrict__ off_s, uint4 *__restrict__ key,
                                                      double *__restrict__ val, rg_window *__restrict__ win,
                                                      const double *__restrict__ theta_tab) {
    const i64 p = e_first + (i64)blockIdx.x * blockDim.x + threadIdx.x;
    int dr_min = 0x7fffffff, dr_max = -0x7fffffff, dj_max = 0;
    if (p < e_last && (p < p_first || p >= p_last)) {
        key[p] = make_uint4((u32)h.npix, (u32)h.npix, (u32)h.npix, (u32)h.npix);
    } else if (p < e_last) {
        const double *orig = orig_s - p_first, *off = off_s - 3 * p_first;
        const double m = orig[p];
        uint4 kk = make_uint4((u32)h.npix, (u32)h.npix, (u32)h.npix, (u32)h.npix);
        double vv[4] = {0.0, 0.0, 0.0, 0.0};
        if (m > 0) {  // HealpixRunner.py:335
            const double o3[3] = {off[3 * p], off[3 * p + 1], off[3 * p + 2]};
            i64 cp[4], ir1; double w[4];
            bfh_displaced_deposits_r(h, p, o3, cp, w, ir1, theta_tab);
            kk = make_uint4((u32)cp[0], (u32)cp[1], (u32)cp[2], (u32)cp[3]);
#pragma unroll
            for (int k = 0; k < 4; ++k) vv[k] = w[k] * m;
            // window of this source: its destinations lie on the rings around the displaced point
            i64 r64, j64;
            bfh_pix2ring(h, p, r64, j64);
            const int r_s = (int)r64, j_s = (int)j64 - 1;
            u32 sp_s; int nr_s;
            rg_ring(h, r_s, sp_s, nr_s);
            const int last = 4 * (int)h.nside - 1;
            const int ring2[2] = {max((int)ir1, 1), min((int)ir1 + 1, last)};
#pragma unroll
            for (int g = 0; g < 2; ++g) {
                const int r_d = ring2[g];
                dr_min = min(dr_min, r_d - r_s);
                dr_max = max(dr_max, r_d - r_s);
                if (rg_polar(h, r_d)) continue;
                u32 sp_d; int nr_d;
                rg_ring(h, r_d, sp_d, nr_d);
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int j_d = (int)((u32)cp[2 * g + e] - sp_d);
                    int dj = j_s - rg_project(h, j_d, nr_d, nr_s);  // wrapped into [0, nr_s / 2]
                    if (dj < 0) dj = -dj;
                    if (2 * dj > nr_s) dj = nr_s - dj;
                    dj_max = max(dj_max, dj);
                }
            }
        }
        key[p] = kk;
        reinterpret_cast<double2 *>(val)[2 * p] = make_double2(vv[0], vv[1]);
        reinterpret_cast<double2 *>(val)[2 * p + 1] = make_double2(vv[2], vv[3]);
    }
    for (int s = 16; s > 0; s >>= 1) {
        dr_min = min(dr_min, __shfl_xor_sync(0xffffffffu, dr_min, s));
        dr_max = max(dr_max, __shfl_xor_sync(0xffffffffu, dr_max, s));
        dj_max = max(dj_max, __shfl_xor_sync(0xffffffffu, dj_max, s));
    }
    if ((threadIdx.x & 31) == 0 && dr_max >= dr_min) {
        // look before the atomic: after the first few warps almost nobody widens the window, and millions of atomics on
        // one address would serialise in L2
        const volatile rg_window *cur = win;
        if (dr_min < cur->dr_min) atomicMin(&win->dr_min, dr_min);
        if (dr_max > cur->dr_max) atomicMax(&win->dr_max, dr_max);
        if (dj_max > cur->dj_max) atomicMax(&win->dj_max, dj_max);
    }
}

// the deposits of sources q0 .. q1 (one ring segment, ascending) that name destination d
// (fetching the keys of a segment four at a time before looking at them was measured SLOWER: 7.66 vs 6.25 ms)
__device__ __forceinline__ void rg_scan(const uint4 *__restrict__ key, const double *__restrict__ val, u32 q0, u32 q1,
                                        u32 d, double &s, bool &any) {
    for (u32 q = q0; q <= q1; ++q) {
        const uint4 k = __ldg(key + q);
        if (k.x == d || k.y == d || k.z == d || k.w == d) {  // (about one source in six)
            const double *v = val + 4 * (size_t)q;
            if (k.x == d) { s = any ? s + v[0] : v[0]; any = true; }
            if (k.y == d) { s = any ? s + v[1] : v[1]; any = true; }
            if (k.z == d) { s = any ? s + v[2] : v[2]; any = true; }
            if (k.w == d) { s = any ? s + v[3] : v[3]; any = true; }
        }
    }
}

__global__ void __launch_bounds__(256) rg_gather_kernel(bf_hpx h, int dr_min, int dr_max, int P, i64 d_first,
                                                        i64 d_last, const uint4 *__restrict__ key,
                                                        const double *__restrict__ val, double *__restrict__ out) {
    const i64 d64 = d_first + (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (d64 >= d_last) return;
    const u32 d = (u32)d64;
    i64 r64, j64;
    bfh_pix2ring(h, d64, r64, j64);
    const int r_d = (int)r64, j_d = (int)j64 - 1;
    u32 sp_d; int nr_d;
    rg_ring(h, r_d, sp_d, nr_d);
    const bool whole = rg_polar(h, r_d);
    const int r0 = max(r_d - dr_max, 1), r1 = min(r_d - dr_min, 4 * (int)h.nside - 1);  // source rings, ascending
    double s = 0.0;
    bool any = false;
    for (int r_s = r0; r_s <= r1; ++r_s) {
        u32 sp; int nr_s;
        rg_ring(h, r_s, sp, nr_s);
        if (whole || 2 * P + 1 >= nr_s) {
            rg_scan(key, val, sp, sp + (u32)(nr_s - 1), d, s, any);
            continue;
        }
        const int p0 = rg_project(h, j_d, nr_d, nr_s);
        const int lo = p0 - P, hi = p0 + P;  // hi - lo + 1 < nr_s: at most one end wraps
        if (lo < 0) {
            rg_scan(key, val, sp, sp + (u32)hi, d, s, any);
            rg_scan(key, val, sp + (u32)(lo + nr_s), sp + (u32)(nr_s - 1), d, s, any);
        } else if (hi >= nr_s) {
            rg_scan(key, val, sp, sp + (u32)(hi - nr_s), d, s, any);
            rg_scan(key, val, sp + (u32)lo, sp + (u32)(nr_s - 1), d, s, any);
        } else {
            rg_scan(key, val, sp + (u32)lo, sp + (u32)hi, d, s, any);
        }
    }
    out[d] = any ? s : 0.0;  // every destination of the range is written: the map needs no zero fill
}

extern "C" size_t bf_shell_regrid_gather_ws_bytes(int64_t nside) {
    if (nside <= 0) return 0;
    return (size_t)(12 * nside * nside) * 48 + 512 + 8 * (size_t)(4 * nside + 2);
}

// workspace layout: [window | keys uint4[npix] | values double[4 npix] | ring co-latitudes double[4 nside + 1]]
static int rg_layout(int64_t nside, void *ws, size_t ws_bytes, const char *who, bf_hpx &h, rg_window *&win, uint4 *&key,
                     double *&val) {
    BF_REQUIRE(nside > 0, "nside");
    h = bf_hpx_make(nside);
    BF_REQUIRE(h.npix < 0xffffffffll, "nside too large for 32-bit destination keys");
    size_t need = bf_shell_regrid_gather_ws_bytes(nside);
    if (!ws || ws_bytes < need) {
        bf_set_error("%s: workspace %zu < %zu bytes", who, ws_bytes, need);
        return BF_E_WORKSPACE;
    }
    BF_REQUIRE((((uintptr_t)ws) & 255) == 0, "workspace must be 256-byte aligned");
    char *base = (char *)ws;
    win = (rg_window *)base;
    key = (uint4 *)(base + 256);
    val = (double *)(base + 256 + (size_t)h.npix * 16);
    return BF_OK;
}

// ring co-latitude table: behind the values, [4 nside + 1] doubles
static double *rg_theta_tab(const bf_hpx &h, void *ws) { return (double *)((char *)ws + 512 + (size_t)h.npix * 48); }

// sources pix_first <= p < pix_last; orig_map / pix_offsets are that slice
static int rg_emit(int64_t nside, int64_t pix_first, int64_t pix_last, const double *orig_map, const double *pix_offsets,
                   void *ws, size_t ws_bytes, int *window_host, int margin_rings, void *stream);

extern "C" int bf_shell_regrid_gather_emit_range(int64_t nside, int64_t pix_first, int64_t pix_last, const double *orig_map,
                                                 const double *pix_offsets, void *ws, size_t ws_bytes, int *window_host,
                                                 void *stream) {
    return rg_emit(nside, pix_first, pix_last, orig_map, pix_offsets, ws, ws_bytes, window_host,
                   BF_SHELL_GATHER_MARGIN_RINGS, stream);
}

// The pipelined form (engine: pair stage, emit, gather and read-back proceed ring-band group by group): the keys of the
// group's own sources only -- the neighbouring groups' keys, already written, must survive.
extern "C" int bf_shell_regrid_gather_emit_part(int64_t nside, int64_t pix_first, int64_t pix_last, const double *orig_map,
                                                const double *pix_offsets, void *ws, size_t ws_bytes, int *window_host,
                                                void *stream) {
    return rg_emit(nside, pix_first, pix_last, orig_map, pix_offsets, ws, ws_bytes, window_host, 0, stream);
}

static int rg_emit(int64_t nside, int64_t pix_first, int64_t pix_last, const double *orig_map, const double *pix_offsets,
                   void *ws, size_t ws_bytes, int *window_host, int margin_rings, void *stream) {
    BF_REQUIRE(orig_map && pix_offsets && window_host, "null argument");
    bf_hpx h; rg_window *win; uint4 *key; double *val;
    int rc = rg_layout(nside, ws, ws_bytes, "bf_shell_regrid_gather_emit", h, win, key, val);
    if (rc) return rc;
    BF_REQUIRE(pix_first >= 0 && pix_first <= pix_last && pix_last <= h.npix, "pixel range");
    const i64 margin = (i64)margin_rings * 4 * nside;
    const i64 e_first = (pix_first - margin > 0) ? pix_first - margin : 0;
    const i64 e_last = (pix_last + margin < h.npix) ? pix_last + margin : h.npix;
    cudaStream_t s = (cudaStream_t)stream;
    rg_init_kernel<<<1, 1, 0, s>>>(win);
    BF_LAUNCH_OK();
    double *theta_tab = rg_theta_tab(h, ws);
    rg_theta_kernel<<<(unsigned)((4 * nside + 1 + 255) / 256), 256, 0, s>>>(h, theta_tab);
    BF_LAUNCH_OK();
    if (e_last > e_first) {
        rg_emit_kernel<<<(unsigned)((e_last - e_first + 255) / 256), 256, 0, s>>>(h, pix_first, pix_last, e_first, e_last,
                                                                                 orig_map, pix_offsets, key, val, win,
                                                                                 theta_tab);
        BF_LAUNCH_OK();
    }
    rg_window w;
    BF_CUDA_OK(cudaMemcpyAsync(&w, win, sizeof(w), cudaMemcpyDeviceToHost, s));
    BF_CUDA_OK(cudaStreamSynchronize(s));
    if (w.dr_max < w.dr_min) {  // nothing to deposit (orig <= 0 everywhere): an empty window
        window_host[0] = 1; window_host[1] = 0; window_host[2] = 0;
        return BF_OK;
    }
    window_host[0] = w.dr_min; window_host[1] = w.dr_max; window_host[2] = w.dj_max;
    const long long cost = (long long)(w.dr_max - w.dr_min + 1) * (2LL * w.dj_max + 1);
    const int reach = (w.dr_max > -w.dr_min ? w.dr_max : -w.dr_min) + 1;  // rings between a source and its destination
    const bool part = margin_rings > 0 && (pix_first > 0 || pix_last < h.npix);
    if (cost > BF_SHELL_GATHER_MAX_WINDOW || (part && 2 * reach > BF_SHELL_GATHER_MARGIN_RINGS)) {
        bf_set_error("bf_shell_regrid_gather: window of %d rings x %d pixels exceeds %d sources per destination",
                     w.dr_max - w.dr_min + 1, 2 * w.dj_max + 1, BF_SHELL_GATHER_MAX_WINDOW);
        return BF_E_WINDOW;
    }
    return BF_OK;
}

extern "C" int bf_shell_regrid_gather_emit(int64_t nside, const double *orig_map, const double *pix_offsets, void *ws,
                                           size_t ws_bytes, int *window_host, void *stream) {
    BF_REQUIRE(nside > 0, "nside");
    return bf_shell_regrid_gather_emit_range(nside, 0, 12 * nside * nside, orig_map, pix_offsets, ws, ws_bytes,
                                             window_host, stream);
}

extern "C" int bf_shell_regrid_gather_range(int64_t nside, const void *ws, size_t ws_bytes, const int *window_host,
                                            int64_t dest_first, int64_t dest_last, double *new_map, void *stream) {
    BF_REQUIRE(window_host && new_map, "null argument");
    bf_hpx h; rg_window *win; uint4 *key; double *val;
    int rc = rg_layout(nside, const_cast<void *>(ws), ws_bytes, "bf_shell_regrid_gather_range", h, win, key, val);
    if (rc) return rc;
    BF_REQUIRE(dest_first >= 0 && dest_first <= dest_last && dest_last <= h.npix, "destination range");
    if (dest_last == dest_first) return BF_OK;
    const i64 n = dest_last - dest_first;
    if (window_host[1] < window_host[0]) {  // no deposits at all: the range is still written
        BF_CUDA_OK(cudaMemsetAsync(new_map + dest_first, 0, sizeof(double) * (size_t)n, (cudaStream_t)stream));
        return BF_OK;
    }
    rg_gather_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        h, window_host[0], window_host[1], window_host[2], dest_first, dest_last, key, val, new_map);
    BF_LAUNCH_OK();
    return BF_OK;
}

extern "C" int bf_shell_regrid_gather(int64_t nside, const double *orig_map, const double *pix_offsets, double *new_map,
                                      void *ws, size_t ws_bytes, int *window_host, void *stream) {
    BF_REQUIRE(nside > 0 && orig_map && pix_offsets && new_map, "null / nside");
    int win[3] = {1, 0, 0};
    int rc = bf_shell_regrid_gather_emit(nside, orig_map, pix_offsets, ws, ws_bytes, win, stream);
    const bool empty = win[1] < win[0];
    if (window_host) { window_host[0] = empty ? 0 : win[0]; window_host[1] = empty ? 0 : win[1]; window_host[2] = win[2]; }
    if (rc) return rc;
    return bf_shell_regrid_gather_range(nside, ws, ws_bytes, win, 0, 12 * nside * nside, new_map, stream);
}
