// bf_shell_bin.cu -- halo binning into sky cells (north_star kernel work (1)).
//
// For every halo the exact query_disc ring walk (HealpixRunner.py:306 / :432 semantics, bf_healpix.cuh) is
// run once per ring by one lane; the columns its pixel segments touch are OR-ed over the 8 rings of a band
// into a 64-bit mask, and one (cell, halo) key is emitted per set bit.  Keys are unique; they are written in ONE
// pass at blocks taken from a cursor, then radix-sorted by (cell, halo) with CUB (which makes the layout
// deterministic whatever order the blocks were taken in), and cell boundaries are marked.  After this, cell c's halos are keys[cell_begin[c] .. cell_end[c]) in catalogue
// order -- the order in which the reference's serial loop would have touched those pixels.
#include <cub/cub.cuh>

#include "bf_shell_cells.cuh"

size_t bf_bin_cub_bytes(i64 nh, i64 max_entries) {
    size_t a = 0, b = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, a, (i64 *)nullptr, (i64 *)nullptr, (int)(nh + 1));
    cub::DoubleBuffer<u64> db(nullptr, nullptr);
    cub::DeviceRadixSort::SortKeys(nullptr, b, db, (int)(max_entries > 0 ? max_entries : 1), 0, 64);
    return (a > b ? a : b) + 256;
}

__device__ __forceinline__ u64 seg_mask(int s, int e, int wc) {  // columns touched by [s, e)
    if (e <= s) return 0ull;
    int c0 = s / wc, c1 = (e - 1) / wc;
    u64 hi = (c1 >= 63) ? ~0ull : ((1ull << (c1 + 1)) - 1ull);
    return hi & ~((1ull << c0) - 1ull);
}

// One pass: every halo's keys go to a block of the key array taken from a global cursor (one atomic per 32 rings).
// Where a block lands depends on scheduling, but the keys are unique and are radix-sorted next, so the result does
// not.  Blocks that would pass `max_entries` are counted but not written: the caller learns the exact size from the
// cursor and comes back with a larger workspace.  "Tiny" halos (baryonify mode, < 4 pixel centres in the disc,
// HealpixRunner.py:309-310) emit no keys and are appended to tiny_list instead.
__global__ void __launch_bounds__(128) bin_kernel(bf_hpx h, bf_cells cg, int mode, bf_shell_halos S, i64 j_first,
                                                  i64 j_last, i64 band_lo, i64 band_hi, double resol, i64 max_entries,
                                                  unsigned long long *__restrict__ cursor, u64 *__restrict__ keys,
                                                  int *__restrict__ tiny_list, int *__restrict__ tiny_count) {
    i64 j = j_first + (((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    int lane = threadIdx.x & 31;
    if (j >= j_last) return;
    double v[3] = {S.vec[3 * j], S.vec[3 * j + 1], S.vec[3 * j + 2]};
    double radius = S.radius[j];
    bf_disc d = bfh_disc_setup(h, v, radius);
    i64 ring_lo = (d.north_end > 0) ? 1 : d.irmin;
    i64 ring_hi = (d.south_begin < h.npix) ? 4 * h.nside - 1 : d.irmax;
    if (ring_lo < 1) ring_lo = 1;
    if (ring_hi > 4 * h.nside - 1) ring_hi = 4 * h.nside - 1;

    if (mode == 1 && radius < 3.0 * resol) {  // HealpixRunner.py:309-310 candidates
        i64 cnt = 0;
        for (i64 iz = ring_lo + lane; iz <= ring_hi; iz += 32) {
            i64 nr; int s0, e0, s1, e1;
            cnt += bf_disc_ring_rel(h, d, iz, nr, s0, e0, s1, e1);
        }
        for (int s = 16; s > 0; s >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, s);
        if (cnt < 4) {
            if (lane == 0) tiny_list[atomicAdd(tiny_count, 1)] = (int)j;
            return;
        }
    }

    i64 b_first = (ring_lo - 1) / BF_BAND_RINGS, b_last = (ring_hi - 1) / BF_BAND_RINGS;
    if (b_first < band_lo) b_first = band_lo;
    if (b_last > band_hi - 1) b_last = band_hi - 1;
    for (i64 b0 = b_first; b0 <= b_last; b0 += 4) {  // 4 bands = 32 rings per pass, one ring per lane
        i64 band = b0 + (lane >> 3);
        i64 iz = band * BF_BAND_RINGS + 1 + (lane & 7);
        u64 m = 0;
        if (band <= b_last && iz >= ring_lo && iz <= ring_hi) {
            i64 nr; int s0, e0, s1, e1;
            if (bf_disc_ring_rel(h, d, iz, nr, s0, e0, s1, e1) > 0) m = seg_mask(s0, e0, cg.wc) | seg_mask(s1, e1, cg.wc);
        }
        m |= __shfl_xor_sync(0xffffffffu, m, 1);
        m |= __shfl_xor_sync(0xffffffffu, m, 2);
        m |= __shfl_xor_sync(0xffffffffu, m, 4);
        int c = ((lane & 7) == 0) ? __popcll(m) : 0;
        // exclusive prefix of the 4 band counts (lanes 0, 8, 16, 24)
        int c0 = __shfl_sync(0xffffffffu, c, 0), c1 = __shfl_sync(0xffffffffu, c, 8);
        int c2 = __shfl_sync(0xffffffffu, c, 16), c3 = __shfl_sync(0xffffffffu, c, 24);
        const int tot = c0 + c1 + c2 + c3;
        if (tot == 0) continue;
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(cursor, (unsigned long long)tot);
        base = __shfl_sync(0xffffffffu, base, 0);
        if ((i64)base + tot <= max_entries && (lane & 7) == 0 && c > 0) {
            int g = lane >> 3;
            i64 o = (i64)base + ((g > 0) ? c0 : 0) + ((g > 1) ? c1 : 0) + ((g > 2) ? c2 : 0);
            while (m) {
                int col = __ffsll((long long)m) - 1;
                m &= m - 1;
                keys[o++] = ((u64)(band * BF_MAX_COLS + col) << 32) | (u64)j;
            }
        }
    }
}

__global__ void cell_bounds_kernel(i64 n, const u64 *__restrict__ keys, int *__restrict__ cell_begin,
                                   int *__restrict__ cell_end) {
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    u64 c = keys[i] >> 32;
    if (i == 0 || (keys[i - 1] >> 32) != c) cell_begin[c] = (int)i;
    if (i == n - 1 || (keys[i + 1] >> 32) != c) cell_end[c] = (int)(i + 1);
}

extern "C" int64_t bf_shell_num_bands(int64_t nside) { return nside > 0 ? bf_cells_make(nside).nbands : 0; }

extern "C" int bf_shell_band_pixel_range(int64_t nside, int64_t band_lo, int64_t band_hi, int64_t *first, int64_t *last) {
    BF_REQUIRE(nside > 0 && first && last, "nside / outputs");
    bf_cells c = bf_cells_make(nside);
    BF_REQUIRE(band_lo >= 0 && band_lo <= band_hi && band_hi <= c.nbands, "band range");
    auto ring_start = [&](i64 ring) -> i64 {  // first pixel of ring (ring may be 4*nside = one past the end)
        i64 npix = 12 * nside * nside, ncap = 2 * nside * (nside - 1);
        if (ring >= 4 * nside) return npix;
        if (ring < nside) return 2 * ring * (ring - 1);
        if (ring < 3 * nside) return ncap + (ring - nside) * 4 * nside;
        i64 nr = 4 * nside - ring;
        return npix - 2 * nr * (nr + 1);
    };
    *first = ring_start(band_lo * BF_BAND_RINGS + 1);
    *last = ring_start(band_hi * BF_BAND_RINGS + 1);
    return BF_OK;
}

extern "C" size_t bf_shell_bin_ws_bytes(int64_t nside, int64_t nh, int64_t max_entries) {
    if (nside <= 0 || nh < 0 || max_entries < 0) return 0;
    return bf_bin_ws_layout(nullptr, nside, nh, max_entries).total;
}

static int bin_check(int64_t nside, int mode, const bf_shell_halos *S, int64_t band_lo, int64_t band_hi,
                     int64_t max_entries, void *ws, size_t ws_bytes, bf_bin_ws &w) {
    BF_REQUIRE(nside > 0 && (mode == 0 || mode == 1), "nside / mode");
    BF_REQUIRE(S && S->nh >= 0 && S->nh < (1ll << 31), "halo record");
    BF_REQUIRE(S->nh == 0 || (S->vec && S->radius), "null per-halo array");
    BF_REQUIRE(ws, "null workspace");
    bf_cells cg = bf_cells_make(nside);
    BF_REQUIRE(band_lo >= 0 && band_lo <= band_hi && band_hi <= cg.nbands, "band range");
    BF_REQUIRE(max_entries >= 0 && max_entries < (1ll << 31), "max_entries must fit int32");
    w = bf_bin_ws_layout(ws, nside, S->nh, max_entries);
    if (ws_bytes < w.total) {
        bf_set_error("bf_shell_bin: workspace %zu < %zu bytes", ws_bytes, w.total);
        return BF_E_WORKSPACE;
    }
    return BF_OK;
}

// The binning in three steps, so that a caller whose per-halo arrays arrive piece by piece (the host prepares the
// catalogue in runs of consecutive halos) can bin every run as soon as it is on the device: _begin resets the key
// cursor, _add emits the keys of halos j_first <= j < j_last (the arrays of `S` are the whole catalogue's; entries beyond
// j_last need not be valid yet), _end sorts and marks the cell boundaries.  bf_shell_bin_halos = begin, add(0, nh), end.
extern "C" int bf_shell_bin_begin(int64_t nside, int mode, const bf_shell_halos *S, int64_t band_lo, int64_t band_hi,
                                  int64_t max_entries, void *ws, size_t ws_bytes, void *stream) {
    bf_bin_ws w;
    int rc = bin_check(nside, mode, S, band_lo, band_hi, max_entries, ws, ws_bytes, w);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    BF_CUDA_OK(cudaMemsetAsync(w.cell_begin, 0, sizeof(int) * (size_t)w.ncells, s));
    BF_CUDA_OK(cudaMemsetAsync(w.cell_end, 0, sizeof(int) * (size_t)w.ncells, s));
    BF_CUDA_OK(cudaMemsetAsync(w.tiny_count, 0, sizeof(int), s));
    BF_CUDA_OK(cudaMemsetAsync(w.counts, 0, sizeof(unsigned long long), s));  // counts[0] holds the key cursor
    return BF_OK;
}

extern "C" int bf_shell_bin_add(int64_t nside, int mode, const bf_shell_halos *S, int64_t j_first, int64_t j_last,
                                int64_t band_lo, int64_t band_hi, int64_t max_entries, void *ws, size_t ws_bytes,
                                void *stream) {
    bf_bin_ws w;
    int rc = bin_check(nside, mode, S, band_lo, band_hi, max_entries, ws, ws_bytes, w);
    if (rc) return rc;
    BF_REQUIRE(j_first >= 0 && j_first <= j_last && j_last <= S->nh, "halo range");
    if (j_last == j_first) return BF_OK;
    bf_hpx h = bf_hpx_make(nside);
    const double resol = sqrt(4.0 * BFH_PI / (double)h.npix);
    const unsigned blocks = (unsigned)(((j_last - j_first) * 32 + 127) / 128);
    bin_kernel<<<blocks, 128, 0, (cudaStream_t)stream>>>(h, bf_cells_make(nside), mode, *S, j_first, j_last, band_lo,
                                                        band_hi, resol, max_entries, (unsigned long long *)w.counts,
                                                        w.keys, w.tiny_list, w.tiny_count);
    BF_LAUNCH_OK();
    return BF_OK;
}

extern "C" int bf_shell_bin_end(int64_t nside, int mode, const bf_shell_halos *S, int64_t band_lo, int64_t band_hi,
                                int64_t max_entries, void *ws, size_t ws_bytes, int64_t *nentries_host, void *stream) {
    BF_REQUIRE(nentries_host, "null outputs");
    bf_bin_ws w;
    int rc = bin_check(nside, mode, S, band_lo, band_hi, max_entries, ws, ws_bytes, w);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    *nentries_host = 0;
    i64 total = 0;
    BF_CUDA_OK(cudaMemcpyAsync(&total, w.counts, sizeof(i64), cudaMemcpyDeviceToHost, s));
    BF_CUDA_OK(cudaStreamSynchronize(s));
    *nentries_host = total;
    if (total > max_entries) {
        bf_set_error("bf_shell_bin: %lld entries exceed max_entries %lld", (long long)total, (long long)max_entries);
        return BF_E_WORKSPACE;
    }
    if (total == 0) return BF_OK;
    int cell_bits = 1;
    while (((i64)1 << cell_bits) < w.ncells) ++cell_bits;
    cub::DoubleBuffer<u64> db(w.keys, w.keys_alt);
    size_t cb = w.cub_bytes;
    BF_CUDA_OK(cub::DeviceRadixSort::SortKeys(w.cub_tmp, cb, db, (int)total, 0, 32 + cell_bits, s));
    if (db.Current() != w.keys)
        BF_CUDA_OK(cudaMemcpyAsync(w.keys, db.Current(), sizeof(u64) * (size_t)total, cudaMemcpyDeviceToDevice, s));
    cell_bounds_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(total, w.keys, w.cell_begin, w.cell_end);
    BF_LAUNCH_OK();
    return BF_OK;
}

extern "C" int bf_shell_bin_halos(int64_t nside, int mode, const bf_shell_halos *S, int64_t band_lo, int64_t band_hi,
                                  int64_t max_entries, void *ws, size_t ws_bytes, int64_t *nentries_host,
                                  void *stream) {
    BF_REQUIRE(nentries_host && ws, "null outputs");
    int rc = bf_shell_bin_begin(nside, mode, S, band_lo, band_hi, max_entries, ws, ws_bytes, stream);
    if (rc) return rc;
    rc = bf_shell_bin_add(nside, mode, S, 0, S->nh, band_lo, band_hi, max_entries, ws, ws_bytes, stream);
    if (rc) return rc;
    return bf_shell_bin_end(nside, mode, S, band_lo, band_hi, max_entries, ws, ws_bytes, nentries_host, stream);
}
