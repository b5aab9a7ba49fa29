// bf_shell_cells.cuh -- sky-cell geometry and workspace layout shared by the halo binning kernels
// (bf_shell_bin.cu) and the pixel-owner gather kernel (bf_shell_gather.cu).
//
// cell = (band, col): band = 8 consecutive rings (band b holds rings 8b+1 .. 8b+8), col = pixels
// [col*Wc, (col+1)*Wc) of each of those rings (by in-ring index; rings shorter than col*Wc simply have
// no pixels in that cell).  Wc = 32*CW is chosen so that the longest ring has at most 64 columns: the
// set of columns a disc touches in a band then fits one 64-bit mask.
#pragma once
#include "bf_healpix.cuh"

#define BF_BAND_RINGS 8
#define BF_GATHER_NR_MAX 8192  // longest radial axis the gather kernel stages (24 B per node + 2 KB of ln table)
#define BF_GATHER_AUX_BYTES (2048 + 24 * BF_GATHER_NR_MAX + 16)
#define BF_MAX_COLS 64

struct bf_cells {
    int cw;       // 32-pixel strips per column
    int wc;       // column width in pixels (32*cw)
    int ncol;     // columns of the longest ring (<= 64)
    i64 nbands;   // ceil((4*nside-1)/8)
};

static inline bf_cells bf_cells_make(i64 nside) {
    bf_cells c;
    i64 nrmax = 4 * nside;
    i64 cw = (nrmax + (i64)BF_MAX_COLS * 32 - 1) / ((i64)BF_MAX_COLS * 32);
    if (cw < 1) cw = 1;
    c.cw = (int)cw;
    c.wc = 32 * c.cw;
    c.ncol = (int)((nrmax + c.wc - 1) / c.wc);
    c.nbands = (4 * nside - 1 + BF_BAND_RINGS - 1) / BF_BAND_RINGS;
    return c;
}

// workspace of the binning stage (all offsets 256-byte aligned)
struct bf_bin_ws {
    i64 *counts;        // [nh]   entries per halo
    i64 *offsets;       // [nh+1] exclusive scan
    u64 *keys;          // [max_entries]  (cell << 32) | halo, sorted
    u64 *keys_alt;      // [max_entries]  radix-sort double buffer
    int *cell_begin;    // [ncells]
    int *cell_end;      // [ncells]
    int *tiny_list;     // [nh]   halos diverted to the < 4-pixel fallback (baryonify)
    int *tiny_count;    // [1]
    void *aux;          // [BF_GATHER_AUX_BYTES] per-call constant tables of the gather kernel (ln table, radial axis)
    void *cub_tmp;
    size_t cub_bytes;
    i64 ncells;
    size_t total;
};

static inline size_t bf_al256(size_t v) { return (v + 255) & ~(size_t)255; }

size_t bf_bin_cub_bytes(i64 nh, i64 max_entries);  // defined in bf_shell_bin.cu

static inline bf_bin_ws bf_bin_ws_layout(void *base, i64 nside, i64 nh, i64 max_entries) {
    bf_cells c = bf_cells_make(nside);
    bf_bin_ws w;
    char *p = (char *)base;
    size_t o = 0;
    w.ncells = c.nbands * BF_MAX_COLS;
    // fixed-size arrays first: their addresses depend on (nside, nh) only, so the gather stage can find
    // them without knowing max_entries
    w.counts = (i64 *)(p + o);      o += bf_al256(sizeof(i64) * (size_t)(nh + 1));
    w.offsets = (i64 *)(p + o);     o += bf_al256(sizeof(i64) * (size_t)(nh + 2));
    w.cell_begin = (int *)(p + o);  o += bf_al256(sizeof(int) * (size_t)w.ncells);
    w.cell_end = (int *)(p + o);    o += bf_al256(sizeof(int) * (size_t)w.ncells);
    w.tiny_list = (int *)(p + o);   o += bf_al256(sizeof(int) * (size_t)(nh + 1));
    w.tiny_count = (int *)(p + o);  o += 256;
    w.aux = (void *)(p + o);        o += bf_al256(BF_GATHER_AUX_BYTES);
    w.keys = (u64 *)(p + o);        o += bf_al256(sizeof(u64) * (size_t)(max_entries + 1));
    w.keys_alt = (u64 *)(p + o);    o += bf_al256(sizeof(u64) * (size_t)(max_entries + 1));
    w.cub_bytes = bf_bin_cub_bytes(nh, max_entries);
    w.cub_tmp = (void *)(p + o);    o += bf_al256(w.cub_bytes);
    w.total = o;
    return w;
}

// ring-relative pixel segment(s) of disc d in ring iz, including whole-cap rings; returns count.
// Segments are [s0, e0) and [s1, e1) in in-ring index (0-based), first one starting lower.
__device__ __forceinline__ i64 bf_disc_ring_rel(const bf_hpx &h, const bf_disc &d, i64 iz, i64 &nr, int &s0, int &e0,
                                                int &s1, int &e1) {
    i64 sp; bool sh;
    bfh_ring_info_small(h, iz, sp, nr, sh);
    s0 = e0 = s1 = e1 = 0;
    if (sp + nr <= d.north_end || sp >= d.south_begin) { e0 = (int)nr; return nr; }  // whole ring (pole in disc)
    if (iz < d.irmin || iz > d.irmax) return 0;
    i64 a0, a1, b0, b1;
    i64 c = bfh_disc_ring(h, d, iz, a0, a1, b0, b1);
    if (a1 > a0) { s0 = (int)(a0 - sp); e0 = (int)(a1 - sp); }
    if (b1 > b0) { s1 = (int)(b0 - sp); e1 = (int)(b1 - sp); }
    return c;
}
