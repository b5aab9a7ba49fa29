// bf_grid_gather.cu -- periodic 2-D / 3-D grid runners, deterministic CELL-OWNER variant.
//
// Replaces BaryonifyGrid.process's halo loop (Runners/Map2DRunner.py:476-580) and PaintProfilesGrid.process's
// halo loop (:714-810).  The halo-centric kernel (bf_grid.cu) adds every (halo, cell) pair into the map with fp64
// atomics: 3 RED per pair, which is what bounds it (16e9 pairs at 1024^3 -> 4.9e10 atomics, 389 ms), and the sum
// order depends on arrival.  Here the loop is inverted like the shell gather kernel:
//   1. bf_grid_bin_halos: every halo's cut-out box (pick_indices :394-423: W = Nsize cells per axis from
//      argmin|bins - x_j| - W/2, periodic) is intersected with tiles of 8^3 (3-D) / 16x32 (2-D) cells; one
//      (tile << 32 | halo) key per overlapped tile, written at an offset from an exclusive scan (no atomics), radix-sorted
//      so that a tile's halos are contiguous and in catalogue order.
//   2. bf_grid_gather: one CTA per tile, one thread per cell.  The tile's halos are staged 32 at a time in shared
//      memory; a thread tests box membership with integer compares, evaluates the pair with the reference's cut-out
//      coordinates (linspace stretch :499, meshgrid 'xy' pairing :510-522) and sums in registers in catalogue order,
//      i.e. the order of the reference's serial loop.  One read-modify-write per cell at the end, no atomics,
//      bit-reproducible.
// The pair set is exactly the halo-centric one (every cell of the box is a pair; the radius cut only zeroes values).
#include <cub/cub.cuh>

#include "bf_common.cuh"
#include "bf_grid_common.cuh"

namespace {

constexpr int GG_THREADS = 512;
constexpr int GG_CHUNK = 32;  // halos staged per round

template <int DIM> struct tile_shape;
template <> struct tile_shape<2> { static constexpr int A = 16, B = 32, C = 1; };
template <> struct tile_shape<3> { static constexpr int A = 8, B = 8, C = 8; };

struct gh_rec {       // one prepared halo, 80 bytes
    int lo[3];        // first cut-out cell per grid axis (unwrapped, may be negative): cen - W/2
    int W;            // cut-out cells per axis (= Nsize), 0 = halo skipped (Nsize < 2)
    int j, N;         // catalogue index; Nsize (the linspace length, = W for the reference's even Nsize)
    double dd[3];     // bins[cen] - x_j per POSITION component (x, y, z)
    double step;      // linspace step W / (W - 1)
    double cut, shift;
    const double *row;
};
static_assert(sizeof(gh_rec) == 80, "gh_rec layout");

struct gg_hdr {  // device-resident header at the start of the workspace
    const unsigned long long *sorted_keys;
    long long nentries;
};

struct gg_ws {
    gg_hdr *hdr;
    gh_rec *recs;
    long long *cnt, *offs;  // [nh + 1]
    int *tstart, *tend;     // [ntiles]
    double *axis;           // ln r grid [nr], 1 / (g[k+1] - g[k]) [nr], ln table [2 * BF_LOGTAB]: filled by bf_grid_gather
    unsigned long long *keys, *keys_alt;
    void *cub_tmp;
    size_t cub_bytes, total;
};

__host__ __device__ inline int tiles_per_axis(int npix, int T) { return (npix + T - 1) / T; }

constexpr int GG_MAX_NR = 8192;  // longest radial axis staged in shared memory

static gg_ws gg_layout(void *base, int dim, i64 npix, i64 nh, i64 max_entries) {
    gg_ws w;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return (char *)base + o; };
    const int ta = tiles_per_axis((int)npix, dim == 2 ? tile_shape<2>::A : tile_shape<3>::A),
              tb = tiles_per_axis((int)npix, dim == 2 ? tile_shape<2>::B : tile_shape<3>::B),
              tc = dim == 3 ? tiles_per_axis((int)npix, tile_shape<3>::C) : 1;
    const size_t nt = (size_t)ta * tb * tc;
    w.hdr = (gg_hdr *)take(sizeof(gg_hdr));
    w.recs = (gh_rec *)take(sizeof(gh_rec) * (size_t)(nh + 1));
    w.cnt = (long long *)take(sizeof(long long) * (size_t)(nh + 1));
    w.offs = (long long *)take(sizeof(long long) * (size_t)(nh + 1));
    w.tstart = (int *)take(sizeof(int) * nt);
    w.tend = (int *)take(sizeof(int) * nt);
    w.axis = (double *)take(sizeof(double) * (2 * GG_MAX_NR + 2 * BF_LOGTAB));
    w.keys = (unsigned long long *)take(sizeof(unsigned long long) * (size_t)max_entries);
    w.keys_alt = (unsigned long long *)take(sizeof(unsigned long long) * (size_t)max_entries);
    size_t c1 = 0, c2 = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, c1, (long long *)nullptr, (long long *)nullptr, (int)(nh + 1));
    cub::DeviceRadixSort::SortKeys(nullptr, c2, (unsigned long long *)nullptr, (unsigned long long *)nullptr,
                                   (int)std::max<i64>(max_entries, 1), 0, 64);
    w.cub_bytes = std::max(c1, c2);
    w.cub_tmp = take(w.cub_bytes);
    w.total = off;
    return w;
}

// circular run of tiles covered by W cells starting at cell lo (any sign) on an axis of npix cells, tile edge T:
// first tile and number of distinct tiles
__device__ __forceinline__ void tile_run(int lo, int W, int npix, int T, int &t0, int &n) {
    int l = lo % npix;
    if (l < 0) l += npix;
    const int nt = tiles_per_axis(npix, T);
    t0 = l / T;
    if (l + W <= npix) n = (l + W - 1) / T - t0 + 1;
    else n = min(nt, (nt - t0) + ((l + W - npix - 1) / T + 1));
}

template <int DIM>
__global__ void __launch_bounds__(256) gg_prep_kernel(int npix, const double *__restrict__ bins, bf_grid_halos G, int nr,
                                                      gh_rec *__restrict__ recs, long long *__restrict__ cnt,
                                                      int *__restrict__ too_wide) {
    using TS = tile_shape<DIM>;
    i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (j > G.nh) return;
    if (j == G.nh) { cnt[j] = 0; return; }
    gh_rec r;
    i64 N = G.Nsize[j];
    r.W = (N < 2) ? 0 : (int)(2 * (N / 2));
    r.j = (int)j; r.N = (int)N;
    r.lo[2] = 0; r.dd[2] = 0.0;
    if (r.W > npix) { atomicExch(too_wide, 1); r.W = 0; }
    long long n = (r.W > 0) ? 1 : 0;
    const int T[3] = {TS::A, TS::B, TS::C};
#pragma unroll
    for (int m = 0; m < DIM; ++m) {
        double xj = G.pos[DIM * j + m];
        i64 cen = argmin_abs(bins, npix, xj);
        r.dd[m] = bins[cen] - xj;
        r.lo[m] = (int)cen - r.W / 2;
        if (r.W > 0) {
            int t0, c;
            tile_run(r.lo[m], r.W, npix, T[m], t0, c);
            n *= c;
        }
    }
    r.step = (N > 1) ? ((double)N / 2.0 - (-(double)N / 2.0)) / (double)(N - 1) : 0.0;
    r.cut = G.cut[j];
    r.shift = G.lnr_shift ? G.lnr_shift[j] : 0.0;
    r.row = G.rows + j * (i64)nr;
    recs[j] = r;
    cnt[j] = n;
}

// one warp per halo: entry q of the halo -> tile (ta, tb[, tc]) of its circular runs
template <int DIM>
__global__ void __launch_bounds__(256) gg_emit_kernel(int npix, i64 nh, const gh_rec *__restrict__ recs,
                                                      const long long *__restrict__ offs,
                                                      unsigned long long *__restrict__ keys) {
    using TS = tile_shape<DIM>;
    const i64 j = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (j >= nh) return;
    const gh_rec r = recs[j];
    if (r.W <= 0) return;
    const int nta = tiles_per_axis(npix, TS::A), ntb = tiles_per_axis(npix, TS::B),
              ntc = (DIM == 3) ? tiles_per_axis(npix, TS::C) : 1;
    int a0, na, b0, nb, c0 = 0, nc = 1;
    tile_run(r.lo[0], r.W, npix, TS::A, a0, na);
    tile_run(r.lo[1], r.W, npix, TS::B, b0, nb);
    if (DIM == 3) tile_run(r.lo[2], r.W, npix, TS::C, c0, nc);
    const long long base = offs[j], total = (long long)na * nb * nc;
    for (long long q = lane; q < total; q += 32) {
        int qa = (int)(q / ((long long)nb * nc)), rem = (int)(q - (long long)qa * nb * nc), qb = rem / nc, qc = rem - qb * nc;
        int ta = (a0 + qa) % nta, tb = (b0 + qb) % ntb, tc = (c0 + qc) % ntc;
        unsigned long long tile = ((unsigned long long)ta * ntb + tb) * ntc + tc;
        keys[base + q] = (tile << 32) | (unsigned long long)(unsigned)j;
    }
}

__global__ void __launch_bounds__(256) gg_bounds_kernel(long long n, const unsigned long long *__restrict__ keys,
                                                        int *__restrict__ tstart, int *__restrict__ tend, gg_hdr *hdr) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { hdr->sorted_keys = keys; hdr->nentries = n; }
    if (i >= n) return;
    unsigned t = (unsigned)(keys[i] >> 32);
    if (i == 0 || (unsigned)(keys[i - 1] >> 32) != t) tstart[t] = (int)i;
    if (i == n - 1 || (unsigned)(keys[i + 1] >> 32) != t) tend[t] = (int)(i + 1);
}

// ln r grid, reciprocal cell widths and the ln table, once per call (every CTA then only copies them)
__global__ void __launch_bounds__(256) gg_axis_kernel(bf_raxis ax, double *__restrict__ axis) {
    const int t = threadIdx.x, nr = ax.nr;
    for (int k = t; k < nr; k += 256) axis[k] = ax.g[k];
    for (int k = t; k < nr - 1; k += 256) axis[nr + k] = 1.0 / (ax.g[k + 1] - ax.g[k]);
    if (t == 0) axis[2 * nr - 1] = 0.0;
    bf_logtab_init(reinterpret_cast<double2 *>(axis + 2 * nr), t, 256);
}

struct gg_args {
    int npix;
    int a_first, a_last;  // rows (first grid axis) this call owns; tile_first = a_first / tile edge
    int tile_first;
    double res, inv_res;
    const double *axis;
    bf_raxis ax;
    const gh_rec *recs;
    const gg_hdr *hdr;
    const int *tstart, *tend;
    const double *Rmat;
    double *out;
    u64 *npairs;
};

template <int DIM, int MODE>  // MODE 0 = baryonify, 1 = paint
__global__ void __launch_bounds__(GG_THREADS, 2) grid_gather_kernel(gg_args A) {
    using TS = tile_shape<DIM>;
    static_assert(TS::A * TS::B * TS::C == GG_THREADS, "one cell per thread");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tile = blockIdx.x + A.tile_first;
    const int s0 = A.tstart[tile], s1 = A.tend[tile];
    if (s0 >= s1) return;  // no halo reaches this tile
    gh_rec *chunk = reinterpret_cast<gh_rec *>(smem_raw);
    double2 *logtab = reinterpret_cast<double2 *>(smem_raw + sizeof(gh_rec) * GG_CHUNK);
    double *g_s = reinterpret_cast<double *>(logtab + BF_LOGTAB);
    double *ig_s = g_s + A.ax.nr;
    const int t = threadIdx.x, npix = A.npix, nr = A.ax.nr;
    for (int k = t; k < nr; k += GG_THREADS) {
        g_s[k] = __ldg(A.axis + k);
        ig_s[k] = __ldg(A.axis + nr + k);
    }
    for (int k = t; k < 2 * BF_LOGTAB; k += GG_THREADS) reinterpret_cast<double *>(logtab)[k] = __ldg(A.axis + 2 * nr + k);
    bf_raxis ax = A.ax;
    ax.g = g_s;
    // this thread's cell
    const int ntb = tiles_per_axis(npix, TS::B), ntc = (DIM == 3) ? tiles_per_axis(npix, TS::C) : 1;
    int tta, ttb, ttc = 0;
    if (DIM == 3) { ttc = tile % ntc; ttb = (tile / ntc) % ntb; tta = tile / (ntc * ntb); }
    else { ttb = tile % ntb; tta = tile / ntb; }
    int la, lb, lc = 0;
    if (DIM == 3) { lc = t % TS::C; lb = (t / TS::C) % TS::B; la = t / (TS::C * TS::B); }
    else { lb = t % TS::B; la = t / TS::B; }
    const int ca = tta * TS::A + la, cb = ttb * TS::B + lb, cc = ttc * TS::C + lc;
    const bool active = ca >= A.a_first && ca < A.a_last && cb < npix && (DIM == 2 || cc < npix);
    const unsigned long long *keys = A.hdr->sorted_keys;
    const double res = A.res;
    double acc[3] = {0.0, 0.0, 0.0};
    unsigned my_pairs = 0;
    for (int base = s0; base < s1; base += GG_CHUNK) {
        const int n = min(GG_CHUNK, s1 - base);
        __syncthreads();
        if (t < n * 10) {  // 80-byte records as 10 eight-byte words
            const int h = t / 10, wd = t - h * 10;
            const unsigned j = (unsigned)keys[base + h];
            reinterpret_cast<double *>(chunk + h)[wd] = reinterpret_cast<const double *>(A.recs + j)[wd];
        }
        __syncthreads();
        if (!active) continue;
        for (int i = 0; i < n; ++i) {
            const gh_rec &H = chunk[i];
            const int W = H.W;
            int ia = ca - H.lo[0];
            ia += (ia < 0) ? npix : 0;
            ia -= (ia >= npix) ? npix : 0;
            int ib = cb - H.lo[1];
            ib += (ib < 0) ? npix : 0;
            ib -= (ib >= npix) ? npix : 0;
            int ic = 0;
            if (DIM == 3) {
                ic = cc - H.lo[2];
                ic += (ic < 0) ? npix : 0;
                ic -= (ic >= npix) ? npix : 0;
            }
            if (ia >= W || ib >= W || ic >= W) continue;
            ++my_pairs;
            // np.linspace(-W/2, W/2, W)[i] * res + (bins[cen] - x_j); x pairs with the SECOND index (meshgrid 'xy')
            const int NL = H.N;
            const double start = -(double)NL / 2.0, stop = (double)NL / 2.0, step = H.step;
            const double gx = ((ib == NL - 1) ? stop : ((double)ib * step + start)) * res + H.dd[0];
            const double gy = ((ia == NL - 1) ? stop : ((double)ia * step + start)) * res + H.dd[1];
            const double gz = (DIM == 3) ? ((ic == NL - 1) ? stop : ((double)ic * step + start)) * res + H.dd[2] : 0.0;
            const double r2 = (DIM == 3) ? gx * gx + gy * gy + gz * gz : gx * gx + gy * gy;
            double rt2 = r2;
            if (DIM == 2 && A.Rmat) {  // build_Rmat shear (:525-530 / :758-763): the model sees |coords @ Rmat|
                const double *R = A.Rmat + 4 * (i64)H.j;
                const double ex = gx * R[0] + gy * R[2], ey = gx * R[1] + gy * R[3];
                rt2 = ex * ex + ey * ey;
            }
            const double x = fma(0.5, bf_fastlog(logtab, rt2), -H.shift);
            double v = nan("");
            if (x >= ax.g0 && x <= ax.glast) {
                int k = (int)((x - ax.g0) * ax.inv_dx);
                k = max(0, min(k, nr - 2));
                double gk = g_s[k];
                if (x < gk || (x >= g_s[k + 1] && k < nr - 2)) {  // rounding at a node / non-uniform axis
                    k = bf_find_cell(ax, x);
                    gk = g_s[k];
                }
                const double tt = (x - gk) * ig_s[k];
                v = __ldg(H.row + k) * (1.0 - tt) + __ldg(H.row + k + 1) * tt;
            }
            // r < cut (BaryonCorrection.py:381, Map2DRunner.py:804) decided on the squares; only a radius within
            // rounding of the cut pays the square root the reference compares with
            const double c2 = H.cut * H.cut;
            bool inside = rt2 < c2;
            if (fabs(rt2 - c2) <= 1e-12 * c2) inside = sqrt(rt2) < H.cut;
            if (MODE == 0) {
                v = inside ? v : 0.0;  // BaryonCorrection.py:381-382
                // :534 / :577: offset = d / res along the unit vector (round radius also for sheared halos)
                const double q = (v * A.inv_res) * rsqrt(r2);
                acc[0] += q * gx;
                acc[1] += q * gy;
                if (DIM == 3) acc[2] += q * gz;
            } else {
                v = exp(v);
                if (bf_finite(v) && inside && v != 0.0) acc[0] += v;  // :803-810
            }
        }
    }
    if (active) {
        const i64 cell = (DIM == 2) ? (i64)ca * npix + cb : ((i64)ca * npix + cb) * npix + cc;
        if (MODE == 0) {
#pragma unroll
            for (int m = 0; m < DIM; ++m) A.out[DIM * cell + m] += acc[m];
        } else {
            A.out[cell] += acc[0];
        }
    }
    if (A.npairs) {
        for (int s = 16; s > 0; s >>= 1) my_pairs += __shfl_down_sync(0xffffffffu, my_pairs, s);
        if ((t & 31) == 0 && my_pairs) atomicAdd(A.npairs, (u64)my_pairs);
    }
}

static size_t gather_smem(int nr) { return sizeof(gh_rec) * GG_CHUNK + sizeof(double2) * BF_LOGTAB + sizeof(double) * 2 * (size_t)nr; }

template <int DIM, int MODE>
static int launch(const gg_args &A, unsigned ntiles, cudaStream_t s) {
    size_t smem = gather_smem(A.ax.nr);
    auto kern = grid_gather_kernel<DIM, MODE>;
    BF_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<ntiles, GG_THREADS, smem, s>>>(A);
    BF_LAUNCH_OK();
    return BF_OK;
}

}  // namespace

extern "C" size_t bf_grid_bin_ws_bytes(int dim, int64_t npix, int64_t nh, int64_t max_entries) {
    if ((dim != 2 && dim != 3) || npix < 1 || nh < 0 || max_entries < 0) return 0;
    return gg_layout(nullptr, dim, npix, nh, max_entries).total;
}

extern "C" int bf_grid_bin_halos(int dim, int64_t npix, const double *bins, const bf_grid_halos *G, int64_t nr,
                                 int64_t max_entries, void *ws, size_t ws_bytes, int64_t *nentries_host, void *stream) {
    BF_REQUIRE(dim == 2 || dim == 3, "dim must be 2 or 3");
    BF_REQUIRE(npix >= 5 && npix < (1 << 24), "5 <= npix < 2^24");
    BF_REQUIRE(bins && G && G->nh >= 0 && G->nh < (1ll << 31) && nentries_host, "bins / halo record / nentries_host");
    BF_REQUIRE(G->nh == 0 || (G->pos && G->Nsize && G->cut && G->rows), "null per-halo array");
    BF_REQUIRE(max_entries >= 0 && max_entries < (1ll << 31), "0 <= max_entries < 2^31");
    gg_ws w = gg_layout(ws, dim, npix, G->nh, max_entries);
    if (!ws || ws_bytes < w.total) {
        bf_set_error("bf_grid_bin_halos: workspace %zu < %zu bytes", ws_bytes, w.total);
        return BF_E_WORKSPACE;
    }
    cudaStream_t s = (cudaStream_t)stream;
    const i64 nh = G->nh;
    const int ta = tiles_per_axis((int)npix, dim == 2 ? tile_shape<2>::A : tile_shape<3>::A),
              tb = tiles_per_axis((int)npix, dim == 2 ? tile_shape<2>::B : tile_shape<3>::B),
              tc = dim == 3 ? tiles_per_axis((int)npix, tile_shape<3>::C) : 1;
    const size_t nt = (size_t)ta * tb * tc;
    BF_REQUIRE(nt < (1ull << 31), "too many tiles");
    int *flag = (int *)w.hdr + 6;  // spare bytes of the (256-byte padded) header slot
    BF_CUDA_OK(cudaMemsetAsync(w.hdr, 0, 64, s));
    BF_CUDA_OK(cudaMemsetAsync(w.tstart, 0, sizeof(int) * nt, s));
    BF_CUDA_OK(cudaMemsetAsync(w.tend, 0, sizeof(int) * nt, s));
    unsigned hb = (unsigned)((nh + 1 + 255) / 256);
    if (dim == 2) gg_prep_kernel<2><<<hb, 256, 0, s>>>((int)npix, bins, *G, (int)nr, w.recs, w.cnt, flag);
    else gg_prep_kernel<3><<<hb, 256, 0, s>>>((int)npix, bins, *G, (int)nr, w.recs, w.cnt, flag);
    BF_LAUNCH_OK();
    size_t cb = w.cub_bytes;
    BF_CUDA_OK(cub::DeviceScan::ExclusiveSum(w.cub_tmp, cb, w.cnt, w.offs, (int)(nh + 1), s));
    long long total = 0;
    int too_wide = 0;
    BF_CUDA_OK(cudaMemcpyAsync(&total, w.offs + nh, sizeof(long long), cudaMemcpyDeviceToHost, s));
    BF_CUDA_OK(cudaMemcpyAsync(&too_wide, flag, sizeof(int), cudaMemcpyDeviceToHost, s));
    BF_CUDA_OK(cudaStreamSynchronize(s));
    *nentries_host = total;
    if (too_wide) {
        bf_set_error("bf_grid_bin_halos: a cut-out is wider than the grid (Nsize > npix); use the halo-centric kernels");
        return BF_E_WINDOW;
    }
    if (total > max_entries) {
        bf_set_error("bf_grid_bin_halos: %lld (tile, halo) entries exceed max_entries %lld", total, (long long)max_entries);
        return BF_E_WORKSPACE;
    }
    if (total == 0) return BF_OK;
    unsigned wb = (unsigned)((nh * 32 + 255) / 256);
    if (dim == 2) gg_emit_kernel<2><<<wb, 256, 0, s>>>((int)npix, nh, w.recs, w.offs, w.keys);
    else gg_emit_kernel<3><<<wb, 256, 0, s>>>((int)npix, nh, w.recs, w.offs, w.keys);
    BF_LAUNCH_OK();
    int hbits = 1, tbits = 1;
    while ((1ll << hbits) < nh) ++hbits;
    while ((1ull << tbits) < nt) ++tbits;
    cb = w.cub_bytes;
    // halo bits only up to hbits, tile bits from 32: sort the two fields in two passes' worth of digits
    BF_CUDA_OK(cub::DeviceRadixSort::SortKeys(w.cub_tmp, cb, w.keys, w.keys_alt, (int)total, 0, 32 + tbits, s));
    (void)hbits;
    gg_bounds_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(total, w.keys_alt, w.tstart, w.tend, w.hdr);
    BF_LAUNCH_OK();
    return BF_OK;
}

extern "C" int bf_grid_gather(int dim, int mode, int64_t npix, double res, const bf_grid_halos *G, const bf_table *T,
                              int64_t max_entries, void *ws, double *out, unsigned long long *npairs, void *stream) {
    return bf_grid_gather_rows(dim, mode, npix, res, G, T, max_entries, ws, out, 0, npix, npairs, stream);
}

extern "C" int bf_grid_gather_rows(int dim, int mode, int64_t npix, double res, const bf_grid_halos *G,
                                   const bf_table *T, int64_t max_entries, void *ws, double *out, int64_t a_first,
                                   int64_t a_last, unsigned long long *npairs, void *stream) {
    BF_REQUIRE(dim == 2 || dim == 3, "dim must be 2 or 3");
    BF_REQUIRE(a_first >= 0 && a_first <= a_last && a_last <= npix, "0 <= a_first <= a_last <= npix");
    if (a_first == a_last) return BF_OK;
    BF_REQUIRE(mode == 0 || mode == 1, "mode: 0 baryonify, 1 paint");
    BF_REQUIRE(npix >= 5 && out && res > 0 && ws && G, "npix/out/res/ws");
    BF_REQUIRE(bf_check_table(T), "bad table");
    BF_REQUIRE(dim == 2 || G->Rmat == nullptr, "ellipticity (Rmat) is only defined for 2-D grids");
    if (G->nh == 0) return BF_OK;
    gg_ws w = gg_layout(ws, dim, npix, G->nh, max_entries);
    gg_args A;
    A.npix = (int)npix;
    A.res = res;
    A.inv_res = 1.0 / res;
    A.ax = bf_make_raxis(T);
    BF_REQUIRE(A.ax.nr <= GG_MAX_NR, "radial axis too long for the shared-memory staging");
    A.axis = w.axis;
    A.recs = w.recs; A.hdr = w.hdr; A.tstart = w.tstart; A.tend = w.tend;
    A.Rmat = G->Rmat; A.out = out; A.npairs = npairs;
    const int ta = tiles_per_axis((int)npix, dim == 2 ? tile_shape<2>::A : tile_shape<3>::A),
              tb = tiles_per_axis((int)npix, dim == 2 ? tile_shape<2>::B : tile_shape<3>::B),
              tc = dim == 3 ? tiles_per_axis((int)npix, tile_shape<3>::C) : 1;
    const int TA = dim == 2 ? tile_shape<2>::A : tile_shape<3>::A;
    const int t_first = (int)(a_first / TA), t_last = (int)((a_last + TA - 1) / TA);
    (void)ta;
    A.a_first = (int)a_first; A.a_last = (int)a_last;
    A.tile_first = t_first * tb * tc;
    const unsigned nt = (unsigned)((size_t)(t_last - t_first) * tb * tc);
    cudaStream_t s = (cudaStream_t)stream;
    gg_axis_kernel<<<1, 256, 0, s>>>(A.ax, w.axis);
    BF_LAUNCH_OK();
    if (dim == 2) return mode == 0 ? launch<2, 0>(A, nt, s) : launch<2, 1>(A, nt, s);
    return mode == 0 ? launch<3, 0>(A, nt, s) : launch<3, 1>(A, nt, s);
}
