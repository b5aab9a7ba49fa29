// bf_grid.cu -- periodic 2-D / 3-D grid runners: per-halo cut-out evaluation (atomic variant) and the
// CIC-equivalent re-binning.
//
// Replaces BaryonifyGrid.process's halo loop (Runners/Map2DRunner.py:476-580), PaintProfilesGrid.process's
// halo loop (:714-810) and BaryonifyGrid's tail (:583-607 with regrid_pixels_2D/_3D, :14-163).
// Reference quirks kept on purpose (they change results above 1e-5):
//   * cut-out coordinates are np.linspace(-N/2, N/2, N)*res, i.e. spacing res*N/(N-1)   (:499, :735)
//   * meshgrid(indexing='xy') vs inds[x_inds][:, y_inds]: at flattened cut-out element (a, b[, c]) the map
//     cell is (x_inds[a], y_inds[b][, z_inds[c]]) while the coordinates are (x[b]+dx, x[a]+dy[, x[c]+dz])
//     with dx = bins[x_cen]-x_j, dy = bins[y_cen]-y_j                                     (:510-522)
//   * BaryonifyGrid does NOT zero non-finite values per halo: NaN accumulates into the cell and the cell's
//     whole offset is zeroed before re-binning (:591, :601); PaintProfilesGrid masks per halo (:803-807).
#include "bf_common.cuh"
#include "bf_grid_common.cuh"

#define GR_THREADS 256

template <int DIM, int MODE>  // MODE 0 = baryonify, 1 = paint
__global__ void __launch_bounds__(GR_THREADS) grid_atomic_kernel(i64 npix, const double *__restrict__ bins,
                                                                 double res, bf_grid_halos G, bf_raxis ax,
                                                                 double *__restrict__ out, u64 *npairs) {
    u64 my_pairs = 0;
    for (i64 j = blockIdx.x; j < G.nh; j += gridDim.x) {
        i64 N = G.Nsize[j];
        if (N < 2) continue;
        i64 w = N / 2, W = 2 * w;
        i64 cen[3] = {0, 0, 0};
        double dd[3] = {0, 0, 0};
#pragma unroll
        for (int m = 0; m < DIM; ++m) {
            double xj = G.pos[DIM * j + m];
            cen[m] = argmin_abs(bins, npix, xj);
            dd[m] = bins[cen[m]] - xj;
        }
        const double *row = G.rows + j * (i64)ax.nr;
        double cut = G.cut[j], shift = (MODE == 0) ? G.lnr_shift[j] : 0.0;
        const bool ell = (DIM == 2) && (G.Rmat != nullptr);  // build_Rmat shear (:525-530 / :758-763)
        double R00 = 1, R01 = 0, R10 = 0, R11 = 1;
        if (ell) { R00 = G.Rmat[4 * j]; R01 = G.Rmat[4 * j + 1]; R10 = G.Rmat[4 * j + 2]; R11 = G.Rmat[4 * j + 3]; }
        i64 total = (DIM == 2) ? W * W : W * W * W;
        for (i64 k = threadIdx.x; k < total; k += GR_THREADS) {
            i64 ia, ib, ic = 0;
            if (DIM == 2) { ia = k / W; ib = k - ia * W; }
            else { ia = k / (W * W); i64 rem = k - ia * W * W; ib = rem / W; ic = rem - ib * W; }
            double gx = cut_coord(ib, N, res) + dd[0];
            double gy = cut_coord(ia, N, res) + dd[1];
            double gz = (DIM == 3) ? cut_coord(ic, N, res) + dd[2] : 0.0;
            double r = (DIM == 3) ? sqrt(gx * gx + gy * gy + gz * gz) : sqrt(gx * gx + gy * gy);
            i64 cell = wrap_once(cen[0] - w + ia, npix) * npix + wrap_once(cen[1] - w + ib, npix);
            if (DIM == 3) cell = cell * npix + wrap_once(cen[2] - w + ic, npix);
            double rt = r;  // radius handed to the model; the unit vector below keeps the round radius (:520-522)
            if (ell) {
                double ex = gx * R00 + gy * R10, ey = gx * R01 + gy * R11;  // coords @ Rmat
                rt = sqrt(ex * ex + ey * ey);
            }
            double v = bf_row_lerp(ax, row, log(rt) - shift);
            if (MODE == 0) {
                v = (rt < cut) ? v : 0.0;  // BaryonCorrection.py:381-382
                double offset = v / res;  // :534 / :577
                atomicAdd(out + DIM * cell + 0, offset * (gx / r));
                atomicAdd(out + DIM * cell + 1, offset * (gy / r));
                if (DIM == 3) atomicAdd(out + DIM * cell + 2, offset * (gz / r));
            } else {
                v = exp(v);
                if (bf_finite(v) && (rt < cut) && v != 0.0) atomicAdd(out + cell, v);  // :803-810
            }
            ++my_pairs;
        }
    }
    for (int s = 16; s > 0; s >>= 1) my_pairs += __shfl_down_sync(0xffffffffu, my_pairs, s);
    if ((threadIdx.x & 31) == 0 && npairs && my_pairs) atomicAdd(npairs, my_pairs);
}

template <int MODE>
static int launch_grid(int dim, int64_t npix, const double *bins, double res, const bf_grid_halos *G,
                       const bf_table *T, double *out, u64 *npairs, void *stream) {
    BF_REQUIRE(dim == 2 || dim == 3, "dim must be 2 or 3");
    BF_REQUIRE(npix >= 5, "npix must be >= 5");
    BF_REQUIRE(bins && out && res > 0, "bins/out/res");
    BF_REQUIRE(bf_check_table(T), "bad table");
    BF_REQUIRE(G && G->nh >= 0, "halo record");
    if (G->nh == 0) return BF_OK;
    BF_REQUIRE(G->pos && G->Nsize && G->cut && G->rows && (MODE == 1 || G->lnr_shift), "null per-halo array");
    BF_REQUIRE(dim == 2 || G->Rmat == nullptr, "ellipticity (Rmat) is only defined for 2-D grids");
    i64 grid = (i64)bf_sm_count() * 8;
    if (grid > G->nh) grid = G->nh;
    cudaStream_t s = (cudaStream_t)stream;
    bf_raxis ax = bf_make_raxis(T);
    if (dim == 2) grid_atomic_kernel<2, MODE><<<(unsigned)grid, GR_THREADS, 0, s>>>(npix, bins, res, *G, ax, out, npairs);
    else grid_atomic_kernel<3, MODE><<<(unsigned)grid, GR_THREADS, 0, s>>>(npix, bins, res, *G, ax, out, npairs);
    BF_LAUNCH_OK();
    return BF_OK;
}

extern "C" int bf_grid_baryonify_atomic(int dim, int64_t npix, const double *bins, double res, const bf_grid_halos *G,
                                        const bf_table *T, double *offsets, unsigned long long *npairs, void *stream) {
    return launch_grid<0>(dim, npix, bins, res, G, T, offsets, npairs, stream);
}

extern "C" int bf_grid_paint_atomic(int dim, int64_t npix, const double *bins, double res, const bf_grid_halos *G,
                                    const bf_table *T, double *out, unsigned long long *npairs, void *stream) {
    return launch_grid<1>(dim, npix, bins, res, G, T, out, npairs, stream);
}

// ------------------------------------------------------------------------------------------
// Re-binning.  regrid_pixels_2D/_3D scan a 5-wide neighbourhood for the overlap of the displaced
// unit cell with the periodic grid; for N >= 5 the only non-empty overlaps are the 2^D cells
// floor(x), floor(x)+1 per axis with lengths (floor+1)-x and (x+1)-(floor+1): a CIC deposit.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ double pymod(double a, double n) {  // python float %, n > 0
    double r = fmod(a, n);
    if (r < 0) r += n;
    return r;
}

template <int DIM>
__global__ void __launch_bounds__(256) grid_regrid_kernel(i64 N, const double *__restrict__ off,
                                                          const double *__restrict__ orig, double *__restrict__ out) {
    i64 ncell = (DIM == 2) ? N * N : N * N * N;
    i64 cell = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= ncell) return;
    i64 a, b, c = 0;
    if (DIM == 2) { a = cell / N; b = cell - a * N; }
    else { a = cell / (N * N); i64 rem = cell - a * N * N; b = rem / N; c = rem - b * N; }
    double o[3] = {0, 0, 0};
#pragma unroll
    for (int m = 0; m < DIM; ++m) {
        double v = off[DIM * cell + m];
        o[m] = bf_finite(v) ? v : 0.0;  // :591 / :601
    }
    // x_grid = b, y_grid = a, z_grid = c (meshgrid 'xy'); grid[i, j, k] <- (y, x, z)
    double s[3] = {pymod(o[1] + (double)a, (double)N), pymod(o[0] + (double)b, (double)N),
                   (DIM == 3) ? pymod(o[2] + (double)c, (double)N) : 0.0};
    i64 i0[3]; double w0[3], w1[3];
#pragma unroll
    for (int m = 0; m < DIM; ++m) {
        i0[m] = (i64)s[m];
        w0[m] = (double)(i0[m] + 1) - s[m];          // min(j+1, x_end) - max(j, x_start), j = floor
        w1[m] = (s[m] + 1.0) - (double)(i0[m] + 1);  // j = floor + 1
    }
    double val = orig[cell];
#pragma unroll
    for (int ca = 0; ca < 2; ++ca) {
        double wa = ca ? w1[0] : w0[0];
        if (!(wa > 0)) continue;
        i64 ia = (i0[0] + ca) % N;
#pragma unroll
        for (int cb = 0; cb < 2; ++cb) {
            double wb = cb ? w1[1] : w0[1];
            if (!(wb > 0)) continue;
            i64 ib = (i0[1] + cb) % N;
            if (DIM == 2) {
                atomicAdd(out + ia * N + ib, (wb * wa) * val);  // overlap_area = dx*dy, dx along axis 1
            } else {
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                    double wc = cc ? w1[2] : w0[2];
                    if (!(wc > 0)) continue;
                    i64 ic = (i0[2] + cc) % N;
                    atomicAdd(out + (ia * N + ib) * N + ic, ((wb * wa) * wc) * val);
                }
            }
        }
    }
}

extern "C" int bf_grid_regrid(int dim, int64_t npix, const double *offsets, const double *orig, double *new_map,
                              void *stream) {
    BF_REQUIRE(dim == 2 || dim == 3, "dim must be 2 or 3");
    BF_REQUIRE(npix >= 5 && offsets && orig && new_map, "npix >= 5 and non-null arrays");
    i64 ncell = (dim == 2) ? npix * npix : npix * npix * npix;
    unsigned blocks = (unsigned)((ncell + 255) / 256);
    if (dim == 2) grid_regrid_kernel<2><<<blocks, 256, 0, (cudaStream_t)stream>>>(npix, offsets, orig, new_map);
    else grid_regrid_kernel<3><<<blocks, 256, 0, (cudaStream_t)stream>>>(npix, offsets, orig, new_map);
    BF_LAUNCH_OK();
    return BF_OK;
}
