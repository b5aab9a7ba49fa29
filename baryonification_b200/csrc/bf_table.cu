// bf_table.cu -- (z, M) collapse of a tabulated model and the 1-D row readout.
//
// Reference: BaryonificationClass._readout (Profiles/BaryonCorrection.py:356-390) and
// TabulatedProfile._readout (utils/Tabulate.py:272-294), both a scipy RegularGridInterpolator
// ('linear', bounds_error=False, fill NaN) evaluated at (ln(1/a), ln M, ln r).  z and M are constant
// per halo, so the 4 (z,M) corners are blended ONCE per halo into an nr-long row (this kernel, HBM
// bound: reads <= 4 rows, writes 1 row per halo) and each halo-pixel pair is a 1-D lerp of that row.
#include "bf_common.cuh"

// scipy find_interval_ascending + clip, for one coordinate; returns -1 for out-of-range / NaN
__device__ __forceinline__ int axis_cell(const double *g, int n, double x, double &t) {
    if (!(x >= g[0] && x <= g[n - 1])) return -1;
    int lo = 0, hi = n - 1;
    if (x == g[n - 1]) lo = n - 2;
    else
        while (hi - lo > 1) {
            int mid = (lo + hi) >> 1;
            if (x < g[mid]) hi = mid; else lo = mid;
        }
    t = (x - g[lo]) / (g[lo + 1] - g[lo]);
    return lo;
}

// one CTA per halo (grid-stride); threads stride over r
__global__ void __launch_bounds__(128) collapse_kernel(bf_table T, i64 nh, const double *__restrict__ z_in,
                                                       const double *__restrict__ M_in, double *__restrict__ rows) {
    for (i64 j = blockIdx.x; j < nh; j += gridDim.x) {
        double tz = 0, tM = 0;
        int iz = axis_cell(T.gz, (int)T.nz, z_in[j], tz);
        int iM = axis_cell(T.gM, (int)T.nM, M_in[j], tM);
        double *out = rows + j * T.nr;
        if (iz < 0 || iM < 0) {
            for (int k = threadIdx.x; k < T.nr; k += blockDim.x) out[k] = nan("");
            continue;
        }
        // weights in scipy's product order: (1*wz)*wM, then *wr at readout
        double w00 = (1.0 - tz) * (1.0 - tM), w01 = (1.0 - tz) * tM, w10 = tz * (1.0 - tM), w11 = tz * tM;
        const double *r00 = T.values + ((i64)iz * T.nM + iM) * T.nr;
        const double *r01 = r00 + T.nr, *r10 = r00 + T.nM * T.nr, *r11 = r10 + T.nr;
        for (int k = threadIdx.x; k < T.nr; k += blockDim.x)
            out[k] = ((r00[k] * w00 + r01[k] * w01) + r10[k] * w10) + r11[k] * w11;
    }
}

extern "C" int bf_table_collapse_zM(const bf_table *t, int64_t nh, const double *z_in, const double *M_in,
                                    double *rows, void *stream) {
    BF_REQUIRE(bf_check_table(t), "bad table");
    BF_REQUIRE(nh >= 0 && (nh == 0 || (z_in && M_in && rows)), "null per-halo array");
    if (nh == 0) return BF_OK;
    int grid = (int)((nh < (i64)bf_sm_count() * 16) ? nh : (i64)bf_sm_count() * 16);
    collapse_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(*t, nh, z_in, M_in, rows);
    BF_LAUNCH_OK();
    return BF_OK;
}

// N-d version: axes (z, M, extras...), r last.  Corner c's bit (naxes-1-a) selects the upper node of axis a, so the
// corners are visited in itertools.product order (first axis slowest), like scipy's _evaluate_linear.
__global__ void __launch_bounds__(128) collapse_nd_kernel(bf_table_nd T, i64 nh, const double *__restrict__ coords,
                                                          double *__restrict__ rows) {
    for (i64 j = blockIdx.x; j < nh; j += gridDim.x) {
        int cell[BF_MAX_AXES];
        double t[BF_MAX_AXES];
        bool oob = false;
        for (int a = 0; a < T.naxes; ++a) {
            cell[a] = axis_cell(T.grid[a], (int)T.n[a], coords[j * T.naxes + a], t[a]);
            if (cell[a] < 0) oob = true;
        }
        double *out = rows + j * T.nr;
        if (oob) {
            for (int k = threadIdx.x; k < T.nr; k += blockDim.x) out[k] = nan("");
            continue;
        }
        const int ncorner = 1 << T.naxes;
        for (int k = threadIdx.x; k < T.nr; k += blockDim.x) {
            double acc = 0.0;
            for (int c = 0; c < ncorner; ++c) {
                double w = 1.0;
                i64 off = 0;
                for (int a = 0; a < T.naxes; ++a) {
                    int bit = (c >> (T.naxes - 1 - a)) & 1;
                    w = w * (bit ? t[a] : 1.0 - t[a]);
                    off = off * T.n[a] + (cell[a] + bit);
                }
                double term = T.values[off * T.nr + k] * w;
                acc = (c == 0) ? term : acc + term;
            }
            out[k] = acc;
        }
    }
}

extern "C" int bf_table_collapse_nd(const bf_table_nd *t, int64_t nh, const double *coords, double *rows, void *stream) {
    BF_REQUIRE(t && t->values && t->naxes >= 1 && t->naxes <= BF_MAX_AXES && t->nr >= 2, "bad N-d table");
    for (int a = 0; a < t->naxes; ++a) BF_REQUIRE(t->grid[a] && t->n[a] >= 2, "N-d table axis needs >= 2 ascending points");
    BF_REQUIRE(nh >= 0 && (nh == 0 || (coords && rows)), "null per-halo array");
    if (nh == 0) return BF_OK;
    int grid = (int)((nh < (i64)bf_sm_count() * 16) ? nh : (i64)bf_sm_count() * 16);
    collapse_nd_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(*t, nh, coords, rows);
    BF_LAUNCH_OK();
    return BF_OK;
}

__global__ void readout_kernel(bf_raxis ax, const double *__restrict__ row, i64 n, const double *__restrict__ r,
                               double lnr_shift, double cut, int mode, double *__restrict__ out) {
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double v = bf_row_lerp(ax, row, log(r[i]) - lnr_shift);
    if (mode == 1) v = exp(v);
    else v = (r[i] < cut) ? v : 0.0;
    out[i] = v;
}

extern "C" int bf_table_readout(const bf_table *t, const double *row, int64_t n, const double *r, double lnr_shift,
                                double cut, int mode, double *out, void *stream) {
    BF_REQUIRE(bf_check_table(t), "bad table");
    BF_REQUIRE(n >= 0 && (n == 0 || (row && r && out)), "null array");
    if (n == 0) return BF_OK;
    readout_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(bf_make_raxis(t), row, n, r, lnr_shift,
                                                                                   cut, mode, out);
    BF_LAUNCH_OK();
    return BF_OK;
}
