// bf_pair.cuh -- evaluation of ONE halo-pixel pair, literal to the reference's arithmetic.
#pragma once
#include "bf_common.cuh"

struct bf_halo_rec {  // per-halo constants broadcast to every pair of that halo
    double c[3];      // vec_j
    double D, a, cut, lnr_shift;
    const double *row;
};

// PaintProfilesShell body, HealpixRunner.py:435-442: returns Paint (non-finite -> 0)
__device__ __forceinline__ double bf_pair_paint_shell(const bf_halo_rec &H, const bf_raxis &ax, const double v[3]) {
    double d0 = v[0] * H.D - H.c[0] * H.D, d1 = v[1] * H.D - H.c[1] * H.D, d2 = v[2] * H.D - H.c[2] * H.D;
    double r_sep = sqrt(d0 * d0 + d1 * d1 + d2 * d2);
    double P = exp(bf_row_lerp(ax, H.row, log(r_sep / H.a)));
    return bf_finite(P) ? P : 0.0;
}

// BaryonifyShell body, HealpixRunner.py:314-328: off[3] = shift of the pixel's unit vector
__device__ __forceinline__ void bf_pair_baryonify_shell(const bf_halo_rec &H, const bf_raxis &ax, const double v[3],
                                                        double off[3]) {
    double pos[3] = {v[0] * H.D, v[1] * H.D, v[2] * H.D};
    double diff[3] = {pos[0] - H.c[0] * H.D, pos[1] - H.c[1] * H.D, pos[2] - H.c[2] * H.D};
    double r_sep = sqrt(diff[0] * diff[0] + diff[1] * diff[1] + diff[2] * diff[2]);
    double r = r_sep / H.a;                                   // physical -> comoving (:321)
    double d = bf_row_lerp(ax, H.row, log(r) - H.lnr_shift);  // BaryonCorrection.py:374-379
    d = (r < H.cut) ? d : 0.0;                                // BaryonCorrection.py:381-382
    double offset = d * H.a;
    double nw[3];
#pragma unroll
    for (int m = 0; m < 3; ++m) {
        double o = offset * (diff[m] / r_sep);  // :322
        if (!bf_finite(o)) o = 0.0;             // :323
        nw[m] = pos[m] + o;                     // :326
    }
    double nrm = sqrt(nw[0] * nw[0] + nw[1] * nw[1] + nw[2] * nw[2]);
#pragma unroll
    for (int m = 0; m < 3; ++m) off[m] = nw[m] / nrm - v[m];  // :327-328
}
