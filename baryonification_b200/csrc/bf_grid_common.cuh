// bf_grid_common.cuh -- cut-out geometry shared by the halo-centric (bf_grid.cu) and the cell-owner
// (bf_grid_gather.cu) grid kernels: Runners/Map2DRunner.py:494-522, :540-560, pick_indices :394-423.
#pragma once
#include "bf_common.cuh"

__device__ __forceinline__ double cut_coord(i64 i, i64 N, double res) {  // np.linspace(-N/2, N/2, N)[i]*res
    double start = -(double)N / 2.0, stop = (double)N / 2.0;
    double step = (stop - start) / (double)(N - 1);
    double y = (i == N - 1) ? stop : ((double)i * step + start);
    return y * res;
}

__device__ __forceinline__ i64 wrap_once(i64 i, i64 N) {  // pick_indices (:394-423)
    if (i < 0) i += N;
    if (i >= N) i -= N;
    return i;
}

// np.argmin(np.abs(bins - x)): first index of the minimum.  bins are ascending, so a bisection lands
// next to it and the exact |.| comparison over a +-2 window reproduces the first-minimum rule.
__device__ __forceinline__ i64 argmin_abs(const double *__restrict__ bins, i64 n, double x) {
    i64 lo = 0, hi = n - 1;
    while (hi - lo > 1) {
        i64 mid = (lo + hi) >> 1;
        if (bins[mid] <= x) lo = mid; else hi = mid;
    }
    i64 b0 = (lo - 2 > 0) ? lo - 2 : 0, b1 = (hi + 2 < n - 1) ? hi + 2 : n - 1;
    i64 best = b0;
    double bv = fabs(bins[b0] - x);
    for (i64 i = b0 + 1; i <= b1; ++i) {
        double v = fabs(bins[i] - x);
        if (v < bv) { bv = v; best = i; }
    }
    return best;
}

