// bf_snapshot.cu -- particle-snapshot baryonification as a deterministic particle-centric gather.
//
// Replaces BaryonifySnapshot (Runners/SnapshotRunner.py): the periodic KD-tree build (:92-97), the per-halo
// query_ball_point + displacement loop (:205-247) and the apply / wrap pass (:250-260).
// Instead of "for each halo: find its particles, scatter-add", every particle looks up the halos whose
// query ball can reach it: halos are binned into a periodic cell list with cell size >= max R_q, so the
// 3^D neighbouring cells hold every candidate; membership is the KD-tree's own criterion (minimum-image
// distance^2 <= R_q^2).  Contributions are summed in registers in a fixed order (cell order, then
// catalogue order inside a cell): no atomics, bit-reproducible, one coalesced read and one coalesced
// write per particle coordinate -- the streaming HBM bound of SURVEY.md section 8d (88 B / particle).
#include "bf_common.cuh"

#define SN_THREADS 256

struct snap_cells {
    int ng, ncell, dim;
    double inv_cs;  // ng / L
};

__device__ __forceinline__ int cell_coord(double x, const snap_cells &C) {
    int g = (int)floor(x * C.inv_cs);
    return max(0, min(g, C.ng - 1));
}

__global__ void snap_count_kernel(snap_cells C, i64 nh, const double *__restrict__ pos, int *__restrict__ count) {
    i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nh) return;
    int c = 0;
    for (int m = 0; m < C.dim; ++m) c = c * C.ng + cell_coord(pos[C.dim * j + m], C);
    atomicAdd(count + c, 1);
}

// exclusive scan of count[0..n) into start[0..n], single block
__global__ void __launch_bounds__(1024) snap_scan_kernel(int n, const int *__restrict__ count, int *__restrict__ start) {
    __shared__ int part[1024];
    int t = threadIdx.x, per = (n + 1023) / 1024;
    int b = t * per, e = min(n, b + per), s = 0;
    for (int i = b; i < e; ++i) s += count[i];
    part[t] = s;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {
        int add = (t >= d) ? part[t - d] : 0;
        __syncthreads();
        part[t] += add;
        __syncthreads();
    }
    int run = part[t] - s;
    for (int i = b; i < e; ++i) { start[i] = run; run += count[i]; }
    if (t == 1023) start[n] = part[1023];
}

__global__ void snap_fill_kernel(snap_cells C, i64 nh, const double *__restrict__ pos, const int *__restrict__ start,
                                 int *__restrict__ cursor, int *__restrict__ list) {
    i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nh) return;
    int c = 0;
    for (int m = 0; m < C.dim; ++m) c = c * C.ng + cell_coord(pos[C.dim * j + m], C);
    list[start[c] + atomicAdd(cursor + c, 1)] = (int)j;
}

// catalogue order inside each cell (makes the summation order independent of the atomics above)
__global__ void snap_sort_kernel(int ncell, const int *__restrict__ start, int *__restrict__ list) {
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncell) return;
    int b = start[c], e = start[c + 1];
    for (int i = b + 1; i < e; ++i) {
        int v = list[i], k = i - 1;
        while (k >= b && list[k] > v) { list[k + 1] = list[k]; --k; }
        list[k + 1] = v;
    }
}

__device__ __forceinline__ double wrap_half(double dx, double L) {  // :124-125, :152-153
    if (dx > L / 2) dx = dx - L;
    if (dx < -L / 2) dx = dx + L;
    return dx;
}

__device__ __forceinline__ double wrap_box(double v, double L) {  // :259-260
    if (v > L) v = v - L;
    if (v < 0) v = v + L;
    return v;
}

template <int DIM>
__global__ void __launch_bounds__(SN_THREADS) snap_gather_kernel(
    snap_cells C, i64 np, const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ z,
    double L, double a, bf_snap_halos H, bf_raxis ax, const int *__restrict__ start, const int *__restrict__ list,
    double *__restrict__ xo, double *__restrict__ yo, double *__restrict__ zo, u64 *npairs) {
    i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    u64 my_pairs = 0;
    if (p < np) {
        double P[3] = {x[p], y[p], (DIM == 3) ? z[p] : 0.0};
        int g[3] = {cell_coord(P[0], C), cell_coord(P[1], C), (DIM == 3) ? cell_coord(P[2], C) : 0};
        double tot[3] = {0.0, 0.0, 0.0};
        const int span = (C.ng >= 3) ? 3 : C.ng;  // ng < 3: visit every cell once instead of +-1 with repeats
        for (int ia = 0; ia < span; ++ia)
            for (int ib = 0; ib < span; ++ib)
                for (int ic = 0; ic < ((DIM == 3) ? span : 1); ++ic) {
                    int ca = (C.ng >= 3) ? (g[0] + ia - 1 + C.ng) % C.ng : ia;
                    int cb = (C.ng >= 3) ? (g[1] + ib - 1 + C.ng) % C.ng : ib;
                    int cc = (DIM == 3) ? ((C.ng >= 3) ? (g[2] + ic - 1 + C.ng) % C.ng : ic) : 0;
                    int cell = (DIM == 3) ? (ca * C.ng + cb) * C.ng + cc : ca * C.ng + cb;
                    for (int s = start[cell]; s < start[cell + 1]; ++s) {
                        int j = list[s];
                        double dx[3] = {0, 0, 0}, dsq = 0.0;
#pragma unroll
                        for (int m = 0; m < DIM; ++m) {
                            dx[m] = wrap_half(P[m] - H.pos[DIM * j + m], L);  // :235-238
                            dsq += dx[m] * dx[m];
                        }
                        double R = H.Rq[j];
                        if (!(dsq <= R * R)) continue;  // query_ball_point (:219 / :234)
                        double d = sqrt(dsq);
                        double v = bf_row_lerp(ax, H.rows + (i64)j * ax.nr, log(d) - H.lnr_shift[j]);
                        v = (d < H.cut[j]) ? v : 0.0;  // BaryonCorrection.py:381-382
                        double off = v * a;            // :228 / :245  (x a_j, reproduced as is)
                        if (!bf_finite(off)) off = 0.0;
#pragma unroll
                        for (int m = 0; m < DIM; ++m) tot[m] += off * (dx[m] / d);  // :247 (0/0 -> NaN kept)
                        ++my_pairs;
                    }
                }
        xo[p] = wrap_box(P[0] + tot[0], L);
        yo[p] = wrap_box(P[1] + tot[1], L);
        if (DIM == 3) zo[p] = wrap_box(P[2] + tot[2], L);
    }
    for (int s = 16; s > 0; s >>= 1) my_pairs += __shfl_down_sync(0xffffffffu, my_pairs, s);
    if ((threadIdx.x & 31) == 0 && npairs && my_pairs) atomicAdd(npairs, my_pairs);
}

static snap_cells make_cells(int dim, double L, double Rq_max) {
    snap_cells C;
    C.dim = dim;
    int ngmax = (dim == 3) ? 96 : 1024;
    double n = (Rq_max > 0) ? floor(L / Rq_max) : (double)ngmax;
    if (!(n >= 1)) n = 1;
    if (n > ngmax) n = ngmax;
    C.ng = (int)n;
    C.ncell = (dim == 3) ? C.ng * C.ng * C.ng : C.ng * C.ng;
    C.inv_cs = (double)C.ng / L;
    return C;
}

static size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

extern "C" size_t bf_snap_ws_bytes(int dim, int64_t nh, double L, double Rq_max) {
    if ((dim != 2 && dim != 3) || nh < 0 || !(L > 0)) return 0;
    snap_cells C = make_cells(dim, L, Rq_max);
    return 3 * align256(sizeof(int) * ((size_t)C.ncell + 1)) + align256(sizeof(int) * (size_t)(nh + 1));
}

extern "C" int bf_snap_displace(int dim, int64_t np, const double *x, const double *y, const double *z, double L,
                                double a, const bf_snap_halos *H, const bf_table *T, double Rq_max, double *xo,
                                double *yo, double *zo, unsigned long long *npairs, void *ws, size_t ws_bytes,
                                void *stream) {
    BF_REQUIRE(dim == 2 || dim == 3, "dim must be 2 or 3");
    BF_REQUIRE(np >= 0 && L > 0, "np / L");
    BF_REQUIRE(bf_check_table(T), "bad table");
    BF_REQUIRE(H && H->nh >= 0 && H->nh < (1ll << 31), "halo record");
    if (np == 0) return BF_OK;
    BF_REQUIRE(x && y && xo && yo && (dim == 2 || (z && zo)), "null particle array");
    BF_REQUIRE(H->nh == 0 || (H->pos && H->Rq && H->cut && H->lnr_shift && H->rows), "null per-halo array");
    size_t need = bf_snap_ws_bytes(dim, H->nh, L, Rq_max);
    if (ws_bytes < need || !ws) {
        bf_set_error("bf_snap_displace: workspace %zu < %zu bytes", ws_bytes, need);
        return BF_E_WORKSPACE;
    }
    cudaStream_t s = (cudaStream_t)stream;
    snap_cells C = make_cells(dim, L, Rq_max);
    size_t csz = align256(sizeof(int) * ((size_t)C.ncell + 1));
    char *w = (char *)ws;
    int *count = (int *)w, *start = (int *)(w + csz), *cursor = (int *)(w + 2 * csz), *list = (int *)(w + 3 * csz);
    BF_CUDA_OK(cudaMemsetAsync(count, 0, csz, s));
    BF_CUDA_OK(cudaMemsetAsync(cursor, 0, csz, s));
    BF_CUDA_OK(cudaMemsetAsync(start, 0, csz, s));
    if (H->nh > 0) {
        unsigned hb = (unsigned)((H->nh + 255) / 256);
        snap_count_kernel<<<hb, 256, 0, s>>>(C, H->nh, H->pos, count);
        BF_LAUNCH_OK();
        snap_scan_kernel<<<1, 1024, 0, s>>>(C.ncell, count, start);
        BF_LAUNCH_OK();
        snap_fill_kernel<<<hb, 256, 0, s>>>(C, H->nh, H->pos, start, cursor, list);
        BF_LAUNCH_OK();
        snap_sort_kernel<<<(unsigned)((C.ncell + 255) / 256), 256, 0, s>>>(C.ncell, start, list);
        BF_LAUNCH_OK();
    }
    unsigned pb = (unsigned)((np + SN_THREADS - 1) / SN_THREADS);
    bf_raxis ax = bf_make_raxis(T);
    if (dim == 2)
        snap_gather_kernel<2><<<pb, SN_THREADS, 0, s>>>(C, np, x, y, z, L, a, *H, ax, start, list, xo, yo, zo, npairs);
    else
        snap_gather_kernel<3><<<pb, SN_THREADS, 0, s>>>(C, np, x, y, z, L, a, *H, ax, start, list, xo, yo, zo, npairs);
    BF_LAUNCH_OK();
    return BF_OK;
}
