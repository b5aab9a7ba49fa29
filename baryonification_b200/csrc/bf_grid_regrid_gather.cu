// bf_grid_regrid_gather.cu -- deterministic re-binning of periodic grids as a DESTINATION gather.
//
// Replaces BaryonifyGrid's tail (Runners/Map2DRunner.py:583-607 with regrid_pixels_2D/_3D, :14-163): every source
// cell deposits its mass onto the 2^D cells around its displaced position (the CIC decomposition derived in
// bf_grid.cu).  The reference's numba loops run over the source cells in ascending flat index and add into the new
// map one deposit at a time, so a destination's value is the left-to-right sum of its deposits in ascending SOURCE
// order.  bf_grid_regrid_sorted reproduces that order with a counting sort of all 2^D * N^D deposits (25 B of
// workspace per deposit: 215 GB at 1024^3).  Here the loop is inverted instead: a first pass measures, per axis, the
// range of floor(displaced position) - cell over the whole grid; a destination can then only receive from the
// sources inside a window of w = hi - lo + 2 cells per axis, and it visits them in ascending (periodic-aware) source
// order, recomputing each source's deposit with exactly the arithmetic of the scatter kernels.  Same bits as the
// sorted variant, no workspace, no atomics, one coalesced write per cell.
//
// A CTA owns a TA x TB x TC tile of destinations and stages the displaced positions + masses of the
// (T + w - 1)^D sources around it in shared memory once (fmod and the offset reads are paid ~2x per cell instead
// of w^D x); the w^D-long inner loop is integer compares on the staged floor(position), and only the <= 2^D
// sources that do deposit here pay the fp64 weight arithmetic.
#include <climits>

#include "bf_common.cuh"

namespace {

__device__ __forceinline__ double pymod_g(double a, double n) {  // python float %, n > 0
    if (a >= 0 && a < n) return a;  // fmod of an in-range value is the value itself
    double r = fmod(a, n);
    if (r < 0) r += n;
    return r;
}

__device__ __forceinline__ int wrap_i(int v, int N) {  // v in [-N, 2N) -> [0, N)
    if (v < 0) v += N;
    if (v >= N) v -= N;
    return v;
}

// displaced position of source cell (a, b, c), per grid axis (axis 0 pairs with offset component 1: meshgrid 'xy')
template <int DIM>
__device__ __forceinline__ void source_position(int N, int a, int b, int c, const double *__restrict__ off, double s[3]) {
    i64 cell = (DIM == 2) ? (i64)a * N + b : ((i64)a * N + b) * N + c;
    double o[3] = {0, 0, 0};
#pragma unroll
    for (int m = 0; m < DIM; ++m) {
        double v = __ldg(off + DIM * cell + m);
        o[m] = bf_finite(v) ? v : 0.0;  // Map2DRunner.py:591 / :601
    }
    s[0] = pymod_g(o[1] + (double)a, (double)N);
    s[1] = pymod_g(o[0] + (double)b, (double)N);
    s[2] = (DIM == 3) ? pymod_g(o[2] + (double)c, (double)N) : 0.0;
}

__global__ void window_init_kernel(int *mm) {
    if (threadIdx.x < 3) mm[threadIdx.x] = INT_MAX;
    else if (threadIdx.x < 6) mm[threadIdx.x] = INT_MIN;
}

// mm[0..2] = min, mm[3..5] = max over all cells of (floor(position) - cell coordinate), min-image, per grid axis
template <int DIM>
__global__ void __launch_bounds__(256) window_kernel(int N, const double *__restrict__ off, i64 cell_first,
                                                     i64 cell_last, int *__restrict__ mm) {
    int lo[3] = {INT_MAX, INT_MAX, INT_MAX}, hi[3] = {INT_MIN, INT_MIN, INT_MIN};
    for (i64 cell = cell_first + (i64)blockIdx.x * blockDim.x + threadIdx.x; cell < cell_last;
         cell += (i64)gridDim.x * blockDim.x) {
        int a, b, c = 0;
        if (DIM == 2) { a = (int)(cell / N); b = (int)(cell - (i64)a * N); }
        else { a = (int)(cell / ((i64)N * N)); i64 rem = cell - (i64)a * N * N; b = (int)(rem / N); c = (int)(rem - (i64)b * N); }
        double s[3];
        source_position<DIM>(N, a, b, c, off, s);
        const int co[3] = {a, b, c};
#pragma unroll
        for (int m = 0; m < DIM; ++m) {
            int i0 = (int)s[m];
            if (i0 >= N) i0 -= N;
            int d = i0 - co[m];
            if (2 * d > N) d -= N;
            else if (2 * d < -N) d += N;
            lo[m] = min(lo[m], d);
            hi[m] = max(hi[m], d);
        }
    }
#pragma unroll
    for (int m = 0; m < DIM; ++m) {
#pragma unroll
        for (int k = 16; k; k >>= 1) {
            lo[m] = min(lo[m], __shfl_xor_sync(0xffffffffu, lo[m], k));
            hi[m] = max(hi[m], __shfl_xor_sync(0xffffffffu, hi[m], k));
        }
        if ((threadIdx.x & 31) == 0) {
            atomicMin(mm + m, lo[m]);
            atomicMax(mm + 3 + m, hi[m]);
        }
    }
}

struct gg_args {
    int N;
    const double *off, *orig;
    double *out;
    int hi[3];  // source coordinate of window slot k for destination d: d - hi - 1 + k
    int w[3];   // window length per axis (1 for the unused third axis of a 2-D grid)
    int a_first, a_last;  // destination rows (first grid axis) written by this call
};

// shared-memory staging, structure of arrays over the E = EA*EB*EC sources around the tile:
// double pos[DIM][E], double mass[E], int cellidx[DIM][E]  (cellidx = floor(pos), in [0, N))
__host__ __device__ inline size_t stage_bytes(int dim, size_t E) { return E * ((dim + 1) * sizeof(double) + dim * sizeof(int)); }

template <int DIM, int TA, int TB, int TC>
__global__ void __launch_bounds__(TA * TB * TC) grid_gather_regrid_kernel(gg_args A) {
    extern __shared__ double stage_s[];
    const int N = A.N;
    const int EA = TA + A.w[0] - 1, EB = TB + A.w[1] - 1, EC = (DIM == 3) ? TC + A.w[2] - 1 : 1;
    const int E = EA * EB * EC;
    double *pos_s = stage_s, *mass_s = stage_s + DIM * E;
    int *idx_s = (int *)(stage_s + (DIM + 1) * E);
    const int oa = A.a_first + ((DIM == 3) ? blockIdx.z : blockIdx.y) * TA, ob = ((DIM == 3) ? blockIdx.y : blockIdx.x) * TB,
              oc = (DIM == 3) ? blockIdx.x * TC : 0;
    const int tid = threadIdx.x;
    // ---- stage the sources around the tile
    int fa = (oa - A.hi[0] - 1) % N, fb = (ob - A.hi[1] - 1) % N, fc = (DIM == 3) ? (oc - A.hi[2] - 1) % N : 0;
    if (fa < 0) fa += N;
    if (fb < 0) fb += N;
    if (fc < 0) fc += N;
    for (int e = tid; e < E; e += TA * TB * TC) {
        int ea = e / (EB * EC), r = e - ea * (EB * EC), eb = r / EC, ec = r - eb * EC;
        int a = (fa + ea) % N, b = (fb + eb) % N, c = (DIM == 3) ? (fc + ec) % N : 0;
        double s[3];
        source_position<DIM>(N, a, b, c, A.off, s);
        i64 cell = (DIM == 2) ? (i64)a * N + b : ((i64)a * N + b) * N + c;
#pragma unroll
        for (int m = 0; m < DIM; ++m) {
            // a position that rounded up to exactly N deposits (weight 1) on cell N % N = 0 and (weight 0) on cell 1:
            // the same deposits as position 0
            double p = (s[m] >= (double)N) ? 0.0 : s[m];
            pos_s[m * E + e] = p;
            idx_s[m * E + e] = (int)p;
        }
        mass_s[e] = __ldg(A.orig + cell);
    }
    __syncthreads();
    // ---- one destination per thread (the contiguous axis is the fastest thread index)
    int ta, tb, tc = 0;
    if (DIM == 3) { tc = tid % TC; tb = (tid / TC) % TB; ta = tid / (TC * TB); }
    else { tb = tid % TB; ta = tid / TB; }
    const int da = oa + ta, db = ob + tb, dc = oc + tc;
    if (da >= A.a_last || db >= N || (DIM == 3 && dc >= N)) return;
    // ascending source coordinate along an axis: the window is a run of w consecutive coordinates mod N; when it
    // crosses N its wrapped tail (coordinates 0, 1, ...) comes first
    const int wa = A.w[0], wb = A.w[1], wc = A.w[2];
    const int na = max(0, wrap_i(da - A.hi[0] - 1, N) + wa - N), nb = max(0, wrap_i(db - A.hi[1] - 1, N) + wb - N),
              nc = (DIM == 3) ? max(0, wrap_i(dc - A.hi[2] - 1, N) + wc - N) : 0;
    double acc = 0.0;
    bool any = false;
    for (int kka = 0; kka < wa; ++kka) {
        int ka = kka + wa - na;
        if (ka >= wa) ka -= wa;
        for (int kkb = 0; kkb < wb; ++kkb) {
            int kb = kkb + wb - nb;
            if (kb >= wb) kb -= wb;
            const int row = ((ta + ka) * EB + (tb + kb)) * EC + tc;
            for (int kkc = 0; kkc < wc; ++kkc) {
                int kc = kkc + wc - nc;
                if (kc >= wc) kc -= wc;
                const int e = row + kc;
                // which of the source's two cells per axis (floor, floor + 1) is this destination, if any
                int ia = idx_s[e], ib = idx_s[E + e], ic = (DIM == 3) ? idx_s[2 * E + e] : 0;
                int fa2 = da - ia, fb2 = db - ib, fc2 = (DIM == 3) ? dc - ic : 0;
                fa2 += (fa2 < 0) ? N : 0;
                fb2 += (fb2 < 0) ? N : 0;
                fc2 += (fc2 < 0) ? N : 0;
                if ((unsigned)(fa2 | fb2 | fc2) > 1u) continue;
                double pa = pos_s[e], pb = pos_s[E + e];
                double w_a = fa2 ? (pa + 1.0) - (double)(ia + 1) : (double)(ia + 1) - pa;
                double w_b = fb2 ? (pb + 1.0) - (double)(ib + 1) : (double)(ib + 1) - pb;
                if (!(w_a > 0) || !(w_b > 0)) continue;
                double v;
                if (DIM == 3) {
                    double pc = pos_s[2 * E + e];
                    double w_c = fc2 ? (pc + 1.0) - (double)(ic + 1) : (double)(ic + 1) - pc;
                    if (!(w_c > 0)) continue;
                    v = ((w_b * w_a) * w_c) * mass_s[e];
                } else {
                    v = (w_b * w_a) * mass_s[e];
                }
                acc = any ? acc + v : v;  // the first deposit is taken as is, like the segment head of the sorted variant
                any = true;
            }
        }
    }
    if (any) {
        i64 dst = (DIM == 2) ? (i64)da * N + db : ((i64)da * N + db) * N + dc;
        A.out[dst] += acc;
    }
}

template <int DIM, int TA, int TB, int TC>
int launch_gather(const gg_args &A, cudaStream_t s) {
    const int EA = TA + A.w[0] - 1, EB = TB + A.w[1] - 1, EC = (DIM == 3) ? TC + A.w[2] - 1 : 1;
    size_t smem = stage_bytes(DIM, (size_t)EA * EB * EC);
    if (smem > 200 * 1024) return BF_E_WINDOW;
    auto kern = grid_gather_regrid_kernel<DIM, TA, TB, TC>;
    BF_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid;
    const int rows = A.a_last - A.a_first;
    if (DIM == 3) grid = dim3((A.N + TC - 1) / TC, (A.N + TB - 1) / TB, (rows + TA - 1) / TA);
    else grid = dim3((A.N + TB - 1) / TB, (rows + TA - 1) / TA, 1);
    kern<<<grid, TA * TB * TC, smem, s>>>(A);
    BF_LAUNCH_OK();
    return BF_OK;
}

}  // namespace

extern "C" size_t bf_grid_regrid_gather_ws_bytes(void) { return 8 * sizeof(int); }

extern "C" int bf_grid_regrid_window(int dim, int64_t npix, const double *offsets, int64_t a_first, int64_t a_last,
                                     void *ws, size_t ws_bytes, int *lohi_host, void *stream) {
    BF_REQUIRE(dim == 2 || dim == 3, "dim must be 2 or 3");
    BF_REQUIRE(npix >= 5 && npix < (1 << 30) && offsets && lohi_host, "5 <= npix < 2^30 and non-null arrays");
    BF_REQUIRE(a_first >= 0 && a_first <= a_last && a_last <= npix, "0 <= a_first <= a_last <= npix");
    if (!ws || ws_bytes < 8 * sizeof(int)) {
        bf_set_error("bf_grid_regrid_window: workspace %zu < %zu bytes", ws_bytes, 8 * sizeof(int));
        return BF_E_WORKSPACE;
    }
    cudaStream_t s = (cudaStream_t)stream;
    int *mm = (int *)ws;
    const int N = (int)npix;
    const i64 rowlen = (dim == 2) ? (i64)N : (i64)N * N, c0 = a_first * rowlen, c1 = a_last * rowlen;
    window_init_kernel<<<1, 32, 0, s>>>(mm);
    BF_LAUNCH_OK();
    if (c1 > c0) {
        unsigned blocks = (unsigned)std::min<i64>((c1 - c0 + 255) / 256, 148 * 16);
        if (dim == 2) window_kernel<2><<<blocks, 256, 0, s>>>(N, offsets, c0, c1, mm);
        else window_kernel<3><<<blocks, 256, 0, s>>>(N, offsets, c0, c1, mm);
        BF_LAUNCH_OK();
    }
    BF_CUDA_OK(cudaMemcpyAsync(lohi_host, mm, 6 * sizeof(int), cudaMemcpyDeviceToHost, s));
    BF_CUDA_OK(cudaStreamSynchronize(s));
    for (int m = dim; m < 3; ++m) lohi_host[m] = lohi_host[3 + m] = 0;
    return BF_OK;
}

extern "C" int bf_grid_regrid_gather_rows(int dim, int64_t npix, const double *offsets, const double *orig,
                                          double *new_map, const int *lohi_host, int64_t a_first, int64_t a_last,
                                          int *window_host, void *stream) {
    BF_REQUIRE(dim == 2 || dim == 3, "dim must be 2 or 3");
    BF_REQUIRE(npix >= 5 && npix < (1 << 30) && offsets && orig && new_map && lohi_host,
               "5 <= npix < 2^30 and non-null arrays");
    BF_REQUIRE(a_first >= 0 && a_first <= a_last && a_last <= npix, "0 <= a_first <= a_last <= npix");
    const int N = (int)npix;
    gg_args A;
    A.N = N; A.off = offsets; A.orig = orig; A.out = new_map;
    A.a_first = (int)a_first; A.a_last = (int)a_last;
    i64 vol = 1;
    for (int m = 0; m < 3; ++m) {
        if (m < dim) { A.hi[m] = lohi_host[3 + m]; A.w[m] = lohi_host[3 + m] - lohi_host[m] + 2; }
        else { A.hi[m] = 0; A.w[m] = 1; }
        if (window_host) window_host[m] = A.w[m];
        if (A.w[m] > N || A.w[m] < 2 - (m >= dim)) {
            bf_set_error("bf_grid_regrid_gather: displacement window %d does not fit the grid", A.w[m]);
            return BF_E_WINDOW;
        }
        vol *= A.w[m];
    }
    if (vol > (dim == 2 ? BF_GRID_GATHER_MAX_WINDOW_2D : BF_GRID_GATHER_MAX_WINDOW_3D)) {
        bf_set_error("bf_grid_regrid_gather: %lld sources per cell exceed the gather limit; use bf_grid_regrid_sorted",
                     (long long)vol);
        return BF_E_WINDOW;
    }
    if (a_first == a_last) return BF_OK;
    cudaStream_t s = (cudaStream_t)stream;
    return (dim == 2) ? launch_gather<2, 16, 32, 1>(A, s) : launch_gather<3, 4, 8, 16>(A, s);
}

extern "C" int bf_grid_regrid_gather(int dim, int64_t npix, const double *offsets, const double *orig, double *new_map,
                                     void *ws, size_t ws_bytes, int *window_host, void *stream) {
    int lohi[6];
    int rc = bf_grid_regrid_window(dim, npix, offsets, 0, npix, ws, ws_bytes, lohi, stream);
    if (rc != BF_OK) return rc;
    return bf_grid_regrid_gather_rows(dim, npix, offsets, orig, new_map, lohi, 0, npix, window_host, stream);
}
