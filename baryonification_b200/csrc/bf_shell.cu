// bf_shell.cu -- HEALPix shell runners, halo-centric ATOMIC variant + re-binning.
//
// Replaces the serial per-halo loops of PaintProfilesShell.process (Runners/HealpixRunner.py:418-445)
// and BaryonifyShell.process (:291-331), and the global re-binning stage (:333-341 + regrid_pixels_hpix
// :13-67).  One CTA owns one halo at a time (dynamic queue, largest discs first when `order` is given):
//   1. 64 rings at a time, one thread per ring runs the query_disc ring walk (bf_healpix.cuh) and
//      writes the ring's pixel segment(s) to shared memory;
//   2. a block scan flattens the segments so that all 256 threads evaluate pairs with full lanes;
//   3. each pair: pix2vec -> chord distance -> row lerp -> fp64 RED.ADD into the map / offsets.
// The deterministic gather variant lives in bf_shell_gather.cu; this one is the "atomic variant
// benchmarked alongside" and the cross-check for it.
#include "bf_healpix.cuh"
#include "bf_pair.cuh"

#define SH_THREADS 256
#define SH_RCH 64  // rings per chunk -> 2*SH_RCH segments

struct shell_smem {
    i64 seg_start[2 * SH_RCH];
    int seg_pref[2 * SH_RCH + 1];
    i64 ring_id[SH_RCH];
    int scan_tmp[SH_THREADS];
    i64 halo;
    unsigned long long total;
};

__device__ __forceinline__ int block_excl_scan(int v, int *tmp, int &total) {
    // Hillis-Steele over SH_THREADS ints in shared memory
    int t = threadIdx.x;
    tmp[t] = v;
    __syncthreads();
    for (int s = 1; s < SH_THREADS; s <<= 1) {
        int add = (t >= s) ? tmp[t - s] : 0;
        __syncthreads();
        tmp[t] += add;
        __syncthreads();
    }
    total = tmp[SH_THREADS - 1];
    int excl = tmp[t] - v;
    __syncthreads();
    return excl;
}

template <int MODE>  // 0 = paint, 1 = baryonify
__device__ __forceinline__ void eval_pixel(const bf_hpx &h, const bf_halo_rec &H, const bf_raxis &ax, i64 pix,
                                           const double v[3], double *__restrict__ out) {
    if (MODE == 0) {
        double P = bf_pair_paint_shell(H, ax, v);
        if (P != 0.0) atomicAdd(out + pix, P);
    } else {
        double off[3];
        bf_pair_baryonify_shell(H, ax, v, off);
        atomicAdd(out + 3 * pix + 0, off[0]);
        atomicAdd(out + 3 * pix + 1, off[1]);
        atomicAdd(out + 3 * pix + 2, off[2]);
    }
}

template <int MODE>
__global__ void __launch_bounds__(SH_THREADS) shell_atomic_kernel(bf_hpx h, bf_shell_halos S, bf_raxis ax,
                                                                  double *__restrict__ out, u64 *npairs,
                                                                  unsigned int *queue, double resol) {
    __shared__ shell_smem sm;
    const int t = threadIdx.x;
    u64 my_pairs = 0;
    for (;;) {
        if (t == 0) sm.halo = (i64)atomicAdd(queue, 1u);
        __syncthreads();
        i64 q = sm.halo;
        __syncthreads();
        if (q >= S.nh) break;
        i64 j = S.order ? S.order[q] : q;
        bf_halo_rec H;
        H.c[0] = S.vec[3 * j]; H.c[1] = S.vec[3 * j + 1]; H.c[2] = S.vec[3 * j + 2];
        H.D = S.D[j]; H.a = S.a[j];
        H.cut = (MODE == 1) ? S.cut[j] : 0.0;
        H.lnr_shift = (MODE == 1) ? S.lnr_shift[j] : 0.0;
        H.row = S.rows + j * (i64)ax.nr;
        double radius = S.radius[j];
        bf_disc d = bfh_disc_setup(h, H.c, radius);

        // ---- BaryonifyShell only: "< 4 pixels -> 4 interpolation neighbours" (:309-310) ----
        if (MODE == 1 && radius < 3.0 * resol) {
            // count the disc first (tiny discs only: at most a handful of rings)
            i64 cnt = 0;
            for (i64 iz = d.irmin + t; iz <= d.irmax; iz += SH_THREADS) {
                i64 a0, a1, b0, b1;
                cnt += bfh_disc_ring(h, d, iz, a0, a1, b0, b1);
            }
            int tot;
            block_excl_scan((int)cnt, sm.scan_tmp, tot);
            i64 total = tot + d.north_end + (h.npix - d.south_begin);
            if (total < 4) {
                if (t < 4) {
                    i64 p[4]; double w[4], v[3];
                    bfh_interp_lonlat(h, S.lon[j], S.lat[j], p, w);
                    bfh_pix2vec(h, p[t], v);
                    eval_pixel<MODE>(h, H, ax, p[t], v, out);
                    ++my_pairs;
                }
                continue;
            }
        }

        // ---- whole polar caps (north/south pole inside the disc) ----
        for (i64 p = t; p < d.north_end; p += SH_THREADS) {
            double v[3]; bfh_pix2vec(h, p, v);
            eval_pixel<MODE>(h, H, ax, p, v, out); ++my_pairs;
        }
        for (i64 p = d.south_begin + t; p < h.npix; p += SH_THREADS) {
            double v[3]; bfh_pix2vec(h, p, v);
            eval_pixel<MODE>(h, H, ax, p, v, out); ++my_pairs;
        }

        // ---- ring walk, SH_RCH rings per chunk ----
        for (i64 iz0 = d.irmin; iz0 <= d.irmax; iz0 += SH_RCH) {
            int len = 0; i64 start = 0;
            // thread t handles segment t: ring t>>1, part t&1 (threads of one ring recompute the walk)
            if (t < 2 * SH_RCH) {
                i64 iz = iz0 + (t >> 1);
                if (iz <= d.irmax) {
                    i64 a0, a1, b0, b1;
                    bfh_disc_ring(h, d, iz, a0, a1, b0, b1);
                    start = (t & 1) ? b0 : a0;
                    len = (int)((t & 1) ? (b1 - b0) : (a1 - a0));
                    if (!(t & 1)) sm.ring_id[t >> 1] = iz;
                }
            }
            int total;
            int excl = block_excl_scan(len, sm.scan_tmp, total);
            if (t < 2 * SH_RCH) { sm.seg_start[t] = start; sm.seg_pref[t] = excl; }
            if (t == 0) sm.seg_pref[2 * SH_RCH] = total;
            __syncthreads();
            for (int k = t; k < total; k += SH_THREADS) {
                int lo = 0, hi = 2 * SH_RCH;  // seg_pref[lo] <= k < seg_pref[hi]
                while (hi - lo > 1) {
                    int mid = (lo + hi) >> 1;
                    if (sm.seg_pref[mid] <= k) lo = mid; else hi = mid;
                }
                i64 pix = sm.seg_start[lo] + (k - sm.seg_pref[lo]);
                i64 iz = sm.ring_id[lo >> 1], sp, nr; bool sh;
                bfh_ring_info_small(h, iz, sp, nr, sh);
                double z, phi, sth, v[3]; bool have;
                bfh_ringpix2loc(h, iz, pix - sp + 1, z, phi, sth, have);
                bfh_loc2vec(z, phi, sth, have, v);
                eval_pixel<MODE>(h, H, ax, pix, v, out);
                ++my_pairs;
            }
            __syncthreads();
        }
    }
    // pair count: warp reduce then one atomic per warp
    for (int s = 16; s > 0; s >>= 1) my_pairs += __shfl_down_sync(0xffffffffu, my_pairs, s);
    if ((t & 31) == 0 && npairs && my_pairs) atomicAdd(npairs, my_pairs);
}

static int check_shell(const bf_shell_halos *S, int mode) {
    if (!S || S->nh < 0) return 0;
    if (S->nh == 0) return 1;
    if (!S->vec || !S->radius || !S->D || !S->a || !S->rows) return 0;
    if (mode == 1 && (!S->lon || !S->lat || !S->cut || !S->lnr_shift)) return 0;
    return 1;
}

template <int MODE>
static int launch_shell(int64_t nside, const bf_shell_halos *S, const bf_table *T, double *out, u64 *npairs,
                        void *stream) {
    BF_REQUIRE(nside > 0, "nside");
    BF_REQUIRE(bf_check_table(T), "bad table");
    BF_REQUIRE(check_shell(S, MODE), "bad halo record");
    BF_REQUIRE(out != nullptr, "out is NULL");
    if (S->nh == 0) return BF_OK;
    cudaStream_t s = (cudaStream_t)stream;
    static unsigned int *queue = nullptr;  // 4-byte work-queue head (allocated once per process)
    if (!queue) BF_CUDA_OK(cudaMalloc(&queue, sizeof(unsigned int)));
    BF_CUDA_OK(cudaMemsetAsync(queue, 0, sizeof(unsigned int), s));
    bf_hpx h = bf_hpx_make(nside);
    double resol = sqrt(4.0 * BFH_PI / (double)h.npix);
    i64 grid = (i64)bf_sm_count() * 8;
    if (grid > S->nh) grid = S->nh;
    shell_atomic_kernel<MODE><<<(unsigned)grid, SH_THREADS, 0, s>>>(h, *S, bf_make_raxis(T), out, npairs, queue, resol);
    BF_LAUNCH_OK();
    return BF_OK;
}

extern "C" int bf_shell_paint_atomic(int64_t nside, const bf_shell_halos *S, const bf_table *T, double *map,
                                     unsigned long long *npairs, void *stream) {
    return launch_shell<0>(nside, S, T, map, npairs, stream);
}

extern "C" int bf_shell_baryonify_atomic(int64_t nside, const bf_shell_halos *S, const bf_table *T, double *pix_offsets,
                                         unsigned long long *npairs, void *stream) {
    return launch_shell<1>(nside, S, T, pix_offsets, npairs, stream);
}

// ------------------------------------------------------------------------------------------
// Re-binning: new_vec = pix2vec(p) + offsets[p]; vec2ang(lonlat) -> degrees; get_interp_weights
// (lonlat) -> 4 neighbours; new_map[nbr] += w * orig[p] for pixels with orig > 0.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) shell_regrid_kernel(bf_hpx h, i64 pix_first, i64 pix_last,
                                                           const double *__restrict__ orig,
                                                           const double *__restrict__ off, double *__restrict__ out) {
    i64 q = (i64)blockIdx.x * blockDim.x + threadIdx.x;  // local index into orig / off
    i64 p = pix_first + q;
    if (p >= pix_last) return;
    double m = orig[q];
    if (!(m > 0)) return;  // :335
    double v[3];
    bfh_pix2vec(h, p, v);
    double x = v[0] + off[3 * q], y = v[1] + off[3 * q + 1], z = v[2] + off[3 * q + 2];
    // healpy.vec2ang(lonlat=True): theta = arccos(z/|v|), phi = arctan2(y, x) (+2pi if < 0), in degrees
    double dnorm = sqrt(x * x + y * y + z * z);
    double theta = acos(z / dnorm);
    double phi = atan2(y, x);
    if (phi < 0) phi += BFH_TWOPI;
    double lon = phi * BFH_RAD2DEG, lat = 90.0 - theta * BFH_RAD2DEG;
    i64 cp[4]; double w[4];
    bfh_interp_lonlat(h, lon, lat, cp, w);
#pragma unroll
    for (int k = 0; k < 4; ++k) atomicAdd(out + cp[k], w[k] * m);
}

extern "C" int bf_shell_regrid(int64_t nside, const double *orig_map, const double *pix_offsets, double *new_map,
                               void *stream) {
    return bf_shell_regrid_range(nside, 0, 12 * nside * nside, orig_map, pix_offsets, new_map, stream);
}

extern "C" int bf_shell_regrid_range(int64_t nside, int64_t pix_first, int64_t pix_last, const double *orig_map,
                                     const double *pix_offsets, double *new_map, void *stream) {
    BF_REQUIRE(nside > 0 && orig_map && pix_offsets && new_map, "null / nside");
    bf_hpx h = bf_hpx_make(nside);
    BF_REQUIRE(pix_first >= 0 && pix_first <= pix_last && pix_last <= h.npix, "pixel range");
    i64 n = pix_last - pix_first;
    if (n == 0) return BF_OK;
    shell_regrid_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(h, pix_first, pix_last, orig_map,
                                                                                      pix_offsets, new_map);
    BF_LAUNCH_OK();
    return BF_OK;
}
