// bf_snapshot.cu -- particle-snapshot baryonification as a deterministic particle-centric gather.
//
// Replaces BaryonifySnapshot (Runners/SnapshotRunner.py): the periodic KD-tree build (:92-97), the per-halo
// query_ball_point + displacement loop (:205-247) and the apply / wrap pass (:250-260).
//
// Instead of "for each halo: find its particles, scatter-add", every particle looks up the halos whose query
// ball can reach it.  Halos are binned into a FINE periodic grid (cell ~ one mean query radius): a halo is
// listed in every cell its ball overlaps (halo binning into grid tiles, north_star kernel work (1)), the
// (cell, halo) keys are radix-sorted by cell so that each cell's list is contiguous and in catalogue order.
// A particle then reads ONE cell list (most are empty: query balls cover a few per cent of the box) and tests
// membership with the KD-tree's own criterion (minimum-image distance^2 <= R_q^2).
//
// The particle kernel (snap_gather_kernel below) is a per-warp pipeline without block barriers:
//   1. stream: read the records of 2 x 32 particles, look up their cells; particles no halo reaches are written
//      straight back, the others are parked in a shared-memory queue;
//   2. whenever 32 particles are parked, one lane each: test the candidate halos (the first 4 with independent
//      loads), a warp scan gives every hit a slot in a second queue;
//   3. the pair queue is evaluated densely (sqrt, log, table lerp, divisions run with full lanes although only
//      ~1 particle in 10 has a pair) and every particle adds ITS pairs in slot order = catalogue order: the
//      reference's summation order (tot_offsets[inds] += ... for j in range(N_halo)), no atomics, bit-reproducible.
// HBM traffic: one coalesced read and one coalesced write of the particle records (32 B each way for the
// reference's (M, x, y, z) float64 records; SURVEY.md section 8d counts 88 B / particle for the reference's
// separate offsets array); the randomly accessed tables (4 B / cell, 4 B / entry, 32 B / halo) are kept small
// enough to live in L2 and the records are streamed with evict-first hints.
#include <cub/cub.cuh>
#include <stdlib.h>

#include "bf_common.cuh"

#define SN_THREADS 256
#define SN_NG_MAX_3D 256
#define SN_NG_MAX_2D 4096

struct snap_grid {
    int ng, dim;
    i64 ncell;
    double inv_cs, cs, L;
};

static snap_grid make_grid(int dim, double L, i64 ng) {
    snap_grid G;
    G.dim = dim;
    G.ng = (int)ng;
    G.ncell = (dim == 3) ? ng * ng * ng : ng * ng;
    G.inv_cs = (double)ng / L;
    G.cs = L / (double)ng;
    G.L = L;
    return G;
}

__device__ __forceinline__ int snap_cell_coord(double x, const snap_grid &G) {
    int g = __double2int_rd(x * G.inv_cs);  // floor; x = L (allowed by the single wrap, :259-260) lands in cell 0
    if ((unsigned)g >= (unsigned)G.ng) { g %= G.ng; if (g < 0) g += G.ng; }
    return g;
}

struct snap_hdr {  // first 256 bytes of the workspace, written by bf_snap_bin_halos, read by the particle kernel
    const int *entj;  // [nentries] halo index of every (cell, halo) entry, cell-sorted, catalogue order inside a cell;
                      // the last entry of a cell is stored as ~j
};

struct snap_ws {
    snap_hdr *hdr;
    i64 *counts, *offsets;
    int *cell_tab;     // [ncell] 0 = no halo reaches the cell; -(j+1) = only halo j; (s << 6 | min(n, 63)) = list entj[s .. s+n)
    double2 *hpar;     // [nh] (cut, lnr_shift)
    double4 *hrec;     // [nh] (x, y, z, R_q): one 32-byte sector per membership test, L2-resident (3.2 MB at 1e5 halos)
    u64 *keys, *keys_alt;
    int *entj;
    void *cub_tmp;
    size_t cub_bytes, total;
};

static size_t sn_al256(size_t v) { return (v + 255) & ~(size_t)255; }

static snap_ws snap_layout(void *base, const snap_grid &G, i64 nh, i64 max_entries) {
    snap_ws w;
    char *p = (char *)base;
    size_t o = 0;
    w.hdr = (snap_hdr *)(p + o);     o += 256;
    w.counts = (i64 *)(p + o);       o += sn_al256(sizeof(i64) * (size_t)(nh + 1));
    w.offsets = (i64 *)(p + o);      o += sn_al256(sizeof(i64) * (size_t)(nh + 2));
    w.cell_tab = (int *)(p + o);     o += sn_al256(sizeof(int) * (size_t)(G.ncell + 1));
    w.hpar = (double2 *)(p + o);     o += sn_al256(sizeof(double2) * (size_t)(nh + 1));
    w.hrec = (double4 *)(p + o);     o += sn_al256(sizeof(double4) * (size_t)(nh + 1));
    w.keys = (u64 *)(p + o);         o += sn_al256(sizeof(u64) * (size_t)(max_entries + 1));
    w.keys_alt = (u64 *)(p + o);     o += sn_al256(sizeof(u64) * (size_t)(max_entries + 1));
    w.entj = (int *)(p + o);         o += sn_al256(sizeof(int) * (size_t)(max_entries + 1));
    size_t a = 0, b = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, a, (i64 *)nullptr, (i64 *)nullptr, (int)(nh + 1));
    cub::DoubleBuffer<u64> db(nullptr, nullptr);
    cub::DeviceRadixSort::SortKeys(nullptr, b, db, (int)(max_entries > 0 ? max_entries : 1), 0, 64);
    w.cub_bytes = (a > b ? a : b) + 256;
    w.cub_tmp = (void *)(p + o);     o += sn_al256(w.cub_bytes);
    w.total = o;
    return w;
}

// ---------------------------------------------------------------------------------------------------
// halo binning: one warp per halo enumerates the cells of its ball's bounding box (unwrapped indices), keeps
// those whose box is within R_q of the centre (slightly inflated: the test must never drop a cell that holds
// a particle the exact membership test would accept) and emits (cell << 32 | halo).
// ---------------------------------------------------------------------------------------------------
template <int DIM, bool FILL>
__global__ void __launch_bounds__(128) snap_bin_kernel(snap_grid G, bf_snap_halos H, i64 *__restrict__ counts,
                                                       const i64 *__restrict__ offsets, u64 *__restrict__ keys,
                                                       double4 *__restrict__ hrec, double2 *__restrict__ hpar) {
    i64 j = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (j >= H.nh) return;
    double p[3] = {H.pos[DIM * j], H.pos[DIM * j + 1], (DIM == 3) ? H.pos[DIM * j + 2] : 0.0};
    double R = H.Rq[j];
    if (!FILL && lane == 0) {
        hrec[j] = make_double4(p[0], p[1], p[2], R);
        hpar[j] = make_double2(H.cut ? H.cut[j] : 0.0, H.lnr_shift ? H.lnr_shift[j] : 0.0);
    }
    if (!(R >= 0.0)) { if (!FILL && lane == 0) counts[j] = 0; return; }  // NaN / negative radius: no particles
    i64 lo[3] = {0, 0, 0}, n[3] = {1, 1, 1};
    bool full[3] = {false, false, false};
#pragma unroll
    for (int m = 0; m < DIM; ++m) {
        i64 g0 = (i64)floor((p[m] - R) * G.inv_cs - 1e-9);
        i64 g1 = (i64)floor((p[m] + R) * G.inv_cs + 1e-9);
        lo[m] = g0;
        n[m] = g1 - g0 + 1;
        if (n[m] >= G.ng) { lo[m] = 0; n[m] = G.ng; full[m] = true; }
    }
    const i64 nbox = n[0] * n[1] * n[2];
    const double R2 = R * R * (1.0 + 1e-9) + 1e-300;
    i64 total = 0;
    const i64 base = FILL ? offsets[j] : 0;
    for (i64 b0 = 0; b0 < nbox; b0 += 32) {
        i64 idx = b0 + lane;
        bool hit = false;
        i64 cell = 0;
        if (idx < nbox) {
            i64 i0 = idx / (n[1] * n[2]), rem = idx - i0 * (n[1] * n[2]);
            i64 i1 = rem / n[2], i2 = rem - i1 * n[2];
            i64 u[3] = {lo[0] + i0, lo[1] + i1, lo[2] + i2};
            double dsq = 0.0;
#pragma unroll
            for (int m = 0; m < DIM; ++m) {
                if (!full[m]) {
                    double a = (double)u[m] * G.cs - p[m], b = p[m] - (double)(u[m] + 1) * G.cs;
                    double d = fmax(0.0, fmax(a, b)) - 1e-9 * G.cs;
                    if (d > 0.0) dsq += d * d;
                }
                i64 g = u[m] % G.ng;
                if (g < 0) g += G.ng;
                cell = cell * G.ng + g;
            }
            hit = dsq <= R2;
        }
        unsigned bal = __ballot_sync(0xffffffffu, hit);
        if (FILL && hit) keys[base + total + __popc(bal & ((1u << lane) - 1u))] = ((u64)cell << 32) | (u64)j;
        total += __popc(bal);
    }
    if (!FILL && lane == 0) counts[j] = total;
}

// after the sort: the cell table and the per-entry halo indices.  The tables a particle touches at random
// (cell_tab 4 B/cell, entj 4 B/entry, hrec 32 B/halo) are small enough to stay L2-resident while the particle
// records stream through with evict-first hints; a cell reached by ONE halo (the common case) names it in the
// cell table itself, which saves a dependent load.
__global__ void snap_entries_kernel(i64 n, const u64 *__restrict__ keys, int *__restrict__ cell_tab,
                                    int *__restrict__ entj) {
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    u64 k = keys[i];
    u64 c = k >> 32;
    int j = (int)(k & 0xffffffffull);
    bool first = (i == 0) || (keys[i - 1] >> 32) != c, last = (i == n - 1) || (keys[i + 1] >> 32) != c;
    if (first) {
        if (last) cell_tab[c] = -(j + 1);
        else {
            int nn = 1;  // list length, saturating at 63 (the kernel follows the ~j sentinel beyond SN_K entries)
            while (nn < 63 && i + nn < n && (keys[i + nn] >> 32) == c) ++nn;
            cell_tab[c] = (int)(i << 6) | nn;
        }
    }
    entj[i] = last ? ~j : j;
}

__global__ void snap_hdr_kernel(snap_hdr *hdr, const int *entj) { hdr->entj = entj; }

// ---------------------------------------------------------------------------------------------------
// particle kernel
// ---------------------------------------------------------------------------------------------------
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

// A warp streams a chunk of SN_TILES x 32 consecutive particles.  Particles whose cell no halo reaches (the
// majority) are written straight back out; the others are parked in a per-warp shared-memory queue and processed
// 32 at a time, so that the candidate tests run with every lane busy although only ~1 particle in 3 has a
// candidate; their hits go to a second queue that is evaluated densely (sqrt, log, table lerp, divisions), and
// every particle adds ITS pairs in slot order = catalogue order.  No block barrier anywhere: warps never wait for
// each other, which is what hides the record -> cell -> halo latency chain.
#ifndef SN_TILES
#define SN_TILES 16
#endif
#ifndef SN_MINBLOCKS
#define SN_MINBLOCKS 3
#endif
#define SN_CHUNK (32 * SN_TILES)       // particles per warp
#define SN_TILE (SN_CHUNK * (SN_THREADS / 32))  // particles per block
#define SN_AQ 96   // parked particles per warp (drained whenever 32 are waiting: <= 31 left over + 2 tiles)
#define SN_WQ 64   // queued pairs per warp per round
#define SN_K 4     // candidates handled with independent loads before the sentinel-terminated tail loop

struct snap_args {
    snap_grid G;
    i64 np;
    const double *x, *y, *z;  // SoA input (REC = false)
    double *xo, *yo, *zo;
    const double4 *rec;       // (M, x, y, z) records (REC = true)
    double4 *rec_out;
    double a;
    bf_snap_halos H;
    bf_raxis ax;
    const int *cell_tab;
    const double4 *hrec;
    const double2 *hpar;
    const snap_hdr *hdr;
    u64 *npairs;
    u64 *pair_list;    // optional (bf_snap_pair_list): (particle << 32 | halo) of every pair, in arrival order
    u64 *pair_cursor;
    i64 pair_cap;
    int wide;          // records in and out are 32-byte aligned: one 256-bit store per record (STG.256, sm_100)
};

struct __align__(16) snap_q {  // one queued pair: min-image separation in, displacement vector out (in place)
    double v[3];
    int j, src;
};

struct __align__(8) snap_act {  // one parked particle
    double M, x, y, z;
    int tab, idx;  // cell-table word, particle index relative to the warp's chunk
};

struct snap_smem {
    snap_act act[SN_THREADS / 32][SN_AQ];
    snap_q q[SN_THREADS / 32][SN_WQ];
};

__device__ __forceinline__ double4 ldg_rec(const double4 *p) {  // one 32-byte sector through the read-only path
    const double2 *q = reinterpret_cast<const double2 *>(p);
    double2 a = __ldg(q), b = __ldg(q + 1);
    return make_double4(a.x, a.y, b.x, b.y);
}

template <int DIM>
__device__ __forceinline__ bool snap_member(const double P[3], const double4 h, double L, double dx[3]) {
    dx[0] = wrap_half(P[0] - h.x, L);  // :235-238
    dx[1] = wrap_half(P[1] - h.y, L);
    double dsq = dx[0] * dx[0] + dx[1] * dx[1];
    dx[2] = 0.0;
    if (DIM == 3) { dx[2] = wrap_half(P[2] - h.z, L); dsq += dx[2] * dx[2]; }
    return dsq <= h.w * h.w;  // query_ball_point (:219 / :234)
}

template <int DIM, bool REC>
__device__ __forceinline__ void snap_store(const snap_args &A, i64 p, double M, double nx, double ny, double nz) {
    if (REC) {
        if (A.wide) {  // one full 32-byte sector per lane, evict-first
            asm volatile("st.global.cs.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(A.rec_out + p), "d"(M), "d"(nx), "d"(ny), "d"(nz)
                         : "memory");
        } else {
            double2 *o = reinterpret_cast<double2 *>(A.rec_out + p);
            __stcs(o, make_double2(M, nx));
            __stcs(o + 1, make_double2(ny, nz));
        }
    } else {
        __stcs(A.xo + p, nx); __stcs(A.yo + p, ny);
        if (DIM == 3) __stcs(A.zo + p, nz);
    }
}

// snap_act::idx -> particle index.  The chunked kernel numbers a warp's particles from the start of its chunk; the
// streaming kernel numbers them along the warp's own sequence of tiles (tile n of warp gw is tile
// (gw + (n / SS_CHUNK_TILES) W) SS_CHUNK_TILES + n % SS_CHUNK_TILES of the array: chunks are dealt round robin).
#define SS_CHUNK_TILES 16
struct snap_map_chunk {
    i64 chunk0;
    __device__ __forceinline__ i64 operator()(int idx) const { return chunk0 + idx; }
};
struct snap_map_stream {
    i64 gw, W;
    __device__ __forceinline__ i64 tile(int n) const {
        return (gw + (i64)(n / SS_CHUNK_TILES) * W) * SS_CHUNK_TILES + (n % SS_CHUNK_TILES);
    }
    __device__ __forceinline__ i64 operator()(int idx) const { return tile(idx >> 5) * 32 + (idx & 31); }
};

// Process the first `nact` (<= 32) parked particles of this warp, one per lane.  Returns the number of pairs.
template <int DIM, bool REC, class MAP>
__device__ __forceinline__ int snap_drain(const snap_args &A, const int *__restrict__ entj, const snap_act *act,
                                          snap_q *wq, int nact, const MAP &mp, int lane) {
    const double L = A.G.L;
    const bool on = lane < nact;
    double P[3] = {0.0, 0.0, 0.0}, M = 0.0;
    int tab = 0, idx = 0;
    if (on) {
        const snap_act &e = act[lane];
        M = e.M; P[0] = e.x; P[1] = e.y; P[2] = e.z; tab = e.tab; idx = e.idx;
    }
    // candidate halos: the first SN_K with independent loads, the rest through the sentinel-terminated list
    int jj[SN_K], n4 = 0, tail = -1, hm = 0, cnt = 0;
#pragma unroll
    for (int k = 0; k < SN_K; ++k) jj[k] = 0;
    if (tab < 0) { jj[0] = -tab - 1; n4 = 1; }
    else if (tab > 0) {
        const int start = tab >> 6, nn = tab & 63;
        n4 = min(nn, SN_K);
#pragma unroll
        for (int k = 0; k < SN_K; ++k)
            if (k < n4) { int e = __ldg(entj + start + k); jj[k] = (e < 0) ? ~e : e; }
        if (nn > SN_K) tail = start + SN_K;
    }
#pragma unroll
    for (int k = 0; k < SN_K; ++k)
        if (k < n4) {
            double dx[3];
            if (snap_member<DIM>(P, ldg_rec(A.hrec + jj[k]), L, dx)) { hm |= 1 << k; ++cnt; }
        }
    for (int pos = tail; pos >= 0;) {
        int e = __ldg(entj + pos);
        const bool last = e < 0;
        double dx[3];
        cnt += snap_member<DIM>(P, ldg_rec(A.hrec + (last ? ~e : e)), L, dx) ? 1 : 0;
        pos = last ? -1 : pos + 1;
    }
    // warp exclusive scan of the hit counts -> queue slots
    int inc = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int v = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += v;
    }
    const int base = inc - cnt, nq = __shfl_sync(0xffffffffu, inc, 31);
    double tot[3] = {0.0, 0.0, 0.0};
    for (int r0 = 0; r0 < nq; r0 += SN_WQ) {  // one round unless > SN_WQ pairs are waiting
        if (cnt > 0 && base < r0 + SN_WQ && base + cnt > r0) {
            int k = base;
#pragma unroll
            for (int kk = 0; kk < SN_K; ++kk)
                if (hm >> kk & 1) {
                    if (k >= r0 && k < r0 + SN_WQ) {
                        double dx[3];
                        snap_member<DIM>(P, ldg_rec(A.hrec + jj[kk]), L, dx);
                        snap_q &e = wq[k - r0];
                        e.v[0] = dx[0]; e.v[1] = dx[1]; e.v[2] = dx[2];
                        e.j = jj[kk]; e.src = lane;
                    }
                    ++k;
                }
            for (int pos = tail; pos >= 0;) {
                int ej = __ldg(entj + pos);
                const bool last = ej < 0;
                const int j = last ? ~ej : ej;
                double dx[3];
                if (snap_member<DIM>(P, ldg_rec(A.hrec + j), L, dx)) {
                    if (k >= r0 && k < r0 + SN_WQ) {
                        snap_q &e = wq[k - r0];
                        e.v[0] = dx[0]; e.v[1] = dx[1]; e.v[2] = dx[2];
                        e.j = j; e.src = lane;
                    }
                    ++k;
                }
                pos = last ? -1 : pos + 1;
            }
        }
        __syncwarp();
        const int nround = min(SN_WQ, nq - r0);
        for (int q = lane; q < nround; q += 32) {  // dense evaluation: lane q evaluates pair q
            snap_q &e = wq[q];
            const int j = e.j;
            if (A.pair_list) {  // the KD-tree's answer, for the parity tests: which particle met which halo
                const u64 slot = atomicAdd(A.pair_cursor, 1ull);
                if ((i64)slot < A.pair_cap) A.pair_list[slot] = ((u64)mp(act[e.src].idx) << 32) | (u64)j;
            }
            double dx[3] = {e.v[0], e.v[1], e.v[2]};
            double dsq = dx[0] * dx[0] + dx[1] * dx[1];
            if (DIM == 3) dsq += dx[2] * dx[2];
            double d = sqrt(dsq);
            const double2 hp = __ldg(A.hpar + j);  // (cut, lnr_shift)
            double v = bf_row_lerp(A.ax, A.H.rows + (i64)j * A.ax.nr, log(d) - hp.y);
            v = (d < hp.x) ? v : 0.0;  // BaryonCorrection.py:381-382
            double off = v * A.a;      // :228 / :245  (x a_j, reproduced as is)
            if (!bf_finite(off)) off = 0.0;
#pragma unroll
            for (int m = 0; m < 3; ++m) e.v[m] = (m < DIM) ? off * (dx[m] / d) : 0.0;  // :247 (0/0 -> NaN kept)
        }
        __syncwarp();
        const int k0 = max(base, r0), k1 = min(base + cnt, r0 + SN_WQ);
        for (int k = k0; k < k1; ++k) {  // this particle's pairs, in catalogue order
            tot[0] += wq[k - r0].v[0];
            tot[1] += wq[k - r0].v[1];
            if (DIM == 3) tot[2] += wq[k - r0].v[2];
        }
        __syncwarp();
    }
    if (on)
        snap_store<DIM, REC>(A, mp(idx), M, wrap_box(P[0] + tot[0], L), wrap_box(P[1] + tot[1], L),
                             (DIM == 3) ? wrap_box(P[2] + tot[2], L) : P[2]);
    return nq;
}

template <int DIM, bool REC>
__global__ void __launch_bounds__(SN_THREADS, SN_MINBLOCKS) snap_gather_kernel(snap_args A) {
    __shared__ snap_smem sm;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const i64 chunk0 = ((i64)blockIdx.x * (SN_THREADS / 32) + wid) * SN_CHUNK;  // this WARP's particles
    if (chunk0 >= A.np) return;
    snap_act *act = sm.act[wid];
    snap_q *wq = sm.q[wid];
    const int *__restrict__ entj = A.hdr->entj;
    int nwait = 0, npairs = 0;
    for (int t0 = 0; t0 < SN_TILES; t0 += 2) {  // two tiles per pass: independent loads in flight
        double P[2][3], M[2];
        int tab[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const i64 p = chunk0 + (t0 + u) * 32 + lane;
            P[u][0] = P[u][1] = P[u][2] = 0.0; M[u] = 0.0;
            if (p < A.np) {
                if (REC) {
                    const double2 *r = reinterpret_cast<const double2 *>(A.rec + p);
                    double2 r0 = __ldcs(r), r1 = __ldcs(r + 1);  // streamed once: evict-first, L2 is for the tables
                    M[u] = r0.x; P[u][0] = r0.y; P[u][1] = r1.x; P[u][2] = r1.y;
                } else {
                    P[u][0] = __ldcs(A.x + p); P[u][1] = __ldcs(A.y + p);
                    if (DIM == 3) P[u][2] = __ldcs(A.z + p);
                }
            }
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const i64 p = chunk0 + (t0 + u) * 32 + lane;
            tab[u] = 0;
            if (p < A.np) {
                int c = snap_cell_coord(P[u][0], A.G) * A.G.ng + snap_cell_coord(P[u][1], A.G);
                if (DIM == 3) c = c * A.G.ng + snap_cell_coord(P[u][2], A.G);
                tab[u] = __ldg(A.cell_tab + c);
            }
        }
        // cells reached by exactly one halo (most non-empty cells): decide membership here, with all lanes' loads in
        // flight together, so that only particles that do move (or sit in a multi-halo cell) are parked
#pragma unroll
        for (int u = 0; u < 2; ++u)
            if (tab[u] < 0) {
                double dx[3];
                if (!snap_member<DIM>(P[u], ldg_rec(A.hrec + (-tab[u] - 1)), A.G.L, dx)) tab[u] = 0;
            }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const i64 p = chunk0 + (t0 + u) * 32 + lane;
            const bool park = tab[u] != 0;
            if (p < A.np && !park)  // no halo reaches this cell: the particle does not move (single wrap still applies)
                snap_store<DIM, REC>(A, p, M[u], wrap_box(P[u][0], A.G.L), wrap_box(P[u][1], A.G.L),
                                     (DIM == 3) ? wrap_box(P[u][2], A.G.L) : P[u][2]);
            const unsigned bal = __ballot_sync(0xffffffffu, park);
            if (park) {
                snap_act &e = act[nwait + __popc(bal & ((1u << lane) - 1u))];
                e.M = M[u]; e.x = P[u][0]; e.y = P[u][1]; e.z = P[u][2];
                e.tab = tab[u]; e.idx = (t0 + u) * 32 + lane;
            }
            nwait += __popc(bal);
        }
        __syncwarp();
        // drain whenever a full warp's worth is waiting (every lane busy), and whatever is left at the very end
        const bool flush = t0 + 2 >= SN_TILES;
        while (nwait >= 32 || (flush && nwait > 0)) {
            const int take = min(nwait, 32);
            npairs += snap_drain<DIM, REC>(A, entj, act, wq, take, snap_map_chunk{chunk0}, lane);
            __syncwarp();
            snap_act keep[2];
            const int left = nwait - take;
#pragma unroll
            for (int h = 0; h < 2; ++h)
                if (lane + 32 * h < left) keep[h] = act[take + lane + 32 * h];
            __syncwarp();
#pragma unroll
            for (int h = 0; h < 2; ++h)
                if (lane + 32 * h < left) act[lane + 32 * h] = keep[h];
            nwait = left;
            __syncwarp();
        }
    }
    if (lane == 0 && A.npairs && npairs) atomicAdd(A.npairs, (u64)npairs);
}

// ---------------------------------------------------------------------------------------------------
// streaming particle kernel for (M, x, y, z) records: the same per-warp pipeline (stream / park / drain), fed by TMA.
//
// The chunked kernel above is latency bound (ncu: long scoreboard 10 cycles per issue, 35 % of the warp slots
// active, 3.2 TB/s of DRAM traffic): a warp's record loads, the dependent cell-table word and the dependent halo
// record are three round trips to L2 / HBM in a row, and nothing is in flight for the warp while it drains its
// parked particles.  Here every warp owns a ring of SS_D shared-memory tiles of 32 records (1 KB each) that lane 0
// keeps full with bulk asynchronous copies (cp.async.bulk global -> shared, completion counted in bytes on one
// mbarrier per slot, L2 evict-first hint): SS_D KB per warp are in flight at all times, also while the warp is
// draining, and they hold no registers.  The dependent loads are software-pipelined over the ring:
//     A(n + 2)  wait for tile n + 2, read its coordinates from shared memory, issue the cell-table load
//     B(n + 1)  cell-table word of tile n + 1 has arrived: issue the load of the one halo record of single-halo cells
//     C(n)      membership, write-through of the particles no halo reaches (one 256-bit store per record), parking
//               of the others, re-issue of the slot with tile n + SS_D, drains
// so a load is consumed one loop iteration after it was issued.  Warps take chunks of SS_CHUNK_TILES tiles round
// robin over the whole grid (persistent CTAs, one wave), which keeps the DRAM front compact and balances the
// halo-rich regions over all warps.  Results are the chunked kernel's bit for bit: the same membership test, the
// same drain, the same summation order.
// ---------------------------------------------------------------------------------------------------
#define SS_WARPS 4
#define SS_THREADS (32 * SS_WARPS)
#ifndef SS_D
#define SS_D 4
#endif
#ifndef SS_MINBLOCKS
#define SS_MINBLOCKS 4
#endif
#define SS_AQ 64  // parked particles per warp: <= 31 left over + one tile

struct __align__(128) ss_smem {
    double4 ring[SS_WARPS][SS_D][32];
    snap_act act[SS_WARPS][SS_AQ];
    snap_q q[SS_WARPS][SN_WQ];
    u64 bar[SS_WARPS][SS_D];
};

__device__ __forceinline__ void ss_tma_load(void *dst_smem, const void *src_gmem, unsigned bytes, u64 *bar, u64 policy) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bf_smem_u32(bar)), "r"(bytes) : "memory");
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
            bf_smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(bf_smem_u32(bar)), "l"(policy)
        : "memory");
}

template <int DIM>
__global__ void __launch_bounds__(SS_THREADS, SS_MINBLOCKS) snap_stream_kernel(snap_args A) {
    __shared__ ss_smem sm;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const snap_map_stream mp{(i64)blockIdx.x * SS_WARPS + wid, (i64)gridDim.x * SS_WARPS};
    const i64 ntiles = (A.np + 31) >> 5;
    if (mp.tile(0) >= ntiles) return;
    double4(*ring)[32] = sm.ring[wid];
    u64 *bar = sm.bar[wid];
    snap_act *act = sm.act[wid];
    snap_q *wq = sm.q[wid];
    const int *__restrict__ entj = A.hdr->entj;
    const double L = A.G.L;
    u64 policy;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
    if (lane == 0)
        for (int s = 0; s < SS_D; ++s) bf_mbar_init(&bar[s], 1);
    __syncwarp();

    // tile T of the array is tile n of this warp's sequence; the tile `ahead` places further on is T + ahead inside the
    // chunk and one round of chunks later beyond it (ahead <= SS_CHUNK_TILES)
    const i64 round_skip = (mp.W - 1) * SS_CHUNK_TILES;
    auto tile_ahead = [&](i64 T, int r, int ahead) -> i64 { return T + ahead + ((r + ahead >= SS_CHUNK_TILES) ? round_skip : 0); };
    auto issue = [&](i64 T, int slot) {  // lane 0: bulk copy of tile T into a ring slot
        if (lane == 0 && T < ntiles) {
            const i64 left = A.np - T * 32;
            ss_tma_load(ring[slot], A.rec + T * 32, (unsigned)(left < 32 ? left : 32) * 32u, &bar[slot], policy);
        }
    };
    auto stage_a = [&](i64 T, int slot, unsigned phase) -> int {  // cell-table word of this lane's particle of tile T
        if (T >= ntiles) return 0;
        bf_mbar_wait(&bar[slot], phase);
        if (T * 32 + lane >= A.np) return 0;
        const double2 *r = reinterpret_cast<const double2 *>(&ring[slot][lane]);
        const double2 r0 = r[0], r1 = r[1];
        int c = snap_cell_coord(r0.y, A.G) * A.G.ng + snap_cell_coord(r1.x, A.G);
        if (DIM == 3) c = c * A.G.ng + snap_cell_coord(r1.y, A.G);
        return __ldg(A.cell_tab + c);  // (ld.global.nc.L1::no_allocate here measured SLOWER: 4.19 vs 3.27 ms, the words then miss L2)
    };

    i64 T = mp.tile(0);  // tile of stage C
    int r = 0;           // its place in the chunk
    static_assert(SS_D >= 3 && SS_D <= SS_CHUNK_TILES, "ring depth");
    for (int n = 0; n < SS_D; ++n) issue(tile_ahead(T, 0, n), n);
    int tab_a = stage_a(T, 0, 0u), tab_b = tab_a;
    double4 h_b = make_double4(0.0, 0.0, 0.0, 0.0);
    if (tab_b < 0) h_b = ldg_rec(A.hrec + (-tab_b - 1));
    tab_a = stage_a(tile_ahead(T, 0, 1), 1, 0u);
    int slot = 0, slot_a = 2;        // ring slots of stage C and of stage A (two tiles ahead)
    unsigned phase_a = 0u;           // parity of stage A's slot: flips every time slot_a wraps
    int nwait = 0, npairs = 0;
    for (int n = 0; T < ntiles; ++n) {
        int tab = tab_b;
        const double4 h = h_b;
        tab_b = tab_a;  // B(n + 1)
        if (tab_b < 0) h_b = ldg_rec(A.hrec + (-tab_b - 1));
        tab_a = stage_a(tile_ahead(T, r, 2), slot_a, phase_a);  // A(n + 2)
        // C(n)
        const i64 p = T * 32 + lane;
        const bool valid = p < A.np;
        double P[3] = {0.0, 0.0, 0.0}, M = 0.0;
        if (valid) {
            const double2 *rr = reinterpret_cast<const double2 *>(&ring[slot][lane]);
            const double2 r0 = rr[0], r1 = rr[1];
            M = r0.x; P[0] = r0.y; P[1] = r1.x; P[2] = r1.y;
        }
        if (tab < 0) {  // single-halo cell: decided here, only particles that do move are parked
            double dx[3];
            if (!snap_member<DIM>(P, h, L, dx)) tab = 0;
        }
        const bool park = tab != 0;
        if (valid && !park) {
            // the single wrap (:259-260) only ever acts on a coordinate outside [0, L]: one test for the three of them
            const double lo = fmin(fmin(P[0], P[1]), (DIM == 3) ? P[2] : P[0]), hi = fmax(fmax(P[0], P[1]), (DIM == 3) ? P[2] : P[0]);
            if (lo < 0.0 || hi > L) {
                P[0] = wrap_box(P[0], L); P[1] = wrap_box(P[1], L);
                if (DIM == 3) P[2] = wrap_box(P[2], L);
            }
            snap_store<DIM, true>(A, p, M, P[0], P[1], P[2]);
        }
        const unsigned bal = __ballot_sync(0xffffffffu, park);
        if (park) {
            snap_act &e = act[nwait + __popc(bal & ((1u << lane) - 1u))];
            e.M = M; e.x = P[0]; e.y = P[1]; e.z = P[2];
            e.tab = tab; e.idx = n * 32 + lane;
        }
        nwait += __popc(bal);
        __syncwarp();                              // every lane has consumed its record of the slot (stored or parked)
        issue(tile_ahead(T, r, SS_D), slot);       // ... so the slot can take the tile SS_D ahead
        // advance the cursors
        T = tile_ahead(T, r, 1);
        r = (r + 1 == SS_CHUNK_TILES) ? 0 : r + 1;
        slot = (slot + 1 == SS_D) ? 0 : slot + 1;
        if (++slot_a == SS_D) { slot_a = 0; phase_a ^= 1u; }
        const bool flush = T >= ntiles;
        while (nwait >= 32 || (flush && nwait > 0)) {
            const int take = min(nwait, 32);
            npairs += snap_drain<DIM, true>(A, entj, act, wq, take, mp, lane);
            __syncwarp();
            const int left = nwait - take;  // <= 31
            snap_act keep;
            if (lane < left) keep = act[take + lane];
            __syncwarp();
            if (lane < left) act[lane] = keep;
            nwait = left;
            __syncwarp();
        }
    }
    if (lane == 0 && A.npairs && npairs) atomicAdd(A.npairs, (u64)npairs);
}

static int ss_resident_blocks(int dim) {  // CTAs of the streaming kernel one SM holds (queried once per process)
    static int nb[2] = {0, 0};
    int &v = nb[dim - 2];
    if (v == 0) {
        int n = 0;
        cudaError_t e = (dim == 2) ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, snap_stream_kernel<2>, SS_THREADS, 0)
                                   : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, snap_stream_kernel<3>, SS_THREADS, 0);
        v = (e == cudaSuccess && n > 0) ? n : SS_MINBLOCKS;
    }
    return v;
}

// ---------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------
extern "C" int64_t bf_snap_grid_cells(int dim, double L, double Rq_mean) {
    if ((dim != 2 && dim != 3) || !(L > 0)) return 0;
    i64 cap = (dim == 3) ? SN_NG_MAX_3D : SN_NG_MAX_2D;
    if (!(Rq_mean > 0)) return cap;
    double n = floor(L / Rq_mean);
    if (!(n >= 1)) n = 1;
    if (n > (double)cap) n = (double)cap;
    return (i64)n;
}

extern "C" size_t bf_snap_ws_bytes(int dim, int64_t nh, double L, int64_t ng, int64_t max_entries) {
    if ((dim != 2 && dim != 3) || nh < 0 || !(L > 0) || ng < 1 || max_entries < 0) return 0;
    if (ng > ((dim == 3) ? SN_NG_MAX_3D : SN_NG_MAX_2D)) return 0;
    return snap_layout(nullptr, make_grid(dim, L, ng), nh, max_entries).total;
}

extern "C" int bf_snap_bin_halos(int dim, double L, int64_t ng, const bf_snap_halos *H, int64_t max_entries, void *ws,
                                 size_t ws_bytes, int64_t *nentries_host, void *stream) {
    BF_REQUIRE(dim == 2 || dim == 3, "dim must be 2 or 3");
    BF_REQUIRE(L > 0 && ng >= 1 && ng <= ((dim == 3) ? SN_NG_MAX_3D : SN_NG_MAX_2D), "L / ng");
    BF_REQUIRE(H && H->nh >= 0 && H->nh < (1ll << 31), "halo record");
    BF_REQUIRE(H->nh == 0 || (H->pos && H->Rq), "null per-halo array");
    BF_REQUIRE(ws && nentries_host, "null outputs");
    BF_REQUIRE(max_entries >= 0, "max_entries");
    if (max_entries >= (1ll << 25)) max_entries = (1ll << 25) - 1;  // (more can never be used: BF_E_CAPACITY below)
    snap_grid G = make_grid(dim, L, ng);
    snap_ws w = snap_layout(ws, G, H->nh, max_entries);
    if (ws_bytes < w.total) {
        bf_set_error("bf_snap_bin_halos: workspace %zu < %zu bytes", ws_bytes, w.total);
        return BF_E_WORKSPACE;
    }
    cudaStream_t s = (cudaStream_t)stream;
    BF_CUDA_OK(cudaMemsetAsync(w.cell_tab, 0, sizeof(int) * (size_t)(G.ncell + 1), s));
    snap_hdr_kernel<<<1, 1, 0, s>>>(w.hdr, w.entj);
    BF_LAUNCH_OK();
    BF_CUDA_OK(cudaMemsetAsync(w.counts, 0, sizeof(i64) * (size_t)(H->nh + 1), s));
    *nentries_host = 0;
    if (H->nh == 0) return BF_OK;
    unsigned blocks = (unsigned)((H->nh * 32 + 127) / 128);
    if (dim == 2) snap_bin_kernel<2, false><<<blocks, 128, 0, s>>>(G, *H, w.counts, nullptr, nullptr, w.hrec, w.hpar);
    else snap_bin_kernel<3, false><<<blocks, 128, 0, s>>>(G, *H, w.counts, nullptr, nullptr, w.hrec, w.hpar);
    BF_LAUNCH_OK();
    size_t cb = w.cub_bytes;
    BF_CUDA_OK(cub::DeviceScan::ExclusiveSum(w.cub_tmp, cb, w.counts, w.offsets, (int)(H->nh + 1), s));
    i64 total = 0;
    BF_CUDA_OK(cudaMemcpyAsync(&total, w.offsets + H->nh, sizeof(i64), cudaMemcpyDeviceToHost, s));
    BF_CUDA_OK(cudaStreamSynchronize(s));
    *nentries_host = total;
    if (total >= (1ll << 25)) {
        bf_set_error("bf_snap_bin_halos: %lld (cell, halo) entries do not fit the cell table; use a smaller ng", (long long)total);
        return BF_E_CAPACITY;
    }
    if (total > max_entries) {
        bf_set_error("bf_snap_bin_halos: %lld entries exceed max_entries %lld", (long long)total, (long long)max_entries);
        return BF_E_WORKSPACE;
    }
    if (total == 0) return BF_OK;
    if (dim == 2) snap_bin_kernel<2, true><<<blocks, 128, 0, s>>>(G, *H, w.counts, w.offsets, w.keys, w.hrec, w.hpar);
    else snap_bin_kernel<3, true><<<blocks, 128, 0, s>>>(G, *H, w.counts, w.offsets, w.keys, w.hrec, w.hpar);
    BF_LAUNCH_OK();
    int cell_bits = 1;
    while (((i64)1 << cell_bits) < G.ncell) ++cell_bits;
    cub::DoubleBuffer<u64> db(w.keys, w.keys_alt);
    cb = w.cub_bytes;
    // stable sort on the cell bits only: keys were emitted halo by halo, so every cell's list stays in catalogue order
    BF_CUDA_OK(cub::DeviceRadixSort::SortKeys(w.cub_tmp, cb, db, (int)total, 32, 32 + cell_bits, s));
    snap_entries_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(total, db.Current(), w.cell_tab, w.entj);
    BF_LAUNCH_OK();
    return BF_OK;
}

static int snap_launch(int dim, int64_t np, snap_args &A, double L, int64_t ng, const bf_snap_halos *H, const bf_table *T,
                       const void *ws, bool rec, void *stream) {
    BF_REQUIRE(dim == 2 || dim == 3, "dim must be 2 or 3");
    BF_REQUIRE(np >= 0 && L > 0, "np / L");
    BF_REQUIRE(ng >= 1 && ng <= ((dim == 3) ? SN_NG_MAX_3D : SN_NG_MAX_2D), "ng");
    BF_REQUIRE(bf_check_table(T), "bad table");
    BF_REQUIRE(H && H->nh >= 0 && H->nh < (1ll << 31), "halo record");
    BF_REQUIRE(H->nh == 0 || (H->pos && H->Rq && H->cut && H->lnr_shift && H->rows), "null per-halo array");
    BF_REQUIRE(ws, "workspace (from bf_snap_bin_halos) is NULL");
    if (np == 0) return BF_OK;
    A.G = make_grid(dim, L, ng);
    snap_ws w = snap_layout(const_cast<void *>(ws), A.G, H->nh, 0);  // the fixed part does not depend on max_entries
    A.np = np;
    A.H = *H;
    A.ax = bf_make_raxis(T);
    A.cell_tab = w.cell_tab;
    A.hrec = w.hrec;
    A.hpar = w.hpar;
    A.hdr = w.hdr;
    unsigned pb = (unsigned)((np + SN_TILE - 1) / SN_TILE);
    cudaStream_t s = (cudaStream_t)stream;
    A.wide = rec && ((((uintptr_t)A.rec | (uintptr_t)A.rec_out) & 31) == 0);
    static const bool chunked = getenv("BF_SNAP_CHUNKED") != nullptr;  // tuning / cross-check: the round-1 kernel
    if (rec && !chunked && !A.pair_list) {
        const i64 nchunks = (np + 32 * SS_CHUNK_TILES - 1) / (32 * SS_CHUNK_TILES);
        i64 blocks = (nchunks + SS_WARPS - 1) / SS_WARPS;
        const i64 wave = (i64)bf_sm_count() * ss_resident_blocks(dim);  // persistent CTAs: one wave
        if (blocks > wave) blocks = wave;
        if (dim == 2) snap_stream_kernel<2><<<(unsigned)blocks, SS_THREADS, 0, s>>>(A);
        else snap_stream_kernel<3><<<(unsigned)blocks, SS_THREADS, 0, s>>>(A);
        BF_LAUNCH_OK();
        return BF_OK;
    }
    if (dim == 2) {
        if (rec) snap_gather_kernel<2, true><<<pb, SN_THREADS, 0, s>>>(A);
        else snap_gather_kernel<2, false><<<pb, SN_THREADS, 0, s>>>(A);
    } else {
        if (rec) snap_gather_kernel<3, true><<<pb, SN_THREADS, 0, s>>>(A);
        else snap_gather_kernel<3, false><<<pb, SN_THREADS, 0, s>>>(A);
    }
    BF_LAUNCH_OK();
    return BF_OK;
}

extern "C" int bf_snap_displace(int dim, int64_t np, const double *x, const double *y, const double *z, double L,
                                double a, const bf_snap_halos *H, const bf_table *T, int64_t ng, const void *ws,
                                double *xo, double *yo, double *zo, unsigned long long *npairs, void *stream) {
    BF_REQUIRE(np == 0 || (x && y && xo && yo && (dim == 2 || (z && zo))), "null particle array");
    snap_args A;
    A.x = x; A.y = y; A.z = z; A.xo = xo; A.yo = yo; A.zo = zo;
    A.rec = nullptr; A.rec_out = nullptr;
    A.a = a;
    A.npairs = npairs;
    A.pair_list = nullptr; A.pair_cursor = nullptr; A.pair_cap = 0;
    return snap_launch(dim, np, A, L, ng, H, T, ws, false, stream);
}

extern "C" int bf_snap_pair_list(int dim, int64_t np, const double *x, const double *y, const double *z, double L,
                                 double a, const bf_snap_halos *H, const bf_table *T, int64_t ng, const void *ws,
                                 double *xo, double *yo, double *zo, unsigned long long *pairs, int64_t pair_cap,
                                 unsigned long long *pair_count, void *stream) {
    BF_REQUIRE(np == 0 || (x && y && xo && yo && (dim == 2 || (z && zo))), "null particle array");
    BF_REQUIRE(pairs && pair_count && pair_cap >= 0 && np < (1ll << 32), "pair list");
    snap_args A;
    A.x = x; A.y = y; A.z = z; A.xo = xo; A.yo = yo; A.zo = zo;
    A.rec = nullptr; A.rec_out = nullptr;
    A.a = a;
    A.npairs = nullptr;
    A.pair_list = pairs; A.pair_cursor = pair_count; A.pair_cap = pair_cap;
    BF_CUDA_OK(cudaMemsetAsync(pair_count, 0, sizeof(unsigned long long), (cudaStream_t)stream));
    return snap_launch(dim, np, A, L, ng, H, T, ws, false, stream);
}

extern "C" int bf_snap_displace_records(int dim, int64_t np, const void *cat, double L, double a,
                                        const bf_snap_halos *H, const bf_table *T, int64_t ng, const void *ws,
                                        void *cat_out, unsigned long long *npairs, void *stream) {
    BF_REQUIRE(np == 0 || (cat && cat_out), "null particle records");
    BF_REQUIRE((((uintptr_t)cat | (uintptr_t)cat_out) & 15) == 0, "particle records must be 16-byte aligned");
    snap_args A;
    A.x = A.y = A.z = nullptr; A.xo = A.yo = A.zo = nullptr;
    A.rec = (const double4 *)cat; A.rec_out = (double4 *)cat_out;
    A.a = a;
    A.npairs = npairs;
    A.pair_list = nullptr; A.pair_cursor = nullptr; A.pair_cap = 0;
    return snap_launch(dim, np, A, L, ng, H, T, ws, true, stream);
}
