// bf_shell_gather.cu -- pixel-owner ("gather") evaluation of the Shell runners: the deterministic hot path.
//
// Replaces the per-halo bodies of PaintProfilesShell.process (Runners/HealpixRunner.py:418-445) and
// BaryonifyShell.process (:291-331).  One CTA owns one sky cell (8 rings x Wc pixels, bf_shell_cells.cuh),
// one warp owns one ring of it, one lane owns one pixel of the current 32-pixel strip.  The cell's halo list
// (catalogue order, from bf_shell_bin.cu) is staged in shared memory in chunks; per chunk
//   1. thread c sets up candidate c's disc constants once (query_disc prologue);
//   2. warp w turns them into ring-relative pixel ranges for ITS ring, 32 candidates at a time -- this is
//      hp.query_disc's ring walk, bit for bit, so the (halo, pixel) pair set equals the reference's;
//   3. every lane walks the candidate list with its own cursor, skipping (integer compares) to the next
//      halo whose range holds its pixel, and evaluates that pair; lanes of a warp work on DIFFERENT halos at
//      the same time, so the fp64 pipe stays full even when discs are narrower than the strip.
// Each pixel's contributions are summed in registers in a fixed order (per candidate chunk: the whole-strip candidates
// in catalogue order, then the partly covering ones in catalogue order) and written once: no atomics, no
// read-modify-write traffic, bit-reproducible.
//
// Pair arithmetic (fp64).  With delta = v - c (unit vectors), |delta|^2 = d2:
//   r_tab = |delta| D / a  ->  ln r_tab = 0.5 ln d2 + ln(D/a)          (no sqrt for the table coordinate)
//   paint:      P = exp(lerp(row, ln r_tab))
//   baryonify:  the reference moves pos = D v by (d a) along delta/|delta| and renormalises.  With
//               e = d (a/D) / |delta| the new unit vector is (v + e delta) / sqrt(1 + q), q = e d2 (1 + e)
//               (v.delta = d2/2 for unit vectors), so  shift = v g + e (1 + g) delta,  g = (1+q)^(-1/2) - 1
//               evaluated by its series (|q| < 1e-3): algebraically identical to HealpixRunner.py:321-328 and
//               free of the reference's 1e-16 cancellation noise.
#include "bf_shell_cells.cuh"

#define GT_WARPS BF_BAND_RINGS
#define GT_THREADS (32 * GT_WARPS)
#ifndef GT_CMAX
#define GT_CMAX 128
#endif
#ifndef GT_MINBLOCKS
#define GT_MINBLOCKS 4
#endif
#ifndef GT_FAST_NP
#define GT_FAST_NP 2  // pairs evaluated side by side in the whole-strip loop: the chains' latencies overlap (54.3 vs 57.1 ms)
#endif

struct gather_args {
    const void *aux;     // [ln table | (g, 1/dg) pairs | g] prepared once per call (gather_aux_kernel)
    unsigned aux_bytes;  // multiple of 16
    bf_hpx h;
    bf_cells cg;
    bf_shell_halos S;
    bf_raxis ax;
    const u64 *keys;
    const int *cell_begin, *cell_end;
    i64 band_lo;
    double *out;
    i64 pix_base;
    int accumulate;
    u64 *npairs;
};

// Pair constants of one staged candidate: (c0, c1 | c2, ln(D/a) | cut^2, a/D | row pointer, -), read with four 16-byte
// shared-memory loads.  The stride of 10 doubles (80 B = 20 banks) spreads the records of 8 different candidates over all
// 32 banks, so the per-lane reads of the partial path rarely conflict; the whole-strip path reads ONE record per warp
// (a broadcast).
#define GT_CSTRIDE 10
struct gather_smem {
    double cand[GT_CMAX][GT_CSTRIDE];
    double phi[GT_CMAX], z0[GT_CMAX], xa[GT_CMAX], cosr[GT_CMAX];
    int irmin[GT_CMAX], irmax[GT_CMAX], wn[GT_CMAX], ws[GT_CMAX];
    int2 rng[GT_WARPS][GT_CMAX];
    unsigned hits[GT_WARPS][GT_CMAX / 32][32];  // [warp][candidate word][lane]: bit c = candidate 32*word+c hits my pixel
    unsigned full[GT_WARPS][GT_CMAX / 32];      // [warp][candidate word]: bit c = candidate 32*word+c covers the WHOLE strip
    u64 mbar, mbar_pad;                         // completion barrier of the TMA bulk copy of the constant tables
};

// bits a..e (inclusive) of a 32-bit word, clipped to the word
__device__ __forceinline__ unsigned bits_between(int a, int e) {
    if (e < 0 || a > 31 || e < a) return 0u;
    a = max(a, 0);
    e = min(e, 31);
    return (0xffffffffu >> (31 - e)) & (0xffffffffu << a);
}

// lanes (pixels b .. b+31 of the ring) covered by a ring-relative range; r.x > r.y encodes the wrapped
// pair [0, r.y] U [r.x, nr) and (INT_MAX, -1) the empty set
__device__ __forceinline__ unsigned lanes_of_range(int2 r, int b) {
    if (r.x <= r.y) return bits_between(r.x - b, r.y - b);
    return bits_between(-b, r.y - b) | bits_between(r.x - b, 31);
}

// 32x32 bit-matrix transpose across a warp: lane c gives row c, lane l receives column l
__device__ __forceinline__ unsigned warp_transpose32(unsigned x, int lane) {
    unsigned m = 0x0000ffffu;
#pragma unroll
    for (int j = 16; j != 0; j >>= 1) {
        unsigned y = __shfl_xor_sync(0xffffffffu, x, j);
        if ((lane & j) == 0) x ^= (((x >> j) ^ y) & m) << j;
        else x ^= ((y >> j) ^ x) & m;
        m ^= m << (j >> 1);
    }
    return x;
}

// ln(x) as bf_fastlog with the log1p series cut at degree 5: |r| <= 2^-8, so the truncation r^6/6 < 6e-16 stays at the
// rounding level of the table coordinate 0.5 ln d2 + ln(D/a) (whose magnitude is 1..10).
__device__ __forceinline__ double gather_log(const double2 *__restrict__ tab, double x) {
    const int hi = __double2hiint(x);
    const int e = (hi >> 20) - 1023;
    const double m = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(x));
    const double2 t = tab[(hi >> 13) & (BF_LOGTAB - 1)];
    const double r = fma(m, t.x, -1.0);
    double p = fma(r, bf_kc[1], -0.25);
    p = fma(r, p, bf_kc[2]);
    p = fma(r, p, -0.5);
    p = fma(r, p, 1.0);
    return fma((double)e, bf_kc[3], fma(r, p, t.y));
}

// One (pixel, candidate) pair, added to the pixel's running sums.  paint: acc[0] += P.  baryonify: the unit-vector
// shift of the pair is v g + e (1 + g) delta; v is the pixel's own vector, so acc[3] collects sum g and acc[0..2] collect
// sum e (1 + g) delta (one FMA per component) -- the pixel's shift is acc[0..2] + v acc[3], formed once at the end.
// fp64 instructions are what bounds this kernel, so tests that do not need fp64 arithmetic (finite?, |q| > 1e-3,
// 0 <= t < 1 - 2^-50) look at the bit patterns with integer instructions instead.
// NP pairs of the same pixel at once (NP = 2 in the whole-strip loop): the pairs' dependent fp64 chains (ln series,
// reciprocal square root, (1+q)^-1/2 series) are independent of each other, so writing them side by side lets the
// scheduler fill one chain's latency with the other's instructions; the rare exact-search and |q| > 1e-3 routes are
// taken once for the group.
template <int MODE, int NP>
__device__ __forceinline__ void gather_pairs(const gather_smem &sm, const int c[NP], const double v[3], const bf_raxis &ax,
                                             const double2 *__restrict__ gi_s, const double2 *__restrict__ logtab,
                                             double acc[4]) {
    const double2 *__restrict__ cq[NP];
    double dx[NP], dy[NP], dz[NP], d2[NP], x[NP], tt[NP];
    int k[NP];
    bool inr[NP];
    bool fast = true;
#pragma unroll
    for (int u = 0; u < NP; ++u) {
        cq[u] = reinterpret_cast<const double2 *>(sm.cand[c[u]]);
        const double2 q0 = cq[u][0], q1 = cq[u][1];  // (c0, c1), (c2, ln(D/a))
        dx[u] = v[0] - q0.x; dy[u] = v[1] - q0.y; dz[u] = v[2] - q1.x;
        d2[u] = fma(dz[u], dz[u], fma(dy[u], dy[u], dx[u] * dx[u]));
        x[u] = fma(0.5, gather_log(logtab, d2[u]), q1.y);
        // Table cell: uniform-grid guess, accepted when 0 <= tt < 1 - 2^-21, which PROVES g[k] <= x < g[k+1] (tt >= 0 iff
        // x >= g[k]; x >= g[k+1] would give tt >= 1 - 4 ulp) and with it that x lies inside the table.  For tt >= 0 the
        // order of doubles is the order of their bit patterns; negative tt and NaN have larger patterns, so one
        // unsigned compare of the HIGH WORDS decides (the 5e-7 of the pairs with 1 - 2^-21 <= tt < 1 take the exact
        // route needlessly and get the same cell).  Everything else -- rounding at a node, a non-uniform axis,
        // x == glast, x outside the table (fill_value NaN) -- takes the exact route.
        k[u] = (int)((x[u] - ax.g0) * ax.inv_dx);
        k[u] = max(0, min(k[u], ax.nr - 2));
        const double2 gi = gi_s[k[u]];
        tt[u] = (x[u] - gi.x) * gi.y;
        inr[u] = true;
        fast = fast && (unsigned)__double2hiint(tt[u]) < 0x3fefffffu;  // (high word of 1 - 2^-21: one 32-bit compare)
    }
    if (!fast) {
#pragma unroll
        for (int u = 0; u < NP; ++u)
            if ((unsigned)__double2hiint(tt[u]) >= 0x3fefffffu) {
                inr[u] = x[u] >= ax.g0 && x[u] <= ax.glast;
                if (inr[u]) {
                    k[u] = bf_find_cell(ax, x[u]);
                    const double2 gi = gi_s[k[u]];
                    tt[u] = (x[u] - gi.x) * gi.y;
                }
            }
    }
    double r0[NP], r1[NP];
#pragma unroll
    for (int u = 0; u < NP; ++u) {
        const double *row = reinterpret_cast<const double *>(__double_as_longlong(cq[u][3].x)) + k[u];
        r0[u] = __ldg(row);
        r1[u] = __ldg(row + 1);
    }
    if (MODE == 0) {
#pragma unroll
        for (int u = 0; u < NP; ++u) {
            // the two-product form keeps scipy's propagation through zero weights (0 * -inf = NaN -> 0, Tabulate.py:237)
            const double val = fma(r1[u], tt[u], r0[u] * (1.0 - tt[u]));
            const double P = exp(inr[u] ? val : nan(""));
            if ((__double2hiint(P) & 0x7ff00000) != 0x7ff00000) acc[0] += P;  // finite (HealpixRunner.py:442)
        }
    } else {
        double e[NP], q[NP], gg[NP];
        bool big = false;
#pragma unroll
        for (int u = 0; u < NP; ++u) {
            const double2 q2 = cq[u][2];  // (cut^2, a/D)
            const double sc = q2.y * bf_rsqrt(d2[u]);
            const bool use = inr[u] && d2[u] < q2.x;  // NaN fill / BaryonCorrection.py:381-382: both end as 0
            // (an inf or NaN anywhere in the cell gives a non-finite e in either form of the lerp, which ends as 0 below)
            e[u] = fma(tt[u], r1[u] - r0[u], r0[u]) * sc;
            if (!use || (__double2hiint(e[u]) & 0x7ff00000) == 0x7ff00000) e[u] = 0.0;  // HealpixRunner.py:323
            const double ed = e[u] * d2[u];
            q[u] = fma(e[u], ed, ed);  // e d2 (1 + e)
            gg[u] = q[u] * fma(q[u], fma(q[u], fma(q[u], bf_kc[5], bf_kc[6]), bf_kc[4]), -0.5);  // (1+q)^-1/2 - 1
            big = big || (__double2hiint(q[u]) & 0x7fffffff) > 0x3f50624d;
        }
        if (big) {  // |q| > 1e-3: pixel (almost) on a halo centre
#pragma unroll
            for (int u = 0; u < NP; ++u)
                if ((__double2hiint(q[u]) & 0x7fffffff) > 0x3f50624d) gg[u] = rsqrt(1.0 + q[u]) - 1.0;
        }
#pragma unroll
        for (int u = 0; u < NP; ++u) {  // (pair u before pair u + 1: the candidates' order)
            const double ef = fma(e[u], gg[u], e[u]);
            acc[3] += gg[u];
            acc[0] = fma(ef, dx[u], acc[0]);
            acc[1] = fma(ef, dy[u], acc[1]);
            acc[2] = fma(ef, dz[u], acc[2]);
        }
    }
}

template <int MODE>
__device__ __forceinline__ void gather_pair(const gather_smem &sm, int c, const double v[3], const bf_raxis &ax,
                                            const double2 *__restrict__ gi_s, const double2 *__restrict__ logtab,
                                            double acc[4]) {
    const int cc[1] = {c};
    gather_pairs<MODE, 1>(sm, cc, v, ax, gi_s, logtab, acc);
}

template <int MODE>
__global__ void __launch_bounds__(GT_THREADS, GT_MINBLOCKS) gather_kernel(gather_args A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    gather_smem &sm = *reinterpret_cast<gather_smem *>(smem_raw);
    // dynamic shared memory after the struct: ln table [BF_LOGTAB] double2, then the radial axis as (g[k], 1 / (g[k+1] - g[k]))
    // pairs so that one 16-byte load serves the lerp, then the plain axis for the (rare) exact search
    double2 *logtab = reinterpret_cast<double2 *>(smem_raw + sizeof(gather_smem));
    double2 *gi_s = logtab + BF_LOGTAB;
    double *g_s = reinterpret_cast<double *>(gi_s + A.ax.nr);

    const bf_hpx &h = A.h;
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const i64 band = A.band_lo + blockIdx.x / BF_MAX_COLS;
    const int col = blockIdx.x % BF_MAX_COLS;
    if (col >= A.cg.ncol) return;
    const i64 cellid = band * BF_MAX_COLS + col;
    const int beg = A.cell_begin[cellid], end = A.cell_end[cellid];

    // this warp's ring
    const i64 iz = band * BF_BAND_RINGS + 1 + w;
    i64 sp = 0, nr = 0; bool shifted = false;
    const bool ring_ok = iz <= 4 * h.nside - 1;
    if (ring_ok) bfh_ring_info_small(h, iz, sp, nr, shifted);
    const int ip0 = col * A.cg.wc;
    const int nstrip = (ring_ok && ip0 < nr) ? (int)min((i64)A.cg.cw, (nr - ip0 + 31) / 32) : 0;
    const int ncomp = (MODE == 0) ? 1 : 3;

    if (end <= beg) {  // empty cell: the pixels are still owned by this call
        if (!A.accumulate)
            for (int s = 0; s < nstrip; ++s) {
                int ip = ip0 + s * 32 + lane;
                if (ip < nr)
                    for (int m = 0; m < ncomp; ++m) A.out[(sp + ip - A.pix_base) * ncomp + m] = 0.0;
            }
        return;
    }

    // stage the constant tables: ONE asynchronous bulk copy (TMA) issued by thread 0, completion on an mbarrier
    if (t == 0) bf_mbar_init(&sm.mbar, 1);
    __syncthreads();
    if (t == 0) bf_tma_load_1d(logtab, A.aux, A.aux_bytes, &sm.mbar);
    bf_mbar_wait(&sm.mbar, 0);
    bf_raxis ax = A.ax;
    ax.g = g_s;

    unsigned my_pairs32 = 0;  // per lane and cell: far below 2^32
    for (int cb = beg; cb < end; cb += GT_CMAX) {
        const int nc = min(GT_CMAX, end - cb);
        __syncthreads();
        if (t < nc) {  // stage candidate t: pair constants + query_disc prologue
            i64 j = (i64)(A.keys[cb + t] & 0xffffffffull);
            double c[3] = {A.S.vec[3 * j], A.S.vec[3 * j + 1], A.S.vec[3 * j + 2]};
            double D = A.S.D[j], a = A.S.a[j];
            double *cr = sm.cand[t];
            cr[0] = c[0]; cr[1] = c[1]; cr[2] = c[2];
            cr[3] = log(D / a) - ((MODE == 1) ? A.S.lnr_shift[j] : 0.0);
            double ct = (MODE == 1) ? A.S.cut[j] * (a / D) : 0.0;
            cr[4] = ct * ct;
            cr[5] = a / D;
            cr[6] = __longlong_as_double((long long)(A.S.rows + j * (i64)A.ax.nr));
            bf_disc d = bfh_disc_setup(h, c, A.S.radius[j]);
            sm.phi[t] = d.phi; sm.z0[t] = d.z0; sm.xa[t] = d.xa; sm.cosr[t] = d.cosr;
            sm.irmin[t] = (int)d.irmin; sm.irmax[t] = (int)d.irmax;
            sm.wn[t] = (d.north_end > 0) ? (d.all ? (int)(4 * h.nside) : (int)d.irmin - 1) : 0;  // rings <= wn whole
            sm.ws[t] = (d.south_begin < h.npix) ? (int)d.irmax + 1 : 0x7fffffff;              // rings >= ws whole
        }
        __syncthreads();
        if (nstrip > 0) {
            // ring-relative ranges of every candidate for this warp's ring
            for (int c = lane; c < nc; c += 32) {
                int2 r = make_int2(0x7fffffff, -1);
                if ((int)iz <= sm.wn[c] || (int)iz >= sm.ws[c]) r = make_int2(0, (int)nr - 1);
                else if ((int)iz >= sm.irmin[c] && (int)iz <= sm.irmax[c]) {
                    bf_disc d;
                    d.phi = sm.phi[c]; d.z0 = sm.z0[c]; d.xa = sm.xa[c]; d.cosr = sm.cosr[c];
                    i64 a0, a1, b0, b1;
                    bfh_disc_ring(h, d, iz, a0, a1, b0, b1);
                    bool ha = a1 > a0, hb = b1 > b0;
                    if (ha && hb) r = make_int2((int)(b0 - sp), (int)(a1 - sp) - 1);  // wrapped: [0,hi] U [lo,nr)
                    else if (ha) r = make_int2((int)(a0 - sp), (int)(a1 - sp) - 1);
                    else if (hb) r = make_int2((int)(b0 - sp), (int)(b1 - sp) - 1);
                }
                sm.rng[w][c] = r;
            }
            __syncwarp();
            for (int s = 0; s < nstrip; ++s) {
                __syncwarp();
                const int ip = ip0 + s * 32 + lane;
                const bool valid = ip < nr;
                double v[3] = {0, 0, 1};
                if (valid) {
                    double z, phi, sth; bool have;
                    bfh_ringpix2loc(h, iz, (i64)ip + 1, z, phi, sth, have);
                    bfh_loc2vec(z, phi, sth, have, v);
                }
                double acc[4] = {0.0, 0.0, 0.0, 0.0};
                // hit matrix of this strip: lane c builds candidate c's lane mask from its range.  Candidates that cover
                // the WHOLE strip (the common case: discs are wider than 32 pixels) are set aside in a warp-uniform word
                // and evaluated by all 32 lanes together -- no transpose, no vote, no bit scan, and the candidate's
                // constants are one broadcast read.  The remaining masks are transposed, so every lane ends up with one
                // bit per candidate for ITS pixel.
                const int nw = (nc + 31) >> 5;
                const unsigned vmask = __ballot_sync(0xffffffffu, valid);  // lanes that hold a pixel (ring ends: caps)
                for (int kw = 0; kw < nw; ++kw) {
                    const int c = kw * 32 + lane;
                    unsigned m = (c < nc) ? (lanes_of_range(sm.rng[w][c], ip0 + s * 32) & vmask) : 0u;
                    const unsigned fw = __ballot_sync(0xffffffffu, m == 0xffffffffu);
                    if (m == 0xffffffffu) m = 0u;
                    const bool any_part = __any_sync(0xffffffffu, m != 0u);
                    sm.hits[w][kw][lane] = any_part ? warp_transpose32(m, lane) : 0u;
                    if (lane == 0) sm.full[w][kw] = fw;
                }
                __syncwarp();
                // whole-strip candidates: one candidate at a time for the whole warp
                for (int kw = 0; kw < nw; ++kw) {
                    unsigned fm = sm.full[w][kw];
                    my_pairs32 += __popc(fm);
#if GT_FAST_NP == 2
                    while (fm & (fm - 1u)) {  // two or more left: a group of two
                        int cc[2];
                        cc[0] = kw * 32 + __ffs((int)fm) - 1;
                        fm &= fm - 1u;
                        cc[1] = kw * 32 + __ffs((int)fm) - 1;
                        fm &= fm - 1u;
                        gather_pairs<MODE, 2>(sm, cc, v, ax, gi_s, logtab, acc);
                    }
#endif
                    while (fm) {
                        const int c = kw * 32 + __ffs((int)fm) - 1;
                        fm &= fm - 1u;
                        gather_pair<MODE>(sm, c, v, ax, gi_s, logtab, acc);
                    }
                }
                // partial candidates: every lane pops ITS bits (lanes work on different candidates at the same time;
                // lanes without a pixel have no bits)
                int kw = 0;
                unsigned hm = sm.hits[w][0][lane];
                my_pairs32 += __popc(hm);
                for (;;) {
                    while (hm == 0u && kw + 1 < nw) {
                        hm = sm.hits[w][++kw][lane];
                        my_pairs32 += __popc(hm);
                    }
                    const bool act = hm != 0u;
                    if (!__any_sync(0xffffffffu, act)) break;
                    if (act) {
                        const int c = kw * 32 + __ffs((int)hm) - 1;
                        hm &= hm - 1u;
                        gather_pair<MODE>(sm, c, v, ax, gi_s, logtab, acc);
                    }
                }
                double acc0 = acc[0], acc1 = 0.0, acc2 = 0.0;
                if (MODE == 1) {
                    acc0 = fma(v[0], acc[3], acc[0]);
                    acc1 = fma(v[1], acc[3], acc[1]);
                    acc2 = fma(v[2], acc[3], acc[2]);
                }
                if (valid) {
                    double *o = A.out + (sp + ip - A.pix_base) * ncomp;
                    const bool add = A.accumulate || cb > beg;
                    if (MODE == 0) o[0] = add ? o[0] + acc0 : acc0;
                    else {
                        o[0] = add ? o[0] + acc0 : acc0;
                        o[1] = add ? o[1] + acc1 : acc1;
                        o[2] = add ? o[2] + acc2 : acc2;
                    }
                }
            }
        }
    }
    u64 my_pairs = my_pairs32;
    for (int s = 16; s > 0; s >>= 1) my_pairs += __shfl_down_sync(0xffffffffu, my_pairs, s);
    if (lane == 0 && A.npairs && my_pairs) atomicAdd(A.npairs, my_pairs);
}

// The tables every CTA of the gather kernel stages: ln table, the radial axis as (g[k], 1 / (g[k+1] - g[k])) pairs, the
// plain axis.  Computed once per call (the same expressions the kernel used to evaluate per CTA).
__global__ void __launch_bounds__(256) gather_aux_kernel(bf_raxis ax, double2 *__restrict__ logtab,
                                                         double2 *__restrict__ gi, double *__restrict__ g) {
    const int t = threadIdx.x;
    bf_logtab_init(logtab, t, 256);
    for (int k = t; k < ax.nr; k += 256) {
        const double gk = ax.g[k];
        g[k] = gk;
        gi[k] = make_double2(gk, (k < ax.nr - 1) ? 1.0 / (ax.g[k + 1] - gk) : 0.0);
    }
    if (t == 0 && (ax.nr & 1)) g[ax.nr] = 0.0;  // pad to a multiple of 16 bytes
}

// Baryonify only: discs with < 4 pixels use the 4 interpolation neighbours (HealpixRunner.py:309-310).  These
// halos were diverted by the binning stage; they are applied on top of the gathered offsets with fp64
// atomics (4 pixels each; order only matters if >= 3 of them share a pixel).
#include "bf_pair.cuh"
__device__ __forceinline__ void tiny_one(const bf_hpx &h, const bf_shell_halos &S, const bf_raxis &ax, i64 j, int k,
                                         i64 pix_first, i64 pix_last, double *out, i64 pix_base, u64 *npairs) {
    bf_halo_rec H;
    H.c[0] = S.vec[3 * j]; H.c[1] = S.vec[3 * j + 1]; H.c[2] = S.vec[3 * j + 2];
    H.D = S.D[j]; H.a = S.a[j]; H.cut = S.cut[j]; H.lnr_shift = S.lnr_shift[j];
    H.row = S.rows + j * (i64)ax.nr;
    i64 p[4]; double wgt[4], v[3], off[3];
    bfh_interp_lonlat(h, S.lon[j], S.lat[j], p, wgt);
    if (p[k] < pix_first || p[k] >= pix_last) return;  // another shard owns this pixel
    bfh_pix2vec(h, p[k], v);
    bf_pair_baryonify_shell(H, ax, v, off);
    for (int m = 0; m < 3; ++m) atomicAdd(out + 3 * (p[k] - pix_base) + m, off[m]);
    if (npairs) atomicAdd(npairs, 1ull);
}

__global__ void tiny_kernel(bf_hpx h, bf_shell_halos S, bf_raxis ax, const int *__restrict__ tiny_list,
                            const int *__restrict__ tiny_count, i64 pix_first, i64 pix_last, double *out, i64 pix_base,
                            u64 *npairs) {
    const int k = threadIdx.x & 3, ntiny = *tiny_count;  // (few or none: the grid is small and strides over the list)
    for (int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 2; i < ntiny; i += (gridDim.x * blockDim.x) >> 2)
        tiny_one(h, S, ax, tiny_list[i], k, pix_first, pix_last, out, pix_base, npairs);
}

extern "C" int bf_shell_gather(int64_t nside, int mode, const bf_shell_halos *S, const bf_table *T, int64_t band_lo,
                               int64_t band_hi, const void *ws, double *out, int64_t pix_base, int accumulate,
                               unsigned long long *npairs, void *stream) {
    BF_REQUIRE(nside > 0 && (mode == 0 || mode == 1), "nside / mode");
    BF_REQUIRE(bf_check_table(T), "bad table");
    BF_REQUIRE(S && S->nh >= 0 && ws && out, "null argument");
    bf_cells cg = bf_cells_make(nside);
    BF_REQUIRE(band_lo >= 0 && band_lo <= band_hi && band_hi <= cg.nbands, "band range");
    if (band_hi == band_lo) return BF_OK;
    BF_REQUIRE(S->nh == 0 || (S->vec && S->radius && S->D && S->a && S->rows), "null per-halo array");
    BF_REQUIRE(mode == 0 || S->nh == 0 || (S->cut && S->lnr_shift && S->lon && S->lat), "null per-halo array");
    cudaStream_t s = (cudaStream_t)stream;
    bf_bin_ws w = bf_bin_ws_layout(const_cast<void *>(ws), nside, S->nh, 0);  // fixed part only
    gather_args A;
    A.h = bf_hpx_make(nside);
    A.cg = cg;
    A.S = *S;
    A.ax = bf_make_raxis(T);
    A.keys = w.keys;
    A.cell_begin = w.cell_begin;
    A.cell_end = w.cell_end;
    A.band_lo = band_lo;
    A.out = out;
    A.pix_base = pix_base;
    A.accumulate = accumulate;
    A.npairs = npairs;
    static_assert(sizeof(gather_smem) % 16 == 0, "the ln table behind the struct must stay 16-byte aligned");
    BF_REQUIRE(T->nr <= BF_GATHER_NR_MAX, "table r-axis too long for the shared-memory stage");
    const size_t aux_bytes = sizeof(double2) * BF_LOGTAB + 16 * (size_t)T->nr + ((8 * (size_t)T->nr + 15) & ~(size_t)15);
    size_t smem = sizeof(gather_smem) + aux_bytes;
    BF_REQUIRE(smem <= 200 * 1024, "table r-axis too long for the shared-memory stage");
    {
        char *aux = (char *)w.aux;
        gather_aux_kernel<<<1, 256, 0, s>>>(A.ax, (double2 *)aux, (double2 *)(aux + sizeof(double2) * BF_LOGTAB),
                                            (double *)(aux + sizeof(double2) * BF_LOGTAB + 16 * (size_t)T->nr));
        BF_LAUNCH_OK();
        A.aux = w.aux;
        A.aux_bytes = (unsigned)aux_bytes;
    }
    unsigned grid = (unsigned)((band_hi - band_lo) * BF_MAX_COLS);
    if (mode == 0) {
        BF_CUDA_OK(cudaFuncSetAttribute(gather_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        gather_kernel<0><<<grid, GT_THREADS, smem, s>>>(A);
    } else {
        BF_CUDA_OK(cudaFuncSetAttribute(gather_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        gather_kernel<1><<<grid, GT_THREADS, smem, s>>>(A);
    }
    BF_LAUNCH_OK();
    if (mode == 1 && S->nh > 0) {
        int64_t first = 0, last = 0;
        int rc = bf_shell_band_pixel_range(nside, band_lo, band_hi, &first, &last);
        if (rc) return rc;
        const i64 want = (S->nh * 4 + 127) / 128;
        tiny_kernel<<<(unsigned)(want < 296 ? want : 296), 128, 0, s>>>(A.h, *S, A.ax, w.tiny_list, w.tiny_count, first, last,
                                                                        out, pix_base, npairs);
        BF_LAUNCH_OK();
    }
    return BF_OK;
}
