// bf_pk.cu -- the Delta P(k) summary of a baryonified particle set.
//
// The reference has this only as notebook code (examples/10_Reproduce_Schneider_deltaPk.ipynb): an NGP mass
// assignment of the (optionally folded) particle positions (cell 1, `numba_histogram3d`), np.fft.fftn,
// |delta_k|^2, and a spherical average in Nk linear bins from k_f = 2 pi / L to the Nyquist wavenumber
// (cell 12: `kinds = floor((k - kbins[0]) / (kbins[1] - kbins[0]))`, `np.bincount`; cells 15-21: fold with
// `% Lbox`, S(k) = P_baryon / P_dmo).  Here: a deposit kernel, cuFFT (through torch.fft.rfftn: library FFT, as
// BASELINE.json's config 5 says) and a hand-written, DETERMINISTIC k-binning kernel that reads the half-complex
// spectrum once.
#include "bf_common.cuh"

// ---------------------------------------------------------------------------------------------------
// NGP deposit: counts[ix][iy][iz] += 1 for ix = int((x % Lf) / (Lf / N)).  Counts are whole numbers, so fp64
// atomic adds are exact and the grid is bit-reproducible (and bit-identical to the reference's int64 histogram).
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ int pk_cell(double x, double Lf, double bw, int N) {
    double r = fmod(x, Lf);  // numpy's float %: fmod, then shifted into [0, Lf)
    if (r != 0.0 && r < 0.0) r += Lf;
    int i = (int)(r / bw);
    return max(0, min(i, N - 1));  // (r / bw can round up to N; numba would index out of bounds there)
}

template <bool REC>
__global__ void __launch_bounds__(256) pk_deposit_kernel(i64 np, const double *__restrict__ x, const double *__restrict__ y,
                                                         const double *__restrict__ z, const double4 *__restrict__ rec,
                                                         double Lf, double bw, int N, double *__restrict__ grid) {
    i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= np) return;
    double px, py, pz;
    if (REC) {
        const double2 *r = reinterpret_cast<const double2 *>(rec + p);
        double2 a = __ldcs(r), b = __ldcs(r + 1);
        px = a.y; py = b.x; pz = b.y;
    } else {
        px = __ldcs(x + p); py = __ldcs(y + p); pz = __ldcs(z + p);
    }
    i64 c = ((i64)pk_cell(px, Lf, bw, N) * N + pk_cell(py, Lf, bw, N)) * N + pk_cell(pz, Lf, bw, N);
    atomicAdd(grid + c, 1.0);
}

extern "C" int bf_pk_ngp_deposit(int64_t np, const double *x, const double *y, const double *z, const void *records,
                                 double Lfold, int64_t ngrid, double *grid, void *stream) {
    BF_REQUIRE(np >= 0 && ngrid >= 2 && ngrid <= 4096 && Lfold > 0 && grid, "np / ngrid / Lfold / grid");
    BF_REQUIRE(np == 0 || records || (x && y && z), "particle coordinates");
    BF_REQUIRE(!records || (((uintptr_t)records) & 15) == 0, "particle records must be 16-byte aligned");
    if (np == 0) return BF_OK;
    double bw = (Lfold - 0.0) / (double)ngrid;  // bin_width = (max_vals - min_vals) / bins
    unsigned blocks = (unsigned)((np + 255) / 256);
    if (records)
        pk_deposit_kernel<true><<<blocks, 256, 0, (cudaStream_t)stream>>>(np, nullptr, nullptr, nullptr,
                                                                            (const double4 *)records, Lfold, bw, (int)ngrid, grid);
    else
        pk_deposit_kernel<false><<<blocks, 256, 0, (cudaStream_t)stream>>>(np, x, y, z, nullptr, Lfold, bw, (int)ngrid, grid);
    BF_LAUNCH_OK();
    return BF_OK;
}

// ---------------------------------------------------------------------------------------------------
// k-binning of the half-complex spectrum F[i][j][l], l = 0 .. N/2 (torch.fft.rfftn layout, complex128).
// Mode (i, j, l) with 0 < l < N/2 stands for itself and its conjugate partner (weight 2).  Bin index exactly as
// the notebook computes it: klin[m] = m * val (np.fft.fftfreq), k = sqrt((kx^2 + kz^2) + ky^2), floor((k-k0)/dk).
// Deterministic reduction: every warp owns a private set of bins in shared memory; lanes that hit the same bin in
// the same step are combined by the lowest such lane in lane order (__match_any_sync), so no atomics are needed;
// warps -> block and blocks -> result are added in index order.
// ---------------------------------------------------------------------------------------------------
#define PK_THREADS 256
#define PK_WARPS (PK_THREADS / 32)

struct pk_args {
    const double2 *F;
    int N, nh;  // grid size, N/2 + 1
    double val, k0, dk;
    int nk;
    i64 nmodes, per_block;
    int mem_n[3], mem_ax[3];  // memory order (slowest first): extent and LOGICAL axis (0 = i, 1 = j, 2 = l) of each level
    int off[3];               // global index of the piece's first element along each LOGICAL axis (slab-decomposed FFT)
    double *partial;          // [nblocks][3][nk]  (power, k, count)
};

__global__ void __launch_bounds__(PK_THREADS) pk_bin_kernel(pk_args A) {
    extern __shared__ double sm[];  // [PK_WARPS][3][nk]
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    for (int i = t; i < PK_WARPS * 3 * A.nk; i += PK_THREADS) sm[i] = 0.0;
    __syncthreads();
    double *mine = sm + (size_t)w * 3 * A.nk;
    const i64 b0 = (i64)blockIdx.x * A.per_block, b1 = min(A.nmodes, b0 + A.per_block);
    const int half = A.N / 2;
    for (i64 base = b0 + (i64)w * 32; base < b1; base += PK_THREADS) {  // all lanes of a warp iterate together
        const i64 m = base + lane;
        int bin = -1;
        double pw = 0.0, kk = 0.0, wt = 0.0;
        if (m < b1) {
            int idx[3];  // m runs in MEMORY order (cuFFT's output of a real 3-D transform is usually a permuted view)
            const int c2 = (int)(m % A.mem_n[2]);
            const i64 r01 = m / A.mem_n[2];
            idx[A.mem_ax[2]] = c2;
            idx[A.mem_ax[1]] = (int)(r01 % A.mem_n[1]);
            idx[A.mem_ax[0]] = (int)(r01 / A.mem_n[1]);
            const int i = idx[0] + A.off[0], j = idx[1] + A.off[1], l = idx[2] + A.off[2];
            const double kx = (double)((i < (A.N + 1) / 2) ? i : i - A.N) * A.val;
            const double ky = (double)((j < (A.N + 1) / 2) ? j : j - A.N) * A.val;
            const double kz = (double)((l < (A.N + 1) / 2) ? l : l - A.N) * A.val;  // l = N/2 -> -N/2 (same square)
            kk = sqrt((kx * kx + kz * kz) + ky * ky);
            const double q = floor((kk - A.k0) / A.dk);
            if (q >= 0.0 && q < (double)A.nk) {
                bin = (int)q;
                const double2 f = A.F[m];
                wt = ((A.N % 2 == 0) ? (l > 0 && l < half) : (l > 0)) ? 2.0 : 1.0;
                pw = (f.x * f.x + f.y * f.y) * wt;  // (conj(F) * F).real
                kk *= wt;
            }
        }
        const unsigned grp = __match_any_sync(0xffffffffu, bin);
        if (bin >= 0) {
            const int leader = __ffs((int)grp) - 1;
            double sp = 0.0, sk = 0.0, sc = 0.0;
            for (unsigned g = grp; g; g &= g - 1) {  // members in lane order
                const int src = __ffs((int)g) - 1;
                sp += __shfl_sync(grp, pw, src);
                sk += __shfl_sync(grp, kk, src);
                sc += __shfl_sync(grp, wt, src);
            }
            if (lane == leader) {
                mine[bin] += sp;
                mine[A.nk + bin] += sk;
                mine[2 * A.nk + bin] += sc;
            }
        }
        __syncwarp();
    }
    __syncthreads();
    for (int i = t; i < 3 * A.nk; i += PK_THREADS) {
        double s = 0.0;
        for (int ww = 0; ww < PK_WARPS; ++ww) s += sm[(size_t)ww * 3 * A.nk + i];
        A.partial[(size_t)blockIdx.x * 3 * A.nk + i] = s;
    }
}

__global__ void pk_finish_kernel(int nblocks, int nk, const double *__restrict__ partial, double *__restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 3 * nk) return;
    double s = 0.0;
    for (int b = 0; b < nblocks; ++b) s += partial[(size_t)b * 3 * nk + i];
    out[i] = s;
}

static int pk_blocks(i64 nmodes) {
    i64 want = (i64)bf_sm_count() * 8;
    i64 nb = (nmodes + PK_THREADS - 1) / PK_THREADS;
    return (int)((nb < want) ? (nb < 1 ? 1 : nb) : want);
}

static size_t pk_ws_bytes(i64 nmodes, i64 nk) { return sizeof(double) * 3 * (size_t)nk * (size_t)pk_blocks(nmodes) + 256; }

extern "C" size_t bf_pk_bin_ws_bytes(int64_t ngrid, int64_t nk) {
    if (ngrid < 2 || nk < 1) return 0;
    return pk_ws_bytes(ngrid * ngrid * (ngrid / 2 + 1), nk);
}

extern "C" size_t bf_pk_bin_part_ws_bytes(const int64_t *ext_host, int64_t nk) {
    if (!ext_host || nk < 1 || ext_host[0] < 0 || ext_host[1] < 0 || ext_host[2] < 0) return 0;
    return pk_ws_bytes(ext_host[0] * ext_host[1] * ext_host[2], nk);
}

// A rectangular PIECE of the half-complex spectrum: logical extents ext[3] (axes i, j, l), first global index off[3].
// The slab-decomposed transform leaves every rank with all i, a range of j and all l; the per-rank [3][nk] sums are
// then added across ranks (sharding.ShardedPowerSpectrum).  The whole spectrum is the piece with off = 0.
extern "C" int bf_pk_bin_part(int64_t ngrid, const void *spectrum, const int64_t *strides_host, const int64_t *ext_host,
                              const int64_t *off_host, double val, double k0, double dk, int64_t nk,
                              double *out /* [3][nk]: sum |F|^2, sum k, count */, void *ws, size_t ws_bytes,
                              void *stream) {
    BF_REQUIRE(ngrid >= 2 && ngrid <= 4096 && out && ext_host && off_host, "ngrid / out / ext / off");
    BF_REQUIRE(nk >= 1 && nk <= 768 && dk > 0, "nk must be in [1, 768], dk > 0");
    const i64 nhalf = ngrid / 2 + 1;
    i64 ext[3] = {ext_host[0], ext_host[1], ext_host[2]}, full[3] = {ngrid, ngrid, nhalf};
    for (int a = 0; a < 3; ++a)
        BF_REQUIRE(ext[a] >= 0 && off_host[a] >= 0 && off_host[a] + ext[a] <= full[a], "piece outside the spectrum");
    size_t need = bf_pk_bin_part_ws_bytes(ext_host, nk);
    if (!ws || ws_bytes < need) {
        bf_set_error("bf_pk_bin: workspace %zu < %zu bytes", ws_bytes, need);
        return BF_E_WORKSPACE;
    }
    pk_args A;
    A.F = (const double2 *)spectrum;
    A.N = (int)ngrid;
    A.nh = (int)nhalf;
    A.val = val; A.k0 = k0; A.dk = dk;
    A.nk = (int)nk;
    A.nmodes = ext[0] * ext[1] * ext[2];
    cudaStream_t s = (cudaStream_t)stream;
    if (A.nmodes == 0) {
        BF_CUDA_OK(cudaMemsetAsync(out, 0, sizeof(double) * 3 * (size_t)nk, s));
        return BF_OK;
    }
    BF_REQUIRE(spectrum && (((uintptr_t)spectrum) & 15) == 0, "spectrum must be 16-byte aligned complex128");
    {   // memory order of the three logical axes from their element strides (NULL = C order); must be dense
        i64 st[3] = {ext[1] * ext[2], ext[2], 1};
        if (strides_host) { st[0] = strides_host[0]; st[1] = strides_host[1]; st[2] = strides_host[2]; }
        int ord[3] = {0, 1, 2};
        for (int a = 0; a < 3; ++a)
            for (int b = a + 1; b < 3; ++b)
                if (st[ord[b]] > st[ord[a]] || (st[ord[b]] == st[ord[a]] && ext[ord[b]] > ext[ord[a]])) {
                    int t = ord[a]; ord[a] = ord[b]; ord[b] = t;
                }
        // (an axis of extent 1 may carry any stride: it never moves)
        bool dense = (ext[ord[2]] == 1 || st[ord[2]] == 1) && (ext[ord[1]] == 1 || st[ord[1]] == ext[ord[2]]) &&
                     (ext[ord[0]] == 1 || st[ord[0]] == ext[ord[2]] * ext[ord[1]]);
        BF_REQUIRE(dense, "spectrum must be a dense permutation of its logical extents");
        for (int d = 0; d < 3; ++d) { A.mem_ax[d] = ord[d]; A.mem_n[d] = (int)ext[ord[d]]; }
    }
    for (int a = 0; a < 3; ++a) A.off[a] = (int)off_host[a];
    int nb = pk_blocks(A.nmodes);
    A.per_block = ((A.nmodes + nb - 1) / nb + 31) / 32 * 32;
    A.partial = (double *)ws;
    size_t smem = sizeof(double) * PK_WARPS * 3 * (size_t)nk;
    BF_CUDA_OK(cudaFuncSetAttribute(pk_bin_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    pk_bin_kernel<<<nb, PK_THREADS, smem, s>>>(A);
    BF_LAUNCH_OK();
    pk_finish_kernel<<<(unsigned)((3 * nk + 127) / 128), 128, 0, s>>>(nb, (int)nk, A.partial, out);
    BF_LAUNCH_OK();
    return BF_OK;
}

extern "C" int bf_pk_bin(int64_t ngrid, const void *spectrum, const int64_t *strides_host, double val, double k0,
                         double dk, int64_t nk, double *out /* [3][nk]: sum |F|^2, sum k, count */, void *ws,
                         size_t ws_bytes, void *stream) {
    BF_REQUIRE(ngrid >= 2 && ngrid <= 4096 && spectrum && out, "ngrid / spectrum / out");
    const int64_t ext[3] = {ngrid, ngrid, ngrid / 2 + 1}, off[3] = {0, 0, 0};
    return bf_pk_bin_part(ngrid, spectrum, strides_host, ext, off, val, k0, dk, nk, out, ws, ws_bytes, stream);
}

// ---------------------------------------------------------------------------------------------------
// bf_pk_bin without a caller workspace and without any allocation (no counterpart in the reference; added
// for callers that cannot allocate: CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc).  The per-block partial sums [blocks][3][nk] that bf_pk_bin keeps in the caller's workspace live in a
// file-scope device array of fixed size instead; every GPU that loads the library holds these 8 MiB until exit.
// In HOST code the array's name is only a stand-in for the symbol, not a device address: the address is asked for once per
// device, with that device current (cudaGetSymbolAddress), and that pointer is what the kernels get.
// All calls on a device share the array: calls that overlap on the GPU (different streams) race.  Nothing in it has to
// start at zero -- every block writes all of its 3 * nk partial sums before the finishing kernel reads them -- so a call
// does not depend on what the previous one left behind.
// ---------------------------------------------------------------------------------------------------
#define PK_STATIC_DOUBLES (1 << 20)  // 8 MiB
#define PK_STATIC_MAX_NK 256
#define PK_STATIC_MAX_DEVICES 64
static __device__ __align__(256) double pk_static_scratch[PK_STATIC_DOUBLES];
static double *pk_static_ptr[PK_STATIC_MAX_DEVICES];  // device address of the array, per device (0 = not asked yet)

extern "C" size_t bf_pk_bin_static_max_bytes(void) { return sizeof(double) * (size_t)PK_STATIC_DOUBLES; }

extern "C" int bf_pk_bin_static(int64_t ngrid, const void *spectrum, const int64_t *strides_host, double val, double k0,
                                double dk, int64_t nk, double *out /* [3][nk] */, void *stream) {
    BF_REQUIRE(ngrid >= 2 && ngrid <= 4096 && spectrum && out, "ngrid / spectrum / out");
    if (nk < 1 || nk > PK_STATIC_MAX_NK) {
        bf_set_error("bf_pk_bin_static: nk = %lld is outside [1, %d], the most the built-in scratch array is sized for; "
                     "use bf_pk_bin with a workspace", (long long)nk, PK_STATIC_MAX_NK);
        return BF_E_INVALID;
    }
    const size_t need = bf_pk_bin_ws_bytes(ngrid, nk), have = bf_pk_bin_static_max_bytes();
    if (need == 0 || need > have) {
        bf_set_error("bf_pk_bin_static: ngrid = %lld, nk = %lld need %zu bytes of scratch on this device, the built-in array "
                     "holds %zu; use bf_pk_bin with a workspace", (long long)ngrid, (long long)nk, need, have);
        return BF_E_WORKSPACE;
    }
    int dev = 0;
    BF_CUDA_OK(cudaGetDevice(&dev));
    BF_REQUIRE(dev >= 0 && dev < PK_STATIC_MAX_DEVICES, "device ordinal beyond the per-device table");
    if (!pk_static_ptr[dev]) {  // set-up, once per device, with that device current
        double *p = nullptr;
        BF_CUDA_OK(cudaGetSymbolAddress((void **)&p, pk_static_scratch));
        BF_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned NULL");
        pk_static_ptr[dev] = p;
    }
    return bf_pk_bin(ngrid, spectrum, strides_host, val, k0, dk, nk, out, pk_static_ptr[dev], have, stream);
}
