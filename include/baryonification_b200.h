/*
 * baryonification_b200.h -- C ABI of the B200-native halo displacement-and-painting path.
 *
 * Drop-in boundary for the per-halo loops of BaryonForge's Runners.  The reference has no FFI
 * (it is pure Python); each entry point cites the reference lines whose work it replaces
 * (paths relative to BaryonForge/).  INTEGRATION.md shows the ctypes binding a maintainer adds.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer unless the name ends in _host; sizes are element counts;
 *  - `stream` is a cudaStream_t passed as void*; all work is stream-ordered, nothing synchronises
 *    unless documented; the library never allocates or frees caller memory: scratch comes from a
 *    caller workspace whose size the matching *_ws_bytes() function returns;
 *  - return value: 0 = ok, >0 = cudaError_t, <0 = BF_E_* ; bf_last_error() gives a thread-local
 *    message; nothing throws across the boundary;
 *  - all floating point is IEEE double, all pixel / cell / particle indices are int64.
 */
#ifndef BARYONIFICATION_B200_H
#define BARYONIFICATION_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BF_OK 0
#define BF_E_INVALID (-1)   /* bad argument */
#define BF_E_WORKSPACE (-2) /* workspace too small */
#define BF_E_CAPACITY (-4)  /* bf_snap_bin_halos: >= 2^25 (cell, halo) entries at this `ng`; call again with fewer cells */
#define BF_E_WINDOW (-3)    /* bf_{grid,shell}_regrid_gather: displacement window too wide, use the _sorted variant */

/* `mode` of the pair entry points.  The two runner families historically number it differently -- use the names:
 * shell entries (bf_shell_bin_halos, bf_shell_gather): 0 = paint, 1 = baryonify;
 * grid entries (bf_grid_gather, bf_grid_gather_rows):  0 = baryonify, 1 = paint. */
#define BF_SHELL_MODE_PAINT 0
#define BF_SHELL_MODE_BARYONIFY 1
#define BF_GRID_MODE_BARYONIFY 0
#define BF_GRID_MODE_PAINT 1

const char *bf_last_error(void);
int bf_version(void);
int bf_device_sm_count(int *sm_count_host);
/* Kernel launches of this library since it was loaded (CUB's internal kernels not included). */
unsigned long long bf_launch_count(void);
/* Page-lock a caller's host array where it lies / undo it (cudaHostRegister / cudaHostUnregister with the runtime's error
 * state left clean on failure).  bf_host_pin: 0 = registered now, 1 = the range is already page-locked (nothing to undo),
 * BF_E_INVALID = not possible.  No counterpart in the reference (numpy arrays are pageable): plumbing of the runners'
 * upload path, which registers inputs >= 64 MiB that are processed a second time (parameter sweeps). */
int bf_host_pin(void *ptr, size_t bytes);
int bf_host_unpin(void *ptr);
/* Throughput probe: `blocks` CTAs of 256 threads, each thread runs 8 independent chains of `iters` fused multiply-adds
 * in fp64; out is [blocks * 256] doubles (scratch), *dfma_per_thread_host = 8 * iters.  Time it with events. */
int bf_probe_fp64(int blocks, int iters, double *out, unsigned long long *dfma_per_thread_host, void *stream);

/* A (ln(1+z), ln M, ln r) table resident on the device.  `values` holds d (displacement,
 * Profiles/BaryonCorrection.py:315) or ln(profile) (utils/Tabulate.py:237-238). */
typedef struct {
    const double *gz, *gM, *gr; /* ascending axes */
    const double *values;       /* [nz][nM][nr], C order */
    int64_t nz, nM, nr;
    double lnr_first, lnr_last; /* gr[0], gr[nr-1] (host copies, so no launch needs a read-back) */
} bf_table;

/* ---- table readout (BaryonCorrection.py:356-390, Tabulate.py:272-294, scipy RGI 'linear') ---- */
/* rows[j][:] = blend of the 4 (z,M) corners for halo j; rows of out-of-table halos are NaN. */
int bf_table_collapse_zM(const bf_table *tab_host, int64_t nh, const double *z_in, const double *M_in,
                         double *rows /* [nh][nr] */, void *stream);
/* Tables with extra property axes (`p_keys`: other_params of BaryonCorrection.py:205-221 and
 * ParamTabulatedProfile, Tabulate.py:524-561; the per-halo values are catalogue columns, HealpixRunner.py:425).
 * The host moves the r axis last: values is [n[0]][n[1]]...[n[naxes-1]][nr] with axes (z, M, extras...), C order.
 * coords[j][a] is halo j's coordinate on axis a; rows[j][:] = N-linear blend of the 2^naxes corners (scipy RGI
 * weights, multiplied in axis order), NaN when any coordinate is outside its axis. */
#define BF_MAX_AXES 6
typedef struct {
    int naxes;
    const double *grid[BF_MAX_AXES]; /* ascending axes */
    int64_t n[BF_MAX_AXES];
    const double *values;
    int64_t nr;
} bf_table_nd;
int bf_table_collapse_nd(const bf_table_nd *tab_host, int64_t nh, const double *coords /* [nh][naxes] */,
                         double *rows /* [nh][nr] */, void *stream);
/* out[i] = readout of halo `row` at radius r[i]; mode 0: linear table with cut (r < cut else 0),
 * mode 1: exp(log table).  lnr_shift is subtracted from ln r (Rdelta_sampling). */
int bf_table_readout(const bf_table *tab_host, const double *row, int64_t n, const double *r,
                     double lnr_shift, double cut, int mode, double *out, void *stream);

/* ---- HEALPix RING geometry (healpy call sites HealpixRunner.py:303-337,429-433) ---- */
int bf_healpix_ang2pix(int64_t nside, int64_t n, const double *theta, const double *phi, int64_t *pix, void *stream);
int bf_healpix_pix2ang(int64_t nside, int64_t n, const int64_t *pix, double *theta, double *phi, void *stream);
int bf_healpix_pix2vec(int64_t nside, int64_t n, const int64_t *pix, double *vec /* [n][3] */, void *stream);
int bf_healpix_ang2vec_lonlat(int64_t n, const double *lon, const double *lat, double *vec, void *stream);
int bf_healpix_interp_weights_lonlat(int64_t nside, int64_t n, const double *lon, const double *lat,
                                     int64_t *pix /* [n][4] */, double *wgt /* [n][4] */, void *stream);
/* query_disc(inclusive=False, nest=False) for n discs at once: counts[j] = #pixels of disc j.
 * If pix != NULL, disc j's ascending pixel list is written at pix[offsets[j] ...]. */
int bf_healpix_query_disc(int64_t nside, int64_t n, const double *vec /* [n][3] */, const double *radius,
                          int64_t *counts, const int64_t *offsets, int64_t *pix, void *stream);

/* ---- per-halo record of the Shell runners (HealpixRunner.py:293-305, 420-431) ---- */
typedef struct {
    const double *vec;       /* [nh][3] hp.ang2vec(ra, dec, lonlat=True) */
    const double *lon, *lat; /* ra, dec in degrees (for the <4-pixel fallback, :309-310) */
    const double *radius;    /* R_j * epsilon_max / D_j  [rad] */
    const double *D, *a;     /* D_A(z_j) [physical Mpc], scale factor */
    const double *cut;       /* model.epsilon_max * R_model_j (comoving); unused for paint */
    const double *lnr_shift; /* ln R_model_j if Rdelta_sampling else 0; unused for paint */
    const double *rows;      /* [nh][nr] from bf_table_collapse_zM */
    const int64_t *order;    /* optional processing order for the halo-centric kernels (e.g. radius-
                                descending for load balance), or NULL for catalogue order */
    int64_t nh;
} bf_shell_halos;

/* PaintProfilesShell per-halo loop (HealpixRunner.py:418-445): map[pix] += P(r).  Atomic variant. */
int bf_shell_paint_atomic(int64_t nside, const bf_shell_halos *h_host, const bf_table *tab_host,
                          double *map, unsigned long long *npairs, void *stream);
/* BaryonifyShell per-halo loop (HealpixRunner.py:291-331): pix_offsets[pix][3] += unit-vector shift. */
int bf_shell_baryonify_atomic(int64_t nside, const bf_shell_halos *h_host, const bf_table *tab_host,
                              double *pix_offsets, unsigned long long *npairs, void *stream);
/* ---- deterministic gather variant: halo binning into sky cells + pixel-owner evaluation ----
 * The sphere is cut into cells of 8 consecutive rings x one column of `Wc` consecutive in-ring pixel
 * indices (Wc = 32*CW, at most 64 columns per ring).  bf_shell_bin_halos lists, for every cell, the halos
 * whose query_disc footprint touches it, in catalogue order (this replaces hp.query_disc's role of
 * deciding WHICH pixels a halo updates, HealpixRunner.py:306/:432).  bf_shell_gather then gives every
 * pixel to one lane, which walks its cell's list in catalogue order and sums in registers: no atomics,
 * bit-reproducible, and the same per-pixel summation order as the reference's serial halo loop.
 * Only cells of ring bands [band_lo, band_hi) are binned / evaluated (multi-GPU ring sharding);
 * pass 0 and bf_shell_num_bands(nside) for the whole sphere. */
int64_t bf_shell_num_bands(int64_t nside);
/* pixel range [first, last) covered by ring bands [band_lo, band_hi) */
int bf_shell_band_pixel_range(int64_t nside, int64_t band_lo, int64_t band_hi, int64_t *first_host, int64_t *last_host);
size_t bf_shell_bin_ws_bytes(int64_t nside, int64_t nh, int64_t max_entries);
/* mode 0 = paint, 1 = baryonify (baryonify diverts discs with < 4 pixels to the tiny-halo list).
 * On return *nentries_host = number of (cell, halo) entries; BF_E_WORKSPACE if it exceeds max_entries
 * (call again with a workspace sized for *nentries_host).  Synchronises `stream` once. */
int bf_shell_bin_halos(int64_t nside, int mode, const bf_shell_halos *h_host, int64_t band_lo, int64_t band_hi,
                       int64_t max_entries, void *ws, size_t ws_bytes, int64_t *nentries_host, void *stream);
/* The same in three steps, for per-halo arrays that arrive on the device piece by piece (the host prepares the catalogue
 * in runs of consecutive halos): _begin resets the key cursor, _add emits the keys of halos j_first <= j < j_last (the
 * arrays of `h_host` are the whole catalogue's; entries at and beyond j_last need not be valid yet), _end reads the count
 * back (synchronises `stream`), sorts and marks the cell boundaries.  bf_shell_bin_halos = begin, add(0, nh), end; the
 * same workspace and max_entries in all three calls. */
int bf_shell_bin_begin(int64_t nside, int mode, const bf_shell_halos *h_host, int64_t band_lo, int64_t band_hi,
                       int64_t max_entries, void *ws, size_t ws_bytes, void *stream);
int bf_shell_bin_add(int64_t nside, int mode, const bf_shell_halos *h_host, int64_t j_first, int64_t j_last,
                     int64_t band_lo, int64_t band_hi, int64_t max_entries, void *ws, size_t ws_bytes, void *stream);
int bf_shell_bin_end(int64_t nside, int mode, const bf_shell_halos *h_host, int64_t band_lo, int64_t band_hi,
                     int64_t max_entries, void *ws, size_t ws_bytes, int64_t *nentries_host, void *stream);
/* out: paint -> map[npix]; baryonify -> pix_offsets[npix][3]; both indexed by GLOBAL pixel number minus
 * `pix_base` (the first pixel of band_lo when sharded, else 0).  Pixels of the evaluated bands are
 * OVERWRITTEN (+= when accumulate != 0). */
int bf_shell_gather(int64_t nside, int mode, const bf_shell_halos *h_host, const bf_table *tab_host,
                    int64_t band_lo, int64_t band_hi, const void *ws, double *out, int64_t pix_base, int accumulate,
                    unsigned long long *npairs, void *stream);

/* BaryonifyShell re-binning (HealpixRunner.py:333-341 + regrid_pixels_hpix :13-67). */
int bf_shell_regrid(int64_t nside, const double *orig_map, const double *pix_offsets, double *new_map,
                    void *stream);
/* Same for the pixel slice [pix_first, pix_last) (ring sharding): orig_map / pix_offsets are the LOCAL slices,
 * new_map is the full-sphere accumulator (deposits may land outside the slice: boundary mass). */
int bf_shell_regrid_range(int64_t nside, int64_t pix_first, int64_t pix_last, const double *orig_map,
                          const double *pix_offsets, double *new_map, void *stream);

/* Deterministic variant of the re-binning: every deposit is emitted as a (destination pixel, w * mass) entry,
 * entries are radix-sorted by destination (stable, so a destination's entries stay in source order) and each
 * segment is summed left to right -- the summation order of the reference's serial numba loop
 * (regrid_pixels_hpix, HealpixRunner.py:13-67), bit-reproducible.  new_map is += (zero it first). */
size_t bf_shell_regrid_sorted_ws_bytes(int64_t nside, int64_t npix_local);
int bf_shell_regrid_sorted(int64_t nside, int64_t pix_first, int64_t pix_last, const double *orig_map,
                           const double *pix_offsets, double *new_map, void *ws, size_t ws_bytes, void *stream);
/* The same sums (same deposits, same ascending (source, j) order: bit-identical to bf_shell_regrid_sorted) as a
 * DESTINATION GATHER over the whole sphere: displaced mass lands next to its source, so a first pass writes every
 * pixel's 4 deposits and measures the field's window (range of ring(destination) - ring(source); largest in-ring
 * distance between a source and its destination's projection on the source ring), and every destination then scans the
 * sources inside that window.  No sort, no atomics on the map, 48 B of workspace per pixel.  Destinations within
 * BF_SHELL_GATHER_POLAR rings of a pole scan whole source rings.  window_host[3] (may be NULL) receives
 * (dr_min, dr_max, dj_max).  BF_E_WINDOW when a destination would have to scan more than BF_SHELL_GATHER_MAX_WINDOW
 * sources (fall back to bf_shell_regrid_sorted).  Every pixel of new_map is WRITTEN (0 where nothing lands): no zero
 * fill is needed.  Synchronises `stream` once. */
#define BF_SHELL_GATHER_POLAR 8
#define BF_SHELL_GATHER_MAX_WINDOW 165
size_t bf_shell_regrid_gather_ws_bytes(int64_t nside);
int bf_shell_regrid_gather(int64_t nside, const double *orig_map, const double *pix_offsets, double *new_map, void *ws,
                           size_t ws_bytes, int *window_host, void *stream);
/* The two passes on their own, so that a caller can produce the map in destination ranges and start reading each
 * range back while the next one is computed: _emit writes the deposits into ws and returns the window in
 * window_host[3] (an empty field gives dr_min > dr_max; BF_E_WINDOW as above; synchronises `stream` once), _range
 * writes the sums of destinations dest_first <= d < dest_last to new_map (=; 0 where nothing lands), given that window. */
int bf_shell_regrid_gather_emit(int64_t nside, const double *orig_map, const double *pix_offsets, void *ws,
                                size_t ws_bytes, int *window_host, void *stream);
int bf_shell_regrid_gather_range(int64_t nside, const void *ws, size_t ws_bytes, const int *window_host,
                                 int64_t dest_first, int64_t dest_last, double *new_map, void *stream);
/* _emit for the sources pix_first <= p < pix_last only (orig_map / pix_offsets are that slice): what a ring-band shard
 * runs.  Keys within BF_SHELL_GATHER_MARGIN_RINGS * 4 * nside pixels of the range are set to "no deposit", so that
 * _range over the destinations the shard's sources can reach (its range widened by (max |dr| + 1) * 4 * nside pixels)
 * yields the shard's PARTIAL sums: own destinations complete up to the neighbours' boundary mass, which the shards
 * exchange and add (sharding.exchange_boundary).  BF_E_WINDOW also when the reach exceeds half the margin. */
#define BF_SHELL_GATHER_MARGIN_RINGS 192
int bf_shell_regrid_gather_emit_range(int64_t nside, int64_t pix_first, int64_t pix_last, const double *orig_map,
                                      const double *pix_offsets, void *ws, size_t ws_bytes, int *window_host,
                                      void *stream);
/* _emit for one GROUP of a pipelined whole-sphere run: only the keys of pix_first <= p < pix_last are written (the other
 * groups' keys, already there, survive) and window_host[3] is the window of THIS group's sources.  The caller emits the
 * groups in ascending pixel order and may run _range over destinations all of whose possible sources -- the rings
 * r_d - dr_max ... r_d - dr_min of the UNION of the windows so far -- have been emitted.  Only BF_E_WINDOW for cost. */
int bf_shell_regrid_gather_emit_part(int64_t nside, int64_t pix_first, int64_t pix_last, const double *orig_map,
                                     const double *pix_offsets, void *ws, size_t ws_bytes, int *window_host,
                                     void *stream);

/* ---- Grid runners (Map2DRunner.py) ---- */
typedef struct {
    const double *pos;       /* [nh][dim] halo x,y[,z] (float32 catalogue values widened) */
    const int64_t *Nsize;    /* cut-out width in cells (:494-497 / :731-733); < 2 = skip */
    const double *cut;       /* baryonify: model cut; paint: R_j*epsilon_max (:804) */
    const double *lnr_shift;
    const double *rows;      /* [nh][nr] */
    const double *Rmat;      /* optional [nh][4]: 2x2 shear matrix of build_Rmat (:275-344), row-major; 2-D only.
                                The table radius becomes |(x, y) @ Rmat| (:525-530, :758-763); NULL = round halos */
    int64_t nh;
} bf_grid_halos;

/* BaryonifyGrid per-halo loop (:476-580): offsets[cell][dim] += d/res * hat   (atomic variant) */
int bf_grid_baryonify_atomic(int dim, int64_t npix, const double *bins, double res, const bf_grid_halos *h_host,
                             const bf_table *tab_host, double *offsets, unsigned long long *npairs, void *stream);
/* PaintProfilesGrid per-halo loop (:714-810): out[cell] += P masked by finite & r < cut */
int bf_grid_paint_atomic(int dim, int64_t npix, const double *bins, double res, const bf_grid_halos *h_host,
                         const bf_table *tab_host, double *out, unsigned long long *npairs, void *stream);
/* Deterministic cell-owner variant of the two loops above.  bf_grid_bin_halos intersects every halo's cut-out box
 * (pick_indices :394-423) with tiles of 8^3 (3-D) / 16 x 32 (2-D) cells and leaves, per tile, its halos in catalogue
 * order; bf_grid_gather runs one CTA per tile, one thread per cell, and sums the cell's pairs in registers in that
 * order (the reference's serial order): no atomics, bit-reproducible, out is += once per cell.
 * Protocol as bf_shell_bin_halos: *nentries_host = exact number of (tile, halo) entries, BF_E_WORKSPACE when it exceeds
 * max_entries (call again with a workspace of bf_grid_bin_ws_bytes for it).  BF_E_WINDOW when a cut-out is wider than
 * the grid (Nsize > npix: the reference's fancy-index += then drops duplicates; use the halo-centric kernels).
 * nr = row length of h_host->rows (= the table's nr).  mode 0 = BaryonifyGrid, 1 = PaintProfilesGrid.  The same
 * max_entries must be given to bf_grid_gather.  bf_grid_bin_halos synchronises `stream` once. */
size_t bf_grid_bin_ws_bytes(int dim, int64_t npix, int64_t nh, int64_t max_entries);
int bf_grid_bin_halos(int dim, int64_t npix, const double *bins, const bf_grid_halos *h_host, int64_t nr,
                      int64_t max_entries, void *ws, size_t ws_bytes, int64_t *nentries_host, void *stream);
int bf_grid_gather(int dim, int mode, int64_t npix, double res, const bf_grid_halos *h_host, const bf_table *tab_host,
                   int64_t max_entries, void *ws, double *out, unsigned long long *npairs, void *stream);
/* Slab form for one rank of a multi-GPU run: only the cells of rows a_first <= a < a_last (first grid axis) are
 * evaluated and written; `out` is still the full [npix^dim] array (global indexing).  *npairs counts this slab. */
int bf_grid_gather_rows(int dim, int mode, int64_t npix, double res, const bf_grid_halos *h_host,
                        const bf_table *tab_host, int64_t max_entries, void *ws, double *out, int64_t a_first,
                        int64_t a_last, unsigned long long *npairs, void *stream);
/* BaryonifyGrid tail (:583-607 + regrid_pixels_2D/3D :14-163): CIC-equivalent periodic deposit. */
int bf_grid_regrid(int dim, int64_t npix, const double *offsets, const double *orig, double *new_map, void *stream);

/* deterministic (sort-by-cell + in-order segmented sum) variant; new_map is += */
size_t bf_grid_regrid_sorted_ws_bytes(int dim, int64_t npix);
int bf_grid_regrid_sorted(int dim, int64_t npix, const double *offsets, const double *orig, double *new_map, void *ws,
                          size_t ws_bytes, void *stream);

/* deterministic destination-gather variant (same bits as the sorted one, no per-deposit workspace): a first pass
 * measures the range of floor(displaced position) - cell per axis; every cell then sums the deposits of the sources
 * inside that window in ascending source order.  ws = bf_grid_regrid_gather_ws_bytes() device bytes;
 * window_host[3] (may be NULL) receives the window length per axis.  BF_E_WINDOW when a cell would have to scan more
 * than BF_GRID_GATHER_MAX_WINDOW_{2D,3D} sources (fall back to bf_grid_regrid_sorted).  new_map is +=.
 * Synchronises `stream` once. */
#define BF_GRID_GATHER_MAX_WINDOW_2D 121
#define BF_GRID_GATHER_MAX_WINDOW_3D 343
size_t bf_grid_regrid_gather_ws_bytes(void);
int bf_grid_regrid_gather(int dim, int64_t npix, const double *offsets, const double *orig, double *new_map, void *ws,
                          size_t ws_bytes, int *window_host, void *stream);
/* The two halves of bf_grid_regrid_gather for slab-sharded runs (global indexing, full-size arrays):
 * bf_grid_regrid_window measures lohi_host[0..2] = min, [3..5] = max of floor(displaced position) - cell over the
 * source rows a_first <= a < a_last (ranks combine theirs with MIN / MAX); bf_grid_regrid_gather_rows writes the
 * destination rows a_first <= a < a_last, reading sources from rows a_first - hi_a - 1 ... a_last - 1 - lo_a
 * (periodic), which the caller must have filled (own rows + ghost rows of the neighbouring slabs). */
int bf_grid_regrid_window(int dim, int64_t npix, const double *offsets, int64_t a_first, int64_t a_last, void *ws,
                          size_t ws_bytes, int *lohi_host, void *stream);
int bf_grid_regrid_gather_rows(int dim, int64_t npix, const double *offsets, const double *orig, double *new_map,
                               const int *lohi_host, int64_t a_first, int64_t a_last, int *window_host, void *stream);

/* ---- Snapshot runner (SnapshotRunner.py) ---- */
typedef struct {
    const double *pos; /* [nh][dim] */
    const double *Rq;  /* clip(eps*R_j/a, 0, L/2) (:213-215) */
    const double *cut, *lnr_shift, *rows;
    int64_t nh;
} bf_snap_halos;

/* The periodic KD-tree (SnapshotRunner.py:92-97) is replaced by a fine periodic halo grid: every halo is listed in
 * each cell its query ball overlaps, lists are in catalogue order.  `ng` = cells per side;
 * bf_snap_grid_cells() proposes one cell per mean query radius (capped at 256 in 3-D, 4096 in 2-D). */
int64_t bf_snap_grid_cells(int dim, double L, double Rq_mean);
size_t bf_snap_ws_bytes(int dim, int64_t nh, double L, int64_t ng, int64_t max_entries);
/* Same protocol as bf_shell_bin_halos: *nentries_host = exact number of (cell, halo) entries; BF_E_WORKSPACE when
 * it exceeds max_entries (call again with a workspace sized for it); BF_E_CAPACITY when the count reaches 2^25, the
 * limit of the packed cell table (call again with a smaller `ng`: the list shrinks ~ ng^dim).  Synchronises `stream` once. */
int bf_snap_bin_halos(int dim, double L, int64_t ng, const bf_snap_halos *h_host, int64_t max_entries, void *ws,
                      size_t ws_bytes, int64_t *nentries_host, void *stream);
/* Periodic ball query + displacement (SnapshotRunner.py:205-247) as a particle-centric gather over the binned
 * halos, then add + single wrap (:250-260).  x,y,z in / out are separate arrays (z NULL in 2-D). */
int bf_snap_displace(int dim, int64_t np, const double *x, const double *y, const double *z, double L, double a,
                     const bf_snap_halos *h_host, const bf_table *tab_host, int64_t ng, const void *ws, double *xo,
                     double *yo, double *zo, unsigned long long *npairs, void *stream);
/* bf_snap_displace that also reports WHICH pairs it evaluated -- the answer of the KD-tree's query_ball_point
 * (SnapshotRunner.py:219 / :234) turned inside out: pairs[k] = (particle index << 32) | halo index for every (particle,
 * halo) with min-image distance <= R_q, in no particular order (sort to compare); *pair_count (device) = how many there
 * are, of which the first min(count, pair_cap) were stored.  Needs np < 2^32.  For parity tests and diagnostics. */
int bf_snap_pair_list(int dim, int64_t np, const double *x, const double *y, const double *z, double L, double a,
                      const bf_snap_halos *h_host, const bf_table *tab_host, int64_t ng, const void *ws, double *xo,
                      double *yo, double *zo, unsigned long long *pairs, int64_t pair_cap,
                      unsigned long long *pair_count, void *stream);
/* Same on the reference's own particle records: ParticleSnapshot.cat is a structured array of float64
 * (M, x, y, z) (utils/io.py:581-589; z is a filler in 2-D).  cat / cat_out are [np] 32-byte records, 16-byte
 * aligned; cat_out receives M unchanged and the displaced, wrapped coordinates (new_cat, SnapshotRunner.py:250). */
int bf_snap_displace_records(int dim, int64_t np, const void *cat, double L, double a, const bf_snap_halos *h_host,
                             const bf_table *tab_host, int64_t ng, const void *ws, void *cat_out,
                             unsigned long long *npairs, void *stream);

/* ---- ParticleSnapshot.make_map on the device (utils/io.py:622-670) ---- */
/* np.histogramdd(coords, bins = (edges,) * dim, weights = M)[0] with edges = np.linspace(0, L, ngrid + 1) UPLOADED by the
 * caller (device, [ngrid + 1]): bin = searchsorted(edges, x, 'right') - 1, samples on the last edge go to the last bin,
 * samples outside [edges[0], edges[ngrid]] are dropped, and a cell's masses are added in particle order (np.bincount) --
 * stable radix sort by cell + in-order run sums, no atomics: bit-identical to numpy and bit-reproducible.
 * Give x, y, (z,) m or `records` ((M, x, y, z) float64, 16-byte aligned).  grid is [ngrid]^dim float64 (overwritten).
 * Synchronises `stream` once (reads the ends of `edges`).  ngrid^dim < 2^32, np < 2^31. */
size_t bf_make_map_ws_bytes(int64_t np);
int bf_make_map(int dim, int64_t np, const double *x, const double *y, const double *z, const double *m,
                const void *records, const double *edges, int64_t ngrid, double *grid, void *ws, size_t ws_bytes,
                void *stream);

/* ---- Delta P(k) summary (examples/10_Reproduce_Schneider_deltaPk.ipynb, cells 1, 12, 15) ---- */
/* NGP mass assignment of (folded) positions: grid[ix][iy][iz] += 1 with ix = int((x % Lfold) / (Lfold / ngrid))
 * (`numba_histogram3d(Part % Lbox, ...)`).  Give either x, y, z or `records` ((M, x, y, z) float64, 16-byte
 * aligned).  grid is [ngrid]^3 float64, += (zero it first); counts are whole numbers, so the result is exact. */
int bf_pk_ngp_deposit(int64_t np, const double *x, const double *y, const double *z, const void *records, double Lfold,
                      int64_t ngrid, double *grid, void *stream);
/* Spherical average of |F|^2 over `nk` linear bins, F = rfftn(grid) as complex128 [ngrid][ngrid][ngrid/2+1];
 * strides_host = element strides of the three axes (torch.fft returns a permuted dense view; NULL = C order).  val = 1 / (ngrid * d) of np.fft.fftfreq, k0 = kbins[0], dk = kbins[1] - kbins[0].
 * out[0][b] = sum |F|^2, out[1][b] = sum k, out[2][b] = number of modes of bin b, each mirrored mode counted twice
 * (equals np.bincount over the full cube).  Deterministic (no atomics). */
size_t bf_pk_bin_ws_bytes(int64_t ngrid, int64_t nk);
int bf_pk_bin(int64_t ngrid, const void *spectrum, const int64_t *strides_host, double val, double k0, double dk,
              int64_t nk, double *out, void *ws, size_t ws_bytes, void *stream);
/* bf_pk_bin for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * NO counterpart in the reference.  The scratch bf_pk_bin takes as a
 * workspace lives in a file-scope device array of bf_pk_bin_static_max_bytes() = 8 MiB that every GPU loading the library
 * holds until exit; the call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine).  Limits: nk <= 256
 * and bf_pk_bin_ws_bytes(ngrid, nk) <= 8 MiB (true for every ngrid on a GPU of up to 170 SMs) -- BF_E_INVALID /
 * BF_E_WORKSPACE with a message otherwise.  All calls on a device SHARE the array: calls that overlap on the GPU
 * (different streams) race; consecutive calls do not depend on each other (nothing in the array has to start at zero).
 * Same arguments and results as bf_pk_bin (bit-identical on the same device). */
size_t bf_pk_bin_static_max_bytes(void);
int bf_pk_bin_static(int64_t ngrid, const void *spectrum, const int64_t *strides_host, double val, double k0, double dk,
                     int64_t nk, double *out, void *stream);
/* The same over a rectangular PIECE of the spectrum, for the slab-decomposed transform of the N-GPU path (SURVEY.md
 * 8(e), last row): ext_host[3] = logical extents of the piece (axes i, j, l), off_host[3] = global index of its first
 * element, strides_host as above (NULL = C order of the piece).  The [3][nk] sums of all pieces add up to bf_pk_bin's
 * (counts exactly); an empty piece zeroes `out`. */
size_t bf_pk_bin_part_ws_bytes(const int64_t *ext_host, int64_t nk);
int bf_pk_bin_part(int64_t ngrid, const void *spectrum, const int64_t *strides_host, const int64_t *ext_host,
                   const int64_t *off_host, double val, double k0, double dk, int64_t nk, double *out, void *ws,
                   size_t ws_bytes, void *stream);

#ifdef __cplusplus
}
#endif
#endif
