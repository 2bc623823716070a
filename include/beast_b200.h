/*
 * beast_b200 -- C ABI of the B200-native BEAST fitting hot path.
 *
 * This is the drop-in boundary: a plain C interface (pointers + sizes, no torch
 * types) over hand-written sm_100a kernels.  Every entry point is asynchronous
 * on the CUDA stream it is given (pass the raw cudaStream_t as void*; NULL = the
 * legacy default stream), takes caller-owned DEVICE pointers unless a parameter
 * says "host", allocates nothing, and returns 0 on success or a negative
 * bf_status code; bf_last_error() returns a message for the calling thread.
 *
 * The reference has no FFI of its own (it is pure NumPy); each entry point
 * cites the reference code (relative to beast/) whose per-star work it
 * replaces.  The Python side (beast_b200/fit.py) keeps the reference's
 * summary_table_memory / Q_all_memory signatures and calls these through ctypes.
 *
 * Index spaces
 *   "model j"  : row of the weight>0 filtered grid (fit.py:402, g0_indxs), padded
 *                up to a multiple of BF_TILE_MODELS with never-selected rows.
 *   "group"    : BF_GROUP consecutive models; the dense pass reports one minimum
 *                per (star, group).
 *   "chunk"    : 8 consecutive models; the unit of the per-parameter tables.
 */
#ifndef BEAST_B200_H
#define BEAST_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BF_TILE_MODELS 256 /* models per TMA-staged shared-memory tile            */
#define BF_GROUP 64        /* models per group minimum                            */
#define BF_CHUNK 8         /* models per parameter-table chunk                    */
#define BF_STAR_BLOCK 1024 /* stars per thread block of the dense pass            */
/* floats per group record of the pruned dense pass: three boxes (the group's 64 models, its first
 * and its second 32), each mlo[nb], mhi[nb], rlo[nb], c2min; the record padded to 16 B */
#define BF_BOX_FLOATS(nb) ((3 * (3 * (nb) + 1) + 3) / 4 * 4)
#define BF_MAX_BANDS 16   /* array bound; the diagonal likelihood is built for 1..12 bands,
                             the full-covariance one for 2..8, trimming for 1..15 */
#define BF_MAX_PARAMS 128
#define BF_SELECT_SPLIT 16 /* model-axis splits of the group-selection pass       */

typedef enum {
  BF_OK = 0,
  BF_ERR_INVALID = -1,   /* bad argument                                        */
  BF_ERR_CUDA = -2,      /* a CUDA runtime call failed; see bf_last_error()     */
  BF_ERR_UNSUPPORTED = -3,
  BF_ERR_TOO_LARGE = -4
} bf_status;

const char* bf_last_error(void);
int bf_version(void);

/* SM count, shared memory per block (opt-in) and SM clock (kHz) of `device`. */
int bf_device_info(int device, int* sm_count, int* smem_optin_bytes, int* clock_khz);

/* ---------------------------------------------------------------------------
 * Dense pass -- replaces the per-star N_logLikelihood_NM / N_chi2_NM sweep over
 * all models (fitting/fit_metrics/likelihood.py:24-55, 109-167; fit.py:657-668).
 *
 * tiles : [n_tiles][2*nb+1][BF_TILE_MODELS] float.  Row b (<nb) = r_b = s_b/sigma_b,
 *         row nb+b = -a_b = -(sed_b+bias_b)/sigma_b, row 2nb = c2 = -2*(c - c_off)
 *         with c = -lnQ + ln w (star-independent part of lnP).  Padding models
 *         carry c2 = 3e38.  (s_b = per-band scale so that values are O(1).)
 * flux  : [nb][ld_star] float, observed flux / s_b, star-fastest.
 * gmin  : out [n_tiles*BF_TILE_MODELS/BF_GROUP][ld_star] float,
 *         min over the group of  acc = chi2 + c2  (= -2 (lnP - c_off)) in FP32.
 * starmin : in/out [ld_star] uint64, packed (FP32 bits of the minimum << 32 | group that
 *         holds it), atomically lowered over all models; caller initialises every entry to
 *         0x7f800000ffffffff (+inf, no group).  The arg-min group lets the caller order
 *         stars so that neighbours touch the same part of the grid.
 * ld_star must be a multiple of BF_STAR_BLOCK and >= n_stars.
 * ------------------------------------------------------------------------- */
int bf_dense_chi2(int nb, int64_t n_tiles, const float* tiles, int64_t n_stars,
                  int64_t ld_star, const float* flux, float* gmin, uint64_t* starmin,
                  int tiles_per_item, void* stream);

/* The same sweep, leaving out what cannot matter.  A warp of the kernel owns 256 neighbouring
 * stars; it skips a group when a lower bound of acc over the group's models is above the
 * selection threshold of every one of them, and reports +inf for that group; otherwise it
 * evaluates, for each 64 of its stars, the halves of the group their bounds do not rule out.
 * The bounds come from boxes per group,
 *   boxes : [n_tiles*BF_TILE_MODELS/BF_GROUP][BF_BOX_FLOATS(nb)] float, 16-byte aligned; per group
 *           three boxes of 3 nb + 1 floats (all 64 models, the first 32, the last 32):
 *           mlo_b <= a_b / r_b <= mhi_b and rlo_b <= r_b for every model of the box and band b,
 *           c2min <= c2  (rounded outwards), so that
 *           acc >= sum_b rlo_b^2 dist(f_b, [mlo_b, mhi_b])^2 + c2min;
 *   cut   : [ld_star] float, per star a ceiling of its selection threshold on the acc scale
 *           (bf_select_count's tau and error terms applied to any upper bound of the star's
 *           minimum, e.g. the minimum over the coarse tiles); -1 for padding stars, +inf to
 *           keep everything.
 * Groups that bf_select_* would list get the same minimum as from bf_dense_chi2, the others
 * +inf, their true minimum or the minimum over the half evaluated (all at or above the star's
 * ceiling), and starmin is the same, so the lists are identical.  Pays off when neighbouring stars are alike (bf_fit_batch orders them).
 *   evaluated : optional device counter (may be NULL), raised by the number of (star, model)
 *           pairs the sweep did evaluate (padding stars and models included)
 * No counterpart in the reference, which evaluates every model for every star. */
int bf_dense_chi2_pruned(int nb, int64_t n_tiles, const float* tiles, const float* boxes,
                         int64_t n_stars, int64_t ld_star, const float* flux, const float* cut,
                         float* gmin, uint64_t* starmin, int tiles_per_item, uint64_t* evaluated,
                         void* stream);

/* Same sweep for the full-covariance likelihood (N_covar_logLikelihood / N_covar_chi2,
 * fitting/fit_metrics/likelihood.py:58-106, 170-218; selected at fit.py:432-444).
 * tiles : [n_tiles][2*nb + nb(nb-1)/2 + 1][BF_TILE_MODELS] float.  Rows: -m'_b = -(sed_b+bias_b)/s_b,
 *         D_b = icov_diag_b*s_b^2, O_kl = 2*icov_offdiag_kl*s_k*s_l (k<l packed row-major, the order
 *         of observationmodel/noisemodel/trunchen.py:394-400), c2 = -2*(c - c_off) with
 *         c = -nb/2 ln 2pi + q_norm + ln w.  Everything else as bf_dense_chi2.  nb = 2..8. */
int bf_dense_chi2_cov(int nb, int64_t n_tiles, const float* tiles, int64_t n_stars,
                      int64_t ld_star, const float* flux, float* gmin, uint64_t* starmin,
                      int tiles_per_item, void* stream);

/* Suggested tiles_per_item for a device with `sm_count` SMs. */
int bf_dense_tiles_per_item(int64_t n_tiles, int64_t n_stars, int sm_count);

/* ---------------------------------------------------------------------------
 * Group selection -- the conservative superset of fit.py:670
 *   indx = where(lnp - max(lnp) > threshold)
 * at group granularity: group g is kept for star s when
 *   gmin[g][s] < x + 2*(err_a*sqrt(x) + err_b*x),   x = starmin[s] + tau,  tau = -2*threshold.
 * err_a, err_b bound the FP32 rounding error of the dense pass for models with chi2 <= x
 * (err_a = k*6u*sqrt(nb)*max|a|, err_b = k*16u, u = 2^-24; see DESIGN.md "margins"), so every
 * member of the exact sparse set lies in a kept group.
 * counts : out [n_stars][BF_SELECT_SPLIT] uint32 (bf_select_count)
 * offsets: in  [n_stars*BF_SELECT_SPLIT + 1] uint32 exclusive scan of counts
 * list   : out group ids, ascending per star (bf_select_fill)
 * entry_star : optional out, star index of every list position (same length as list)
 * ------------------------------------------------------------------------- */
int bf_select_count(int64_t n_groups, int64_t n_stars, int64_t ld_star, const float* gmin,
                    const uint64_t* starmin, float tau, float err_a, float err_b, uint32_t* counts,
                    void* stream);
int bf_select_fill(int64_t n_groups, int64_t n_stars, int64_t ld_star, const float* gmin,
                   const uint64_t* starmin, float tau, float err_a, float err_b,
                   const uint32_t* offsets, uint32_t* list, uint32_t* entry_star, void* stream);

/* Exclusive prefix sum of n uint32 (out[n] = total).  scratch: >= bf_scan_scratch_elems(n)
 * uint32.  Plumbing for the variable-length per-star lists. */
int64_t bf_scan_scratch_elems(int64_t n);
int bf_exclusive_scan_u32(int64_t n, const uint32_t* in, uint32_t* out, uint32_t* scratch,
                          void* stream);

/* Same prefix sum for callers that cannot supply scratch.  (No counterpart in the reference;
 * this exists for hosts of the library, not for BEAST.)
 *
 * Who needs it: an application that has handed all of a GPU's memory to its own arena and
 * treats any further cudaMalloc as an error, or a plugin slot whose contract forbids allocating
 * (and the implicit device-wide synchronisation that cudaMalloc/cudaFree bring) once the
 * pipeline is running.  Such a caller still owns `in` and `out`, but has nowhere to put the
 * tile sums, so they live in a fixed array inside the library image instead.
 *
 * Limits and rules:
 *  - n <= 65536 * 2048 = 134,217,728 elements; larger inputs are refused with BF_ERR_TOO_LARGE
 *    (use bf_exclusive_scan_u32).  The array costs 256 KiB on every GPU that loads the library,
 *    for the life of the process.
 *  - The first call on each GPU looks up the array's device address (with that GPU current);
 *    make that first call before starting any stream capture.  The GPU that owns `stream` must
 *    be the current one on every call (checked where the driver can tell); the address is
 *    remembered per GPU for the life of the process, so do not call this again on a GPU after
 *    cudaDeviceReset.
 *  - All calls on one GPU share the array.  Its previous contents never matter (each call
 *    overwrites what it later reads), but two calls that overlap on the GPU -- different streams,
 *    or different host threads -- race on it: serialise them.
 *  - Nothing is allocated or freed, on host or device. */
int bf_exclusive_scan_u32_noalloc(int64_t n, const uint32_t* in, uint32_t* out, void* stream);

/* ---------------------------------------------------------------------------
 * Exact (FP64) grid, band-major, ld_model = padded model count:
 *   msb  [nb][ld_model]  seds+bias            (fit.py:463)
 *   ivar [nb][ld_model]  1/error^2            (fit.py:444)
 *   lnq  [ld_model]      n/2 ln 2pi - 1/2 sum ln ivar   (likelihood.py:149-159)
 *   lnw  [ld_model]      ln weight [- ln sum weight]    (fit.py:411-415)
 * lnP = (-lnq - 0.5*chi2) + lnw, evaluated in the reference's operation order.
 * ------------------------------------------------------------------------- */
typedef struct {
  int nb;
  int64_t n_models; /* real (unpadded) filtered models */
  int64_t ld_model;
  const double* msb;
  const double* ivar;
  const double* lnq;
  const double* lnw;
  /* full covariance (optional, NULL = diagonal noise model): [nb(nb-1)/2][ld_model]
   * 2*icov_offdiag packed k<l row-major; then `ivar` holds icov_diag and `lnq` holds
   * nb/2 ln 2pi - q_norm, so that lnP = (-lnq - 0.5*chi2) + lnw in both cases
   * (likelihood.py:212-216). */
  const double* cov_off;
} bf_exact_grid;

/* Exact per-star maximum of lnP and its first index (fit.py:670, 684, 719, 727):
 * evaluates only groups whose FP32 minimum is within the error bound of the star's minimum
 * (the exact argmax cannot be anywhere else).
 * flux64 [n_stars][nb] double.  pmax [n_stars] double, pmax_j [n_stars] int64 (filtered index,
 * -1 when the star has no finite lnP). */
int bf_exact_max(const bf_exact_grid* grid, int64_t n_stars, int64_t ld_star,
                 const double* flux64, const float* gmin, const uint64_t* starmin, float err_a,
                 float err_b, const uint32_t* offsets, const uint32_t* list, double* pmax,
                 int64_t* pmax_j, void* stream);

/* ---------------------------------------------------------------------------
 * Per-parameter tables for the marginal PDFs / expectations, chunk-major:
 *   bin_tab [ld_model/8][n_q][8] uint8   0-based pdf1d bin of each model
 *                                        (np.digitize result, pdf1d.py:90)
 *   q_tab   [ld_model/8][n_q][8] int32   parameter value of each model (fit.py:733-737) in
 *                                        32-bit fixed point: q = q_lin[2k] + q_lin[2k+1] * fix,
 *                                        i.e. a rounding of 2^-33 of the parameter's range;
 *                                        sums are accumulated in FP64
 *   q_lin   [n_q][2] double              offset and scale of that encoding
 *   bin_off [n_q+1] int32                start of parameter k's bins in a star's pdf row
 *   bin_val [bin_off[n_q]] double        bin centres (pdf1d.bin_vals)
 *   q_const [n_q] uint8, may be NULL     1: all models of a bin of parameter k share one value
 *                                        (grid axes, unique-value bins).  Its expectation is then
 *                                        q_lin[2k] + sum_b pdf_b * bin_q_b, taken from the
 *                                        histogram, and q_tab is not read for it (the same
 *                                        sum, regrouped by bin)
 *   bin_q   [bin_off[n_q]] double        that shared value minus q_lin[2k], per bin (read
 *                                        where q_const is 1)
 * ------------------------------------------------------------------------- */
typedef struct {
  int n_q;
  int total_bins;
  int max_bins; /* largest bin count of any parameter (<= 256) */
  const uint8_t* bin_tab;
  const int32_t* q_tab;
  const double* q_lin;
  const int32_t* bin_off;
  const double* bin_val;
  const uint8_t* q_const;
  const double* bin_q;
  /* Class split for the split histogram kernel (optional: n_fast = n_slow = 0 when absent).
   * "slow" parameters have one value per chunk (everything that depends on the stellar model
   * only, when the dust axes run innermost), so they take one histogram update and one
   * expectation term per chunk, with the chunk's weight sum; "fast" parameters take one per
   * model.  n_fast + n_slow = n_q, each <= 32; fast_pad / slow_pad = 16 or 32 table lanes.
   *   fast_k   [n_fast] int32               parameter index of fast slot i (slow_k alike)
   *   fast_bin [ld_model/8][fast_pad][8] u8 bins of the chunk's models
   *   fast_q   [ld_model/8][2][fast_pad][4] int32   fixed-point values (q_tab encoding), models
   *                                         0..3 of all parameters, then models 4..7
   *   slow_bin [ld_model/8][slow_pad] u8,  slow_q [ld_model/8][slow_pad] int32
   *   slow_run [ld_model/64] u8             bit c: chunk c of the group starts a run of chunks
   *                                         with equal slow values (bit 0 always set)
   *   fast_rng [ld_model/8][fast_pad] u16   optional (may be NULL): smallest bin | largest bin << 8
   *                                         of the chunk's 8 models, what the histogram kernel
   *                                         sizes its private copies from */
  int n_fast, n_slow, fast_pad, slow_pad;
  const int32_t* fast_k;
  const int32_t* slow_k;
  const uint8_t* fast_bin;
  const int32_t* fast_q;
  const uint8_t* slow_bin;
  const int32_t* slow_q;
  const uint8_t* slow_run;
  const uint16_t* fast_rng;
} bf_param_tables;

typedef struct {
  double* total_log_norm; /* [n_stars]            fit.py:716                       */
  double* chi2min;        /* [n_stars]            fit.py:725                       */
  int64_t* chi2min_j;     /* [n_stars] filtered   fit.py:726                       */
  uint32_t* n_sparse;     /* [n_stars]  size of the sparse set                     */
  double* exp_vals;       /* [n_stars][n_q]       fit.py:743, common.py:126-167    */
  double* pdf1d;          /* [n_stars][total_bins] fit.py:746-748, pdf1d.py:119-148*/
  double* per_vals;       /* [n_stars][n_q][3]    fit.py:749-755, common.py:48-123 */
  uint32_t* member_mask;  /* [n_list][BF_GROUP/32] membership bits per listed group */
  uint32_t* member_prefix;/* [n_list] members of the star before this group        */
} bf_star_outputs;

/* Scratch of the sparse pass, one slot per listed group of the stars being processed
 * (slot = list position - entry_base). */
typedef struct {
  double* weights;      /* [n_entries][BF_GROUP] exp(lnP - Pmax) of members, 0 otherwise */
  double* part_wsum;    /* [n_entries] sum of the group's weights                        */
  double* part_chi2;    /* [n_entries] smallest member chi2 of the group                 */
  int64_t* part_chi2_j; /* [n_entries] its model index (-1: no member)                   */
  int32_t* star_flags;  /* [n_stars] or NULL.  Set by the histogram kernel: 1 = it has also
                           written this star's expectations (warps without a private
                           histogram did the sums), so the expectation kernel skips the star */
  uint32_t* group_order;/* [n_entries] or NULL: the list positions of this range ordered by group
                           id (bf_group_order).  When given (diagonal noise model, <= 12 bands)
                           the exact weights are evaluated group by group, so that a group's
                           model rows are fetched once for all the stars that list it */
  uint32_t* chunk_refs; /* [n_entries][16][2] or NULL: room for the per-star lists of chunks
                           that hold a member and of slow-value runs that hold one, built and
                           consumed by the split histogram kernel */
  double* chunk_wsum;   /* [n_entries][8] or NULL: at the first chunk of every run of equal slow
                           values (params->slow_run) the weight sum of the run, 0 elsewhere;
                           written by the group-ordered weights kernel */
  double* exp_part;     /* [n_entries][fast_pad] or NULL: sum_m w_m * fix_m[k] of each listed
                           group for the fast parameters (FP64 tensor-core products in the
                           group-ordered weights kernel).  The split histogram kernel runs when
                           group_order, chunk_refs, chunk_wsum and exp_part are all given and
                           `params` carries the class split */
} bf_accum_scratch;

/* Order the list positions [entry_base, entry_base + n_entries) by group id (counting sort
 * with device atomics; the order inside a group is not defined and does not matter).
 * work : >= bf_group_order_work_elems(n_groups) uint32 of scratch.  group_order: out [n_entries]. */
int64_t bf_group_order_work_elems(int64_t n_groups);
int bf_group_order(int64_t n_groups, const uint32_t* list, int64_t entry_base, int64_t n_entries,
                   uint32_t* work, uint32_t* group_order, void* stream);

/* Sparse pass: exact sparse set, weights, normalisation, chi2min, expectations, 1D PDFs
 * and p16/p50/p84 (fit.py:670-755) for stars [star_begin, star_begin + n_stars) of the
 * batch-wide arrays passed in.
 * Three launches: (1) one warp per listed group evaluates the exact lnP of its 64 models and
 * stores membership bits, weights and per-group partial sums; (2) one block per star builds the
 * histograms and percentiles -- with full-width private histograms per warp, or, for wide bin
 * tables (200 bins), with histograms sized by the star's occupied bin ranges, in which case
 * warps left without a histogram also sum the star's expectations; (3) one block per star sums
 * w*q into the expectations of the stars and parameters not yet done.  Results do not depend
 * on which histogram kernel runs beyond the summation order (tests/test_gpu_fullsize.py).
 * entry_star : [list length] star index of every list position (from bf_select_fill).
 * star_order : optional [n_stars] int32 permutation of 0..n_stars-1; block i of the histogram
 *              kernel works on star star_begin + star_order[i] (callers put the stars with the
 *              longest lists first, so that the last blocks of the launch are short).
 * entry_base / n_entries : list range of those stars,
 *              [offsets[star_begin*SPLIT], offsets[(star_begin+n_stars)*SPLIT]), covered by `scratch`.
 * percentiles_host : [3] host doubles in percent.
 * aux_stream : optional second stream; when given and the full-width histogram kernel is used,
 *              the expectation kernel runs on it next to the histogram kernel (ordered by
 *              events; `stream` has joined it on return).  The two events that order the streams
 *              are created on the first such call and kept per GPU: one caller at a time per GPU
 *              may use aux_stream (callers on different GPUs are independent).
 * The first call on a GPU also configures the kernels' shared-memory limits (cudaFuncSetAttribute). */
int bf_exact_accumulate(const bf_exact_grid* grid, const bf_param_tables* params,
                        int64_t star_begin, int64_t n_stars, const double* flux64, const double* pmax,
                        double threshold, const uint32_t* offsets, const uint32_t* list,
                        const uint32_t* entry_star, const int32_t* star_order,
                        int64_t entry_base, int64_t n_entries,
                        const double* percentiles_host, const bf_accum_scratch* scratch,
                        const bf_star_outputs* out, void* stream, void* aux_stream);

/* Sparse lnP record (fit.py:693-709): for star s, entries [rec_off[s], rec_off[s+1]) of
 * `ranks` hold 0-based positions inside the star's ascending sparse set; outputs the filtered
 * model index, lnP and chi2 of each (float32 like the reference stores them). */
int bf_sparse_gather(const bf_exact_grid* grid, int64_t n_stars, const double* flux64,
                     const uint32_t* offsets, const uint32_t* list, const uint32_t* member_mask,
                     const uint32_t* member_prefix, const int64_t* rec_off, const uint32_t* ranks,
                     int64_t* out_j, float* out_lnp, float* out_chi2, void* stream);

/* Device-side stand-in for np.random.choice(indx, lnp_npts, replace=False) (fit.py:694-700):
 * star s gets rec_off[s+1]-rec_off[s] = min(n_sparse[s], lnp_npts) distinct pseudo-random ranks
 * (identity when the whole set is kept).  Same distribution as the reference, different
 * random stream; the host path with NumPy's stream stays available in engine.py. */
int bf_sparse_sample_ranks(int64_t n_stars, const uint32_t* n_sparse, const int64_t* rec_off,
                           uint64_t seed, int64_t star_base, uint32_t* ranks, void* stream);

/* ---------------------------------------------------------------------------
 * 2D PDFs (fitting/pdf2d.py:122-148, fit.py:758-762).
 *   pair_bin [n_pairs][ld_model] int32  flattened (b1*nb2+b2) bin, -1 = outside
 *   pair_off [n_pairs+1] int64          start of pair p inside a star's pdf2d row
 *   pdf2d    [n_stars][pair_off[n_pairs]] double, must be zero on entry
 * ------------------------------------------------------------------------- */
int bf_pdf2d_accumulate(const bf_exact_grid* grid, int n_pairs, const int32_t* pair_bin,
                        const int64_t* pair_off, int64_t n_stars, const double* flux64,
                        const double* pmax, const double* total_log_norm, double threshold,
                        const uint32_t* offsets, const uint32_t* list, double* pdf2d,
                        void* stream);

/* The same 2D PDFs from what bf_exact_accumulate has just left behind for the list range
 * [entry_base, entry_base + n_entries): the members' weights exp(lnP - Pmax) in
 * scratch->weights, the membership bits, and total_log_norm (1 / sum of weights =
 * exp(Pmax - total_log_norm), fit.py:716).  No model is evaluated again, so this is the call
 * to use right after each bf_exact_accumulate; bf_pdf2d_accumulate needs neither scratch nor
 * masks.  pdf2d rows of the stars concerned must be zero before their first range. */
int bf_pdf2d_from_weights(int64_t ld_model, int n_pairs, const int32_t* pair_bin,
                          const int64_t* pair_off, const double* pmax, const double* total_log_norm,
                          const uint32_t* list, const uint32_t* entry_star,
                          const uint32_t* member_mask, const double* weights, int64_t entry_base,
                          int64_t n_entries, double* pdf2d, void* stream);

/* ---------------------------------------------------------------------------
 * The whole fit of one batch of stars behind one call: the star loop of Q_all_memory
 * (fitting/fit.py:637-762) -- likelihood of every model, prior, sparse selection, normalisation,
 * best fit, chi2 minimum, expectations, 1D PDFs, percentiles, optional 2D PDFs and sparse lnP
 * records -- for n_stars stars against the grid described by `plan`.  It runs the entry points
 * above in sequence (plus the ordering, layout and copy kernels between them) inside a
 * caller-provided workspace; nothing is allocated.  The host synchronises the stream once (the
 * list sizes decide the memory layout), twice with sparse_mode 1.
 *
 * plan    : the resident grid (device pointers; the struct itself is host memory)
 *   tiles / tiles_coarse : see bf_dense_chi2; tiles_coarse (a subset of the models, e.g. every
 *                   64th, in the same layout; may be NULL) is swept first: its minima are the
 *                   ceilings of the pruned sweep, and the stars are put in an order in which
 *                   neighbours are alike (sort_band) -- results come back in the caller's order
 *   inv_band_scale [nb] : 1 / s_b (device), turns flux64 into the FP32 fluxes of the dense pass
 *   boxes         : see bf_dense_chi2_pruned; used when tiles_coarse is given too (its minima
 *                   bound the selection thresholds), diagonal likelihood only
 *   sort_band     : with tiles_coarse, the stars are fitted in the order of their flux in this
 *                   band (the one that tells stars apart best: neighbours then rule out the same
 *                   groups); -1 orders them by the group of their coarse minimum instead
 *   pair_bin / pair_off / pair_len : 2D PDF tables (bf_pdf2d_accumulate), pair_len = pair_off[n_pairs]
 *   force_legacy  : 1 = do not use the class-split kernels even if `params` carries the split
 *   max_range_entries : the sparse pass works on ranges of stars whose lists fit the scratch;
 *                   this caps a range further (0 = no cap; a star's own list is never cut)
 * flux64  : [n_stars][nb] observed fluxes, device memory, or host memory with flux_on_host = 1
 * percentiles_host : [3] host doubles, in percent
 * sparse_mode : 0 no sparse records; 1 all members of every star's sparse set (ascending model
 *               index); 2 min(|S|, lnp_npts) members drawn by the keyed permutation of
 *               bf_sparse_sample_ranks with (seed, star_base + star index) -- fit.py:693-709
 * out     : device pointers, one row per star in the caller's order; pdf2d, n_groups_selected and
 *           the sparse_* members may be NULL when not wanted.  sparse_off [n_stars + 1] gives the
 *           record range of every star inside sparse_j / sparse_lnp / sparse_chi2
 *           (sparse_capacity records of room; mode 2 needs n_stars * lnp_npts)
 * state   : optional out; what a follow-up call needs (bf_fit_sparse_records), valid until the
 *           workspace is written again
 * Returns BF_ERR_TOO_LARGE when the workspace (or sparse_capacity) is too small;
 * bf_fit_workspace_bytes gives a size from the expected total of listed groups.
 * ------------------------------------------------------------------------- */
typedef struct {
  int nb;
  int full_cov;
  int64_t n_groups;
  int64_t n_tiles;
  const float* tiles;
  int64_t n_tiles_coarse;
  const float* tiles_coarse;
  const double* inv_band_scale;
  float err_a, err_b;
  bf_exact_grid grid;
  bf_param_tables params;
  int n_pairs;
  const int32_t* pair_bin;
  const int64_t* pair_off;
  int64_t pair_len;
  int sm_count;
  int force_legacy;
  int64_t max_range_entries; /* 0, or a cap on the listed groups the sparse pass takes at a time */
  const float* boxes;        /* group boxes for bf_dense_chi2_pruned, or NULL: sweep everything */
  int sort_band;             /* order the stars by their flux in this band; -1: by coarse arg-min group */
} bf_fit_plan;

typedef struct {
  double* pmax;             /* fit.py:727 */
  int64_t* pmax_j;          /* fit.py:719 (filtered model index) */
  double* total_log_norm;
  double* chi2min;
  int64_t* chi2min_j;
  uint32_t* n_sparse;
  int64_t* n_groups_selected;
  double* exp_vals;         /* [n_stars][n_q] */
  double* pdf1d;            /* [n_stars][total_bins] */
  double* per_vals;         /* [n_stars][n_q][3] */
  double* pdf2d;            /* [n_stars][pair_len] or NULL */
  int64_t* sparse_off;
  int64_t* sparse_j;
  float* sparse_lnp;
  float* sparse_chi2;
  int64_t sparse_capacity;
} bf_fit_outputs;

typedef struct {
  const double* flux64;           /* fluxes in the order the batch was fitted in */
  const uint32_t* offsets;
  const uint32_t* list;
  const uint32_t* member_mask;
  const uint32_t* member_prefix;
  const int32_t* perm;            /* fit order -> caller's order, NULL = the same */
  const uint32_t* n_sparse_sorted;
  int64_t n_list;                 /* listed groups of the batch */
  int64_t n_records;              /* sparse records wanted when sparse_capacity was too small */
  int64_t launches;               /* kernels launched by the call */
  int split;                      /* 1: the class-split kernels ran */
  int n_ranges;                   /* ranges of stars the sparse pass was cut into */
  int64_t dense_evaluated;        /* (star, model) pairs the dense sweep evaluated; n_stars padded to
                                     BF_STAR_BLOCK times the padded model count when nothing was skipped */
  int want_timing;                /* in: 1 = fill stage_ms (the call then synchronises the stream
                                     after every stage; for measurements) */
  float stage_ms[12];             /* device time per stage: presort, dense, select, exact_max,
                                     group_order, exact_weights, histograms, expectations, pdf2d,
                                     reorder, sparse_records */
} bf_fit_state;

int64_t bf_fit_workspace_bytes(const bf_fit_plan* plan, int64_t n_stars, int64_t list_entries_hint);
int bf_fit_batch(const bf_fit_plan* plan, int64_t n_stars, const double* flux64, int flux_on_host,
                 double threshold, const double* percentiles_host, int sparse_mode, int64_t lnp_npts,
                 uint64_t seed, int64_t star_base, void* workspace, size_t workspace_bytes,
                 const bf_fit_outputs* out, bf_fit_state* state, void* stream, void* aux_stream);
/* Sparse records for ranks drawn by the caller (np.random.choice on the host, fit.py:696) after a
 * bf_fit_batch with sparse_mode 0: rec_off [n_stars + 1] and ranks in the caller's star order;
 * ranks == NULL takes positions 0 .. rec_off[s+1]-rec_off[s]-1 of every star's sparse set. */
int bf_fit_sparse_records(const bf_fit_plan* plan, const bf_fit_state* state, int64_t n_stars,
                          const int64_t* rec_off, const uint32_t* ranks, int64_t* out_j,
                          float* out_lnp, float* out_chi2, void* stream);

/* ---------------------------------------------------------------------------
 * Subgrid merge (tools/subgridding_tools.py:392-527), device part.  Each rank holds the
 * results of its own subgrid for the same stars; after the caller all-gathers
 * total_log_norm / Pmax into [n_sub][n_stars] arrays these kernels produce the merged rows.
 * ------------------------------------------------------------------------- */
/* weight_g = exp(tln_g - max_g tln_g); merged_tln = ln sum_g weight_g + max (:394-406, 513) */
int bf_merge_weights(int n_sub, int64_t n_stars, const double* tln_all, double* weight_all,
                     double* merged_tln, void* stream);
/* in place: rows[s][:] *= weight[s]   (the scaling that feeds the cross-rank sum, :421-426) */
int bf_scale_rows(int64_t n_stars, int64_t row_len, const double* weight, double* rows,
                  void* stream);
/* dst[s][:] = (accumulate ? dst[s][:] : 0) + weight[s] * src[s][:]: sums the weighted rows of
 * the subgrids one process holds before the cross-rank reduce-scatter (:421-426, :462-470) */
int bf_axpy_rows(int64_t n_stars, int64_t row_len, const double* weight, const double* src,
                 double* dst, int accumulate, void* stream);
/* after the cross-rank sum: normalise each parameter's pdf to unit sum where positive
 * (:428-434) and recompute the percentiles from the merged pdf (:472-491). */
int bf_merge_finalize(const bf_param_tables* params, int64_t n_stars, double* pdf1d,
                      const double* percentiles_host, double* per_vals, void* stream);

/* ---------------------------------------------------------------------------
 * Grid trimming, the step before the fit (fitting/trim_grid.py:53-118).
 * Inputs are the reference's own arrays, [n_models][nb] row-major doubles.
 * bf_trim_flags : pattern[m] bit b = model m survives band b's test
 *                   sed+bias + sigma_fac*error >= min_data[b]  and  sed+bias - sigma_fac*error <= max_data[b]
 *                 (:98-109), bit 31 = completeness > 0 in every band (:84-88);
 *                 hist[(pattern & 0xffff) | complete << nb] += 1, hist has 2^(nb+1) zeroed entries.
 *                 min_data / max_data are DEVICE arrays of nb doubles.
 * The caller resolves the reference's "skip a band that would leave nothing" rule (:110-111)
 * from the histogram, then
 * bf_trim_select : keep[m] = (pattern[m] & required) == required
 * bf_trim_compact: indices[offsets[m]] = m for kept models (offsets = exclusive scan of keep),
 *                  i.e. the reference's ascending `indxs`.
 * ------------------------------------------------------------------------- */
int bf_trim_flags(int nb, int64_t n_models, const double* seds, const double* bias,
                  const double* unc, const double* completeness, double sigma_fac,
                  const double* min_data, const double* max_data, uint32_t* pattern,
                  uint32_t* hist, void* stream);
int bf_trim_select(int64_t n_models, const uint32_t* pattern, uint32_t required, uint32_t* keep,
                   void* stream);
int bf_trim_compact(int64_t n_models, const uint32_t* keep, const uint32_t* offsets,
                    int64_t* indices, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* BEAST_B200_H */
