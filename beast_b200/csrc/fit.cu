// bf_fit_batch: the whole star loop of the reference for one batch of stars
// (beast/fitting/fit.py:637-762) behind one C call.
//
//   presort (coarse sweep, stars ordered by arg-min group)  ->  dense FP32 sweep  ->  group
//   selection + prefix sums  ->  exact FP64 maximum  ->  per range of stars that fits the scratch:
//   group order, exact weights, histograms / percentiles / expectations [, 2D PDFs]  ->  sparse
//   lnP records  ->  results put back in input order.
//
// Everything lives in a caller-provided workspace (no allocation here).  The group-minimum map
// is taken from the top of the workspace and released after the exact maximum; everything else
// grows from the bottom, so that the per-range scratch can reuse the map's memory.
// Two host synchronisations per call: the list sizes after the prefix sum, and (sparse_mode
// "all" only) the number of sparse records.

#include <cub/device/device_radix_sort.cuh>

#include <cstdlib>
#include <vector>

#include "accum_common.cuh"

namespace bf {

// sparse.cu
int launch_sparse_records(const bf_exact_grid& grid, int64_t n_stars, const double* flux64,
                          const uint32_t* offsets, const uint32_t* list, const uint32_t* member_mask,
                          const uint32_t* member_prefix, const uint32_t* n_sparse_sorted,
                          const int32_t* perm, const int64_t* out_off, const uint32_t* ranks,
                          int64_t lnp_npts, uint64_t seed, int64_t star_base, int64_t* out_j,
                          float* out_lnp, float* out_chi2, cudaStream_t st);

namespace {

struct Arena {
  char* base;
  size_t lo, hi;  // free space is [lo, hi)
  bool ok = true;
  template <typename T>
  T* take(size_t n) {
    const size_t bytes = (n * sizeof(T) + 255) & ~static_cast<size_t>(255);
    if (lo + bytes > hi) {
      ok = false;
      return nullptr;
    }
    T* p = reinterpret_cast<T*>(base + lo);
    lo += bytes;
    return p;
  }
  template <typename T>
  T* take_top(size_t n) {
    const size_t bytes = (n * sizeof(T) + 255) & ~static_cast<size_t>(255);
    if (lo + bytes > hi) {
      ok = false;
      return nullptr;
    }
    hi -= bytes;
    return reinterpret_cast<T*>(base + hi);
  }
};

// device time of the stages of a call, for state->want_timing (synchronises after every stage)
struct StageTimer {
  bool on;
  cudaStream_t st;
  float* ms;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  void begin() {
    if (!on) return;
    if (!e0) {
      cudaEventCreate(&e0);
      cudaEventCreate(&e1);
    }
    cudaEventRecord(e0, st);
  }
  void end(int stage) {
    if (!on) return;
    cudaEventRecord(e1, st);
    cudaEventSynchronize(e1);
    float t = 0.f;
    cudaEventElapsedTime(&t, e0, e1);
    ms[stage] += t;
  }
  ~StageTimer() {
    if (e0) {
      cudaEventDestroy(e0);
      cudaEventDestroy(e1);
    }
  }
};

constexpr int kT = 256;
inline unsigned blocks_for(int64_t n) { return static_cast<unsigned>((n + kT - 1) / kT); }

// flux32[b][s] = float(flux64[s][b] / s_b); columns s >= n stay zero
__global__ void flux_prepare_kernel(int nb, int64_t n, int64_t ld, const double* __restrict__ flux64,
                                    const double* __restrict__ inv_scale, float* __restrict__ flux32) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= ld * nb) return;
  const int b = static_cast<int>(i / ld);
  const int64_t s = i % ld;
  flux32[i] = s < n ? static_cast<float>(flux64[s * nb + b] * inv_scale[b]) : 0.0f;
}

__global__ void starmin_init_kernel(int64_t ld, unsigned long long* starmin) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < ld) starmin[i] = 0x7f800000ffffffffULL;  // (+inf, no group)
}

__global__ void argmin_keys_kernel(int64_t n, const unsigned long long* __restrict__ starmin,
                                   uint32_t* keys, int32_t* vals) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  keys[i] = static_cast<uint32_t>(starmin[i]);
  vals[i] = static_cast<int32_t>(i);
}

// Ceiling of every star's selection threshold, in sorted order, from the coarse sweep's minima
// (a minimum over some of the models bounds the minimum over all of them from above).
__global__ void prune_cut_kernel(int64_t n, int64_t ld, const int32_t* __restrict__ perm,
                                 const unsigned long long* __restrict__ starmin, float tau, float err_a,
                                 float err_b, float* __restrict__ cut) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= ld) return;
  cut[i] = i < n ? select_threshold(starmin_value(starmin[perm[i]]), tau, err_a, err_b) : -1.0f;
}

// sort key = the star's flux in one band (IEEE bits made monotone; NaN last)
__global__ void flux_keys_kernel(int64_t n, int nb, int band, const double* __restrict__ flux,
                                 uint32_t* keys, int32_t* vals) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float v = static_cast<float>(flux[i * nb + band]);
  const uint32_t u = __float_as_uint(v);
  keys[i] = v != v ? 0xffffffffu : ((u & 0x80000000u) ? ~u : (u | 0x80000000u));
  vals[i] = static_cast<int32_t>(i);
}

__global__ void iota_kernel(int64_t n, int32_t* vals) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) vals[i] = static_cast<int32_t>(i);
}

// dst[i][:] = src[perm[i]][:]   (rows of `row` doubles)
__global__ void gather_rows_kernel(int64_t n, int row, const int32_t* __restrict__ perm,
                                   const double* __restrict__ src, double* __restrict__ dst) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n * row) return;
  dst[i] = src[static_cast<int64_t>(perm[i / row]) * row + i % row];
}

// dst[perm[i]][:] = src[i][:]   (rows of `row` words of W bytes; perm == nullptr: identity)
template <typename W>
__global__ void scatter_rows_kernel(int64_t n, int64_t row, const int32_t* __restrict__ perm,
                                    const W* __restrict__ src, W* __restrict__ dst) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n * row) return;
  const int64_t r = i / row;
  dst[(perm ? static_cast<int64_t>(perm[r]) : r) * row + i % row] = src[i];
}

// per-star list offsets (every BF_SELECT_SPLIT-th entry of the split offsets) and list lengths
__global__ void star_offsets_kernel(int64_t n, const uint32_t* __restrict__ offsets,
                                    uint32_t* __restrict__ off_star, int64_t* __restrict__ n_listed) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i > n) return;
  const uint32_t o = offsets[i * BF_SELECT_SPLIT];
  off_star[i] = o;
  if (i < n) n_listed[i] = static_cast<int64_t>(offsets[(i + 1) * BF_SELECT_SPLIT]) - o;
}

__global__ void listed_keys_kernel(int64_t n, const int64_t* __restrict__ n_listed, uint32_t* keys) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) keys[i] = static_cast<uint32_t>(n_listed[i]);
}

// number of sparse records of every star, in input order
__global__ void record_counts_kernel(int64_t n, const uint32_t* __restrict__ n_sparse, int64_t npts,
                                     uint32_t* __restrict__ counts) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t m = n_sparse[i];
  counts[i] = (npts > 0 && static_cast<int64_t>(m) > npts) ? static_cast<uint32_t>(npts) : m;
}

__global__ void widen_kernel(int64_t n, const uint32_t* __restrict__ src, int64_t* __restrict__ dst) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src[i];
}

size_t sort_temp_bytes(int64_t n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, static_cast<const uint32_t*>(nullptr),
                                  static_cast<uint32_t*>(nullptr), static_cast<const int32_t*>(nullptr),
                                  static_cast<int32_t*>(nullptr), static_cast<int>(n));
  return bytes + 256;
}

int64_t pad_to(int64_t n, int64_t m) { return (n + m - 1) / m * m; }

// bytes the per-range scratch needs per listed group
size_t scratch_per_entry(const bf_fit_plan& p, bool split) {
  size_t b = BF_GROUP * 8 + 8 + 8 + 8;  // weights, part_wsum, part_chi2, part_chi2_j
  if (split) b += 4 + 128 + 64 + 8 * static_cast<size_t>(p.params.fast_pad > 0 ? p.params.fast_pad : 16);
  return b;
}

bool plan_uses_split(const bf_fit_plan& p) {
  return p.params.n_fast + p.params.n_slow == p.params.n_q && p.params.n_q > 0 &&
         p.grid.cov_off == nullptr && p.grid.nb <= 8 && !p.force_legacy;
}

}  // namespace
}  // namespace bf

extern "C" int64_t bf_fit_workspace_bytes(const bf_fit_plan* plan, int64_t n_stars,
                                          int64_t list_entries_hint) {
  using namespace bf;
  if (!plan || n_stars <= 0) return 0;
  const bf_fit_plan& p = *plan;
  const int64_t ld = pad_to(n_stars, BF_STAR_BLOCK);
  const int nq = p.params.n_q;
  size_t b = 0;
  b += 2 * static_cast<size_t>(n_stars) * p.nb * 8;            // flux copies
  b += static_cast<size_t>(p.nb) * ld * 4;                      // flux32
  b += static_cast<size_t>(p.n_tiles) * (BF_TILE_MODELS / BF_GROUP) * ld * 4;  // group minima
  b += static_cast<size_t>(ld) * 12 + 256;                      // starmin, selection ceilings, counter
  b += (2 * static_cast<size_t>(n_stars) * BF_SELECT_SPLIT + 2) * 4 + bf_scan_scratch_elems(n_stars * BF_SELECT_SPLIT) * 4;
  b += static_cast<size_t>(n_stars) * (4 * 4 + 4 + 8 + 4) + 2 * sort_temp_bytes(n_stars);
  // results in sorted order
  b += static_cast<size_t>(n_stars) * (8 * 5 + 4 + 8);
  b += static_cast<size_t>(n_stars) * (static_cast<size_t>(nq) * 4 + p.params.total_bins) * 8;
  b += static_cast<size_t>(n_stars) * static_cast<size_t>(p.pair_len) * 8;
  b += static_cast<size_t>(list_entries_hint) * 20;              // list, entry_star, masks, prefix
  b += static_cast<size_t>(bf_group_order_work_elems(p.n_groups)) * 4;
  // scratch: room for at least a third of the list at a time
  b += static_cast<size_t>(list_entries_hint / 3 + p.n_groups) * scratch_per_entry(p, plan_uses_split(p));
  return static_cast<int64_t>(b + (64 << 10));
}

extern "C" int bf_fit_batch(const bf_fit_plan* plan, int64_t n_stars, const double* flux64,
                            int flux_on_host, double threshold, const double* percentiles_host,
                            int sparse_mode, int64_t lnp_npts, uint64_t seed, int64_t star_base,
                            void* workspace, size_t workspace_bytes, const bf_fit_outputs* out,
                            bf_fit_state* state, void* stream, void* aux_stream) {
  using namespace bf;
  BF_REQUIRE(plan && out && percentiles_host && flux64 && workspace, "bf_fit_batch: null argument");
  const bf_fit_plan& p = *plan;
  BF_REQUIRE(n_stars > 0 && n_stars < (1LL << 31), "bf_fit_batch: bad star count");
  BF_REQUIRE(p.nb >= 1 && p.nb <= BF_MAX_BANDS && p.n_tiles > 0 && p.tiles && p.inv_band_scale,
             "bf_fit_batch: incomplete plan");
  BF_REQUIRE(p.n_groups == p.n_tiles * (BF_TILE_MODELS / BF_GROUP), "bf_fit_batch: n_groups does not match n_tiles");
  BF_REQUIRE(out->pmax && out->pmax_j && out->total_log_norm && out->chi2min && out->chi2min_j &&
                 out->n_sparse && out->exp_vals && out->pdf1d && out->per_vals,
             "bf_fit_batch: null output pointer");
  BF_REQUIRE(sparse_mode >= 0 && sparse_mode <= 2, "bf_fit_batch: sparse_mode must be 0, 1 or 2");
  BF_REQUIRE(sparse_mode == 0 || (out->sparse_off && out->sparse_j && out->sparse_lnp && out->sparse_chi2),
             "bf_fit_batch: sparse records wanted but no room given");
  BF_REQUIRE(sparse_mode != 2 || (lnp_npts > 0 && out->sparse_capacity >= n_stars * lnp_npts),
             "bf_fit_batch: sparse_capacity must hold n_stars * lnp_npts records");
  BF_REQUIRE(!(out->pdf2d) || (p.n_pairs > 0 && p.pair_bin && p.pair_off && p.pair_len > 0),
             "bf_fit_batch: 2D PDFs wanted but the plan has no pair tables");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nb = p.nb, nq = p.params.n_q, tb = p.params.total_bins;
  const int64_t n = n_stars, ld = pad_to(n, BF_STAR_BLOCK);
  const bool split = plan_uses_split(p);
  int64_t launches = 0;

  StageTimer tm{state != nullptr && state->want_timing != 0, st, state ? state->stage_ms : nullptr};
  if (tm.on) {
    for (float& v : state->stage_ms) v = 0.f;
    g_accum_timer = AccumTimer();
    g_accum_timer.on = true;
  }
  struct TimerOff {
    ~TimerOff() { g_accum_timer.on = false; }
  } timer_off;
  Arena ar{static_cast<char*>(workspace), 0, workspace_bytes & ~static_cast<size_t>(255)};
  BF_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "bf_fit_batch: workspace must be 256-byte aligned");
#define BF_WS(ptr) do { if (!(ptr) || !ar.ok) { set_error("bf_fit_batch: workspace of %zu bytes is too small (bf_fit_workspace_bytes)", workspace_bytes); return BF_ERR_TOO_LARGE; } } while (0)

  // ---- fluxes on the device, FP32 copy ----------------------------------------------------
  double* d_flux = ar.take<double>(n * nb);
  double* d_flux_sorted = ar.take<double>(n * nb);
  float* flux32 = ar.take<float>(static_cast<size_t>(nb) * ld);
  unsigned long long* starmin = ar.take<unsigned long long>(ld);
  BF_WS(starmin);
  if (flux_on_host)
    BF_CUDA_TRY(cudaMemcpyAsync(d_flux, flux64, static_cast<size_t>(n) * nb * 8, cudaMemcpyHostToDevice, st));
  const double* flux_in = flux_on_host ? d_flux : flux64;

  // ---- presort: coarse sweep, stars ordered by the group of their coarse minimum ----------------
  int32_t* perm = nullptr;  // sorted position -> input position
  int sm_count = p.sm_count;
  if (sm_count <= 0) {
    int dev = 0;
    BF_CUDA_TRY(cudaGetDevice(&dev));
    BF_CUDA_TRY(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
  }
  const bool cov = p.full_cov != 0;
  const float tau = static_cast<float>(-2.0 * threshold);
  const float* cut = nullptr;  // set by the presort when the dense sweep may skip groups
  uint64_t* evaluated = nullptr;
  auto dense = [&](int64_t n_tiles, const float* tiles, float* gmin, bool prune) -> int {
    starmin_init_kernel<<<blocks_for(ld), kT, 0, st>>>(ld, starmin);
    int tpi = bf_dense_tiles_per_item(n_tiles, n, sm_count);
    launches += 2;
    // (the pruned sweep's blocks differ widely in what they have to evaluate: four times as many,
    //  shorter items even them out -- dense stage 44.7 -> 41.7 ms per C2 step, no gain beyond)
    if (prune && tpi >= 16) tpi /= 4;
    if (prune)
      return bf_dense_chi2_pruned(nb, n_tiles, tiles, p.boxes, n, ld, flux32, cut, gmin,
                                  reinterpret_cast<uint64_t*>(starmin), tpi, evaluated, stream);
    return cov ? bf_dense_chi2_cov(nb, n_tiles, tiles, n, ld, flux32, gmin, reinterpret_cast<uint64_t*>(starmin), tpi, stream)
               : bf_dense_chi2(nb, n_tiles, tiles, n, ld, flux32, gmin, reinterpret_cast<uint64_t*>(starmin), tpi, stream);
  };
  const double* flux_run = flux_in;
  tm.begin();
  if (p.n_tiles_coarse > 0 && p.tiles_coarse && n > 1) {
    flux_prepare_kernel<<<blocks_for(ld * nb), kT, 0, st>>>(nb, n, ld, flux_in, p.inv_band_scale, flux32);
    float* cg = ar.take_top<float>(static_cast<size_t>(p.n_tiles_coarse) * (BF_TILE_MODELS / BF_GROUP) * ld);
    BF_WS(cg);
    int rc = dense(p.n_tiles_coarse, p.tiles_coarse, cg, false);
    if (rc != BF_OK) return rc;
    ar.hi = workspace_bytes & ~static_cast<size_t>(255);  // the coarse map is not needed any more
    uint32_t* keys = ar.take<uint32_t>(n);
    uint32_t* keys_out = ar.take<uint32_t>(n);
    int32_t* vals = ar.take<int32_t>(n);
    perm = ar.take<int32_t>(n);
    size_t tbytes = sort_temp_bytes(n);
    void* temp = ar.take<char>(tbytes);
    BF_WS(temp);
    static const char* sort_env = std::getenv("BF_SORT_BAND");
    const int sort_band = sort_env ? std::atoi(sort_env) : p.sort_band;
    if (sort_band >= 0 && sort_band < nb)
      flux_keys_kernel<<<blocks_for(n), kT, 0, st>>>(n, nb, sort_band, flux_in, keys, vals);
    else
      argmin_keys_kernel<<<blocks_for(n), kT, 0, st>>>(n, starmin, keys, vals);
    BF_CUDA_TRY(cub::DeviceRadixSort::SortPairs(temp, tbytes, keys, keys_out, vals, perm, static_cast<int>(n), 0, 32, st));
    gather_rows_kernel<<<blocks_for(n * nb), kT, 0, st>>>(n, nb, perm, flux_in, d_flux_sorted);
    flux_run = d_flux_sorted;
    launches += 5;
    static const bool no_prune = std::getenv("BF_NO_PRUNE") != nullptr;
    if (p.boxes && !cov && nb <= 12 && !no_prune) {
      float* c = ar.take<float>(ld);
      evaluated = ar.take<uint64_t>(1);
      BF_WS(evaluated);
      BF_CUDA_TRY(cudaMemsetAsync(evaluated, 0, sizeof(uint64_t), st));
      prune_cut_kernel<<<blocks_for(ld), kT, 0, st>>>(n, ld, perm, starmin, tau, p.err_a, p.err_b, c);
      cut = c;
      ++launches;
    }
  }
  tm.end(0);
  tm.begin();
  flux_prepare_kernel<<<blocks_for(ld * nb), kT, 0, st>>>(nb, n, ld, flux_run, p.inv_band_scale, flux32);
  BF_CHECK_LAUNCH("flux_prepare_kernel");

  // ---- dense sweep, group selection -----------------------------------------------------------
  float* gmin = ar.take_top<float>(static_cast<size_t>(p.n_groups) * ld);
  BF_WS(gmin);
  int rc = dense(p.n_tiles, p.tiles, gmin, cut != nullptr);
  if (rc != BF_OK) return rc;
  tm.end(1);
  tm.begin();
  const int64_t nc = n * BF_SELECT_SPLIT;
  uint32_t* counts = ar.take<uint32_t>(nc);
  uint32_t* offsets = ar.take<uint32_t>(nc + 1);
  uint32_t* scan_tmp = ar.take<uint32_t>(bf_scan_scratch_elems(nc > n ? nc : n));
  uint32_t* off_star = ar.take<uint32_t>(n + 1);
  int64_t* n_listed = ar.take<int64_t>(n);
  BF_WS(n_listed);
  rc = bf_select_count(p.n_groups, n, ld, gmin, reinterpret_cast<uint64_t*>(starmin), tau, p.err_a, p.err_b, counts, stream);
  if (rc != BF_OK) return rc;
  rc = bf_exclusive_scan_u32(nc, counts, offsets, scan_tmp, stream);
  if (rc != BF_OK) return rc;
  star_offsets_kernel<<<blocks_for(n + 1), kT, 0, st>>>(n, offsets, off_star, n_listed);
  launches += 5;
  std::vector<uint32_t> off_host(n + 1);
  BF_CUDA_TRY(cudaMemcpyAsync(off_host.data(), off_star, (n + 1) * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
  uint64_t evaluated_host = static_cast<uint64_t>(ld) * static_cast<uint64_t>(p.n_groups) * BF_GROUP;
  if (evaluated)
    BF_CUDA_TRY(cudaMemcpyAsync(&evaluated_host, evaluated, sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
  BF_CUDA_TRY(cudaStreamSynchronize(st));  // the list sizes decide the layout of everything below
  const int64_t total = off_host[n];
  const int64_t tot1 = total > 0 ? total : 1;
  uint32_t* list = ar.take<uint32_t>(tot1);
  uint32_t* entry_star = ar.take<uint32_t>(tot1);
  uint32_t* mask = ar.take<uint32_t>(tot1 * (BF_GROUP / 32));
  uint32_t* prefix = ar.take<uint32_t>(tot1);
  BF_WS(prefix);
  rc = bf_select_fill(p.n_groups, n, ld, gmin, reinterpret_cast<uint64_t*>(starmin), tau, p.err_a, p.err_b, offsets, list, entry_star, stream);
  if (rc != BF_OK) return rc;

  // ---- results in sorted order (scattered back at the end) --------------------------------
  const bool direct = perm == nullptr;  // no presort: write straight into the caller's arrays
  double* s_pmax = direct ? out->pmax : ar.take<double>(n);
  int64_t* s_pmax_j = direct ? out->pmax_j : ar.take<int64_t>(n);
  double* s_tln = direct ? out->total_log_norm : ar.take<double>(n);
  double* s_chi2 = direct ? out->chi2min : ar.take<double>(n);
  int64_t* s_chi2_j = direct ? out->chi2min_j : ar.take<int64_t>(n);
  uint32_t* s_nsp = direct ? out->n_sparse : ar.take<uint32_t>(n);
  double* s_exp = direct ? out->exp_vals : ar.take<double>(static_cast<size_t>(n) * nq);
  double* s_per = direct ? out->per_vals : ar.take<double>(static_cast<size_t>(n) * nq * 3);
  double* s_pdf = direct ? out->pdf1d : ar.take<double>(static_cast<size_t>(n) * tb);
  double* s_pdf2d = nullptr;
  if (out->pdf2d) s_pdf2d = direct ? out->pdf2d : ar.take<double>(static_cast<size_t>(n) * p.pair_len);
  BF_WS(s_pdf);
  if (out->pdf2d) {
    BF_WS(s_pdf2d);
    BF_CUDA_TRY(cudaMemsetAsync(s_pdf2d, 0, static_cast<size_t>(n) * p.pair_len * 8, st));
  }

  tm.end(2);
  tm.begin();
  rc = bf_exact_max(&p.grid, n, ld, flux_run, gmin, reinterpret_cast<uint64_t*>(starmin), p.err_a, p.err_b, offsets, list, s_pmax, s_pmax_j, stream);
  if (rc != BF_OK) return rc;
  launches += 2;
  tm.end(3);
  ar.hi = workspace_bytes & ~static_cast<size_t>(255);  // the group minima are done with

  // ---- the sparse pass, one range of stars at a time ----------------------------------------------
  uint32_t* okeys = ar.take<uint32_t>(n);
  uint32_t* okeys_out = ar.take<uint32_t>(n);
  int32_t* oidx = ar.take<int32_t>(n);
  int32_t* order = ar.take<int32_t>(n);
  int32_t* flags = ar.take<int32_t>(n);
  size_t otemp_bytes = sort_temp_bytes(n);
  void* otemp = ar.take<char>(otemp_bytes);
  uint32_t* gwork = split ? ar.take<uint32_t>(bf_group_order_work_elems(p.n_groups)) : nullptr;
  BF_WS(otemp);
  if (split) BF_WS(gwork);
  const size_t per_entry = scratch_per_entry(p, split);
  const size_t room = ar.hi > ar.lo + 4096 ? ar.hi - ar.lo - 4096 : 0;
  int64_t max_entries = static_cast<int64_t>(room / (per_entry + 8));  // (+8: alignment slack)
  int64_t longest = 0;
  for (int64_t i = 0; i < n; ++i) {
    const int64_t len = static_cast<int64_t>(off_host[i + 1]) - off_host[i];
    if (len > longest) longest = len;
  }
  if (max_entries < longest || max_entries < 1) {
    set_error("bf_fit_batch: workspace of %zu bytes leaves room for %lld listed groups at a time, one star lists %lld",
              workspace_bytes, static_cast<long long>(max_entries), static_cast<long long>(longest));
    return BF_ERR_TOO_LARGE;
  }
  if (p.max_range_entries > 0 && max_entries > p.max_range_entries)
    max_entries = p.max_range_entries > longest ? p.max_range_entries : longest;
  if (max_entries > total) max_entries = tot1;
  bf_accum_scratch scr = {};
  scr.weights = ar.take<double>(static_cast<size_t>(max_entries) * BF_GROUP);
  scr.part_wsum = ar.take<double>(max_entries);
  scr.part_chi2 = ar.take<double>(max_entries);
  scr.part_chi2_j = ar.take<int64_t>(max_entries);
  scr.star_flags = flags;
  if (split) {
    scr.group_order = ar.take<uint32_t>(max_entries);
    scr.chunk_refs = ar.take<uint32_t>(static_cast<size_t>(max_entries) * 32);
    scr.chunk_wsum = ar.take<double>(static_cast<size_t>(max_entries) * 8);
    scr.exp_part = ar.take<double>(static_cast<size_t>(max_entries) * (p.params.fast_pad > 0 ? p.params.fast_pad : 16));
    BF_WS(scr.exp_part);
  }
  BF_WS(scr.part_chi2_j);
  bf_star_outputs so = {s_tln, s_chi2, s_chi2_j, s_nsp, s_exp, s_pdf, s_per, mask, prefix};
  listed_keys_kernel<<<blocks_for(n), kT, 0, st>>>(n, n_listed, okeys);
  launches += 1;
  int n_ranges = 0;
  for (int64_t a0 = 0; a0 < n;) {
    ++n_ranges;
    int64_t b0 = a0 + 1;
    while (b0 < n && static_cast<int64_t>(off_host[b0 + 1]) - off_host[a0] <= max_entries) ++b0;
    const int64_t base = off_host[a0], cnt = static_cast<int64_t>(off_host[b0]) - off_host[a0];
    const int64_t m = b0 - a0;
    // one block per star in the histogram kernel: longest lists first
    iota_kernel<<<blocks_for(m), kT, 0, st>>>(m, oidx);
    BF_CUDA_TRY(cub::DeviceRadixSort::SortPairsDescending(otemp, otemp_bytes, okeys + a0, okeys_out, oidx, order,
                                                          static_cast<int>(m), 0, 32, st));
    launches += 2;
    tm.begin();
    if (split && cnt > 0) {
      rc = bf_group_order(p.n_groups, list, base, cnt, gwork, scr.group_order, stream);
      if (rc != BF_OK) return rc;
      launches += 5;
    }
    tm.end(4);
    rc = bf_exact_accumulate(&p.grid, &p.params, a0, m, flux_run, s_pmax, threshold, offsets, list, entry_star, order, base, cnt,
                             percentiles_host, &scr, &so, stream, aux_stream);
    if (rc != BF_OK) return rc;
    launches += split ? 2 : 3;
    tm.begin();
    if (s_pdf2d && cnt > 0) {
      rc = bf_pdf2d_from_weights(p.grid.ld_model, p.n_pairs, p.pair_bin, p.pair_off, s_pmax, s_tln, list, entry_star, mask,
                                 scr.weights, base, cnt, s_pdf2d, stream);
      if (rc != BF_OK) return rc;
      launches += 1;
    }
    tm.end(8);
    a0 = b0;
  }
  if (tm.on) {
    state->stage_ms[5] = g_accum_timer.weights_ms;
    state->stage_ms[6] = g_accum_timer.hist_ms;
    state->stage_ms[7] = g_accum_timer.exp_ms;
  }
  tm.begin();

  // ---- back to input order ----------------------------------------------------------------
  if (!direct) {
    auto scatter8 = [&](const void* src, void* dst, int64_t row) {
      scatter_rows_kernel<unsigned long long><<<blocks_for(n * row), kT, 0, st>>>(
          n, row, perm, static_cast<const unsigned long long*>(src), static_cast<unsigned long long*>(dst));
      ++launches;
    };
    scatter8(s_pmax, out->pmax, 1);
    scatter8(s_pmax_j, out->pmax_j, 1);
    scatter8(s_tln, out->total_log_norm, 1);
    scatter8(s_chi2, out->chi2min, 1);
    scatter8(s_chi2_j, out->chi2min_j, 1);
    scatter_rows_kernel<uint32_t><<<blocks_for(n), kT, 0, st>>>(n, 1, perm, s_nsp, out->n_sparse);
    scatter8(s_exp, out->exp_vals, nq);
    scatter8(s_per, out->per_vals, static_cast<int64_t>(nq) * 3);
    scatter8(s_pdf, out->pdf1d, tb);
    if (s_pdf2d) scatter8(s_pdf2d, out->pdf2d, p.pair_len);
    if (out->n_groups_selected) scatter8(n_listed, out->n_groups_selected, 1);
    ++launches;
  } else if (out->n_groups_selected) {
    BF_CUDA_TRY(cudaMemcpyAsync(out->n_groups_selected, n_listed, n * 8, cudaMemcpyDeviceToDevice, st));
  }
  BF_CHECK_LAUNCH("scatter_rows_kernel");
  tm.end(9);
  tm.begin();

  // ---- sparse lnP records (fit.py:693-709), written in input order ----------------------------
  if (sparse_mode != 0) {
    uint32_t* rc_counts = ar.take<uint32_t>(n);
    uint32_t* rc_off = ar.take<uint32_t>(n + 1);
    BF_WS(rc_off);
    record_counts_kernel<<<blocks_for(n), kT, 0, st>>>(n, out->n_sparse, sparse_mode == 2 ? lnp_npts : 0, rc_counts);
    rc = bf_exclusive_scan_u32(n, rc_counts, rc_off, scan_tmp, stream);
    if (rc != BF_OK) return rc;
    widen_kernel<<<blocks_for(n + 1), kT, 0, st>>>(n + 1, rc_off, out->sparse_off);
    launches += 5;
    if (sparse_mode == 1) {
      uint32_t nrec = 0;
      BF_CUDA_TRY(cudaMemcpyAsync(&nrec, rc_off + n, 4, cudaMemcpyDeviceToHost, st));
      BF_CUDA_TRY(cudaStreamSynchronize(st));
      if (static_cast<int64_t>(nrec) > out->sparse_capacity) {
        set_error("bf_fit_batch: %u sparse records, room for %lld", nrec, static_cast<long long>(out->sparse_capacity));
        if (state) state->n_records = nrec;
        return BF_ERR_TOO_LARGE;
      }
    }
    rc = launch_sparse_records(p.grid, n, flux_run, offsets, list, mask, prefix, s_nsp, perm, out->sparse_off,
                               nullptr, sparse_mode == 2 ? lnp_npts : 0, seed, star_base, out->sparse_j,
                               out->sparse_lnp, out->sparse_chi2, st);
    if (rc != BF_OK) return rc;
    launches += 1;
  }
  tm.end(10);
  if (state) {
    state->flux64 = flux_run;
    state->offsets = offsets;
    state->list = list;
    state->member_mask = mask;
    state->member_prefix = prefix;
    state->perm = perm;
    state->n_sparse_sorted = s_nsp;
    state->n_list = total;
    state->launches = launches;
    state->split = split ? 1 : 0;
    state->n_ranges = n_ranges;
    state->dense_evaluated = static_cast<int64_t>(evaluated_host);
  }
#undef BF_WS
  return BF_OK;
}

extern "C" int bf_fit_sparse_records(const bf_fit_plan* plan, const bf_fit_state* state, int64_t n_stars,
                                     const int64_t* rec_off, const uint32_t* ranks, int64_t* out_j,
                                     float* out_lnp, float* out_chi2, void* stream) {
  using namespace bf;
  BF_REQUIRE(plan && state && rec_off && out_j && out_lnp && out_chi2,
             "bf_fit_sparse_records: null argument");
  BF_REQUIRE(n_stars > 0 && state->offsets && state->list, "bf_fit_sparse_records: no fit state");
  return launch_sparse_records(plan->grid, n_stars, state->flux64, state->offsets, state->list,
                               state->member_mask, state->member_prefix, state->n_sparse_sorted,
                               state->perm, rec_off, ranks, 0, 0, 0, out_j, out_lnp, out_chi2,
                               static_cast<cudaStream_t>(stream));
}
