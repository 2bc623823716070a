// Sparse lnP records and 2D PDFs: both walk the membership bits the sparse pass left
// behind (member_mask / member_prefix) and re-evaluate the exact lnP of the members.
//
//   bf_sparse_gather    -> the (idx, lnp, chi2) triples the reference appends to
//                          save_lnp_vals (beast/fitting/fit.py:693-709)
//   bf_pdf2d_accumulate -> pdf2d.gen2d for every parameter pair
//                          (beast/fitting/pdf2d.py:122-148, fit.py:758-762)

#include "exact.cuh"

namespace bf {

constexpr int kGatherThreads = 128;

__global__ void __launch_bounds__(kGatherThreads)
sparse_gather_kernel(bf_exact_grid grid, const double* __restrict__ flux64,
                     const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ list,
                     const uint32_t* __restrict__ member_mask,
                     const uint32_t* __restrict__ member_prefix,
                     const int64_t* __restrict__ rec_off, const uint32_t* __restrict__ ranks,
                     int64_t* __restrict__ out_j, float* __restrict__ out_lnp,
                     float* __restrict__ out_chi2) {
  __shared__ double sflux[BF_MAX_BANDS];
  const int64_t s = blockIdx.x;
  if (threadIdx.x < grid.nb) sflux[threadIdx.x] = flux64[s * grid.nb + threadIdx.x];
  __syncthreads();
  const uint32_t e0 = offsets[s * BF_SELECT_SPLIT];
  const uint32_t e1 = offsets[(s + 1) * BF_SELECT_SPLIT];
  const int64_t r0 = rec_off[s], r1 = rec_off[s + 1];
  for (int64_t r = r0 + threadIdx.x; r < r1; r += kGatherThreads) {
    const uint32_t rank = ranks ? ranks[r] : static_cast<uint32_t>(r - r0);
    // last entry whose exclusive prefix is <= rank
    uint32_t lo = e0, hi = e1;  // invariant: prefix[lo] <= rank, answer in [lo, hi)
    while (hi - lo > 1) {
      const uint32_t mid = lo + (hi - lo) / 2;
      if (member_prefix[mid] <= rank) lo = mid; else hi = mid;
    }
    uint32_t local = rank - member_prefix[lo];
    int64_t j = -1;
#pragma unroll
    for (int h = 0; h < kWordsPerGroup; ++h) {
      const uint32_t word = member_mask[static_cast<size_t>(lo) * kWordsPerGroup + h];
      const uint32_t c = __popc(word);
      if (j < 0) {
        if (local < c) {
          const int bit = __fns(word, 0, static_cast<int>(local) + 1);
          j = static_cast<int64_t>(list[lo]) * BF_GROUP + h * 32 + bit;
        } else {
          local -= c;
        }
      }
    }
    double lnp = 0.0, chi2 = 0.0;
    if (j >= 0) exact_lnp(grid, sflux, j, lnp, chi2);
    out_j[r] = j;
    out_lnp[r] = static_cast<float>(lnp);
    out_chi2[r] = static_cast<float>(chi2);
  }
}

// ---- device-side random subsample of each star's sparse set ---------------------------
// The reference draws np.random.choice(indx, lnp_npts, replace=False) per star
// (fit.py:694-700), which shuffles the whole set on the host: O(|S|) per star.  Here each
// star gets min(|S|, npts) *distinct* pseudo-random ranks from a keyed 4-round Feistel
// permutation of [0, |S|) (cycle walking), O(npts) per star and reproducible from the seed.
__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}

__device__ __forceinline__ uint32_t feistel_perm(uint32_t i, uint32_t n, uint32_t key) {
  // domain 2^(2h) >= n
  int bits = 32 - __clz(n - 1 > 0 ? n - 1 : 1);
  const int h = (bits + 1) / 2;
  const uint32_t hm = (1u << h) - 1u;
  uint32_t x = i;
  do {
    uint32_t l = x >> h, r = x & hm;
#pragma unroll
    for (int round = 0; round < 4; ++round) {
      const uint32_t f = mix32(r ^ (key + 0x9e3779b9u * (round + 1))) & hm;
      const uint32_t nl = r;
      r = l ^ f;
      l = nl;
    }
    x = (l << h) | r;
  } while (x >= n);
  return x;
}

__global__ void sample_ranks_kernel(int64_t n_stars, const uint32_t* __restrict__ n_sparse,
                                    const int64_t* __restrict__ rec_off, uint64_t seed,
                                    int64_t star_base, uint32_t* __restrict__ ranks) {
  const int64_t s = blockIdx.x;
  const uint32_t n = n_sparse[s];
  const int64_t r0 = rec_off[s], r1 = rec_off[s + 1];
  const uint32_t k = static_cast<uint32_t>(r1 - r0);
  const uint64_t sid = static_cast<uint64_t>(star_base + s);
  const uint32_t key = mix32(static_cast<uint32_t>(seed) ^ mix32(static_cast<uint32_t>(sid) * 0x9e3779b9u +
                                                               static_cast<uint32_t>(seed >> 32)));
  for (uint32_t i = threadIdx.x; i < k; i += blockDim.x)
    ranks[r0 + i] = (k >= n) ? i : feistel_perm(i, n, key);
}

// Sparse records of a whole batch whose stars were fitted in another order than the caller's:
// block i works on the i-th star of the fit order, input star perm[i] (perm == nullptr: the same
// order); its records go to out_off[input star] of the output arrays.  The ranks are the ones
// given (positions out_off[.] + r of `ranks`), or all of 0..n-1, or -- lnp_npts > 0 -- the first
// min(n, lnp_npts) values of the keyed permutation of sample_ranks_kernel.
__global__ void __launch_bounds__(kGatherThreads)
sparse_records_kernel(bf_exact_grid grid, const double* __restrict__ flux64,
                      const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ list,
                      const uint32_t* __restrict__ member_mask,
                      const uint32_t* __restrict__ member_prefix,
                      const uint32_t* __restrict__ n_sparse_sorted, const int32_t* __restrict__ perm,
                      const int64_t* __restrict__ out_off, const uint32_t* __restrict__ ranks,
                      int64_t lnp_npts, uint64_t seed, int64_t star_base,
                      int64_t* __restrict__ out_j, float* __restrict__ out_lnp,
                      float* __restrict__ out_chi2) {
  __shared__ double sflux[BF_MAX_BANDS];
  const int64_t s = blockIdx.x;
  const int64_t s_in = perm ? perm[s] : s;
  if (threadIdx.x < grid.nb) sflux[threadIdx.x] = flux64[s * grid.nb + threadIdx.x];
  __syncthreads();
  const uint32_t e0 = offsets[s * BF_SELECT_SPLIT];
  const uint32_t e1 = offsets[(s + 1) * BF_SELECT_SPLIT];
  const uint32_t n = n_sparse_sorted[s];
  const int64_t r0 = out_off[s_in];
  const uint32_t k = static_cast<uint32_t>(out_off[s_in + 1] - r0);
  const uint64_t sid = static_cast<uint64_t>(star_base + s_in);
  const uint32_t key = mix32(static_cast<uint32_t>(seed) ^ mix32(static_cast<uint32_t>(sid) * 0x9e3779b9u +
                                                               static_cast<uint32_t>(seed >> 32)));
  for (uint32_t r = threadIdx.x; r < k; r += kGatherThreads) {
    uint32_t rank = r;
    if (ranks) rank = ranks[r0 + r];
    else if (lnp_npts > 0 && k < n) rank = feistel_perm(r, n, key);
    uint32_t lo = e0, hi = e1;  // invariant: prefix[lo] <= rank, answer in [lo, hi)
    while (hi - lo > 1) {
      const uint32_t mid = lo + (hi - lo) / 2;
      if (member_prefix[mid] <= rank) lo = mid; else hi = mid;
    }
    uint32_t local = rank - member_prefix[lo];
    int64_t j = -1;
#pragma unroll
    for (int h = 0; h < kWordsPerGroup; ++h) {
      const uint32_t word = member_mask[static_cast<size_t>(lo) * kWordsPerGroup + h];
      const uint32_t c = __popc(word);
      if (j < 0) {
        if (local < c) {
          const int bit = __fns(word, 0, static_cast<int>(local) + 1);
          j = static_cast<int64_t>(list[lo]) * BF_GROUP + h * 32 + bit;
        } else {
          local -= c;
        }
      }
    }
    double lnp = 0.0, chi2 = 0.0;
    if (j >= 0) exact_lnp(grid, sflux, j, lnp, chi2);
    out_j[r0 + r] = j;
    out_lnp[r0 + r] = static_cast<float>(lnp);
    out_chi2[r0 + r] = static_cast<float>(chi2);
  }
}

int launch_sparse_records(const bf_exact_grid& grid, int64_t n_stars, const double* flux64,
                          const uint32_t* offsets, const uint32_t* list, const uint32_t* member_mask,
                          const uint32_t* member_prefix, const uint32_t* n_sparse_sorted,
                          const int32_t* perm, const int64_t* out_off, const uint32_t* ranks,
                          int64_t lnp_npts, uint64_t seed, int64_t star_base, int64_t* out_j,
                          float* out_lnp, float* out_chi2, cudaStream_t st) {
  sparse_records_kernel<<<static_cast<unsigned>(n_stars), kGatherThreads, 0, st>>>(
      grid, flux64, offsets, list, member_mask, member_prefix, n_sparse_sorted, perm, out_off, ranks,
      lnp_npts, seed, star_base, out_j, out_lnp, out_chi2);
  BF_CHECK_LAUNCH("sparse_records_kernel");
  return BF_OK;
}

// one warp per listed group; weights are already normalised: exp(lnp - total_log_norm)
constexpr int kPdf2dWarps = 8;

__global__ void __launch_bounds__(kPdf2dWarps * 32)
pdf2d_kernel(bf_exact_grid grid, int n_pairs, const int32_t* __restrict__ pair_bin,
             const int64_t* __restrict__ pair_off, const double* __restrict__ flux64,
             const double* __restrict__ pmax, const double* __restrict__ tln, double threshold,
             const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ list,
             double* __restrict__ pdf2d) {
  __shared__ double sflux[BF_MAX_BANDS];
  const int64_t s = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x < grid.nb) sflux[threadIdx.x] = flux64[s * grid.nb + threadIdx.x];
  __syncthreads();
  const double M = pmax[s];
  const double norm = tln[s];
  const uint32_t e0 = offsets[s * BF_SELECT_SPLIT];
  const uint32_t e1 = offsets[(s + 1) * BF_SELECT_SPLIT];
  const int64_t row_len = pair_off[n_pairs];
  double* row = pdf2d + s * row_len;
  for (uint32_t e = e0 + warp; e < e1; e += kPdf2dWarps) {
    const uint32_t g = list[e];
#pragma unroll
    for (int h = 0; h < kWordsPerGroup; ++h) {
      const int64_t j = static_cast<int64_t>(g) * BF_GROUP + h * 32 + lane;
      if (j >= grid.n_models) continue;
      double lnp, chi2;
      exact_lnp(grid, sflux, j, lnp, chi2);
      if (!((lnp - M) > threshold)) continue;
      const double w = exp(lnp - norm);
      for (int p = 0; p < n_pairs; ++p) {
        const int32_t b = pair_bin[static_cast<int64_t>(p) * grid.ld_model + j];
        if (b >= 0) atomicAdd(row + pair_off[p] + b, w);
      }
    }
  }
}

// Same sums from the weights the sparse pass has just left in its scratch (exp(lnP - Pmax) of the
// members, membership bits): no model is evaluated again.  One warp per listed group.
__global__ void __launch_bounds__(kPdf2dWarps * 32)
pdf2d_weights_kernel(int64_t ld_model, int n_pairs, const int32_t* __restrict__ pair_bin,
                     const int64_t* __restrict__ pair_off, const double* __restrict__ pmax,
                     const double* __restrict__ tln, const uint32_t* __restrict__ list,
                     const uint32_t* __restrict__ entry_star, const uint32_t* __restrict__ member_mask,
                     const double* __restrict__ weights, int64_t entry_base, int64_t n_entries,
                     double* __restrict__ pdf2d) {
  const int lane = threadIdx.x & 31;
  const int64_t row_len = pair_off[n_pairs];
  for (int64_t slot = static_cast<int64_t>(blockIdx.x) * kPdf2dWarps + (threadIdx.x >> 5);
       slot < n_entries; slot += static_cast<int64_t>(gridDim.x) * kPdf2dWarps) {
    const uint32_t e = static_cast<uint32_t>(entry_base + slot);
    const uint32_t m0 = member_mask[static_cast<size_t>(e) * kWordsPerGroup];
    const uint32_t m1 = member_mask[static_cast<size_t>(e) * kWordsPerGroup + 1];
    if ((m0 | m1) == 0) continue;
    const int64_t s = entry_star[e];
    // weights are relative to Pmax; 1 / (sum of weights) = exp(Pmax - total_log_norm) (fit.py:716)
    const double inv = exp(pmax[s] - tln[s]);
    double* row = pdf2d + s * row_len;
    const int64_t j0 = static_cast<int64_t>(list[e]) * BF_GROUP;
#pragma unroll
    for (int h = 0; h < kWordsPerGroup; ++h) {
      if (!(((h ? m1 : m0) >> lane) & 1u)) continue;
      const int64_t j = j0 + h * 32 + lane;
      const double w = weights[slot * BF_GROUP + h * 32 + lane] * inv;
      for (int p = 0; p < n_pairs; ++p) {
        const int32_t b = pair_bin[static_cast<int64_t>(p) * ld_model + j];
        if (b >= 0) atomicAdd(row + pair_off[p] + b, w);
      }
    }
  }
}

}  // namespace bf

extern "C" int bf_pdf2d_from_weights(int64_t ld_model, int n_pairs, const int32_t* pair_bin,
                                     const int64_t* pair_off, const double* pmax,
                                     const double* total_log_norm, const uint32_t* list,
                                     const uint32_t* entry_star, const uint32_t* member_mask,
                                     const double* weights, int64_t entry_base, int64_t n_entries,
                                     double* pdf2d, void* stream) {
  using namespace bf;
  BF_REQUIRE(n_pairs >= 1 && ld_model > 0, "bf_pdf2d_from_weights: bad sizes");
  BF_REQUIRE(entry_base >= 0 && n_entries >= 0, "bf_pdf2d_from_weights: bad entry range");
  BF_REQUIRE(pair_bin && pair_off && pmax && total_log_norm && list && entry_star && member_mask &&
                 weights && pdf2d,
             "bf_pdf2d_from_weights: null pointer");
  if (n_entries == 0) return BF_OK;
  const int64_t blocks = (n_entries + kPdf2dWarps - 1) / kPdf2dWarps;
  int dev = 0, sms = 0;
  BF_CUDA_TRY(cudaGetDevice(&dev));
  BF_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int64_t cap = static_cast<int64_t>(sms) * 64;  // grid-stride kernel: 64 blocks per SM
  pdf2d_weights_kernel<<<static_cast<unsigned>(blocks < cap ? blocks : cap), kPdf2dWarps * 32, 0,
                         static_cast<cudaStream_t>(stream)>>>(
      ld_model, n_pairs, pair_bin, pair_off, pmax, total_log_norm, list, entry_star, member_mask,
      weights, entry_base, n_entries, pdf2d);
  BF_CHECK_LAUNCH("pdf2d_weights_kernel");
  return BF_OK;
}

extern "C" int bf_sparse_gather(const bf_exact_grid* grid, int64_t n_stars, const double* flux64,
                                const uint32_t* offsets, const uint32_t* list,
                                const uint32_t* member_mask, const uint32_t* member_prefix,
                                const int64_t* rec_off, const uint32_t* ranks, int64_t* out_j,
                                float* out_lnp, float* out_chi2, void* stream) {
  using namespace bf;
  BF_REQUIRE(grid && grid->nb >= 1 && grid->nb <= BF_MAX_BANDS, "bf_sparse_gather: bad grid");
  BF_REQUIRE(n_stars > 0 && n_stars < (1LL << 31), "bf_sparse_gather: bad n_stars");
  BF_REQUIRE(flux64 && offsets && list && member_mask && member_prefix && rec_off && out_j &&
                 out_lnp && out_chi2,
             "bf_sparse_gather: null pointer");
  sparse_gather_kernel<<<static_cast<unsigned>(n_stars), kGatherThreads, 0,
                         static_cast<cudaStream_t>(stream)>>>(
      *grid, flux64, offsets, list, member_mask, member_prefix, rec_off, ranks, out_j, out_lnp,
      out_chi2);
  BF_CHECK_LAUNCH("sparse_gather_kernel");
  return BF_OK;
}

extern "C" int bf_pdf2d_accumulate(const bf_exact_grid* grid, int n_pairs, const int32_t* pair_bin,
                                   const int64_t* pair_off, int64_t n_stars, const double* flux64,
                                   const double* pmax, const double* total_log_norm,
                                   double threshold, const uint32_t* offsets, const uint32_t* list,
                                   double* pdf2d, void* stream) {
  using namespace bf;
  BF_REQUIRE(grid && grid->nb >= 1 && grid->nb <= BF_MAX_BANDS, "bf_pdf2d_accumulate: bad grid");
  BF_REQUIRE(n_pairs >= 1, "bf_pdf2d_accumulate: n_pairs must be >= 1");
  BF_REQUIRE(n_stars > 0 && n_stars < (1LL << 31), "bf_pdf2d_accumulate: bad n_stars");
  BF_REQUIRE(pair_bin && pair_off && flux64 && pmax && total_log_norm && offsets && list && pdf2d,
             "bf_pdf2d_accumulate: null pointer");
  pdf2d_kernel<<<static_cast<unsigned>(n_stars), kPdf2dWarps * 32, 0,
                 static_cast<cudaStream_t>(stream)>>>(
      *grid, n_pairs, pair_bin, pair_off, flux64, pmax, total_log_norm, threshold, offsets, list,
      pdf2d);
  BF_CHECK_LAUNCH("pdf2d_kernel");
  return BF_OK;
}

extern "C" int bf_sparse_sample_ranks(int64_t n_stars, const uint32_t* n_sparse,
                                      const int64_t* rec_off, uint64_t seed, int64_t star_base,
                                      uint32_t* ranks, void* stream) {
  using namespace bf;
  BF_REQUIRE(n_stars > 0 && n_stars < (1LL << 31), "bf_sparse_sample_ranks: bad n_stars");
  BF_REQUIRE(n_sparse && rec_off && ranks, "bf_sparse_sample_ranks: null pointer");
  sample_ranks_kernel<<<static_cast<unsigned>(n_stars), 128, 0, static_cast<cudaStream_t>(stream)>>>(
      n_stars, n_sparse, rec_off, seed, star_base, ranks);
  BF_CHECK_LAUNCH("sample_ranks_kernel");
  return BF_OK;
}
