// Grid trimming (the step before the fit): which models can possibly match the observed
// flux range?  Reference: beast/fitting/trim_grid.py:53-118.
//
//   keep model m  <=>  completeness_mb > 0 in every band                         (:84-88)
//                 and, band after band, [val - f*sigma, val + f*sigma] overlaps
//                     [min_data_b, max_data_b], val = sed + bias                 (:98-111)
//
// with the reference's quirk that a band whose test would leave nothing is skipped (:110-111).
// To honour that without a host round trip per band, one kernel records every model's pass/fail
// bit pattern and a histogram over the 2^nb patterns; the host resolves the skip sequence from
// the (tiny) histogram and a second pass compacts the surviving indices in ascending order.
// Arithmetic is done without FMA contraction so the comparisons see the reference's values.

#include "common.cuh"

namespace bf {

constexpr uint32_t kTrimComplete = 1u << 31;

__global__ void __launch_bounds__(256)
trim_flags_kernel(int nb, int64_t n_models, const double* __restrict__ seds,
                  const double* __restrict__ bias, const double* __restrict__ unc,
                  const double* __restrict__ compl_, double sigma_fac, const double* __restrict__ lo,
                  const double* __restrict__ hi, uint32_t* __restrict__ pattern,
                  uint32_t* __restrict__ hist) {
  const int64_t m = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const bool valid = m < n_models;
  uint32_t p = kTrimComplete;
  for (int b = 0; valid && b < nb; ++b) {
    const int64_t i = m * nb + b;
    if (!(compl_[i] > 0.0)) p &= ~kTrimComplete;
    const double val = __dadd_rn(seds[i], bias[i]);
    const double w = __dmul_rn(sigma_fac, unc[i]);
    const double down = __dsub_rn(val, w), up = __dadd_rn(val, w);
    if (up >= lo[b] && down <= hi[b]) p |= 1u << b;
  }
  if (valid) pattern[m] = p;
  // few distinct patterns exist, so neighbouring models mostly share one: count per warp first
  const uint32_t bin = valid ? ((p & 0xffffu) | ((p & kTrimComplete) ? (1u << nb) : 0u)) : 0xffffffffu;
  const uint32_t peers = __match_any_sync(0xffffffffu, bin);
  if (valid && (threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(hist + bin, __popc(peers));
}

__global__ void __launch_bounds__(256)
trim_select_kernel(int64_t n_models, const uint32_t* __restrict__ pattern, uint32_t required,
                   uint32_t* __restrict__ keep) {
  const int64_t m = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (m < n_models) keep[m] = ((pattern[m] & required) == required) ? 1u : 0u;
}

__global__ void __launch_bounds__(256)
trim_compact_kernel(int64_t n_models, const uint32_t* __restrict__ keep,
                    const uint32_t* __restrict__ offsets, int64_t* __restrict__ indices) {
  const int64_t m = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (m < n_models && keep[m]) indices[offsets[m]] = m;
}

}  // namespace bf

extern "C" int bf_trim_flags(int nb, int64_t n_models, const double* seds, const double* bias,
                             const double* unc, const double* completeness, double sigma_fac,
                             const double* min_data, const double* max_data, uint32_t* pattern,
                             uint32_t* hist, void* stream) {
  using namespace bf;
  BF_REQUIRE(nb >= 1 && nb <= 15, "bf_trim_flags: nb=%d out of range (1..15: one pattern bit per band)", nb);
  BF_REQUIRE(n_models > 0 && (n_models + 255) / 256 < (1LL << 31), "bf_trim_flags: bad n_models");
  BF_REQUIRE(seds && bias && unc && completeness && min_data && max_data && pattern && hist,
             "bf_trim_flags: null pointer");
  trim_flags_kernel<<<static_cast<unsigned>((n_models + 255) / 256), 256, 0,
                      static_cast<cudaStream_t>(stream)>>>(
      nb, n_models, seds, bias, unc, completeness, sigma_fac, min_data, max_data, pattern, hist);
  BF_CHECK_LAUNCH("trim_flags_kernel");
  return BF_OK;
}

extern "C" int bf_trim_select(int64_t n_models, const uint32_t* pattern, uint32_t required,
                              uint32_t* keep, void* stream) {
  using namespace bf;
  BF_REQUIRE(n_models > 0 && (n_models + 255) / 256 < (1LL << 31), "bf_trim_select: bad n_models");
  BF_REQUIRE(pattern && keep, "bf_trim_select: null pointer");
  trim_select_kernel<<<static_cast<unsigned>((n_models + 255) / 256), 256, 0,
                       static_cast<cudaStream_t>(stream)>>>(n_models, pattern, required, keep);
  BF_CHECK_LAUNCH("trim_select_kernel");
  return BF_OK;
}

extern "C" int bf_trim_compact(int64_t n_models, const uint32_t* keep, const uint32_t* offsets,
                               int64_t* indices, void* stream) {
  using namespace bf;
  BF_REQUIRE(n_models > 0 && (n_models + 255) / 256 < (1LL << 31), "bf_trim_compact: bad n_models");
  BF_REQUIRE(keep && offsets && indices, "bf_trim_compact: null pointer");
  trim_compact_kernel<<<static_cast<unsigned>((n_models + 255) / 256), 256, 0,
                        static_cast<cudaStream_t>(stream)>>>(n_models, keep, offsets, indices);
  BF_CHECK_LAUNCH("trim_compact_kernel");
  return BF_OK;
}
