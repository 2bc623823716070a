// Exact (FP64) pass over the groups the dense pass selected.
//
// What the reference does per star after its likelihood sweep
// (beast/fitting/fit.py:665-755) happens here, on the selected groups only:
//   exact lnP and chi2 per model      likelihood.py:149-165, fit.py:665-668
//   max / first argmax                fit.py:670, 684, 719, 727-728
//   sparse set  lnp - max > threshold fit.py:670
//   weights, weight_sum, total_log_norm   fit.py:685-690, 716
//   chi2min (+ index)                 fit.py:725-726
//   expectations                      fit.py:743, fit_metrics/common.py:126-167
//   1D PDFs                           fit.py:746-748, pdf1d.py:119-148
//   p16/p50/p84                       fit.py:749-755, fit_metrics/common.py:48-123
// lnP is evaluated in the reference's operation order, without FMA contraction:
//   chi2 = sum_b ((f_b - m_b) * (f_b - m_b)) * ivar_b ;  lnP = (-lnQ - 0.5 chi2) + ln w

#include "exact.cuh"

namespace bf {

// ---------------------------------------------------------------------------------------
// K3: exact maximum over the groups whose FP32 minimum is within the FP32 error bound of the star's.
// one warp per star
// ---------------------------------------------------------------------------------------
constexpr int kMaxWarps = 4;

__global__ void __launch_bounds__(kMaxWarps * 32)
exact_max_kernel(bf_exact_grid grid, int64_t n_stars, int64_t ld_star,
                 const double* __restrict__ flux64, const float* __restrict__ gmin,
                 const unsigned long long* __restrict__ starmin, float err_a, float err_b,
                 const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ list,
                 double* __restrict__ pmax, int64_t* __restrict__ pmax_j) {
  __shared__ double sflux_all[kMaxWarps][BF_MAX_BANDS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t s = static_cast<int64_t>(blockIdx.x) * kMaxWarps + warp;
  if (s >= n_stars) return;
  double* sflux = sflux_all[warp];
  if (lane < grid.nb) sflux[lane] = flux64[s * grid.nb + lane];
  __syncwarp();

  const uint32_t e0 = offsets[s * BF_SELECT_SPLIT];
  const uint32_t e1 = offsets[(s + 1) * BF_SELECT_SPLIT];
  const float thr = select_threshold(starmin_value(starmin[s]), 0.0f, err_a, err_b);

  double best = -1.7976931348623157e308;
  int64_t best_j = -1;
  for (uint32_t base = e0; base < e1; base += 32) {
    const uint32_t e = base + lane;
    bool tight = false;
    uint32_t gid = 0;
    if (e < e1) {
      gid = list[e];
      tight = gmin[static_cast<int64_t>(gid) * ld_star + s] < thr;
    }
    uint32_t todo = __ballot_sync(0xffffffffu, tight);
    while (todo) {
      const int src = __ffs(todo) - 1;
      todo &= todo - 1;
      const uint32_t g = __shfl_sync(0xffffffffu, gid, src);
#pragma unroll
      for (int h = 0; h < kWordsPerGroup; ++h) {
        const int64_t j = static_cast<int64_t>(g) * BF_GROUP + h * 32 + lane;
        if (j < grid.n_models) {
          double lnp, chi2;
          exact_lnp(grid, sflux, j, lnp, chi2);
          if (is_finite(lnp) && (lnp > best || (lnp == best && j < best_j))) {
            best = lnp;
            best_j = j;
          }
        }
      }
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    const double ob = __shfl_xor_sync(0xffffffffu, best, d);
    const int64_t oj = __shfl_xor_sync(0xffffffffu, best_j, d);
    if (oj >= 0 && (best_j < 0 || ob > best || (ob == best && oj < best_j))) {
      best = ob;
      best_j = oj;
    }
  }
  if (lane == 0) {
    pmax[s] = best_j >= 0 ? best : __longlong_as_double(0x7ff8000000000000LL);
    pmax_j[s] = best_j;
  }
}

}  // namespace bf

extern "C" int bf_exact_max(const bf_exact_grid* grid, int64_t n_stars, int64_t ld_star,
                            const double* flux64, const float* gmin, const uint64_t* starmin,
                            float err_a, float err_b, const uint32_t* offsets,
                            const uint32_t* list,
                            double* pmax, int64_t* pmax_j, void* stream) {
  using namespace bf;
  BF_REQUIRE(grid && grid->nb >= 1 && grid->nb <= BF_MAX_BANDS, "bf_exact_max: bad grid");
  BF_REQUIRE(n_stars > 0 && ld_star >= n_stars, "bf_exact_max: bad sizes");
  BF_REQUIRE(flux64 && gmin && starmin && offsets && list && pmax && pmax_j,
             "bf_exact_max: null pointer");
  const unsigned blocks = static_cast<unsigned>((n_stars + kMaxWarps - 1) / kMaxWarps);
  exact_max_kernel<<<blocks, kMaxWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
      *grid, n_stars, ld_star, flux64, gmin, reinterpret_cast<const unsigned long long*>(starmin),
      err_a, err_b, offsets, list, pmax, pmax_j);
  BF_CHECK_LAUNCH("exact_max_kernel");
  return BF_OK;
}
