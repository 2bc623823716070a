// Weighted percentiles of a binned pdf, the way the reference computes them
// (beast/fitting/fit_metrics/common.py:112-121): knots w_k = (cumsum_k - pdf_k/2) / sum,
// result = np.interp(p/100, w, bin centres).  One warp per (star, parameter); the CDF is a
// warp-parallel prefix scan over the bins.
#pragma once

#include "common.cuh"

namespace bf {

constexpr int kMaxBinsPerParam = 256;  // bin indices are bytes

// np.interp(x, cdf, centres) for one x, evaluated by a whole warp (lanes stride the bins)
__device__ __forceinline__ double warp_interp(const double* cdf, const double* centres, int nbins,
                                              double x, int lane) {
  double res = 0.0;
  bool found = false;
  for (int i = lane; i < nbins; i += 32) {
    const double wi = cdf[i];
    if (wi <= x && (i == nbins - 1 || cdf[i + 1] > x)) {
      found = true;
      if (i == nbins - 1 || wi == x) {
        res = centres[i];
      } else {
        const double slope = (centres[i + 1] - centres[i]) / (cdf[i + 1] - wi);
        res = slope * (x - wi) + centres[i];
      }
    }
  }
  const uint32_t who = __ballot_sync(0xffffffffu, found);
  if (who == 0) return centres[0];  // x below the first knot: np.interp clamps to fp[0]
  return __shfl_sync(0xffffffffu, res, __ffs(who) - 1);
}

// pdf (normalised, in shared or global memory) -> three percentiles; whole warp cooperates.
__device__ __forceinline__ void warp_percentiles(const double* pdf, const double* centres,
                                                 int nbins, double* cdf, double p0, double p1,
                                                 double p2, double* out3, int lane) {
  double part = 0.0;
  for (int i = lane; i < nbins; i += 32) part += pdf[i];
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) part += __shfl_xor_sync(0xffffffffu, part, d);
  const double total = part;
  if (!(total > 0.0)) {  // all-zero pdf: the reference stores zeros (fit.py:754-755)
    if (lane < 3) out3[lane] = 0.0;
    return;
  }
  double carry = 0.0;
  for (int base = 0; base < nbins; base += 32) {
    const int i = base + lane;
    const double p = i < nbins ? pdf[i] : 0.0;
    double inc = p;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const double t = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += t;
    }
    const double aw = carry + inc;
    if (i < nbins) cdf[i] = (aw - 0.5 * p) / total;
    carry = __shfl_sync(0xffffffffu, aw, 31);
  }
  __syncwarp();
  const double r0 = warp_interp(cdf, centres, nbins, p0 * 0.01, lane);
  const double r1 = warp_interp(cdf, centres, nbins, p1 * 0.01, lane);
  const double r2 = warp_interp(cdf, centres, nbins, p2 * 0.01, lane);
  if (lane == 0) {
    out3[0] = r0;
    out3[1] = r1;
    out3[2] = r2;
  }
  __syncwarp();
}

}  // namespace bf
