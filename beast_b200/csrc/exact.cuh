// Exact FP64 likelihood of one model, shared by the exact-pass kernels.
#pragma once

#include "common.cuh"

namespace bf {

constexpr int kWordsPerGroup = BF_GROUP / 32;

// exact lnP / chi2 of model j for the star whose flux sits in shared memory
__device__ __forceinline__ void exact_lnp(const bf_exact_grid& g, const double* sflux, int64_t j,
                                          double& lnp, double& chi2) {
  double acc = 0.0;
  if (g.cov_off == nullptr) {
    const double* pm = g.msb + j;  // walk the band-major arrays: one pointer add per band
    const double* pv = g.ivar + j;
#pragma unroll 2
    for (int b = 0; b < g.nb; ++b) {
      const double d = __dsub_rn(sflux[b], __ldg(pm));
      const double dd = __dmul_rn(d, d);
      acc = __dadd_rn(acc, __dmul_rn(dd, __ldg(pv)));
      pm += g.ld_model;
      pv += g.ld_model;
    }
  } else {
    // N_covar_chi2 (likelihood.py:88-106): diagonal terms, then for each k the packed
    // off-diagonal row dotted with the remaining residuals, times the k-th residual
    double d[BF_MAX_BANDS];
    for (int b = 0; b < g.nb; ++b) {
      d[b] = __dsub_rn(sflux[b], g.msb[static_cast<int64_t>(b) * g.ld_model + j]);
      acc = __dadd_rn(acc, __dmul_rn(__dmul_rn(d[b], d[b]),
                                     g.ivar[static_cast<int64_t>(b) * g.ld_model + j]));
    }
    int m = 0;
    for (int k = 0; k < g.nb - 1; ++k) {
      double t = 0.0;
      for (int l = k + 1; l < g.nb; ++l, ++m)
        t = __dadd_rn(t, __dmul_rn(g.cov_off[static_cast<int64_t>(m) * g.ld_model + j], d[l]));
      acc = __dadd_rn(acc, __dmul_rn(t, d[k]));
    }
  }
  chi2 = acc;
  lnp = __dadd_rn(__dsub_rn(-g.lnq[j], __dmul_rn(0.5, acc)), g.lnw[j]);
}

// Diagonal likelihood with the band count known at compile time: the loads of all bands are
// independent of the arithmetic, so the compiler issues them together (same operation order as
// exact_lnp, bit-identical results).
template <int NB>
__device__ __forceinline__ void exact_lnp_fixed(const bf_exact_grid& g, const double* sflux, int64_t j,
                                                double& lnp, double& chi2) {
  double m[NB], v[NB];
#pragma unroll
  for (int b = 0; b < NB; ++b) {
    m[b] = __ldg(g.msb + static_cast<int64_t>(b) * g.ld_model + j);
    v[b] = __ldg(g.ivar + static_cast<int64_t>(b) * g.ld_model + j);
  }
  const double lq = __ldg(g.lnq + j), lw = __ldg(g.lnw + j);
  double acc = 0.0;
#pragma unroll
  for (int b = 0; b < NB; ++b) {
    const double d = __dsub_rn(sflux[b], m[b]);
    acc = __dadd_rn(acc, __dmul_rn(__dmul_rn(d, d), v[b]));
  }
  chi2 = acc;
  lnp = __dadd_rn(__dsub_rn(-lq, __dmul_rn(0.5, acc)), lw);
}

__device__ __forceinline__ bool is_finite(double x) { return fabs(x) <= 1.7976931348623157e308; }

}  // namespace bf
