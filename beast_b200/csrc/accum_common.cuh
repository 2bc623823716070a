// Device helpers shared by the histogram kernels of the sparse pass (accumulate.cu, split.cu).
#pragma once

#include "exact.cuh"
#include "percentile.cuh"

namespace bf {

// expectation of a parameter whose bins each hold one value: q_mid + sum_b H_b (q_b - q_mid) / sum w
// (fit.py:743 regrouped by bin; bin_q holds q_b - q_mid, q_mid = q_lin[2k], so that a
// single-valued parameter comes out exactly); eq holds this lane's part of the sum
__device__ __forceinline__ void store_bin_expectation(double eq, double wtot, bool empty,
                                                      double q_mid, double* dst, int lane) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) eq += __shfl_xor_sync(0xffffffffu, eq, d);
  if (lane == 0) *dst = empty ? __longlong_as_double(0x7ff8000000000000LL) : q_mid + eq / wtot;
}

// per-star reductions over the group partials, fixed order: weight sum (returned), chi2
// minimum and its model (fit.py:689, 725-726), total_log_norm (fit.py:731)
__device__ __forceinline__ double star_totals(uint32_t e0, uint32_t e1, int64_t entry_base,
                                              const bf_accum_scratch& sc, double M, int64_t s,
                                              double* red, long long* redj,
                                              const bf_star_outputs& out) {
  const int nw = blockDim.x >> 5;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double wpart = 0.0, cpart = 1.7976931348623157e308;
  long long cj = -1;
  for (uint32_t e = e0 + threadIdx.x; e < e1; e += blockDim.x) {
    const int64_t slot = static_cast<int64_t>(e) - entry_base;
    wpart += sc.part_wsum[slot];
    const long long oj = sc.part_chi2_j[slot];
    const double oc = sc.part_chi2[slot];
    if (oj >= 0 && (cj < 0 || oc < cpart || (oc == cpart && oj < cj))) {
      cpart = oc;
      cj = oj;
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    wpart += __shfl_xor_sync(0xffffffffu, wpart, d);
    const double oc = __shfl_xor_sync(0xffffffffu, cpart, d);
    const long long oj = __shfl_xor_sync(0xffffffffu, cj, d);
    if (oj >= 0 && (cj < 0 || oc < cpart || (oc == cpart && oj < cj))) {
      cpart = oc;
      cj = oj;
    }
  }
  if (lane == 0) {
    red[warp] = wpart;
    red[nw + warp] = cpart;
    redj[warp] = cj;
  }
  __syncthreads();
  double wtot = 0.0;
  for (int w = 0; w < nw; ++w) wtot += red[w];
  const bool empty = !(wtot > 0.0);
  const double qnan = __longlong_as_double(0x7ff8000000000000LL);
  if (threadIdx.x == 0) {
    double c = 1.7976931348623157e308;
    long long bj = -1;
    for (int w = 0; w < nw; ++w) {
      if (redj[w] >= 0 && (bj < 0 || red[nw + w] < c || (red[nw + w] == c && redj[w] < bj))) {
        c = red[nw + w];
        bj = redj[w];
      }
    }
    out.total_log_norm[s] = empty ? qnan : M + log(wtot);
    out.chi2min[s] = empty ? qnan : c;
    out.chi2min_j[s] = bj;
  }
  return wtot;
}

// percentiles of the finished pdfs (one warp per parameter, fit_metrics/common.py:112-121)
// and the exclusive prefix of the per-group member counts over this star's list
__device__ __forceinline__ void star_percentiles_and_prefix(const bf_param_tables& prm, int64_t s,
                                                            uint32_t e0, uint32_t e1, double p0,
                                                            double p1, double p2, double* cdf_area,
                                                            uint32_t* sscan,
                                                            const bf_star_outputs& out) {
  const int nw = blockDim.x >> 5;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = prm.n_q, tb = prm.total_bins;
  double* cdf = cdf_area + warp * kMaxBinsPerParam;
  for (int kq = warp; kq < nq; kq += nw) {
    const int b0 = prm.bin_off[kq];
    const int nbk = prm.bin_off[kq + 1] - b0;
    warp_percentiles(out.pdf1d + s * tb + b0, prm.bin_val + b0, nbk, cdf, p0, p1, p2,
                     out.per_vals + (s * nq + kq) * 3, lane);
  }

  // ---- per-group member counts -> exclusive prefix over this star's list -----------------
  uint32_t carry = 0;
  for (uint32_t base = e0; base < e1; base += blockDim.x) {
    const uint32_t e = base + threadIdx.x;
    const uint32_t v = e < e1 ? out.member_prefix[e] : 0u;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += t;
    }
    __syncthreads();
    if (lane == 31) sscan[warp] = inc;
    __syncthreads();
    uint32_t woff = 0, tot = 0;
    for (int w = 0; w < nw; ++w) {
      if (w < warp) woff += sscan[w];
      tot += sscan[w];
    }
    if (e < e1) out.member_prefix[e] = carry + woff + inc - v;
    carry += tot;
  }
  if (threadIdx.x == 0) out.n_sparse[s] = carry;
}


// Per-kernel device times of bf_exact_accumulate for a caller that asked for them
// (bf_fit_batch with state->want_timing): the calling thread switches it on, every call then
// synchronises and adds its kernels' times.
struct AccumTimer {
  bool on = false;
  float weights_ms = 0.f, hist_ms = 0.f, exp_ms = 0.f;
};
extern thread_local AccumTimer g_accum_timer;

// ---- the split sparse pass (split.cu), launched from bf_exact_accumulate --------------------
// group-ordered exact weights kernel: same outputs as exact_weights_kernel
int launch_weights_grouped(const bf_exact_grid& grid, const double* flux64, const double* pmax,
                           double threshold, const uint32_t* list, const uint32_t* entry_star,
                           int64_t entry_base, int64_t n_entries, const bf_accum_scratch& sc,
                           const int32_t* fast_q, int fast_pad, const uint8_t* slow_run,
                           const bf_star_outputs& out, int sm_count, cudaStream_t st);
// class-split histogram kernel: histograms, percentiles and all expectations
bool split_hist_supported(const bf_param_tables& prm, size_t smem_budget);
int launch_split_hist(const bf_param_tables& prm, const double* pmax, const uint32_t* offsets,
                      const uint32_t* list, const int32_t* star_order, int64_t star_begin,
                      int64_t n_stars, int64_t entry_base, int64_t n_entries,
                      const bf_accum_scratch& sc, double p0, double p1, double p2,
                      const bf_star_outputs& out, size_t smem_budget, cudaStream_t st);

}  // namespace bf
