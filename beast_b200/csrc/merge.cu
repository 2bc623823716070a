// Device side of the subgrid merge (beast/tools/subgridding_tools.py:392-527).
//
// Each rank fitted the same stars against its own subgrid with do_not_normalize=True.
// The relative weight of subgrid g for a star is exp(total_log_norm_g - max_g)
// (:394-406).  The caller all-gathers total_log_norm (NCCL), bf_merge_weights turns it
// into weights and the merged total_log_norm (:513-514), bf_scale_rows applies a rank's
// weight to its pdf1d / expectation rows just before the cross-rank sum (:421-426,
// :462-470), and bf_merge_finalize renormalises the summed pdfs (:428-434) and
// recomputes the percentiles from them (:472-491).

#include "percentile.cuh"

namespace bf {

__global__ void merge_weights_kernel(int n_sub, int64_t n_stars, const double* __restrict__ tln,
                                     double* __restrict__ weight, double* __restrict__ merged) {
  const int64_t s = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (s >= n_stars) return;
  // np.argmax semantics: NaN wins (first NaN), otherwise first maximum
  double mx = tln[s];
  for (int g = 1; g < n_sub; ++g) {
    const double v = tln[static_cast<int64_t>(g) * n_stars + s];
    if (!(mx != mx) && (v > mx || v != v)) mx = v;
  }
  double sum = 0.0;
  for (int g = 0; g < n_sub; ++g) {
    const double w = exp(tln[static_cast<int64_t>(g) * n_stars + s] - mx);
    weight[static_cast<int64_t>(g) * n_stars + s] = w;
    sum += w;
  }
  merged[s] = log(sum) + mx;
}

__global__ void scale_rows_kernel(int64_t n_stars, int64_t row_len,
                                  const double* __restrict__ weight, double* __restrict__ rows) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n_stars * row_len) return;
  rows[i] *= weight[i / row_len];
}

// dst[s][:] = (accumulate ? dst[s][:] : 0) + weight[s] * src[s][:]  -- a subgrid's rows, scaled
// by its weight, join the sum that then crosses the ranks (:421-426, :462-470)
__global__ void axpy_rows_kernel(int64_t n_stars, int64_t row_len, const double* __restrict__ weight,
                                 const double* __restrict__ src, double* __restrict__ dst,
                                 int accumulate) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n_stars * row_len) return;
  const double w = weight[i / row_len];
  // a subgrid whose weight is exactly 0 must not turn the sum into NaN (inf * 0)
  const double v = w != 0.0 ? w * src[i] : 0.0;
  dst[i] = accumulate ? dst[i] + v : v;
}

constexpr int kFinWarps = 4;

// one block per star, one warp per parameter
__global__ void __launch_bounds__(kFinWarps * 32)
merge_finalize_kernel(bf_param_tables prm, double* __restrict__ pdf1d, double p0, double p1,
                      double p2, double* __restrict__ per_vals) {
  __shared__ double cdf[kFinWarps][kMaxBinsPerParam];
  const int64_t s = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int k = warp; k < prm.n_q; k += kFinWarps) {
    const int b0 = prm.bin_off[k];
    const int nb = prm.bin_off[k + 1] - b0;
    double* row = pdf1d + s * prm.total_bins + b0;
    double part = 0.0;
    for (int i = lane; i < nb; i += 32) part += row[i];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) part += __shfl_xor_sync(0xffffffffu, part, d);
    if (part > 0.0) {
      for (int i = lane; i < nb; i += 32) row[i] /= part;
    }
    __syncwarp();
    warp_percentiles(row, prm.bin_val + b0, nb, cdf[warp], p0, p1, p2,
                     per_vals + (s * prm.n_q + k) * 3, lane);
  }
}

}  // namespace bf

extern "C" int bf_merge_weights(int n_sub, int64_t n_stars, const double* tln_all,
                                double* weight_all, double* merged_tln, void* stream) {
  using namespace bf;
  BF_REQUIRE(n_sub >= 1 && n_stars > 0, "bf_merge_weights: bad sizes");
  BF_REQUIRE(tln_all && weight_all && merged_tln, "bf_merge_weights: null pointer");
  const unsigned blocks = static_cast<unsigned>((n_stars + 255) / 256);
  merge_weights_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      n_sub, n_stars, tln_all, weight_all, merged_tln);
  BF_CHECK_LAUNCH("merge_weights_kernel");
  return BF_OK;
}

extern "C" int bf_scale_rows(int64_t n_stars, int64_t row_len, const double* weight, double* rows,
                             void* stream) {
  using namespace bf;
  BF_REQUIRE(n_stars > 0 && row_len > 0, "bf_scale_rows: bad sizes");
  BF_REQUIRE(weight && rows, "bf_scale_rows: null pointer");
  const int64_t n = n_stars * row_len;
  BF_REQUIRE((n + 255) / 256 < (1LL << 31), "bf_scale_rows: too large");
  scale_rows_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0,
                      static_cast<cudaStream_t>(stream)>>>(n_stars, row_len, weight, rows);
  BF_CHECK_LAUNCH("scale_rows_kernel");
  return BF_OK;
}

extern "C" int bf_merge_finalize(const bf_param_tables* params, int64_t n_stars, double* pdf1d,
                                 const double* percentiles_host, double* per_vals, void* stream) {
  using namespace bf;
  BF_REQUIRE(params && params->n_q >= 1 && params->bin_off && params->bin_val,
             "bf_merge_finalize: bad params");
  BF_REQUIRE(n_stars > 0 && n_stars < (1LL << 31), "bf_merge_finalize: bad n_stars");
  BF_REQUIRE(pdf1d && percentiles_host && per_vals, "bf_merge_finalize: null pointer");
  merge_finalize_kernel<<<static_cast<unsigned>(n_stars), kFinWarps * 32, 0,
                          static_cast<cudaStream_t>(stream)>>>(
      *params, pdf1d, percentiles_host[0], percentiles_host[1], percentiles_host[2], per_vals);
  BF_CHECK_LAUNCH("merge_finalize_kernel");
  return BF_OK;
}

extern "C" int bf_axpy_rows(int64_t n_stars, int64_t row_len, const double* weight, const double* src,
                            double* dst, int accumulate, void* stream) {
  using namespace bf;
  BF_REQUIRE(n_stars > 0 && row_len > 0, "bf_axpy_rows: bad sizes");
  BF_REQUIRE(weight && src && dst, "bf_axpy_rows: null pointer");
  const int64_t n = n_stars * row_len;
  BF_REQUIRE((n + 255) / 256 < (1LL << 31), "bf_axpy_rows: too large");
  axpy_rows_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      n_stars, row_len, weight, src, dst, accumulate);
  BF_CHECK_LAUNCH("axpy_rows_kernel");
  return BF_OK;
}
