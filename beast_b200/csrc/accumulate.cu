// K4: the sparse pass.  One thread block per star, warps take the star's selected groups
// round-robin.  Two phases per group:
//   A  lane <-> model : exact lnP, membership (lnp - max > threshold, fit.py:670), weight
//      exp(lnp - max) (fit.py:685), running weight sum and chi2 minimum (fit.py:689, 725-726);
//      membership bits go to member_mask, weights to a per-warp shared buffer.
//   B  lane <-> parameter : for every 8-model chunk that has a member, each lane reads its
//      parameter's 8 bin bytes (one coalesced 256 B row per warp) and adds the weights into
//      its own row of a per-warp private histogram -- no atomics, no conflicts, and a fixed
//      summation order (results are bit-reproducible run to run).  Expectation sums
//      sum w*q accumulate in registers (fit.py:743).
// After the loop the private histograms are merged in warp order, normalised by the weight
// sum (fit.py:690), and one warp per parameter turns the pdf into p16/p50/p84 with a
// warp-parallel CDF scan that reproduces np.interp on (cumsum - w/2)/sum
// (fit_metrics/common.py:112-121).

#include "exact.cuh"
#include "percentile.cuh"

namespace bf {

constexpr int kMaxRounds = BF_MAX_PARAMS / 32;

struct AccumSmem {
  // layout (doubles unless noted), nw = warps per block:
  //   hist   [nw][total_bins]
  //   wbuf   [nw][BF_GROUP]
  //   cdf    [nw][kMaxBinsPerParam]
  //   rexp   [nw][n_q]
  //   rw     [nw], rchi [nw], rj [nw] (int64)
  //   sflux  [BF_MAX_BANDS]
  //   scan   [nw] uint32 (+pad)
  static __host__ __device__ size_t bytes(int nw, int total_bins, int n_q) {
    size_t d = static_cast<size_t>(nw) * total_bins + static_cast<size_t>(nw) * BF_GROUP +
               static_cast<size_t>(nw) * kMaxBinsPerParam + static_cast<size_t>(nw) * n_q +
               3 * static_cast<size_t>(nw) + BF_MAX_BANDS + 8;
    return d * sizeof(double);
  }
};

__global__ void __launch_bounds__(256)
accumulate_kernel(bf_exact_grid grid, bf_param_tables prm, const double* __restrict__ flux64,
                  const double* __restrict__ pmax, double threshold,
                  const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ list,
                  double p0, double p1, double p2, bf_star_outputs out) {
  extern __shared__ __align__(16) double smem_d[];
  const int nw = blockDim.x >> 5;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = prm.n_q, tb = prm.total_bins;

  double* hist = smem_d;                                   // [nw][tb]
  double* wbuf = hist + static_cast<size_t>(nw) * tb;      // [nw][BF_GROUP]
  double* cdf = wbuf + nw * BF_GROUP;                      // [nw][kMaxBinsPerParam]
  double* rexp = cdf + nw * kMaxBinsPerParam;              // [nw][nq]
  double* rw = rexp + nw * nq;                             // [nw]
  double* rchi = rw + nw;                                  // [nw]
  long long* rj = reinterpret_cast<long long*>(rchi + nw); // [nw]
  double* sflux = reinterpret_cast<double*>(rj + nw);      // [BF_MAX_BANDS]
  uint32_t* sscan = reinterpret_cast<uint32_t*>(sflux + BF_MAX_BANDS);  // [16]

  const int64_t s = blockIdx.x;
  for (int i = threadIdx.x; i < nw * tb; i += blockDim.x) hist[i] = 0.0;
  if (threadIdx.x < grid.nb) sflux[threadIdx.x] = flux64[s * grid.nb + threadIdx.x];
  __syncthreads();

  const double M = pmax[s];
  const uint32_t e0 = offsets[s * BF_SELECT_SPLIT];
  const uint32_t e1 = offsets[(s + 1) * BF_SELECT_SPLIT];

  double* myhist = hist + static_cast<size_t>(warp) * tb;
  double* wb = wbuf + warp * BF_GROUP;

  int boff[kMaxRounds];
  double eacc[kMaxRounds];
#pragma unroll
  for (int r = 0; r < kMaxRounds; ++r) {
    const int k = r * 32 + lane;
    boff[r] = k < nq ? prm.bin_off[k] : 0;
    eacc[r] = 0.0;
  }

  double wsum = 0.0;
  double chimin = 1.7976931348623157e308;
  long long chij = -1;

  for (uint32_t e = e0 + warp; e < e1; e += nw) {
    const uint32_t g = list[e];
    uint32_t mw[kWordsPerGroup];
    // ---- phase A: lane <-> model -------------------------------------------------
#pragma unroll
    for (int h = 0; h < kWordsPerGroup; ++h) {
      const int64_t j = static_cast<int64_t>(g) * BF_GROUP + h * 32 + lane;
      bool member = false;
      double w = 0.0;
      if (j < grid.n_models) {
        double lnp, chi2;
        exact_lnp(grid, sflux, j, lnp, chi2);
        member = (lnp - M) > threshold;  // NaN compares false, like np.where (fit.py:670)
        if (member) {
          w = exp(lnp - M);
          wsum += w;
          if (chi2 < chimin) {  // ascending j per lane: first minimum wins within the lane
            chimin = chi2;
            chij = j;
          }
        }
      }
      wb[h * 32 + lane] = w;
      mw[h] = __ballot_sync(0xffffffffu, member);
    }
    if (lane == 0) {
      uint32_t cnt = 0;
#pragma unroll
      for (int h = 0; h < kWordsPerGroup; ++h) {
        out.member_mask[static_cast<size_t>(e) * kWordsPerGroup + h] = mw[h];
        cnt += __popc(mw[h]);
      }
      out.member_prefix[e] = cnt;  // turned into an exclusive prefix below
    }
    __syncwarp();
    // ---- phase B: lane <-> parameter ---------------------------------------------
#pragma unroll
    for (int h = 0; h < kWordsPerGroup; ++h) {
      if (mw[h] == 0) continue;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        const uint32_t byte = (mw[h] >> (8 * c)) & 0xffu;
        if (byte == 0) continue;
        const int64_t ch = static_cast<int64_t>(g) * (BF_GROUP / BF_CHUNK) + h * 4 + c;
        const double* wc = wb + h * 32 + c * 8;
#pragma unroll
        for (int r = 0; r < kMaxRounds; ++r) {
          const int k = r * 32 + lane;
          if (r * 32 < nq && k < nq) {
            const size_t row = (static_cast<size_t>(ch) * nq + k) * BF_CHUNK;
            const uint2 bb = *reinterpret_cast<const uint2*>(prm.bin_tab + row);
            const unsigned long long bits =
                static_cast<unsigned long long>(bb.x) | (static_cast<unsigned long long>(bb.y) << 32);
            const double* qp = prm.q_tab + row;
            double* hrow = myhist + boff[r];
#pragma unroll
            for (int i = 0; i < BF_CHUNK; ++i) {
              if ((byte >> i) & 1u) {
                const double w = wc[i];
                const int b = static_cast<int>((bits >> (8 * i)) & 0xffull);
                hrow[b] += w;
                eacc[r] += w * qp[i];
              }
            }
          }
        }
      }
    }
    __syncwarp();
  }

  // ---- block-level merges, all in fixed order ------------------------------------
#pragma unroll
  for (int r = 0; r < kMaxRounds; ++r) {
    const int k = r * 32 + lane;
    if (k < nq) rexp[warp * nq + k] = eacc[r];
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    wsum += __shfl_xor_sync(0xffffffffu, wsum, d);
    const double oc = __shfl_xor_sync(0xffffffffu, chimin, d);
    const long long oj = __shfl_xor_sync(0xffffffffu, chij, d);
    if (oj >= 0 && (chij < 0 || oc < chimin || (oc == chimin && oj < chij))) {
      chimin = oc;
      chij = oj;
    }
  }
  if (lane == 0) {
    rw[warp] = wsum;
    rchi[warp] = chimin;
    rj[warp] = chij;
  }
  __syncthreads();

  double wtot = 0.0;
  for (int w = 0; w < nw; ++w) wtot += rw[w];
  const bool empty = !(wtot > 0.0);
  const double qnan = __longlong_as_double(0x7ff8000000000000LL);

  if (threadIdx.x == 0) {
    double c = 1.7976931348623157e308;
    long long cj = -1;
    for (int w = 0; w < nw; ++w) {
      if (rj[w] >= 0 && (cj < 0 || rchi[w] < c || (rchi[w] == c && rj[w] < cj))) {
        c = rchi[w];
        cj = rj[w];
      }
    }
    out.total_log_norm[s] = empty ? qnan : M + log(wtot);
    out.chi2min[s] = empty ? qnan : c;
    out.chi2min_j[s] = cj;
  }
  for (int k = threadIdx.x; k < nq; k += blockDim.x) {
    double e = 0.0;
    for (int w = 0; w < nw; ++w) e += rexp[w * nq + k];
    out.exp_vals[s * nq + k] = empty ? qnan : e / wtot;
  }
  // merged, normalised pdf lands in warp 0's histogram and in the output row
  const double inv = empty ? 0.0 : 1.0 / wtot;
  for (int i = threadIdx.x; i < tb; i += blockDim.x) {
    double v = 0.0;
    for (int w = 0; w < nw; ++w) v += hist[static_cast<size_t>(w) * tb + i];
    v *= inv;
    out.pdf1d[s * tb + i] = v;
    hist[i] = v;  // thread i only ever touches column i: no hazard
  }
  __syncthreads();
  for (int k = warp; k < nq; k += nw) {
    const int b0 = prm.bin_off[k];
    const int nb = prm.bin_off[k + 1] - b0;
    warp_percentiles(hist + b0, prm.bin_val + b0, nb, cdf + warp * kMaxBinsPerParam, p0, p1, p2,
                     out.per_vals + (s * nq + k) * 3, lane);
  }

  // ---- per-group member counts -> exclusive prefix over this star's list ------------
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

}  // namespace bf

extern "C" int bf_exact_accumulate(const bf_exact_grid* grid, const bf_param_tables* params,
                                   int64_t n_stars, const double* flux64, const double* pmax,
                                   double threshold, const uint32_t* offsets, const uint32_t* list,
                                   const double* percentiles_host, const bf_star_outputs* out,
                                   void* stream) {
  using namespace bf;
  BF_REQUIRE(grid && params && out && percentiles_host, "bf_exact_accumulate: null struct");
  BF_REQUIRE(grid->nb >= 1 && grid->nb <= BF_MAX_BANDS, "bf_exact_accumulate: bad band count");
  BF_REQUIRE(params->n_q >= 1 && params->n_q <= BF_MAX_PARAMS,
             "bf_exact_accumulate: n_q=%d out of range (1..%d)", params->n_q, BF_MAX_PARAMS);
  BF_REQUIRE(params->total_bins >= params->n_q, "bf_exact_accumulate: bad total_bins");
  BF_REQUIRE(n_stars > 0 && n_stars < (1LL << 31), "bf_exact_accumulate: bad n_stars");
  BF_REQUIRE(flux64 && pmax && offsets && list, "bf_exact_accumulate: null pointer");
  BF_REQUIRE(out->total_log_norm && out->chi2min && out->chi2min_j && out->n_sparse &&
                 out->exp_vals && out->pdf1d && out->per_vals && out->member_mask &&
                 out->member_prefix,
             "bf_exact_accumulate: null output pointer");

  int dev = 0;
  BF_CUDA_TRY(cudaGetDevice(&dev));
  static int smem_limit[64] = {};
  if (dev >= 64) {
    set_error("bf_exact_accumulate: device ordinal %d not supported", dev);
    return BF_ERR_UNSUPPORTED;
  }
  if (smem_limit[dev] == 0) {
    int optin = 0;
    BF_CUDA_TRY(cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    BF_CUDA_TRY(cudaFuncSetAttribute(accumulate_kernel,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    smem_limit[dev] = optin;
  }
  // widest block whose private histograms fit; prefer two resident blocks per SM
  int nw = 8;
  const size_t budget = static_cast<size_t>(smem_limit[dev]);
  while (nw > 1 && AccumSmem::bytes(nw, params->total_bins, params->n_q) > budget / 2) nw >>= 1;
  const size_t smem = AccumSmem::bytes(nw, params->total_bins, params->n_q);
  if (smem > budget) {
    set_error("bf_exact_accumulate: %d bins over %d parameters need %zu B of shared memory (> %zu)",
              params->total_bins, params->n_q, smem, budget);
    return BF_ERR_TOO_LARGE;
  }
  accumulate_kernel<<<static_cast<unsigned>(n_stars), nw * 32, smem,
                      static_cast<cudaStream_t>(stream)>>>(
      *grid, *params, flux64, pmax, threshold, offsets, list, percentiles_host[0],
      percentiles_host[1], percentiles_host[2], *out);
  BF_CHECK_LAUNCH("accumulate_kernel");
  return BF_OK;
}
