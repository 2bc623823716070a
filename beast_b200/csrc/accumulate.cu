// The sparse pass (K4), two launches.
//
// (1) exact_weights_kernel -- one warp per listed group, lane <-> model: exact lnP,
//     membership (lnp - max > threshold, beast/fitting/fit.py:670), weight exp(lnp - max)
//     (fit.py:685), per-group weight sum and chi2 minimum (fit.py:689, 725-726).  Perfectly
//     load balanced and latency tolerant (tens of thousands of independent warps).
// (2) hist_kernel -- one block per star, lane <-> parameter: every 8-model chunk that has
//     a member is one coalesced row read (8 bin bytes per lane);
//     each lane adds the members' weights into its own column of a per-warp private
//     histogram laid out [bin][lane] (bank = lane: conflict free, no atomics, fixed summation
//     order, bit-reproducible).  The next group's rows are fetched while the current one is
//     processed.  exp_kernel does the same walk for sum w*q (fit.py:743) without any
//     shared-memory state, so it runs at full occupancy.  Private
//     histograms are merged in warp order, normalised by the weight sum (fit.py:690), and
//     one warp per parameter turns the pdf into p16/p50/p84 (fit_metrics/common.py:112-121).

#include "exact.cuh"
#include "percentile.cuh"

namespace bf {

constexpr int kMaxRounds = BF_MAX_PARAMS / 32;
constexpr int kWeightWarps = 8;

__global__ void __launch_bounds__(kWeightWarps * 32)
exact_weights_kernel(bf_exact_grid grid, int64_t n_stars, const double* __restrict__ flux64,
                     const double* __restrict__ pmax, double threshold,
                     const uint32_t* __restrict__ list, const uint32_t* __restrict__ entry_star,
                     int64_t entry_base, int64_t n_entries, bf_accum_scratch sc,
                     uint32_t* __restrict__ member_mask, uint32_t* __restrict__ member_prefix) {
  const int lane = threadIdx.x & 31;
  const int64_t slot = static_cast<int64_t>(blockIdx.x) * kWeightWarps + (threadIdx.x >> 5);
  if (slot >= n_entries) return;
  const uint32_t e = static_cast<uint32_t>(entry_base + slot);
  const int64_t s = entry_star[e];
  const double* sflux = flux64 + s * grid.nb;
  const double M = pmax[s];
  const uint32_t g = list[e];

  double wsum = 0.0;
  double chimin = 1.7976931348623157e308;
  long long chij = -1;
  uint32_t cnt = 0;
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
        if (chi2 < chimin) {
          chimin = chi2;
          chij = j;
        }
      }
    }
    sc.weights[slot * BF_GROUP + h * 32 + lane] = w;
    const uint32_t mw = __ballot_sync(0xffffffffu, member);
    cnt += __popc(mw);
    if (lane == 0) member_mask[static_cast<size_t>(e) * kWordsPerGroup + h] = mw;
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
    sc.part_wsum[slot] = wsum;
    sc.part_chi2[slot] = chimin;
    sc.part_chi2_j[slot] = chij;
    member_prefix[e] = cnt;  // turned into an exclusive prefix by hist_kernel
  }
}


// shared-memory layout of hist_kernel (doubles), nw = warps per block:
//   hist [nw][rounds][max_bins][32]   private, interleaved: bank = lane
//   cdf  [nw][kMaxBinsPerParam]
//   wbuf [nw][BF_GROUP]   weights of the group a warp is working on
//   red  [nw] (+ int64 [nw]) for the block reductions, scan [nw] uint32
struct HistSmem {
  static __host__ __device__ size_t bytes(int nw, int rounds, int max_bins, int n_q) {
    size_t d = static_cast<size_t>(nw) * rounds * max_bins * 32 +
               static_cast<size_t>(nw) * kMaxBinsPerParam + static_cast<size_t>(nw) * BF_GROUP +
               3 * static_cast<size_t>(nw) + 8;
    return d * sizeof(double);
  }
};

// ---------------------------------------------------------------------------------------
// exp_kernel: sum_m w_m * q_k[m] for every parameter (fit.py:743), lane <-> parameter.
// No shared-memory state, so many warps stay resident and hide the row-load latency.
// Writes the raw sums; hist_kernel divides by the weight sum.
// ---------------------------------------------------------------------------------------
constexpr int kExpWarps = 8;

__global__ void __launch_bounds__(kExpWarps * 32)
exp_kernel(bf_param_tables prm, int rounds, const uint32_t* __restrict__ offsets,
           const uint32_t* __restrict__ list, const uint32_t* __restrict__ member_mask,
           int64_t star_begin, int64_t entry_base, const double* __restrict__ weights,
           double* __restrict__ exp_raw) {
  __shared__ double red[kExpWarps][BF_MAX_PARAMS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = prm.n_q;
  const int64_t s = star_begin + blockIdx.x;
  const uint32_t e0 = offsets[s * BF_SELECT_SPLIT];
  const uint32_t e1 = offsets[(s + 1) * BF_SELECT_SPLIT];
  for (int r = 0; r < rounds; ++r) {
    const int k = r * 32 + lane;
    const bool live = k < nq;
    const int kk = live ? k : 0;
    double eacc = 0.0;
    for (uint32_t e = e0 + warp; e < e1; e += kExpWarps) {
      const uint32_t m0 = member_mask[static_cast<size_t>(e) * kWordsPerGroup];
      const uint32_t m1 = member_mask[static_cast<size_t>(e) * kWordsPerGroup + 1];
      if ((m0 | m1) == 0) continue;
      const double* wp = weights + (static_cast<int64_t>(e) - entry_base) * BF_GROUP;
      const int64_t ch0 = static_cast<int64_t>(list[e]) * (BF_GROUP / BF_CHUNK);
#pragma unroll
      for (int c = 0; c < BF_GROUP / BF_CHUNK; ++c) {
        const uint32_t byte = ((c < 4 ? m0 : m1) >> (8 * (c & 3))) & 0xffu;
        if (byte == 0) continue;
        const size_t row = (static_cast<size_t>(ch0 + c) * nq + kk) * BF_CHUNK;
        const int4* qp = reinterpret_cast<const int4*>(prm.q_tab + row);
        const double2* wq = reinterpret_cast<const double2*>(wp + c * BF_CHUNK);
        const int4 qa = __ldg(qp), qb = __ldg(qp + 1);
        const double2 w01 = __ldg(wq), w23 = __ldg(wq + 1), w45 = __ldg(wq + 2), w67 = __ldg(wq + 3);
        // non-members carry weight 0
        eacc += w01.x * static_cast<double>(qa.x);
        eacc += w01.y * static_cast<double>(qa.y);
        eacc += w23.x * static_cast<double>(qa.z);
        eacc += w23.y * static_cast<double>(qa.w);
        eacc += w45.x * static_cast<double>(qb.x);
        eacc += w45.y * static_cast<double>(qb.y);
        eacc += w67.x * static_cast<double>(qb.z);
        eacc += w67.y * static_cast<double>(qb.w);
      }
    }
    if (live) red[warp][k] = eacc;
  }
  __syncthreads();
  for (int k = threadIdx.x; k < nq; k += blockDim.x) {
    double e = 0.0;
#pragma unroll
    for (int w = 0; w < kExpWarps; ++w) e += red[w][k];
    exp_raw[s * nq + k] = e;
  }
}

// one listed group as the histogram loop sees it
struct EntryRows {
  uint32_t m0, m1;                                   // membership bits
  unsigned long long bits[BF_GROUP / BF_CHUNK];      // this lane's 8 bin bytes of each chunk
  double w0, w1;                                     // weights of models lane, lane+32
};

__device__ __forceinline__ void load_entry(const bf_param_tables& prm, const uint32_t* list,
                                           const uint32_t* member_mask, const double* weights,
                                           int64_t entry_base, uint32_t e, int k, int lane,
                                           EntryRows& r) {
  r.m0 = member_mask[static_cast<size_t>(e) * kWordsPerGroup];
  r.m1 = member_mask[static_cast<size_t>(e) * kWordsPerGroup + 1];
  const double* wp = weights + (static_cast<int64_t>(e) - entry_base) * BF_GROUP;
  r.w0 = __ldg(wp + lane);
  r.w1 = __ldg(wp + 32 + lane);
  const int64_t ch0 = static_cast<int64_t>(list[e]) * (BF_GROUP / BF_CHUNK);
#pragma unroll
  for (int c = 0; c < BF_GROUP / BF_CHUNK; ++c) {
    const size_t row = (static_cast<size_t>(ch0 + c) * prm.n_q + k) * BF_CHUNK;
    const uint2 bb = __ldg(reinterpret_cast<const uint2*>(prm.bin_tab + row));
    r.bits[c] = static_cast<unsigned long long>(bb.x) | (static_cast<unsigned long long>(bb.y) << 32);
  }
}

__global__ void __launch_bounds__(512, 1)
hist_kernel(bf_param_tables prm, int rounds, const double* __restrict__ pmax,
            const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ list,
            const int32_t* __restrict__ star_order, int64_t star_begin, int64_t entry_base,
            bf_accum_scratch sc,
            double p0, double p1, double p2, bf_star_outputs out) {
  extern __shared__ __align__(16) double smem_d[];
  const int nw = blockDim.x >> 5;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = prm.n_q, tb = prm.total_bins, mb = prm.max_bins;
  const size_t hist_per_warp = static_cast<size_t>(rounds) * mb * 32;

  double* hist = smem_d;
  double* cdf = hist + nw * hist_per_warp;
  double* red_w = cdf + nw * kMaxBinsPerParam;
  double* red = red_w + nw * BF_GROUP;
  long long* redj = reinterpret_cast<long long*>(red + 2 * nw);
  uint32_t* sscan = reinterpret_cast<uint32_t*>(redj + nw);

  const int64_t s =
      star_begin + (star_order ? star_order[blockIdx.x] : static_cast<int64_t>(blockIdx.x));
  for (size_t i = threadIdx.x; i < nw * hist_per_warp; i += blockDim.x) hist[i] = 0.0;
  __syncthreads();

  const double M = pmax[s];
  const uint32_t e0 = offsets[s * BF_SELECT_SPLIT];
  const uint32_t e1 = offsets[(s + 1) * BF_SELECT_SPLIT];
  double* myhist = hist + warp * hist_per_warp + lane;

  double* wbuf = red_w + warp * BF_GROUP;  // this warp's copy of the current group's weights
  for (int r = 0; r < rounds; ++r) {
    const int k = r * 32 + lane;
    const bool live = k < nq;
    const int kk = live ? k : 0;
    double* hr = myhist + static_cast<size_t>(r) * mb * 32;
    uint32_t e = e0 + warp;
    EntryRows cur, nxt;
    if (e < e1) load_entry(prm, list, out.member_mask, sc.weights, entry_base, e, kk, lane, nxt);
    while (e < e1) {
      cur = nxt;
      const uint32_t en = e + nw;
      // the next group's rows travel while this one is accumulated
      if (en < e1) load_entry(prm, list, out.member_mask, sc.weights, entry_base, en, kk, lane, nxt);
      if ((cur.m0 | cur.m1) != 0) {
        wbuf[lane] = cur.w0;
        wbuf[32 + lane] = cur.w1;
        __syncwarp();
#pragma unroll
        for (int c = 0; c < BF_GROUP / BF_CHUNK; ++c) {
          const uint32_t byte = ((c < 4 ? cur.m0 : cur.m1) >> (8 * (c & 3))) & 0xffu;
          if (byte == 0) continue;
#pragma unroll
          for (int i = 0; i < BF_CHUNK; ++i) {
            if ((byte >> i) & 1u) {  // warp-uniform
              const double w = wbuf[c * BF_CHUNK + i];
              if (live) {
                const int b = static_cast<int>((cur.bits[c] >> (8 * i)) & 0xffull);
                hr[b * 32] += w;
              }
            }
          }
        }
        __syncwarp();
      }
      e = en;
    }
  }

  // ---- per-star reductions over the group partials, fixed order ------------------------
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
  for (int k = threadIdx.x; k < nq; k += blockDim.x)  // exp_kernel left sum w * fix here
    out.exp_vals[s * nq + k] = empty ? qnan : prm.q_lin[2 * k] + prm.q_lin[2 * k + 1] * (out.exp_vals[s * nq + k] / wtot);

  // ---- merge private histograms (warp order), normalise, percentiles ---------------------
  const double inv = empty ? 0.0 : 1.0 / wtot;
  for (int k = warp; k < nq; k += nw) {
    const int b0 = prm.bin_off[k];
    const int nbk = prm.bin_off[k + 1] - b0;
    const size_t col = static_cast<size_t>(k >> 5) * mb * 32 + (k & 31);
    double* row = out.pdf1d + s * tb + b0;
    for (int b = lane; b < nbk; b += 32) {
      double v = 0.0;
      for (int w = 0; w < nw; ++w) v += hist[w * hist_per_warp + col + static_cast<size_t>(b) * 32];
      row[b] = v * inv;
    }
    __syncwarp();
    warp_percentiles(row, prm.bin_val + b0, nbk, cdf + warp * kMaxBinsPerParam, p0, p1, p2,
                     out.per_vals + (s * nq + k) * 3, lane);
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

}  // namespace bf

extern "C" int bf_exact_accumulate(const bf_exact_grid* grid, const bf_param_tables* params,
                                   int64_t star_begin, int64_t n_stars, const double* flux64,
                                   const double* pmax, double threshold, const uint32_t* offsets,
                                   const uint32_t* list,
                                   const uint32_t* entry_star, const int32_t* star_order,
                                   int64_t entry_base, int64_t n_entries, const double* percentiles_host,
                                   const bf_accum_scratch* scratch, const bf_star_outputs* out,
                                   void* stream) {
  using namespace bf;
  BF_REQUIRE(grid && params && out && scratch && percentiles_host, "bf_exact_accumulate: null struct");
  BF_REQUIRE(grid->nb >= 1 && grid->nb <= BF_MAX_BANDS, "bf_exact_accumulate: bad band count");
  BF_REQUIRE(params->n_q >= 1 && params->n_q <= BF_MAX_PARAMS,
             "bf_exact_accumulate: n_q=%d out of range (1..%d)", params->n_q, BF_MAX_PARAMS);
  BF_REQUIRE(params->total_bins >= params->n_q && params->max_bins >= 1 &&
                 params->max_bins <= kMaxBinsPerParam,
             "bf_exact_accumulate: bad bin counts");
  BF_REQUIRE(n_stars > 0 && n_stars < (1LL << 31) && star_begin >= 0,
             "bf_exact_accumulate: bad star range");
  BF_REQUIRE(entry_base >= 0 && n_entries >= 0, "bf_exact_accumulate: bad entry range");
  BF_REQUIRE(flux64 && pmax && offsets && list && entry_star, "bf_exact_accumulate: null pointer");
  BF_REQUIRE(scratch->weights && scratch->part_wsum && scratch->part_chi2 && scratch->part_chi2_j,
             "bf_exact_accumulate: null scratch pointer");
  BF_REQUIRE(out->total_log_norm && out->chi2min && out->chi2min_j && out->n_sparse &&
                 out->exp_vals && out->pdf1d && out->per_vals && out->member_mask &&
                 out->member_prefix,
             "bf_exact_accumulate: null output pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);

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
    BF_CUDA_TRY(cudaFuncSetAttribute(hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    smem_limit[dev] = optin;
  }
  if (n_entries > 0) {
    const int64_t blocks = (n_entries + kWeightWarps - 1) / kWeightWarps;
    BF_REQUIRE(blocks < (1LL << 31), "bf_exact_accumulate: too many list entries in one call");
    exact_weights_kernel<<<static_cast<unsigned>(blocks), kWeightWarps * 32, 0, st>>>(
        *grid, n_stars, flux64, pmax, threshold, list, entry_star, entry_base, n_entries, *scratch,
        out->member_mask, out->member_prefix);
    BF_CHECK_LAUNCH("exact_weights_kernel");
  }
  const int rounds = (params->n_q + 31) / 32;
  exp_kernel<<<static_cast<unsigned>(n_stars), kExpWarps * 32, 0, st>>>(
      *params, rounds, offsets, list, out->member_mask, star_begin, entry_base, scratch->weights,
      out->exp_vals);
  BF_CHECK_LAUNCH("exp_kernel");
  // widest block whose private histograms fit; prefer >= 2 resident blocks per SM
  const size_t budget = static_cast<size_t>(smem_limit[dev]);
  int nw = 8;
  while (nw > 1 && HistSmem::bytes(nw, rounds, params->max_bins, params->n_q) > budget / 2) nw >>= 1;
  if (nw < 4) {  // too few warps for two blocks: take the whole SM with one block instead
    nw = 16;
    while (nw > 1 && HistSmem::bytes(nw, rounds, params->max_bins, params->n_q) > budget) --nw;
  }
  const size_t smem = HistSmem::bytes(nw, rounds, params->max_bins, params->n_q);
  if (smem > budget) {
    set_error("bf_exact_accumulate: %d bins x %d parameter rounds need %zu B of shared memory (> %zu)",
              params->max_bins, rounds, smem, budget);
    return BF_ERR_TOO_LARGE;
  }
  hist_kernel<<<static_cast<unsigned>(n_stars), nw * 32, smem, st>>>(
      *params, rounds, pmax, offsets, list, star_order, star_begin, entry_base, *scratch,
      percentiles_host[0],
      percentiles_host[1], percentiles_host[2], *out);
  BF_CHECK_LAUNCH("hist_kernel");
  return BF_OK;
}
