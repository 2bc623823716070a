// The sparse pass (K4), three launches.
//
// (1) exact_weights_kernel<NB> -- one warp per listed group, lane <-> model: exact lnP,
//     membership (lnp - max > threshold, beast/fitting/fit.py:670), weight exp(lnp - max)
//     (fit.py:685), per-group weight sum and chi2 minimum (fit.py:689, 725-726).  Tens of
//     thousands of independent warps; the kernel waits on HBM, so the band count is a
//     template parameter and all rows of a model are requested before the first is used.
// (2) the histograms, one block per star, lane <-> parameter: every 8-model chunk that has
//     a member is one coalesced row read (8 bin bytes per lane); each lane adds the weights
//     into its own column of a per-warp private histogram (no atomics, fixed summation
//     order, bit-reproducible); private histograms are merged in warp order, normalised by
//     the weight sum (fit.py:690), and one warp per parameter turns the pdf into
//     p16/p50/p84 (fit_metrics/common.py:112-121).  Two kernels:
//       hist_kernel         full-width columns [bin][lane] (bank = lane), as many warps as fit
//                           (16 at 50 bins, 4 at 200);
//       hist_window_kernel  for wide bin tables: columns sized by the star's occupied bin
//                           ranges, up to 20 warps, models added four at a time; warps left
//                           without a histogram do the star's expectation sums.
// (3) exp_kernel -- sum w*q (fit.py:743) for the stars and parameters the histogram kernel
//     has not covered: no shared-memory state, full occupancy.  Parameters whose bins each
//     hold a single value get their expectation from the histogram instead.

#include <cstdlib>

#include "accum_common.cuh"

namespace bf {

thread_local AccumTimer g_accum_timer;

constexpr int kWeightWarps = 8;
constexpr int kStageDoubles = BF_GROUP;  // per-warp staging: the current group's weights
constexpr int kSideBlocksPerSm = 64;  // grid cap of the grid-stride kernels, in blocks per SM

template <int NB>  // NB > 0: diagonal likelihood with that many bands, loads issued together
__global__ void __launch_bounds__(kWeightWarps * 32, NB > 0 ? 3 : 4)
exact_weights_kernel(bf_exact_grid grid, int64_t n_stars, const double* __restrict__ flux64,
                     const double* __restrict__ pmax, double threshold,
                     const uint32_t* __restrict__ list, const uint32_t* __restrict__ entry_star,
                     int64_t entry_base, int64_t n_entries, bf_accum_scratch sc,
                     uint32_t* __restrict__ member_mask, uint32_t* __restrict__ member_prefix) {
  const int lane = threadIdx.x & 31;
  // grid-stride over the groups
  for (int64_t slot = static_cast<int64_t>(blockIdx.x) * kWeightWarps + (threadIdx.x >> 5);
       slot < n_entries; slot += static_cast<int64_t>(gridDim.x) * kWeightWarps) {
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
      if constexpr (NB > 0) exact_lnp_fixed<NB>(grid, sflux, j, lnp, chi2);
      else exact_lnp(grid, sflux, j, lnp, chi2);
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
}


// shared-memory layout of hist_kernel (doubles), nw = warps per block:
//   hist [nw][rounds][max_bins][32]   private per warp, interleaved: bank = lane
//   wbuf [nw][BF_GROUP]               weights of the group a warp is working on
//   red  [2 nw] (+ int64 [nw]) block reductions, scan [nw] uint32
// after the main loop the histogram area is reused as the percentile scratch (cdf [nw][256])
struct HistSmem {
  static __host__ __device__ size_t bytes(int nw, int rounds, int max_bins) {
    size_t h = static_cast<size_t>(nw) * rounds * max_bins * 32;
    const size_t c = static_cast<size_t>(nw) * kMaxBinsPerParam;
    if (h < c) h = c;
    return (h + static_cast<size_t>(nw) * kStageDoubles + 3 * static_cast<size_t>(nw) + 8) * sizeof(double);
  }
};

// ---------------------------------------------------------------------------------------
// exp_kernel: sum_m w_m * q_k[m] for every parameter (fit.py:743), lane <-> parameter.
// No shared-memory state, so many warps stay resident and hide the row-load latency.
// Also sums the per-group weight partials, so it writes the finished expectations.
// ---------------------------------------------------------------------------------------
constexpr int kExpWarps = 8;

__global__ void __launch_bounds__(kExpWarps * 32)
exp_kernel(bf_param_tables prm, int rounds, const uint32_t* __restrict__ offsets,
           const uint32_t* __restrict__ list, const uint32_t* __restrict__ member_mask,
           int64_t star_begin, int64_t n_stars, int64_t entry_base,
           const double* __restrict__ weights, const double* __restrict__ part_wsum,
           const int32_t* __restrict__ star_flags, double* __restrict__ exp_vals) {
  extern __shared__ __align__(16) double exp_smem[];  // [kExpWarps][nq] + [kExpWarps]
  double* red = exp_smem;
  double* wred = exp_smem + kExpWarps * prm.n_q;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = prm.n_q;
  for (int64_t sb = blockIdx.x; sb < n_stars; sb += gridDim.x) {
  if (star_flags && star_flags[sb]) continue;  // done by the histogram kernel
  const int64_t s = star_begin + sb;
  const uint32_t e0 = offsets[s * BF_SELECT_SPLIT];
  const uint32_t e1 = offsets[(s + 1) * BF_SELECT_SPLIT];
  for (int r = 0; r < rounds; ++r) {
    const int k = r * 32 + lane;
    // parameters whose expectation comes from the histogram are not read here
    const bool live = k < nq && !(prm.q_const && prm.q_const[k]);
    if (!__any_sync(0xffffffffu, live)) continue;
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
        const double2* wq = reinterpret_cast<const double2*>(wp + c * BF_CHUNK);
        const double2 w01 = __ldg(wq), w23 = __ldg(wq + 1), w45 = __ldg(wq + 2), w67 = __ldg(wq + 3);
        if (live) {  // idle lanes touch no row: a row is one 32-byte sector
          const size_t row = (static_cast<size_t>(ch0 + c) * nq + kk) * BF_CHUNK;
          const int4* qp = reinterpret_cast<const int4*>(prm.q_tab + row);
          const int4 qa = __ldg(qp), qb = __ldg(qp + 1);
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
    }
    if (live) red[warp * nq + k] = eacc;
  }
  // the star's weight sum, from the per-group partials (fixed order)
  double wpart = 0.0;
  for (uint32_t e = e0 + threadIdx.x; e < e1; e += blockDim.x)
    wpart += part_wsum[static_cast<int64_t>(e) - entry_base];
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) wpart += __shfl_xor_sync(0xffffffffu, wpart, d);
  if (lane == 0) wred[warp] = wpart;
  __syncthreads();
  double wtot = 0.0;
#pragma unroll
  for (int w = 0; w < kExpWarps; ++w) wtot += wred[w];
  for (int k = threadIdx.x; k < nq; k += blockDim.x) {
    if (prm.q_const && prm.q_const[k]) continue;  // written by the histogram kernel
    double e = 0.0;
#pragma unroll
    for (int w = 0; w < kExpWarps; ++w) e += red[w * nq + k];
    // q = q_lin[2k] + q_lin[2k+1] * fix  ->  Exp = offset + scale * sum(w fix) / sum(w)
    exp_vals[s * nq + k] = wtot > 0.0 ? prm.q_lin[2 * k] + prm.q_lin[2 * k + 1] * (e / wtot)
                                      : __longlong_as_double(0x7ff8000000000000LL);
  }
  __syncthreads();  // red / wred are reused by the next star
  }
}

// ---------------------------------------------------------------------------------------
// hist_kernel.  One block per star, lane <-> parameter, warps split the star's groups.
// ---------------------------------------------------------------------------------------
struct EntryRows {
  uint32_t m0, m1;                               // membership bits
  unsigned long long bits[BF_GROUP / BF_CHUNK];  // this lane's 8 bin bytes of each chunk
  double w0, w1;                                 // weights of models lane, lane+32
};

// rows of a group whose list entry (group id g, membership words) is already in registers:
// no load here depends on another load
template <bool MASKED = false>
__device__ __forceinline__ void load_entry(const bf_param_tables& prm, const double* weights,
                                           int64_t entry_base, uint32_t e, uint32_t g, uint32_t m0,
                                           uint32_t m1, int k, int lane, EntryRows& r,
                                           uint32_t lane_mask = 0xffffffffu) {
  r.m0 = m0;
  r.m1 = m1;
  const double* wp = weights + (static_cast<int64_t>(e) - entry_base) * BF_GROUP;
  r.w0 = __ldg(wp + lane);
  r.w1 = __ldg(wp + 32 + lane);
  // row of chunk c of this group = base + c * (n_q * 8) bytes: one 64-bit product per group
  const uint8_t* base = prm.bin_tab +
      (static_cast<size_t>(g) * (BF_GROUP / BF_CHUNK) * prm.n_q + k) * BF_CHUNK;
  const uint32_t stride = static_cast<uint32_t>(prm.n_q) * BF_CHUNK;
#pragma unroll
  for (int c = 0; c < BF_GROUP / BF_CHUNK; ++c) {
    const uint32_t byte = ((c < 4 ? m0 : m1) >> (8 * (c & 3))) & 0xffu;
    if (byte != 0) {  // warp-uniform: chunks without a member are never read
      uint2 bb = __ldg(reinterpret_cast<const uint2*>(base + c * stride));
      if (MASKED) {  // lanes without a parameter see bin 0 everywhere
        bb.x &= lane_mask;
        bb.y &= lane_mask;
      }
      r.bits[c] = static_cast<unsigned long long>(bb.x) | (static_cast<unsigned long long>(bb.y) << 32);
    }
  }
}

template <bool BATCH>
__global__ void __launch_bounds__(512, 1)
hist_kernel(bf_param_tables prm, int rounds, const double* __restrict__ pmax,
            const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ list,
            const int32_t* __restrict__ star_order, int64_t star_begin, int64_t entry_base,
            bf_accum_scratch sc, double p0, double p1, double p2, bf_star_outputs out) {
  extern __shared__ __align__(16) double smem_d[];
  const int nw = blockDim.x >> 5;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = prm.n_q, tb = prm.total_bins, mb = prm.max_bins;
  const size_t hist_per_warp = static_cast<size_t>(rounds) * mb * 32;
  size_t harea = nw * hist_per_warp;
  if (harea < static_cast<size_t>(nw) * kMaxBinsPerParam) harea = static_cast<size_t>(nw) * kMaxBinsPerParam;

  double* hist = smem_d;
  double* wbuf_all = hist + harea;
  double* red = wbuf_all + nw * kStageDoubles;
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
  double* wbuf = wbuf_all + warp * kStageDoubles;

  for (int r = 0; r < rounds; ++r) {
    const int k = r * 32 + lane;
    const bool live = k < nq;
    const int kk = live ? k : 0;
    double* hr = myhist + static_cast<size_t>(r) * mb * 32;
    const uint32_t hr_s = static_cast<uint32_t>(__cvta_generic_to_shared(hr));
    uint32_t e = e0 + warp;
    EntryRows cur, nxt;
    // three-deep software pipeline: list entry (group id + membership words) two groups
    // ahead, bin rows + weights one group ahead, accumulation on the current group.
    // (Deeper register-ring prefetching and a forwarding read-modify-write pipeline were
    // measured: no gain at 200 bins, a loss at 50 bins where the extra registers cost a
    // resident block -- the kernel is bound by single-warp issue, see DESIGN.md.)
    uint32_t g1 = 0, a1 = 0, b1 = 0, g2 = 0, a2 = 0, b2 = 0;
    if (e < e1) {
      g1 = list[e];
      a1 = out.member_mask[static_cast<size_t>(e) * kWordsPerGroup];
      b1 = out.member_mask[static_cast<size_t>(e) * kWordsPerGroup + 1];
      load_entry(prm, sc.weights, entry_base, e, g1, a1, b1, kk, lane, nxt);
      if (e + nw < e1) {
        g2 = list[e + nw];
        a2 = out.member_mask[static_cast<size_t>(e + nw) * kWordsPerGroup];
        b2 = out.member_mask[static_cast<size_t>(e + nw) * kWordsPerGroup + 1];
      }
    }
    while (e < e1) {
      cur = nxt;
      const uint32_t en = e + nw;
      if (en < e1) {
        load_entry(prm, sc.weights, entry_base, en, g2, a2, b2, kk, lane, nxt);
        const uint32_t en2 = en + nw;
        if (en2 < e1) {
          g2 = list[en2];
          a2 = out.member_mask[static_cast<size_t>(en2) * kWordsPerGroup];
          b2 = out.member_mask[static_cast<size_t>(en2) * kWordsPerGroup + 1];
        }
      }
      if ((cur.m0 | cur.m1) != 0) {
        wbuf[lane] = cur.w0;
        wbuf[32 + lane] = cur.w1;
        __syncwarp();
#pragma unroll
        for (int c = 0; c < BF_GROUP / BF_CHUNK; ++c) {
          const uint32_t byte = ((c < 4 ? cur.m0 : cur.m1) >> (8 * (c & 3))) & 0xffu;
          if (byte == 0) continue;
          if constexpr (!BATCH) {
#pragma unroll
            for (int i = 0; i < BF_CHUNK; ++i) {
              if ((byte >> i) & 1u) {  // warp-uniform
                const int b = static_cast<int>((cur.bits[c] >> (8 * i)) & 0xffull);
                hr[b * 32] += wbuf[c * BF_CHUNK + i];
              }
            }
            continue;
          }
          // the chunk's 8 models four at a time (non-members carry weight 0; lanes beyond n_q
          // own a column too, never read back): four loads in flight, models sharing a bin take
          // the earlier weights along, stores in order -- see add-four in hist_window_kernel
#pragma unroll
          for (int i0 = 0; i0 < BF_CHUNK; i0 += 4) {
            uint32_t a[4];
            double h[4], sw[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint32_t b = static_cast<uint32_t>((cur.bits[c] >> (8 * (i0 + j))) & 0xffull);
              a[j] = hr_s + b * (32 * 8);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) asm volatile("ld.shared.f64 %0, [%1];" : "=d"(h[j]) : "r"(a[j]));
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              sw[j] = wbuf[c * BF_CHUNK + i0 + j];
#pragma unroll
              for (int i = 0; i < j; ++i)
                if (a[i] == a[j]) sw[j] += wbuf[c * BF_CHUNK + i0 + i];
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              h[j] += sw[j];
              asm volatile("st.shared.f64 [%0], %1;" ::"r"(a[j]), "d"(h[j]));
            }
          }
        }
        __syncwarp();
      }
      e = en;
    }
  }

  const double wtot = star_totals(e0, e1, entry_base, sc, M, s, red, redj, out);
  const bool empty = !(wtot > 0.0);

  // ---- merge private histograms (warp order), normalise --------------------------------
  const double inv = empty ? 0.0 : 1.0 / wtot;
  for (int kq = warp; kq < nq; kq += nw) {
    const int b0 = prm.bin_off[kq];
    const int nbk = prm.bin_off[kq + 1] - b0;
    const size_t col = static_cast<size_t>(kq >> 5) * mb * 32 + (kq & 31);
    double* row = out.pdf1d + s * tb + b0;
    const bool from_bins = prm.q_const && prm.q_const[kq];
    double eq = 0.0;
    for (int b = lane; b < nbk; b += 32) {
      double v = 0.0;
      for (int w = 0; w < nw; ++w) v += hist[w * hist_per_warp + col + static_cast<size_t>(b) * 32];
      row[b] = v * inv;
      if (from_bins) eq += v * prm.bin_q[b0 + b];
    }
    if (from_bins)
      store_bin_expectation(eq, wtot, empty, prm.q_lin[2 * kq], out.exp_vals + s * nq + kq, lane);
  }
  __syncthreads();  // histograms are consumed: their memory becomes the percentile scratch
  star_percentiles_and_prefix(prm, s, e0, e1, p0, p1, p2, hist, sscan, out);
}

// ---------------------------------------------------------------------------------------
// hist_window_kernel: the histogram kernel for wide bin tables.  A star's members fall in a
// narrow range of bins of most parameters, so the private histograms only need the occupied
// range [min bin, max bin] of every parameter.  A first walk over the star's bin rows finds
// the ranges (16-bit SIMD min/max, no shared-memory traffic); the private copies are then
// sized by the ranges (see the layout step) instead of 32 x max_bins doubles per warp, and
// as many of the block's 20 warps as fit work on the star (4 with the full-width layout at
// 200 bins).
// Shared memory (doubles): hist [hist_doubles] | wbuf [20][64] | red, redj, scan | epart | int
// tables wmin, wmax, plane, lparam, lcol [rounds*32] | tmpw [32] | misc [16].
// ---------------------------------------------------------------------------------------
constexpr int kWindowWarps = 20;
constexpr int kMinExpWarps = 10;  // idle warps from which a block does the expectation sums itself

struct WindowSmem {
  static __host__ __device__ size_t bytes(int hist_doubles, int rounds) {
    return (static_cast<size_t>(hist_doubles) + kWindowWarps * kStageDoubles + 4 * kWindowWarps +
            static_cast<size_t>(kWindowWarps) * rounds * 32) * sizeof(double) +
           (5 * static_cast<size_t>(rounds) * 32 + 32 + 16) * sizeof(int);
  }
};

// The accumulation loop of hist_window_kernel for one warp and one round of 32 parameters:
// groups e, e + step, ... of the star; col[bin * STRIDE] is this lane's histogram column.
// Same three-deep software pipeline as hist_kernel.
template <int STRIDE>
__device__ __forceinline__ void accumulate_columns(const bf_param_tables& prm, const double* weights,
                                                   int64_t entry_base, const uint32_t* __restrict__ list,
                                                   const uint32_t* __restrict__ member_mask, uint32_t e,
                                                   uint32_t e1, uint32_t step, int kk, uint32_t lmask,
                                                   int lane, double* col, double* wbuf) {
  // 32-bit shared-window address of the column: the update below is written with explicit
  // ld.shared / st.shared so that the compiler neither rebuilds the base per member nor
  // orders the (distinct) weight loads behind the histogram stores
  const uint32_t col_s = static_cast<uint32_t>(__cvta_generic_to_shared(col));
  EntryRows cur, nxt;
  uint32_t g1 = 0, a1 = 0, b1 = 0, g2 = 0, a2 = 0, b2 = 0;
  if (e < e1) {
    g1 = list[e];
    a1 = member_mask[static_cast<size_t>(e) * kWordsPerGroup];
    b1 = member_mask[static_cast<size_t>(e) * kWordsPerGroup + 1];
    load_entry<true>(prm, weights, entry_base, e, g1, a1, b1, kk, lane, nxt, lmask);
    if (e + step < e1) {
      g2 = list[e + step];
      a2 = member_mask[static_cast<size_t>(e + step) * kWordsPerGroup];
      b2 = member_mask[static_cast<size_t>(e + step) * kWordsPerGroup + 1];
    }
  }
  while (e < e1) {
    cur = nxt;
    const uint32_t en = e + step;
    if (en < e1) {
      load_entry<true>(prm, weights, entry_base, en, g2, a2, b2, kk, lane, nxt, lmask);
      const uint32_t en2 = en + step;
      if (en2 < e1) {
        g2 = list[en2];
        a2 = member_mask[static_cast<size_t>(en2) * kWordsPerGroup];
        b2 = member_mask[static_cast<size_t>(en2) * kWordsPerGroup + 1];
      }
    }
    if ((cur.m0 | cur.m1) != 0) {
      wbuf[lane] = cur.w0;
      wbuf[32 + lane] = cur.w1;
      __syncwarp();
#pragma unroll
      for (int c = 0; c < BF_GROUP / BF_CHUNK; ++c) {
        const uint32_t byte = ((c < 4 ? cur.m0 : cur.m1) >> (8 * (c & 3))) & 0xffu;
        if (byte == 0) continue;
        // All 8 models of the chunk are added, four at a time: non-members carry weight 0 and
        // their bins lie inside the ranges too (pass 1 looked at whole chunks).  The four
        // loads are issued together; a model that shares its bin with earlier ones of the
        // four takes their weights along (s_j below, computed while the loads are in flight),
        // so the stores, issued in order, leave the complete sum whichever comes last.
#pragma unroll
        for (int i0 = 0; i0 < BF_CHUNK; i0 += 4) {
          uint32_t a[4];
          double h[4], sw[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint32_t b = static_cast<uint32_t>((cur.bits[c] >> (8 * (i0 + j))) & 0xffull);
            a[j] = col_s + b * (STRIDE * 8);
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) asm volatile("ld.shared.f64 %0, [%1];" : "=d"(h[j]) : "r"(a[j]));
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            sw[j] = wbuf[c * BF_CHUNK + i0 + j];
#pragma unroll
            for (int i = 0; i < j; ++i)
              if (a[i] == a[j]) sw[j] += wbuf[c * BF_CHUNK + i0 + i];
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            h[j] += sw[j];
            asm volatile("st.shared.f64 [%0], %1;" ::"r"(a[j]), "d"(h[j]));
          }
        }
      }
      __syncwarp();
    }
    e = en;
  }
}

// bin rows of a listed group for pass 1 of hist_window_kernel (no weights)
struct BinRows {
  uint32_t m0, m1;
  uint2 b[BF_GROUP / BF_CHUNK];
};

__device__ __forceinline__ void load_bin_rows(const bf_param_tables& prm, uint32_t g, uint32_t m0,
                                              uint32_t m1, int k, BinRows& r) {
  r.m0 = m0;
  r.m1 = m1;
  const uint8_t* base = prm.bin_tab +
      (static_cast<size_t>(g) * (BF_GROUP / BF_CHUNK) * prm.n_q + k) * BF_CHUNK;
  const uint32_t stride = static_cast<uint32_t>(prm.n_q) * BF_CHUNK;
#pragma unroll
  for (int c = 0; c < BF_GROUP / BF_CHUNK; ++c) {
    const uint32_t byte = ((c < 4 ? m0 : m1) >> (8 * (c & 3))) & 0xffu;
    if (byte != 0) r.b[c] = __ldg(reinterpret_cast<const uint2*>(base + c * stride));
  }
}

// Expectation sums of one warp inside hist_window_kernel: sum_m w_m * fix_m for every
// parameter that is not taken from the histogram (fit.py:743), lane <-> parameter, over the
// groups e, e + step, ... of the star.  The rows of a group are requested together.
__device__ __forceinline__ void expectation_rows(const bf_param_tables& prm, int rounds,
                                                 const double* __restrict__ weights,
                                                 int64_t entry_base, const uint32_t* __restrict__ list,
                                                 const uint32_t* __restrict__ member_mask,
                                                 uint32_t e_first, uint32_t e1, uint32_t step,
                                                 int lane, double* epart_w) {
  const int nq = prm.n_q;
  for (int r = 0; r < rounds; ++r) {
    const int k = r * 32 + lane;
    const bool live = k < nq && !(prm.q_const && prm.q_const[k]);
    if (!__any_sync(0xffffffffu, live)) continue;
    const int kk = live ? k : 0;
    double eacc = 0.0;
    uint32_t e = e_first;
    uint32_t g1 = 0, a1 = 0, b1 = 0;
    if (e < e1) {
      g1 = list[e];
      a1 = member_mask[static_cast<size_t>(e) * kWordsPerGroup];
      b1 = member_mask[static_cast<size_t>(e) * kWordsPerGroup + 1];
    }
    while (e < e1) {
      const uint32_t g = g1, m0 = a1, m1 = b1;
      const uint32_t en = e + step;
      if (en < e1) {
        g1 = list[en];
        a1 = member_mask[static_cast<size_t>(en) * kWordsPerGroup];
        b1 = member_mask[static_cast<size_t>(en) * kWordsPerGroup + 1];
      }
      const double* wp = weights + (static_cast<int64_t>(e) - entry_base) * BF_GROUP;
      const int32_t* qbase =
          prm.q_tab + (static_cast<size_t>(g) * (BF_GROUP / BF_CHUNK) * nq + kk) * BF_CHUNK;
      if ((m0 | m1) != 0) {  // warp-uniform
        // the rows of all chunks with a member are requested before the first one is used
        int4 qa[8], qb[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          qa[c] = make_int4(0, 0, 0, 0);
          qb[c] = make_int4(0, 0, 0, 0);
          if ((((c < 4 ? m0 : m1) >> (8 * (c & 3))) & 0xffu) != 0 && live) {  // idle lanes touch no row
            const int4* qp = reinterpret_cast<const int4*>(qbase + static_cast<size_t>(c) * nq * BF_CHUNK);
            qa[c] = __ldg(qp);
            qb[c] = __ldg(qp + 1);
          }
        }
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          if ((((c < 4 ? m0 : m1) >> (8 * (c & 3))) & 0xffu) == 0) continue;
          const double2* wq = reinterpret_cast<const double2*>(wp + c * BF_CHUNK);
          const double2 w01 = __ldg(wq), w23 = __ldg(wq + 1), w45 = __ldg(wq + 2), w67 = __ldg(wq + 3);
          // non-members carry weight 0
          eacc += w01.x * static_cast<double>(qa[c].x);
          eacc += w01.y * static_cast<double>(qa[c].y);
          eacc += w23.x * static_cast<double>(qa[c].z);
          eacc += w23.y * static_cast<double>(qa[c].w);
          eacc += w45.x * static_cast<double>(qb[c].x);
          eacc += w45.y * static_cast<double>(qb[c].y);
          eacc += w67.x * static_cast<double>(qb[c].z);
          eacc += w67.y * static_cast<double>(qb[c].w);
        }
      }
      e = en;
    }
    if (live) epart_w[k] = eacc;
  }
}

__global__ void __launch_bounds__(kWindowWarps * 32, 1)
hist_window_kernel(bf_param_tables prm, int rounds, int hist_doubles,
                   const double* __restrict__ pmax, const uint32_t* __restrict__ offsets,
                   const uint32_t* __restrict__ list, const int32_t* __restrict__ star_order,
                   int64_t star_begin, int64_t entry_base, bf_accum_scratch sc, double p0,
                   double p1, double p2, bf_star_outputs out) {
  extern __shared__ __align__(16) double smem_d[];
  constexpr int NW = kWindowWarps;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = prm.n_q, tb = prm.total_bins;
  const int npad = rounds * 32;

  double* hist = smem_d;
  double* wbuf_all = hist + hist_doubles;
  double* red = wbuf_all + NW * kStageDoubles;
  long long* redj = reinterpret_cast<long long*>(red + 2 * NW);
  uint32_t* sscan = reinterpret_cast<uint32_t*>(redj + NW);
  double* epart = red + 4 * NW;  // [NW][npad] expectation sums of the warps that do them
  int* wmin = reinterpret_cast<int*>(epart + NW * npad);
  int* wmax = wmin + npad;
  int* plane = wmax + npad;    // lane that works on parameter k (within its round of 32)
  int* lparam = plane + npad;  // parameter of a lane, -1: none
  int* lcol = lparam + npad;   // a lane's column: offset of (absolute) bin 0 within a copy
  int* tmpw = lcol + npad;     // [32] range widths by lane
  int* misc = tmpw + 32;       // [0] doubles of one copy of every round, [1] copies per round,
                               // [4+r] doubles of a copy of round r, [8+r] the same summed over earlier rounds

  const int64_t s =
      star_begin + (star_order ? star_order[blockIdx.x] : static_cast<int64_t>(blockIdx.x));
  for (int i = threadIdx.x; i < npad; i += blockDim.x) {
    wmin[i] = 255;
    wmax[i] = 0;
  }
  __syncthreads();

  const double M = pmax[s];
  const uint32_t e0 = offsets[s * BF_SELECT_SPLIT];
  const uint32_t e1 = offsets[(s + 1) * BF_SELECT_SPLIT];

  // ---- pass 1: occupied bin range of every parameter ------------------------------------
  for (int r = 0; r < rounds; ++r) {
    const int k = r * 32 + lane;
    const bool live = k < nq;
    const int kk = live ? k : 0;
    uint32_t mn2 = 0x00ff00ffu, mx2 = 0;
    uint32_t e = e0 + warp;
    // same three-deep pipeline as the accumulation: list entry two groups ahead, bin rows one
    BinRows cur, nxt;
    uint32_t g2 = 0, a2 = 0, b2 = 0;
    nxt.m0 = nxt.m1 = 0;
    if (e < e1) {
      load_bin_rows(prm, list[e], out.member_mask[static_cast<size_t>(e) * kWordsPerGroup],
                    out.member_mask[static_cast<size_t>(e) * kWordsPerGroup + 1], kk, nxt);
      if (e + NW < e1) {
        g2 = list[e + NW];
        a2 = out.member_mask[static_cast<size_t>(e + NW) * kWordsPerGroup];
        b2 = out.member_mask[static_cast<size_t>(e + NW) * kWordsPerGroup + 1];
      }
    }
    while (e < e1) {
      cur = nxt;
      const uint32_t en = e + NW;
      if (en < e1) {
        load_bin_rows(prm, g2, a2, b2, kk, nxt);
        const uint32_t en2 = en + NW;
        if (en2 < e1) {
          g2 = list[en2];
          a2 = out.member_mask[static_cast<size_t>(en2) * kWordsPerGroup];
          b2 = out.member_mask[static_cast<size_t>(en2) * kWordsPerGroup + 1];
        }
      }
#pragma unroll
      for (int c = 0; c < BF_GROUP / BF_CHUNK; ++c) {
        const uint32_t byte = ((c < 4 ? cur.m0 : cur.m1) >> (8 * (c & 3))) & 0xffu;
        if (byte == 0) continue;
        // every model of a chunk with a member counts (a superset of the members' range
        // is all that is needed): bytes -> two 16-bit lanes, native 3-input SIMD min/max
        const uint2 bb = cur.b[c];
        const uint32_t xe = __byte_perm(bb.x, 0, 0x4240), xo = __byte_perm(bb.x, 0, 0x4341);
        const uint32_t ye = __byte_perm(bb.y, 0, 0x4240), yo = __byte_perm(bb.y, 0, 0x4341);
        mn2 = __vimin3_u16x2(__vimin3_u16x2(mn2, xe, xo), ye, yo);
        mx2 = __vimax3_u16x2(__vimax3_u16x2(mx2, xe, xo), ye, yo);
      }
      e = en;
    }
    const int mn = min(mn2 & 0xffffu, mn2 >> 16);
    const int mx = max(mx2 & 0xffffu, mx2 >> 16);
    if (live) {
      atomicMin(wmin + k, mn);
      atomicMax(wmax + k, mx);
    }
  }
  __syncthreads();

  // ---- lay out a private copy --------------------------------------------------------------
  // 16 stacks of doubles, interleaved (stack j holds slots j, j+16, j+32, ...).  Lanes are
  // re-assigned to parameters by range width: lanes 0..15 take the 16 widest ranges in
  // decreasing order, lanes 16..31 the others in increasing order, and lane 16+j stacks its
  // range on top of lane j's.  A half-warp always touches 16 different stacks, so the 64-bit
  // accesses are conflict free whatever the bins are, and the copy holds
  // 16 * max_j (width_j + width_{31-j}) doubles -- close to the sum of the widths.
  // Lanes without a parameter get a one-bin range of their own.
  // With more than 32 parameters every round of 32 gets its own copies and its own warps
  // (n per round), so that a warp walks the star's groups once.
  if (warp == 0) {
    int total = 0;
    for (int r = 0; r < rounds; ++r) {
      const int k = r * 32 + lane;
      const bool occupied = k < nq && wmin[k] <= wmax[k];
      const int mn = occupied ? wmin[k] : 0;
      const int w = occupied ? wmax[k] - wmin[k] + 1 : 1;
      int rank = 0;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const int wj = __shfl_sync(0xffffffffu, w, j);
        rank += (wj > w || (wj == w && j < lane)) ? 1 : 0;
      }
      const int slot = rank < 16 ? rank : 47 - rank;
      tmpw[slot] = w;
      __syncwarp();
      const int below = slot >= 16 ? tmpw[slot - 16] : 0;
      const int depth = __reduce_max_sync(0xffffffffu, lane < 16 ? tmpw[lane] + tmpw[lane + 16] : 0);
      __syncwarp();
      lparam[r * 32 + slot] = k < nq ? k : -1;
      lcol[r * 32 + slot] = (below - mn) * 16 + (slot & 15);
      plane[k] = slot;
      if (lane == 0) {
        misc[4 + r] = depth * 16;
        misc[8 + r] = total;
      }
      total += depth * 16;
    }
    if (lane == 0) {
      const int fit = hist_doubles / total, most = NW / rounds;
      int n = fit < most ? fit : most;
      // several rounds make the expectation kernel expensive (it walks the star once per
      // round): rather give up a copy or two per round and keep kMinExpWarps warps for the
      // expectation sums in this block
      const int n_in = (NW - kMinExpWarps) / rounds;
      if (rounds > 1 && sc.star_flags != nullptr && n > n_in && n < 2 * n_in) n = n_in;
      misc[0] = total;
      misc[1] = n;
    }
  }
  __syncthreads();
  const int copy = misc[0], npr = misc[1];  // doubles of one copy of every round, copies per round
  const int nwe = npr * rounds;             // warps that accumulate
  // A star with broad ranges leaves warps without a private copy: from kMinExpWarps on they
  // take the star's expectation sums, and the expectation kernel skips the star.
  const int n_exp = NW - nwe;
  const bool exp_here = sc.star_flags != nullptr && n_exp >= kMinExpWarps;
  if (threadIdx.x == 0 && sc.star_flags) sc.star_flags[s - star_begin] = exp_here ? 1 : 0;
  for (int i = threadIdx.x; i < npr * copy; i += blockDim.x) hist[i] = 0.0;
  __syncthreads();

  // ---- pass 2: accumulate; warp w works on round w / npr, copy w % npr -----------------------
  if (warp < nwe) {
    const int r = warp / npr, i = warp - r * npr;
    const int k = lparam[r * 32 + lane];
    const bool live = k >= 0;  // the others read bin 0 everywhere and own a one-bin range
    accumulate_columns<16>(prm, sc.weights, entry_base, list, out.member_mask, e0 + i, e1, npr,
                           live ? k : 0, live ? 0xffffffffu : 0u, lane,
                           hist + npr * misc[8 + r] + i * misc[4 + r] + lcol[r * 32 + lane],
                           wbuf_all + warp * kStageDoubles);
  } else if (exp_here) {
    expectation_rows(prm, rounds, sc.weights, entry_base, list, out.member_mask,
                     e0 + (warp - nwe), e1, n_exp, lane, epart + (warp - nwe) * npad);
  }

  const double wtot = star_totals(e0, e1, entry_base, sc, M, s, red, redj, out);
  const bool empty = !(wtot > 0.0);
  if (exp_here) {
    // q = q_lin[2k] + q_lin[2k+1] * fix  ->  Exp = offset + scale * sum(w fix) / sum(w)
    for (int k = threadIdx.x; k < nq; k += blockDim.x) {
      if (prm.q_const && prm.q_const[k]) continue;  // written with the merged histogram below
      double e = 0.0;
      for (int w = 0; w < n_exp; ++w) e += epart[w * npad + k];
      out.exp_vals[s * nq + k] = empty ? __longlong_as_double(0x7ff8000000000000LL)
                                       : prm.q_lin[2 * k] + prm.q_lin[2 * k + 1] * (e / wtot);
    }
  }

  // ---- merge the private copies (warp order), normalise; bins outside the range are 0 -----
  const double inv = empty ? 0.0 : 1.0 / wtot;
  for (int kq = warp; kq < nq; kq += NW) {
    const int b0 = prm.bin_off[kq];
    const int nbk = prm.bin_off[kq + 1] - b0;
    const int mn = wmin[kq];
    const int width = wmax[kq] >= mn ? wmax[kq] - mn + 1 : 0;
    const int copy_r = misc[4 + (kq >> 5)];
    const double* src = hist + npr * misc[8 + (kq >> 5)] + lcol[(kq & ~31) + plane[kq]];
    double* row = out.pdf1d + s * tb + b0;
    const bool from_bins = prm.q_const && prm.q_const[kq];
    double eq = 0.0;
    for (int b = lane; b < nbk; b += 32) {
      double v = 0.0;
      if (b >= mn && b < mn + width)
        for (int w = 0; w < npr; ++w) v += src[w * copy_r + b * 16];
      row[b] = v * inv;
      if (from_bins) eq += v * prm.bin_q[b0 + b];
    }
    if (from_bins)
      store_bin_expectation(eq, wtot, empty, prm.q_lin[2 * kq], out.exp_vals + s * nq + kq, lane);
  }
  __syncthreads();  // histograms are consumed: their memory becomes the percentile scratch
  star_percentiles_and_prefix(prm, s, e0, e1, p0, p1, p2, hist, sscan, out);
}

}  // namespace bf

extern "C" int bf_exact_accumulate(const bf_exact_grid* grid, const bf_param_tables* params,
                                   int64_t star_begin, int64_t n_stars, const double* flux64,
                                   const double* pmax, double threshold, const uint32_t* offsets,
                                   const uint32_t* list, const uint32_t* entry_star,
                                   const int32_t* star_order, int64_t entry_base, int64_t n_entries,
                                   const double* percentiles_host, const bf_accum_scratch* scratch,
                                   const bf_star_outputs* out, void* stream, void* aux_stream) {
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
  // BF_TIMING=1: synchronous per-kernel timing of this call on stderr (diagnostics only)
  static const bool env_timing = getenv("BF_TIMING") != nullptr;
  const bool timing = env_timing || g_accum_timer.on;
  cudaEvent_t tev[5] = {};
  if (timing) {
    for (auto& e : tev) cudaEventCreate(&e);
    cudaEventRecord(tev[0], st);
  }

  int dev = 0;
  BF_CUDA_TRY(cudaGetDevice(&dev));
  static int smem_limit[64] = {};
  static int sm_count[64] = {};
  if (dev >= 64) {
    set_error("bf_exact_accumulate: device ordinal %d not supported", dev);
    return BF_ERR_UNSUPPORTED;
  }
  if (smem_limit[dev] == 0) {
    int optin = 0;
    BF_CUDA_TRY(cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    BF_CUDA_TRY(cudaFuncSetAttribute(hist_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    BF_CUDA_TRY(cudaFuncSetAttribute(hist_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    // same L1/shared split as the histogram kernel, or the two cannot share an SM
    BF_CUDA_TRY(cudaFuncSetAttribute(exp_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                     cudaSharedmemCarveoutMaxShared));
    BF_CUDA_TRY(cudaFuncSetAttribute(hist_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                     cudaSharedmemCarveoutMaxShared));
    BF_CUDA_TRY(cudaFuncSetAttribute(hist_kernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                     cudaSharedmemCarveoutMaxShared));
    BF_CUDA_TRY(cudaFuncSetAttribute(hist_window_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
    BF_CUDA_TRY(cudaFuncSetAttribute(hist_window_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                     cudaSharedmemCarveoutMaxShared));
    BF_CUDA_TRY(cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev));
    smem_limit[dev] = optin;
  }
  // The class-split histogram kernel needs the class tables, room for the chunk lists and the
  // by-products of the group-ordered weights kernel (chunk sums, expectation partials of the
  // fast class); BF_HIST=full|window and BF_WEIGHTS=star select the other kernels.
  const char* hist_env = getenv("BF_HIST");
  const char* wenv = getenv("BF_WEIGHTS");
  const bool grouped = scratch->group_order != nullptr && grid->cov_off == nullptr && grid->nb <= 8 &&
                       !(wenv && wenv[0] == 's') && !getenv("BF_WEIGHTS_GENERIC");
  const bool want_split = grouped && scratch->chunk_refs && scratch->chunk_wsum &&
                          (scratch->exp_part || params->n_fast == 0) &&
                          split_hist_supported(*params, static_cast<size_t>(smem_limit[dev])) &&
                          !(hist_env && (hist_env[0] == 'f' || hist_env[0] == 'w'));
  if (n_entries > 0) {
    const int64_t blocks = (n_entries + kWeightWarps - 1) / kWeightWarps;
    BF_REQUIRE(blocks < (1LL << 31), "bf_exact_accumulate: too many list entries in one call");
    const int64_t cap = static_cast<int64_t>(sm_count[dev]) * kSideBlocksPerSm;
    const unsigned wgrid = static_cast<unsigned>(blocks < cap ? blocks : cap);
    // diagonal likelihood: one instantiation per band count (all band loads in flight at once,
    // 1.4x faster than the runtime loop); full covariance and > 12 bands take the generic one
    int nbk = grid->cov_off == nullptr && grid->nb <= 12 ? grid->nb : 0;
    if (getenv("BF_WEIGHTS_GENERIC")) nbk = 0;  // measurements
    // with the list positions ordered by group the rows of a group are fetched once for all
    // its stars (two models' rows per lane: up to 8 bands fit the register file without spills)
    if (grouped) {
      const int rc = launch_weights_grouped(*grid, flux64, pmax, threshold, list, entry_star,
                                            entry_base, n_entries, *scratch,
                                            want_split ? params->fast_q : nullptr,
                                            want_split && params->n_fast > 0 ? params->fast_pad : 0,
                                            want_split ? params->slow_run : nullptr, *out,
                                            sm_count[dev], st);
      if (rc != BF_OK) return rc;
    } else {
#define BF_WEIGHTS_CASE(N)                                                                      \
  case N:                                                                                       \
    exact_weights_kernel<N><<<wgrid, kWeightWarps * 32, 0, st>>>(                               \
        *grid, n_stars, flux64, pmax, threshold, list, entry_star, entry_base, n_entries,       \
        *scratch, out->member_mask, out->member_prefix);                                        \
    break;
    switch (nbk) {
      BF_WEIGHTS_CASE(1) BF_WEIGHTS_CASE(2) BF_WEIGHTS_CASE(3) BF_WEIGHTS_CASE(4)
      BF_WEIGHTS_CASE(5) BF_WEIGHTS_CASE(6) BF_WEIGHTS_CASE(7) BF_WEIGHTS_CASE(8)
      BF_WEIGHTS_CASE(9) BF_WEIGHTS_CASE(10) BF_WEIGHTS_CASE(11) BF_WEIGHTS_CASE(12)
      default:
        BF_WEIGHTS_CASE(0)
    }
#undef BF_WEIGHTS_CASE
    BF_CHECK_LAUNCH("exact_weights_kernel");
    }
  }
  if (timing) cudaEventRecord(tev[1], st);
  const int rounds = (params->n_q + 31) / 32;
  // widest block whose private histograms fit; two resident blocks per SM when they are small
  const size_t budget = static_cast<size_t>(smem_limit[dev]);
  int nw = 16;
  while (nw > 1 && HistSmem::bytes(nw, rounds, params->max_bins) + 4096 > budget) --nw;
  // with 16 warps on an SM the one-member-at-a-time update is as fast; with fewer, adding four
  // models at a time pays (100 bins: 682 -> 658 ms per C2 step, 50 bins: 611 vs 617 ms)
  const bool batch4 = nw < 16 && !getenv("BF_HIST_SERIAL");
  if (nw == 16 && HistSmem::bytes(8, rounds, params->max_bins) + 2048 <= budget / 2) nw = 8;
  const size_t smem = HistSmem::bytes(nw, rounds, params->max_bins);
  if (smem + 1024 > budget) {
    set_error("bf_exact_accumulate: %d bins x %d parameters need %zu B of shared memory (> %zu)",
              params->max_bins, params->n_q, smem, budget);
    return BF_ERR_TOO_LARGE;
  }
  // wide bin tables leave the full-width layout few warps: use the windowed kernel there
  // (BF_HIST=full|window overrides the choice, for measurements and tests)
  const size_t wbytes_total = budget - 1024;
  const int hist_doubles = static_cast<int>(
      (wbytes_total - WindowSmem::bytes(0, rounds)) / sizeof(double)) & ~1;
  // worst case of one private copy: every range as wide as the widest bin table
  const bool window_fits = hist_doubles >= 32 * params->max_bins * rounds &&
                           hist_doubles >= kWindowWarps * kMaxBinsPerParam;
  bool use_window = window_fits && nw < 8;
  if (hist_env && hist_env[0] == 'f') use_window = false;
  if (hist_env && hist_env[0] == 'w') use_window = window_fits;
  // the class-split kernel needs the class tables, room for the chunk lists and, like the
  // windowed kernel, one worst-case private copy per class; it does every expectation itself
  if (want_split) {
    const int rc = launch_split_hist(*params, pmax, offsets, list, star_order, star_begin, n_stars,
                                     entry_base, n_entries, *scratch, percentiles_host[0], percentiles_host[1],
                                     percentiles_host[2], *out, budget, st);
    if (rc != BF_OK) return rc;
    if (timing) {
      cudaEventRecord(tev[2], st);
      cudaStreamSynchronize(st);
      float w = 0, h = 0;
      cudaEventElapsedTime(&w, tev[0], tev[1]);
      cudaEventElapsedTime(&h, tev[1], tev[2]);
      g_accum_timer.weights_ms += w;
      g_accum_timer.hist_ms += h;
      if (env_timing)
        fprintf(stderr, "[bf timing] stars=%lld entries=%lld weights=%.2f ms (%s) split hist=%.2f ms\n",
                static_cast<long long>(n_stars), static_cast<long long>(n_entries), w,
                scratch->group_order ? "grouped" : "by star", h);
      for (auto& e : tev) cudaEventDestroy(e);
    }
    return BF_OK;
  }
  BF_REQUIRE(params->bin_tab && params->q_tab,
             "bf_exact_accumulate: the per-model tables bin_tab / q_tab are needed (no class split "
             "usable for this call)");
  if (timing && (g_accum_timer.on || getenv("BF_TIMING")[0] == '2')) aux_stream = nullptr;  // each kernel alone
  // the windowed kernel tells the expectation kernel which stars it has done: same stream
  if (use_window) aux_stream = nullptr;
  cudaStream_t sx = aux_stream ? static_cast<cudaStream_t>(aux_stream) : st;
  static cudaEvent_t ev_weights[64] = {}, ev_exp[64] = {};
  if (aux_stream) {  // the weights are ready at this point of `stream`
    if (!ev_weights[dev]) {
      BF_CUDA_TRY(cudaEventCreateWithFlags(&ev_weights[dev], cudaEventDisableTiming));
      BF_CUDA_TRY(cudaEventCreateWithFlags(&ev_exp[dev], cudaEventDisableTiming));
    }
    BF_CUDA_TRY(cudaEventRecord(ev_weights[dev], st));
  }
  if (use_window) {
    hist_window_kernel<<<static_cast<unsigned>(n_stars), kWindowWarps * 32,
                         WindowSmem::bytes(hist_doubles, rounds), st>>>(
        *params, rounds, hist_doubles, pmax, offsets, list, star_order, star_begin, entry_base,
        *scratch, percentiles_host[0], percentiles_host[1], percentiles_host[2], *out);
    BF_CHECK_LAUNCH("hist_window_kernel");
  } else {
    auto hk = batch4 ? hist_kernel<true> : hist_kernel<false>;
    hk<<<static_cast<unsigned>(n_stars), nw * 32, smem, st>>>(
        *params, rounds, pmax, offsets, list, star_order, star_begin, entry_base, *scratch,
        percentiles_host[0], percentiles_host[1], percentiles_host[2], *out);
    BF_CHECK_LAUNCH("hist_kernel");
  }
  if (timing) cudaEventRecord(tev[2], st);
  // The expectation kernel is bound by L2 bandwidth and the histogram kernel by shared-memory
  // latency, and together they fit on an SM; with an auxiliary stream they run side by side
  // (the histogram blocks are launched first so that they claim their shared memory).
  if (aux_stream) BF_CUDA_TRY(cudaStreamWaitEvent(sx, ev_weights[dev], 0));
  if (timing) cudaEventRecord(tev[3], sx);
  {
    const int64_t cap = static_cast<int64_t>(sm_count[dev]) * kSideBlocksPerSm;
    exp_kernel<<<static_cast<unsigned>(n_stars < cap ? n_stars : cap), kExpWarps * 32,
                 (kExpWarps * params->n_q + kExpWarps) * sizeof(double), sx>>>(
        *params, rounds, offsets, list, out->member_mask, star_begin, n_stars, entry_base,
        scratch->weights, scratch->part_wsum, use_window ? scratch->star_flags : nullptr,
        out->exp_vals);
    BF_CHECK_LAUNCH("exp_kernel");
  }
  if (aux_stream) BF_CUDA_TRY(cudaEventRecord(ev_exp[dev], sx));
  if (aux_stream) BF_CUDA_TRY(cudaStreamWaitEvent(st, ev_exp[dev], 0));
  if (timing) {
    cudaEventRecord(tev[4], sx);
    cudaStreamSynchronize(sx);
    cudaStreamSynchronize(st);
    float w = 0, h = 0, x = 0, tot = 0;
    cudaEventElapsedTime(&w, tev[0], tev[1]);
    cudaEventElapsedTime(&h, tev[1], tev[2]);
    cudaEventElapsedTime(&x, tev[3], tev[4]);
    cudaEventElapsedTime(&tot, tev[0], tev[4]);
    g_accum_timer.weights_ms += w;
    g_accum_timer.hist_ms += h;
    g_accum_timer.exp_ms += x;
    if (env_timing)
      fprintf(stderr, "[bf timing] stars=%lld entries=%lld weights=%.2f ms hist=%.2f ms exp=%.2f ms "
              "(exp end - start of call = %.2f ms, hist warps=%d%s)\n",
              static_cast<long long>(n_stars), static_cast<long long>(n_entries), w, h, x, tot,
              use_window ? kWindowWarps : nw, use_window ? " windowed" : "");
    for (auto& e : tev) cudaEventDestroy(e);
  }
  return BF_OK;
}
