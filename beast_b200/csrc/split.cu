// The split sparse pass: the two kernels bf_exact_accumulate prefers when the caller supplies
// what they need (scratch->group_order, scratch->chunk_refs, the class tables of bf_param_tables).
//
// (A) exact_weights_grouped_kernel<NB> -- the same arithmetic and the same outputs as
//     exact_weights_kernel (beast/fitting/fit.py:670, 685, 689, 725-726), but the list positions
//     are visited in group order (bf_group_order).  A warp keeps the FP64 rows of a group's 64
//     models in registers and serves every star of its stretch of the order with them: about a
//     thousand stars of a batch list the same group, so the 7 KB of rows come out of HBM once
//     per stretch instead of once per (star, group).
//
// (B) split_hist_kernel -- histograms, percentiles and expectations (fit.py:732-755), one block
//     per star.  Parameters come in two classes (gridprep.FitSetup._class_split):
//       fast: one shared-memory update and one expectation term per model.  With <= 16 fast
//             parameters a warp works on two chunks at a time (half-warp <-> chunk, lane <->
//             parameter), so every update instruction has all 32 lanes busy;
//       slow: one value per chunk -> one update and one expectation term per chunk, with the
//             chunk's weight sum, i.e. an eighth of the work.
//     The block first writes the list of the star's chunks that hold a member (chunk id + place
//     of its 8 weights), finds every parameter's occupied bin range over those chunks, lays out
//     private histogram copies sized by the ranges (16 interleaved stacks, conflict free for
//     64-bit accesses, as in hist_window_kernel), and then fast warps and slow warps walk the
//     chunk list.  Sums are in fixed order: results are bit-reproducible.

#include <cstdlib>

#include "accum_common.cuh"

namespace bf {

// =========================================================================================
// bf_group_order: counting sort of list positions by group id
// =========================================================================================
__global__ void __launch_bounds__(256)
group_count_kernel(const uint32_t* __restrict__ list, int64_t entry_base, int64_t n_entries,
                   uint32_t* __restrict__ counts) {
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n_entries;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x)
    atomicAdd(counts + list[entry_base + i], 1u);
}

__global__ void __launch_bounds__(256)
group_scatter_kernel(const uint32_t* __restrict__ list, int64_t entry_base, int64_t n_entries,
                     const uint32_t* __restrict__ goff, uint32_t* __restrict__ cursor,
                     uint32_t* __restrict__ order) {
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n_entries;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const uint32_t g = list[entry_base + i];
    const uint32_t pos = atomicAdd(cursor + g, 1u);
    order[goff[g] + pos] = static_cast<uint32_t>(entry_base + i);
  }
}

// =========================================================================================
// (A) group-ordered exact weights
// =========================================================================================
constexpr int kGroupedWarps = 8;
constexpr int kOrderStretch = 128;  // positions of the group order one warp walks in one go

// Per-warp shared memory of the grouped kernel when it also forms the expectation partials of
// the fast parameters: the weights of up to 8 list positions (A operand) and the group's
// fixed-point parameter values (B operand) of the FP64 tensor-core products below.
constexpr int kWtLd = BF_GROUP + 4;  // row pitch of the weight tile: conflict-free fragments
template <int NT>
struct GroupedSmem {
  static constexpr int kLdq = NT * 8 + 4;
  static constexpr size_t kWarpBytes =
      NT == 0 ? 0 : 8 * kWtLd * sizeof(double) + BF_GROUP * kLdq * sizeof(int32_t) + 8 * sizeof(int32_t);
};

// D[8 list positions][8 parameters] += A[8][4 models] * B[4 models][8 parameters]  (FP64 tensor core)
__device__ __forceinline__ void dmma_8x8x4(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

// sum_m w_m * fix_m[k] (fit.py:743) for the pending list positions of a warp: rows of wt are
// list positions, qg holds the group's 64 x (NT*8) fixed-point values
template <int NT>
__device__ __forceinline__ void flush_expectations(const double* wt, const int32_t* qg,
                                                   const int32_t* slots, int pending,
                                                   double* __restrict__ exp_part, int lane) {
  constexpr int kLdq = GroupedSmem<NT>::kLdq;
  __syncwarp();
  double acc[NT][2];
#pragma unroll
  for (int n = 0; n < NT; ++n) acc[n][0] = acc[n][1] = 0.0;
  const int r = lane >> 2, c = lane & 3;
#pragma unroll 4
  for (int t = 0; t < BF_GROUP / 4; ++t) {
    const double a = wt[r * kWtLd + 4 * t + c];
#pragma unroll
    for (int n = 0; n < NT; ++n) {
      const double b = static_cast<double>(qg[(4 * t + c) * kLdq + n * 8 + r]);
      dmma_8x8x4(acc[n][0], acc[n][1], a, b);
    }
  }
  if (r < pending) {
    double* dst = exp_part + static_cast<size_t>(slots[r]) * (NT * 8) + c * 2;
#pragma unroll
    for (int n = 0; n < NT; ++n)
      *reinterpret_cast<double2*>(dst + n * 8) = make_double2(acc[n][0], acc[n][1]);
  }
  __syncwarp();
}

// NT: 8-parameter tiles of the fast class whose expectation partials are formed here (0: none)
template <int NB, int NT>
__global__ void __launch_bounds__(kGroupedWarps * 32, 2)
exact_weights_grouped_kernel(bf_exact_grid grid, const double* __restrict__ flux64,
                             const double* __restrict__ pmax, double threshold,
                             const uint32_t* __restrict__ list,
                             const uint32_t* __restrict__ entry_star,
                             const uint32_t* __restrict__ order, int64_t entry_base,
                             int64_t n_entries, bf_accum_scratch sc,
                             const int32_t* __restrict__ fast_q,
                             const uint8_t* __restrict__ slow_run,
                             uint32_t* __restrict__ member_mask,
                             uint32_t* __restrict__ member_prefix) {
  extern __shared__ __align__(16) unsigned char grouped_smem[];
  const int lane = threadIdx.x & 31;
  constexpr int kLdq = GroupedSmem<NT>::kLdq;
  unsigned char* mine = grouped_smem + (threadIdx.x >> 5) * GroupedSmem<NT>::kWarpBytes;
  double* wt = reinterpret_cast<double*>(mine);
  int32_t* qg = reinterpret_cast<int32_t*>(wt + 8 * kWtLd);
  int32_t* slots = qg + BF_GROUP * kLdq;
  const int64_t n_stretch = (n_entries + kOrderStretch - 1) / kOrderStretch;
  for (int64_t st = static_cast<int64_t>(blockIdx.x) * kGroupedWarps + (threadIdx.x >> 5);
       st < n_stretch; st += static_cast<int64_t>(gridDim.x) * kGroupedWarps) {
    const int64_t i0 = st * kOrderStretch;
    const int64_t i1 = i0 + kOrderStretch < n_entries ? i0 + kOrderStretch : n_entries;
    uint32_t cur_g = 0xffffffffu;
    double m0[NB], v0[NB], m1[NB], v1[NB];
    double lq0 = 0.0, lw0 = 0.0, lq1 = 0.0, lw1 = 0.0;
    bool ok0 = false, ok1 = false;
    int64_t j0 = 0;
    int pending = 0;  // list positions whose weights wait in wt for the expectation products
    uint32_t run_bits = 0xffu;  // chunks of the group that start a run of equal slow values
    for (int64_t blk = i0; blk < i1; blk += 32) {
      const int cnt = static_cast<int>(i1 - blk < 32 ? i1 - blk : 32);
      // every lane fetches the description of one list position of this block of 32
      uint32_t my_e = 0, my_s = 0, my_g = 0;
      double my_M = 0.0;
      if (lane < cnt) {
        my_e = order[blk + lane];
        my_s = entry_star[my_e];
        my_g = list[my_e];
        my_M = pmax[my_s];
      }
      for (int t = 0; t < cnt; ++t) {
        const uint32_t e = __shfl_sync(0xffffffffu, my_e, t);
        const uint32_t s = __shfl_sync(0xffffffffu, my_s, t);
        const uint32_t g = __shfl_sync(0xffffffffu, my_g, t);
        const double M = __shfl_sync(0xffffffffu, my_M, t);
        double f[NB];
        const double* sflux = flux64 + static_cast<int64_t>(s) * NB;
#pragma unroll
        for (int b = 0; b < NB; ++b) f[b] = __ldg(sflux + b);
        if (g != cur_g) {  // warp-uniform: the rows of the next group
          if constexpr (NT > 0) {
            if (pending) flush_expectations<NT>(wt, qg, slots, pending, sc.exp_part, lane);
            pending = 0;
            // fast_q rows of the group: [8 chunks][2][pad][4] int32 -> qg[model][parameter]
            const int4* src = reinterpret_cast<const int4*>(fast_q) +
                              static_cast<size_t>(g) * (BF_GROUP / 4) * (NT * 8);
#pragma unroll
            for (int it = 0; it < NT * 4; ++it) {
              const int v = lane + 32 * it;
              const int4 q4 = __ldg(src + v);
              const int quad = v / (NT * 8), p = v % (NT * 8);
              qg[(quad * 4 + 0) * kLdq + p] = q4.x;
              qg[(quad * 4 + 1) * kLdq + p] = q4.y;
              qg[(quad * 4 + 2) * kLdq + p] = q4.z;
              qg[(quad * 4 + 3) * kLdq + p] = q4.w;
            }
          }
          cur_g = g;
          run_bits = slow_run ? (__ldg(slow_run + g) | 1u) : 0xffu;
          j0 = static_cast<int64_t>(g) * BF_GROUP + lane;
          ok0 = j0 < grid.n_models;
          ok1 = j0 + 32 < grid.n_models;
#pragma unroll
          for (int b = 0; b < NB; ++b) {
            m0[b] = __ldg(grid.msb + static_cast<int64_t>(b) * grid.ld_model + j0);
            v0[b] = __ldg(grid.ivar + static_cast<int64_t>(b) * grid.ld_model + j0);
            m1[b] = __ldg(grid.msb + static_cast<int64_t>(b) * grid.ld_model + j0 + 32);
            v1[b] = __ldg(grid.ivar + static_cast<int64_t>(b) * grid.ld_model + j0 + 32);
          }
          lq0 = __ldg(grid.lnq + j0);
          lw0 = __ldg(grid.lnw + j0);
          lq1 = __ldg(grid.lnq + j0 + 32);
          lw1 = __ldg(grid.lnw + j0 + 32);
        }
        // the reference's operation order (likelihood.py:47-55, 149-165; fit.py:668), as in
        // exact_lnp_fixed
        double a0 = 0.0, a1 = 0.0;
#pragma unroll
        for (int b = 0; b < NB; ++b) {
          const double d0 = __dsub_rn(f[b], m0[b]);
          const double d1 = __dsub_rn(f[b], m1[b]);
          a0 = __dadd_rn(a0, __dmul_rn(__dmul_rn(d0, d0), v0[b]));
          a1 = __dadd_rn(a1, __dmul_rn(__dmul_rn(d1, d1), v1[b]));
        }
        const double lnp0 = __dadd_rn(__dsub_rn(-lq0, __dmul_rn(0.5, a0)), lw0);
        const double lnp1 = __dadd_rn(__dsub_rn(-lq1, __dmul_rn(0.5, a1)), lw1);
        const bool mem0 = ok0 && (lnp0 - M) > threshold;  // NaN compares false (fit.py:670)
        const bool mem1 = ok1 && (lnp1 - M) > threshold;
        const uint32_t mw0 = __ballot_sync(0xffffffffu, mem0);
        const uint32_t mw1 = __ballot_sync(0xffffffffu, mem1);
        // one exp per lane that has a member; a second one only where a lane has two
        double w0 = 0.0, w1 = 0.0;
        if (mw0 | mw1) {
          if (mem0 || mem1) {
            const double r = exp((mem0 ? lnp0 : lnp1) - M);
            if (mem0) w0 = r;
            else w1 = r;
          }
          if (mw0 & mw1) {
            if (mem0 && mem1) w1 = exp(lnp1 - M);
          }
        }
        const int64_t slot = static_cast<int64_t>(e) - entry_base;
        sc.weights[slot * BF_GROUP + lane] = w0;
        sc.weights[slot * BF_GROUP + 32 + lane] = w1;
        if constexpr (NT > 0) {
          wt[pending * kWtLd + lane] = w0;
          wt[pending * kWtLd + 32 + lane] = w1;
          if (lane == 0) slots[pending] = static_cast<int32_t>(slot);
        }
        double wsum;
        if (run_bits == 1u || !sc.chunk_wsum) {
          // one run of equal slow values (or no run sums wanted): the group's weight sum only
          wsum = w0 + w1;
#pragma unroll
          for (int d = 16; d > 0; d >>= 1) wsum += __shfl_xor_sync(0xffffffffu, wsum, d);
          if (sc.chunk_wsum && lane < 8) sc.chunk_wsum[slot * 8 + lane] = lane == 0 ? wsum : 0.0;
        } else {
          // weight sums of the 8 chunks (8 neighbouring lanes each), then of the group
          double s0 = w0, s1 = w1;
#pragma unroll
          for (int d = 1; d < 8; d <<= 1) {
            s0 += __shfl_xor_sync(0xffffffffu, s0, d);
            s1 += __shfl_xor_sync(0xffffffffu, s1, d);
          }
          // lane c (< 8) takes chunk c's sum, then adds the later chunks of its run: the sum of
          // a run of equal slow values sits at the run's first chunk, 0 at the others
          const double t0 = __shfl_sync(0xffffffffu, s0, (lane & 3) * 8);
          const double t1 = __shfl_sync(0xffffffffu, s1, (lane & 3) * 8);
          const double cs = (lane & 4) ? t1 : t0;
          const int c = lane & 7;
          const uint32_t later = run_bits >> (c + 1);
          const int end = later ? c + __ffs(later) : 8;
          double run = cs;
#pragma unroll
          for (int d = 1; d < 8; ++d) {
            const double v = __shfl_down_sync(0xffffffffu, cs, d, 8);
            if (c + d < end) run += v;
          }
          if (lane < 8) sc.chunk_wsum[slot * 8 + c] = ((run_bits >> c) & 1u) ? run : 0.0;
          wsum = s0 + s1;
          wsum += __shfl_xor_sync(0xffffffffu, wsum, 8);
          wsum += __shfl_xor_sync(0xffffffffu, wsum, 16);
        }
        // chi2 minimum of the members, lowest index on ties (fit.py:725-726): each lane's better
        // model first, then three 32-bit warp reductions (chi2 >= 0 orders like its bit pattern)
        const bool has = mem0 || mem1;
        const bool take1 = mem1 && (!mem0 || a1 < a0);
        const unsigned long long cu =
            static_cast<unsigned long long>(__double_as_longlong(take1 ? a1 : a0));
        const uint32_t chi_hi = has ? static_cast<uint32_t>(cu >> 32) : 0xffffffffu;
        const uint32_t min_hi = __reduce_min_sync(0xffffffffu, chi_hi);
        const bool tie_hi = has && chi_hi == min_hi;
        const uint32_t chi_lo = tie_hi ? static_cast<uint32_t>(cu) : 0xffffffffu;
        const uint32_t min_lo = __reduce_min_sync(0xffffffffu, chi_lo);
        const uint32_t cand = (tie_hi && chi_lo == min_lo) ? static_cast<uint32_t>(take1 ? lane + 32 : lane) : 64u;
        const uint32_t min_idx = __reduce_min_sync(0xffffffffu, cand);
        if (lane == 0) {
          double cm = 1.7976931348623157e308;
          long long cj = -1;
          if (min_idx < 64u) {
            cm = __longlong_as_double(static_cast<long long>((static_cast<unsigned long long>(min_hi) << 32) | min_lo));
            cj = static_cast<long long>(g) * BF_GROUP + min_idx;
          }
          *reinterpret_cast<uint2*>(member_mask + static_cast<size_t>(e) * kWordsPerGroup) =
              make_uint2(mw0, mw1);
          sc.part_wsum[slot] = wsum;
          sc.part_chi2[slot] = cm;
          sc.part_chi2_j[slot] = cj;
          member_prefix[e] = __popc(mw0) + __popc(mw1);
        }
        if constexpr (NT > 0) {
          if (++pending == 8) {
            flush_expectations<NT>(wt, qg, slots, pending, sc.exp_part, lane);
            pending = 0;
          }
        }
      }
    }
    if constexpr (NT > 0) {
      if (pending) flush_expectations<NT>(wt, qg, slots, pending, sc.exp_part, lane);
    }
  }
}

template <int NB, int NT>
static int launch_grouped_inst(const bf_exact_grid& grid, const double* flux64, const double* pmax,
                               double threshold, const uint32_t* list, const uint32_t* entry_star,
                               int64_t entry_base, int64_t n_entries, const bf_accum_scratch& sc,
                               const int32_t* fast_q, const uint8_t* slow_run,
                               const bf_star_outputs& out, unsigned gridx, cudaStream_t st) {
  const size_t smem = kGroupedWarps * GroupedSmem<NT>::kWarpBytes;
  static bool configured[64] = {};
  int dev = 0;
  BF_CUDA_TRY(cudaGetDevice(&dev));
  if (smem > 48 * 1024 && dev >= 0 && dev < 64 && !configured[dev]) {
    BF_CUDA_TRY(cudaFuncSetAttribute(exact_weights_grouped_kernel<NB, NT>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     static_cast<int>(smem)));
    configured[dev] = true;
  }
  exact_weights_grouped_kernel<NB, NT><<<gridx, kGroupedWarps * 32, smem, st>>>(
      grid, flux64, pmax, threshold, list, entry_star, sc.group_order, entry_base, n_entries, sc,
      fast_q, slow_run, out.member_mask, out.member_prefix);
  BF_CHECK_LAUNCH("exact_weights_grouped_kernel");
  return BF_OK;
}

// fast_pad: 0 = weights only; 16 / 32 = also the expectation partials of the fast class
// (needs sc.exp_part and prm_fast_q)
int launch_weights_grouped(const bf_exact_grid& grid, const double* flux64, const double* pmax,
                           double threshold, const uint32_t* list, const uint32_t* entry_star,
                           int64_t entry_base, int64_t n_entries, const bf_accum_scratch& sc,
                           const int32_t* fast_q, int fast_pad, const uint8_t* slow_run,
                           const bf_star_outputs& out, int sm_count, cudaStream_t st) {
  const int64_t n_stretch = (n_entries + kOrderStretch - 1) / kOrderStretch;
  const int64_t blocks = (n_stretch + kGroupedWarps - 1) / kGroupedWarps;
  const int64_t cap = static_cast<int64_t>(sm_count) * 64;
  const unsigned gridx = static_cast<unsigned>(blocks < cap ? blocks : cap);
#define BF_GROUPED_CASE(N)                                                                        \
  case N:                                                                                         \
    if (fast_pad == 16)                                                                           \
      return launch_grouped_inst<N, 2>(grid, flux64, pmax, threshold, list, entry_star,           \
                                       entry_base, n_entries, sc, fast_q, slow_run, out, gridx, st);        \
    if (fast_pad == 32)                                                                           \
      return launch_grouped_inst<N, 4>(grid, flux64, pmax, threshold, list, entry_star,           \
                                       entry_base, n_entries, sc, fast_q, slow_run, out, gridx, st);        \
    return launch_grouped_inst<N, 0>(grid, flux64, pmax, threshold, list, entry_star, entry_base, \
                                     n_entries, sc, fast_q, slow_run, out, gridx, st);
  switch (grid.nb) {
    BF_GROUPED_CASE(1) BF_GROUPED_CASE(2) BF_GROUPED_CASE(3) BF_GROUPED_CASE(4)
    BF_GROUPED_CASE(5) BF_GROUPED_CASE(6) BF_GROUPED_CASE(7) BF_GROUPED_CASE(8)
    default:
      break;
  }
#undef BF_GROUPED_CASE
  set_error("exact_weights_grouped_kernel: %d bands not instantiated", grid.nb);
  return BF_ERR_UNSUPPORTED;
}

// =========================================================================================
// (B) split histogram kernel
// =========================================================================================
constexpr int kSplitWarpsMax = 20;
// chunks of a star's list worked through at a time (C2: 54 ms per 33,792 stars with or without
// segments of this size; C5, whose subgrid posteriors are broad: 2.0 s -> 1.6 s per 16,384 stars)
constexpr uint32_t kSegmentChunks = 16384;
constexpr int64_t kNarrowListedGroups = 2000;  // mean listed groups per star below which two blocks share an SM

// shared memory: hist [hist_doubles] | red [2 NW] | redj [NW] | sscan (NW u32, as NW/2 doubles)
//   | epart [NW][32] | int tables (see kernel); sized for the largest block
constexpr int kSplitInts = 3 * 64 + 2 * 64 + 128 + 64 + 16;
struct SplitSmem {
  static __host__ __device__ size_t bytes(int hist_doubles) {
    return (static_cast<size_t>(hist_doubles) + 4 * kSplitWarpsMax + kSplitWarpsMax * 32 + 64) * sizeof(double) +
           kSplitInts * sizeof(int);
  }
};

// lay out one private copy of a class: lane roles, columns and the copy size (in doubles).
// SLOTS == 1: up to 32 parameters, one chunk per instruction (the hist_window_kernel layout);
// SLOTS == 2: up to 16 parameters, lanes 0..15 <-> chunk slot 0 with the parameters in order of
// decreasing range width, lanes 16..31 <-> chunk slot 1 in increasing order; lane 16+j stacks
// its range on top of lane j's.  A half-warp always touches 16 different stacks.
// wmin/wmax: ranges by class-local parameter; lparam/lcol by lane; plane[slot*32 + p] = lane.
__device__ __forceinline__ int layout_class(int slots, int n, const int* wmin, const int* wmax,
                                            int* lparam, int* lcol, int* plane, int* tmpw,
                                            int lane) {
  const int p = slots == 2 ? (lane & 15) : lane;
  const int sl = slots == 2 ? (lane >> 4) : 0;
  const bool valid = p < n;
  const bool occupied = valid && wmin[p] <= wmax[p];
  const int mn = occupied ? wmin[p] : 0;
  const int w = occupied ? wmax[p] - wmin[p] + 1 : 1;
  int rank = 0;
  const int np = slots == 2 ? 16 : 32;
  for (int j = 0; j < np; ++j) {
    const int wj = __shfl_sync(0xffffffffu, w, j);
    rank += (wj > w || (wj == w && j < p)) ? 1 : 0;
  }
  int target;
  if (slots == 2) target = sl == 0 ? rank : 16 + (15 - rank);
  else target = rank < 16 ? rank : 47 - rank;
  tmpw[target] = w;
  __syncwarp();
  const int below = target >= 16 ? tmpw[target - 16] : 0;
  const int depth = __reduce_max_sync(0xffffffffu, lane < 16 ? tmpw[lane] + tmpw[lane + 16] : 0);
  __syncwarp();
  lparam[target] = valid ? p : -1;
  lcol[target] = (below - mn) * 16 + (target & 15);
  plane[sl * 32 + p] = target;
  return depth * 16;
}

struct FastRows {
  uint2 bins;
  double w[8];
};

// eight zero weights: what a lane without a chunk of its own reads (tail of a star's chunk list)
__device__ __align__(16) const double g_zero_weights[8] = {0, 0, 0, 0, 0, 0, 0, 0};

// two updates of one lane's column: both loads in flight; when the two models share a bin the
// second takes the first one's weight along and its store, issued second, leaves the full sum.
// (Four at a time, as in hist_window_kernel, costs 44 instructions per four models -- most of
// them for the six pairwise bin comparisons -- against 24 here; this kernel has few warps per
// SM when the private copies are large and is bound by instruction issue, not by the chain.)
__device__ __forceinline__ void add_two(uint32_t col_s, uint32_t b0, uint32_t b1, double w0, double w1,
                                        uint32_t bin_stride = 128u) {
  const uint32_t a0 = col_s + b0 * bin_stride, a1 = col_s + b1 * bin_stride;
  double h0, h1;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(h0) : "r"(a0));
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(h1) : "r"(a1));
  asm("{\n .reg .pred p;\n setp.eq.u32 p, %1, %2;\n @p add.f64 %0, %0, %3;\n}\n"
      : "+d"(w1)
      : "r"(a0), "r"(a1), "d"(w0));
  h0 += w0;
  h1 += w1;
  asm volatile("st.shared.f64 [%0], %1;" ::"r"(a0), "d"(h0));
  asm volatile("st.shared.f64 [%0], %1;" ::"r"(a1), "d"(h1));
}

// (a lane without a parameter has bin_mask = 0 and adds into the one bin of its column: applied
// here as a bin stride of 0, not where the bins are loaded -- an operation on a freshly loaded
// register makes the warp wait for the load on the spot, which is what the pipeline must avoid)
__device__ __forceinline__ void add_chunk(uint32_t col_s, const FastRows& r, uint32_t bin_mask) {
  const uint32_t bx = r.bins.x, by = r.bins.y, st = bin_mask & 128u;
  add_two(col_s, bx & 0xffu, (bx >> 8) & 0xffu, r.w[0], r.w[1], st);
  add_two(col_s, (bx >> 16) & 0xffu, bx >> 24, r.w[2], r.w[3], st);
  add_two(col_s, by & 0xffu, (by >> 8) & 0xffu, r.w[4], r.w[5], st);
  add_two(col_s, (by >> 16) & 0xffu, by >> 24, r.w[6], r.w[7], st);
}

// A fast warp: chunks first*SLOTS + h, then every nfw*SLOTS-th, of the star's member chunks.
// (Pulling the rows into L2 twelve steps ahead from the fast warps themselves, with
// prefetch.global.L2, was measured: no faster for narrow stars, slower for broad ones.)
// Software pipeline in registers: the chunk reference is read four steps ahead, the rows (8 bin
// bytes + 8 weights) three steps ahead of their use, so that three steps' rows are in flight
// per warp.  bins_lane = fast_bin + this lane's parameter, bin_mask = 0 for a lane without a
// parameter (it then adds into the one-bin column it owns); a lane whose slot has run past the
// end of the list re-reads the last chunk's bins with zero weights.
template <int SLOTS>
__device__ __forceinline__ void fast_accumulate(const uint8_t* __restrict__ bins_lane, uint32_t row_bytes,
                                                uint32_t bin_mask, const double* __restrict__ weights,
                                                const uint2* refs, uint32_t n_mc, uint32_t first,
                                                uint32_t nfw, int lane, double* col,
                                                volatile uint32_t* progress) {
  const uint32_t h = SLOTS == 2 ? static_cast<uint32_t>(lane >> 4) : 0u;
  const uint32_t step = nfw * SLOTS;
  const uint32_t col_s = static_cast<uint32_t>(__cvta_generic_to_shared(col));
  const uint32_t start = first * SLOTS;  // warp-uniform position of slot 0
  if (start >= n_mc) return;
  const uint32_t n_steps = (n_mc - start + step - 1) / step;
  const uint32_t last = n_mc - 1;
  uint32_t i = start + h;  // this lane's position of the step whose reference is read next
  // (a position past the end of the list reads the last chunk's reference, and the row load,
  //  which keeps its own running position, turns it into zero weights: no flag travels with the
  //  reference, and nothing operates on a reference before it is used as an address)
  auto next_ref = [&]() {
    const uint2 r = refs[i <= last ? i : last];
    i += step;
    return r;
  };
  uint32_t il = start + h;  // position of the step whose rows are loaded next
  auto load = [&](uint2 ref, FastRows& r) {
    const double2* wp = il <= last
                            ? reinterpret_cast<const double2*>(weights + static_cast<size_t>(ref.y) * BF_CHUNK)
                            : reinterpret_cast<const double2*>(g_zero_weights);
    il += step;
    const double2 a = __ldg(wp), b = __ldg(wp + 1), c = __ldg(wp + 2), d = __ldg(wp + 3);
    const uint2 bb = __ldg(reinterpret_cast<const uint2*>(bins_lane + static_cast<size_t>(ref.x) * row_bytes));
    r.bins = bb;
    r.w[0] = a.x; r.w[1] = a.y; r.w[2] = b.x; r.w[3] = b.y;
    r.w[4] = c.x; r.w[5] = c.y; r.w[6] = d.x; r.w[7] = d.y;
  };
  // Three row sets in rotation, each refilled (for three steps later) right after it has been
  // consumed, and two sets of three chunk references in turn, requested a whole round before
  // they are needed as addresses: no register copy and no address ever waits for a load that
  // was issued less than three steps earlier.
  FastRows r0, r1, r2;
  uint2 qa0, qa1, qa2, qb0, qb1, qb2;
  load(next_ref(), r0);
  load(next_ref(), r1);
  load(next_ref(), r2);
  qa0 = next_ref();
  qa1 = next_ref();
  qa2 = next_ref();
  for (uint32_t t = 0; t < n_steps; t += 6) {
    if (progress && lane == 0) *progress = start + t * step;  // (the prefetching warps follow it)
    qb0 = next_ref();
    qb1 = next_ref();
    qb2 = next_ref();
    add_chunk(col_s, r0, bin_mask);
    load(qa0, r0);
    if (t + 1 < n_steps) {
      add_chunk(col_s, r1, bin_mask);
      load(qa1, r1);
    }
    if (t + 2 < n_steps) {
      add_chunk(col_s, r2, bin_mask);
      load(qa2, r2);
    }
    if (t + 3 >= n_steps) break;
    qa0 = next_ref();
    qa1 = next_ref();
    qa2 = next_ref();
    add_chunk(col_s, r0, bin_mask);
    load(qb0, r0);
    if (t + 4 < n_steps) {
      add_chunk(col_s, r1, bin_mask);
      load(qb1, r1);
    }
    if (t + 5 < n_steps) {
      add_chunk(col_s, r2, bin_mask);
      load(qb2, r2);
    }
  }
}


constexpr int kSlowBatch = 4;  // runs per lane and step of a slow warp
struct SlowRefs {
  uint2 ref[kSlowBatch];
  uint32_t valid;  // bit u: reference u exists
};
struct SlowRows {
  double ws[kSlowBatch];
  uint32_t bin[kSlowBatch];
  int q[kSlowBatch];
};

__device__ __forceinline__ void load_slow_refs(const uint2* refs, uint32_t n_sc, uint32_t base,
                                               uint32_t step, uint32_t h, SlowRefs& r) {
  r.valid = 0;
#pragma unroll
  for (int u = 0; u < kSlowBatch; ++u) {
    const uint32_t i = base + u * step + h;
    const bool ok = base + u * step < n_sc && i < n_sc;
    r.valid |= ok ? 1u << u : 0u;
    r.ref[u] = refs[ok ? i : n_sc - 1];
  }
}

__device__ __forceinline__ void load_slow_rows(const bf_param_tables& prm,
                                               const double* __restrict__ chunk_wsum,
                                               const SlowRefs& f, int ks, bool live, bool live_q,
                                               SlowRows& r) {
#pragma unroll
  for (int u = 0; u < kSlowBatch; ++u) {
    r.ws[u] = ((f.valid >> u) & 1u) ? __ldg(chunk_wsum + f.ref[u].y) : 0.0;
    r.bin[u] = live ? __ldg(prm.slow_bin + static_cast<size_t>(f.ref[u].x) * prm.slow_pad + ks) : 0u;
    r.q[u] = live_q ? __ldg(prm.slow_q + static_cast<size_t>(f.ref[u].x) * prm.slow_pad + ks) : 0;
  }
}

// A slow warp: one update per run of equal slow values with the run's weight sum (written by the
// grouped weights kernel), four runs per lane and step.  The references of a step are
// requested two steps, its values one step before the shared-memory work that uses them.
// Returns this lane's sum of wsum * fix (fit.py:743).
template <int SLOTS>
__device__ __forceinline__ double slow_accumulate(const bf_param_tables& prm,
                                                  const double* __restrict__ chunk_wsum,
                                                  const uint2* refs, uint32_t n_sc, uint32_t first,
                                                  uint32_t nsw, int ks, bool live, bool live_q,
                                                  int lane, double* col) {
  const uint32_t h = SLOTS == 2 ? static_cast<uint32_t>(lane >> 4) : 0u;
  const uint32_t step = nsw * SLOTS;
  const uint32_t col_s = static_cast<uint32_t>(__cvta_generic_to_shared(col));
  double eacc = 0.0;
  uint32_t base = first * SLOTS;
  if (base >= n_sc) return 0.0;
  const uint32_t stride = kSlowBatch * step;
  auto consume = [&](const SlowRows& r) {
#pragma unroll
    for (int u = 0; u < kSlowBatch; u += 2) add_two(col_s, r.bin[u], r.bin[u + 1], r.ws[u], r.ws[u + 1]);
    if (live_q) {
#pragma unroll
      for (int u = 0; u < kSlowBatch; ++u) eacc += r.ws[u] * static_cast<double>(r.q[u]);
    }
  };
  // steps k (rows in ra), k+1 (rows in rb, requested), k+2 (references in fa, requested)
  SlowRefs fa, fb;
  SlowRows ra, rb;
  load_slow_refs(refs, n_sc, base, step, h, fa);
  load_slow_rows(prm, chunk_wsum, fa, ks, live, live_q, ra);
  load_slow_refs(refs, n_sc, base + stride, step, h, fb);
  load_slow_rows(prm, chunk_wsum, fb, ks, live, live_q, rb);
  load_slow_refs(refs, n_sc, base + 2 * stride, step, h, fa);
  while (base < n_sc) {
    load_slow_refs(refs, n_sc, base + 3 * stride, step, h, fb);
    consume(ra);
    load_slow_rows(prm, chunk_wsum, fa, ks, live, live_q, ra);
    if (base + stride >= n_sc) break;
    load_slow_refs(refs, n_sc, base + 4 * stride, step, h, fa);
    consume(rb);
    load_slow_rows(prm, chunk_wsum, fb, ks, live, live_q, rb);
    base += 2 * stride;
  }
  return eacc;
}

// Warps without a private copy run ahead of the fast warps and pull the rows those will need
// (a chunk's 8 weights and its bin bytes) into L2: the weights are read once per star, so
// without this every row request of a fast warp goes to HBM.  kPrefetchAhead bounds the distance
// in chunks (the chunks in flight over all SMs must stay well inside L2).
constexpr uint32_t kPrefetchAhead = 1024;

__device__ __forceinline__ void prefetch_l2(const void* p) {
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}

__device__ __forceinline__ void prefetch_rows(const bf_param_tables& prm, const double* weights,
                                              const double* chunk_wsum, const uint2* refs,
                                              uint32_t n_mc, const uint2* srefs, uint32_t n_sc,
                                              uint32_t first, uint32_t n_warps, int lane,
                                              const volatile uint32_t* progress) {
  const uint32_t row_bytes = static_cast<uint32_t>(prm.fast_pad) * BF_CHUNK;
  for (uint32_t base = first * 32; base < n_mc; base += n_warps * 32) {
    while (base >= *progress + kPrefetchAhead) __nanosleep(200);  // warp-uniform
    const uint32_t i = base + lane;
    if (i < n_mc) {
      const uint2 ref = refs[i];
      prefetch_l2(weights + static_cast<size_t>(ref.y) * BF_CHUNK);
      const uint8_t* row = prm.fast_bin + static_cast<size_t>(ref.x) * row_bytes;
      prefetch_l2(row);
      if (row_bytes > 128) prefetch_l2(row + 128);
    }
  }
  // the slow warp's rows (run sums, bin bytes, values): a fraction of the above, no throttle
  for (uint32_t i = first * 32 + lane; i < n_sc; i += n_warps * 32) {
    const uint2 ref = srefs[i];
    prefetch_l2(chunk_wsum + ref.y);
    prefetch_l2(prm.slow_bin + static_cast<size_t>(ref.x) * prm.slow_pad);
    prefetch_l2(prm.slow_q + static_cast<size_t>(ref.x) * prm.slow_pad);
  }
}

// occupied bin range of the class parameters over the star's member chunks (four chunks per
// lane in flight: this walk waits on L2 only)
template <int SLOTS, bool FAST>
__device__ __forceinline__ void class_ranges(const bf_param_tables& prm, const uint2* refs,
                                             uint32_t n_mc, int warp, int nw, int lane, int n,
                                             int* wmin, int* wmax) {
  const uint32_t h = SLOTS == 2 ? static_cast<uint32_t>(lane >> 4) : 0u;
  const int p = SLOTS == 2 ? (lane & 15) : lane;
  const bool live = p < n;
  if (!live) return;
  uint32_t mn2 = 0x00ff00ffu, mx2 = 0;
  const uint32_t step = static_cast<uint32_t>(nw) * SLOTS;
  for (uint32_t base = static_cast<uint32_t>(warp) * SLOTS + h; base < n_mc; base += 4 * step) {
    uint32_t chunk[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const uint32_t i = base + u * step;
      chunk[u] = refs[i < n_mc ? i : base].x;  // (a repeated chunk changes nothing)
    }
    if (FAST && prm.fast_rng != nullptr) {  // the chunks' own bin ranges: 2 bytes instead of 8
      uint32_t v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = __ldg(prm.fast_rng + static_cast<size_t>(chunk[u]) * prm.fast_pad + p);
      const uint32_t lo01 = (v[0] & 0xffu) | ((v[1] & 0xffu) << 16), lo23 = (v[2] & 0xffu) | ((v[3] & 0xffu) << 16);
      const uint32_t hi01 = (v[0] >> 8) | ((v[1] >> 8) << 16), hi23 = (v[2] >> 8) | ((v[3] >> 8) << 16);
      mn2 = __vimin3_u16x2(mn2, lo01, lo23);
      mx2 = __vimax3_u16x2(mx2, hi01, hi23);
    } else if (FAST) {
      uint2 bb[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        bb[u] = __ldg(reinterpret_cast<const uint2*>(
            prm.fast_bin + (static_cast<size_t>(chunk[u]) * prm.fast_pad + p) * BF_CHUNK));
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t xe = __byte_perm(bb[u].x, 0, 0x4240), xo = __byte_perm(bb[u].x, 0, 0x4341);
        const uint32_t ye = __byte_perm(bb[u].y, 0, 0x4240), yo = __byte_perm(bb[u].y, 0, 0x4341);
        mn2 = __vimin3_u16x2(__vimin3_u16x2(mn2, xe, xo), ye, yo);
        mx2 = __vimax3_u16x2(__vimax3_u16x2(mx2, xe, xo), ye, yo);
      }
    } else {
      uint32_t b[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        b[u] = __ldg(prm.slow_bin + static_cast<size_t>(chunk[u]) * prm.slow_pad + p);
      const uint32_t p01 = b[0] | (b[1] << 16), p23 = b[2] | (b[3] << 16);
      mn2 = __vimin3_u16x2(mn2, p01, p23);
      mx2 = __vimax3_u16x2(mx2, p01, p23);
    }
  }
  atomicMin(wmin + p, static_cast<int>(min(mn2 & 0xffffu, mn2 >> 16)));
  atomicMax(wmax + p, static_cast<int>(max(mx2 & 0xffffu, mx2 >> 16)));
}

template <int FS, int SS>  // chunk slots per instruction of the fast / slow class (1 or 2)
__global__ void __launch_bounds__(kSplitWarpsMax * 32, 1)
split_hist_kernel(bf_param_tables prm, int hist_doubles, uint32_t seg_chunks, const double* __restrict__ pmax,
                  const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ list,
                  const int32_t* __restrict__ star_order, int64_t star_begin, int64_t entry_base,
                  bf_accum_scratch sc, double p0, double p1, double p2, bf_star_outputs out) {
  extern __shared__ __align__(16) double smem_d[];
  const int NW = blockDim.x >> 5;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = prm.n_q, tb = prm.total_bins, nf = prm.n_fast, nsl = prm.n_slow;

  double* hist = smem_d;
  double* red = hist + hist_doubles;
  long long* redj = reinterpret_cast<long long*>(red + 2 * kSplitWarpsMax);
  uint32_t* sscan = reinterpret_cast<uint32_t*>(redj + kSplitWarpsMax);
  double* epart = red + 4 * kSplitWarpsMax;  // [NW][32]
  double* esum = epart + kSplitWarpsMax * 32;  // [64] expectation sums by parameter
  int* wmin = reinterpret_cast<int*>(esum + 64);  // [64]: fast 0..31, slow 32..63
  int* wmax = wmin + 64;
  int* kcls = wmax + 64;       // [64] by parameter: class-local index, +32 for slow (n_q <= 64)
  int* lparam = kcls + 64;     // [2][32] by class and lane: class-local parameter or -1
  int* lcol = lparam + 64;     // [2][32]
  int* plane = lcol + 64;      // [2][64] by class, slot*32 + parameter: lane
  int* tmpw = plane + 128;     // [2][32]
  int* misc = tmpw + 64;       // [1] fast copy, [2] slow copy, [3] fast warps, [4] slow warps

  const int64_t s =
      star_begin + (star_order ? star_order[blockIdx.x] : static_cast<int64_t>(blockIdx.x));
  const uint32_t e0 = offsets[s * BF_SELECT_SPLIT];
  const uint32_t e1 = offsets[(s + 1) * BF_SELECT_SPLIT];
  const double M = pmax[s];
  for (int i = threadIdx.x; i < nf; i += blockDim.x) kcls[prm.fast_k[i]] = i;
  for (int i = threadIdx.x; i < nsl; i += blockDim.x) kcls[prm.slow_k[i]] = 32 + i;

  // ---- the star's chunks that hold a member (chunk id, place of its 8 weights) and its runs of
  // equal slow values that hold one (first chunk of the run, place of the run's weight sum) ----
  uint2* refs = reinterpret_cast<uint2*>(sc.chunk_refs) + (static_cast<size_t>(e0) - entry_base) * 16;
  uint2* srefs = refs + static_cast<size_t>(e1 - e0) * 8;
  uint32_t carry_f = 0, carry_s = 0;
  for (uint32_t eb = e0; eb < e1; eb += blockDim.x) {
    const uint32_t e = eb + threadIdx.x;
    uint32_t occ = 0, socc = 0, g = 0;
    if (e < e1) {
      const uint2 mm = *reinterpret_cast<const uint2*>(out.member_mask + static_cast<size_t>(e) * kWordsPerGroup);
      g = list[e];
#pragma unroll
      for (int c = 0; c < 8; ++c)
        if ((((c < 4 ? mm.x : mm.y) >> (8 * (c & 3))) & 0xffu) != 0) occ |= 1u << c;
      if (occ && nsl > 0) {
        uint32_t rb = prm.slow_run ? (prm.slow_run[g] | 1u) : 0xffu;
        while (rb) {
          const int c = __ffs(rb) - 1;
          rb &= rb - 1;
          const uint32_t end = rb ? __ffs(rb) - 1 : 8;
          if (occ & ((1u << end) - (1u << c))) socc |= 1u << c;
        }
      }
    }
    const uint32_t v = __popc(occ) | (__popc(socc) << 16);
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
    for (int w = 0; w < NW; ++w) {
      if (w < warp) woff += sscan[w];
      tot += sscan[w];
    }
    const uint32_t ex = woff + inc - v;
    uint32_t pos = carry_f + (ex & 0xffffu), spos = carry_s + (ex >> 16);
    const uint32_t wslot = static_cast<uint32_t>((static_cast<int64_t>(e) - entry_base) * 8);
    while (occ) {
      const int c = __ffs(occ) - 1;
      occ &= occ - 1;
      refs[pos++] = make_uint2(g * 8u + c, wslot + c);
    }
    while (socc) {
      const int c = __ffs(socc) - 1;
      socc &= socc - 1;
      srefs[spos++] = make_uint2(g * 8u + c, wslot + c);
    }
    carry_f += tot & 0xffffu;
    carry_s += tot >> 16;
  }
  const uint32_t n_mc = carry_f, n_sc = carry_s;
  for (int i = threadIdx.x; i < 64; i += blockDim.x) esum[i] = 0.0;
  if (threadIdx.x == 0) {
    misc[10] = 0;
    if (sc.star_flags) sc.star_flags[s - star_begin] = 1;
  }

  // A star with a long chunk list is worked through in segments of the list.  The list is in
  // model order, so a segment covers a narrow stretch of the stellar axes: its bin ranges are
  // narrower than the whole star's, the private copies smaller, and more warps get one -- a
  // broad star as a whole leaves room for 8 fast warps of 20.  Every segment adds its merged
  // copies to the star's pdf rows (in global memory, not normalised yet).
  const uint32_t n_seg = n_mc > seg_chunks + seg_chunks / 2 ? (n_mc + seg_chunks / 2) / seg_chunks : 1u;
  const uint32_t seg_len = (n_mc + n_seg - 1) / n_seg;
  for (uint32_t seg = 0; seg < n_seg; ++seg) {
  const uint32_t f0 = seg * seg_len;
  const uint32_t nf_seg = n_mc - f0 < seg_len ? n_mc - f0 : seg_len;
  for (int i = threadIdx.x; i < 64; i += blockDim.x) {
    wmin[i] = 255;
    wmax[i] = 0;
  }
  if (threadIdx.x == 0) {
    // the slow runs of this segment: those that start before the next segment's first chunk
    uint32_t hi = n_sc;
    if (seg + 1 < n_seg) {
      const uint32_t next_chunk = refs[f0 + nf_seg].x;
      uint32_t lo = misc[10];
      while (lo < hi) {
        const uint32_t mid = lo + (hi - lo) / 2;
        if (srefs[mid].x < next_chunk) lo = mid + 1; else hi = mid;
      }
      hi = lo;
    }
    misc[9] = misc[10];
    misc[10] = hi;
  }
  __syncthreads();  // refs (global) and the tables above are visible to the whole block
  const uint32_t s0 = misc[9], ns_seg = misc[10] - misc[9];
  const uint2* frefs = refs + f0;
  const uint2* sseg = srefs + s0;

  // ---- occupied bin ranges -----------------------------------------------------------------
  if (nf > 0) class_ranges<FS, true>(prm, frefs, nf_seg, warp, NW, lane, nf, wmin, wmax);
  if (nsl > 0) class_ranges<SS, false>(prm, sseg, ns_seg, warp, NW, lane, nsl, wmin + 32, wmax + 32);
  __syncthreads();

  // ---- layout: warp 0 the fast class, warp 1 the slow class ----------------------------------
  if (warp == 0) {
    const int c = layout_class(FS, nf, wmin, wmax, lparam, lcol, plane, tmpw, lane);
    if (lane == 0) misc[1] = c;
  } else if (warp == 1) {
    const int c = layout_class(SS, nsl, wmin + 32, wmax + 32, lparam + 32, lcol + 32, plane + 64,
                               tmpw + 32, lane);
    if (lane == 0) misc[2] = c;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const int cf = misc[1], cs = misc[2];
    int F = 0, S = 0;
    if (nf == 0) {
      S = hist_doubles / cs;
      if (S > NW) S = NW;
    } else if (nsl == 0) {
      F = hist_doubles / cf;
    } else {
      // a slow warp moves a sixteenth of what a fast warp moves per chunk: one is enough up to
      // a dozen fast warps
      S = 1;
      F = (hist_doubles - cs) / cf;
      if (F > 12 && (hist_doubles - 2 * cs) / cf >= 12 && NW > 15) {
        S = 2;
        F = (hist_doubles - 2 * cs) / cf;
      }
      if (F < 1) F = 1;  // cannot happen: the host checked the worst case
    }
    // one warp of ten always stays free for the expectation partials of the fast class
    const int X = NW >= 16 ? 2 : 1;
    if (nf > 0 && F > NW - S - X) F = NW - S - X;
    misc[3] = F;
    misc[4] = S;
    misc[8] = 0;  // chunks of the segment the first fast warp has passed
  }
  __syncthreads();
  const int cf = misc[1], cs = misc[2], F = misc[3], S = misc[4];
  volatile uint32_t* progress = reinterpret_cast<volatile uint32_t*>(misc + 8);
  const int used = F * cf + S * cs;
  for (int i = threadIdx.x; i < used; i += blockDim.x) hist[i] = 0.0;
  __syncthreads();

  // ---- accumulate ---------------------------------------------------------------------------
  double eacc = 0.0;
  if (nf_seg > 0) {
    if (warp < F) {
      const int kf = lparam[lane];
      fast_accumulate<FS>(prm.fast_bin + (kf >= 0 ? kf : 0) * BF_CHUNK,
                          static_cast<uint32_t>(prm.fast_pad) * BF_CHUNK, kf >= 0 ? 0xffffffffu : 0u,
                          sc.weights, frefs, nf_seg, warp, F, lane, hist + warp * cf + lcol[lane],
                          warp == 0 ? progress : nullptr);
      if (warp == 0 && lane == 0) *progress = 0xffffffffu - kPrefetchAhead;  // release the followers
    } else if (warp < F + S) {
      const int ks = lparam[32 + lane];
      const bool live = ks >= 0;
      const bool live_q = live && !(prm.q_const && prm.q_const[prm.slow_k[live ? ks : 0]]);
      eacc = slow_accumulate<SS>(prm, sc.chunk_wsum, sseg, ns_seg, warp - F, S, live ? ks : 0, live,
                                 live_q, lane, hist + F * cf + (warp - F) * cs + lcol[32 + lane]);
    } else if (nf > 0) {
      if (seg == 0) {
        // warps without a private copy add up the expectation partials of the fast class that
        // the grouped weights kernel left per list position: lane <-> (parameter, part)
        const int pad = prm.fast_pad;
        const int per = 32 / pad;  // list positions per warp instruction
        const int p = lane % pad, sub = lane / pad;
        const uint32_t stride = static_cast<uint32_t>(NW - F - S) * per;
        const double* src = sc.exp_part + p;
        double acc8[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) acc8[u] = 0.0;
        for (uint32_t e = e0 + (warp - F - S) * per + sub; e < e1; e += 8 * stride) {
#pragma unroll
          for (int u = 0; u < 8; ++u) {  // eight rows in flight per lane
            const uint32_t eu = e + u * stride;
            if (eu < e1) acc8[u] += __ldg(src + (static_cast<size_t>(eu) - entry_base) * pad);
          }
        }
        eacc = ((acc8[0] + acc8[1]) + (acc8[2] + acc8[3])) + ((acc8[4] + acc8[5]) + (acc8[6] + acc8[7]));
      }
      if (F > 0 && NW - F - S >= 4)  // (with fewer idle warps the fast warps cover the latency)
        prefetch_rows(prm, sc.weights, sc.chunk_wsum, frefs, nf_seg, sseg, ns_seg, warp - F - S,
                      NW - F - S, lane, progress);
    }
  }
  epart[warp * 32 + lane] = eacc;
  __syncthreads();

  // ---- this segment's expectation sums (fit.py:743), by parameter ------------------------------
  const int nx = NW - F - S;
  for (int k = threadIdx.x; k < nq; k += blockDim.x) {
    if (prm.q_const && prm.q_const[k]) continue;
    const int cl = kcls[k];
    double e = 0.0;
    if (cl >= 32) {
      const int p = cl & 31;
      for (int w = 0; w < S; ++w)
        for (int sl = 0; sl < SS; ++sl) e += epart[(F + w) * 32 + plane[64 + sl * 32 + p]];
    } else if (seg == 0) {
      const int pad = prm.fast_pad;
      for (int w = 0; w < nx; ++w)
        for (int sub = 0; sub < 32 / pad; ++sub) e += epart[(F + S + w) * 32 + sub * pad + cl];
    }
    esum[k] += e;
  }

  // ---- merge the private copies (fixed order) into the star's rows; the first segment writes
  // every bin (0 outside its ranges), the later ones add inside theirs ---------------------------
  for (int kq = warp; kq < nq; kq += NW) {
    const int b0 = prm.bin_off[kq];
    const int nbk = prm.bin_off[kq + 1] - b0;
    const int cl = kcls[kq];
    const bool slow = cl >= 32;
    const int p = cl & 31;
    const int mn = wmin[cl];
    const int width = wmax[cl] >= mn ? wmax[cl] - mn + 1 : 0;
    const int slots = slow ? SS : FS;
    const int ncopy = slow ? S : F, csize = slow ? cs : cf;
    const double* area = hist + (slow ? F * cf : 0);
    const int* lc = lcol + (slow ? 32 : 0);
    const int* pl = plane + (slow ? 64 : 0);
    double* row = out.pdf1d + s * tb + b0;
    for (int b = lane; b < nbk; b += 32) {
      const bool inside = b >= mn && b < mn + width;
      if (!inside && seg > 0) continue;
      double v = 0.0;
      if (inside)
        for (int w = 0; w < ncopy; ++w)
          for (int sl = 0; sl < slots; ++sl) v += area[w * csize + lc[pl[sl * 32 + p]] + b * 16];
      row[b] = seg == 0 ? v : row[b] + v;
    }
  }
  __syncthreads();  // the copies and the layout tables are free for the next segment
  }  // segments

  const double wtot = star_totals(e0, e1, entry_base, sc, M, s, red, redj, out);  // syncs
  const bool empty = !(wtot > 0.0);
  for (int k = threadIdx.x; k < nq; k += blockDim.x) {
    if (prm.q_const && prm.q_const[k]) continue;
    out.exp_vals[s * nq + k] = empty ? __longlong_as_double(0x7ff8000000000000LL)
                                     : prm.q_lin[2 * k] + prm.q_lin[2 * k + 1] * (esum[k] / wtot);
  }
  // ---- normalise (fit.py:690); expectations of the parameters whose bins hold one value ------
  const double inv = empty ? 0.0 : 1.0 / wtot;
  for (int kq = warp; kq < nq; kq += NW) {
    const int b0 = prm.bin_off[kq];
    const int nbk = prm.bin_off[kq + 1] - b0;
    double* row = out.pdf1d + s * tb + b0;
    const bool from_bins = prm.q_const && prm.q_const[kq];
    double eq = 0.0;
    for (int b = lane; b < nbk; b += 32) {
      const double v = row[b];
      row[b] = v * inv;
      if (from_bins) eq += v * prm.bin_q[b0 + b];
    }
    if (from_bins)
      store_bin_expectation(eq, wtot, empty, prm.q_lin[2 * kq], out.exp_vals + s * nq + kq, lane);
  }
  __syncthreads();  // histograms are consumed: their memory becomes the percentile scratch
  star_percentiles_and_prefix(prm, s, e0, e1, p0, p1, p2, hist, sscan, out);
}

static int split_hist_doubles(size_t smem_budget) {
  if (smem_budget < SplitSmem::bytes(0) + 2048) return 0;
  return static_cast<int>((smem_budget - 1024 - SplitSmem::bytes(0)) / sizeof(double)) & ~1;
}

// worst case of one copy of each class: 16 stacks of two full-width ranges; the percentile
// scratch of every warp must fit too
static bool split_fits(const bf_param_tables& prm, size_t smem, int warps) {
  const int hd = split_hist_doubles(smem);
  return hd >= 2 * 32 * prm.max_bins && hd >= warps * kMaxBinsPerParam;
}

bool split_hist_supported(const bf_param_tables& prm, size_t smem_budget) {
  if (prm.n_fast + prm.n_slow != prm.n_q || prm.n_q > 64) return false;
  if (prm.n_fast < 0 || prm.n_fast > 32 || prm.n_slow < 0 || prm.n_slow > 32) return false;
  if ((prm.n_fast > 0 && !(prm.fast_k && prm.fast_bin && prm.fast_q)) ||
      (prm.n_slow > 0 && !(prm.slow_k && prm.slow_bin && prm.slow_q)))
    return false;
  if ((prm.fast_pad != 16 && prm.fast_pad != 32) || (prm.slow_pad != 16 && prm.slow_pad != 32)) return false;
  if (prm.n_fast > prm.fast_pad || prm.n_slow > prm.slow_pad) return false;
  return split_fits(prm, smem_budget, kSplitWarpsMax);
}

int launch_split_hist(const bf_param_tables& prm, const double* pmax, const uint32_t* offsets,
                      const uint32_t* list, const int32_t* star_order, int64_t star_begin,
                      int64_t n_stars, int64_t entry_base, int64_t n_entries,
                      const bf_accum_scratch& sc, double p0, double p1, double p2,
                      const bf_star_outputs& out, size_t smem_budget, cudaStream_t st) {
  // One block of 20 warps per SM, or two of 10 warps (when a worst-case star fits half the shared
  // memory) for a range of stars with short lists: a narrow star cannot use 20 warps' worth of
  // private copies, and with two blocks one star's serial phases (lists, ranges, merge,
  // percentiles) overlap the other's accumulation.  Measured per range of one C2 batch (listed
  // groups per star: 12,800 / 3,000 / 610): 27.9 / 13.8 / 11.5 ms with one block, 35.3 / 13.6 / 9.5
  // with two -- a broad star then has half the copies.  BF_SPLIT_WARPS=20 | 1x forces either.
  const size_t half = (smem_budget + 1024) / 2 - 1024;
  int warps = kSplitWarpsMax;
  const bool narrow = n_stars > 0 && n_entries / n_stars < kNarrowListedGroups;
  const char* env = getenv("BF_SPLIT_WARPS");
  const bool want_half = env ? env[0] == '1' : narrow;
  if (want_half && split_fits(prm, half, kSplitWarpsMax / 2)) warps = kSplitWarpsMax / 2;
  const int hd = split_hist_doubles(warps == kSplitWarpsMax ? smem_budget : half);
  uint32_t seg_chunks = kSegmentChunks;  // BF_SEG=<chunks>: measurements
  if (const char* env = getenv("BF_SEG")) seg_chunks = static_cast<uint32_t>(strtoul(env, nullptr, 10));
  if (seg_chunks < 256) seg_chunks = 256;
  const size_t smem = SplitSmem::bytes(hd);
  const int fs = prm.fast_pad == 16 ? 2 : 1, ss = prm.slow_pad == 16 ? 2 : 1;
  static bool configured[64][4] = {};
  int dev = 0;
  BF_CUDA_TRY(cudaGetDevice(&dev));
#define BF_SPLIT_CASE(A, B)                                                                    \
  if (fs == A && ss == B) {                                                                    \
    if (dev < 64 && !configured[dev][(A - 1) * 2 + (B - 1)]) {                                 \
      BF_CUDA_TRY(cudaFuncSetAttribute(split_hist_kernel<A, B>,                                \
                                       cudaFuncAttributeMaxDynamicSharedMemorySize,            \
                                       static_cast<int>(smem_budget)));                        \
      BF_CUDA_TRY(cudaFuncSetAttribute(split_hist_kernel<A, B>,                                \
                                       cudaFuncAttributePreferredSharedMemoryCarveout,         \
                                       cudaSharedmemCarveoutMaxShared));                       \
      configured[dev][(A - 1) * 2 + (B - 1)] = true;                                           \
    }                                                                                          \
    split_hist_kernel<A, B><<<static_cast<unsigned>(n_stars), warps * 32, smem, st>>>(         \
        prm, hd, seg_chunks, pmax, offsets, list, star_order, star_begin, entry_base, sc, p0, p1, p2, out); \
  }
  BF_SPLIT_CASE(1, 1)
  BF_SPLIT_CASE(1, 2)
  BF_SPLIT_CASE(2, 1)
  BF_SPLIT_CASE(2, 2)
#undef BF_SPLIT_CASE
  BF_CHECK_LAUNCH("split_hist_kernel");
  return BF_OK;
}

}  // namespace bf

extern "C" int64_t bf_group_order_work_elems(int64_t n_groups) {
  if (n_groups <= 0) return 1;
  // counts [n_groups], offsets [n_groups + 1], scan tile sums
  return 2 * n_groups + 1 + bf_scan_scratch_elems(n_groups);
}

extern "C" int bf_group_order(int64_t n_groups, const uint32_t* list, int64_t entry_base,
                              int64_t n_entries, uint32_t* work, uint32_t* group_order,
                              void* stream) {
  using namespace bf;
  BF_REQUIRE(n_groups > 0 && entry_base >= 0 && n_entries >= 0, "bf_group_order: bad sizes");
  BF_REQUIRE(list && work && group_order, "bf_group_order: null pointer");
  BF_REQUIRE(entry_base + n_entries < (1LL << 32), "bf_group_order: list positions exceed 32 bits");
  if (n_entries == 0) return BF_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  uint32_t* counts = work;
  uint32_t* goff = work + n_groups;
  uint32_t* tiles = goff + n_groups + 1;
  BF_CUDA_TRY(cudaMemsetAsync(counts, 0, static_cast<size_t>(n_groups) * sizeof(uint32_t), st));
  const int64_t want = (n_entries + 255) / 256;
  const unsigned blocks = static_cast<unsigned>(want < 8192 ? want : 8192);  // grid-stride
  group_count_kernel<<<blocks, 256, 0, st>>>(list, entry_base, n_entries, counts);
  BF_CHECK_LAUNCH("group_count_kernel");
  const int rc = bf_exclusive_scan_u32(n_groups, counts, goff, tiles, stream);
  if (rc != BF_OK) return rc;
  BF_CUDA_TRY(cudaMemsetAsync(counts, 0, static_cast<size_t>(n_groups) * sizeof(uint32_t), st));
  group_scatter_kernel<<<blocks, 256, 0, st>>>(list, entry_base, n_entries, goff, counts, group_order);
  BF_CHECK_LAUNCH("group_scatter_kernel");
  return BF_OK;
}
