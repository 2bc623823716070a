// Dense pass: FP32 chi2 + prior constant for every (star, model), reduced to one
// minimum per (star, group of BF_GROUP models).
//
// Replaces the per-star full-grid sweep of the reference
// (beast/fitting/fit_metrics/likelihood.py:24-55, 109-167 and beast/fitting/fit.py:657-668):
//   acc(s, m) = sum_b ( f_sb * r_mb - a_mb )^2 + c2_m   =  -2 ( lnP(s, m) - c_off )
// with r = 1/sigma, a = (sed+bias)/sigma, c2 = -2 (-lnQ + ln w - c_off).
//
// Mapping: each thread keeps 2*NPAIR stars in registers as packed f32x2 pairs
// (FFMA2: one instruction = 2 star x 1 model FMAs); a thread block owns BF_STAR_BLOCK
// stars.  Grid tiles (BF_TILE_MODELS models x (2*nb+1) rows, 13 KB at 6 bands) are staged
// into shared memory by 1-D bulk async copies (TMA, UBLKCP) through a 4-stage
// mbarrier ring, and every value read from a tile is a warp-wide broadcast that is
// reused by all 1024 stars of the block.  Per (star, model): 2*nb FFMA2 halves
// (= 2*nb FP32 FMA) + 1/2 FMNMX3; no exp, no divergent code.
//
// This pass is a *filter*: the exact FP64 pass (exact.cu) re-evaluates every model of
// every selected group, so FP32 rounding here only has to stay inside the selection
// margin ("margins").
//
// Pruned sweep (PRUNE): with a per-star ceiling `cut` on the selection threshold (from the
// coarse sweep: the minimum over a subset of the models is an upper bound of the minimum) and a
// box around each group's models (per band: the range of model fluxes m = a / r and the smallest
// r; the smallest c2),
//   acc(s, m) = sum_b r_mb^2 (f_sb - m_mb)^2 + c2_m >= sum_b rlo_b^2 dist(f_sb, [mlo_b, mhi_b])^2 + c2min
// a warp skips a group that none of its 256 stars can select and reports +inf for it; within a
// group it evaluates only those of its four star pairs (64 stars each) that hold such a star, and
// for those only the halves of the group (32 models, with boxes of their own) that they meet.  The
// group lists that follow are the same ones: a skipped group's minimum is above the star's
// selection threshold, and the group of the star's minimum is never skipped.

#include "common.cuh"

namespace bf {

constexpr int kDenseThreads = 128;
constexpr int kNPair = BF_STAR_BLOCK / kDenseThreads / 2;  // star pairs per thread (4)

// COV = full-covariance likelihood (N_covar_chi2, fit_metrics/likelihood.py:58-106): the tile
// then carries, per model, -m'_b = -(sed+bias)/s_b, D_b = icov_diag_b s_b^2, the packed upper
// triangle O_kl = 2 icov_offdiag_kl s_k s_l (k<l, row-major) and c2, and
//   acc = sum_b t_b^2 D_b + sum_k t_k sum_{l>k} O_kl t_l + c2,   t_b = f_b/s_b - m'_b.
template <int NB, bool COV, bool PRUNE = false>
struct DenseSmem {
  static constexpr int kRows = COV ? 2 * NB + NB * (NB - 1) / 2 + 1 : 2 * NB + 1;
  static constexpr int kStages = COV ? 3 : 4;
  static constexpr int kPrefetch = COV ? 1 : 2;  // tiles in flight ahead of the consumer
  static constexpr uint32_t kTileBytes = kRows * BF_TILE_MODELS * sizeof(float);
  // group boxes of a tile: per group 3 x (mlo[NB], mhi[NB], rlo[NB], c2min) -- the group and its
  // two halves --, padded to 16 bytes
  static constexpr int kBoxFloats = BF_BOX_FLOATS(NB);
  static constexpr uint32_t kBoxBytes = PRUNE ? (BF_TILE_MODELS / BF_GROUP) * kBoxFloats * sizeof(float) : 0;
  static constexpr uint32_t kCutBytes = PRUNE ? BF_STAR_BLOCK * sizeof(float) : 0;
  static constexpr uint32_t kBytes =
      kStages * (kTileBytes + kBoxBytes) + kCutBytes + 2 * kStages * sizeof(uint64_t);
};

// Stars of a thread: star0 + warp * 256 + q * 32 + lane, q = 0..2*NPAIR-1 (pair p = (2p, 2p+1)):
// a warp owns 256 consecutive stars -- neighbours after the presort, which is what lets a whole
// warp skip a group.
template <int NB, bool COV, bool PRUNE>
__global__ void __launch_bounds__(kDenseThreads, COV ? 2 : 3)
dense_chi2_kernel(const float* __restrict__ tiles, int64_t n_tiles, int tiles_per_item,
                  const float* __restrict__ flux, int64_t ld_star, float* __restrict__ gmin,
                  unsigned long long* __restrict__ starmin, const float* __restrict__ boxes,
                  const float* __restrict__ cut, unsigned long long* __restrict__ evaluated) {
  constexpr int TM = BF_TILE_MODELS;
  using Smem = DenseSmem<NB, COV, PRUNE>;
  constexpr int ROWS = Smem::kRows;
  constexpr int kDenseStages = Smem::kStages;
  constexpr int kDensePrefetch = Smem::kPrefetch;
  constexpr uint32_t TILE_BYTES = Smem::kTileBytes;
  constexpr uint32_t BOX_BYTES = Smem::kBoxBytes;
  constexpr int BOXF = Smem::kBoxFloats;
  constexpr int GROUPS_PER_TILE = TM / BF_GROUP;

  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* stage_buf = reinterpret_cast<float*>(smem_raw);
  float* box_buf = reinterpret_cast<float*>(smem_raw + kDenseStages * TILE_BYTES);
  float* cut_s = reinterpret_cast<float*>(smem_raw + kDenseStages * (TILE_BYTES + BOX_BYTES));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_raw + kDenseStages * (TILE_BYTES + BOX_BYTES) +
                                                   Smem::kCutBytes);
  uint64_t* empty_bar = full_bar + kDenseStages;

  const int tid = threadIdx.x;
  const int lane = tid & 31;
  const int wbase = (tid >> 5) * (2 * kNPair * 32) + lane;  // first star of this thread in the block
  const int64_t star0 = static_cast<int64_t>(blockIdx.y) * BF_STAR_BLOCK;
  const int64_t t_begin = static_cast<int64_t>(blockIdx.x) * tiles_per_item;
  const int nt = static_cast<int>(min(static_cast<int64_t>(tiles_per_item), n_tiles - t_begin));

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < kDenseStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], kDenseThreads / 32);
    }
    mbar_fence_init();
  }
  __syncthreads();

  const float* tile_src = tiles + t_begin * static_cast<int64_t>(ROWS * TM);
  const float* box_src = PRUNE ? boxes + t_begin * static_cast<int64_t>(GROUPS_PER_TILE * BOXF) : nullptr;
  auto load_tile = [&](int j, int js) {  // tile j of this item (and its group boxes) -> stage js
    mbar_expect_tx(&full_bar[js], TILE_BYTES + BOX_BYTES);
    tma_load_1d(stage_buf + js * (ROWS * TM), tile_src + static_cast<int64_t>(j) * (ROWS * TM),
                TILE_BYTES, &full_bar[js]);
    if constexpr (PRUNE)
      tma_load_1d(box_buf + js * (GROUPS_PER_TILE * BOXF),
                  box_src + static_cast<int64_t>(j) * (GROUPS_PER_TILE * BOXF), BOX_BYTES, &full_bar[js]);
  };
  if (tid == 0) {
#pragma unroll
    for (int j = 0; j < kDensePrefetch; ++j) {
      if (j < nt) load_tile(j, j);
    }
  }

  f32x2 f[kNPair][NB];
#pragma unroll
  for (int p = 0; p < kNPair; ++p) {
#pragma unroll
    for (int b = 0; b < NB; ++b) {
      const float* fb = flux + static_cast<int64_t>(b) * ld_star + star0 + wbase;
      f[p][b] = pack2(fb[(2 * p) * 32], fb[(2 * p + 1) * 32]);
    }
  }
  const float kInf = __int_as_float(0x7f800000);
  // Pruned: a first test per (warp, group) in one band only -- the band in which this warp's 256
  // stars lie closest together (the one the caller ordered them by, found here from the fluxes):
  // the interval [wlo, whi] of their fluxes in band pb against the group's box in that band and
  // the largest of their ceilings.  Six instructions rule out most groups before any star is
  // looked at.
  float wlo = kInf, whi = -kInf, cutmax = -kInf;
  int pb = 0;
  if constexpr (PRUNE) {
    float cq[2 * kNPair];
#pragma unroll
    for (int q = 0; q < 2 * kNPair; ++q) {
      cq[q] = cut[star0 + wbase + q * 32];
      cut_s[wbase + q * 32] = cq[q];
      cutmax = fmaxf(cutmax, cq[q]);
    }
    float best = kInf;
#pragma unroll
    for (int b = 0; b < NB; ++b) {
      float lo = kInf, hi = -kInf;
#pragma unroll
      for (int p = 0; p < kNPair; ++p) {
        float fl, fh;
        unpack2(f[p][b], fl, fh);
        if (cq[2 * p] >= 0.0f) {  // (padding stars carry -1 and take no part)
          lo = fminf(lo, fl);
          hi = fmaxf(hi, fl);
        }
        if (cq[2 * p + 1] >= 0.0f) {
          lo = fminf(lo, fh);
          hi = fmaxf(hi, fh);
        }
      }
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) {
        lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, d));
        hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, d));
      }
      const float spread = (hi - lo) / fmaxf(fabsf(hi) + fabsf(lo), 1e-30f);
      if (spread < best) {  // (a NaN spread -- no star, or infinite fluxes -- never wins)
        best = spread;
        pb = b;
        wlo = lo;
        whi = hi;
      }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) cutmax = fmaxf(cutmax, __shfl_xor_sync(0xffffffffu, cutmax, d));
    __syncwarp();  // a thread reads back only what it wrote; the fence keeps the order explicit
  }

  float rmin[2 * kNPair], smin[2 * kNPair];
  unsigned int sgrp[2 * kNPair];  // group that holds the running minimum
#pragma unroll
  for (int q = 0; q < 2 * kNPair; ++q) {
    rmin[q] = kInf;
    smin[q] = kInf;
    sgrp[q] = 0xffffffffu;
  }

  unsigned n_eval = 0;
  for (int i = 0; i < nt; ++i) {
    const int stage = i % kDenseStages;
    // producer side (one thread): refill the stage that was consumed two iterations ago
    if (tid == 0) {
      const int j = i + kDensePrefetch;
      if (j < nt) {
        const int js = j % kDenseStages;
        const int use = j / kDenseStages;
        if (use > 0) mbar_wait(&empty_bar[js], (use - 1) & 1);
        load_tile(j, js);
      }
    }
    mbar_wait(&full_bar[stage], (i / kDenseStages) & 1);

    const float* sb = stage_buf + stage * (ROWS * TM);
#pragma unroll 1
    for (int g = 0; g < GROUPS_PER_TILE; ++g) {
      const int64_t gidx = (t_begin + i) * GROUPS_PER_TILE + g;
      float* gp = gmin + gidx * ld_star + star0 + wbase;
      unsigned need_mask = (1u << kNPair) - 1;
      if constexpr (PRUNE) {
        // (n_eval: star pairs x half groups this warp evaluates, for the caller's statistics)
        // can any star of this warp select a model of the group?  lower bound of acc per star
        // A group's record: the box of its 64 models, then the boxes of its two halves.
        const float* bx = box_buf + stage * (GROUPS_PER_TILE * BOXF) + g * BOXF;
        auto pair_needs = [&](const float* box, int p) {  // does a star of pair p get below its ceiling?
          float lo = box[3 * NB], hi = lo;
#pragma unroll
          for (int b = 0; b < NB; ++b) {
            float fl, fh;
            unpack2(f[p][b], fl, fh);
            const float mlo = box[b], mhi = box[NB + b], rlo = box[2 * NB + b];
            const float dl = fmaxf(fmaxf(mlo - fl, fl - mhi), 0.0f) * rlo;
            const float dh = fmaxf(fmaxf(mlo - fh, fh - mhi), 0.0f) * rlo;
            lo = fmaf(dl, dl, lo);
            hi = fmaf(dh, dh, hi);
          }
          // 1 - 2^-16 covers the FP32 rounding of both this bound and the accumulators below
          const bool need = lo * 0.99998474f < cut_s[wbase + (2 * p) * 32] ||
                            hi * 0.99998474f < cut_s[wbase + (2 * p + 1) * 32];
          return __any_sync(0xffffffffu, need) != 0;
        };
        need_mask = 0;
        const float dw = fmaxf(fmaxf(bx[pb] - whi, wlo - bx[NB + pb]), 0.0f) * bx[2 * NB + pb];
        if (fmaf(dw, dw, bx[3 * NB]) * 0.99998474f < cutmax) {  // warp-uniform
#pragma unroll
          for (int p = 0; p < kNPair; ++p) {
            if (pair_needs(bx, p)) need_mask |= 1u << p;  // the pair's 64 stars meet the group
          }
        }
        if (need_mask == 0) {
#pragma unroll
          for (int q = 0; q < 2 * kNPair; ++q) gp[q * 32] = kInf;
          continue;
        }
        // the pairs that do: which half of the group (bits p: first 32 models, kNPair + p: second)
        unsigned halves = 0;
#pragma unroll
        for (int p = 0; p < kNPair; ++p) {
          if ((need_mask >> p) & 1) {  // warp-uniform
            if (pair_needs(bx + (3 * NB + 1), p)) halves |= 1u << p;
            if (pair_needs(bx + 2 * (3 * NB + 1), p)) halves |= 1u << (kNPair + p);
          }
        }
        need_mask = halves;
        n_eval += __popc(need_mask);
      }
      if constexpr (!COV) {
#pragma unroll 1
        for (int hf = 0; hf < 2; ++hf) {
        // (pruned: the pairs that meet this half of the group; a half nobody meets is not walked)
        const unsigned live = PRUNE ? (need_mask >> (hf * kNPair)) & ((1u << kNPair) - 1) : need_mask;
        if (PRUNE && live == 0) continue;
        const int m0 = g * BF_GROUP + hf * (BF_GROUP / 2);
#pragma unroll 2
        for (int m = m0; m < m0 + BF_GROUP / 2; m += 2) {
          float2 r[NB], na[NB];
#pragma unroll
          for (int b = 0; b < NB; ++b) {
            r[b] = *reinterpret_cast<const float2*>(sb + b * TM + m);
            na[b] = *reinterpret_cast<const float2*>(sb + (NB + b) * TM + m);
          }
          const float2 c2 = *reinterpret_cast<const float2*>(sb + 2 * NB * TM + m);
#pragma unroll
          for (int p = 0; p < kNPair; ++p) {
            if (PRUNE && !((live >> p) & 1)) continue;  // warp-uniform
            f32x2 acc0 = pack2(c2.x, c2.x);
            f32x2 acc1 = pack2(c2.y, c2.y);
#pragma unroll
            for (int b = 0; b < NB; ++b) {
              const f32x2 t0 = fma2(f[p][b], pack2(r[b].x, r[b].x), pack2(na[b].x, na[b].x));
              const f32x2 t1 = fma2(f[p][b], pack2(r[b].y, r[b].y), pack2(na[b].y, na[b].y));
              acc0 = fma2(t0, t0, acc0);
              acc1 = fma2(t1, t1, acc1);
            }
            float a0l, a0h, a1l, a1h;
            unpack2(acc0, a0l, a0h);
            unpack2(acc1, a1l, a1h);
            rmin[2 * p] = min3(rmin[2 * p], a0l, a1l);
            rmin[2 * p + 1] = min3(rmin[2 * p + 1], a0h, a1h);
          }
        }
        }
      } else {
        constexpr int NP = NB * (NB - 1) / 2;
        const f32x2 one = pack2(1.0f, 1.0f), zero = pack2(0.0f, 0.0f);
#pragma unroll 1
        for (int m = g * BF_GROUP; m < (g + 1) * BF_GROUP; ++m) {
          float nm[NB], dg[NB], off[NP > 0 ? NP : 1];
#pragma unroll
          for (int b = 0; b < NB; ++b) {
            nm[b] = sb[b * TM + m];
            dg[b] = sb[(NB + b) * TM + m];
          }
#pragma unroll
          for (int q = 0; q < NP; ++q) off[q] = sb[(2 * NB + q) * TM + m];
          const float c2 = sb[(2 * NB + NP) * TM + m];
#pragma unroll
          for (int p = 0; p < kNPair; ++p) {
            f32x2 t[NB];
#pragma unroll
            for (int b = 0; b < NB; ++b) t[b] = fma2(f[p][b], one, pack2(nm[b], nm[b]));
            f32x2 acc = pack2(c2, c2);
#pragma unroll
            for (int b = 0; b < NB; ++b) {
              const f32x2 u = fma2(t[b], pack2(dg[b], dg[b]), zero);
              acc = fma2(u, t[b], acc);
            }
            int q = 0;
#pragma unroll
            for (int k = 0; k < NB - 1; ++k) {
              f32x2 v = zero;
#pragma unroll
              for (int l = k + 1; l < NB; ++l, ++q) v = fma2(t[l], pack2(off[q], off[q]), v);
              acc = fma2(v, t[k], acc);
            }
            float al, ah;
            unpack2(acc, al, ah);
            rmin[2 * p] = fminf(rmin[2 * p], al);
            rmin[2 * p + 1] = fminf(rmin[2 * p + 1], ah);
          }
        }
      }
#pragma unroll
      for (int q = 0; q < 2 * kNPair; ++q) {
        gp[q * 32] = rmin[q];
        if (rmin[q] < smin[q]) {
          smin[q] = rmin[q];
          sgrp[q] = static_cast<unsigned int>(gidx);
        }
        rmin[q] = kInf;
      }
    }
    __syncwarp();
    if ((tid & 31) == 0) mbar_arrive(&empty_bar[stage]);
  }

  if constexpr (PRUNE) {
    if (evaluated != nullptr && lane == 0 && n_eval > 0)  // (star, model) pairs: 64 stars x half a group each
      atomicAdd(evaluated, static_cast<unsigned long long>(n_eval) * (64 * (BF_GROUP / 2)));
  }
  // acc >= 0 always (chi2 >= 0, c2 >= 0), so the IEEE bit pattern orders like the value;
  // packed as (value bits << 32 | group) one 64-bit atomicMin keeps the arg-min group too
#pragma unroll
  for (int q = 0; q < 2 * kNPair; ++q) {
    if (smin[q] < kInf) {
      const unsigned long long packed =
          (static_cast<unsigned long long>(__float_as_uint(smin[q])) << 32) | sgrp[q];
      atomicMin(starmin + star0 + wbase + q * 32, packed);
    }
  }
}

template <int NB, bool COV, bool PRUNE = false>
static int launch_dense(int64_t n_tiles, const float* tiles, int64_t ld_star, const float* flux,
                        float* gmin, unsigned long long* starmin, int tiles_per_item,
                        cudaStream_t st, const float* boxes = nullptr, const float* cut = nullptr,
                        unsigned long long* evaluated = nullptr) {
  auto kern = dense_chi2_kernel<NB, COV, PRUNE>;
  static bool configured[64] = {};
  int dev = 0;
  BF_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 64 && !configured[dev]) {
    BF_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     DenseSmem<NB, COV, PRUNE>::kBytes));
    configured[dev] = true;
  }
  const int64_t n_items = (n_tiles + tiles_per_item - 1) / tiles_per_item;
  dim3 grid(static_cast<unsigned>(n_items), static_cast<unsigned>(ld_star / BF_STAR_BLOCK));
  kern<<<grid, kDenseThreads, DenseSmem<NB, COV, PRUNE>::kBytes, st>>>(
      tiles, n_tiles, tiles_per_item, flux, ld_star, gmin, starmin, boxes, cut, evaluated);
  BF_CHECK_LAUNCH("dense_chi2_kernel");
  return BF_OK;
}

}  // namespace bf

extern "C" int bf_dense_tiles_per_item(int64_t n_tiles, int64_t n_stars, int sm_count) {
  if (n_tiles <= 0 || n_stars <= 0 || sm_count <= 0) return 1;
  const int64_t star_blocks = (n_stars + BF_STAR_BLOCK - 1) / BF_STAR_BLOCK;
  // aim for ~12 waves of 3 resident blocks per SM, at least 4 tiles per item
  const int64_t want_items = 12LL * 3 * sm_count;
  int64_t slices = (want_items + star_blocks - 1) / star_blocks;
  if (slices < 1) slices = 1;
  int64_t tpi = (n_tiles + slices - 1) / slices;
  if (tpi < 4) tpi = 4;
  if (tpi > n_tiles) tpi = n_tiles;
  return static_cast<int>(tpi);
}

static int dense_entry(bool cov, int nb, int64_t n_tiles, const float* tiles, int64_t n_stars,
                       int64_t ld_star, const float* flux, float* gmin, uint64_t* starmin,
                       int tiles_per_item, void* stream) {
  using namespace bf;
  BF_REQUIRE(nb >= 1 && nb <= BF_MAX_BANDS, "bf_dense_chi2: nb=%d out of range", nb);
  BF_REQUIRE(n_tiles > 0 && n_stars > 0, "bf_dense_chi2: empty problem");
  BF_REQUIRE(ld_star >= n_stars && ld_star % BF_STAR_BLOCK == 0,
             "bf_dense_chi2: ld_star must be a multiple of %d and >= n_stars", BF_STAR_BLOCK);
  BF_REQUIRE(tiles_per_item > 0, "bf_dense_chi2: tiles_per_item must be positive");
  BF_REQUIRE(tiles && flux && gmin && starmin, "bf_dense_chi2: null pointer");
  BF_REQUIRE(ld_star / BF_STAR_BLOCK <= 65535, "bf_dense_chi2: too many stars per call");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  unsigned long long* sm = reinterpret_cast<unsigned long long*>(starmin);
  if (!cov) {
    switch (nb) {
#define BF_CASE(N) \
  case N:          \
    return launch_dense<N, false>(n_tiles, tiles, ld_star, flux, gmin, sm, tiles_per_item, st);
      BF_CASE(1) BF_CASE(2) BF_CASE(3) BF_CASE(4) BF_CASE(5) BF_CASE(6) BF_CASE(7) BF_CASE(8)
      BF_CASE(9) BF_CASE(10) BF_CASE(11) BF_CASE(12)
#undef BF_CASE
      default:
        set_error("bf_dense_chi2: nb=%d not instantiated (1..12)", nb);
        return BF_ERR_UNSUPPORTED;
    }
  }
  switch (nb) {
#define BF_CASE(N) \
  case N:          \
    return launch_dense<N, true>(n_tiles, tiles, ld_star, flux, gmin, sm, tiles_per_item, st);
    BF_CASE(2) BF_CASE(3) BF_CASE(4) BF_CASE(5) BF_CASE(6) BF_CASE(7) BF_CASE(8)
#undef BF_CASE
    default:
      set_error("bf_dense_chi2_cov: nb=%d not instantiated (2..8)", nb);
      return BF_ERR_UNSUPPORTED;
  }
}

extern "C" int bf_dense_chi2(int nb, int64_t n_tiles, const float* tiles, int64_t n_stars,
                             int64_t ld_star, const float* flux, float* gmin, uint64_t* starmin,
                             int tiles_per_item, void* stream) {
  return dense_entry(false, nb, n_tiles, tiles, n_stars, ld_star, flux, gmin, starmin,
                     tiles_per_item, stream);
}

extern "C" int bf_dense_chi2_cov(int nb, int64_t n_tiles, const float* tiles, int64_t n_stars,
                                 int64_t ld_star, const float* flux, float* gmin,
                                 uint64_t* starmin, int tiles_per_item, void* stream) {
  return dense_entry(true, nb, n_tiles, tiles, n_stars, ld_star, flux, gmin, starmin,
                     tiles_per_item, stream);
}

extern "C" int bf_dense_chi2_pruned(int nb, int64_t n_tiles, const float* tiles, const float* boxes,
                                    int64_t n_stars, int64_t ld_star, const float* flux,
                                    const float* cut, float* gmin, uint64_t* starmin,
                                    int tiles_per_item, uint64_t* evaluated, void* stream) {
  using namespace bf;
  BF_REQUIRE(nb >= 1 && nb <= 12, "bf_dense_chi2_pruned: nb=%d not instantiated (1..12)", nb);
  BF_REQUIRE(n_tiles > 0 && n_stars > 0, "bf_dense_chi2_pruned: empty problem");
  BF_REQUIRE(ld_star >= n_stars && ld_star % BF_STAR_BLOCK == 0,
             "bf_dense_chi2_pruned: ld_star must be a multiple of %d and >= n_stars", BF_STAR_BLOCK);
  BF_REQUIRE(tiles_per_item > 0, "bf_dense_chi2_pruned: tiles_per_item must be positive");
  BF_REQUIRE(tiles && boxes && flux && cut && gmin && starmin, "bf_dense_chi2_pruned: null pointer");
  BF_REQUIRE((reinterpret_cast<uintptr_t>(boxes) & 15) == 0, "bf_dense_chi2_pruned: boxes must be 16-byte aligned");
  BF_REQUIRE(ld_star / BF_STAR_BLOCK <= 65535, "bf_dense_chi2_pruned: too many stars per call");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  unsigned long long* sm = reinterpret_cast<unsigned long long*>(starmin);
  switch (nb) {
#define BF_CASE(N) \
  case N:          \
    return launch_dense<N, false, true>(n_tiles, tiles, ld_star, flux, gmin, sm, tiles_per_item, st, boxes, cut, \
                                        reinterpret_cast<unsigned long long*>(evaluated));
    BF_CASE(1) BF_CASE(2) BF_CASE(3) BF_CASE(4) BF_CASE(5) BF_CASE(6) BF_CASE(7) BF_CASE(8)
    BF_CASE(9) BF_CASE(10) BF_CASE(11) BF_CASE(12)
#undef BF_CASE
  }
  return BF_ERR_UNSUPPORTED;
}
