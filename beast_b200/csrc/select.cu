// Group selection and the prefix sum that lays out the per-star group lists.
//
// Conservative, group-granular version of the reference's sparse selection
//   indx = where(lnp - max(lnp[isfinite]) > threshold)          (beast/fitting/fit.py:670)
// A group survives for a star when its FP32 minimum is within tau of the star's
// FP32 minimum; tau carries the threshold plus the FP32 error margin, so every model
// of the exact sparse set lives in a surviving group.  The exact FP64 test happens in
// exact.cu.

#include <atomic>

#include "common.cuh"

namespace bf {

constexpr int kSelThreads = 128;

// thread <-> star (coalesced along the star axis of gmin), blockIdx.y <-> model-axis split
template <bool FILL>
__global__ void __launch_bounds__(kSelThreads)
select_kernel(int64_t n_groups, int64_t n_stars, int64_t ld_star, const float* __restrict__ gmin,
              const unsigned long long* __restrict__ starmin, float tau, float err_a,
              float err_b,
              uint32_t* __restrict__ counts,
              const uint32_t* __restrict__ offsets, uint32_t* __restrict__ list,
              uint32_t* __restrict__ entry_star) {
  const int64_t s = static_cast<int64_t>(blockIdx.x) * kSelThreads + threadIdx.x;
  if (s >= n_stars) return;
  const int c = blockIdx.y;
  const int64_t per = (n_groups + BF_SELECT_SPLIT - 1) / BF_SELECT_SPLIT;
  const int64_t g0 = c * per;
  const int64_t g1 = min(n_groups, g0 + per);
  const float thr = select_threshold(starmin_value(starmin[s]), tau, err_a, err_b);
  const float* p = gmin + s;
  uint32_t n = 0;
  uint32_t pos = FILL ? offsets[s * BF_SELECT_SPLIT + c] : 0u;
  int64_t g = g0;
  for (; g + 8 <= g1; g += 8) {
    float v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = __ldg(p + (g + u) * ld_star);
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      if (v[u] < thr) {
        if (FILL) {
          list[pos] = static_cast<uint32_t>(g + u);
          if (entry_star) entry_star[pos] = static_cast<uint32_t>(s);
          ++pos;
        }
        ++n;
      }
    }
  }
  for (; g < g1; ++g) {
    if (__ldg(p + g * ld_star) < thr) {
      if (FILL) {
        list[pos] = static_cast<uint32_t>(g);
        if (entry_star) entry_star[pos] = static_cast<uint32_t>(s);
        ++pos;
      }
      ++n;
    }
  }
  if (!FILL) counts[s * BF_SELECT_SPLIT + c] = n;
}

// ---- exclusive scan of uint32: tile sums -> scan of tile sums -> apply ----------------
constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;  // 2048

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* total) {
  __shared__ uint32_t warp_sums[kScanThreads / 32];
  __shared__ uint32_t block_total;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint32_t inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += t;
  }
  if (lane == 31) warp_sums[w] = inc;
  __syncthreads();
  if (w == 0) {
    uint32_t ws = lane < kScanThreads / 32 ? warp_sums[lane] : 0u;
    uint32_t winc = ws;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      uint32_t t = __shfl_up_sync(0xffffffffu, winc, d);
      if (lane >= d) winc += t;
    }
    if (lane < kScanThreads / 32) warp_sums[lane] = winc - ws;
    if (lane == 31) block_total = winc;
  }
  __syncthreads();
  const uint32_t r = inc - v + warp_sums[w];
  *total = block_total;
  __syncthreads();
  return r;
}

__global__ void __launch_bounds__(kScanThreads)
scan_tile_sums(int64_t n, const uint32_t* __restrict__ in, uint32_t* __restrict__ tile_sums) {
  const int64_t base = static_cast<int64_t>(blockIdx.x) * kScanTile;
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    const int64_t idx = base + i * kScanThreads + threadIdx.x;
    if (idx < n) s += in[idx];
  }
  uint32_t total;
  block_exclusive_scan(s, &total);
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

// single block: in-place exclusive scan of the tile sums, carry across chunks
__global__ void __launch_bounds__(kScanThreads)
scan_of_sums(int64_t n_tiles, uint32_t* __restrict__ tile_sums, uint32_t* __restrict__ grand_total) {
  uint32_t carry = 0;
  for (int64_t base = 0; base < n_tiles; base += kScanThreads) {
    const int64_t idx = base + threadIdx.x;
    const uint32_t v = idx < n_tiles ? tile_sums[idx] : 0u;
    uint32_t total;
    const uint32_t ex = block_exclusive_scan(v, &total);
    if (idx < n_tiles) tile_sums[idx] = ex + carry;
    carry += total;
  }
  if (threadIdx.x == 0) *grand_total = carry;
}

__global__ void __launch_bounds__(kScanThreads)
scan_apply(int64_t n, const uint32_t* __restrict__ in, const uint32_t* __restrict__ tile_sums,
           uint32_t* __restrict__ out) {
  const int64_t base = static_cast<int64_t>(blockIdx.x) * kScanTile + threadIdx.x * kScanItems;
  uint32_t v[kScanItems];
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    v[i] = (base + i < n) ? in[base + i] : 0u;
    s += v[i];
  }
  uint32_t total;
  uint32_t ex = block_exclusive_scan(s, &total) + tile_sums[blockIdx.x];
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    if (base + i < n) out[base + i] = ex;
    ex += v[i];
  }
}

// ---- scratch for bf_exclusive_scan_u32_noalloc ------------------------------------------
// One tile sum per 2048 input elements.  Lives in the module image: every GPU that loads this
// library carries these 256 KiB until the process exits.
constexpr int64_t kStaticScanTiles = 65536;
static __device__ __align__(256) uint32_t g_scan_tile_sums[kStaticScanTiles];

static int run_scan(int64_t n, const uint32_t* in, uint32_t* out, uint32_t* scratch,
                    cudaStream_t st) {
  const int64_t n_tiles = (n + kScanTile - 1) / kScanTile;
  scan_tile_sums<<<static_cast<unsigned>(n_tiles), kScanThreads, 0, st>>>(n, in, scratch);
  BF_CHECK_LAUNCH("scan_tile_sums");
  scan_of_sums<<<1, kScanThreads, 0, st>>>(n_tiles, scratch, out + n);
  BF_CHECK_LAUNCH("scan_of_sums");
  scan_apply<<<static_cast<unsigned>(n_tiles), kScanThreads, 0, st>>>(n, in, scratch, out);
  BF_CHECK_LAUNCH("scan_apply");
  return BF_OK;
}

}  // namespace bf

extern "C" int bf_select_count(int64_t n_groups, int64_t n_stars, int64_t ld_star,
                               const float* gmin, const uint64_t* starmin, float tau,
                               float err_a, float err_b, uint32_t* counts, void* stream) {
  using namespace bf;
  BF_REQUIRE(n_groups > 0 && n_stars > 0 && ld_star >= n_stars, "bf_select_count: bad sizes");
  BF_REQUIRE(gmin && starmin && counts, "bf_select_count: null pointer");
  dim3 grid(static_cast<unsigned>((n_stars + kSelThreads - 1) / kSelThreads), BF_SELECT_SPLIT);
  select_kernel<false><<<grid, kSelThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      n_groups, n_stars, ld_star, gmin, reinterpret_cast<const unsigned long long*>(starmin), tau,
      err_a, err_b, counts, nullptr, nullptr, nullptr);
  BF_CHECK_LAUNCH("select_kernel<count>");
  return BF_OK;
}

extern "C" int bf_select_fill(int64_t n_groups, int64_t n_stars, int64_t ld_star,
                              const float* gmin, const uint64_t* starmin, float tau,
                              float err_a, float err_b, const uint32_t* offsets,
                              uint32_t* list, uint32_t* entry_star, void* stream) {
  using namespace bf;
  BF_REQUIRE(n_groups > 0 && n_stars > 0 && ld_star >= n_stars, "bf_select_fill: bad sizes");
  BF_REQUIRE(gmin && starmin && offsets && list, "bf_select_fill: null pointer");
  dim3 grid(static_cast<unsigned>((n_stars + kSelThreads - 1) / kSelThreads), BF_SELECT_SPLIT);
  select_kernel<true><<<grid, kSelThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      n_groups, n_stars, ld_star, gmin, reinterpret_cast<const unsigned long long*>(starmin), tau,
      err_a, err_b, nullptr, offsets, list, entry_star);
  BF_CHECK_LAUNCH("select_kernel<fill>");
  return BF_OK;
}

extern "C" int64_t bf_scan_scratch_elems(int64_t n) {
  if (n <= 0) return 1;
  return (n + bf::kScanTile - 1) / bf::kScanTile;
}

extern "C" int bf_exclusive_scan_u32(int64_t n, const uint32_t* in, uint32_t* out,
                                     uint32_t* scratch, void* stream) {
  using namespace bf;
  BF_REQUIRE(n > 0, "bf_exclusive_scan_u32: n must be positive");
  BF_REQUIRE(in && out && scratch, "bf_exclusive_scan_u32: null pointer");
  BF_REQUIRE((n + kScanTile - 1) / kScanTile < (1LL << 31), "bf_exclusive_scan_u32: n too large");
  return run_scan(n, in, out, scratch, static_cast<cudaStream_t>(stream));
}

/* Prefix sum for callers that cannot give the library any scratch memory -- see the header.
 *
 * `g_scan_tile_sums` above is only a host-side handle for the device array: passing that name
 * to a kernel from host code would hand the kernel a host address.  The real device address
 * is looked up with cudaGetSymbolAddress, once per GPU, while that GPU is the current device,
 * and remembered per device ordinal. */
extern "C" int bf_exclusive_scan_u32_noalloc(int64_t n, const uint32_t* in, uint32_t* out,
                                             void* stream) {
  using namespace bf;
  BF_REQUIRE(n > 0, "bf_exclusive_scan_u32_noalloc: n must be positive");
  BF_REQUIRE(in && out, "bf_exclusive_scan_u32_noalloc: null pointer");
  const int64_t n_tiles = (n + kScanTile - 1) / kScanTile;
  if (n_tiles > kStaticScanTiles) {
    set_error("bf_exclusive_scan_u32_noalloc: n=%lld exceeds the built-in scratch (at most %lld "
              "elements); use bf_exclusive_scan_u32 with caller-provided scratch instead",
              static_cast<long long>(n), static_cast<long long>(kStaticScanTiles * kScanTile));
    return BF_ERR_TOO_LARGE;
  }
  int dev = 0;
  BF_CUDA_TRY(cudaGetDevice(&dev));
  // one slot per device ordinal; atomic because several host threads may make their first call
  // at once (they would all store the same address)
  static std::atomic<uint32_t*> scratch_of_device[64];
  if (dev < 0 || dev >= 64) {
    set_error("bf_exclusive_scan_u32_noalloc: device ordinal %d not supported", dev);
    return BF_ERR_UNSUPPORTED;
  }
  if (stream != nullptr) {
    // the array that is looked up belongs to the current GPU: a stream of another GPU would run
    // the kernels there with this GPU's address
    int sdev = dev;
    const cudaError_t e = cudaStreamGetDevice(static_cast<cudaStream_t>(stream), &sdev);
    if (e == cudaSuccess && sdev != dev) {
      set_error("bf_exclusive_scan_u32_noalloc: the stream belongs to GPU %d but GPU %d is current; "
                "make the stream's GPU current for every call", sdev, dev);
      return BF_ERR_INVALID;
    }
    if (e != cudaSuccess) (void)cudaGetLastError();  // (older drivers: no such query; rule stays documented)
  }
  uint32_t* scratch = scratch_of_device[dev].load(std::memory_order_acquire);
  if (scratch == nullptr) {
    BF_CUDA_TRY(cudaGetSymbolAddress(reinterpret_cast<void**>(&scratch), g_scan_tile_sums));
    scratch_of_device[dev].store(scratch, std::memory_order_release);
  }
  // every tile sum of this call is written by the first kernel before the second reads it,
  // so nothing left over from an earlier call has to be cleared
  return run_scan(n, in, out, scratch, static_cast<cudaStream_t>(stream));
}
