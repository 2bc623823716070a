// Error plumbing and device queries of the C ABI (include/beast_b200.h).

#include "common.cuh"

namespace bf {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("%s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
  return BF_ERR_CUDA;
}

}  // namespace bf

extern "C" const char* bf_last_error(void) { return bf::g_err; }

extern "C" int bf_version(void) { return 100; }

extern "C" int bf_device_info(int device, int* sm_count, int* smem_optin_bytes, int* clock_khz) {
  using namespace bf;
  cudaDeviceProp prop;
  BF_CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    set_error("bf_device_info: device %d is sm_%d%d; this library is built for sm_100a only",
              device, prop.major, prop.minor);
    return BF_ERR_UNSUPPORTED;
  }
  if (sm_count) *sm_count = prop.multiProcessorCount;
  if (smem_optin_bytes) *smem_optin_bytes = static_cast<int>(prop.sharedMemPerBlockOptin);
  if (clock_khz) {
    int khz = 0;
    BF_CUDA_TRY(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, device));
    *clock_khz = khz;
  }
  return BF_OK;
}
