// api.cu - error plumbing, version and device info of libb200tts.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace b2 {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("%s: CUDA error %d (%s)", what, static_cast<int>(e), cudaGetErrorString(e));
  return B2TTS_ECUDA;
}

static unsigned long long g_launches = 0;
void count_launch() { __atomic_add_fetch(&g_launches, 1ull, __ATOMIC_RELAXED); }

int sm_count() {
  // per-device constant cache (the only global state the library keeps)
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

}  // namespace b2

extern "C" {

int b2tts_version(void) { return 2; }

const char* b2tts_last_error(void) { return b2::g_err; }

unsigned long long b2tts_launch_count(void) { return __atomic_load_n(&b2::g_launches, __ATOMIC_RELAXED); }

int b2tts_device_info(int* sm_count, int* max_smem_bytes, int* cc_major, int* cc_minor) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return b2::cuda_fail(e, "b2tts_device_info");
  int v = 0;
  if (sm_count) {
    e = cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return b2::cuda_fail(e, "b2tts_device_info");
    *sm_count = v;
  }
  if (max_smem_bytes) {
    e = cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (e != cudaSuccess) return b2::cuda_fail(e, "b2tts_device_info");
    *max_smem_bytes = v;
  }
  if (cc_major) {
    e = cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMajor, dev);
    if (e != cudaSuccess) return b2::cuda_fail(e, "b2tts_device_info");
    *cc_major = v;
  }
  if (cc_minor) {
    e = cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMinor, dev);
    if (e != cudaSuccess) return b2::cuda_fail(e, "b2tts_device_info");
    *cc_minor = v;
  }
  return B2TTS_OK;
}

}  // extern "C"
