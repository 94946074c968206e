// ABI bookkeeping: version, thread-local last error, device queries.
#include <stdarg.h>
#include <stdlib.h>

#include <mutex>

#include "common.h"

namespace gcnb {

std::string& last_error_ref() {
  static thread_local std::string s;
  return s;
}

int set_error(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  last_error_ref() = buf;
  return code;
}

int pointer_device(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return -1;
  }
  if (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) return a.device;
  return -1;
}

int sm_count(int device) {
  static std::mutex mu;
  static int cache[64];
  static bool init = false;
  std::lock_guard<std::mutex> lk(mu);
  if (!init) {
    for (int& c : cache) c = 0;
    init = true;
  }
  if (device < 0 || device >= 64) return 0;
  if (!cache[device]) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) {
      cudaGetLastError();
      return 0;
    }
    cache[device] = v;
    // Scratch of the kernels (prepared W images, partial sums) comes from the device's default stream-ordered pool.
    // Its release threshold is 0 by default: every synchronisation hands the memory back to the driver and the
    // next call pays for mapping it again (~1 ms for the 258 MB of partials the tall-K product once used).
    // Keep up to GCNB_POOL_KEEP_MB (default 1024; 0 = leave the pool alone) cached instead.
    const char* keep_env = getenv("GCNB_POOL_KEEP_MB");
    uint64_t keep = (uint64_t)(keep_env ? atoll(keep_env) : 1024) << 20;
    cudaMemPool_t pool;
    if (keep > 0 && cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
      uint64_t cur = 0;
      if (cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &cur) == cudaSuccess && cur < keep)
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
    cudaGetLastError();
  }
  return cache[device];
}

int require_gpu() {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0) {
    cudaGetLastError();
    return set_error(GCNB_ERR_NOGPU,
                     "no CUDA device visible (%s): gcn_b200 has no CPU fallback",
                     e == cudaSuccess ? "count = 0" : cudaGetErrorString(e));
  }
  return GCNB_OK;
}

}  // namespace gcnb

extern "C" {

int gcnb_abi_version(void) { return GCNB_ABI_VERSION; }

const char* gcnb_last_error(void) { return gcnb::last_error_ref().c_str(); }

int gcnb_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

}  // extern "C"
