// Shared helpers of libgcn_b200.so: error reporting, device guard, cache-hinted loads.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

#include "../../include/gcn_b200.h"

namespace gcnb {

std::string& last_error_ref();
int set_error(int code, const char* fmt, ...);

#define GCNB_CHECK_CUDA(expr)                                                              \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess)                                                                 \
      return gcnb::set_error(GCNB_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,                \
                             cudaGetErrorString(_e), __FILE__, __LINE__);                  \
  } while (0)

#define GCNB_REQUIRE(cond, ...)                                   \
  do {                                                            \
    if (!(cond)) return gcnb::set_error(GCNB_ERR_ARG, __VA_ARGS__); \
  } while (0)

// Makes `device` current for the scope and restores the caller's device afterwards (torch keeps
// its own notion of the current device; never leave it changed).
struct DeviceGuard {
  int prev = -1;
  bool changed = false;
  cudaError_t status = cudaSuccess;
  explicit DeviceGuard(int device) {
    status = cudaGetDevice(&prev);
    if (status == cudaSuccess && prev != device) {
      status = cudaSetDevice(device);
      changed = (status == cudaSuccess);
    }
  }
  ~DeviceGuard() {
    if (changed) cudaSetDevice(prev);
  }
};

// device of a device pointer, or -1
int pointer_device(const void* p);
// cached per-device SM count (the reference queries cudaGetDeviceProperties on every launch,
// flexspmm.cu:506-508)
int sm_count(int device);
// GCNB_ERR_NOGPU unless a compute-capability-10.x device is present
int require_gpu();

}  // namespace gcnb
