// Shared host-side helpers for libse_b200.so (status codes, last-error string, launch checks).
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include "../../include/se_b200.h"

namespace se {

inline char* last_error_buf() {
  static thread_local char buf[512] = "";
  return buf;
}

inline int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(last_error_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}

// kernels launched by this library since load (every launch goes through check_launch)
extern std::atomic<int64_t> g_launch_count;

inline int check_launch(const char* what) {
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(SE_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
  return SE_OK;
}

#define SE_REQUIRE(cond, ...) \
  do { if (!(cond)) return se::fail(SE_ERR_BAD_ARG, __VA_ARGS__); } while (0)
#define SE_CUDA(call) \
  do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
         return se::fail(SE_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); } while (0)

// Per-device host-side state.  cudaFuncSetAttribute (the > 48 KB dynamic shared memory opt-in) and the SM count
// belong to a DEVICE, not to the process: a flag set while cuda:0 was current says nothing about cuda:1
// (model.to('cuda:1'), nn.DataParallel, tests that touch two GPUs in one process).  `get()` is the slot of the
// device that is current on the calling thread.  Concurrent first calls may both run the (idempotent) set-up.
constexpr int kMaxDevices = 64;
inline int current_device_slot() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) dev = 0;
  return dev;
}
template <typename T> struct PerDevice {
  T v[kMaxDevices];
  explicit PerDevice(T init) { for (int i = 0; i < kMaxDevices; ++i) v[i] = init; }
  T& get() { return v[current_device_slot()]; }
};

inline int num_sms() {
  static PerDevice<int> count(0);
  int& n = count.get();
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
  }
  return n;
}

inline unsigned grid_1d(int64_t total, int threads = 256, int waves = 16) {
  int64_t blocks = (total + threads - 1) / threads, cap = (int64_t)num_sms() * waves;
  return (unsigned)(blocks < cap ? blocks : cap);
}

template <typename T> __host__ __device__ inline T ceil_div(T a, T b) { return (a + b - 1) / b; }

}  // namespace se
