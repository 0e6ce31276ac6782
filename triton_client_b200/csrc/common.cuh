// Shared host/device helpers for libtcb200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>

#include "../../include/tcb200.h"

namespace tcb {

// ---- thread-local error text ------------------------------------------------------------------
char* last_error_buf();
int set_error(int code, const char* fmt, ...);
extern std::atomic<uint64_t> g_kernel_launches;
inline void count_launch(uint64_t n = 1) { g_kernel_launches.fetch_add(n, std::memory_order_relaxed); }

#define TCB_CUDA(expr)                                                                       \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess)                                                                   \
      return ::tcb::set_error(TCB_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,                  \
                              cudaGetErrorString(_e), __FILE__, __LINE__);                   \
  } while (0)

#define TCB_CHECK_ARG(cond, ...)                                              \
  do {                                                                        \
    if (!(cond)) return ::tcb::set_error(TCB_ERR_INVALID_ARG, __VA_ARGS__);   \
  } while (0)

// Make `device` current for the scope and restore the previous device afterwards (the
// reference's create/__del__ do the same dance: cuda_shared_memory/__init__.py:128-147).
struct DeviceGuard {
  int prev = -1;
  cudaError_t err = cudaSuccess;
  explicit DeviceGuard(int device) {
    err = cudaGetDevice(&prev);
    if (err == cudaSuccess && prev != device) err = cudaSetDevice(device);
    else if (err == cudaSuccess) prev = -1;  // nothing to restore
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

#define TCB_DEVICE_GUARD(device)                                                             \
  ::tcb::DeviceGuard _guard(device);                                                         \
  if (_guard.err != cudaSuccess)                                                             \
    return ::tcb::set_error(TCB_ERR_CUDA, "cudaSetDevice(%d) failed: %s", (int)(device),     \
                            cudaGetErrorString(_guard.err));

inline int sm_count(int device) {
  static int cached[64] = {0};
  if (device >= 0 && device < 64 && cached[device]) return cached[device];
  int n = 148;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) n = 148;
  if (device >= 0 && device < 64) cached[device] = n;
  return n;
}

// pinned staging for large copies from/to pageable host memory (staging.cu)
bool staging_applies(const void* host_ptr, uint64_t bytes);
bool is_pageable_host(const void* p);
bool is_device_ptr(const void* p);
int staged_h2d(int device, void* dst_dev, const void* src_host, uint64_t bytes, cudaStream_t s);
int staged_d2h(int device, void* dst_host, const void* src_dev, uint64_t bytes, cudaStream_t s);

static inline uint64_t round_up(uint64_t x, uint64_t a) { return (x + a - 1) / a * a; }

}  // namespace tcb
