// CUDA shared-memory regions, streams and staging buffers behind the C ABI (include/tcb200.h).
//
// Behaviour mirrors /root/reference/src/python/library/tritonclient/utils/cuda_shared_memory/
// __init__.py:107-149 (create), :152-170 (raw handle), :173-239 (set), :242-276 (get) and
// _utils.py:67-128 (region object, stream, device restore), with the handle style of libcshm
// (utils/shared_memory/shared_memory.cc:76-148).
#include <mutex>
#include <string>

#include "common.cuh"

namespace tcb {

std::atomic<uint64_t> g_kernel_launches{0};

char* last_error_buf() {
  static thread_local char buf[512] = {0};
  return buf;
}

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(last_error_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}

struct Region {
  std::string name;
  void* base = nullptr;
  uint64_t byte_size = 0;
  int device = 0;
  cudaIpcMemHandle_t ipc;
};

}  // namespace tcb

using tcb::Region;

extern "C" {

const char* TcbLastError(void) { return tcb::last_error_buf(); }
const char* TcbVersion(void) { return "tcb200 0.1 sm_100a"; }
uint64_t TcbKernelLaunchCount(void) { return tcb::g_kernel_launches.load(); }

int TcbRegionCreate(const char* triton_shm_name, uint64_t byte_size, int device_id, void** region) {
  TCB_CHECK_ARG(region != nullptr, "TcbRegionCreate: region out-pointer is NULL");
  *region = nullptr;
  TCB_DEVICE_GUARD(device_id);
  Region* r = new (std::nothrow) Region();
  if (!r) return tcb::set_error(TCB_ERR_NOMEM, "TcbRegionCreate: out of host memory");
  r->name = triton_shm_name ? triton_shm_name : "";
  r->byte_size = byte_size;
  r->device = device_id;
  memset(&r->ipc, 0, sizeof(r->ipc));
  if (byte_size > 0) {
    cudaError_t e = cudaMalloc(&r->base, byte_size);
    if (e != cudaSuccess) {
      delete r;
      return tcb::set_error(TCB_ERR_CUDA, "cudaMalloc(%llu) failed: %s",
                            (unsigned long long)byte_size, cudaGetErrorString(e));
    }
    e = cudaIpcGetMemHandle(&r->ipc, r->base);
    if (e != cudaSuccess) {
      cudaFree(r->base);
      delete r;
      return tcb::set_error(TCB_ERR_CUDA, "cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
    }
  }
  *region = r;
  return TCB_OK;
}

int TcbRegionInfo(void* region, uint64_t* base_addr, uint64_t* byte_size, int* device_id,
                  unsigned char ipc_handle[64]) {
  TCB_CHECK_ARG(region != nullptr, "TcbRegionInfo: region is NULL");
  Region* r = static_cast<Region*>(region);
  if (base_addr) *base_addr = reinterpret_cast<uint64_t>(r->base);
  if (byte_size) *byte_size = r->byte_size;
  if (device_id) *device_id = r->device;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
  if (ipc_handle) memcpy(ipc_handle, &r->ipc, 64);
  return TCB_OK;
}

int TcbRegionSet(void* region, uint64_t offset, const void* src, uint64_t byte_size, void* stream) {
  TCB_CHECK_ARG(region != nullptr, "TcbRegionSet: region is NULL");
  Region* r = static_cast<Region*>(region);
  if (offset > r->byte_size || byte_size > r->byte_size - offset)
    return tcb::set_error(TCB_ERR_OUT_OF_RANGE,
                          "TcbRegionSet: offset %llu + size %llu exceeds region of %llu bytes",
                          (unsigned long long)offset, (unsigned long long)byte_size,
                          (unsigned long long)r->byte_size);
  if (byte_size == 0) return TCB_OK;
  TCB_CHECK_ARG(src != nullptr, "TcbRegionSet: src is NULL");
  TCB_DEVICE_GUARD(r->device);
  if (tcb::staging_applies(src, byte_size))  // pageable numpy memory: pinned ring + parallel memcpy
    return tcb::staged_h2d(r->device, static_cast<char*>(r->base) + offset, src, byte_size,
                           static_cast<cudaStream_t>(stream));
  TCB_CUDA(cudaMemcpyAsync(static_cast<char*>(r->base) + offset, src, byte_size, cudaMemcpyDefault,
                           static_cast<cudaStream_t>(stream)));
  return TCB_OK;
}

int TcbRegionGet(void* region, uint64_t offset, void* dst, uint64_t byte_size, void* stream) {
  TCB_CHECK_ARG(region != nullptr, "TcbRegionGet: region is NULL");
  Region* r = static_cast<Region*>(region);
  if (offset > r->byte_size || byte_size > r->byte_size - offset)
    return tcb::set_error(TCB_ERR_OUT_OF_RANGE,
                          "TcbRegionGet: offset %llu + size %llu exceeds region of %llu bytes",
                          (unsigned long long)offset, (unsigned long long)byte_size,
                          (unsigned long long)r->byte_size);
  TCB_DEVICE_GUARD(r->device);
  if (byte_size > 0) {
    TCB_CHECK_ARG(dst != nullptr, "TcbRegionGet: dst is NULL");
    if (tcb::staging_applies(dst, byte_size))
      return tcb::staged_d2h(r->device, dst, static_cast<char*>(r->base) + offset, byte_size,
                             static_cast<cudaStream_t>(stream));
    TCB_CUDA(cudaMemcpyAsync(dst, static_cast<char*>(r->base) + offset, byte_size,
                             cudaMemcpyDefault, static_cast<cudaStream_t>(stream)));
  }
  TCB_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  return TCB_OK;
}

int TcbRegionDestroy(void* region) {
  TCB_CHECK_ARG(region != nullptr, "TcbRegionDestroy: region is NULL");
  Region* r = static_cast<Region*>(region);
  int rc = TCB_OK;
  if (r->base) {
    tcb::DeviceGuard g(r->device);
    cudaError_t e = g.err == cudaSuccess ? cudaFree(r->base) : g.err;
    if (e != cudaSuccess)
      rc = tcb::set_error(TCB_ERR_CUDA, "cudaFree failed: %s", cudaGetErrorString(e));
  }
  delete r;
  return rc;
}

int TcbGlobalStream(int device_id, void** stream) {
  TCB_CHECK_ARG(stream != nullptr, "TcbGlobalStream: stream out-pointer is NULL");
  TCB_CHECK_ARG(device_id >= 0 && device_id < 64, "TcbGlobalStream: bad device %d", device_id);
  static std::mutex mu;
  static cudaStream_t streams[64] = {nullptr};
  std::lock_guard<std::mutex> lock(mu);
  if (!streams[device_id]) {
    TCB_DEVICE_GUARD(device_id);
    TCB_CUDA(cudaStreamCreate(&streams[device_id]));  // blocking stream, as _utils.py:108
  }
  *stream = streams[device_id];
  return TCB_OK;
}

int TcbStreamSync(int device_id, void* stream) {
  TCB_DEVICE_GUARD(device_id);
  TCB_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  return TCB_OK;
}

int TcbHostAlloc(uint64_t byte_size, void** ptr) {
  TCB_CHECK_ARG(ptr != nullptr, "TcbHostAlloc: ptr is NULL");
  *ptr = nullptr;
  if (byte_size == 0) return TCB_OK;
  cudaError_t e = cudaHostAlloc(ptr, byte_size, cudaHostAllocPortable);
  if (e != cudaSuccess)
    return tcb::set_error(TCB_ERR_NOMEM, "cudaHostAlloc(%llu) failed: %s",
                          (unsigned long long)byte_size, cudaGetErrorString(e));
  return TCB_OK;
}

int TcbHostFree(void* ptr) {
  if (ptr) TCB_CUDA(cudaFreeHost(ptr));
  return TCB_OK;
}

int TcbDeviceAlloc(int device_id, uint64_t byte_size, void** ptr_dev) {
  TCB_CHECK_ARG(ptr_dev != nullptr, "TcbDeviceAlloc: ptr is NULL");
  *ptr_dev = nullptr;
  if (byte_size == 0) return TCB_OK;
  TCB_DEVICE_GUARD(device_id);
  cudaError_t e = cudaMalloc(ptr_dev, byte_size);
  if (e != cudaSuccess)
    return tcb::set_error(TCB_ERR_NOMEM, "cudaMalloc(%llu) failed: %s",
                          (unsigned long long)byte_size, cudaGetErrorString(e));
  return TCB_OK;
}

int TcbDeviceFree(int device_id, void* ptr_dev) {
  if (!ptr_dev) return TCB_OK;
  TCB_DEVICE_GUARD(device_id);
  TCB_CUDA(cudaFree(ptr_dev));
  return TCB_OK;
}

int TcbMemcpy(int device_id, void* dst, const void* src, uint64_t byte_size, void* stream, int sync) {
  TCB_DEVICE_GUARD(device_id);
  if (byte_size > 0) {
    TCB_CHECK_ARG(dst != nullptr && src != nullptr, "TcbMemcpy: NULL pointer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (tcb::staging_applies(src, byte_size) && tcb::is_device_ptr(dst)) {
      const int rc = tcb::staged_h2d(device_id, dst, src, byte_size, s);
      if (rc) return rc;
    } else if (tcb::staging_applies(dst, byte_size) && tcb::is_device_ptr(src)) {
      return tcb::staged_d2h(device_id, dst, src, byte_size, s);  // returns with dst complete
    } else {
      TCB_CUDA(cudaMemcpyAsync(dst, src, byte_size, cudaMemcpyDefault, s));
    }
  }
  if (sync) TCB_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  return TCB_OK;
}

}  // extern "C"
