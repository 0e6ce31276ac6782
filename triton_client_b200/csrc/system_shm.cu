// POSIX system shared-memory regions behind the C ABI (include/tcb200.h), plus the device bridge.
//
// Replaces libcshm, /root/reference/src/python/library/tritonclient/utils/shared_memory/
// shared_memory.cc:76-148 (SharedMemoryRegionCreate / Set / GetSharedMemoryHandleInfo / Destroy):
// shm_open + ftruncate + mmap of the whole object, memcpy in, munmap + shm_unlink. Return codes
// are libcshm's (-2 descriptor, -3 size, -4 mmap, -5 unlink, -6 munmap) so that the Python
// layer's SharedMemoryException texts stay the reference's.
//
// Added for this library: the mapping is page-locked with cudaHostRegister when a CUDA device is
// present, so the kernels' output can be copied into (and inputs out of) the region by DMA at
// PCIe rate without a pageable bounce, and the bounds of every copy are checked (libcshm's
// memcpy is not).
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <string>

#include "common.cuh"

namespace tcb {

struct SysRegion {
  std::string name, key;
  void* base = nullptr;
  uint64_t byte_size = 0;
  int fd = -1;
  bool pinned = false;
};

}  // namespace tcb

using tcb::SysRegion;

extern "C" {

int TcbShmRegionCreate(const char* triton_shm_name, const char* shm_key, uint64_t byte_size,
                       int pin, void** shm_handle) {
  if (!shm_handle || !shm_key) return tcb::set_error(TCB_ERR_INVALID_ARG, "TcbShmRegionCreate: NULL argument");
  *shm_handle = nullptr;
  const int fd = shm_open(shm_key, O_RDWR | O_CREAT, S_IRUSR | S_IWUSR);
  if (fd == -1) return tcb::set_error(-2, "unable to get shared memory descriptor");
  if (ftruncate(fd, (off_t)byte_size) == -1) {
    close(fd);
    return tcb::set_error(-3, "unable to initialize the size");
  }
  void* addr = nullptr;
  if (byte_size > 0) {
    addr = mmap(nullptr, byte_size, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    if (addr == MAP_FAILED) {
      close(fd);
      return tcb::set_error(-4, "unable to read/mmap the shared memory region");
    }
  }
  close(fd);  // libcshm closes the descriptor once the object is mapped
  SysRegion* r = new (std::nothrow) SysRegion();
  if (!r) return tcb::set_error(TCB_ERR_NOMEM, "out of host memory");
  r->name = triton_shm_name ? triton_shm_name : "";
  r->key = shm_key;
  r->base = addr;
  r->byte_size = byte_size;
  r->fd = fd;
  if (pin && byte_size > 0) {  // best effort: without a device the region is simply not page-locked
    if (cudaHostRegister(addr, byte_size, cudaHostRegisterPortable) == cudaSuccess) r->pinned = true;
    else cudaGetLastError();
  }
  *shm_handle = r;
  return TCB_OK;
}

int TcbShmRegionSet(void* shm_handle, uint64_t offset, uint64_t byte_size, const void* data) {
  if (!shm_handle) return tcb::set_error(TCB_ERR_INVALID_ARG, "TcbShmRegionSet: handle is NULL");
  SysRegion* r = static_cast<SysRegion*>(shm_handle);
  if (offset > r->byte_size || byte_size > r->byte_size - offset)
    return tcb::set_error(TCB_ERR_OUT_OF_RANGE, "TcbShmRegionSet: offset %llu + size %llu exceeds region of %llu bytes",
                          (unsigned long long)offset, (unsigned long long)byte_size, (unsigned long long)r->byte_size);
  if (byte_size) memcpy(static_cast<char*>(r->base) + offset, data, byte_size);
  return TCB_OK;
}

int TcbShmRegionInfo(void* shm_handle, void** shm_addr, const char** shm_key, int* shm_fd, uint64_t* offset,
                     uint64_t* byte_size, int* pinned) {
  if (!shm_handle) return tcb::set_error(TCB_ERR_INVALID_ARG, "TcbShmRegionInfo: handle is NULL");
  SysRegion* r = static_cast<SysRegion*>(shm_handle);
  if (shm_addr) *shm_addr = r->base;
  if (shm_key) *shm_key = r->key.c_str();
  if (shm_fd) *shm_fd = r->fd;
  if (offset) *offset = 0;
  if (byte_size) *byte_size = r->byte_size;
  if (pinned) *pinned = r->pinned ? 1 : 0;
  return TCB_OK;
}

// Device <-> region copies on `stream` (to_region != 0: device -> region). Bounds-checked;
// synchronises the stream so that the bytes are in the region (or on the device) on return.
int TcbShmRegionCopyDevice(void* shm_handle, uint64_t offset, void* dev_ptr, uint64_t byte_size, int device_id,
                           int to_region, void* stream) {
  if (!shm_handle) return tcb::set_error(TCB_ERR_INVALID_ARG, "TcbShmRegionCopyDevice: handle is NULL");
  SysRegion* r = static_cast<SysRegion*>(shm_handle);
  if (offset > r->byte_size || byte_size > r->byte_size - offset)
    return tcb::set_error(TCB_ERR_OUT_OF_RANGE, "TcbShmRegionCopyDevice: offset %llu + size %llu exceeds region of %llu bytes",
                          (unsigned long long)offset, (unsigned long long)byte_size, (unsigned long long)r->byte_size);
  if (byte_size == 0) return TCB_OK;
  char* host = static_cast<char*>(r->base) + offset;
  return to_region ? TcbMemcpy(device_id, host, dev_ptr, byte_size, stream, 1)
                   : TcbMemcpy(device_id, dev_ptr, host, byte_size, stream, 1);
}

int TcbShmRegionDestroy(void* shm_handle) {
  if (!shm_handle) return tcb::set_error(TCB_ERR_INVALID_ARG, "TcbShmRegionDestroy: handle is NULL");
  SysRegion* r = static_cast<SysRegion*>(shm_handle);
  int rc = TCB_OK;
  if (r->pinned) {
    if (cudaHostUnregister(r->base) != cudaSuccess) cudaGetLastError();
  }
  if (r->base && munmap(r->base, r->byte_size) == -1) rc = tcb::set_error(-6, "unable to munmap the shared memory region");
  if (rc == TCB_OK && shm_unlink(r->key.c_str()) == -1) rc = tcb::set_error(-5, "unable to unlink the shared memory region");
  delete r;
  return rc;
}

}  // extern "C"
