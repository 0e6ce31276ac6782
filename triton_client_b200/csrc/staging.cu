// Pinned staging for copies between ordinary (pageable) host memory and the device.
//
// The reference hands pageable numpy memory straight to cudaMemcpyAsync
// (/root/reference/src/python/library/tritonclient/utils/cuda_shared_memory/__init__.py:212-231)
// and reads a region back into a pageable ctypes buffer (:266-276); the driver then stages the
// data itself, one thread, chunk by chunk, and the copy runs far below the PCIe rate. Here large
// pageable copies go through a small ring of pinned buffers per device: a few host threads fill
// (or drain) one ring slot with memcpy while the DMA engine moves the previous slot, so the host
// side is parallel and both sides overlap.
//
// Semantics match what callers of cudaMemcpyAsync on pageable memory already rely on: when an
// H2D call returns, the source buffer has been consumed (the DMA of the last slots may still be
// in flight on `stream`); a D2H call returns with the destination complete.
#include <condition_variable>
#include <cstdlib>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

#include "common.cuh"

namespace tcb {

constexpr size_t kStageSlot = 4u << 20;    // bytes per ring slot
constexpr int kStageSlots = 4;             // ring depth (16 MiB of pinned memory per device)
constexpr size_t kStageMin = 2u << 20;     // below this a plain cudaMemcpyAsync is as good

// ---- a tiny worker pool for the host-side memcpy ------------------------------------------------
class CopyPool {
 public:
  static CopyPool& get() {
    static CopyPool* pool = new CopyPool();  // never destroyed: workers may outlive static teardown
    return *pool;
  }
  int workers() const { return (int)threads_.size(); }
  // dst[0, n) = src[0, n), split across the workers and the calling thread
  void copy(void* dst, const void* src, size_t n) {
    const int parts = workers() + 1;
    if (parts == 1 || n < (256u << 10)) { memcpy(dst, src, n); return; }
    const size_t per = ((n + parts - 1) / parts + 63) & ~(size_t)63;
    {
      std::lock_guard<std::mutex> lk(mu_);
      dst_ = static_cast<char*>(dst); src_ = static_cast<const char*>(src); n_ = n; per_ = per;
      pending_ = workers();
      generation_++;
    }
    cv_.notify_all();
    if (n > 0) memcpy(dst, src, std::min(per, n));  // part 0 on the caller
    std::unique_lock<std::mutex> lk(mu_);
    done_.wait(lk, [&] { return pending_ == 0; });
  }

 private:
  CopyPool() {
    int want = 6;
    if (const char* e = getenv("TCB_STAGING_THREADS")) want = atoi(e);
    const int hw = (int)std::thread::hardware_concurrency();
    if (hw > 0) want = std::min(want, std::max(0, hw - 1));
    for (int i = 0; i < want; i++) threads_.emplace_back([this, i] { run(i + 1); });
    for (auto& t : threads_) t.detach();
  }
  void run(int part) {
    uint64_t seen = 0;
    for (;;) {
      char* d; const char* s; size_t n, per;
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return generation_ != seen; });
        seen = generation_;
        d = dst_; s = src_; n = n_; per = per_;
      }
      const size_t lo = per * (size_t)part;
      if (lo < n) memcpy(d + lo, s + lo, std::min(per, n - lo));
      {
        std::lock_guard<std::mutex> lk(mu_);
        if (--pending_ == 0) done_.notify_one();
      }
    }
  }
  std::vector<std::thread> threads_;
  std::mutex mu_;
  std::condition_variable cv_, done_;
  uint64_t generation_ = 0;
  int pending_ = 0;
  char* dst_ = nullptr;
  const char* src_ = nullptr;
  size_t n_ = 0, per_ = 0;
};

struct StageRing {
  char* slot[kStageSlots] = {nullptr};
  cudaEvent_t ev[kStageSlots] = {nullptr};
  bool busy[kStageSlots] = {false};  // ev[i] marks a DMA out of slot i that nobody has waited for yet
  int next = 0;                      // slot the next host-to-device call starts with (oldest DMA first)
  bool ready = false;
};

static std::mutex g_stage_mu;  // one staged copy at a time per process (the ring and the pool are shared)
static StageRing g_rings[64];

static int ring_for(int device, StageRing** out) {
  TCB_CHECK_ARG(device >= 0 && device < 64, "staging: bad device %d", device);
  StageRing& r = g_rings[device];
  if (!r.ready) {
    for (int i = 0; i < kStageSlots; i++) {
      cudaError_t e = cudaHostAlloc((void**)&r.slot[i], kStageSlot, cudaHostAllocPortable);
      if (e == cudaSuccess) e = cudaEventCreateWithFlags(&r.ev[i], cudaEventDisableTiming);
      if (e != cudaSuccess)
        return set_error(TCB_ERR_NOMEM, "staging: pinned ring allocation failed: %s", cudaGetErrorString(e));
    }
    r.ready = true;
  }
  *out = &r;
  return TCB_OK;
}

bool staging_enabled() {
  static const bool on = !(getenv("TCB_STAGING") && atoi(getenv("TCB_STAGING")) == 0);
  return on;
}

// Is `p` ordinary host memory the CUDA runtime knows nothing about?
bool is_pageable_host(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeUnregistered;
}

bool is_device_ptr(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeDevice;
}

bool staging_applies(const void* host_ptr, uint64_t bytes) {
  return staging_enabled() && bytes >= kStageMin && is_pageable_host(host_ptr);
}

// Device current = `device` (caller holds a DeviceGuard).
int staged_h2d(int device, void* dst_dev, const void* src_host, uint64_t bytes, cudaStream_t s) {
  std::lock_guard<std::mutex> lock(g_stage_mu);
  StageRing* r;
  int rc = ring_for(device, &r);
  if (rc) return rc;
  CopyPool& pool = CopyPool::get();
  const char* src = static_cast<const char*>(src_host);
  char* dst = static_cast<char*>(dst_dev);
  uint64_t off = 0;
  for (int k = 0; off < bytes; k++) {
    const int i = (r->next + k) % kStageSlots;
    const size_t n = (size_t)std::min<uint64_t>(kStageSlot, bytes - off);
    // the DMA that last read this slot - of this call or of an earlier one: a call does not drain the
    // ring when it ends (the source has been consumed by then, which is all the caller is promised), so
    // the first slots of the next call fill while the last DMAs of this one are still running. Draining
    // per call cost ~1 ms per call: 31 ms for 1.6 GB in one call, 40 ms in 8 calls, 67 ms in 32.
    if (r->busy[i]) TCB_CUDA(cudaEventSynchronize(r->ev[i]));
    pool.copy(r->slot[i], src + off, n);
    TCB_CUDA(cudaMemcpyAsync(dst + off, r->slot[i], n, cudaMemcpyHostToDevice, s));
    TCB_CUDA(cudaEventRecord(r->ev[i], s));
    r->busy[i] = true;
    off += n;
  }
  r->next = (int)((r->next + (bytes + kStageSlot - 1) / kStageSlot) % kStageSlots);
  return TCB_OK;
}

int staged_d2h(int device, void* dst_host, const void* src_dev, uint64_t bytes, cudaStream_t s) {
  std::lock_guard<std::mutex> lock(g_stage_mu);
  StageRing* r;
  int rc = ring_for(device, &r);
  if (rc) return rc;
  CopyPool& pool = CopyPool::get();
  char* dst = static_cast<char*>(dst_host);
  const char* src = static_cast<const char*>(src_dev);
  const uint64_t chunks = (bytes + kStageSlot - 1) / kStageSlot;
  auto issue = [&](uint64_t k) -> cudaError_t {
    const int i = (int)(k % kStageSlots);
    const size_t n = (size_t)std::min<uint64_t>(kStageSlot, bytes - k * kStageSlot);
    if (r->busy[i]) {  // a host-to-device DMA of an earlier call may still be reading this slot
      cudaError_t w = cudaEventSynchronize(r->ev[i]);
      if (w != cudaSuccess) return w;
      r->busy[i] = false;
    }
    cudaError_t e = cudaMemcpyAsync(r->slot[i], src + k * kStageSlot, n, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaEventRecord(r->ev[i], s);
    return e;
  };
  for (uint64_t k = 0; k < std::min<uint64_t>(chunks, kStageSlots - 1); k++) TCB_CUDA(issue(k));
  for (uint64_t k = 0; k < chunks; k++) {
    if (k + kStageSlots - 1 < chunks) TCB_CUDA(issue(k + kStageSlots - 1));  // keep the DMA engine ahead
    const int i = (int)(k % kStageSlots);
    const size_t n = (size_t)std::min<uint64_t>(kStageSlot, bytes - k * kStageSlot);
    TCB_CUDA(cudaEventSynchronize(r->ev[i]));
    pool.copy(dst + k * kStageSlot, r->slot[i], n);
  }
  return TCB_OK;
}

}  // namespace tcb
