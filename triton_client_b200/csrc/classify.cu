// Field split of classification strings on the device (the step after the
// BYTES decode).
//
// The reference's example turns every element of a classification output ("score:index:label")
// into a Python string character by character and splits it on ':'
// (/root/reference/src/python/examples/image_client.py:196-216). Here the elements stay where
// TcbBytesDeserialize left them (payload + int64 offsets on the device); one warp per element finds
// the first three ':' with a ballot per 32 bytes, so that the host only has to slice three fields
// out of one payload copy - no per-byte Python work, no per-element objects before the print.
#include "common.cuh"

namespace tcb {

// fields[4 i + k], k = 0..2: position of the (k + 1)-th ':' inside element i, or -1 if it has fewer;
// fields[4 i + 3]: the element's length.
__global__ void __launch_bounds__(256)
k_class_fields(const uint8_t* __restrict__ payload, const int64_t* __restrict__ offsets, int64_t n,
               int64_t* __restrict__ fields) {
  const int lane = threadIdx.x & 31;
  const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += warps) {
    const int64_t b = offsets[i], len = offsets[i + 1] - b;
    int64_t p0 = -1, p1 = -1, p2 = -1;
    for (int64_t base = 0; base < len && p2 < 0; base += 32) {
      const int64_t j = base + lane;
      unsigned m = __ballot_sync(0xffffffffu, j < len && payload[b + j] == (uint8_t)':');
      while (m && p2 < 0) {
        const int64_t p = base + (__ffs(m) - 1);
        m &= m - 1;
        if (p0 < 0) p0 = p;
        else if (p1 < 0) p1 = p;
        else p2 = p;
      }
    }
    if (lane == 0) {
      fields[4 * i] = p0; fields[4 * i + 1] = p1; fields[4 * i + 2] = p2; fields[4 * i + 3] = len;
    }
  }
}

}  // namespace tcb

extern "C" int TcbClassificationFields(int device_id, const void* payload_dev, const int64_t* offsets_dev,
                                       int64_t num_elements, int64_t* fields_dev, void* stream) {
  using namespace tcb;
  TCB_CHECK_ARG(num_elements >= 0, "TcbClassificationFields: negative element count");
  if (num_elements == 0) return TCB_OK;
  TCB_CHECK_ARG(offsets_dev && fields_dev, "TcbClassificationFields: NULL argument");
  TCB_DEVICE_GUARD(device_id);
  const int64_t want = (num_elements + 7) / 8;  // eight warps per CTA
  const int ctas = (int)std::min<int64_t>(want, (int64_t)sm_count(device_id) * 16);
  k_class_fields<<<ctas, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint8_t*>(payload_dev), offsets_dev, num_elements, fields_dev);
  count_launch();
  TCB_CUDA(cudaGetLastError());
  return TCB_OK;
}
