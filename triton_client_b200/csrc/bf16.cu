// BF16 wire codec behind the C ABI (include/tcb200.h).
//
// Replaces the per-element Python loops of
// /root/reference/src/python/library/tritonclient/utils/__init__.py:279-323 (serialize_bf16_tensor:
// struct.pack("<f", x)[2:4], i.e. the upper 16 bits of the float32 pattern - truncation, no
// rounding; signalling NaNs come out quiet, see quiet()) and :326-348 (deserialize_bf16_tensor:
// b"\x00\x00" + two bytes -> float32).
//
// Two streaming kernels, HBM-bound: 6 bytes of traffic per element either way. A thread moves
// four elements per step: one 128-bit load + one 64-bit store (pack) or the reverse (expand). Unaligned heads and tails are handled element-wise by the first CTA.
#include <algorithm>
#include <cstdlib>

#include "common.cuh"

namespace tcb {

constexpr int kBfThreads = 256;
constexpr int kBfUnroll = 8;  // 4-element groups in flight per thread

__device__ __forceinline__ uint4 ld_stream_v4(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ void st_stream_v4(void* p, uint4 v) {
  asm volatile("st.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// The reference packs through a Python float (struct.pack("<f", obj)) and unpacks through one
// (struct.unpack("<f", ...)): a double round trip, which turns a signalling NaN into a quiet one
// (mantissa bit 22 set). So 0x7F800001 serialises as 0x7FC0, not as the infinity 0x7F80, and the
// wire value 0x7F81 deserialises as 0x7FC10000. Every other pattern is cut / shifted unchanged.
__device__ __forceinline__ uint32_t quiet(uint32_t f32_bits) {
  return (f32_bits & 0x7fffffffu) > 0x7f800000u ? f32_bits | 0x00400000u : f32_bits;
}
// upper halves of two float32 patterns -> one word (element order = memory order)
__device__ __forceinline__ uint32_t pack2(uint32_t lo, uint32_t hi) {
  return __byte_perm(quiet(lo), quiet(hi), 0x7632);
}

__device__ __forceinline__ uint2 ld_stream_v2(const void* p) {
  uint2 v;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
  return v;
}
__device__ __forceinline__ void st_stream_v2(void* p, uint2 v) {
  asm volatile("st.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}

// A thread moves groups of FOUR elements: 16 bytes of float32 <-> 8 bytes of bfloat16, so that both
// the loads and the stores of a warp are one contiguous run (512 / 256 bytes). (With eight elements
// per thread the expand kernel's two 16-byte stores were 32 bytes apart from lane to lane and every
// store instruction filled only half of each sector it touched: 59 % of the copy rate.)
constexpr int kBfGroup = 4;

// src: n float32 (4-byte aligned), dst: n uint16 (2-byte aligned). `head` elements are peeled so
// that the vector body reads 16-byte aligned float4s; DST_VEC: the body's destination is 8-byte
// aligned too (decided on the host).
template <bool DST_VEC>
__global__ void __launch_bounds__(kBfThreads)
k_bf16_pack(const uint32_t* __restrict__ src, uint16_t* __restrict__ dst, unsigned long long n,
            unsigned int head) {
  const unsigned long long groups = (n - head) / kBfGroup;
  const unsigned long long body_end = head + groups * kBfGroup;
  const uint4* s4 = reinterpret_cast<const uint4*>(src + head);
  const unsigned long long stride = (unsigned long long)gridDim.x * kBfThreads;
  for (unsigned long long g0 = (unsigned long long)blockIdx.x * kBfThreads + threadIdx.x; g0 < groups;
       g0 += stride * kBfUnroll) {
    uint4 a[kBfUnroll];
#pragma unroll
    for (int u = 0; u < kBfUnroll; u++) {
      const unsigned long long g = g0 + u * stride;
      if (g < groups) a[u] = ld_stream_v4(s4 + g);
    }
#pragma unroll
    for (int u = 0; u < kBfUnroll; u++) {
      const unsigned long long g = g0 + u * stride;
      if (g >= groups) continue;
      const uint2 o = make_uint2(pack2(a[u].x, a[u].y), pack2(a[u].z, a[u].w));
      uint16_t* d = dst + head + kBfGroup * g;
      if (DST_VEC) {
        st_stream_v2(d, o);
      } else {  // destination only 2-byte aligned relative to the source: halfword stores
        d[0] = (uint16_t)o.x; d[1] = (uint16_t)(o.x >> 16); d[2] = (uint16_t)o.y; d[3] = (uint16_t)(o.y >> 16);
      }
    }
  }
  if (blockIdx.x == 0) {  // head and tail, element-wise
    for (unsigned long long i = threadIdx.x; i < head; i += kBfThreads) dst[i] = (uint16_t)(quiet(src[i]) >> 16);
    for (unsigned long long i = body_end + threadIdx.x; i < n; i += kBfThreads) dst[i] = (uint16_t)(quiet(src[i]) >> 16);
  }
}

// src: n uint16, dst: n float32 patterns (bf16 << 16). `head` elements peeled so that the body
// reads 8-byte aligned vectors; DST_VEC: the body's destination is 16-byte aligned too.
template <bool DST_VEC>
__global__ void __launch_bounds__(kBfThreads)
k_bf16_expand(const uint16_t* __restrict__ src, uint32_t* __restrict__ dst, unsigned long long n,
              unsigned int head) {
  const unsigned long long groups = (n - head) / kBfGroup;
  const unsigned long long body_end = head + groups * kBfGroup;
  const uint2* s2 = reinterpret_cast<const uint2*>(src + head);
  const unsigned long long stride = (unsigned long long)gridDim.x * kBfThreads;
  for (unsigned long long g0 = (unsigned long long)blockIdx.x * kBfThreads + threadIdx.x; g0 < groups;
       g0 += stride * kBfUnroll) {
    uint2 a[kBfUnroll];
#pragma unroll
    for (int u = 0; u < kBfUnroll; u++) {
      const unsigned long long g = g0 + u * stride;
      if (g < groups) a[u] = ld_stream_v2(s2 + g);
    }
#pragma unroll
    for (int u = 0; u < kBfUnroll; u++) {
      const unsigned long long g = g0 + u * stride;
      if (g >= groups) continue;
      const uint4 o = make_uint4(quiet(a[u].x << 16), quiet(a[u].x & 0xffff0000u), quiet(a[u].y << 16),
                                 quiet(a[u].y & 0xffff0000u));
      uint32_t* d = dst + head + kBfGroup * g;
      if (DST_VEC) {
        st_stream_v4(d, o);
      } else {
        d[0] = o.x; d[1] = o.y; d[2] = o.z; d[3] = o.w;
      }
    }
  }
  if (blockIdx.x == 0) {
    for (unsigned long long i = threadIdx.x; i < head; i += kBfThreads) dst[i] = quiet((uint32_t)src[i] << 16);
    for (unsigned long long i = body_end + threadIdx.x; i < n; i += kBfThreads) dst[i] = quiet((uint32_t)src[i] << 16);
  }
}

static unsigned bf16_grid(int device, unsigned long long groups) {
  // one CTA per 8 K elements up to 64 per SM; grid-stride covers the rest
  const unsigned long long want = (groups + (unsigned long long)kBfThreads * kBfUnroll - 1) /
                                  ((unsigned long long)kBfThreads * kBfUnroll);
  static const int per_sm = getenv("TCB_BF16_CTAS_PER_SM") ? atoi(getenv("TCB_BF16_CTAS_PER_SM")) : 64;  // one pass, no grid-stride tail (8: 0.75-0.80 of the copy rate, 64: 0.83)
  const unsigned long long cap = (unsigned long long)sm_count(device) * per_sm;
  return (unsigned)std::max<unsigned long long>(1, std::min(want, cap));
}

}  // namespace tcb

using namespace tcb;

extern "C" {

int TcbBf16Serialize(int device_id, const void* src_f32_dev, uint64_t n_elements, void* dst_dev,
                     uint64_t dst_capacity, void* stream) {
  if (n_elements == 0) return TCB_OK;
  TCB_CHECK_ARG(src_f32_dev != nullptr && dst_dev != nullptr, "TcbBf16Serialize: NULL pointer");
  TCB_CHECK_ARG(reinterpret_cast<uintptr_t>(src_f32_dev) % 4 == 0, "TcbBf16Serialize: source not 4-byte aligned");
  TCB_CHECK_ARG(reinterpret_cast<uintptr_t>(dst_dev) % 2 == 0, "TcbBf16Serialize: destination not 2-byte aligned");
  if (2 * n_elements > dst_capacity)
    return set_error(TCB_ERR_OUT_OF_RANGE, "TcbBf16Serialize: %llu elements need %llu bytes, capacity %llu",
                     (unsigned long long)n_elements, (unsigned long long)(2 * n_elements),
                     (unsigned long long)dst_capacity);
  TCB_DEVICE_GUARD(device_id);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const uint32_t* src = static_cast<const uint32_t*>(src_f32_dev);
  uint16_t* dst = static_cast<uint16_t*>(dst_dev);
  // elements to peel so that the source body is 16-byte aligned
  unsigned head = (unsigned)(((16 - (reinterpret_cast<uintptr_t>(src) & 15)) & 15) / 4);
  if (head > n_elements) head = (unsigned)n_elements;
  const bool dst_vec = (reinterpret_cast<uintptr_t>(dst + head) & 7) == 0;
  const unsigned grid = bf16_grid(device_id, (n_elements - head) / kBfGroup);
  if (dst_vec) k_bf16_pack<true><<<grid, kBfThreads, 0, s>>>(src, dst, n_elements, head);
  else k_bf16_pack<false><<<grid, kBfThreads, 0, s>>>(src, dst, n_elements, head);
  count_launch();
  TCB_CUDA(cudaGetLastError());
  return TCB_OK;
}

int TcbBf16Deserialize(int device_id, const void* src_dev, uint64_t src_bytes, void* dst_f32_dev,
                       uint64_t dst_capacity, void* stream) {
  if (src_bytes == 0) return TCB_OK;
  if (src_bytes & 1)  // the reference's unpack_from("<2s") raises struct.error on the odd byte
    return set_error(TCB_ERR_TRUNCATED, "TcbBf16Deserialize: %llu bytes is not a whole number of bfloat16 values",
                     (unsigned long long)src_bytes);
  const uint64_t n = src_bytes / 2;
  TCB_CHECK_ARG(src_dev != nullptr && dst_f32_dev != nullptr, "TcbBf16Deserialize: NULL pointer");
  TCB_CHECK_ARG(reinterpret_cast<uintptr_t>(src_dev) % 2 == 0, "TcbBf16Deserialize: source not 2-byte aligned");
  TCB_CHECK_ARG(reinterpret_cast<uintptr_t>(dst_f32_dev) % 4 == 0, "TcbBf16Deserialize: destination not 4-byte aligned");
  if (4 * n > dst_capacity)
    return set_error(TCB_ERR_OUT_OF_RANGE, "TcbBf16Deserialize: %llu elements need %llu bytes, capacity %llu",
                     (unsigned long long)n, (unsigned long long)(4 * n), (unsigned long long)dst_capacity);
  TCB_DEVICE_GUARD(device_id);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const uint16_t* src = static_cast<const uint16_t*>(src_dev);
  uint32_t* dst = static_cast<uint32_t*>(dst_f32_dev);
  unsigned head = (unsigned)(((8 - (reinterpret_cast<uintptr_t>(src) & 7)) & 7) / 2);
  if (head > n) head = (unsigned)n;
  const bool dst_vec = (reinterpret_cast<uintptr_t>(dst + head) & 15) == 0;
  const unsigned grid = bf16_grid(device_id, (n - head) / kBfGroup);
  if (dst_vec) k_bf16_expand<true><<<grid, kBfThreads, 0, s>>>(src, dst, n, head);
  else k_bf16_expand<false><<<grid, kBfThreads, 0, s>>>(src, dst, n, head);
  count_launch();
  TCB_CUDA(cudaGetLastError());
  return TCB_OK;
}

}  // extern "C"
