// Image preprocess: plan (host) + two-pass kernels (device).
//
// Replaces /root/reference/src/python/examples/image_client.py:154-193 `preprocess`: Pillow
// convert + resize(BILINEAR) (8-bit, 22-bit fixed-point coefficients, uint8 intermediate between
// the two separable passes; arithmetic restated in SURVEY.md Appendix A and oracle/), numpy
// astype/scaling (as a 256-entry table per channel supplied by the caller) and the NHWC/NCHW
// ordering, written straight to the destination (a registered CUDA shared-memory region).
//
// This file holds the plan and the general two-pass path: pass 1 writes the uint8 intermediate
// to plan-owned scratch, pass 2 resamples the other axis, applies the table and the layout. It
// handles every geometry (any size, either pass order, identity, single-axis). The fused
// single-kernel path for the common both-axes case lives in preprocess_fused.cu.
#include <algorithm>
#include <cmath>
#include <map>

#include "preprocess.cuh"

namespace tcb {

// ---- host: Pillow's coefficient set-up ----------------------------------------------------------
// precompute_coeffs + normalize_coeffs_8bpc for the BILINEAR filter (support 1.0) over the full
// axis. Double precision, operation order as in Pillow; this translation unit is compiled with
// -ffp-contract=off so no multiply-add is fused.
static inline double tri(double x) {
  if (x < 0.0) x = -x;
  if (x < 1.0) return 1.0 - x;
  return 0.0;
}

int compute_axis_coeffs(int in_size, int out_size, std::vector<int32_t>& bounds,
                        std::vector<int32_t>& kk) {
  const double scale = (double)in_size / (double)out_size;
  const double filterscale = scale < 1.0 ? 1.0 : scale;
  const double support = 1.0 * filterscale;
  const int ksize = (int)std::ceil(support) * 2 + 1;
  const double ss = 1.0 / filterscale;
  bounds.assign((size_t)out_size * 2, 0);
  kk.assign((size_t)out_size * ksize, 0);
  std::vector<double> w((size_t)ksize);
  for (int xx = 0; xx < out_size; xx++) {
    const double center = 0.0 + (xx + 0.5) * scale;
    double ww = 0.0;
    int xmin = (int)(center - support + 0.5);
    if (xmin < 0) xmin = 0;
    int xmax = (int)(center + support + 0.5);
    if (xmax > in_size) xmax = in_size;
    xmax -= xmin;
    for (int x = 0; x < xmax; x++) {
      const double v = tri((x + xmin - center + 0.5) * ss);
      w[x] = v;
      ww += v;
    }
    int32_t* k = &kk[(size_t)xx * ksize];
    for (int x = 0; x < xmax; x++) {
      if (ww != 0.0) w[x] /= ww;
      if (w[x] < 0)
        k[x] = (int)(-0.5 + w[x] * (1 << kPrecisionBits));
      else
        k[x] = (int)(0.5 + w[x] * (1 << kPrecisionBits));
    }
    bounds[xx * 2 + 0] = xmin;
    bounds[xx * 2 + 1] = xmax;
  }
  return ksize;
}

// ---- device helpers -------------------------------------------------------------------------------
__device__ __forceinline__ int clip8(int acc) {
  int v = acc >> kPrecisionBits;  // arithmetic shift, then clamp (Pillow's clip8 lookup)
  return min(max(v, 0), 255);
}

// One pixel of the converted image (image_client.py:161-164): C == 3 is convert("RGB"), C == 1 is
// convert("L") with Pillow's rgb2l weights.
template <int C>
__device__ __forceinline__ void fetch_px(const uint8_t* __restrict__ row, int cin, int x, int (&v)[C]) {
  if (C == 3) {
    if (cin == 1) {
      const int p = row[x];
      v[0] = p; v[1 % C] = p; v[2 % C] = p;
    } else {
      const uint8_t* q = row + (size_t)x * cin;
      v[0] = q[0]; v[1 % C] = q[1]; v[2 % C] = q[2];
    }
  } else {
    if (cin == 1) {
      v[0] = row[x];
    } else {
      const uint8_t* q = row + (size_t)x * cin;
      v[0] = (int)((q[0] * 19595u + q[1] * 38470u + q[2] * 7471u + 0x8000u) >> 16);
    }
  }
}

__device__ __forceinline__ int find_image(const int32_t* __restrict__ blk, int n, int b) {
  int lo = 0, hi = n;  // largest i with blk[i] <= b
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (blk[mid] <= b) lo = mid; else hi = mid;
  }
  return lo;
}

constexpr int kTwoPassThreads = 256;
constexpr int kPass2Px = 8;  // output pixels per thread of pass 2 (amortises the image lookup and descriptor load)

// Pass 1: first separable pass into the uint8 intermediate tmp[p1_h][p1_w][C].
template <int C>
__global__ void __launch_bounds__(kTwoPassThreads)
k_resample_pass1(const ImgDesc* __restrict__ descs, const int32_t* __restrict__ blk, int n,
                 const int32_t* __restrict__ pool, uint8_t* __restrict__ scratch) {
  const int img = find_image(blk, n, blockIdx.x);
  const ImgDesc d = descs[img];
  const int local = (blockIdx.x - blk[img]) * kTwoPassThreads + threadIdx.x;
  if (local >= d.p1_w * d.p1_h) return;
  const int y = local / d.p1_w, x = local - y * d.p1_w;
  int acc[C];
#pragma unroll
  for (int ch = 0; ch < C; ch++) acc[ch] = 1 << (kPrecisionBits - 1);
  int v[C];
  if (d.p1_axis == AXIS_H) {
    const int xmin = pool[d.th.bounds_off + 2 * x], cnt = pool[d.th.bounds_off + 2 * x + 1];
    const int32_t* __restrict__ k = pool + d.th.kk_off + (size_t)x * d.th.ksize;
    const uint8_t* row = d.src + (size_t)y * d.pitch;
    for (int t = 0; t < cnt; t++) {
      fetch_px<C>(row, d.cin, xmin + t, v);
      const int kv = k[t];
#pragma unroll
      for (int ch = 0; ch < C; ch++) acc[ch] += v[ch] * kv;
    }
  } else {
    const int ymin = pool[d.tv.bounds_off + 2 * y], cnt = pool[d.tv.bounds_off + 2 * y + 1];
    const int32_t* __restrict__ k = pool + d.tv.kk_off + (size_t)y * d.tv.ksize;
    for (int t = 0; t < cnt; t++) {
      fetch_px<C>(d.src + (size_t)(ymin + t) * d.pitch, d.cin, x, v);
      const int kv = k[t];
#pragma unroll
      for (int ch = 0; ch < C; ch++) acc[ch] += v[ch] * kv;
    }
  }
  uint8_t* o = scratch + d.tmp_off + ((size_t)y * d.p1_w + x) * C;
#pragma unroll
  for (int ch = 0; ch < C; ch++) o[ch] = (uint8_t)clip8(acc[ch]);
}

template <int ELEM>
__device__ __forceinline__ void store_elem(uint8_t* __restrict__ dst, size_t idx,
                                           const uint8_t* __restrict__ lut_entry) {
  if (ELEM == 1) dst[idx] = *lut_entry;
  if (ELEM == 2) reinterpret_cast<uint16_t*>(dst)[idx] = *reinterpret_cast<const uint16_t*>(lut_entry);
  if (ELEM == 4) reinterpret_cast<uint32_t*>(dst)[idx] = *reinterpret_cast<const uint32_t*>(lut_entry);
  if (ELEM == 8) reinterpret_cast<uint64_t*>(dst)[idx] = *reinterpret_cast<const uint64_t*>(lut_entry);
}

// ---- same-size images (no resample pass at all): value table + layout only, four pixels per thread ----
template <int ELEM> struct ElemOf;
template <> struct ElemOf<1> { using type = uint8_t; };
template <> struct ElemOf<2> { using type = uint16_t; };
template <> struct ElemOf<4> { using type = uint32_t; };
template <> struct ElemOf<8> { using type = unsigned long long; };

// RUN (multiple of 4) consecutive elements v[first..] to a destination aligned to min(16, 4 * ELEM) bytes
template <int ELEM, int RUN, int N>
__device__ __forceinline__ void store_run(uint8_t* __restrict__ p, const typename ElemOf<ELEM>::type (&v)[N], int first) {
#pragma unroll
  for (int k = 0; k < RUN; k += 4) {
    if (ELEM == 1) {
      *reinterpret_cast<uint32_t*>(p + k) = (uint32_t)v[first + k] | ((uint32_t)v[first + k + 1] << 8) |
                                            ((uint32_t)v[first + k + 2] << 16) | ((uint32_t)v[first + k + 3] << 24);
    } else if (ELEM == 2) {
      *reinterpret_cast<uint2*>(p + 2 * k) = make_uint2((uint32_t)v[first + k] | ((uint32_t)v[first + k + 1] << 16),
                                                        (uint32_t)v[first + k + 2] | ((uint32_t)v[first + k + 3] << 16));
    } else if (ELEM == 4) {
      *reinterpret_cast<uint4*>(p + 4 * k) = make_uint4((uint32_t)v[first + k], (uint32_t)v[first + k + 1],
                                                        (uint32_t)v[first + k + 2], (uint32_t)v[first + k + 3]);
    } else {
#pragma unroll
      for (int h = 0; h < 4; h += 2) {
        const unsigned long long a = (unsigned long long)v[first + k + h], b = (unsigned long long)v[first + k + h + 1];
        *reinterpret_cast<uint4*>(p + 8 * (k + h)) = make_uint4((uint32_t)a, (uint32_t)(a >> 32), (uint32_t)b, (uint32_t)(b >> 32));
      }
    }
  }
}

// Pass 2: second pass (or the only pass, or none) + value table + layout, one thread per output
// pixel.
template <int C, int ELEM>
__global__ void __launch_bounds__(kTwoPassThreads)
k_resample_pass2(const ImgDesc* __restrict__ descs, const int32_t* __restrict__ blk, int n,
                 const int32_t* __restrict__ pool, const uint8_t* __restrict__ scratch,
                 const uint8_t* __restrict__ lut, uint8_t* __restrict__ dst, int out_h, int out_w,
                 int layout) {
  // the value table (C x 256 entries) is read straight from global memory: it stays in L1, and
  // copying it to shared memory cost every 256-pixel block three times its own payload
  const int img = find_image(blk, n, blockIdx.x);
  const ImgDesc d = descs[img];
  // input of this pass: the intermediate of pass 1, or the (converted) source image
  const uint8_t* in;
  int in_pitch, in_cin;
  if (d.p1_axis != AXIS_NONE) {
    in = scratch + d.tmp_off; in_pitch = d.p1_w * C; in_cin = C;
  } else {
    in = d.src; in_pitch = d.pitch; in_cin = d.cin;
  }
  uint8_t* o = dst + (size_t)img * C * out_h * out_w * ELEM;
  if (d.p2_axis == AXIS_NONE && (out_w & 3) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
    // nothing to resample in this pass: four consecutive pixels per thread, vector stores
    using E = typename ElemOf<ELEM>::type;
    const E* __restrict__ tab = reinterpret_cast<const E*>(lut);
    const int g0 = (blockIdx.x - blk[img]) * (kTwoPassThreads * kPass2Px / 4) + threadIdx.x;
    const size_t plane = (size_t)out_h * out_w;
#pragma unroll
    for (int r = 0; r < kPass2Px / 4; r++) {
      const int local = 4 * (g0 + r * kTwoPassThreads);
      if (local >= out_h * out_w) return;
      const int y = local / out_w, x = local - y * out_w;
      const uint8_t* row = in + (size_t)y * in_pitch;
      E vals[4 * C];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        int u[C];
        fetch_px<C>(row, in_cin, x + j, u);
#pragma unroll
        for (int ch = 0; ch < C; ch++) {
          const E e = __ldg(tab + ch * 256 + u[ch]);
          if (layout == TCB_LAYOUT_NCHW) vals[ch * 4 + j] = e; else vals[j * C + ch] = e;
        }
      }
      if (layout == TCB_LAYOUT_NCHW) {
#pragma unroll
        for (int ch = 0; ch < C; ch++) store_run<ELEM, 4>(o + ((size_t)ch * plane + local) * ELEM, vals, ch * 4);
      } else {
        store_run<ELEM, 4 * C>(o + (size_t)local * C * ELEM, vals, 0);
      }
    }
    return;
  }
  const int local0 = (blockIdx.x - blk[img]) * (kTwoPassThreads * kPass2Px) + threadIdx.x;
  for (int px = 0; px < kPass2Px; px++) {  // consecutive threads take consecutive pixels in every round
  const int local = local0 + px * kTwoPassThreads;
  if (local >= out_h * out_w) return;
  const int y = local / out_w, x = local - y * out_w;
  int u[C];
  int v[C];
  if (d.p2_axis == AXIS_NONE) {
    fetch_px<C>(in + (size_t)y * in_pitch, in_cin, x, u);
  } else {
    int acc[C];
#pragma unroll
    for (int ch = 0; ch < C; ch++) acc[ch] = 1 << (kPrecisionBits - 1);
    if (d.p2_axis == AXIS_H) {
      const int xmin = pool[d.th.bounds_off + 2 * x], cnt = pool[d.th.bounds_off + 2 * x + 1];
      const int32_t* __restrict__ k = pool + d.th.kk_off + (size_t)x * d.th.ksize;
      const uint8_t* row = in + (size_t)y * in_pitch;
      for (int t = 0; t < cnt; t++) {
        fetch_px<C>(row, in_cin, xmin + t, v);
        const int kv = k[t];
#pragma unroll
        for (int ch = 0; ch < C; ch++) acc[ch] += v[ch] * kv;
      }
    } else {
      const int ymin = pool[d.tv.bounds_off + 2 * y], cnt = pool[d.tv.bounds_off + 2 * y + 1];
      const int32_t* __restrict__ k = pool + d.tv.kk_off + (size_t)y * d.tv.ksize;
      for (int t = 0; t < cnt; t++) {
        fetch_px<C>(in + (size_t)(ymin + t) * in_pitch, in_cin, x, v);
        const int kv = k[t];
#pragma unroll
        for (int ch = 0; ch < C; ch++) acc[ch] += v[ch] * kv;
      }
    }
#pragma unroll
    for (int ch = 0; ch < C; ch++) u[ch] = clip8(acc[ch]);
  }
#pragma unroll
  for (int ch = 0; ch < C; ch++) {
    const size_t idx = layout == TCB_LAYOUT_NCHW ? ((size_t)ch * out_h + y) * out_w + x
                                                 : ((size_t)y * out_w + x) * C + ch;
    store_elem<ELEM>(o, idx, lut + ((size_t)ch * 256 + u[ch]) * ELEM);
  }
  }
}

template <int C>
static void launch_pass2(const Plan& p, void* dst, cudaStream_t s) {
#define TCB_P2(E)                                                                               \
  k_resample_pass2<C, E><<<p.blocks2, kTwoPassThreads, 0, s>>>(                                \
      p.d_desc, p.d_blk2, p.n, p.d_pool, p.d_scratch, p.d_lut, (uint8_t*)dst, p.out_h, p.out_w, \
      p.layout)
  switch (p.elem) {
    case 1: TCB_P2(1); break;
    case 2: TCB_P2(2); break;
    case 4: TCB_P2(4); break;
    default: TCB_P2(8); break;
  }
#undef TCB_P2
}

// ---- host: plan ------------------------------------------------------------------------------------
static void decide_passes(ImgDesc& d, int out_w, int out_h) {
  // PIL/Image.py::Image.resize (Pillow 12.2.0) + Resample.c ImagingResampleInner
  d.p1_axis = AXIS_NONE; d.p2_axis = AXIS_NONE; d.p1_w = 0; d.p1_h = 0;
  if (d.W == out_w && d.H == out_h) return;  // same size: copy
  const bool need_h = out_w != d.W, need_v = out_h != d.H;
  const bool vfirst = (d.H > d.W * 100) && (out_h < d.H);
  if (need_h && need_v) {
    if (vfirst) { d.p1_axis = AXIS_V; d.p1_w = d.W; d.p1_h = out_h; d.p2_axis = AXIS_H; }
    else        { d.p1_axis = AXIS_H; d.p1_w = out_w; d.p1_h = d.H; d.p2_axis = AXIS_V; }
  } else if (need_h) {
    d.p2_axis = AXIS_H;
  } else {
    d.p2_axis = AXIS_V;
  }
}

static int upload(void** dptr, const void* src, size_t bytes) {
  *dptr = nullptr;
  if (bytes == 0) return TCB_OK;
  TCB_CUDA(cudaMalloc(dptr, bytes));
  TCB_CUDA(cudaMemcpy(*dptr, src, bytes, cudaMemcpyHostToDevice));
  return TCB_OK;
}

static void plan_free(Plan* p) {
  if (!p) return;
  DeviceGuard g(p->device);
  cudaFree(p->d_desc); cudaFree(p->d_pool); cudaFree(p->d_lut); cudaFree(p->d_scratch);
  cudaFree(p->d_blk1); cudaFree(p->d_blk2);
  fused_release(*p);
  delete p;
}

static int validate_image(const TcbImage& im, int i) {
  TCB_CHECK_ARG(im.width > 0 && im.height > 0, "image %d: width/height must be positive", i);
  TCB_CHECK_ARG(im.src_channels == 1 || im.src_channels == 3 || im.src_channels == 4,
                "image %d: src_channels must be 1, 3 or 4 (got %d)", i, im.src_channels);
  TCB_CHECK_ARG((int64_t)im.pitch >= (int64_t)im.width * im.src_channels,
                "image %d: pitch %d smaller than a row (%d bytes)", i, im.pitch,
                im.width * im.src_channels);
  TCB_CHECK_ARG((int64_t)im.width * im.src_channels < (1ll << 30) && (int64_t)im.height < (1 << 24),
                "image %d: dimensions too large", i);
  return TCB_OK;
}

}  // namespace tcb

using namespace tcb;

extern "C" {

int TcbResampleCoeffs(int in_size, int out_size, int32_t* bounds, int32_t* kk, int kk_capacity) {
  TCB_CHECK_ARG(in_size > 0 && out_size > 0 && bounds && kk, "TcbResampleCoeffs: bad argument");
  std::vector<int32_t> b, k;
  const int ksize = compute_axis_coeffs(in_size, out_size, b, k);
  if ((int64_t)k.size() > (int64_t)kk_capacity)
    return set_error(TCB_ERR_OUT_OF_RANGE, "TcbResampleCoeffs: kk needs %zu entries", k.size());
  memcpy(bounds, b.data(), b.size() * sizeof(int32_t));
  memcpy(kk, k.data(), k.size() * sizeof(int32_t));
  return ksize;
}

int TcbPreprocessPlanCreate(int device_id, const TcbImage* images, int n_images, int c, int out_h,
                            int out_w, int layout, const void* lut, int elem_size, int path,
                            void** plan) {
  TCB_CHECK_ARG(plan != nullptr, "TcbPreprocessPlanCreate: plan out-pointer is NULL");
  *plan = nullptr;
  TCB_CHECK_ARG(path >= 0 && path <= 3, "TcbPreprocessPlanCreate: path must be 0, 1, 2 or 3");
  TCB_CHECK_ARG(images != nullptr && n_images > 0, "TcbPreprocessPlanCreate: no images");
  TCB_CHECK_ARG(c == 1 || c == 3, "TcbPreprocessPlanCreate: c must be 1 or 3 (got %d)", c);
  TCB_CHECK_ARG(out_h > 0 && out_w > 0 && out_h <= 16384 && out_w <= 16384,
                "TcbPreprocessPlanCreate: bad output size %dx%d", out_w, out_h);
  TCB_CHECK_ARG(layout == TCB_LAYOUT_NHWC || layout == TCB_LAYOUT_NCHW,
                "TcbPreprocessPlanCreate: layout must be 1 (NHWC) or 2 (NCHW)");
  TCB_CHECK_ARG(elem_size == 1 || elem_size == 2 || elem_size == 4 || elem_size == 8,
                "TcbPreprocessPlanCreate: elem_size must be 1, 2, 4 or 8");
  TCB_CHECK_ARG(lut != nullptr, "TcbPreprocessPlanCreate: lut is NULL");
  for (int i = 0; i < n_images; i++) {
    int rc = validate_image(images[i], i);
    if (rc) return rc;
  }
  TCB_DEVICE_GUARD(device_id);
  Plan* p = new (std::nothrow) Plan();
  if (!p) return set_error(TCB_ERR_NOMEM, "out of host memory");
  p->device = device_id; p->n = n_images; p->c = c; p->out_h = out_h; p->out_w = out_w;
  p->layout = layout; p->elem = elem_size;
  p->no_mma = path == 3;

  // coefficient tables, one per distinct (in, out) pair
  std::vector<int32_t> pool;
  std::map<std::pair<int, int>, AxisTable> tables;
  auto get_table = [&](int in_size, int out_size) -> AxisTable {
    auto key = std::make_pair(in_size, out_size);
    auto it = tables.find(key);
    if (it != tables.end()) return it->second;
    std::vector<int32_t> b, k;
    AxisTable t;
    t.ksize = compute_axis_coeffs(in_size, out_size, b, k);
    t.in_size = in_size; t.out_size = out_size;
    t.bounds_off = (int32_t)pool.size();
    pool.insert(pool.end(), b.begin(), b.end());
    while (pool.size() % 4) pool.push_back(0);
    t.kk_off = (int32_t)pool.size();
    pool.insert(pool.end(), k.begin(), k.end());
    while (pool.size() % 4) pool.push_back(0);
    tables[key] = t;
    return t;
  };

  p->h_desc.resize(n_images);
  std::vector<int32_t> blk1(n_images + 1, 0), blk2(n_images + 1, 0);
  uint64_t scratch = 0;
  size_t fused_smem = 0;
  for (int i = 0; i < n_images; i++) {
    ImgDesc& d = p->h_desc[i];
    memset(&d, 0, sizeof(d));
    d.src = reinterpret_cast<const uint8_t*>(images[i].src_dev);
    d.W = images[i].width; d.H = images[i].height; d.pitch = images[i].pitch;
    d.cin = images[i].src_channels;
    decide_passes(d, out_w, out_h);
    if (d.p1_axis == AXIS_H || d.p2_axis == AXIS_H) d.th = get_table(d.W, out_w);
    if (d.p1_axis == AXIS_V || d.p2_axis == AXIS_V) d.tv = get_table(d.H, out_h);
    int band = 0; size_t smem = 0;
    const bool can_fuse = path != 1 && fused_supported(*p, d, &band, &smem);
    if (path >= 2 && !can_fuse) {
      const int fw = d.W, fh = d.H, fc = d.cin, fp = d.pitch;
      delete p;
      return set_error(TCB_ERR_UNSUPPORTED,
                       "TcbPreprocessPlanCreate: image %d (%dx%d, %d ch, pitch %d) does not fit the "
                       "fused kernel", i, fw, fh, fc, fp);
    }
    if (can_fuse) {
      d.fused = 1; d.band_rows = band;
      fused_smem = std::max(fused_smem, smem);
      p->n_fused_images++;
      blk1[i + 1] = blk1[i]; blk2[i + 1] = blk2[i];
      continue;
    }
    if (d.p1_axis != AXIS_NONE) {
      d.tmp_off = scratch;
      scratch += round_up((uint64_t)d.p1_w * d.p1_h * c, 256);
      blk1[i + 1] = blk1[i] + (int32_t)(((int64_t)d.p1_w * d.p1_h + kTwoPassThreads - 1) / kTwoPassThreads);
    } else {
      blk1[i + 1] = blk1[i];
    }
    blk2[i + 1] = blk2[i] + (out_h * out_w + kTwoPassThreads * kPass2Px - 1) / (kTwoPassThreads * kPass2Px);
  }
  p->blocks1 = blk1[n_images]; p->blocks2 = blk2[n_images];
  p->scratch_bytes = scratch; p->pool_words = pool.size();
  p->n_fused_ctas = p->n_fused_images; p->fused_smem = fused_smem;
  if (pool.empty()) pool.push_back(0);

  int rc = upload((void**)&p->d_desc, p->h_desc.data(), sizeof(ImgDesc) * n_images);
  if (!rc) rc = upload((void**)&p->d_pool, pool.data(), pool.size() * sizeof(int32_t));
  if (!rc) rc = upload((void**)&p->d_lut, lut, (size_t)c * 256 * elem_size);
  if (!rc) rc = upload((void**)&p->d_blk1, blk1.data(), blk1.size() * sizeof(int32_t));
  if (!rc) rc = upload((void**)&p->d_blk2, blk2.data(), blk2.size() * sizeof(int32_t));
  if (!rc && p->n_fused_images > 0) rc = fused_prepare(*p);
  if (!rc && scratch > 0) {
    cudaError_t e = cudaMalloc((void**)&p->d_scratch, scratch);
    if (e != cudaSuccess)
      rc = set_error(TCB_ERR_NOMEM, "cudaMalloc(scratch %llu) failed: %s", (unsigned long long)scratch,
                     cudaGetErrorString(e));
  }
  if (rc) { plan_free(p); return rc; }
  *plan = p;
  return TCB_OK;
}

int TcbPreprocessPlanInfo(void* plan, uint64_t* scratch_bytes, uint64_t* out_bytes_per_image,
                          int* n_fused_images) {
  TCB_CHECK_ARG(plan != nullptr, "TcbPreprocessPlanInfo: plan is NULL");
  Plan* p = static_cast<Plan*>(plan);
  if (scratch_bytes) *scratch_bytes = p->scratch_bytes;
  if (out_bytes_per_image) *out_bytes_per_image = (uint64_t)p->c * p->out_h * p->out_w * p->elem;
  if (n_fused_images) *n_fused_images = p->n_fused_images;
  return TCB_OK;
}

int TcbPreprocessPlanRun(void* plan, const TcbImage* images_or_null, void* dst_dev,
                         uint64_t dst_capacity, void* stream) {
  TCB_CHECK_ARG(plan != nullptr, "TcbPreprocessPlanRun: plan is NULL");
  Plan* p = static_cast<Plan*>(plan);
  TCB_CHECK_ARG(dst_dev != nullptr, "TcbPreprocessPlanRun: dst is NULL");
  const uint64_t need = (uint64_t)p->n * p->c * p->out_h * p->out_w * p->elem;
  if (need > dst_capacity)
    return set_error(TCB_ERR_OUT_OF_RANGE, "TcbPreprocessPlanRun: output needs %llu bytes, capacity %llu",
                     (unsigned long long)need, (unsigned long long)dst_capacity);
  TCB_CHECK_ARG(reinterpret_cast<uintptr_t>(dst_dev) % p->elem == 0,
                "TcbPreprocessPlanRun: dst not aligned to the element size");
  TCB_DEVICE_GUARD(p->device);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (images_or_null) {
    // validate everything first, then commit: a failed run must leave the plan's host and device
    // descriptors as they were
    bool changed = false;
    for (int i = 0; i < p->n; i++) {
      const TcbImage& im = images_or_null[i];
      const ImgDesc& d = p->h_desc[i];
      TCB_CHECK_ARG(im.width == d.W && im.height == d.H && im.src_channels == d.cin,
                    "TcbPreprocessPlanRun: image %d geometry differs from the plan", i);
      TCB_CHECK_ARG((int64_t)im.pitch >= (int64_t)im.width * im.src_channels,
                    "TcbPreprocessPlanRun: image %d pitch too small", i);
      const uint8_t* sp = reinterpret_cast<const uint8_t*>(im.src_dev);
      if (sp != d.src || im.pitch != d.pitch) {
        // the fused kernel's eligibility depends on pitch/alignment; re-check
        if (d.fused && !fused_accepts(*p, i, sp, im.pitch))
          return set_error(TCB_ERR_UNSUPPORTED,
                           "TcbPreprocessPlanRun: image %d pointer/pitch does not fit the plan's fused launch; "
                           "create a new plan", i);
        changed = true;
      }
    }
    if (changed) {
      for (int i = 0; i < p->n; i++) {
        p->h_desc[i].src = reinterpret_cast<const uint8_t*>(images_or_null[i].src_dev);
        p->h_desc[i].pitch = images_or_null[i].pitch;
      }
      p->desc_dirty = true;
    }
  }
  if (p->desc_dirty) {  // stays set until both uploads have been enqueued
    // pageable source: cudaMemcpyAsync returns once the host buffer has been staged
    TCB_CUDA(cudaMemcpyAsync(p->d_desc, p->h_desc.data(), sizeof(ImgDesc) * p->n, cudaMemcpyHostToDevice, s));
    int rc = fused_update_sources(*p, s);
    if (rc) return rc;
    p->desc_dirty = false;
  }
  for (int i = 0; i < p->n; i++)
    TCB_CHECK_ARG(p->h_desc[i].src != nullptr, "TcbPreprocessPlanRun: image %d has no source pointer", i);

  if (p->blocks1 > 0) {
    if (p->c == 3)
      k_resample_pass1<3><<<p->blocks1, kTwoPassThreads, 0, s>>>(p->d_desc, p->d_blk1, p->n, p->d_pool, p->d_scratch);
    else
      k_resample_pass1<1><<<p->blocks1, kTwoPassThreads, 0, s>>>(p->d_desc, p->d_blk1, p->n, p->d_pool, p->d_scratch);
    count_launch();
  }
  if (p->blocks2 > 0) {
    if (p->c == 3) launch_pass2<3>(*p, dst_dev, s); else launch_pass2<1>(*p, dst_dev, s);
    count_launch();
  }
  if (p->n_fused_ctas > 0) {
    int rc = fused_launch(*p, dst_dev, s);
    if (rc) return rc;
  }
  TCB_CUDA(cudaGetLastError());
  return TCB_OK;
}

int TcbPreprocessPlanDestroy(void* plan) {
  TCB_CHECK_ARG(plan != nullptr, "TcbPreprocessPlanDestroy: plan is NULL");
  plan_free(static_cast<Plan*>(plan));
  return TCB_OK;
}

}  // extern "C"
