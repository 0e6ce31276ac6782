/* Oracle (TEST INFRASTRUCTURE ONLY - never linked into or called by the product library).
 *
 * Plain-C restatement of the two CPU algorithms on the hot path:
 *
 *  1. Pillow's 8-bit BILINEAR resampler, reached by the reference at
 *     /root/reference/src/python/examples/image_client.py:166
 *     (`sample_img.resize((w, h), Image.BILINEAR)`). Pillow is third-party and absent from
 *     /root/reference; this restates the published algorithm of Pillow 12.2.0
 *     src/libImaging/Resample.c (precompute_coeffs, normalize_coeffs_8bpc, the 8bpc
 *     horizontal/vertical passes), Convert.c (rgb2l) and the pass-order rule of
 *     PIL/Image.py::Image.resize. Written from the behavioural spec in SURVEY.md Appendix A.
 *
 *  2. The BYTES wire format of
 *     /root/reference/src/python/library/tritonclient/utils/__init__.py:193-276
 *     (4-byte LE length + payload per element, logical C order).
 *
 * Build: `make -C oracle` -> oracle/liboracle.so (gcc -O2 -ffp-contract=off: the coefficient
 * set-up must round every double operation separately, as the reference's C does on x86-64).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_PRECISION_BITS 22

static double orc_tri(double x) {
  if (x < 0.0) x = -x;
  if (x < 1.0) return 1.0 - x;
  return 0.0;
}

int orc_ksize(int in_size, int out_size) {
  double scale = (double)in_size / (double)out_size;
  double fs = scale < 1.0 ? 1.0 : scale;
  return (int)ceil(1.0 * fs) * 2 + 1;
}

/* bounds: int32[out*2] (xmin, count); kk: int32[out*ksize]. Returns ksize. */
int orc_precompute_coeffs(int in_size, int out_size, int32_t* bounds, int32_t* kk) {
  double scale = (double)in_size / (double)out_size;
  double filterscale = scale < 1.0 ? 1.0 : scale;
  double support = 1.0 * filterscale;
  int ksize = (int)ceil(support) * 2 + 1;
  double ss = 1.0 / filterscale;
  double* w = (double*)malloc(sizeof(double) * (size_t)ksize);
  if (!w) return -1;
  for (int xx = 0; xx < out_size; xx++) {
    double center = 0.0 + (xx + 0.5) * scale;
    double ww = 0.0;
    int xmin = (int)(center - support + 0.5);
    if (xmin < 0) xmin = 0;
    int xmax = (int)(center + support + 0.5);
    if (xmax > in_size) xmax = in_size;
    xmax -= xmin;
    int32_t* k = kk + (size_t)xx * ksize;
    for (int x = 0; x < xmax; x++) {
      double v = orc_tri((x + xmin - center + 0.5) * ss);
      w[x] = v;
      ww += v;
    }
    for (int x = 0; x < xmax; x++) {
      if (ww != 0.0) w[x] /= ww;
      if (w[x] < 0)
        k[x] = (int)(-0.5 + w[x] * (1 << ORC_PRECISION_BITS));
      else
        k[x] = (int)(0.5 + w[x] * (1 << ORC_PRECISION_BITS));
    }
    for (int x = xmax; x < ksize; x++) k[x] = 0;
    bounds[xx * 2 + 0] = xmin;
    bounds[xx * 2 + 1] = xmax;
  }
  free(w);
  return ksize;
}

static inline uint8_t orc_clip8(int32_t acc) {
  int32_t v = acc >> ORC_PRECISION_BITS; /* arithmetic shift */
  if (v < 0) return 0;
  if (v > 255) return 255;
  return (uint8_t)v;
}

/* src[h][w][c] (row pitch src_pitch bytes) -> dst[h][out_w][c], dense. */
static int orc_pass_h(const uint8_t* src, int h, int w, int c, size_t src_pitch, int out_w,
                      uint8_t* dst) {
  int ksize = orc_ksize(w, out_w);
  int32_t* bounds = (int32_t*)malloc(sizeof(int32_t) * 2 * (size_t)out_w);
  int32_t* kk = (int32_t*)malloc(sizeof(int32_t) * (size_t)out_w * ksize);
  if (!bounds || !kk) return -1;
  orc_precompute_coeffs(w, out_w, bounds, kk);
  for (int y = 0; y < h; y++) {
    const uint8_t* row = src + (size_t)y * src_pitch;
    uint8_t* orow = dst + (size_t)y * out_w * c;
    for (int xx = 0; xx < out_w; xx++) {
      int xmin = bounds[xx * 2], n = bounds[xx * 2 + 1];
      const int32_t* k = kk + (size_t)xx * ksize;
      for (int ch = 0; ch < c; ch++) {
        int32_t acc = 1 << (ORC_PRECISION_BITS - 1);
        for (int x = 0; x < n; x++) acc += (int32_t)row[(size_t)(xmin + x) * c + ch] * k[x];
        orow[xx * c + ch] = orc_clip8(acc);
      }
    }
  }
  free(bounds);
  free(kk);
  return 0;
}

/* src[h][rowbytes] (pitch) -> dst[out_h][rowbytes], dense; rowbytes = w*c. */
static int orc_pass_v(const uint8_t* src, int h, size_t rowbytes, size_t src_pitch, int out_h,
                      uint8_t* dst) {
  int ksize = orc_ksize(h, out_h);
  int32_t* bounds = (int32_t*)malloc(sizeof(int32_t) * 2 * (size_t)out_h);
  int32_t* kk = (int32_t*)malloc(sizeof(int32_t) * (size_t)out_h * ksize);
  if (!bounds || !kk) return -1;
  orc_precompute_coeffs(h, out_h, bounds, kk);
  for (int yy = 0; yy < out_h; yy++) {
    int ymin = bounds[yy * 2], n = bounds[yy * 2 + 1];
    const int32_t* k = kk + (size_t)yy * ksize;
    uint8_t* orow = dst + (size_t)yy * rowbytes;
    for (size_t i = 0; i < rowbytes; i++) {
      int32_t acc = 1 << (ORC_PRECISION_BITS - 1);
      for (int y = 0; y < n; y++) acc += (int32_t)src[(size_t)(ymin + y) * src_pitch + i] * k[y];
      orow[i] = orc_clip8(acc);
    }
  }
  free(bounds);
  free(kk);
  return 0;
}

/* Image.resize((out_w, out_h), BILINEAR) for src[h][w][c] u8 with row pitch; dst dense
 * [out_h][out_w][c]. Horizontal first, u8 intermediate, then vertical; tall-narrow images
 * (h > 100*w and out_h < h) go vertical first (PIL/Image.py::Image.resize). */
int orc_resize_bilinear_u8(const uint8_t* src, int h, int w, int c, size_t src_pitch, int out_w,
                           int out_h, uint8_t* dst) {
  size_t row = (size_t)w * c;
  if (w == out_w && h == out_h) {
    for (int y = 0; y < h; y++) memcpy(dst + (size_t)y * row, src + (size_t)y * src_pitch, row);
    return 0;
  }
  int vfirst = (h > w * 100) && (out_h < h);
  int rc = 0;
  if (vfirst) {
    uint8_t* tmp = (uint8_t*)malloc((size_t)out_h * row);
    if (!tmp) return -1;
    rc = orc_pass_v(src, h, row, src_pitch, out_h, tmp);
    if (!rc) {
      if (out_w != w)
        rc = orc_pass_h(tmp, out_h, w, c, row, out_w, dst);
      else
        memcpy(dst, tmp, (size_t)out_h * row);
    }
    free(tmp);
    return rc;
  }
  if (out_w != w && out_h != h) {
    uint8_t* tmp = (uint8_t*)malloc((size_t)h * out_w * c);
    if (!tmp) return -1;
    rc = orc_pass_h(src, h, w, c, src_pitch, out_w, tmp);
    if (!rc) rc = orc_pass_v(tmp, h, (size_t)out_w * c, (size_t)out_w * c, out_h, dst);
    free(tmp);
    return rc;
  }
  if (out_w != w) return orc_pass_h(src, h, w, c, src_pitch, out_w, dst);
  return orc_pass_v(src, h, row, src_pitch, out_h, dst);
}

/* Convert.c rgb2l on packed pixels with cin channels (3 or 4); dst dense [h][w]. */
void orc_rgb_to_l(const uint8_t* src, int h, int w, int cin, size_t src_pitch, uint8_t* dst) {
  for (int y = 0; y < h; y++) {
    const uint8_t* r = src + (size_t)y * src_pitch;
    for (int x = 0; x < w; x++) {
      uint32_t v = (uint32_t)r[x * cin] * 19595u + (uint32_t)r[x * cin + 1] * 38470u +
                   (uint32_t)r[x * cin + 2] * 7471u + 0x8000u;
      dst[(size_t)y * w + x] = (uint8_t)(v >> 16);
    }
  }
}

/* ------------------------------------------------------------------------------------------ */
/* BYTES wire format, utils/__init__.py:193-246: offsets = exclusive scan of (len + 4). */
int64_t orc_bytes_serialize(const uint8_t* blob, const int64_t* lens, int64_t n, uint8_t* out) {
  int64_t src = 0, dst = 0;
  for (int64_t i = 0; i < n; i++) {
    uint32_t l = (uint32_t)lens[i];
    out[dst + 0] = (uint8_t)(l & 0xff);
    out[dst + 1] = (uint8_t)((l >> 8) & 0xff);
    out[dst + 2] = (uint8_t)((l >> 16) & 0xff);
    out[dst + 3] = (uint8_t)((l >> 24) & 0xff);
    memcpy(out + dst + 4, blob + src, (size_t)l);
    dst += 4 + (int64_t)l;
    src += (int64_t)l;
  }
  return dst;
}

/* utils/__init__.py:249-276 (max_elems < 0: to end of buffer) and
 * cuda_shared_memory/__init__.py:305-323 (max_elems >= 0: stop after that many).
 * offsets: int64[capacity+1]; blob: at least nbytes. Returns element count, or
 * -1 truncated header, -2 truncated payload, -3 capacity exceeded. */
int64_t orc_bytes_deserialize(const uint8_t* stream, int64_t nbytes, int64_t max_elems,
                              int64_t capacity, int64_t* offsets, uint8_t* blob) {
  int64_t pos = 0, n = 0, out = 0;
  offsets[0] = 0;
  while (pos < nbytes) {
    if (max_elems >= 0 && n >= max_elems) break;
    if (pos + 4 > nbytes) return -1;
    uint32_t l = (uint32_t)stream[pos] | ((uint32_t)stream[pos + 1] << 8) |
                 ((uint32_t)stream[pos + 2] << 16) | ((uint32_t)stream[pos + 3] << 24);
    pos += 4;
    if (pos + (int64_t)l > nbytes) return -2;
    if (n >= capacity) return -3;
    memcpy(blob + out, stream + pos, (size_t)l);
    out += (int64_t)l;
    pos += (int64_t)l;
    n++;
    offsets[n] = out;
  }
  return n;
}
