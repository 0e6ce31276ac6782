// Shared declarations for the preprocess kernels (two-pass and fused).
#pragma once
#include <array>
#include <map>
#include <vector>

#include "common.cuh"

namespace tcb {

constexpr int kPrecisionBits = 22;  // Pillow Resample.c: PRECISION_BITS = 32 - 8 - 2

enum Axis : int32_t { AXIS_NONE = 0, AXIS_H = 1, AXIS_V = 2 };

// One resampling table (one axis, one (in, out) pair) inside the plan's int32 pool.
struct AxisTable {
  int32_t bounds_off;  // pool index of bounds[out*2] = (first tap, tap count)
  int32_t kk_off;      // pool index of kk[out*ksize]
  int32_t ksize;
  int32_t in_size, out_size;
};

// Per-image descriptor on the device.
struct ImgDesc {
  const uint8_t* src;
  int32_t W, H, pitch, cin;
  int32_t p1_axis, p2_axis;  // pass order decided on the host (Pillow's rule)
  int32_t p1_w, p1_h;        // geometry of the uint8 intermediate written by pass 1
  AxisTable th, tv;          // horizontal / vertical tables (valid when that axis is resampled)
  unsigned long long tmp_off;  // byte offset of this image's intermediate in the scratch buffer
  int32_t fused;               // 1: handled by the fused kernel, skipped by the two-pass kernels
  int32_t band_rows;           // fused kernel: output rows per CTA
};

// launch geometry of one image in the fused kernel (preprocess_fused.cu)
struct FusedGeom {
  int nw_class = 0, nwv = 0, wq = 0, stage_bytes = 0, threads = 0;
  int mma_nk = 0, mma_wp = 0;  // IMMA variant (nw_class >= 100): k-steps per tile, bytes per planar line
  size_t smem = 0;
  bool ok = false;
};

struct Plan {
  int device = 0;
  int n = 0, c = 0, out_h = 0, out_w = 0, layout = 0, elem = 0;
  std::vector<ImgDesc> h_desc;
  ImgDesc* d_desc = nullptr;
  int32_t* d_pool = nullptr;
  size_t pool_words = 0;
  uint8_t* d_lut = nullptr;      // c * 256 * elem bytes
  uint8_t* d_scratch = nullptr;  // two-pass intermediates
  uint64_t scratch_bytes = 0;
  // work lists (block -> image) for the two-pass kernels
  int32_t* d_blk1 = nullptr;  // prefix of pass-1 block counts, n+1 entries
  int32_t* d_blk2 = nullptr;  // prefix of pass-2 block counts
  int blocks1 = 0, blocks2 = 0;
  // fused kernel work list
  int32_t* d_fused_work = nullptr;  // pairs (image, first output row)
  int n_fused_ctas = 0, n_fused_images = 0;
  size_t fused_smem = 0;
  bool no_mma = false;      // path 3: fused dp4a kernel only
  bool desc_dirty = false;  // host descriptors changed, device copies not yet re-uploaded
  // fused-kernel geometry per distinct (W, H, channels, pitch): a batch of thousands of images has a
  // handful of distinct sizes, and the coefficient set-up behind the answer is not free
  std::map<std::array<int, 4>, FusedGeom> geom_cache;
  void* fused = nullptr;  // FusedPlan (preprocess_fused.cu), owned by the plan: no process-global state
};

// Pillow precompute_coeffs + normalize_coeffs_8bpc (host, double precision).
int compute_axis_coeffs(int in_size, int out_size, std::vector<int32_t>& bounds,
                        std::vector<int32_t>& kk);

// fused kernel (preprocess_fused.cu)
bool fused_supported(Plan& p, const ImgDesc& d, int* band_rows, size_t* smem_bytes);
int fused_prepare(Plan& p);                                // after h_desc is final
int fused_update_sources(const Plan& p, cudaStream_t s);   // after src/pitch changed
// can image i of the plan take (src, pitch) without a new plan? (alignment class, staging size)
bool fused_accepts(Plan& p, int i, const uint8_t* src, int pitch);
int fused_launch(const Plan& p, void* dst, cudaStream_t stream);
void fused_release(Plan& p);

}  // namespace tcb
