// Fused single-kernel preprocess (sm_100a): PIL-exact BILINEAR resample (H pass -> uint8 -> V pass
// -> uint8) + value table + NHWC/NCHW store, one launch per coefficient-width class.
//
// Replaces image_client.py:154-193 (/root/reference/src/python/examples/image_client.py) for the
// common case "both axes resampled, horizontal pass first"; every other geometry takes the
// two-pass kernels in preprocess.cu. Arithmetic is the one restated in oracle/ (Pillow
// Resample.c): acc = 2^21 + sum(px * k), k 22-bit fixed point, clip8(acc >> 22).
//
// Per CTA = (image, band of output rows). Source rows stream through the CTA four at a time:
//   1. TMA bulk copy (cp.async.bulk + mbarrier complete_tx) of 4 interleaved source rows into a
//      shared-memory stage; issued by one thread as soon as the previous group has been
//      de-interleaved, so it lands while the CTA runs the horizontal and vertical passes.
//   2. de-interleave to planar, 4-row-interleaved words P4[ch][word][row] (PRMT byte shuffles).
//   3. horizontal pass with dp4a: thread = output column xx; its coefficients live in registers
//      as three byte planes (k = k0 + k1<<8 + k2<<16) aligned to the 4-pixel word grid; one
//      128-bit LDS brings the same 4 pixels of all 4 rows; 12 dp4a per load.
//      The four uint8 results (rows 4g..4g+3) are packed into one word of a small ring tmp4[g].
//   4. vertical pass, again dp4a (the 4 rows of a word are 4 vertical taps), as soon as the last
//      row group an output row needs has arrived; then the 256-entry table and a vector store
//      straight to the destination (the registered region).
// HBM traffic is the source once (+ the band halo) and the output once.
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdlib>
#include <map>
#include <type_traits>

#include "preprocess.cuh"

namespace tcb {

constexpr int kFusedMaxThreads = 512;
constexpr size_t kFusedMaxSmem = 200 * 1024;
static const int kNwClasses[] = {2, 3, 4, 5, 6, 8, 10, 12, 16};

// ---- PTX helpers ------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// 1-D TMA: global -> shared, completion signalled on the mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void tma_load_1d(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void st_global_v4(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void st_global_v2(void* p, uint32_t a, uint32_t b) {
  asm volatile("st.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(a), "r"(b) : "memory");
}

// fused-kernel view of one image (lives in the plan's descriptor array next to ImgDesc)
struct FusedDesc {
  const uint8_t* src;
  int32_t W, H, pitch, cin;
  int32_t nw;        // horizontal coefficient words per output column (class)
  int32_t nwv;       // vertical coefficient words per output row
  int32_t hc_off;    // pool: packed H coefficients [out_w][nw][3]
  int32_t hw0_off;   // pool: first source word of every output column [out_w]
  int32_t hperm_off; // pool: output column of every thread [threads]
  int32_t vc_off;    // pool: packed V coefficients [out_h][nwv][4] (k0, k1, k2, 0), 16-byte aligned
  int32_t vg_off;    // pool: (first group, last group) of every output row [out_h][2]
  int32_t vb_off;    // pool: Pillow bounds of the vertical axis [out_h][2]
  int32_t band_rows;
  int32_t wq;        // words per planar row incl. padding
  int32_t stage_bytes;
  int32_t aligned4;  // rows start 4-byte aligned: the de-interleave reads them without funnel shifts
  // IMMA variant (k_preprocess_fused_mma), valid when mma_nk > 0
  int32_t mma_nk;     // 32-pixel k-steps per tile of 8 output columns
  int32_t mma_wp;     // bytes per planar line in shared memory (multiple of 16, == 16 mod 128)
  int32_t mma_b_off;  // pool: B fragments [n-tile][k-step][plane][lane][2]
  int32_t mma_s_off;  // pool: first source pixel of every n-tile's window [n-tiles] (multiple of 16)
  int32_t vq_off;     // pool: output rows complete after 8-row step s (last group <= 2s + 1) [ceil(H / 8)]
};

__device__ __forceinline__ uint32_t fetch_conv(const uint8_t* row, int cin, int C, int x, int ch) {
  if (C == 3) {
    if (cin == 1) return row[x];
    return row[x * cin + ch];
  }
  if (cin == 1) return row[x];
  const uint8_t* q = row + x * cin;
  return (q[0] * 19595u + q[1] * 38470u + q[2] * 7471u + 0x8000u) >> 16;
}

// Physical granule index of logical granule w in a P4 plane. Identity: the horizontal pass reads
// granules ~2 apart from consecutive lanes (1080p -> 224), which is a 2-way bank conflict for any
// 8-lane phase; padding and XOR swizzles were simulated and measured and do not beat it.
__host__ __device__ __forceinline__ int p4_phys(int w) { return w; }

// NW: coefficient words per output column; C: output channels; BIG: more than 256 output columns
// (one CTA per SM). Per row group: wait for its TMA, de-interleave it (raw -> P4), barrier, start
// the next group's TMA, horizontal pass (P4 -> tmp4 ring) and the vertical pass of the output
// rows whose last group landed earlier, barrier. One raw stage and one P4 buffer keep the CTA at
// ~60 KB for 1080p so that three CTAs share an SM: the phases of different CTAs overlap (measured
// 0.522 ms on cfg3 against 0.552 ms for the double-buffered two-CTA version).
template <int NW, int C, bool BIG>
__global__ void __launch_bounds__(BIG ? 512 : 256, BIG ? 1 : 3)
k_preprocess_fused(const FusedDesc* __restrict__ descs, const int4* __restrict__ items,
                   const int32_t* __restrict__ cta_first, const int32_t* __restrict__ pool,
                   const uint8_t* __restrict__ lut, uint8_t* __restrict__ dst, int out_h, int out_w,
                   int layout, int elem, int flags) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, nthreads = blockDim.x;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  uint32_t tma_phase = 0;  // parity of the next completion of bars[0]; runs on across work items
  // Persistent CTA: the host cut the (image, output row) space of this launch into gridDim.x
  // contiguous shares of equal estimated cost; a share is one to a few items (image, row span).
  for (int item = cta_first[blockIdx.x]; item < cta_first[blockIdx.x + 1]; item++) {
  const int4 wi = items[item];
  const int img = wi.x, yy0 = wi.y;
  const FusedDesc d = descs[img];
  const int yy1 = wi.z;
  const int L = out_w * C;   // words per tmp4 slot = elements per output row
  const int ng = d.nwv + 1;  // ring slots
  const int wqp = d.wq;      // physical granules per P4 plane (padding included)

  // shared memory: barriers | raw stage | P4 | tmp4 ring
  uint8_t* raw = smem + 64;
  uint4* P4 = reinterpret_cast<uint4*>(raw + (size_t)d.stage_bytes);  // [C][wqp]
  uint32_t* tmp4 = reinterpret_cast<uint32_t*>(P4 + (size_t)C * wqp); // [ng][L]

  const int32_t* vb = pool + d.vb_off;
  const int ys = vb[2 * yy0], ye = vb[2 * (yy1 - 1)] + vb[2 * (yy1 - 1) + 1];  // source rows [ys, ye)
  const int g_first = ys >> 2, g_last = (ye - 1) >> 2;
  const int n_groups = g_last - g_first + 1;

  auto issue = [&](int i) {  // thread 0: TMA for the i-th row group of this band into the raw stage
    const int g = g_first + i, s = 0;
    const int r1 = min(4 * g + 4, d.H);
    const uintptr_t start = reinterpret_cast<uintptr_t>(d.src) + (size_t)(4 * g) * d.pitch;
    const uintptr_t end = reinterpret_cast<uintptr_t>(d.src) + (size_t)(r1 - 1) * d.pitch + (size_t)d.W * d.cin;
    const uintptr_t a0 = start & ~(uintptr_t)15, a1 = (end + 15) & ~(uintptr_t)15;
    const uint32_t bytes = (uint32_t)(a1 - a0);
    mbar_expect_tx(&bars[s], bytes);
    tma_load_1d(raw + (size_t)s * d.stage_bytes, reinterpret_cast<const void*>(a0), bytes, &bars[s]);
  };

  // Items of one CTA may belong to images of different sizes, i.e. different shared-memory
  // layouts: nobody may still be reading the previous item's ring when this item's TMA lands.
  if (item != cta_first[blockIdx.x]) __syncthreads();
  if (tid == 0) issue(0);

  // ---- per-thread constants ------------------------------------------------------------------------
  const bool nchw = layout == TCB_LAYOUT_NCHW;
  const int xx = pool[d.hperm_off + tid];  // output column of this thread (>= out_w: none), see lane_columns()
  uint32_t kc[NW][3];
  int pw0 = 0;  // first P4 granule of this column (p4_phys is the identity: the NW granules are consecutive)
#pragma unroll
  for (int j = 0; j < NW; j++) { kc[j][0] = 0; kc[j][1] = 0; kc[j][2] = 0; }
  if (xx < out_w) {
    const int w0 = pool[d.hw0_off + xx];
    const int32_t* hc = pool + d.hc_off + (size_t)xx * NW * 3;
#pragma unroll
    for (int j = 0; j < NW; j++) {
      kc[j][0] = hc[j * 3 + 0]; kc[j][1] = hc[j * 3 + 1]; kc[j][2] = hc[j * 3 + 2];
    }
    pw0 = w0;
  }
  // first tmp4 word of this thread's column (+ ch * tmp_chs), or -1: no column. The test in the main
  // loop reads this register, which the horizontal pass needs anyway, instead of keeping xx alive.
  const int tmp_idx0 = xx < out_w ? (nchw ? xx : xx * C) : -1;
  const int tmp_chs = nchw ? out_w : 1;
  const int nwords = (d.W + 3) >> 2;
  const bool fast3 = d.cin == 3 && C == 3;  // word-wise de-interleave (rows of any alignment)
  const bool fast1 = d.cin == 1 && C == 1;
  const bool rows_al4 = d.aligned4 != 0;
  // vertical pass: units of 4 consecutive output elements; unit -> (line, x4) fixed per thread
  const bool vec4 = (flags & 1) && (nchw ? (out_w % 4 == 0) : (L % 4 == 0));  // bit 0: dst is 16-byte aligned
  const int upl = nchw ? out_w / 4 : L / 4;  // units per line (vec4 only)
  const int n_units = vec4 ? (nchw ? C * upl : upl) : 0;
  const int32_t* vg = pool + d.vg_off;
  uint8_t* dimg = dst + (size_t)img * C * out_h * out_w * elem;
  int yy_next = yy0;
  // (first group, last group) of the next output row to finish, kept in registers so that the
  // per-phase test does not wait for a global load
  int vg_first = vg[2 * yy0], vg_last = vg[2 * yy0 + 1];
  const int my_ln = (nchw && upl > 0) ? tid / upl : 0;  // this thread's first unit, hoisted out of the row loop
  const int my_x4 = (nchw && upl > 0) ? tid - my_ln * upl : tid;

  // de-interleave of the i-th group (raw stage -> P4)
  auto deinterleave = [&](int i) {
    const int g = g_first + i;
    const uint8_t* stage = raw;
    const uint32_t lead = (uint32_t)((reinterpret_cast<uintptr_t>(d.src) + (size_t)(4 * g) * d.pitch) & 15);
    const int nrows = min(4, d.H - 4 * g);
    uint4* Pb = P4;
    const uint8_t* r0 = stage + lead;
    if (fast3) {
      // AL: rows start on 4-byte boundaries (compile-time, so the aligned loop carries no trace of
      // the other one); otherwise (odd pitch: width not a multiple of 4) aligned words are read
      // and funnel-shifted into place.
      auto loop3 = [&](auto AL) {
        for (int w = tid; w < nwords; w += nthreads) {
          uint32_t R[4], G[4], B[4];
#pragma unroll
          for (int r = 0; r < 4; r++) {
            uint32_t x0 = 0, x1 = 0, x2 = 0;
            if (r < nrows) {
              const uint8_t* bp = r0 + r * d.pitch + 12 * w;
              if (decltype(AL)::value) {
                const uint32_t* rp = reinterpret_cast<const uint32_t*>(bp);
                x0 = rp[0]; x1 = rp[1]; x2 = rp[2];  // R0 G0 B0 R1 | G1 B1 R2 G2 | B2 R3 G3 B3
              } else {
                const uint32_t sh = (smem_u32(bp) & 3u) * 8u;
                const uint32_t* rp = reinterpret_cast<const uint32_t*>(bp - (sh >> 3));
                const uint32_t y0 = rp[0], y1 = rp[1], y2 = rp[2], y3 = rp[3];
                x0 = __funnelshift_r(y0, y1, sh); x1 = __funnelshift_r(y1, y2, sh); x2 = __funnelshift_r(y2, y3, sh);
              }
            }
            R[r] = __byte_perm(__byte_perm(x0, x1, 0x0630), x2, 0x5210);
            G[r] = __byte_perm(__byte_perm(x0, x1, 0x0741), x2, 0x6210);
            B[r] = __byte_perm(__byte_perm(x0, x1, 0x0052), x2, 0x7410);
          }
          const int pw = p4_phys(w);
          Pb[pw] = make_uint4(R[0], R[1], R[2], R[3]);
          Pb[(C > 1 ? 1 : 0) * wqp + pw] = make_uint4(G[0], G[1], G[2], G[3]);
          Pb[(C > 2 ? 2 : 0) * wqp + pw] = make_uint4(B[0], B[1], B[2], B[3]);
        }
      };
      if (rows_al4) loop3(std::true_type{}); else loop3(std::false_type{});
    } else if (fast1) {
      for (int w = tid; w < nwords; w += nthreads) {
        uint32_t v[4];
#pragma unroll
        for (int r = 0; r < 4; r++) {
          v[r] = 0u;
          if (r < nrows) {
            const uint8_t* bp = r0 + r * d.pitch + 4 * w;
            if (rows_al4) {
              v[r] = *reinterpret_cast<const uint32_t*>(bp);
            } else {
              const uint32_t sh = (smem_u32(bp) & 3u) * 8u;
              const uint32_t* rp = reinterpret_cast<const uint32_t*>(bp - (sh >> 3));
              v[r] = __funnelshift_r(rp[0], rp[1], sh);
            }
          }
        }
        Pb[p4_phys(w)] = make_uint4(v[0], v[1], v[2], v[3]);
      }
    } else {  // any channel conversion / rows not 4-byte aligned: byte-wise
      for (int it = tid; it < nwords * C; it += nthreads) {
        const int ch = it / nwords, w = it - ch * nwords;
        uint32_t v[4];
#pragma unroll
        for (int r = 0; r < 4; r++) {
          uint32_t word = 0;
          if (r < nrows) {
            const uint8_t* row = r0 + (size_t)r * d.pitch;
#pragma unroll
            for (int bb = 0; bb < 4; bb++) {
              const int x = 4 * w + bb;
              if (x < d.W) word |= fetch_conv(row, d.cin, C, x, ch) << (8 * bb);
            }
          }
          v[r] = word;
        }
        Pb[ch * wqp + p4_phys(w)] = make_uint4(v[0], v[1], v[2], v[3]);
      }
    }
  };

  // vertical pass of output row yy (all groups it needs are in the ring). One unit = 4 consecutive
  // output elements (one 16-byte tmp4 load per group, 128-bit store for 4-byte elements). Everything
  // that depends only on the thread (unit -> line/column, table rows, destination offset) is
  // computed once, above the row loop.
  int v_i0 = 0, v_c0 = 0, v_c1 = 0, v_c2 = 0, v_c3 = 0;
  size_t v_e0 = 0;              // element offset of the unit at yy = 0
  const size_t v_erow = nchw ? (size_t)out_w : (size_t)L;  // element offset increment per output row
  if (vec4 && tid < n_units) {
    if (nchw) {
      v_i0 = my_ln * out_w + 4 * my_x4;
      v_e0 = (size_t)my_ln * out_h * out_w + 4 * my_x4;
      v_c0 = v_c1 = v_c2 = v_c3 = my_ln * 256;
    } else {
      v_i0 = 4 * tid;
      v_e0 = (size_t)v_i0;
      if (C == 3) {
        const int c0 = tid % 3;  // (4 * tid) % 3
        const int c1 = c0 == 2 ? 0 : c0 + 1, c2 = c1 == 2 ? 0 : c1 + 1;
        v_c0 = c0 * 256; v_c1 = c1 * 256; v_c2 = c2 * 256; v_c3 = c0 * 256;
      }
    }
  }
  auto vertical = [&](int yy, int g_now, int slot_now) {
    const int gv0 = vg_first;
    const uint4* vc = reinterpret_cast<const uint4*>(pool + d.vc_off) + (size_t)yy * d.nwv;  // [nwv] x (k0, k1, k2, -)
    // ring slot of the row's first group: the group being written now sits in slot_now and the row
    // reaches back at most nwv = ng - 1 groups
    int slot0 = slot_now - (g_now - gv0);
    if (slot0 < 0) slot0 += ng;
    if (vec4) {
      if (tid < n_units) {  // first (normally only) unit of this thread: hoisted constants
        uint32_t a[4][3];
#pragma unroll
        for (int v = 0; v < 4; v++) { a[v][0] = 1u << (kPrecisionBits - 1); a[v][1] = 0; a[v][2] = 0; }
        // tap words: unrolled for the common counts (the generic loop cost ~45 instructions per word)
        const uint32_t* tbase = tmp4 + v_i0;
        auto tap = [&](int j, int slot) {
          const uint4 kk = __ldg(vc + j);
          const uint32_t k0 = kk.x, k1 = kk.y, k2 = kk.z;
          const uint4 px = *reinterpret_cast<const uint4*>(tbase + slot * L);
          a[0][0] = __dp4a(px.x, k0, a[0][0]); a[0][1] = __dp4a(px.x, k1, a[0][1]); a[0][2] = __dp4a(px.x, k2, a[0][2]);
          a[1][0] = __dp4a(px.y, k0, a[1][0]); a[1][1] = __dp4a(px.y, k1, a[1][1]); a[1][2] = __dp4a(px.y, k2, a[1][2]);
          a[2][0] = __dp4a(px.z, k0, a[2][0]); a[2][1] = __dp4a(px.z, k1, a[2][1]); a[2][2] = __dp4a(px.z, k2, a[2][2]);
          a[3][0] = __dp4a(px.w, k0, a[3][0]); a[3][1] = __dp4a(px.w, k1, a[3][1]); a[3][2] = __dp4a(px.w, k2, a[3][2]);
        };
        auto taps = [&](auto NWVc) {
          constexpr int NWV = decltype(NWVc)::value;
#pragma unroll
          for (int j = 0; j < NWV; j++) {
            int slot = slot0 + j;
            if (slot >= ng) slot -= ng;
            tap(j, slot);
          }
        };
        switch (d.nwv) {
          case 2: taps(std::integral_constant<int, 2>{}); break;
          case 3: taps(std::integral_constant<int, 3>{}); break;
          case 4: taps(std::integral_constant<int, 4>{}); break;
          default: {
            int slot = slot0;
            for (int j = 0; j < d.nwv; j++) {
              tap(j, slot);
              if (++slot == ng) slot = 0;
            }
          }
        }
        uint32_t u[4];
#pragma unroll
        for (int v = 0; v < 4; v++) u[v] = (a[v][0] + (a[v][1] << 8) + (a[v][2] << 16)) >> kPrecisionBits;  // <= 255, see coeffs_never_clip
        const size_t e0 = v_e0 + (size_t)yy * v_erow;
        if (elem == 4) {
          const uint32_t* l32 = reinterpret_cast<const uint32_t*>(lut);
          st_global_v4(dimg + e0 * 4, __ldg(l32 + v_c0 + u[0]), __ldg(l32 + v_c1 + u[1]), __ldg(l32 + v_c2 + u[2]),
                       __ldg(l32 + v_c3 + u[3]));
        } else if (elem == 2) {
          const uint16_t* l16 = reinterpret_cast<const uint16_t*>(lut);
          const uint32_t lo = __ldg(l16 + v_c0 + u[0]) | ((uint32_t)__ldg(l16 + v_c1 + u[1]) << 16);
          const uint32_t hi = __ldg(l16 + v_c2 + u[2]) | ((uint32_t)__ldg(l16 + v_c3 + u[3]) << 16);
          st_global_v2(dimg + e0 * 2, lo, hi);
        } else if (elem == 1) {
          const uint32_t w = __ldg(lut + v_c0 + u[0]) | ((uint32_t)__ldg(lut + v_c1 + u[1]) << 8) |
                             ((uint32_t)__ldg(lut + v_c2 + u[2]) << 16) | ((uint32_t)__ldg(lut + v_c3 + u[3]) << 24);
          *reinterpret_cast<uint32_t*>(dimg + e0) = w;
        } else {
          const uint2* l64 = reinterpret_cast<const uint2*>(lut);
          const uint2 q0 = __ldg(l64 + v_c0 + u[0]), q1 = __ldg(l64 + v_c1 + u[1]);
          const uint2 q2 = __ldg(l64 + v_c2 + u[2]), q3 = __ldg(l64 + v_c3 + u[3]);
          st_global_v4(dimg + e0 * 8, q0.x, q0.y, q1.x, q1.y);
          st_global_v4(dimg + e0 * 8 + 16, q2.x, q2.y, q3.x, q3.y);
        }
      }
      // further units (only when a row has more than blockDim units): word by word below
      for (int i = 4 * nthreads + 4 * tid; i < 4 * n_units; i += 4 * nthreads) {
#pragma unroll 1
        for (int v = 0; v < 4; v++) {
          const int iw = i + v;
          uint32_t a0 = 1u << (kPrecisionBits - 1), a1 = 0, a2 = 0;
          int slot = slot0;
          for (int j = 0; j < d.nwv; j++) {
            const uint32_t px = tmp4[(size_t)slot * L + iw];
            const uint4 kk = __ldg(vc + j);
            a0 = __dp4a(px, kk.x, a0); a1 = __dp4a(px, kk.y, a1); a2 = __dp4a(px, kk.z, a2);
            if (++slot == ng) slot = 0;
          }
          const uint32_t u = (a0 + (a1 << 8) + (a2 << 16)) >> kPrecisionBits;
          size_t e; int ch;
          if (nchw) { ch = iw / out_w; e = ((size_t)ch * out_h + yy) * out_w + (iw - ch * out_w); }
          else { ch = iw % C; e = (size_t)yy * L + iw; }
          const uint8_t* le = lut + ((size_t)ch * 256 + u) * elem;
          if (elem == 4) reinterpret_cast<uint32_t*>(dimg)[e] = __ldg(reinterpret_cast<const uint32_t*>(le));
          else if (elem == 2) reinterpret_cast<uint16_t*>(dimg)[e] = __ldg(reinterpret_cast<const uint16_t*>(le));
          else if (elem == 1) dimg[e] = __ldg(le);
          else reinterpret_cast<unsigned long long*>(dimg)[e] = __ldg(reinterpret_cast<const unsigned long long*>(le));
        }
      }
    } else {
      for (int i = tid; i < L; i += nthreads) {
        uint32_t a0 = 1u << (kPrecisionBits - 1), a1 = 0, a2 = 0;
        int slot = slot0;
        for (int j = 0; j < d.nwv; j++) {
          const uint32_t px = tmp4[(size_t)slot * L + i];
          const uint4 kk = __ldg(vc + j);
          a0 = __dp4a(px, kk.x, a0); a1 = __dp4a(px, kk.y, a1); a2 = __dp4a(px, kk.z, a2);
          if (++slot == ng) slot = 0;
        }
        const uint32_t u = (a0 + (a1 << 8) + (a2 << 16)) >> kPrecisionBits;
        size_t e;
        int ch;
        if (nchw) { ch = i / out_w; e = ((size_t)ch * out_h + yy) * out_w + (i - ch * out_w); }
        else { ch = i % C; e = (size_t)yy * L + i; }
        const uint8_t* le = lut + ((size_t)ch * 256 + u) * elem;
        if (elem == 4) reinterpret_cast<uint32_t*>(dimg)[e] = __ldg(reinterpret_cast<const uint32_t*>(le));
        else if (elem == 2) reinterpret_cast<uint16_t*>(dimg)[e] = __ldg(reinterpret_cast<const uint16_t*>(le));
        else if (elem == 1) dimg[e] = __ldg(le);
        else reinterpret_cast<unsigned long long*>(dimg)[e] = __ldg(reinterpret_cast<const unsigned long long*>(le));
      }
    }
  };

  // ---- main loop: one raw stage, one P4 buffer, two barriers per row group ----------------------
  // (three CTAs per SM instead of two: the other CTAs' phases fill the barrier gaps)
  int slot = 0;
  for (int i = 0; i <= n_groups; i++) {  // the last iteration only drains the vertical pass
    const int g = g_first + i;
    if (i < n_groups) {
      mbar_wait(&bars[0], tma_phase);
      tma_phase ^= 1u;
      deinterleave(i);
      __syncthreads();  // raw stage drained, P4 complete
      if (tid == 0 && i + 1 < n_groups) issue(i + 1);
      if ((flags >> 1) && tid == 32 && i + 1 + (flags >> 1) < n_groups) {  // the group after that one: towards L2
        const int gp = g_first + i + 1 + (flags >> 1);
        const int r1 = min(4 * gp + 4, d.H);
        const uintptr_t start = reinterpret_cast<uintptr_t>(d.src) + (size_t)(4 * gp) * d.pitch;
        const uintptr_t end = reinterpret_cast<uintptr_t>(d.src) + (size_t)(r1 - 1) * d.pitch + (size_t)d.W * d.cin;
        const uintptr_t a0 = start & ~(uintptr_t)15, a1 = (end + 15) & ~(uintptr_t)15;
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a0), "r"((uint32_t)(a1 - a0)) : "memory");
      }
      if (tmp_idx0 >= 0) {
        const uint4* Pb = P4;
        uint32_t* out = tmp4 + (size_t)slot * L + tmp_idx0;
#pragma unroll
        for (int ch = 0; ch < C; ch++) {
          uint32_t a[4][3];
#pragma unroll
          for (int r = 0; r < 4; r++) { a[r][0] = 1u << (kPrecisionBits - 1); a[r][1] = 0; a[r][2] = 0; }
          const uint4* p = Pb + ch * wqp;
#pragma unroll
          for (int j = 0; j < NW; j++) {
            const uint4 v = p[pw0 + j];
            a[0][0] = __dp4a(v.x, kc[j][0], a[0][0]); a[0][1] = __dp4a(v.x, kc[j][1], a[0][1]); a[0][2] = __dp4a(v.x, kc[j][2], a[0][2]);
            a[1][0] = __dp4a(v.y, kc[j][0], a[1][0]); a[1][1] = __dp4a(v.y, kc[j][1], a[1][1]); a[1][2] = __dp4a(v.y, kc[j][2], a[1][2]);
            a[2][0] = __dp4a(v.z, kc[j][0], a[2][0]); a[2][1] = __dp4a(v.z, kc[j][1], a[2][1]); a[2][2] = __dp4a(v.z, kc[j][2], a[2][2]);
            a[3][0] = __dp4a(v.w, kc[j][0], a[3][0]); a[3][1] = __dp4a(v.w, kc[j][1], a[3][1]); a[3][2] = __dp4a(v.w, kc[j][2], a[3][2]);
          }
          uint32_t u[4];
#pragma unroll
          for (int r = 0; r < 4; r++) u[r] = (a[r][0] + (a[r][1] << 8) + (a[r][2] << 16)) >> kPrecisionBits;  // <= 255, see coeffs_never_clip
          out[ch * tmp_chs] = __byte_perm(__byte_perm(u[0], u[1], 0x0040), __byte_perm(u[2], u[3], 0x0040), 0x5410);
        }
      }
    }
    // vertical pass of the rows whose last group was written in an earlier phase
    while (yy_next < yy1 && vg_last < g) {
      // the next row's bookkeeping is requested before this row's work and consumed after it
      const int nxt = min(yy_next + 1, yy1 - 1);
      const int nf = vg[2 * nxt], nl = vg[2 * nxt + 1];
      vertical(yy_next, g, slot);
      yy_next++;
      vg_first = nf; vg_last = nl;
    }
    if (i < n_groups) {
      __syncthreads();  // P4 free again, tmp4 slot visible
      if (++slot == ng) slot = 0;
    }
  }
  }  // work items of this CTA
}

// =====================================================================================================
// IMMA variant of the fused kernel (RGB -> RGB, rows 16-byte aligned).
//
// The horizontal pass is a banded u8 GEMM: tmp[line][xx] = sum_x P[line][x] * k[x][xx], exact in
// s32 when the 22-bit coefficient is split into three byte planes (k = k0 + k1*2^8 + k2*2^16), which
// is what the dp4a kernel above already does one lane at a time. Here a warp does it with
// mma.sync.m16n8k32 (u8 x u8 -> s32, SASS IMMA.16832.U8.U8): A = 16 lines x 32 source pixels from
// shared memory via ldmatrix, B = 32 pixels x 8 output columns of one coefficient plane, resident in
// registers for the two 8-column tiles every warp owns. One 8-column tile sees a window of NK*32
// source pixels (NK <= 3 for scales up to ~8.6), so a tile costs NK*3 IMMAs per 16 lines instead of
// 8 columns x 16 lines x NK*8... dp4a lane-instructions: about a fifth of the issue slots.
//
// Rows stream through the CTA eight at a time (two 4-row groups = two words of the tmp4 ring):
//   wait TMA(s) | de-interleave raw -> planar lines P[slot][x] (LDS.128, PRMT, STS.128) | barrier |
//   start TMA(s+1) | IMMA pass P -> tmp4 ring | barrier | vertical pass (dp4a) + table + store.
// Lines: slot = r*6 + ch*2 + half holds row 4*half + r of the step, channel ch. A 16-line MMA tile t
// is rows r = 2t, 2t+1 of all six (ch, half) pairs (+ 2 pad lines that alias pair 0, 1), so thread
// (g = lane/4) ends up with the four rows of pair g for two adjacent columns: exactly one tmp4 word
// each, packed with PRMT, no shuffles.
// =====================================================================================================
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void imma_16832(int (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
      : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void imma_16832_first(int (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1, int init) {
  asm volatile(
      "mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%10, %10, %10, %10};"
      : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1), "r"(init));
}

constexpr int kMmaLines = 24;  // 4 rows x 3 channels x 2 halves

// TPW: 8-column tiles per warp. TPW = 2: up to 14 warps, two CTAs per SM (72 registers), the 36
// fragment registers leave no room to interleave the accumulation chains. TPW = 1: up to 28 warps,
// one CTA per SM (72 registers again), 18 fragment registers, both 16-line tiles of a column tile in
// flight at once and ~115 KB of L1 left for tables.
template <int NK, int TPW, int WMAX>
__global__ void __launch_bounds__(WMAX * 32, WMAX <= 16 ? 2 : 1)
k_preprocess_fused_mma(const FusedDesc* __restrict__ descs, const int4* __restrict__ items,
                       const int32_t* __restrict__ cta_first, const int32_t* __restrict__ pool,
                       const uint8_t* __restrict__ lut, uint8_t* __restrict__ dst, int out_h, int out_w,
                       int layout, int elem, int flags) {
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr int C = 3;
  constexpr int kMmaTilesPerWarp = TPW;
  const int tid = threadIdx.x, nthreads = blockDim.x;
  const int lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, tig = lane & 3;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  // raw stages: two when the CTA has an SM to itself (the TMA of step s + 2 starts as soon as step s has
  // been de-interleaved and has a whole step to land), one when two CTAs share the shared memory
  constexpr int NST = TPW == 1 ? 2 : 1;
  uint32_t tma_phase = 0;  // bit i: parity of the next completion of bars[i]; runs on across work items
  const bool nchw = layout == TCB_LAYOUT_NCHW;
  const int L = out_w * C;
  const int ntiles = (out_w + 7) >> 3;
  // ---- vertical-pass constants (depend on the launch only) ---------------------------------------
  const bool vec4 = (flags & 1) && (nchw ? (out_w % 4 == 0) : (L % 4 == 0));  // bit 0: dst is 16-byte aligned
  const int upl = nchw ? out_w / 4 : L / 4;       // units (4 consecutive output elements) per line
  const int n_units = vec4 ? (nchw ? C * upl : upl) : 0;
  const int vb_rows = n_units > 0 ? max(1, nthreads / n_units) : 1;  // output rows handled side by side
  const int vk = n_units > 0 ? tid / n_units : 0;  // which of the side-by-side rows this thread serves (>= vb_rows: none)
  int v_i0 = 0, v_e0 = 0, v_lut = 0;              // first tmp4 word, element offset at yy = 0, packed table rows
  if (n_units > 0) {
    const int vu0 = tid - vk * n_units;
    if (nchw) {
      const int ln = vu0 / upl, x4 = vu0 - ln * upl;
      v_i0 = ln * out_w + 4 * x4;
      v_e0 = v_i0 + ln * (out_h - 1) * out_w;
      v_lut = (ln * 256) * (1 + 1024 + 1024 * 1024);
    } else {
      v_i0 = v_e0 = 4 * vu0;
      const int m3 = vu0 % 3;  // (4 * vu0) % 3
      const int n1 = m3 == 2 ? 0 : m3 + 1, n2 = n1 == 2 ? 0 : n1 + 1;
      v_lut = (m3 * 256) | ((n1 * 256) << 10) | ((n2 * 256) << 20);
    }
  }
  const int v_erow = nchw ? out_w : L;
  // value table and (per item) the row -> (first group, last group) table live in shared memory: with two
  // CTAs of ~110 KB per SM the L1 cache that would otherwise hold them is a few KB
  uint8_t* lut_s = smem + 128;
  int32_t* vg_s = reinterpret_cast<int32_t*>(lut_s + 768 * elem);
  const int vg_bytes = (out_h * 8 + 127) & ~127;
  for (int i = tid; i < 768 * elem / 4; i += nthreads)
    reinterpret_cast<uint32_t*>(lut_s)[i] = __ldg(reinterpret_cast<const uint32_t*>(lut) + i);
  // ---- horizontal-pass constants --------------------------------------------------------------------
  // accumulator ownership: lines g, g+8 of both tiles = rows 0..3 of pair g = (channel g>>1, half g&1)
  const int own_ch = g >> 1, own_half = g & 1;
  const int st_cs = nchw ? 1 : C;  // tmp4 words between neighbouring columns
  int st_base = 0, st_mask = 0;    // first word of this thread's columns (tile q = 0), which of its 4 stores exist
  {
    const int col0 = warp * kMmaTilesPerWarp * 8 + 2 * tig;
    st_base = nchw ? own_ch * out_w + col0 : col0 * C + own_ch;
    if (g < 6)
      for (int q = 0; q < kMmaTilesPerWarp; q++) {
        if (col0 + 8 * q < out_w) st_mask |= 1 << (2 * q);
        if (col0 + 8 * q + 1 < out_w) st_mask |= 2 << (2 * q);
      }
  }

  for (int item = cta_first[blockIdx.x]; item < cta_first[blockIdx.x + 1]; item++) {
    const int4 wi = items[item];
    const int img = wi.x, yy0 = wi.y, yy1 = wi.z;
    const FusedDesc d = descs[img];
    const int ng = d.nwv + 1;
    const int Wp = d.mma_wp;
    // shared memory: barrier | raw stage (8 rows) | P lines | tmp4 ring
    uint8_t* raw = reinterpret_cast<uint8_t*>(vg_s) + vg_bytes;
    uint8_t* P = raw + NST * d.stage_bytes;
    uint32_t* tmp4 = reinterpret_cast<uint32_t*>(P + (size_t)kMmaLines * Wp);
    const uint32_t P_s = smem_u32(P);

    const int32_t* vb = pool + d.vb_off;
    const int ys = vb[2 * yy0], ye = vb[2 * (yy1 - 1)] + vb[2 * (yy1 - 1) + 1];  // source rows [ys, ye)
    const int s_first = ys >> 3, s_last = (ye - 1) >> 3;

    auto issue = [&](int s) {  // thread 0: rows [8s, min(8s + 8, H)) -> raw stage (s - s_first) % NST
      const int r0 = 8 * s, r1 = min(r0 + 8, d.H);
      const int st = (s - s_first) % NST;
      const uint8_t* a0 = d.src + (size_t)r0 * d.pitch;
      const uint32_t bytes = (uint32_t)((r1 - r0 - 1) * d.pitch + ((d.W * 3 + 15) & ~15));
      mbar_expect_tx(&bars[st], bytes);
      tma_load_1d(raw + (size_t)st * d.stage_bytes, a0, bytes, &bars[st]);
    };
    // different images = different shared-memory layouts: see k_preprocess_fused
    if (item != cta_first[blockIdx.x]) __syncthreads();
    if (tid == 0)
      for (int i = 0; i < NST && s_first + i <= s_last; i++) issue(s_first + i);
    for (int i = tid; i < 2 * out_h; i += nthreads) vg_s[i] = pool[d.vg_off + i];  // visible after the step's first barrier

    // ---- per-item constants ---------------------------------------------------------------------
    // B fragments of this warp's two tiles
    uint32_t bfr[kMmaTilesPerWarp][NK][3][2];
    int s_nt[kMmaTilesPerWarp];
#pragma unroll
    for (int q = 0; q < kMmaTilesPerWarp; q++) {
      const int nt = warp * kMmaTilesPerWarp + q;
      s_nt[q] = 0;
#pragma unroll
      for (int ks = 0; ks < NK; ks++)
#pragma unroll
        for (int p = 0; p < 3; p++) { bfr[q][ks][p][0] = 0; bfr[q][ks][p][1] = 0; }
      if (nt < ntiles) {
        s_nt[q] = pool[d.mma_s_off + nt];
        const uint2* bp = reinterpret_cast<const uint2*>(pool + d.mma_b_off) + (size_t)nt * NK * 3 * 32 + lane;
#pragma unroll
        for (int ks = 0; ks < NK; ks++)
#pragma unroll
          for (int p = 0; p < 3; p++) {
            const uint2 v = __ldg(bp + (ks * 3 + p) * 32);
            bfr[q][ks][p][0] = v.x; bfr[q][ks][p][1] = v.y;
          }
      }
    }
    // ldmatrix row address of this lane for tile t: matrix mi = lane >> 3 -> rows (mi & 1) * 8 .., k half mi >> 1
    uint32_t a_addr[2];
    {
      const int mi = lane >> 3, rr = lane & 7;
      const int jj = rr < 6 ? rr : rr - 6;  // pad lines alias pairs 0 and 1 (results unused)
#pragma unroll
      for (int t = 0; t < 2; t++) {
        const int r = 2 * t + (mi & 1);
        a_addr[t] = P_s + (uint32_t)((r * 6 + jj) * Wp + 16 * (mi >> 1));
      }
    }
    // de-interleave work split: item i = (row, 16-pixel chunk), i = tid, tid + T, ...
    const int nch = (d.W + 15) >> 4;
    const int di_r0 = tid / nch, di_m0 = tid - di_r0 * nch;
    const int di_dr = nthreads / nch, di_dm = nthreads - di_dr * nch;
    // tmp4 word of this thread's two columns
    uint8_t* dimg = dst + (size_t)img * C * out_h * out_w * elem;
    int yy_next = yy0;
    int slotA = 0, slotB = ng > 1 ? 1 : 0;  // ring slots of the step's two row groups

    int rows_done = pool[d.vq_off + s_first], rows_done_next = pool[d.vq_off + min(s_first + 1, s_last)];
    for (int s = s_first; s <= s_last; s++) {
      const int nrows = min(8, d.H - 8 * s);
      const int st = (s - s_first) % NST;
      const uint8_t* rawst = raw + (size_t)st * d.stage_bytes;
      mbar_wait(&bars[st], (tma_phase >> st) & 1u);
      tma_phase ^= 1u << st;
      // ---- de-interleave: 16 pixels (48 bytes) of one row -> three 16-byte planar pieces ------------
      {
        int r = di_r0, m = di_m0;
        while (r < nrows) {
          const uint4* sp = reinterpret_cast<const uint4*>(rawst + (size_t)r * d.pitch + 48 * m);
          const uint4 v0 = sp[0], v1 = sp[1], v2 = sp[2];
          const uint32_t x[12] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w, v2.x, v2.y, v2.z, v2.w};
          uint32_t R[4], G[4], B[4];
#pragma unroll
          for (int q = 0; q < 4; q++) {
            const uint32_t x0 = x[3 * q], x1 = x[3 * q + 1], x2 = x[3 * q + 2];  // R0 G0 B0 R1 | G1 B1 R2 G2 | B2 R3 G3 B3
            R[q] = __byte_perm(__byte_perm(x0, x1, 0x0630), x2, 0x5210);
            G[q] = __byte_perm(__byte_perm(x0, x1, 0x0741), x2, 0x6210);
            B[q] = __byte_perm(__byte_perm(x0, x1, 0x0052), x2, 0x7410);
          }
          uint8_t* pl = P + (size_t)(((r & 3) * 6 + (r >> 2)) * Wp + 16 * m);
          *reinterpret_cast<uint4*>(pl) = make_uint4(R[0], R[1], R[2], R[3]);
          *reinterpret_cast<uint4*>(pl + 2 * Wp) = make_uint4(G[0], G[1], G[2], G[3]);
          *reinterpret_cast<uint4*>(pl + 4 * Wp) = make_uint4(B[0], B[1], B[2], B[3]);
          r += di_dr; m += di_dm;
          if (m >= nch) { m -= nch; r++; }
        }
      }
      __syncthreads();  // raw drained, P complete
      if (tid == 0 && s + NST <= s_last) issue(s + NST);  // into the stage that has just been drained
      // flags >> 1: also ask L2 for the rows of the step after that one, so that its bulk copy, issued one
      // step from now with a single step to land, finds them on chip (cfg3: 0.431 -> 0.426 ms; two or more
      // steps ahead is slower: 0.427, 0.432, 0.448 ms for 2, 3, 4)
      if ((flags >> 1) && tid == 32 && s + NST + (flags >> 1) <= s_last) {
        const int r0 = 8 * (s + NST + (flags >> 1)), r1 = min(r0 + 8, d.H);
        const uint32_t bytes = (uint32_t)((r1 - r0 - 1) * d.pitch + ((d.W * 3 + 15) & ~15));
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(d.src + (size_t)r0 * d.pitch), "r"(bytes) : "memory");
      }
      // ---- horizontal pass: IMMA ----------------------------------------------------------------------
      // (tiles beyond the last one compute on all-zero fragments and store nothing: no branch per tile)
      uint32_t* st_ptr = tmp4 + (own_half ? slotB : slotA) * L + st_base;
#pragma unroll
      for (int q = 0; q < kMmaTilesPerWarp; q++) {
        uint32_t w0 = 0, w1 = 0;
#pragma unroll
        for (int t = 0; t < 2; t++) {
          int acc[3][4];
#pragma unroll
          for (int ks = 0; ks < NK; ks++) {
            uint32_t a[4];
            ldsm_x4(a, a_addr[t] + (uint32_t)(s_nt[q] + 32 * ks));
#pragma unroll
            for (int p = 0; p < 3; p++) {
              // the fragments hold 4 * k, plane 0 starts from 4 * 2^21: the sum of the three planes then
              // carries the rounded result (sum >> 22) in its top byte, where PRMT picks it up
              if (ks == 0) imma_16832_first(acc[p], a, bfr[q][ks][p][0], bfr[q][ks][p][1], p == 0 ? (4 << (kPrecisionBits - 1)) : 0);
              else imma_16832(acc[p], a, bfr[q][ks][p][0], bfr[q][ks][p][1]);
            }
          }
          uint32_t u[4];
#pragma unroll
          for (int k = 0; k < 4; k++)  // top byte <= 255, see coeffs_never_clip
            u[k] = (uint32_t)acc[0][k] + ((uint32_t)acc[1][k] << 8) + ((uint32_t)acc[2][k] << 16);
          // rows 2t, 2t+1 (c0/c1 and c2/c3) of columns 2*tig / 2*tig + 1 -> byte lanes 2t, 2t+1
          const uint32_t h0 = __byte_perm(u[0], u[2], 0x0073), h1 = __byte_perm(u[1], u[3], 0x0073);
          w0 = t == 0 ? h0 : __byte_perm(w0, h0, 0x5410);
          w1 = t == 0 ? h1 : __byte_perm(w1, h1, 0x5410);
        }
        // this thread's two columns of pair g: index and guards fixed per thread
        uint32_t* o = st_ptr + q * 8 * st_cs;
        if (st_mask & (1 << (2 * q))) o[0] = w0;
        if (st_mask & (2 << (2 * q))) o[st_cs] = w1;
      }
      __syncthreads();  // P free again, ring slots visible
      // ---- vertical pass of the output rows whose last row group is now in the ring ------------------------
      // Every thread owns at most one unit (4 consecutive output elements) of one of the vb_rows rows
      // handled side by side; everything that depends only on the thread was computed before the item
      // loop (a first version recomputed unit -> line/column/table rows here and spent 40 % of the
      // kernel's instructions in this block).
      const int g_now = 2 * s + 1;
      const int yend = min(rows_done, yy1);  // rows [yy_next, yend) have all their row groups in the ring
      rows_done = rows_done_next;
      if (s < s_last) rows_done_next = __ldg(pool + d.vq_off + min(s + 2, s_last));  // consumed two steps from now
      while (yy_next < yend) {
        int nready = min(vb_rows, yend - yy_next);
        if (vec4) {
          if (vk < nready) {
            const int yy = yy_next + vk;
            const int c0 = v_lut & 1023, c1 = (v_lut >> 10) & 1023, c2 = v_lut >> 20;  // table rows (the fourth equals the first)
            int slot0 = slotB - (g_now - vg_s[2 * yy]);
            if (slot0 < 0) slot0 += ng;
            const uint4* vc = reinterpret_cast<const uint4*>(pool + d.vc_off) + (size_t)yy * d.nwv;
            uint32_t a[4][3];
#pragma unroll
            for (int v = 0; v < 4; v++) { a[v][0] = 1u << (kPrecisionBits - 1); a[v][1] = 0; a[v][2] = 0; }
            const uint32_t* tbase = tmp4 + v_i0;
            auto tap = [&](int j, int slot) {
              const uint4 kk = __ldg(vc + j);
              const uint4 px = *reinterpret_cast<const uint4*>(tbase + slot * L);
              a[0][0] = __dp4a(px.x, kk.x, a[0][0]); a[0][1] = __dp4a(px.x, kk.y, a[0][1]); a[0][2] = __dp4a(px.x, kk.z, a[0][2]);
              a[1][0] = __dp4a(px.y, kk.x, a[1][0]); a[1][1] = __dp4a(px.y, kk.y, a[1][1]); a[1][2] = __dp4a(px.y, kk.z, a[1][2]);
              a[2][0] = __dp4a(px.z, kk.x, a[2][0]); a[2][1] = __dp4a(px.z, kk.y, a[2][1]); a[2][2] = __dp4a(px.z, kk.z, a[2][2]);
              a[3][0] = __dp4a(px.w, kk.x, a[3][0]); a[3][1] = __dp4a(px.w, kk.y, a[3][1]); a[3][2] = __dp4a(px.w, kk.z, a[3][2]);
            };
            auto taps = [&](auto NWVc) {
              constexpr int NWV = decltype(NWVc)::value;
#pragma unroll
              for (int j = 0; j < NWV; j++) {
                int slot = slot0 + j;
                if (slot >= ng) slot -= ng;
                tap(j, slot);
              }
            };
            switch (d.nwv) {
              case 2: taps(std::integral_constant<int, 2>{}); break;
              case 3: taps(std::integral_constant<int, 3>{}); break;
              case 4: taps(std::integral_constant<int, 4>{}); break;
              default: {
                int slot = slot0;
                for (int j = 0; j < d.nwv; j++) {
                  tap(j, slot);
                  if (++slot == ng) slot = 0;
                }
              }
            }
            uint32_t u[4];
#pragma unroll
            for (int v = 0; v < 4; v++) u[v] = (a[v][0] + (a[v][1] << 8) + (a[v][2] << 16)) >> kPrecisionBits;
            const size_t e0 = (size_t)(v_e0 + yy * v_erow);
            if (elem == 4) {
              const uint32_t* l32 = reinterpret_cast<const uint32_t*>(lut_s);
              st_global_v4(dimg + e0 * 4, l32[c0 + u[0]], l32[c1 + u[1]], l32[c2 + u[2]], l32[c0 + u[3]]);
            } else if (elem == 2) {
              const uint16_t* l16 = reinterpret_cast<const uint16_t*>(lut_s);
              const uint32_t lo = l16[c0 + u[0]] | ((uint32_t)l16[c1 + u[1]] << 16);
              const uint32_t hi = l16[c2 + u[2]] | ((uint32_t)l16[c0 + u[3]] << 16);
              st_global_v2(dimg + e0 * 2, lo, hi);
            } else if (elem == 1) {
              const uint32_t w = lut_s[c0 + u[0]] | ((uint32_t)lut_s[c1 + u[1]] << 8) |
                                 ((uint32_t)lut_s[c2 + u[2]] << 16) | ((uint32_t)lut_s[c0 + u[3]] << 24);
              *reinterpret_cast<uint32_t*>(dimg + e0) = w;
            } else {
              const uint2* l64 = reinterpret_cast<const uint2*>(lut_s);
              const uint2 q0 = l64[c0 + u[0]], q1 = l64[c1 + u[1]], q2 = l64[c2 + u[2]], q3 = l64[c0 + u[3]];
              st_global_v4(dimg + e0 * 8, q0.x, q0.y, q1.x, q1.y);
              st_global_v4(dimg + e0 * 8 + 16, q2.x, q2.y, q3.x, q3.y);
            }
          }
        } else {  // destination not 16-byte aligned / odd row length: one element at a time
          nready = 1;
          const int yy = yy_next;
          int slot0 = slotB - (g_now - vg_s[2 * yy]);
          if (slot0 < 0) slot0 += ng;
          const uint4* vc = reinterpret_cast<const uint4*>(pool + d.vc_off) + (size_t)yy * d.nwv;
          for (int i = tid; i < L; i += nthreads) {
            uint32_t a0 = 1u << (kPrecisionBits - 1), a1 = 0, a2 = 0;
            int slot = slot0;
            for (int j = 0; j < d.nwv; j++) {
              const uint32_t px = tmp4[(size_t)slot * L + i];
              const uint4 kk = __ldg(vc + j);
              a0 = __dp4a(px, kk.x, a0); a1 = __dp4a(px, kk.y, a1); a2 = __dp4a(px, kk.z, a2);
              if (++slot == ng) slot = 0;
            }
            const uint32_t u = (a0 + (a1 << 8) + (a2 << 16)) >> kPrecisionBits;
            size_t e;
            int ch;
            if (nchw) { ch = i / out_w; e = ((size_t)ch * out_h + yy) * out_w + (i - ch * out_w); }
            else { ch = i % C; e = (size_t)yy * L + i; }
            const uint8_t* le = lut_s + ((size_t)ch * 256 + u) * elem;
            if (elem == 4) reinterpret_cast<uint32_t*>(dimg)[e] = *reinterpret_cast<const uint32_t*>(le);
            else if (elem == 2) reinterpret_cast<uint16_t*>(dimg)[e] = *reinterpret_cast<const uint16_t*>(le);
            else if (elem == 1) dimg[e] = *le;
            else reinterpret_cast<unsigned long long*>(dimg)[e] = *reinterpret_cast<const unsigned long long*>(le);
          }
        }
        yy_next += nready;
      }
      // next step's groups go to the two slots after slotB
      slotA = slotB + 1; if (slotA >= ng) slotA -= ng;
      slotB = slotA + 1; if (slotB >= ng) slotB -= ng;
    }
  }  // work items of this CTA
}

// ---- host side ---------------------------------------------------------------------------------------
// Packs Pillow's int32 coefficients of one axis into dp4a byte planes on the 4-sample word grid.
// Returns the words per output (max over outputs); first[xx] = first word index.
static int pack_axis(const std::vector<int32_t>& bounds, const std::vector<int32_t>& kk, int ksize, int out,
                     int force_nw, std::vector<int32_t>& packed, std::vector<int32_t>& first,
                     std::vector<int32_t>* last = nullptr) {
  int nw = 1;
  for (int o = 0; o < out; o++) {
    const int xmin = bounds[2 * o], n = bounds[2 * o + 1];
    nw = std::max(nw, ((xmin + n - 1) >> 2) - (xmin >> 2) + 1);
  }
  if (force_nw > 0) {
    if (force_nw < nw) return -nw;
    nw = force_nw;
  }
  packed.assign((size_t)out * nw * 3, 0);
  first.assign(out, 0);
  if (last) last->assign(out, 0);
  for (int o = 0; o < out; o++) {
    const int xmin = bounds[2 * o], n = bounds[2 * o + 1];
    const int w0 = xmin >> 2;
    first[o] = w0;
    if (last) (*last)[o] = (xmin + n - 1) >> 2;
    for (int t = 0; t < n; t++) {
      const uint32_t k = (uint32_t)kk[(size_t)o * ksize + t];
      const int pos = xmin + t - 4 * w0, j = pos >> 2, b = pos & 3;
      for (int p = 0; p < 3; p++) {
        uint32_t& word = reinterpret_cast<uint32_t&>(packed[((size_t)o * nw + j) * 3 + p]);
        word |= ((k >> (8 * p)) & 0xffu) << (8 * b);
      }
    }
  }
  return nw;
}

// Pillow clips every pass to [0, 255]. With non-negative weights (BILINEAR) the clip never acts as
// long as 2^21 + 255 * sum(k) < 2^30 for every output sample, i.e. sum(k) <= 2^22 + 8224; rounding
// moves the sum by at most half a unit per tap. The fused kernel relies on this (no min/max in
// the inner loops); a geometry that violates it takes the two-pass kernels.
static bool coeffs_never_clip(const std::vector<int32_t>& bounds, const std::vector<int32_t>& kk, int ksize, int out) {
  for (int o = 0; o < out; o++) {
    long long sum = 0;
    for (int t = 0; t < bounds[2 * o + 1]; t++) {
      const int32_t k = kk[(size_t)o * ksize + t];
      if (k < 0) return false;
      sum += k;
    }
    if ((1ll << (kPrecisionBits - 1)) + 255ll * sum >= (1ll << 30)) return false;
  }
  return true;
}

// Which output column each thread of the CTA takes in the horizontal pass. Thread = column makes
// consecutive lanes read P4 granules ~scale/4 apart (2.14 for 1080p -> 224), and the 8 lanes that
// share one shared-memory wavefront of a 128-bit load then fall on 4-5 distinct 16-byte bank
// groups: a 2-way conflict on the kernel's dominant shared-memory traffic. No padding or XOR
// swizzle of P4 removes it (the stride is irregular), but the lane -> column map is free: group
// the columns so that every quarter-warp holds 8 columns whose first granules differ mod 8 (all
// NW loads shift them together). The better of this map and the identity is used.
static double lds128_cost(const std::vector<int32_t>& first, const std::vector<int32_t>& perm, int out_w) {
  long long waves = 0, n = 0;
  for (size_t base = 0; base < perm.size(); base += 8) {
    int cnt[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    std::vector<int> seen;
    int worst = 0;
    for (size_t l = base; l < std::min(perm.size(), base + 8); l++) {
      if (perm[l] >= out_w) continue;
      const int g = first[perm[l]];
      if (std::find(seen.begin(), seen.end(), g) != seen.end()) continue;  // same granule: broadcast
      seen.push_back(g);
      worst = std::max(worst, ++cnt[g & 7]);
    }
    if (worst) { waves += worst; n++; }
  }
  return n ? (double)waves / (double)n : 1.0;
}

static std::vector<int32_t> lane_columns(const std::vector<int32_t>& first, int out_w, int threads) {
  std::vector<int32_t> ident(threads), perm;
  for (int t = 0; t < threads; t++) ident[t] = std::min(t, out_w);
  const int nq = (out_w + 7) / 8;
  std::vector<std::vector<int>> quarters(nq), buckets(8);
  std::vector<int> left;
  for (int c = 0; c < out_w; c++) buckets[first[c] & 7].push_back(c);
  for (auto& b : buckets)
    for (size_t i = 0; i < b.size(); i++) {
      if ((int)i < nq) quarters[i].push_back(b[i]); else left.push_back(b[i]);
    }
  for (auto& q : quarters) {
    while (!left.empty() && q.size() < 8) {
      size_t best = 0;
      int best_cnt = 1 << 30;
      for (size_t i = 0; i < left.size(); i++) {
        int cnt = 0;
        for (int x : q) cnt += (first[x] & 7) == (first[left[i]] & 7);
        if (cnt < best_cnt) { best_cnt = cnt; best = i; }
      }
      q.push_back(left[best]);
      left.erase(left.begin() + best);
    }
  }
  for (auto& q : quarters) {
    for (int c : q) perm.push_back(c);
    for (size_t i = q.size(); i < 8; i++) perm.push_back(out_w);  // a short quarter: idle lanes
  }
  if (!left.empty() || (int)perm.size() > threads) return ident;  // does not fit the CTA: keep it simple
  perm.resize(threads, out_w);
  return lds128_cost(first, perm, out_w) < lds128_cost(first, ident, out_w) ? perm : ident;
}

static int nw_class(int nw) {
  for (int c : kNwClasses)
    if (nw <= c) return c;
  return -1;
}

// Source-pixel windows of the IMMA variant: tile nt = output columns [8nt, 8nt + 8) reads the pixels
// [start[nt], start[nt] + 32 * nk), start a multiple of 16 (ldmatrix rows are 16-byte pieces).
struct MmaWindows {
  int ntiles = 0, nk = 0, wp = 0;
  std::vector<int32_t> start;
};

static MmaWindows mma_windows(const std::vector<int32_t>& bounds, int W, int out_w) {
  MmaWindows m;
  m.ntiles = (out_w + 7) / 8;
  m.start.resize(m.ntiles);
  int extent_max = 0, end_max = 0;
  for (int nt = 0; nt < m.ntiles; nt++) {
    const int c0 = nt * 8, c1 = std::min(nt * 8 + 7, out_w - 1);
    int lo = bounds[2 * c0], hi = 0;
    for (int cc = c0; cc <= c1; cc++) {
      lo = std::min(lo, bounds[2 * cc]);
      hi = std::max(hi, bounds[2 * cc] + bounds[2 * cc + 1]);
    }
    m.start[nt] = lo & ~15;
    extent_max = std::max(extent_max, hi - m.start[nt]);
  }
  m.nk = (extent_max + 31) / 32;
  for (int nt = 0; nt < m.ntiles; nt++) end_max = std::max(end_max, m.start[nt] + 32 * m.nk);
  int wp = (int)round_up((uint64_t)std::max(end_max, W), 16);
  while (wp % 128 != 16) wp += 16;  // six consecutive lines of one ldmatrix fall on distinct bank groups
  m.wp = wp;
  return m;
}

static bool fused_geometry(int c, int out_h, int out_w, int elem, const ImgDesc& d, bool allow_mma, FusedGeom* g) {
  if (d.p1_axis != AXIS_H || d.p2_axis != AXIS_V) return false;
  if (!((d.cin == 3 && c == 3) || (d.cin == 1 && c == 1) || (d.cin == 4 && c == 3) ||
        ((d.cin == 3 || d.cin == 4) && c == 1)))
    return false;
  if (out_w > kFusedMaxThreads) return false;
  std::vector<int32_t> b, k, packed, first;
  const int ksh = compute_axis_coeffs(d.W, out_w, b, k);
  if (!coeffs_never_clip(b, k, ksh, out_w)) return false;
  const int nw = pack_axis(b, k, ksh, out_w, 0, packed, first);
  const int cls = nw_class(nw);
  if (cls < 0) return false;
  const int ksv = compute_axis_coeffs(d.H, out_h, b, k);
  if (!coeffs_never_clip(b, k, ksv, out_h)) return false;
  const int nwv = pack_axis(b, k, ksv, out_h, 0, packed, first);
  if (nwv > 24) return false;
  g->nw_class = cls; g->nwv = nwv;
  g->wq = p4_phys(((d.W + 3) >> 2) + cls + 1) + 1;  // physical granules per plane
  g->stage_bytes = (int)round_up((uint64_t)4 * d.pitch + 48, 128);
  g->threads = out_w > 256 ? 512 : 256;
  g->smem = 64 + (size_t)g->stage_bytes + (size_t)c * g->wq * 16 + (size_t)(nwv + 1) * out_w * c * 4 + 64;
  g->mma_nk = 0;
  // IMMA variant: RGB -> RGB, every row 16-byte aligned, windows of at most three 32-pixel k-steps
  static const bool mma_on = !(getenv("TCB_FUSED_MMA") && atoi(getenv("TCB_FUSED_MMA")) == 0);
  if (mma_on && allow_mma && c == 3 && d.cin == 3 && d.pitch % 16 == 0 && reinterpret_cast<uintptr_t>(d.src) % 16 == 0 &&
      out_w <= 512) {
    std::vector<int32_t> bh, kh;
    compute_axis_coeffs(d.W, out_w, bh, kh);
    MmaWindows mw = mma_windows(bh, d.W, out_w);
    int32_t kmax = 0;
    for (int32_t kv : kh) kmax = std::max(kmax, kv);
    // the fragments carry 4 * k in three byte planes: every coefficient must stay below 2^22
    if (mw.nk >= 1 && mw.nk <= 3 && kmax < (1 << kPrecisionBits)) {
      const int stage = (int)round_up((uint64_t)8 * d.pitch + 64, 128);
      static const int tpw_env = getenv("TCB_MMA_TPW") ? atoi(getenv("TCB_MMA_TPW")) : 0;
      // measured on cfg3: two CTAs of 14 warps (TPW 2) 0.428 ms, one CTA of 28 warps (TPW 1) 0.543 ms - the
      // phases of one CTA do not overlap, so the single-CTA form is kept only as an experiment switch
      const int tpw = (tpw_env == 1 && mw.ntiles <= 28) ? 1 : 2;
      const int warps = (mw.ntiles + tpw - 1) / tpw;
      const size_t smem = 128 + 768 * (size_t)elem + round_up((uint64_t)out_h * 8, 128) + (size_t)(tpw == 1 ? 2 : 1) * stage +
                          (size_t)kMmaLines * mw.wp + (size_t)(nwv + 1) * out_w * 3 * 4;
      if (smem <= kFusedMaxSmem) {
        g->mma_nk = mw.nk; g->mma_wp = mw.wp;
        // class = 100 + 4 * k-steps + (one tile per warp: 2) + (more than 14 / 28 warps: 1)
        g->nw_class = 100 + 4 * mw.nk + (tpw == 1 ? 2 : 0) + (tpw == 2 && warps > 14 ? 1 : 0);
        g->stage_bytes = stage;
        g->threads = warps * 32;
        g->smem = smem;
      }
    }
  }
  return g->smem <= kFusedMaxSmem;
}

struct FusedClass {
  int threads = 256;
  size_t smem = 0;                  // dynamic shared memory of the launch (max over the class's images)
  std::vector<int> images;          // plan indices of the images this launch handles
  std::vector<int4> items;          // (image, first output row, end output row, -)
  std::vector<int32_t> cta_first;   // items of CTA b: [cta_first[b], cta_first[b + 1]), relative to items_off
  size_t items_off = 0, cta_off = 0;
};

struct FusedPlan {
  std::vector<FusedDesc> desc;      // one per plan image (valid where fused)
  std::map<int, FusedClass> cls;    // nw class -> launch
  FusedDesc* d_desc = nullptr;
  int4* d_items = nullptr;
  int32_t* d_cta = nullptr;
  int32_t* d_pool = nullptr;
};

static FusedPlan* fused_of(const Plan& p) { return static_cast<FusedPlan*>(p.fused); }

static const FusedGeom& cached_geometry(Plan& p, const ImgDesc& d) {
  // the IMMA variant also needs 16-byte aligned rows: the source alignment is part of the key
  const int al16 = reinterpret_cast<uintptr_t>(d.src) % 16 == 0 ? 256 : 0;
  const std::array<int, 4> key = {d.W, d.H, d.cin | al16, d.pitch};
  auto it = p.geom_cache.find(key);
  if (it == p.geom_cache.end()) {
    FusedGeom g;
    g.ok = fused_geometry(p.c, p.out_h, p.out_w, p.elem, d, !p.no_mma, &g);
    it = p.geom_cache.emplace(key, g).first;
  }
  return it->second;
}

bool fused_supported(Plan& p, const ImgDesc& d, int* band_rows, size_t* smem_bytes) {
  const FusedGeom& g = cached_geometry(p, d);
  if (!g.ok) return false;
  *band_rows = 0;  // the row spans are cut per launch, see partition_class()
  *smem_bytes = g.smem;
  return true;
}

// One template instance per (class, channels, CTA size); everything that needs the function
// pointer (attributes, occupancy, launch) goes through this dispatcher.
template <typename F>
static int with_kernel(int cls, int c, bool big, F&& f) {
#define TCB_K(NWv)                                                                         \
  case NWv:                                                                                \
    if (big) return c == 3 ? f(k_preprocess_fused<NWv, 3, true>) : f(k_preprocess_fused<NWv, 1, true>); \
    return c == 3 ? f(k_preprocess_fused<NWv, 3, false>) : f(k_preprocess_fused<NWv, 1, false>);
  switch (cls) {
    TCB_K(2) TCB_K(3) TCB_K(4) TCB_K(5) TCB_K(6) TCB_K(8) TCB_K(10) TCB_K(12) TCB_K(16)
    // IMMA variant: 100 + 4 * k-steps + (one tile per warp: 2) + (two tiles per warp, more than 14 warps: 1)
    case 104: return f(k_preprocess_fused_mma<1, 2, 14>);
    case 105: return f(k_preprocess_fused_mma<1, 2, 32>);
    case 106: return f(k_preprocess_fused_mma<1, 1, 28>);
    case 108: return f(k_preprocess_fused_mma<2, 2, 14>);
    case 109: return f(k_preprocess_fused_mma<2, 2, 32>);
    case 110: return f(k_preprocess_fused_mma<2, 1, 28>);
    case 112: return f(k_preprocess_fused_mma<3, 2, 14>);
    case 113: return f(k_preprocess_fused_mma<3, 2, 32>);
    case 114: return f(k_preprocess_fused_mma<3, 1, 28>);
    default: return set_error(TCB_ERR_INTERNAL, "bad coefficient class %d", cls);
  }
#undef TCB_K
}

// Cuts the (image, output row) space of one launch into `ctas` contiguous shares of equal
// estimated cost. A uniform batch such as cfg3 (256 x 1080p on 3 x 148 resident CTAs) used to run
// as 1024 equal bands = 2.31 waves, the last one a third full; now every resident CTA gets exactly
// one share and they all finish together. Cost model per output row of an image: its share of the
// source bytes plus kOutCost byte-equivalents per output element (the vertical pass, table and
// store; fitted to the per-resolution timings in DESIGN.md), which is what balances mixed batches.
static void partition_class(const Plan& p, const FusedPlan& fp, FusedClass& fc, int ctas) {
  constexpr double kOutCost = 14.0;
  constexpr int kMinRows = 4;
  const int out_h = p.out_h;
  std::vector<double> row_cost(fc.images.size());
  double total = 0;
  for (size_t k = 0; k < fc.images.size(); k++) {
    const FusedDesc& f = fp.desc[fc.images[k]];
    row_cost[k] = (double)f.H * f.W * f.cin / out_h + kOutCost * p.out_w * p.c;
    total += row_cost[k] * out_h;
  }
  const long long total_rows = (long long)fc.images.size() * out_h;
  ctas = (int)std::max<long long>(1, std::min<long long>(ctas, total_rows / 8));
  fc.items.clear();
  fc.cta_first.assign(1, 0);
  double cum = 0;
  size_t k = 0;
  int y = 0;
  for (int b = 0; b < ctas; b++) {
    const double target = b + 1 == ctas ? total * 2 : total * (b + 1) / ctas;
    const size_t first_item = fc.items.size();
    while (k < fc.images.size() && cum < target) {
      const int left = out_h - y;
      int take = (int)std::ceil((target - cum) / row_cost[k]);
      if (take < kMinRows && fc.items.size() > first_item && y == 0) break;  // not worth a halo: next CTA
      take = std::max(take, std::min(kMinRows, left));
      if (left - take < kMinRows) take = left;  // no sliver for the next CTA either
      fc.items.push_back(make_int4(fc.images[k], y, y + take, 0));
      cum += row_cost[k] * take;
      y += take;
      if (y == out_h) { y = 0; k++; }
    }
    fc.cta_first.push_back((int32_t)fc.items.size());
  }
}

int fused_prepare(Plan& p) {
  FusedPlan* fp = new FusedPlan();
  fp->desc.resize(p.n);
  std::vector<int32_t> pool;
  std::map<std::pair<int, int>, std::array<int32_t, 3>> htab;   // (W, class) -> hc_off, hw0_off, hperm_off
  std::map<int, std::array<int32_t, 5>> vtab;                    // H -> vc_off, vg_off, vb_off, nwv, vq_off
  auto align4 = [&]() { while (pool.size() % 4) pool.push_back(0); };
  for (int i = 0; i < p.n; i++) {
    const ImgDesc& d = p.h_desc[i];
    if (!d.fused) continue;
    const FusedGeom g = cached_geometry(p, d);
    FusedDesc& f = fp->desc[i];
    f.src = d.src; f.W = d.W; f.H = d.H; f.pitch = d.pitch; f.cin = d.cin;
    f.nw = g.nw_class; f.nwv = g.nwv; f.band_rows = 0; f.wq = g.wq;
    f.stage_bytes = g.stage_bytes;
    f.aligned4 = (reinterpret_cast<uintptr_t>(d.src) % 4 == 0) && (d.pitch % 4 == 0);
    f.mma_nk = g.mma_nk; f.mma_wp = g.mma_wp; f.mma_b_off = 0; f.mma_s_off = 0;
    auto hk = std::make_pair(d.W, g.nw_class);
    if (g.mma_nk > 0 && !htab.count(hk)) {
      // B fragments of mma.m16n8k32 (col-major B: lane = 4 * column + k/4, two words = k 0..15 / 16..31
      // quarter `lane & 3`), one set per (tile of 8 columns, k-step, coefficient byte plane)
      std::vector<int32_t> b, k;
      const int ks = compute_axis_coeffs(d.W, p.out_w, b, k);
      const MmaWindows mw = mma_windows(b, d.W, p.out_w);
      std::array<int32_t, 3> e;
      align4(); e[0] = (int32_t)pool.size();
      for (int nt = 0; nt < mw.ntiles; nt++)
        for (int kstep = 0; kstep < mw.nk; kstep++)
          for (int pl = 0; pl < 3; pl++)
            for (int ln = 0; ln < 32; ln++)
              for (int half = 0; half < 2; half++) {
                const int col = nt * 8 + (ln >> 2);
                uint32_t word = 0;
                if (col < p.out_w) {
                  const int xmin = b[2 * col], cnt = b[2 * col + 1];
                  for (int bi = 0; bi < 4; bi++) {
                    const int x = mw.start[nt] + 32 * kstep + 16 * half + 4 * (ln & 3) + bi;
                    const int t = x - xmin;
                    if (t >= 0 && t < cnt) {
                      const uint32_t kv = 4u * (uint32_t)k[(size_t)col * ks + t];  // 4k < 2^24: see the kernel's epilogue
                      word |= ((kv >> (8 * pl)) & 0xffu) << (8 * bi);
                    }
                  }
                }
                pool.push_back((int32_t)word);
              }
      align4(); e[1] = (int32_t)pool.size(); pool.insert(pool.end(), mw.start.begin(), mw.start.end());
      e[2] = 0;
      htab[hk] = e;
    }
    if (g.mma_nk > 0) { f.mma_b_off = htab[hk][0]; f.mma_s_off = htab[hk][1]; }
    if (!htab.count(hk)) {
      std::vector<int32_t> b, k, packed, first;
      const int ks = compute_axis_coeffs(d.W, p.out_w, b, k);
      pack_axis(b, k, ks, p.out_w, g.nw_class, packed, first);
      std::array<int32_t, 3> e;
      align4(); e[0] = (int32_t)pool.size(); pool.insert(pool.end(), packed.begin(), packed.end());
      align4(); e[1] = (int32_t)pool.size(); pool.insert(pool.end(), first.begin(), first.end());
      const std::vector<int32_t> cols = lane_columns(first, p.out_w, g.threads);
      align4(); e[2] = (int32_t)pool.size(); pool.insert(pool.end(), cols.begin(), cols.end());
      htab[hk] = e;
    }
    if (!vtab.count(d.H)) {
      std::vector<int32_t> b, k, packed, first, last;
      const int ks = compute_axis_coeffs(d.H, p.out_h, b, k);
      const int nwv = pack_axis(b, k, ks, p.out_h, 0, packed, first, &last);
      std::array<int32_t, 5> e;
      align4(); e[0] = (int32_t)pool.size();  // [out_h][nwv] x (k0, k1, k2, 0): one 128-bit load per tap word
      for (size_t t = 0; t + 2 < packed.size(); t += 3) {
        pool.push_back(packed[t]); pool.push_back(packed[t + 1]); pool.push_back(packed[t + 2]); pool.push_back(0);
      }
      align4(); e[1] = (int32_t)pool.size();
      for (int o = 0; o < p.out_h; o++) { pool.push_back(first[o]); pool.push_back(last[o]); }
      align4(); e[2] = (int32_t)pool.size(); pool.insert(pool.end(), b.begin(), b.end());
      e[3] = nwv;
      align4(); e[4] = (int32_t)pool.size();
      for (int st = 0, done = 0; st < (d.H + 7) / 8; st++) {
        while (done < p.out_h && last[done] <= 2 * st + 1) done++;
        pool.push_back(done);
      }
      vtab[d.H] = e;
    }
    f.hc_off = htab[hk][0]; f.hw0_off = htab[hk][1]; f.hperm_off = htab[hk][2];
    f.vc_off = vtab[d.H][0]; f.vg_off = vtab[d.H][1]; f.vb_off = vtab[d.H][2]; f.vq_off = vtab[d.H][4];
    FusedClass& fc = fp->cls[g.nw_class];
    fc.images.push_back(i);
    fc.smem = std::max(fc.smem, g.smem);
    fc.threads = g.threads;
  }
  // resident CTAs per launch, then the cost-balanced cut
  static const int per_sm_cap = getenv("TCB_FUSED_CTAS_PER_SM") ? atoi(getenv("TCB_FUSED_CTAS_PER_SM")) : 0;
  const int sms = sm_count(p.device);
  std::vector<int4> allitems;
  std::vector<int32_t> allcta;
  for (auto& kv : fp->cls) {
    FusedClass& fc = kv.second;
    int per_sm = 0;
    const int rc = with_kernel(kv.first, p.c, fc.threads > 256, [&](auto kern) -> int {
      TCB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fc.smem));
      TCB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, fc.threads, fc.smem));
      return TCB_OK;
    });
    if (rc) { delete fp; return rc; }
    per_sm = std::max(per_sm, 1);
    if (getenv("TCB_DEBUG"))
      fprintf(stderr, "[tcb] fused class %d: %zu images, %d threads, %zu B smem, %d CTAs/SM resident\n", kv.first,
              fc.images.size(), fc.threads, fc.smem, per_sm);
    if (per_sm_cap > 0) per_sm = per_sm_cap;
    partition_class(p, *fp, fc, per_sm * sms);
    fc.items_off = allitems.size();
    fc.cta_off = allcta.size();
    allitems.insert(allitems.end(), fc.items.begin(), fc.items.end());
    allcta.insert(allcta.end(), fc.cta_first.begin(), fc.cta_first.end());
  }
  if (pool.empty()) pool.push_back(0);
  cudaError_t e = cudaMalloc((void**)&fp->d_desc, sizeof(FusedDesc) * p.n);
  if (e == cudaSuccess) e = cudaMemcpy(fp->d_desc, fp->desc.data(), sizeof(FusedDesc) * p.n, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMalloc((void**)&fp->d_items, allitems.size() * sizeof(int4));
  if (e == cudaSuccess) e = cudaMemcpy(fp->d_items, allitems.data(), allitems.size() * sizeof(int4), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMalloc((void**)&fp->d_cta, allcta.size() * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMemcpy(fp->d_cta, allcta.data(), allcta.size() * sizeof(int32_t), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMalloc((void**)&fp->d_pool, pool.size() * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMemcpy(fp->d_pool, pool.data(), pool.size() * sizeof(int32_t), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    cudaFree(fp->d_desc); cudaFree(fp->d_items); cudaFree(fp->d_cta); cudaFree(fp->d_pool);
    delete fp;
    return set_error(TCB_ERR_NOMEM, "fused preprocess set-up failed: %s", cudaGetErrorString(e));
  }
  p.fused = fp;
  return TCB_OK;
}

void fused_release(Plan& p) {
  FusedPlan* fp = fused_of(p);
  if (!fp) return;
  cudaFree(fp->d_desc); cudaFree(fp->d_items); cudaFree(fp->d_cta); cudaFree(fp->d_pool);
  delete fp;
  p.fused = nullptr;
}

// A run may hand the plan new source pointers / pitches. The launch geometry (class, shared
// memory) was fixed at creation, so the new pitch must still fit the class's staging area; the
// row-alignment flag and the staging size are per image and simply follow.
bool fused_accepts(Plan& p, int i, const uint8_t* src, int pitch) {
  const FusedPlan* fp = fused_of(p);
  if (!fp) return false;
  ImgDesc t = p.h_desc[i];
  t.src = src; t.pitch = pitch;
  const FusedGeom g = cached_geometry(p, t);
  if (!g.ok) return false;
  const auto it = fp->cls.find(g.nw_class);
  if (getenv("TCB_DEBUG"))
    fprintf(stderr, "[tcb] fused_accepts image %d: class %d (plan %d), smem %zu (launch %zu)\n", i, g.nw_class,
            fp->desc[i].nw, g.smem, it != fp->cls.end() ? it->second.smem : (size_t)0);
  return it != fp->cls.end() && g.nw_class == fp->desc[i].nw && g.smem <= it->second.smem;
}

int fused_update_sources(const Plan& p, cudaStream_t s) {
  FusedPlan* fp = fused_of(p);
  if (!fp) return TCB_OK;
  for (int i = 0; i < p.n; i++)
    if (p.h_desc[i].fused) {
      FusedDesc& f = fp->desc[i];
      f.src = p.h_desc[i].src;
      f.pitch = p.h_desc[i].pitch;
      // follows the pitch (fused_accepts checked the fit)
      f.stage_bytes = f.mma_nk > 0 ? (int)round_up((uint64_t)8 * f.pitch + 64, 128) : (int)round_up((uint64_t)4 * f.pitch + 48, 128);
      f.aligned4 = (reinterpret_cast<uintptr_t>(f.src) % 4 == 0) && (f.pitch % 4 == 0);  // may change from run to run
    }
  TCB_CUDA(cudaMemcpyAsync(fp->d_desc, fp->desc.data(), sizeof(FusedDesc) * p.n, cudaMemcpyHostToDevice, s));
  return TCB_OK;
}

int fused_launch(const Plan& p, void* dst, cudaStream_t s) {
  const FusedPlan* fp = fused_of(p);
  if (!fp) return set_error(TCB_ERR_INTERNAL, "fused plan missing");
  static const int pf = (getenv("TCB_FUSED_L2PF") ? atoi(getenv("TCB_FUSED_L2PF")) : 1) << 1;  // L2 prefetch distance (IMMA kernel)
  const int al = ((reinterpret_cast<uintptr_t>(dst) % 16 == 0) ? 1 : 0) | pf;
  for (auto& kv : fp->cls) {
    const FusedClass& fc = kv.second;
    const int ctas = (int)fc.cta_first.size() - 1;
    const int rc = with_kernel(kv.first, p.c, fc.threads > 256, [&](auto kern) -> int {
      kern<<<ctas, fc.threads, fc.smem, s>>>(fp->d_desc, fp->d_items + fc.items_off, fp->d_cta + fc.cta_off,
                                            fp->d_pool, p.d_lut, (uint8_t*)dst, p.out_h, p.out_w, p.layout,
                                            p.elem, al);
      return TCB_OK;
    });
    if (rc) return rc;
    count_launch();
  }
  return TCB_OK;
}

}  // namespace tcb
