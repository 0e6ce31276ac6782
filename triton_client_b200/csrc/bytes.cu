// BYTES tensor (de)serialisation kernels behind the C ABI (include/tcb200.h).
//
// Wire format (/root/reference/src/python/library/tritonclient/utils/__init__.py:193-276): per
// element, in C order, a little-endian uint32 length followed by the payload bytes.
//
// Serialise  (replaces the np.nditer + struct.pack + b"".join loop, utils/__init__.py:229-242):
//   one kernel; per tile of 1024 elements: lengths -> warp-shuffle block scan -> decoupled
//   look-back for the tile's global payload offset -> coalesced scatter of prefixes + payload.
//   Output-centric: every thread builds 16 aligned destination bytes at a time. Because element i
//   starts at src_off[i] + 4*i, payload bytes move by a multiple of 4, so interior words are
//   plain aligned 32-bit loads.
// Deserialise (replaces the unpack_from loop, utils/__init__.py:270-275 and
//   cuda_shared_memory/__init__.py:312-318): the pointer chase is cut into 8 KiB chunks. Each
//   chunk guesses its entry header by chasing from every byte of a 512-byte look-behind window
//   and voting, then chases its own chain; a verify pass checks entry[k+1] == exit[k] for every
//   chunk, which makes the result exact; any mismatch (elements longer than the window,
//   adversarial payloads) is repaired by a sequential sweep. A scan of per-chunk counts gives
//   element indices; payloads are then gathered with the same output-centric copy.
#include <cstdlib>
#include <utility>
#include <vector>

#include "common.cuh"

namespace tcb {

// =================================================================================================
// small device helpers
// =================================================================================================
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint32_t ldg_u32(const uint8_t* p) {
  return __ldg(reinterpret_cast<const uint32_t*>(p));
}
__device__ __forceinline__ void st_v4(uint8_t* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d)
               : "memory");
}

// status word of the look-back scan: flag in bits 63..62, value in bits 61..0
constexpr unsigned long long kFlagAgg = 1ull << 62;
constexpr unsigned long long kFlagInc = 2ull << 62;
constexpr unsigned long long kValMask = (1ull << 62) - 1;
constexpr unsigned int kSpinLimit = 1u << 24;  // bounded: a lost tile becomes an error, not a hang

// workspace header (in 64-bit words)
enum : int { WS_TICKET = 0, WS_ERROR = 1, WS_TOTAL = 2, WS_HEADER_WORDS = 8 };
enum : unsigned long long { ERR_NONE = 0, ERR_SPIN = 1, ERR_CAPACITY = 2 };

// =================================================================================================
// serialise
// =================================================================================================
constexpr int kSerThreads = 256;
constexpr int kSerItems = 4;
constexpr int kSerTile = kSerThreads * kSerItems;

// Exclusive look-back over tile aggregates (single-pass scan, Merrill & Garland style). Called by
// warp 0 of the CTA that owns `tile`; tiles are handed out by an atomic ticket so every
// predecessor is already running. Returns the exclusive prefix in all lanes.
__device__ unsigned long long lookback_exclusive(unsigned long long* status, unsigned int tile,
                                                 unsigned long long aggregate,
                                                 unsigned long long* ws) {
  const int lane = threadIdx.x & 31;
  if (lane == 0) st_relaxed_u64(&status[tile], (tile == 0 ? kFlagInc : kFlagAgg) | aggregate);
  unsigned long long exclusive = 0;
  if (tile == 0) return 0;
  long long look = (long long)tile - 1;
  while (true) {
    const long long idx = look - lane;
    unsigned long long st = kFlagInc;  // virtual predecessors of tile 0: inclusive 0
    if (idx >= 0) {
      unsigned int spins = 0;
      st = ld_relaxed_u64(&status[idx]);
      while ((st >> 62) == 0) {
        if (++spins > kSpinLimit) {
          atomicMax(&ws[WS_ERROR], (unsigned long long)ERR_SPIN);
          st = kFlagInc;
          break;
        }
        __nanosleep(20);
        st = ld_relaxed_u64(&status[idx]);
      }
    }
    const unsigned inc_mask = __ballot_sync(0xffffffffu, (st >> 62) == 2);
    unsigned long long v = st & kValMask;
    if (inc_mask) {
      const int first = __ffs(inc_mask) - 1;  // nearest predecessor that knows its inclusive prefix
      if (lane > first) v = 0;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    exclusive += v;
    if (inc_mask) break;
    look -= 32;
  }
  if (lane == 0) st_relaxed_u64(&status[tile], kFlagInc | (exclusive + aggregate));
  return exclusive;
}

// Copies the payload of the tile's elements between "blob space" (payloads back to back) and
// "stream space" (each element preceded by its 4-byte length). TO_STREAM = serialise direction.
//   s_off[e]  : blob offset of tile element e relative to the tile's first element (e in 0..nE)
//   blob0     : global blob offset of the tile's first element
//   e0        : global index of the tile's first element
//   s_tab     : kTabSize entries of scratch (element containing every 2^tsh-th output byte)
// Output-centric, 16 aligned output bytes at a time, in two divergence-free phases:
//   A. interior chunks (entirely inside one payload): element found through s_tab + a short walk,
//      4 aligned word loads (payload moves by a multiple of 4 bytes), one 128-bit store;
//   B. boundary chunks (contain a length prefix / an element boundary / a tile edge): each is owned
//      by exactly one element-thread, which assembles it byte-wise; all lanes of a warp are busy
//      with boundary chunks at the same time instead of dragging the slow path through phase A.
constexpr int kTabSize = 1024;

template <bool TO_STREAM>
__device__ __forceinline__ unsigned long long elem_start(const unsigned long long* s_off, int e) {
  return TO_STREAM ? s_off[e] + 4ull * (unsigned)e : s_off[e];  // relative to the tile's output start
}

__device__ __forceinline__ uint32_t bytemask(int n) {  // low n bytes set, n clamped to [0, 4]
  return n <= 0 ? 0u : (n >= 4 ? 0xffffffffu : ((1u << (8 * n)) - 1u));
}
__device__ __forceinline__ uint32_t shl_bytes(uint32_t x, int n) { return n >= 4 ? 0u : x << (8 * n); }
__device__ __forceinline__ uint32_t shr_bytes(uint32_t x, int n) { return n >= 4 ? 0u : x >> (8 * n); }
__device__ __forceinline__ int clamp_u64_to(unsigned long long v, int hi) {
  return v > (unsigned long long)hi ? hi : (int)v;
}
// Word whose lanes selected by `mask` are needed. Aligned case: one 32-bit load (never crosses a
// page). Misaligned case: only the needed bytes are touched.
__device__ __forceinline__ uint32_t load_word(const uint8_t* p, bool word_ok, uint32_t mask = 0xffffffffu,
                                              bool in_smem = false) {
  if (in_smem) {  // p points into this thread's shared-memory slot (always word aligned)
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(static_cast<uint32_t>(__cvta_generic_to_shared(p))));
    return v;
  }
  if (word_ok) return ldg_u32(p);
  uint32_t v = 0;
#pragma unroll
  for (int i = 0; i < 4; i++)
    if (mask & (0xffu << (8 * i))) v |= (uint32_t)p[i] << (8 * i);
  return v;
}

// One aligned 32-bit word of the tile's output at tile-relative offset `rel` (rel % 4 == 0 in
// absolute terms), composed arithmetically instead of byte by byte. `e` must be the element that
// holds byte `rel` and is advanced by the caller. `sp` points at the source byte that corresponds
// to output byte `rel` for an element with global index 0; element E's bytes are 4*(E+1) before
// (serialise) or after (deserialise) that.
//   serialise:   word = [tail of prefix e][payload of e][head of prefix e+1]   (at most these three)
//   deserialise: word = payload bytes of e, e+1, ... (elements may be shorter than a word)
template <bool TO_STREAM, bool WORD_OK>
__device__ __forceinline__ uint32_t compose_word(const unsigned long long* s_off, int nE, int e,
                                                 unsigned long long rel, unsigned long long e0,
                                                 const uint8_t* sp, bool in_smem = false) {
  if (TO_STREAM) {
    const unsigned long long st = elem_start<true>(s_off, e);
    const unsigned long long len = s_off[e + 1] - s_off[e];
    const int r = clamp_u64_to(rel - st, 8);                 // offset of the word inside element e's footprint
    const int n1 = 4 - r;                                    // lanes [0, n1) hold prefix bytes of e (if r < 4)
    const int n2 = clamp_u64_to(st + 4 + len - rel, 4);      // lanes [n1, n2) hold payload of e
    uint32_t w = 0;
    if (r < 4) w = shr_bytes((uint32_t)len, r) & bytemask(n1);
    const uint32_t pm = bytemask(n2) & ~bytemask(n1);
    if (pm) w |= load_word(sp - 4ull * (e0 + e + 1), WORD_OK, pm, in_smem) & pm;
    if (n2 < 4 && e + 1 < nE) w |= shl_bytes((uint32_t)(s_off[e + 2] - s_off[e + 1]), n2);
    return w;
  } else {
    uint32_t w = 0;
    int lane = 0;  // next byte lane to fill
    while (lane < 4 && e < nE) {
      const unsigned long long en = s_off[e + 1];
      const int n = clamp_u64_to(en - rel, 4);  // lanes [lane, n) come from element e
      if (n > lane) {
        const uint32_t m = bytemask(n) & ~bytemask(lane);
        w |= load_word(sp + 4ull * (e0 + e + 1), WORD_OK, m, in_smem) & m;
        lane = n;
      }
      if (lane < 4) e++;
    }
    return w;
  }
}

// Builds one boundary chunk (16 aligned output bytes, of which [A, lim) belong to this tile) and
// stores it. `e` is an element at or before the one that holds the chunk's first in-range byte.
// The source bytes a boundary chunk needs sit in one or two 32-byte sectors; they are fetched once,
// as up to three aligned 16-byte loads into a per-thread shared-memory slot, and the words are then
// composed from the slot. (Composing straight from global memory issued ~10 separate sector
// accesses per chunk and tripled the L2 -> L1 traffic of the whole kernel.)
constexpr int kSlotBytes = 48;

template <bool TO_STREAM, bool WORD_OK>
__device__ __forceinline__ void assemble_chunk(const unsigned long long* s_off, int nE, unsigned long long A,
                                               unsigned long long lim, unsigned long long e0, int e,
                                               unsigned long long q, unsigned long long mis,
                                               const uint8_t* __restrict__ src, uint8_t* __restrict__ vd,
                                               uint8_t* slot) {
  const unsigned long long cb = q << 4;
  const unsigned long long a = max(cb, A), b = min(cb + 16, lim);
  if (a >= b) return;
  const bool full = (a == cb) && (b == cb + 16);
  // element of the first in-range byte, and of the last one
  while (e + 1 < nE && elem_start<TO_STREAM>(s_off, e + 1) <= a - A) e++;
  int eb = e;
  while (eb + 1 < nE && elem_start<TO_STREAM>(s_off, eb + 1) <= b - 1 - A) eb++;
  // source offsets (relative to src) that can be touched: [lo_off, hi_off)
  const uint8_t* sb = src;  // sb[x] = source byte at offset x
  bool staged = false;
  if (WORD_OK) {
    // exact range of source bytes this chunk needs, widened to whole aligned words (compose_word
    // loads aligned words); every 16-byte block fetched below therefore holds a needed byte
    long long lo_off, hi_off;
    if (TO_STREAM) {
      const unsigned long long blob0 = A - mis - 4ull * e0;  // blob offset of the tile's first element
      const unsigned long long ra = a - A, rb = b - 1 - A;
      lo_off = ra < elem_start<true>(s_off, e) + 4 ? (long long)(blob0 + s_off[e])
                                                    : (long long)(a - mis) - 4ll * (long long)(e0 + e + 1);
      hi_off = rb < elem_start<true>(s_off, eb) + 4 ? (long long)(blob0 + s_off[eb])
                                                     : (long long)(b - mis) - 4ll * (long long)(e0 + eb + 1);
    } else {
      lo_off = (long long)(a - mis) + 4ll * (long long)(e0 + e + 1);
      hi_off = (long long)(b - mis) + 4ll * (long long)(e0 + eb + 1);
    }
    lo_off &= ~3ll;
    hi_off = (hi_off + 3) & ~3ll;
    if (lo_off < 0) lo_off = 0;
    const long long base = lo_off & ~15ll;
    if (hi_off > lo_off && hi_off - base <= kSlotBytes &&
        ((reinterpret_cast<uintptr_t>(src) & 15) == 0)) {
      uint4* s4 = reinterpret_cast<uint4*>(slot);
#pragma unroll
      for (int k = 0; k < kSlotBytes / 16; k++)
        if (base + 16 * k < hi_off) s4[k] = __ldg(reinterpret_cast<const uint4*>(src + base) + k);
      sb = slot - base;  // only this thread touches its slot: no barrier needed
      staged = true;
    }
  }
  uint32_t w[4];
#pragma unroll
  for (int wi = 0; wi < 4; wi++) {
    const unsigned long long wa = cb + 4 * wi;
    w[wi] = 0;
    if (wa + 4 <= a || wa >= b) continue;
    // first in-range byte of this word decides the element; for a word that starts before A (head
    // of the tile) that byte is A itself
    const unsigned long long rel = wa >= A ? wa - A : 0;
    while (e + 1 < nE && elem_start<TO_STREAM>(s_off, e + 1) <= rel) e++;
    if (wa >= A) {
      w[wi] = compose_word<TO_STREAM, WORD_OK>(s_off, nE, e, rel, e0, sb + (wa - mis), staged);
    } else {
      // word straddles the tile start: only lanes >= A - wa are ours; compose them byte-wise
      uint32_t v = 0;
      int el = e;
      for (int bi = (int)(A - wa); bi < 4; bi++) {
        const unsigned long long rb = wa + bi - A;
        while (el + 1 < nE && elem_start<TO_STREAM>(s_off, el + 1) <= rb) el++;
        const unsigned long long st = elem_start<TO_STREAM>(s_off, el);
        uint32_t byte;
        if (TO_STREAM && rb < st + 4) byte = (uint32_t)((s_off[el + 1] - s_off[el]) >> (8 * (rb - st))) & 0xffu;
        else byte = src[TO_STREAM ? wa + bi - mis - 4ull * (e0 + el + 1) : wa + bi - mis + 4ull * (e0 + el + 1)];
        v |= byte << (8 * bi);
      }
      w[wi] = v;
    }
  }
  if (full) {
    st_v4(vd + cb, w[0], w[1], w[2], w[3]);
  } else {
#pragma unroll
    for (int bi = 0; bi < 16; bi++) {
      const unsigned long long p = cb + bi;
      if (p >= a && p < b) vd[p] = (uint8_t)(w[bi >> 2] >> (8 * (bi & 3)));
    }
  }
}

// Fast paths of phase B (word-aligned copies only). The common boundary chunk holds exactly one
// boundary: the tail of element e-1, (serialise: the length prefix of e,) the head of element e,
// and both payloads reach the chunk's edges. Its 16 bytes then come from five consecutive aligned
// source words (the two elements' source positions differ by exactly one word) merged with byte
// masks: no search, no per-byte work. Returns false when the chunk is not of that shape.
template <bool WORD_OK>
__device__ __forceinline__ bool fast_prefix_chunk(const unsigned long long* s_off, int nE, int e,
                                                  unsigned long long hs, unsigned long long cb,
                                                  unsigned long long A, unsigned long long lim,
                                                  unsigned long long e0, unsigned long long mis,
                                                  const uint8_t* __restrict__ src, uint8_t* __restrict__ vd) {
  if (!WORD_OK || e == 0 || cb < A || cb + 16 > lim) return false;
  if (hs > cb && A + elem_start<true>(s_off, e - 1) + 4 > cb) return false;   // e-1's payload must cover [cb, hs)
  if (A + elem_start<true>(s_off, e + 1) < cb + 16) return false;             // e's payload must reach the chunk end
  const uint32_t len = (uint32_t)(s_off[e + 1] - s_off[e]);
  const int k = (int)(long long)(hs - cb);                                    // prefix position in the chunk, -3..15
  const uint8_t* B = src + (cb - mis) - 4ull * (e0 + e + 1);                  // source of chunk byte 0 as payload of e
  uint32_t P[5];
#pragma unroll
  for (int j = 0; j < 5; j++) {
    // word j is payload of e for output word j, and payload of e-1 for output word j-1
    const bool need = (j < 4 && 4 * j + 3 >= k + 4) || (j >= 1 && 4 * (j - 1) < k);
    P[j] = need ? ldg_u32(B + 4 * j) : 0u;
  }
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const int t = k - 4 * i;  // prefix start relative to this word
    const uint32_t mA = bytemask(t), mB = bytemask(t + 4);
    const uint32_t ls = t >= 0 ? shl_bytes(len, t) : shr_bytes(len, -t);
    w[i] = (P[i + 1] & mA) | (ls & mB & ~mA) | (P[i] & ~mB);
  }
  st_v4(vd + cb, w[0], w[1], w[2], w[3]);
  return true;
}

template <bool WORD_OK>
__device__ __forceinline__ bool fast_boundary_chunk(const unsigned long long* s_off, int nE, int e,
                                                    unsigned long long be, unsigned long long cb,
                                                    unsigned long long A, unsigned long long lim,
                                                    unsigned long long e0, unsigned long long mis,
                                                    const uint8_t* __restrict__ src, uint8_t* __restrict__ vd) {
  if (!WORD_OK || e == 0 || e >= nE || cb < A || cb + 16 > lim) return false;
  if (A + s_off[e - 1] > cb || A + s_off[e + 1] < cb + 16) return false;  // exactly one boundary inside the chunk
  const int k = (int)(be - cb);                                           // 1..15
  const uint8_t* S = src + (cb - mis) + 4ull * (e0 + e);                  // source of chunk byte 0 as payload of e-1
  uint32_t W[5];
#pragma unroll
  for (int j = 0; j < 5; j++) W[j] = ldg_u32(S + 4 * j);  // [S, S+20) lies between two payload bytes that exist
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const uint32_t m = bytemask(k - 4 * i);
    w[i] = (W[i] & m) | (W[i + 1] & ~m);
  }
  st_v4(vd + cb, w[0], w[1], w[2], w[3]);
  return true;
}

template <bool TO_STREAM, bool WORD_OK>
__device__ __forceinline__ void move_tile(const unsigned long long* s_off, int nE,
                                          unsigned long long blob0, unsigned long long e0,
                                          const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                          unsigned long long dst_capacity, unsigned short* s_tab, uint8_t* slot) {
  // output coordinates are shifted by `mis` so that multiples of 16 are 16-byte aligned addresses
  const unsigned long long mis = reinterpret_cast<uintptr_t>(dst) & 15;
  uint8_t* vd = dst - mis;
  const unsigned long long o0 = TO_STREAM ? blob0 + 4 * e0 : blob0;
  const unsigned long long total = elem_start<TO_STREAM>(s_off, nE);  // tile output bytes
  const unsigned long long o1 = o0 + total;
  const unsigned long long lim = min(o1, dst_capacity) + mis;
  const unsigned long long A = o0 + mis;
  if (lim <= A) return;
  // ---- lookup table: s_tab[t] = element containing output byte (t << tsh) of the tile ------------
  int tsh = 8;
  while ((total >> tsh) >= (unsigned long long)kTabSize) tsh++;
  for (int e = threadIdx.x; e < nE; e += blockDim.x) {
    const unsigned long long s0 = elem_start<TO_STREAM>(s_off, e), s1 = elem_start<TO_STREAM>(s_off, e + 1);
    if (s1 > s0) {
      const unsigned long long t0 = (s0 + (1ull << tsh) - 1) >> tsh, t1 = (s1 - 1) >> tsh;
      for (unsigned long long t = t0; t <= t1; t++) s_tab[t] = (unsigned short)e;
    }
  }
  __syncthreads();
  // ---- phase A: interior chunks, two per thread in flight ------------------------------------------------
  const unsigned long long q0 = (A + 15) >> 4, q1 = lim >> 4;  // chunks fully inside [A, lim)
  auto locate = [&](unsigned long long q, unsigned long long qlim, const uint8_t*& sp) -> bool {  // interior? -> source pointer
    if (q >= qlim) return false;
    const unsigned long long cb = q << 4;
    const unsigned long long rel = cb - A;
    int e = s_tab[rel >> tsh];
    unsigned long long nx = elem_start<TO_STREAM>(s_off, e + 1);
    while (nx <= rel) {  // short walk inside one table block
      e++;
      nx = elem_start<TO_STREAM>(s_off, e + 1);
    }
    if (rel + 16 > nx) return false;
    if (TO_STREAM && rel < elem_start<TO_STREAM>(s_off, e) + 4) return false;
    const unsigned long long E1 = 4ull * (e0 + e + 1);
    sp = src + (TO_STREAM ? cb - mis - E1 : cb - mis + E1);
    return true;
  };
  constexpr int kFly = 4;  // chunks in flight per thread: all loads are issued before the first store
  // The tile is walked in sub-tiles of blockDim.x elements, phase A then phase B for each: phase B
  // re-reads sectors that phase A touched, and with both phases run over the whole 1024-element
  // tile (~136 KB in, ~136 KB out, x 592 resident CTAs > L2) those re-reads all missed L2.
  for (int sub = 0; sub <= nE; sub += blockDim.x) {
  const int sub_end = min(sub + (int)blockDim.x, nE);
  const unsigned long long qs0 = sub == 0 ? q0 : max(q0, (A + elem_start<TO_STREAM>(s_off, sub) + 15) >> 4);
  const unsigned long long qs1 = sub_end == nE ? q1 : min(q1, (A + elem_start<TO_STREAM>(s_off, sub_end)) >> 4);
  for (unsigned long long q = qs0 + threadIdx.x; q < qs1; q += (unsigned long long)kFly * blockDim.x) {
    const uint8_t* sp[kFly];
    bool ok[kFly];
    uint32_t w[kFly][4];
#pragma unroll
    for (int f = 0; f < kFly; f++) {
      sp[f] = src;
      ok[f] = locate(q + (unsigned long long)f * blockDim.x, qs1, sp[f]);
    }
#pragma unroll
    for (int f = 0; f < kFly; f++) {
      if (ok[f]) {
#pragma unroll
        for (int k = 0; k < 4; k++) w[f][k] = load_word(sp[f] + 4 * k, WORD_OK);
      }
    }
#pragma unroll
    for (int f = 0; f < kFly; f++)
      if (ok[f]) st_v4(vd + ((q + (unsigned long long)f * blockDim.x) << 4), w[f][0], w[f][1], w[f][2], w[f][3]);
  }
  // ---- phase B: boundary chunks, one owner thread each ------------------------------------------------
  // Boundary points b_e: for TO_STREAM the prefix bytes [hs_e, hs_e + 4) of every element; otherwise
  // the element start offsets. Chunk q is owned by the first element whose boundary touches it; the
  // tile's head (A) and tail (lim) are the boundaries of pseudo-elements 0 and nE.
  {
    const int e = sub + threadIdx.x;
    if (e > nE) continue;
    if (TO_STREAM) {
      if (e < nE) {
        const unsigned long long hs = A + elem_start<true>(s_off, e);
        const unsigned long long qa = hs >> 4, qb = (hs + 3) >> 4;
        bool own_a = true;
        if (e > 0) {
          const unsigned long long hp = A + elem_start<true>(s_off, e - 1);
          own_a = ((hp + 3) >> 4) != qa;  // the previous prefix does not reach into qa
        }
        if (own_a && !fast_prefix_chunk<WORD_OK>(s_off, nE, e, hs, qa << 4, A, lim, e0, mis, src, vd))
          assemble_chunk<true, WORD_OK>(s_off, nE, A, lim, e0, e > 0 ? e - 1 : 0, qa, mis, src, vd, slot);
        if (qb != qa && !fast_prefix_chunk<WORD_OK>(s_off, nE, e, hs, qb << 4, A, lim, e0, mis, src, vd))
          assemble_chunk<true, WORD_OK>(s_off, nE, A, lim, e0, e, qb, mis, src, vd, slot);
      } else if (lim & 15) {  // partial tail chunk with no prefix of this tile in it
        const unsigned long long qt = lim >> 4;
        const unsigned long long hl = A + elem_start<true>(s_off, nE - 1);
        if (((hl + 3) >> 4) < qt) assemble_chunk<true, WORD_OK>(s_off, nE, A, lim, e0, nE - 1, qt, mis, src, vd, slot);
      }
    } else {
      // boundary point of pseudo-element e: its start (e < nE) or the tile end (e == nE)
      const unsigned long long be = e < nE ? A + s_off[e] : lim;
      if ((be & 15) == 0 || be > lim) continue;  // chunk-aligned boundaries create no boundary chunk
      const unsigned long long q = be >> 4;
      bool own = true;
      if (e > 0) {
        const unsigned long long bp = A + s_off[e - 1];
        own = (bp >> 4) < q || (bp & 15) == 0;  // previous boundary is not strictly inside chunk q
        if ((bp >> 4) == q && (bp & 15) == 0) own = true;
      }
      if (own && !fast_boundary_chunk<WORD_OK>(s_off, nE, e, be, q << 4, A, lim, e0, mis, src, vd))
        assemble_chunk<false, WORD_OK>(s_off, nE, A, lim, e0, e > 0 ? e - 1 : 0, q, mis, src, vd, slot);
    }
  }
  }  // sub-tiles
}

constexpr unsigned long long kTile32Limit = (1ull << 31) - 64;  // tile output bytes the 32-bit variant handles

// ---- 32-bit variant of move_tile ---------------------------------------------------------------
// Same algorithm, but every position is a 32-bit offset from the 16-byte aligned start T of the
// tile's output, element starts come from a precomputed table s_st[] (serialise: start of the
// length prefix, deserialise: start of the payload) and the source is addressed as srcT + offset.
// The 64-bit version spent most of its instructions on 64-bit address arithmetic and on
// recomputing s_off[e] + 4*e. Caller guarantees that the tile's output is shorter than 2^31 bytes.
template <bool WORD_OK>
__device__ __forceinline__ bool fast_prefix_chunk32(const uint32_t* s_st, int e, uint32_t hs, uint32_t cb,
                                                    uint32_t a, uint32_t lm, const uint8_t* __restrict__ srcT,
                                                    uint8_t* __restrict__ vdT) {
  if (!WORD_OK || e == 0 || cb < a || cb + 16 > lm) return false;
  if (hs > cb && s_st[e - 1] + 4 > cb) return false;  // e-1's payload must cover [cb, hs)
  const uint32_t nx = s_st[e + 1];
  if (nx < cb + 16) return false;                     // e's payload must reach the chunk end
  const uint32_t len = nx - hs - 4;
  const int k = (int)hs - (int)cb;                    // prefix position in the chunk, -3..15
  const uint8_t* B = srcT + ((long long)(int)cb - 4ll * (e + 1));  // source of chunk byte 0 as payload of e
  uint32_t P[5];
#pragma unroll
  for (int j = 0; j < 5; j++) {
    const bool need = (j < 4 && 4 * j + 3 >= k + 4) || (j >= 1 && 4 * (j - 1) < k);
    P[j] = need ? ldg_u32(B + 4 * j) : 0u;
  }
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const int t = k - 4 * i;
    const uint32_t mA = bytemask(t), mB = bytemask(t + 4);
    const uint32_t ls = t >= 0 ? shl_bytes(len, t) : shr_bytes(len, -t);
    w[i] = (P[i + 1] & mA) | (ls & mB & ~mA) | (P[i] & ~mB);
  }
  st_v4(vdT + cb, w[0], w[1], w[2], w[3]);
  return true;
}

template <bool WORD_OK>
__device__ __forceinline__ bool fast_boundary_chunk32(const uint32_t* s_st, int nE, int e, uint32_t be, uint32_t cb,
                                                      uint32_t a, uint32_t lm, const uint8_t* __restrict__ srcT,
                                                      uint8_t* __restrict__ vdT) {
  if (!WORD_OK || e == 0 || e >= nE || cb < a || cb + 16 > lm) return false;
  if (s_st[e - 1] > cb || s_st[e + 1] < cb + 16) return false;  // exactly one boundary inside the chunk
  const int k = (int)(be - cb);                                 // 1..15
  const uint8_t* S = srcT + (cb + 4u * (uint32_t)e);            // source of chunk byte 0 as payload of e-1
  uint32_t W[5];
#pragma unroll
  for (int j = 0; j < 5; j++) W[j] = ldg_u32(S + 4 * j);
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const uint32_t m = bytemask(k - 4 * i);
    w[i] = (W[i] & m) | (W[i + 1] & ~m);
  }
  st_v4(vdT + cb, w[0], w[1], w[2], w[3]);
  return true;
}

template <bool TO_STREAM, bool WORD_OK>
__device__ __forceinline__ void move_tile32(const unsigned long long* s_off, uint32_t* s_st, int nE,
                                            unsigned long long blob0, unsigned long long e0,
                                            const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                            unsigned long long dst_capacity, unsigned short* s_tab, uint8_t* slot,
                                            uint32_t* s_queue, unsigned int* s_qn) {
  const unsigned long long mis = reinterpret_cast<uintptr_t>(dst) & 15;
  uint8_t* vd = dst - mis;
  const unsigned long long o0 = TO_STREAM ? blob0 + 4 * e0 : blob0;
  const unsigned long long total = elem_start<TO_STREAM>(s_off, nE);
  const unsigned long long lim = min(o0 + total, dst_capacity) + mis;
  if (threadIdx.x == 0) *s_qn = 0;
  const unsigned long long A = o0 + mis;
  if (lim <= A) return;
  const unsigned long long T = A & ~15ull;
  const uint32_t a = (uint32_t)(A - T), lm = (uint32_t)(lim - T);
  uint8_t* vdT = vd + T;
  // source byte of output byte p (tile coordinates) of element e: srcT + p -/+ 4 * (e + 1)
  const uint8_t* srcT = TO_STREAM ? src + (T - mis) - 4ull * e0 : src + (T - mis) + 4ull * e0;
  int tsh = 8;
  while (((a + (uint32_t)total) >> tsh) >= (uint32_t)kTabSize) tsh++;
  for (int e = threadIdx.x; e <= nE; e += blockDim.x) {
    const uint32_t s0 = a + (uint32_t)elem_start<TO_STREAM>(s_off, e);
    s_st[e] = s0;
    if (e < nE) {
      const uint32_t s1 = a + (uint32_t)elem_start<TO_STREAM>(s_off, e + 1);
      const uint32_t f0 = e == 0 ? 0u : s0;  // element 0 also answers for the bytes before the tile start
      if (s1 > f0) {
        const uint32_t t0 = (f0 + (1u << tsh) - 1) >> tsh, t1 = (s1 - 1) >> tsh;
        for (uint32_t t = t0; t <= t1; t++) s_tab[t] = (unsigned short)e;
      }
    }
  }
  __syncthreads();
  const uint32_t q0 = (a + 15) >> 4, q1 = lm >> 4;  // chunks fully inside the tile's output
  const unsigned long long qT = T >> 4;
  constexpr int kFly = 4;
  for (int sub = 0; sub <= nE; sub += blockDim.x) {
    const int sub_end = min(sub + (int)blockDim.x, nE);
    const uint32_t qs0 = sub == 0 ? q0 : max(q0, (s_st[sub] + 15) >> 4);
    const uint32_t qs1 = sub_end == nE ? q1 : min(q1, s_st[sub_end] >> 4);
    // ---- phase A: chunks inside one payload --------------------------------------------------
    for (uint32_t q = qs0 + threadIdx.x; q < qs1; q += kFly * blockDim.x) {
      int so[kFly];
      bool ok[kFly];
      uint32_t w[kFly][4];
#pragma unroll
      for (int f = 0; f < kFly; f++) {
        const uint32_t qq = q + f * blockDim.x;
        ok[f] = false;
        so[f] = 0;
        if (qq < qs1) {
          const uint32_t cb = qq << 4;
          int e = s_tab[cb >> tsh];
          uint32_t nx = s_st[e + 1];
          while (nx <= cb) nx = s_st[++e + 1];
          ok[f] = cb + 16 <= nx && (!TO_STREAM || cb >= s_st[e] + 4);
          so[f] = TO_STREAM ? (int)cb - 4 * (e + 1) : (int)cb + 4 * (e + 1);
        }
      }
#pragma unroll
      for (int f = 0; f < kFly; f++) {
        if (ok[f]) {
          const uint8_t* sp = srcT + so[f];
#pragma unroll
          for (int k = 0; k < 4; k++) w[f][k] = load_word(sp + 4 * k, WORD_OK);
        }
      }
#pragma unroll
      for (int f = 0; f < kFly; f++)
        if (ok[f]) st_v4(vdT + ((q + f * blockDim.x) << 4), w[f][0], w[f][1], w[f][2], w[f][3]);
    }
    // ---- phase B: boundary chunks of this sub-tile's elements ---------------------------------
    const int e = sub + threadIdx.x;
    if (e > nE) continue;
    if (TO_STREAM) {
      if (e < nE) {
        const uint32_t hs = s_st[e];
        const uint32_t qa = hs >> 4, qb = (hs + 3) >> 4;
        const bool own_a = e == 0 || ((s_st[e - 1] + 3) >> 4) != qa;  // the previous prefix does not reach into qa
        if (own_a && !fast_prefix_chunk32<WORD_OK>(s_st, e, hs, qa << 4, a, lm, srcT, vdT))
          s_queue[atomicAdd(s_qn, 1u)] = qa;
        if (qb != qa && !fast_prefix_chunk32<WORD_OK>(s_st, e, hs, qb << 4, a, lm, srcT, vdT))
          s_queue[atomicAdd(s_qn, 1u)] = qb;
      } else if (lm & 15) {  // partial tail chunk with no prefix of this tile in it
        const uint32_t qt = lm >> 4;
        if (((s_st[nE - 1] + 3) >> 4) < qt) s_queue[atomicAdd(s_qn, 1u)] = qt;
      }
    } else {
      const uint32_t be = e < nE ? s_st[e] : lm;  // start of element e, or the tile end
      if ((be & 15) == 0 || be > lm) continue;    // chunk-aligned boundaries create no boundary chunk
      const uint32_t q = be >> 4;
      bool own = true;
      if (e > 0) {
        const uint32_t bp = s_st[e - 1];
        own = (bp >> 4) < q || (bp & 15) == 0;    // previous boundary is not strictly inside chunk q
      }
      if (own && !fast_boundary_chunk32<WORD_OK>(s_st, nE, e, be, q << 4, a, lm, srcT, vdT))
        s_queue[atomicAdd(s_qn, 1u)] = q;
    }
  }
  // ---- the boundary chunks that are not of the simple shape (several boundaries in 16 bytes, tile
  // edges): queued above and assembled here by consecutive threads. Done in place, the general
  // assembly ran with a handful of active lanes per warp in almost every warp.
  __syncthreads();
  const unsigned int nq = *s_qn;
  for (unsigned int i = threadIdx.x; i < nq; i += blockDim.x) {
    const uint32_t q = s_queue[i];
    const int eh = s_tab[max(q << 4, a) >> tsh];  // an element at or before the chunk's first byte
    assemble_chunk<TO_STREAM, WORD_OK>(s_off, nE, A, lim, e0, eh, qT + q, mis, src, vd, slot);
  }
}

// LEN_SIZE: 4 or 8 bytes per length; FIXED: lengths come from fixed-width rows (trailing NULs
// stripped) and the payload of element i starts at i * width instead of being packed.
template <int LEN_SIZE, bool FIXED, bool WORD_OK>
__global__ void __launch_bounds__(kSerThreads, 4)
k_bytes_serialize(const uint8_t* __restrict__ payload, const void* __restrict__ lengths,
                  unsigned long long n, uint32_t width, uint8_t* __restrict__ dst,
                  unsigned long long dst_capacity, unsigned long long* __restrict__ ws,
                  unsigned long long* __restrict__ total_out) {
  __shared__ unsigned long long s_off[kSerTile + 1];
  __shared__ uint32_t s_st[kSerTile + 1];
  __shared__ uint32_t s_queue[2 * kSerTile + 2];  // deferred boundary chunks (at most two per element + the tail)
  __shared__ unsigned int s_qn;
  __shared__ unsigned long long s_warp[kSerThreads / 32];
  __shared__ unsigned long long s_prefix;
  __shared__ unsigned int s_tile;
  __shared__ unsigned short s_tab[kTabSize + 1];
  __shared__ __align__(16) uint8_t s_slots[kSerThreads * kSlotBytes];
  unsigned long long* status = ws + WS_HEADER_WORDS;
  if (threadIdx.x == 0) s_tile = atomicAdd(reinterpret_cast<unsigned int*>(&ws[WS_TICKET]), 1u);
  __syncthreads();
  const unsigned int tile = s_tile;
  const unsigned long long e0 = (unsigned long long)tile * kSerTile;
  const int nE = (int)min((unsigned long long)kSerTile, n - e0);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  unsigned long long len[kSerItems];
  unsigned long long tsum = 0;
#pragma unroll
  for (int i = 0; i < kSerItems; i++) {
    const int el = threadIdx.x * kSerItems + i;
    unsigned long long l = 0;
    if (el < nE) {
      if (FIXED) {
        const uint8_t* row = payload + (e0 + el) * (unsigned long long)width;
        uint32_t k = width;
        while (k > 0 && row[k - 1] == 0) k--;  // numpy S<k>.item() strips trailing NULs
        l = k;
      } else if (LEN_SIZE == 4) {
        l = reinterpret_cast<const uint32_t*>(lengths)[e0 + el];
      } else {
        l = reinterpret_cast<const unsigned long long*>(lengths)[e0 + el];
      }
    }
    len[i] = l;
    tsum += l;
  }
  // warp-shuffle inclusive scan of the per-thread sums, then across warps
  unsigned long long incl = tsum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  unsigned long long wbase = 0, tile_total = 0;
#pragma unroll
  for (int wdx = 0; wdx < kSerThreads / 32; wdx++) {
    if (wdx < warp) wbase += s_warp[wdx];
    tile_total += s_warp[wdx];
  }
  unsigned long long run = wbase + incl - tsum;
#pragma unroll
  for (int i = 0; i < kSerItems; i++) {
    s_off[threadIdx.x * kSerItems + i] = run;
    run += len[i];
  }
  if (threadIdx.x == kSerThreads - 1) s_off[kSerTile] = run;
  if (warp == 0) {
    const unsigned long long ex = lookback_exclusive(status, tile, tile_total, ws);
    if (lane == 0) s_prefix = ex;
  }
  __syncthreads();
  const unsigned long long blob0 = s_prefix;
  if (e0 + nE == n && threadIdx.x == 0) {
    const unsigned long long total = blob0 + s_off[nE] + 4ull * n;
    ws[WS_TOTAL] = total;
    if (total_out) *total_out = total;
    if (total > dst_capacity) atomicMax(&ws[WS_ERROR], (unsigned long long)ERR_CAPACITY);
  }
  if (FIXED) {
    // rows are not packed: element e's payload lives at (e0+e)*width, i.e. the "blob" coordinate
    // of its first byte is not blob0 + s_off[e]. Byte path with explicit per-element source.
    const unsigned long long o0 = blob0 + 4 * e0;
    const unsigned long long o1 = min(blob0 + s_off[nE] + 4ull * (e0 + nE), dst_capacity);
    for (unsigned long long p = o0 + threadIdx.x; p < o1; p += kSerThreads) {
      const unsigned long long rel = p - o0;
      int lo = 0, hi = nE;
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (s_off[mid] + 4ull * mid <= rel) lo = mid; else hi = mid;
      }
      const unsigned long long es = s_off[lo] + 4ull * lo;
      const unsigned long long l = s_off[lo + 1] - s_off[lo];
      dst[p] = rel < es + 4 ? (uint8_t)(l >> (8 * (rel - es)))
                            : payload[(e0 + lo) * (unsigned long long)width + (rel - es - 4)];
    }
  } else {
    uint8_t* slot = s_slots + threadIdx.x * kSlotBytes;
    if (s_off[nE] + 4ull * nE < kTile32Limit)  // uniform across the CTA
      move_tile32<true, WORD_OK>(s_off, s_st, nE, blob0, e0, payload, dst, dst_capacity, s_tab, slot, s_queue, &s_qn);
    else
      move_tile<true, WORD_OK>(s_off, nE, blob0, e0, payload, dst, dst_capacity, s_tab, slot);
  }
}

// =================================================================================================
// deserialise
// =================================================================================================
constexpr int kDesChunk = 16384;   // bytes of stream per chunk
constexpr int kDesWindow = 512;    // look-behind window for the entry guess
constexpr int kDesWarps = 4;       // warps per CTA of k_des_chase (one lane per chunk)
constexpr int kDesThreads = kDesWarps * 32;
constexpr int kDesListCap = 512;   // header positions recorded per chunk (uint16, chunk-relative)
constexpr unsigned long long kNoGuess = ~0ull;

// header words of the deserialise workspace
enum : int {
  DW_STATUS = 0,        // 0 ok, 1 needs sweep, 2 truncated
  DW_NELEM = 1,         // elements decoded
  DW_KEND = 2,          // last chunk that holds decoded elements (+1)
  DW_FIRST_MISMATCH = 3,
  DW_PAYLOAD = 4,       // payload bytes of the decoded elements
  DW_SWEPT = 5,         // sticky: sweep ran
  DW_HEADER_WORDS = 8
};

struct DesArrays {
  unsigned long long* g;     // entry header position of chunk k (absolute stream offset)
  unsigned long long* x;     // exit: first header position >= end of chunk k
  unsigned long long* base;  // elements before chunk k
  unsigned int* cnt;         // elements whose header lies in chunk k
  unsigned int* err;         // 1: chain of chunk k hit a truncated header/payload after cnt elements
  unsigned short* list;      // [K][kDesListCap] chunk-relative header positions
};

// Little-endian uint32 at an arbitrary byte position of the stream: two aligned word loads (both
// normally L1 hits) and a funnel shift. The second word is only read when the value straddles.
__device__ __forceinline__ uint32_t ldg_u32_unaligned(const uint8_t* __restrict__ base_aligned4,
                                                      unsigned long long pos) {
  const uint32_t* w = reinterpret_cast<const uint32_t*>(base_aligned4) + (pos >> 2);
  const uint32_t lo = __ldg(w);
  const unsigned sh = (unsigned)(pos & 3) * 8;
  if (sh == 0) return lo;
  return __funnelshift_r(lo, __ldg(w + 1), sh);
}

// Speculative chunk chase, two kernels.
//   k_des_vote : one WARP per chunk guesses the chunk's entry header. The 32 lanes chase from every
//      byte of the chunk's 512-byte look-behind window (staged in shared memory with one coalesced
//      round trip) to the chunk start and vote on where they land; the plurality (ties: smallest
//      offset) is the guess. Chunks are independent, so the grid is as wide as the stream is long -
//      the first version ran the 32 votes of a warp's chunks one after the other in front of the
//      chase, and that serial prologue was a third of the whole deserialise.
//   k_des_chase: one LANE per chunk follows the chain of its chunk straight from global memory from
//      the guessed entry, recording header positions: a warp instruction advances 32 independent
//      chains.
// Exactness does not depend on the guess: k_des_verify checks entry[k] == exit[k-1] for all k.
// stream4 = stream rounded down to a 4-byte boundary, bias = stream - stream4 (0..3).
constexpr int kVoteWarps = 8;
constexpr int kWinWords = kDesWindow / 4 + 4;  // look-behind window + 16 bytes

__global__ void __launch_bounds__(kVoteWarps * 32)
k_des_vote(const uint8_t* __restrict__ stream4, unsigned int bias, unsigned long long S,
           unsigned long long K, DesArrays A) {
  __shared__ unsigned int s_votes[kVoteWarps][kDesWindow];
  __shared__ unsigned int s_win[kVoteWarps][kWinWords];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned long long k = (unsigned long long)blockIdx.x * kVoteWarps + warp;
  if (k >= K) return;
  if (k == 0) {
    if (lane == 0) A.g[0] = 0;
    return;
  }
  unsigned int* votes = s_votes[warp];
  unsigned int* win = s_win[warp];
  const unsigned long long c0 = k * kDesChunk;
  const unsigned long long S4 = (S + bias + 3) >> 2;  // words of stream4 that hold stream bytes
  const unsigned long long w0 = (c0 - kDesWindow) >> 2;  // first window word (stream4 is 4-byte aligned)
  for (int i = lane; i < kWinWords; i += 32) {
    const unsigned long long wi = w0 + i;
    win[i] = wi < S4 ? __ldg(reinterpret_cast<const uint32_t*>(stream4) + wi) : 0u;
  }
  for (int i = lane; i < kDesWindow; i += 32) votes[i] = 0;
  __syncwarp();
  // window byte b (0 = stream position c0 - kDesWindow) sits at byte b + bias of win[]
  auto win_len = [&](unsigned int b) {
    const unsigned int q = b + bias;
    const unsigned int lo = win[q >> 2];
    const unsigned int sh = (q & 3) * 8;
    return sh ? __funnelshift_r(lo, win[(q >> 2) + 1], sh) : lo;
  };
  const unsigned long long wbase = c0 - kDesWindow;
  for (int it = 0; it < kDesWindow / 32; it++) {
    unsigned long long pos = wbase + it * 32 + lane;
    bool ok = true;
    while (pos < c0) {
      if (pos + 4 > S) { ok = false; break; }
      const unsigned long long nx = pos + 4 + win_len((unsigned int)(pos - wbase));
      if (nx > S) { ok = false; break; }
      pos = nx;
    }
    if (ok && pos - c0 < kDesWindow) atomicAdd(&votes[pos - c0], 1u);
  }
  __syncwarp();
  unsigned int best = 0;
  for (int i = lane; i < kDesWindow; i += 32) {
    const unsigned int v = votes[i];
    if (v) best = max(best, (v << 16) | (unsigned int)(kDesWindow - 1 - i));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
  if (lane == 0) A.g[k] = best ? c0 + (kDesWindow - 1 - (best & 0xffffu)) : kNoGuess;
}

constexpr unsigned int kErrRechase = 2;  // A.err value: a wide vote replaced this chunk's guess, chase it again

__global__ void __launch_bounds__(kDesThreads)
k_des_chase(const uint8_t* __restrict__ stream4, unsigned int bias, unsigned long long S,
            unsigned long long K, DesArrays A, int only_flagged) {
  const unsigned long long k = (unsigned long long)blockIdx.x * kDesThreads + threadIdx.x;
  if (k >= K) return;
  if (only_flagged && A.err[k] != kErrRechase) return;
  const unsigned long long c0 = k * kDesChunk, c1 = min(c0 + kDesChunk, S);
  const unsigned long long my_g = A.g[k];
  unsigned long long pos = my_g;
  unsigned int n = 0, err = 0;
  unsigned short* list = A.list + k * kDesListCap;
  if (my_g != kNoGuess) {
    while (pos < c1) {
      if (pos + 4 > S) { err = 1; break; }
      const unsigned long long nx = pos + 4 + ldg_u32_unaligned(stream4, pos + bias);
      if (nx > S) { err = 1; break; }
      if (n < kDesListCap) list[n] = (unsigned short)(pos - c0);
      n++;
      pos = nx;
    }
  }
  A.x[k] = pos; A.cnt[k] = n; A.err[k] = err;
}

// Second-chance vote for chunks whose first guess failed verification, with a look-behind window of
// `window` bytes instead of 512: elements longer than the small window leave it without a single
// header, and every such chunk used to fall to the sequential sweep (130 ms for 256 MiB of 1 KiB
// elements). One warp per failed chunk; candidates are read straight from global memory (the window
// does not fit shared memory), chains are cut after kWideHops hops, landing positions are counted in
// a small per-warp hash table. The result is again only a guess - verification decides.
constexpr int kWideHops = 96;
constexpr int kWideSlots = 128;

__global__ void __launch_bounds__(kVoteWarps * 32)
k_des_vote_wide(const uint8_t* __restrict__ stream4, unsigned int bias, unsigned long long S,
                unsigned long long K, DesArrays A, const unsigned long long* __restrict__ hdr,
                unsigned long long window) {
  __shared__ unsigned long long s_key[kVoteWarps][kWideSlots];
  __shared__ unsigned int s_cnt[kVoteWarps][kWideSlots];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned long long k = (unsigned long long)blockIdx.x * kVoteWarps + warp;
  if (k >= K || k == 0 || k < hdr[DW_FIRST_MISMATCH]) return;
  if (A.g[k] == A.x[k - 1] && A.g[k] != kNoGuess) return;  // consistent with its predecessor so far: keep
  unsigned long long* key = s_key[warp];
  unsigned int* cnt = s_cnt[warp];
  for (int i = lane; i < kWideSlots; i += 32) { key[i] = ~0ull; cnt[i] = 0; }
  __syncwarp();
  const unsigned long long c0 = k * kDesChunk;
  const unsigned long long wbase = c0 > window ? c0 - window : 0;
  for (unsigned long long p0 = wbase; p0 < c0; p0 += 32) {
    unsigned long long pos = p0 + lane;
    bool ok = pos < c0;
    for (int hop = 0; ok && pos < c0; hop++) {
      if (hop >= kWideHops || pos + 4 > S) { ok = false; break; }
      const unsigned long long nx = pos + 4 + ldg_u32_unaligned(stream4, pos + bias);
      if (nx > S) { ok = false; break; }
      pos = nx;
    }
    if (ok) {  // count the landing position
      unsigned int h = (unsigned int)((pos * 0x9E3779B97F4A7C15ull) >> 57);  // 7 bits
      for (int probe = 0; probe < kWideSlots; probe++) {
        const unsigned long long prev = atomicCAS(&key[h], ~0ull, pos);
        if (prev == ~0ull || prev == pos) { atomicAdd(&cnt[h], 1u); break; }
        h = (h + 1) & (kWideSlots - 1);
      }
    }
  }
  __syncwarp();
  unsigned int best = 0;
  unsigned long long best_pos = ~0ull;
  for (int i = lane; i < kWideSlots; i += 32) {
    const unsigned int c = cnt[i];
    if (c > best || (c == best && c > 0 && key[i] < best_pos)) { best = c; best_pos = key[i]; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned int ob = __shfl_xor_sync(0xffffffffu, best, o);
    const unsigned long long op = __shfl_xor_sync(0xffffffffu, best_pos, o);
    if (ob > best || (ob == best && op < best_pos)) { best = ob; best_pos = op; }
  }
  if (lane == 0) {
    A.g[k] = best ? best_pos : kNoGuess;
    A.err[k] = kErrRechase;
  }
}

// One chase step straight from global memory (sweep / overflow paths).
__device__ __forceinline__ bool chase_global(const uint8_t* __restrict__ stream, unsigned long long S,
                                             unsigned long long pos, unsigned long long* next) {
  if (pos + 4 > S) return false;
  const uint32_t l = (uint32_t)stream[pos] | ((uint32_t)stream[pos + 1] << 8) |
                     ((uint32_t)stream[pos + 2] << 16) | ((uint32_t)stream[pos + 3] << 24);
  const unsigned long long nx = pos + 4 + l;
  if (nx > S) return false;
  *next = nx;
  return true;
}

// Resolve = three small kernels (the chunk count reaches 262 144 for a 2 GB stream, too many for
// one CTA): verify entry[k] == exit[k-1] and find the first truncated chunk; look-back scan of the
// per-chunk counts over the verified prefix; decide the outcome.
constexpr int kResThreads = 256;
constexpr int kResItems = 8;
constexpr int kResTile = kResThreads * kResItems;
enum : int { DW_FIRST = 3, DW_FERR = 6, DW_TOTAL = 7 };  // header words (DW_FIRST == DW_FIRST_MISMATCH)

__global__ void __launch_bounds__(kResThreads)
k_des_verify(unsigned long long K, DesArrays A, unsigned long long* hdr) {
  const unsigned long long k = (unsigned long long)blockIdx.x * kResThreads + threadIdx.x;
  if (k >= K) return;
  const unsigned long long expect = k == 0 ? 0 : A.x[k - 1];
  if (A.g[k] != expect) atomicMin(&hdr[DW_FIRST], k);   // hdr[DW_FIRST] starts at ~0
  if (A.err[k]) atomicMin(&hdr[DW_FERR], k);
}

// base[k] = number of elements in verified chunks before k (chunks >= first contribute nothing)
__global__ void __launch_bounds__(kResThreads)
k_des_scan(unsigned long long K, DesArrays A, unsigned long long* hdr, unsigned long long* sws) {
  __shared__ unsigned long long s_warp[kResThreads / 32];
  __shared__ unsigned long long s_prefix;
  __shared__ unsigned int s_tile;
  if (threadIdx.x == 0) s_tile = atomicAdd(reinterpret_cast<unsigned int*>(&sws[WS_TICKET]), 1u);
  __syncthreads();
  const unsigned int tile = s_tile;
  const unsigned long long first = min(hdr[DW_FIRST], K);
  const unsigned long long k0 = (unsigned long long)tile * kResTile + (unsigned long long)threadIdx.x * kResItems;
  unsigned int c[kResItems];
  unsigned long long tsum = 0;
#pragma unroll
  for (int i = 0; i < kResItems; i++) {
    const unsigned long long k = k0 + i;
    c[i] = (k < K && k < first) ? A.cnt[k] : 0u;
    tsum += c[i];
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned long long incl = tsum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  unsigned long long wbase = 0, tile_total = 0;
#pragma unroll
  for (int w = 0; w < kResThreads / 32; w++) {
    if (w < warp) wbase += s_warp[w];
    tile_total += s_warp[w];
  }
  if (warp == 0) {
    const unsigned long long ex = lookback_exclusive(sws + WS_HEADER_WORDS, tile, tile_total, sws);
    if (lane == 0) s_prefix = ex;
  }
  __syncthreads();
  unsigned long long run = s_prefix + wbase + incl - tsum;
#pragma unroll
  for (int i = 0; i < kResItems; i++) {
    const unsigned long long k = k0 + i;
    if (k < K) A.base[k] = run;
    run += c[i];
  }
  if (k0 <= K - 1 && K - 1 < k0 + kResItems) hdr[DW_TOTAL] = run;  // the thread that owns the last chunk
}

__global__ void k_des_finalize(unsigned long long S, unsigned long long K, long long max_elements,
                               DesArrays A, unsigned long long* hdr) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const unsigned long long first = min(hdr[DW_FIRST], K);
  const unsigned long long total = hdr[DW_TOTAL];  // elements in the verified prefix [0, first)
  unsigned long long ferr = hdr[DW_FERR];
  if (ferr >= first) ferr = K;                     // only errors inside the verified prefix count
  unsigned long long status, nelem, kend;
  const bool limited = max_elements >= 0;
  const unsigned long long M = limited ? (unsigned long long)max_elements : ~0ull;
  // elements that decode before anything goes wrong inside the verified prefix
  const unsigned long long good = ferr < K ? A.base[ferr] + A.cnt[ferr] : total;
  if (limited && good >= M) {
    unsigned long long lo = 0, hi = first;  // smallest kk with base[kk] + cnt[kk] >= M
    while (lo < hi) {
      const unsigned long long mid = (lo + hi) >> 1;
      if (A.base[mid] + A.cnt[mid] >= M) hi = mid; else lo = mid + 1;
    }
    status = 0; nelem = M; kend = M == 0 ? 0 : lo + 1;
  } else if (ferr < K) {
    status = 2; nelem = good; kend = ferr + 1;  // the true chain runs off the buffer
  } else if (first < K) {
    status = 1; nelem = total; kend = first;    // a guessed entry was wrong: sweep
  } else {
    status = 0; nelem = total; kend = K;
    const unsigned long long end = K ? A.x[K - 1] : 0;
    if (end != S) status = 2;
    if (limited) status = 2;  // fewer than M elements before the end: the reference unpacks past it
  }
  hdr[DW_STATUS] = status; hdr[DW_NELEM] = nelem; hdr[DW_KEND] = kend;
  hdr[DW_FIRST_MISMATCH] = first;
}

// Sequential repair from the first mismatching chunk (whatever the votes could not settle): lane 0
// walks the true chain, re-chasing chunks whose guessed entry was wrong; the chunks that a long
// element skips entirely are filled in by the whole warp at once. Exact by construction.
__global__ void k_des_sweep(const uint8_t* __restrict__ stream, unsigned long long S,
                            unsigned long long K, long long max_elements, DesArrays A,
                            unsigned long long* hdr) {
  if (blockIdx.x != 0 || threadIdx.x >= 32) return;
  const int lane = threadIdx.x;
  unsigned long long k = hdr[DW_FIRST_MISMATCH];
  unsigned long long count = k < K ? A.base[k] : 0;
  unsigned long long e = k == 0 ? 0 : A.x[k - 1];
  const bool limited = max_elements >= 0;
  const unsigned long long M = limited ? (unsigned long long)max_elements : ~0ull;
  bool stop = false;
  if (lane == 0) hdr[DW_SWEPT] = 1;
  while (k < K) {
    const unsigned long long c0 = k * kDesChunk, c1 = min(c0 + kDesChunk, S);
    if (stop || e >= c1) {  // nothing of the true chain starts in chunks [k, k2): all lanes fill them
      const unsigned long long k2 = stop ? K : min(K, e / kDesChunk);
      const unsigned long long kend = max(k2, k + 1);
      for (unsigned long long kk = k + lane; kk < kend; kk += 32) {
        A.g[kk] = e; A.x[kk] = e; A.cnt[kk] = 0; A.err[kk] = 0;
      }
      k = kend;
      continue;
    }
    // lane 0 handles the chunk; the outcome is broadcast
    unsigned long long ne = e, ncount = count;
    int nstop = 0;
    if (lane == 0) {
      if (A.g[k] == e && A.err[k] != kErrRechase) {  // the guess was right after all
        ncount += A.cnt[k];
        if (A.err[k] && !(limited && ncount >= M)) nstop = 1;
        ne = A.x[k];
        if (limited && ncount >= M) nstop = 1;
      } else {
        unsigned long long pos = e;
        unsigned int n = 0, err = 0;
        unsigned short* list = A.list + k * kDesListCap;
        while (pos < c1) {
          unsigned long long nx;
          if (!chase_global(stream, S, pos, &nx)) { err = 1; break; }
          if (n < kDesListCap) list[n] = (unsigned short)(pos - c0);
          n++;
          pos = nx;
          if (limited && count + n >= M) break;
        }
        A.g[k] = e; A.x[k] = pos; A.cnt[k] = n; A.err[k] = err;
        ncount += n;
        ne = pos;
        if (limited && ncount >= M) { nstop = 1; A.err[k] = 0; }
        else if (err) nstop = 1;
      }
    }
    e = __shfl_sync(0xffffffffu, ne, 0);
    count = __shfl_sync(0xffffffffu, ncount, 0);
    stop = __shfl_sync(0xffffffffu, nstop, 0) != 0;
    k++;
  }
}

// Element blob offsets: offsets[j] = header_pos(j) - 4*j for the elements whose header lies in
// chunk k; the chunk holding the last decoded element also writes offsets[N].
__global__ void __launch_bounds__(64)
k_des_offsets(const uint8_t* __restrict__ stream, unsigned long long S, DesArrays A,
              const unsigned long long* __restrict__ hdr, long long* __restrict__ offsets) {
  const unsigned long long k = blockIdx.x;
  const unsigned long long N = hdr[DW_NELEM];
  if (hdr[DW_STATUS] != 0 || k >= hdr[DW_KEND]) return;
  const unsigned long long base = A.base[k];
  if (base >= N) return;
  const unsigned long long c0 = k * kDesChunk;
  const unsigned int n = (unsigned int)min((unsigned long long)A.cnt[k], N - base);
  if (n == 0) return;
  const bool last = base + n == N;
  if (A.cnt[k] <= kDesListCap) {
    const unsigned short* list = A.list + k * kDesListCap;
    for (unsigned int i = threadIdx.x; i < n; i += blockDim.x)
      offsets[base + i] = (long long)(c0 + list[i]) - 4ll * (long long)(base + i);
    if (last && threadIdx.x == 0) {
      unsigned long long nx = 0;
      chase_global(stream, S, c0 + list[n - 1], &nx);
      offsets[N] = (long long)nx - 4ll * (long long)N;
    }
  } else if (threadIdx.x == 0) {  // very short elements: re-chase this chunk
    unsigned long long pos = A.g[k];
    for (unsigned int i = 0; i < n; i++) {
      offsets[base + i] = (long long)pos - 4ll * (long long)(base + i);
      unsigned long long nx = pos;
      chase_global(stream, S, pos, &nx);
      pos = nx;
    }
    if (last) offsets[N] = (long long)pos - 4ll * (long long)N;
  }
}

constexpr int kGatTile = 1024;
constexpr int kGatThreads = 256;

template <bool WORD_OK>
__global__ void __launch_bounds__(kGatThreads, 4)
k_des_gather(const uint8_t* __restrict__ stream, const long long* __restrict__ offsets,
             const unsigned long long* __restrict__ hdr, uint8_t* __restrict__ blob,
             unsigned long long blob_capacity) {
  __shared__ unsigned long long s_off[kGatTile + 1];
  __shared__ uint32_t s_st[kGatTile + 1];
  __shared__ uint32_t s_queue[kGatTile + 2];  // deferred boundary chunks (at most one per element + the tail)
  __shared__ unsigned int s_qn;
  __shared__ unsigned short s_tab[kTabSize + 1];
  __shared__ __align__(16) uint8_t s_slots[kGatThreads * kSlotBytes];
  const unsigned long long N = hdr[DW_NELEM];
  if (hdr[DW_STATUS] != 0) return;
  const unsigned long long e0 = (unsigned long long)blockIdx.x * kGatTile;
  if (e0 >= N) return;
  const int nE = (int)min((unsigned long long)kGatTile, N - e0);
  const unsigned long long b0 = (unsigned long long)offsets[e0];
  for (int i = threadIdx.x; i <= nE; i += kGatThreads)
    s_off[i] = (unsigned long long)offsets[e0 + i] - b0;
  __syncthreads();
  uint8_t* slot = s_slots + threadIdx.x * kSlotBytes;
  if (s_off[nE] < kTile32Limit)  // uniform across the CTA
    move_tile32<false, WORD_OK>(s_off, s_st, nE, b0, e0, stream, blob, blob_capacity, s_tab, slot, s_queue, &s_qn);
  else
    move_tile<false, WORD_OK>(s_off, nE, b0, e0, stream, blob, blob_capacity, s_tab, slot);
}

// Gather for long elements. k_des_gather gives every CTA 1024 ELEMENTS, which is 1 GiB of payload when
// the elements are 1 MiB each: a handful of CTAs then copy the whole stream. Here the work is cut by
// BYTES instead: a CTA owns kLongSeg bytes of the compact payload, finds the first element that
// reaches into its segment with a 32-way search over the offsets, and copies element piece by
// element piece (16 aligned destination bytes per thread and step, source words are 4-byte aligned
// because an element's payload moves by 4 * (index + 1) bytes). Used when the mean element is >= 2 KiB.
constexpr int kLongSeg = 32768;

template <bool WORD_OK>
__global__ void __launch_bounds__(256)
k_des_gather_long(const uint8_t* __restrict__ stream, const long long* __restrict__ offsets,
                  const unsigned long long* __restrict__ hdr, uint8_t* __restrict__ blob,
                  unsigned long long blob_capacity) {
  __shared__ unsigned long long s_first;
  if (hdr[DW_STATUS] != 0) return;
  const unsigned long long N = hdr[DW_NELEM];
  const unsigned long long total = min((unsigned long long)offsets[N], blob_capacity);
  const unsigned long long s0 = (unsigned long long)blockIdx.x * kLongSeg;
  if (s0 >= total) return;
  const unsigned long long s1 = min(s0 + kLongSeg, total);
  if (threadIdx.x < 32) {  // largest e with offsets[e] <= s0: 32-way search
    unsigned long long lo = 0, hi = N;  // answer in [lo, hi)
    while (hi - lo > 1) {
      const unsigned long long step = (hi - lo + 31) / 32;
      const unsigned long long probe = lo + step * (threadIdx.x + 1);
      const bool le = probe < hi && (unsigned long long)offsets[probe] <= s0;
      const unsigned m = __ballot_sync(0xffffffffu, le);
      const int c = __popc(m);  // probes are monotone: the first c lanes are <=
      const unsigned long long nlo = lo + step * c;
      hi = min(hi, lo + step * (c + 1));
      lo = nlo;
    }
    if (threadIdx.x == 0) s_first = lo;
  }
  __syncthreads();
  for (unsigned long long e = s_first; e < N; e++) {
    const unsigned long long o0 = (unsigned long long)offsets[e];
    if (o0 >= s1) break;
    const unsigned long long o1 = (unsigned long long)offsets[e + 1];
    const unsigned long long a = max(o0, s0), b = min(o1, s1);
    if (a >= b) continue;
    uint8_t* d = blob + a;
    const uint8_t* sp = stream + a + 4ull * (e + 1);
    const unsigned long long n = b - a;
    const unsigned long long head = min(n, (unsigned long long)((16 - (reinterpret_cast<uintptr_t>(d) & 15)) & 15));
    if (threadIdx.x < head) d[threadIdx.x] = sp[threadIdx.x];
    const unsigned long long chunks = (n - head) >> 4;
    uint8_t* db = d + head;
    const uint8_t* sb = sp + head;
    for (unsigned long long c = threadIdx.x; c < chunks; c += 256) {
      uint32_t w[4];
#pragma unroll
      for (int k = 0; k < 4; k++) w[k] = load_word(sb + 16 * c + 4 * k, WORD_OK);
      st_v4(db + 16 * c, w[0], w[1], w[2], w[3]);
    }
    const unsigned long long done = head + (chunks << 4);
    if (done + threadIdx.x < n) d[done + threadIdx.x] = sp[done + threadIdx.x];
  }
}

// A stream that fits one chunk (a request's handful of strings): everything in one CTA and one launch.
// The stream is staged in shared memory, thread 0 follows the chain there (the general path chases a
// single chunk with a single lane as well, but from global memory and behind six more launches and
// two host round trips), offsets follow from the header positions alone (element i starts at
// header_i - 4 i of the payload), then all threads move the payload. hdr[DW_STATUS]: 0 ok,
// 2 truncated, 3 offsets capacity, 4 payload capacity; hdr[DW_NELEM], hdr[DW_PAYLOAD] as usual.
constexpr int kTinyThreads = 256;

__global__ void __launch_bounds__(kTinyThreads)
k_des_tiny(const uint8_t* __restrict__ stream, unsigned int S, long long max_elements,
           long long* __restrict__ offsets, unsigned long long offsets_capacity, uint8_t* __restrict__ payload,
           unsigned long long payload_capacity, unsigned long long* __restrict__ hdr) {
  __shared__ __align__(16) uint8_t s_stream[kDesChunk + 16];
  __shared__ unsigned short s_pos[kDesChunk / 4 + 1];
  __shared__ unsigned int s_n, s_status, s_end;
  const int tid = threadIdx.x;
  unsigned int bulk = 0;
  if ((reinterpret_cast<uintptr_t>(stream) & 15) == 0) {
    bulk = S & ~15u;
    for (unsigned int i = tid * 16u; i < bulk; i += kTinyThreads * 16u)
      *reinterpret_cast<uint4*>(s_stream + i) = __ldg(reinterpret_cast<const uint4*>(stream + i));
  }
  for (unsigned int i = bulk + tid; i < S; i += kTinyThreads) s_stream[i] = stream[i];
  __syncthreads();
  if (tid == 0) {
    unsigned int pos = 0, n = 0, status = 0;
    while (pos < S && (max_elements < 0 || (long long)n < max_elements)) {
      if (pos + 4 > S) { status = 2; break; }
      const unsigned int len = (unsigned int)s_stream[pos] | ((unsigned int)s_stream[pos + 1] << 8) |
                               ((unsigned int)s_stream[pos + 2] << 16) | ((unsigned int)s_stream[pos + 3] << 24);
      if (len > S - pos - 4) { status = 2; break; }
      s_pos[n++] = (unsigned short)pos;
      pos += 4 + len;
    }
    if (status == 0 && max_elements > 0 && (long long)n < max_elements) status = 2;  // asked for more than there is
    if (status == 0 && n > offsets_capacity) status = 3;
    if (status == 0 && (unsigned long long)(pos - 4 * n) > payload_capacity) status = 4;
    s_n = n; s_status = status; s_end = pos;
    hdr[DW_STATUS] = status; hdr[DW_NELEM] = n; hdr[DW_PAYLOAD] = pos - 4 * n;
  }
  __syncthreads();
  if (s_status != 0) return;
  const unsigned int n = s_n, end = s_end;
  for (unsigned int i = tid; i <= n; i += kTinyThreads) offsets[i] = (long long)((i < n ? s_pos[i] : end) - 4 * i);
  // payload: one warp per element, a byte per lane and round (elements here are short)
  const int lane = tid & 31, warp = tid >> 5;
  for (unsigned int i = warp; i < n; i += kTinyThreads / 32) {
    const unsigned int a = s_pos[i] + 4u, b = i + 1 < n ? s_pos[i + 1] : end;
    uint8_t* dst = payload + (a - 4u * (i + 1));
    for (unsigned int j = a + lane; j < b; j += 32) dst[j - a] = s_stream[j];
  }
}

static uint64_t des_chunks(uint64_t stream_bytes) { return (stream_bytes + kDesChunk - 1) / kDesChunk; }

}  // namespace tcb

using namespace tcb;

extern "C" {

uint64_t TcbBytesSerializeWorkspace(uint64_t n_elements) {
  const uint64_t tiles = (n_elements + kSerTile - 1) / kSerTile;
  return (WS_HEADER_WORDS + tiles + 1) * 8;
}

uint64_t TcbBytesDeserializeWorkspace(uint64_t stream_bytes) {
  const uint64_t K = des_chunks(stream_bytes) + 1;
  // header + scan status + g, x, base (u64) + cnt, err (u32) + list (u16 * cap), each 256-byte aligned
  return 256 + round_up((WS_HEADER_WORDS + K / kResTile + 2) * 8, 256) + 3 * round_up(K * 8, 256) +
         2 * round_up(K * 4, 256) + round_up(K * kDesListCap * 2, 256);
}

static int serialize_common(int device_id, const void* payload_dev, const void* lengths_dev,
                            int length_size, bool fixed, uint32_t width, uint64_t n, void* dst_dev,
                            uint64_t dst_capacity, void* workspace_dev, uint64_t workspace_bytes,
                            uint64_t* total_dev, uint64_t* total_host, void* stream) {
  TCB_CHECK_ARG(fixed || length_size == 4 || length_size == 8, "TcbBytesSerialize: length_size must be 4 or 8");
  TCB_DEVICE_GUARD(device_id);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (n == 0) {
    if (total_dev) TCB_CUDA(cudaMemsetAsync(total_dev, 0, 8, s));
    if (total_host) { TCB_CUDA(cudaStreamSynchronize(s)); *total_host = 0; }
    return TCB_OK;
  }
  TCB_CHECK_ARG(payload_dev != nullptr || !fixed, "TcbBytesSerialize: payload is NULL");
  TCB_CHECK_ARG(fixed || lengths_dev != nullptr, "TcbBytesSerialize: lengths is NULL");
  TCB_CHECK_ARG(dst_dev != nullptr && workspace_dev != nullptr, "TcbBytesSerialize: NULL dst/workspace");
  const uint64_t need = TcbBytesSerializeWorkspace(n);
  if (workspace_bytes < need)
    return set_error(TCB_ERR_OUT_OF_RANGE, "TcbBytesSerialize: workspace needs %llu bytes, got %llu",
                     (unsigned long long)need, (unsigned long long)workspace_bytes);
  TCB_CHECK_ARG(reinterpret_cast<uintptr_t>(workspace_dev) % 8 == 0, "TcbBytesSerialize: workspace not 8-byte aligned");
  TCB_CHECK_ARG(fixed || reinterpret_cast<uintptr_t>(lengths_dev) % length_size == 0,
                "TcbBytesSerialize: lengths not aligned");
  const uint64_t tiles = (n + kSerTile - 1) / kSerTile;
  TCB_CHECK_ARG(tiles < (1ull << 31), "TcbBytesSerialize: too many elements");
  unsigned long long* ws = static_cast<unsigned long long*>(workspace_dev);
  TCB_CUDA(cudaMemsetAsync(ws, 0, need, s));
  const uint8_t* pay = static_cast<const uint8_t*>(payload_dev);
  uint8_t* dst = static_cast<uint8_t*>(dst_dev);
  const bool word_ok = ((reinterpret_cast<uintptr_t>(pay) ^ reinterpret_cast<uintptr_t>(dst)) & 3) == 0;
  unsigned long long* tot = reinterpret_cast<unsigned long long*>(total_dev);
#define TCB_SER(LS, FX, WO)                                                                     \
  k_bytes_serialize<LS, FX, WO><<<(unsigned)tiles, kSerThreads, 0, s>>>(pay, lengths_dev, n, width, dst, \
                                                                        dst_capacity, ws, tot)
  if (fixed) TCB_SER(4, true, false);
  else if (length_size == 4) { if (word_ok) TCB_SER(4, false, true); else TCB_SER(4, false, false); }
  else { if (word_ok) TCB_SER(8, false, true); else TCB_SER(8, false, false); }
#undef TCB_SER
  count_launch();
  TCB_CUDA(cudaGetLastError());
  if (total_host) {
    unsigned long long h[WS_HEADER_WORDS];
    TCB_CUDA(cudaMemcpyAsync(h, ws, sizeof(h), cudaMemcpyDeviceToHost, s));
    TCB_CUDA(cudaStreamSynchronize(s));
    *total_host = h[WS_TOTAL];
    if (h[WS_ERROR] == ERR_SPIN) return set_error(TCB_ERR_INTERNAL, "TcbBytesSerialize: look-back scan timed out");
    if (h[WS_ERROR] == ERR_CAPACITY)
      return set_error(TCB_ERR_OUT_OF_RANGE, "TcbBytesSerialize: serialized size %llu exceeds capacity %llu",
                       (unsigned long long)h[WS_TOTAL], (unsigned long long)dst_capacity);
  }
  return TCB_OK;
}

int TcbBytesSerialize(int device_id, const void* payload_dev, const void* lengths_dev,
                      int length_size, uint64_t n_elements, void* dst_dev, uint64_t dst_capacity,
                      void* workspace_dev, uint64_t workspace_bytes, uint64_t* total_dev,
                      uint64_t* total_host, void* stream) {
  return serialize_common(device_id, payload_dev, lengths_dev, length_size, false, 0, n_elements, dst_dev,
                          dst_capacity, workspace_dev, workspace_bytes, total_dev, total_host, stream);
}

int TcbBytesSerializeFixedWidth(int device_id, const void* rows_dev, uint32_t width,
                                uint64_t n_elements, void* dst_dev, uint64_t dst_capacity,
                                void* workspace_dev, uint64_t workspace_bytes,
                                uint64_t* total_dev, uint64_t* total_host, void* stream) {
  TCB_CHECK_ARG(width > 0 || n_elements == 0, "TcbBytesSerializeFixedWidth: width must be positive");
  return serialize_common(device_id, rows_dev, nullptr, 4, true, width, n_elements, dst_dev, dst_capacity,
                          workspace_dev, workspace_bytes, total_dev, total_host, stream);
}

// mode 0: the whole call; 1: count only (everything up to the verified element count); 2: fill only
// (offsets + gather, on a workspace that a mode-1 call with the same stream left behind)
enum : int { DES_ALL = 0, DES_COUNT = 1, DES_FILL = 2 };

static int des_impl(int mode, int device_id, const void* stream_dev, uint64_t stream_bytes,
                    int64_t max_elements, int64_t* offsets_dev, uint64_t offsets_capacity,
                    void* payload_dev, uint64_t payload_capacity, void* workspace_dev,
                    uint64_t workspace_bytes, uint64_t* n_elements_host,
                    uint64_t* payload_bytes_host, void* stream) {
  TCB_CHECK_ARG(n_elements_host != nullptr, "TcbBytesDeserialize: n_elements_host is NULL");
  *n_elements_host = 0;
  if (payload_bytes_host) *payload_bytes_host = 0;
  TCB_DEVICE_GUARD(device_id);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  TCB_CHECK_ARG(mode == DES_COUNT || offsets_dev != nullptr, "TcbBytesDeserialize: offsets is NULL");
  if (max_elements == 0 || (stream_bytes == 0 && max_elements < 0)) {
    if (mode == DES_COUNT) return TCB_OK;
    TCB_CUDA(cudaMemsetAsync(offsets_dev, 0, 8, s));
    TCB_CUDA(cudaStreamSynchronize(s));
    return TCB_OK;
  }
  if (stream_bytes == 0)
    return set_error(TCB_ERR_TRUNCATED, "TcbBytesDeserialize: empty buffer but %lld elements requested",
                     (long long)max_elements);
  TCB_CHECK_ARG(stream_dev != nullptr && workspace_dev != nullptr, "TcbBytesDeserialize: NULL stream/workspace");
  if (mode == DES_ALL && stream_bytes <= (uint64_t)kDesChunk && workspace_bytes >= 256 &&
      reinterpret_cast<uintptr_t>(workspace_dev) % 8 == 0) {  // one chunk: one launch, one round trip
    unsigned long long* th = reinterpret_cast<unsigned long long*>(workspace_dev);
    k_des_tiny<<<1, kTinyThreads, 0, s>>>(static_cast<const uint8_t*>(stream_dev), (unsigned int)stream_bytes,
                                          (long long)max_elements, reinterpret_cast<long long*>(offsets_dev),
                                          offsets_capacity, static_cast<uint8_t*>(payload_dev), payload_capacity, th);
    count_launch();
    unsigned long long h[DW_HEADER_WORDS];
    TCB_CUDA(cudaMemcpyAsync(h, th, sizeof(h), cudaMemcpyDeviceToHost, s));
    TCB_CUDA(cudaStreamSynchronize(s));
    if (h[DW_STATUS] == 2)
      return set_error(TCB_ERR_TRUNCATED, "TcbBytesDeserialize: buffer ends inside an element after %llu elements",
                       (unsigned long long)h[DW_NELEM]);
    if (h[DW_STATUS] == 3)
      return set_error(TCB_ERR_OUT_OF_RANGE, "TcbBytesDeserialize: %llu elements exceed offsets capacity %llu",
                       (unsigned long long)h[DW_NELEM], (unsigned long long)offsets_capacity);
    if (h[DW_STATUS] == 4)
      return set_error(TCB_ERR_OUT_OF_RANGE, "TcbBytesDeserialize: payload of %llu bytes exceeds capacity %llu",
                       (unsigned long long)h[DW_PAYLOAD], (unsigned long long)payload_capacity);
    *n_elements_host = h[DW_NELEM];
    if (payload_bytes_host) *payload_bytes_host = h[DW_PAYLOAD];
    return TCB_OK;
  }
  const uint64_t need = TcbBytesDeserializeWorkspace(stream_bytes);
  if (workspace_bytes < need)
    return set_error(TCB_ERR_OUT_OF_RANGE, "TcbBytesDeserialize: workspace needs %llu bytes, got %llu",
                     (unsigned long long)need, (unsigned long long)workspace_bytes);
  TCB_CHECK_ARG(reinterpret_cast<uintptr_t>(workspace_dev) % 256 == 0, "TcbBytesDeserialize: workspace not 256-byte aligned");
  const uint64_t K = des_chunks(stream_bytes);
  TCB_CHECK_ARG(K < (1ull << 31), "TcbBytesDeserialize: stream too large");
  uint8_t* w = static_cast<uint8_t*>(workspace_dev);
  unsigned long long* hdr = reinterpret_cast<unsigned long long*>(w);
  DesArrays A;
  const uint64_t K1 = K + 1;
  unsigned long long* sws = reinterpret_cast<unsigned long long*>(w + 256);  // scan workspace
  const uint64_t sws_bytes = round_up((WS_HEADER_WORDS + K1 / kResTile + 2) * 8, 256);
  uint8_t* p = w + 256 + sws_bytes;
  A.g = reinterpret_cast<unsigned long long*>(p); p += round_up(K1 * 8, 256);
  A.x = reinterpret_cast<unsigned long long*>(p); p += round_up(K1 * 8, 256);
  A.base = reinterpret_cast<unsigned long long*>(p); p += round_up(K1 * 8, 256);
  A.cnt = reinterpret_cast<unsigned int*>(p); p += round_up(K1 * 4, 256);
  A.err = reinterpret_cast<unsigned int*>(p); p += round_up(K1 * 4, 256);
  A.list = reinterpret_cast<unsigned short*>(p);
  const uint8_t* st = static_cast<const uint8_t*>(stream_dev);
  // TCB_TIMING=1: per-phase device times on stderr (development aid; adds event records only)
  static const bool timing = getenv("TCB_TIMING") != nullptr;
  std::vector<std::pair<const char*, cudaEvent_t>> marks;
  auto mark = [&](const char* name) {
    if (!timing) return;
    cudaEvent_t e;
    cudaEventCreate(&e);
    cudaEventRecord(e, s);
    marks.emplace_back(name, e);
  };
  auto report = [&]() {
    if (!timing || marks.size() < 2) return;
    cudaEventSynchronize(marks.back().second);
    for (size_t i = 1; i < marks.size(); i++) {
      float ms = 0;
      cudaEventElapsedTime(&ms, marks[i - 1].second, marks[i].second);
      fprintf(stderr, "[tcb] deserialize %-10s %.3f ms\n", marks[i].first, ms);
    }
    for (auto& m : marks) cudaEventDestroy(m.second);
  };
  mark("start");
  unsigned long long h[DW_HEADER_WORDS];
  if (mode == DES_FILL) {  // the verified header of the preceding count call
    TCB_CUDA(cudaMemcpyAsync(h, hdr, sizeof(h), cudaMemcpyDeviceToHost, s));
    TCB_CUDA(cudaStreamSynchronize(s));
  } else {
  TCB_CUDA(cudaMemsetAsync(hdr, 0, 256, s));
  {
    const unsigned int bias = (unsigned int)(reinterpret_cast<uintptr_t>(st) & 3);
    mark("setup");
    k_des_vote<<<(unsigned)((K + kVoteWarps - 1) / kVoteWarps), kVoteWarps * 32, 0, s>>>(st - bias, bias, stream_bytes, K, A);
    mark("vote");
    k_des_chase<<<(unsigned)((K + kDesThreads - 1) / kDesThreads), kDesThreads, 0, s>>>(st - bias, bias, stream_bytes, K, A, 0);
    mark("chase");
    count_launch(1);
  }
  auto resolve = [&]() -> cudaError_t {
    // first mismatch / first error start at ~0 (atomicMin), the scan workspace at 0
    cudaMemsetAsync(hdr + DW_FIRST, 0xFF, 8, s);
    cudaMemsetAsync(hdr + DW_FERR, 0xFF, 8, s);
    cudaMemsetAsync(sws, 0, sws_bytes, s);
    k_des_verify<<<(unsigned)((K + kResThreads - 1) / kResThreads), kResThreads, 0, s>>>(K, A, hdr);
    k_des_scan<<<(unsigned)((K + kResTile - 1) / kResTile), kResThreads, 0, s>>>(K, A, hdr, sws);
    k_des_finalize<<<1, 32, 0, s>>>(stream_bytes, K, max_elements, A, hdr);
    count_launch(3);
    return cudaGetLastError();
  };
  TCB_CUDA(resolve());
  mark("resolve");
  count_launch(1);
  TCB_CUDA(cudaMemcpyAsync(h, hdr, sizeof(h), cudaMemcpyDeviceToHost, s));
  TCB_CUDA(cudaStreamSynchronize(s));
  // a guess was wrong (elements longer than the 512-byte window, payloads that look like headers):
  // vote again with a wider window on the chunks that do not fit their predecessor, then whatever is
  // still inconsistent is repaired by the sequential sweep
  // (a third round with a 128 KiB window was measured: 20 ms per 256 MiB, slower than the sweep it saves)
  static const unsigned long long kWide[] = {8ull << 10};
  for (int r = 0; r < 1 && h[DW_STATUS] == 1; r++) {
    const unsigned int bias = (unsigned int)(reinterpret_cast<uintptr_t>(st) & 3);
    k_des_vote_wide<<<(unsigned)((K + kVoteWarps - 1) / kVoteWarps), kVoteWarps * 32, 0, s>>>(st - bias, bias, stream_bytes, K, A, hdr, kWide[r]);
    k_des_chase<<<(unsigned)((K + kDesThreads - 1) / kDesThreads), kDesThreads, 0, s>>>(st - bias, bias, stream_bytes, K, A, 1);
    mark("wide vote");
    TCB_CUDA(resolve());
    count_launch(2);
    TCB_CUDA(cudaMemcpyAsync(h, hdr, sizeof(h), cudaMemcpyDeviceToHost, s));
    TCB_CUDA(cudaStreamSynchronize(s));
  }
  if (h[DW_STATUS] == 1) {
    k_des_sweep<<<1, 32, 0, s>>>(st, stream_bytes, K, max_elements, A, hdr);
    mark("sweep");
    TCB_CUDA(resolve());
    mark("resolve2");
    count_launch(1);
    TCB_CUDA(cudaMemcpyAsync(h, hdr, sizeof(h), cudaMemcpyDeviceToHost, s));
    TCB_CUDA(cudaStreamSynchronize(s));
  }
  }  // mode != DES_FILL
  if (h[DW_STATUS] == 2)
    return set_error(TCB_ERR_TRUNCATED, "TcbBytesDeserialize: buffer ends inside an element after %llu elements",
                     (unsigned long long)h[DW_NELEM]);
  if (h[DW_STATUS] != 0)
    return set_error(TCB_ERR_INTERNAL, "TcbBytesDeserialize: chunk chain did not resolve (status %llu)",
                     (unsigned long long)h[DW_STATUS]);
  const uint64_t N = h[DW_NELEM];
  if (mode == DES_COUNT) {
    *n_elements_host = N;
    // every decoded element spends 4 bytes of the stream on its prefix: an upper bound (exact when the
    // whole buffer is decoded) of the payload, for the caller to size its buffers with
    if (payload_bytes_host) *payload_bytes_host = stream_bytes >= 4 * N ? stream_bytes - 4 * N : 0;
    report();
    return TCB_OK;
  }
  if (N > offsets_capacity)
    return set_error(TCB_ERR_OUT_OF_RANGE, "TcbBytesDeserialize: %llu elements exceed offsets capacity %llu",
                     (unsigned long long)N, (unsigned long long)offsets_capacity);
  *n_elements_host = N;
  if (N == 0) {
    TCB_CUDA(cudaMemsetAsync(offsets_dev, 0, 8, s));
    TCB_CUDA(cudaStreamSynchronize(s));
    return TCB_OK;
  }
  mark("host");
  k_des_offsets<<<(unsigned)h[DW_KEND], 64, 0, s>>>(st, stream_bytes, A, hdr, reinterpret_cast<long long*>(offsets_dev));
  mark("offsets");
  count_launch();
  long long total = 0;
  TCB_CUDA(cudaMemcpyAsync(&total, offsets_dev + N, 8, cudaMemcpyDeviceToHost, s));
  TCB_CUDA(cudaStreamSynchronize(s));
  if (payload_bytes_host) *payload_bytes_host = (uint64_t)total;
  if (total > 0) {
    TCB_CHECK_ARG(payload_dev != nullptr, "TcbBytesDeserialize: payload is NULL");
    if ((uint64_t)total > payload_capacity)
      return set_error(TCB_ERR_OUT_OF_RANGE, "TcbBytesDeserialize: payload of %lld bytes exceeds capacity %llu",
                       total, (unsigned long long)payload_capacity);
    uint8_t* blob = static_cast<uint8_t*>(payload_dev);
    const unsigned tiles = (unsigned)((N + kGatTile - 1) / kGatTile);
    const bool word_ok = ((reinterpret_cast<uintptr_t>(st) ^ reinterpret_cast<uintptr_t>(blob)) & 3) == 0;
    if ((uint64_t)total / N >= 2048) {  // long elements: cut the work by bytes, not by elements
      const unsigned segs = (unsigned)(((uint64_t)total + kLongSeg - 1) / kLongSeg);
      if (word_ok)
        k_des_gather_long<true><<<segs, 256, 0, s>>>(st, reinterpret_cast<long long*>(offsets_dev), hdr, blob, payload_capacity);
      else
        k_des_gather_long<false><<<segs, 256, 0, s>>>(st, reinterpret_cast<long long*>(offsets_dev), hdr, blob, payload_capacity);
    } else if (word_ok)
      k_des_gather<true><<<tiles, kGatThreads, 0, s>>>(st, reinterpret_cast<long long*>(offsets_dev), hdr, blob, payload_capacity);
    else
      k_des_gather<false><<<tiles, kGatThreads, 0, s>>>(st, reinterpret_cast<long long*>(offsets_dev), hdr, blob, payload_capacity);
    count_launch();
  }
  mark("gather");
  TCB_CUDA(cudaGetLastError());
  TCB_CUDA(cudaStreamSynchronize(s));
  report();
  return TCB_OK;
}

int TcbBytesDeserialize(int device_id, const void* stream_dev, uint64_t stream_bytes,
                        int64_t max_elements, int64_t* offsets_dev, uint64_t offsets_capacity,
                        void* payload_dev, uint64_t payload_capacity, void* workspace_dev,
                        uint64_t workspace_bytes, uint64_t* n_elements_host,
                        uint64_t* payload_bytes_host, void* stream) {
  return des_impl(DES_ALL, device_id, stream_dev, stream_bytes, max_elements, offsets_dev, offsets_capacity,
                  payload_dev, payload_capacity, workspace_dev, workspace_bytes, n_elements_host,
                  payload_bytes_host, stream);
}

int TcbBytesDeserializeCount(int device_id, const void* stream_dev, uint64_t stream_bytes,
                             int64_t max_elements, void* workspace_dev, uint64_t workspace_bytes,
                             uint64_t* n_elements_host, uint64_t* payload_bound_host, void* stream) {
  return des_impl(DES_COUNT, device_id, stream_dev, stream_bytes, max_elements, nullptr, 0, nullptr, 0,
                  workspace_dev, workspace_bytes, n_elements_host, payload_bound_host, stream);
}

int TcbBytesDeserializeFill(int device_id, const void* stream_dev, uint64_t stream_bytes,
                            int64_t max_elements, int64_t* offsets_dev, uint64_t offsets_capacity,
                            void* payload_dev, uint64_t payload_capacity, void* workspace_dev,
                            uint64_t workspace_bytes, uint64_t* n_elements_host,
                            uint64_t* payload_bytes_host, void* stream) {
  return des_impl(DES_FILL, device_id, stream_dev, stream_bytes, max_elements, offsets_dev, offsets_capacity,
                  payload_dev, payload_capacity, workspace_dev, workspace_bytes, n_elements_host,
                  payload_bytes_host, stream);
}

}  // extern "C"

// =================================================================================================
// Serialise without allocating anything (see TcbBytesSerializeNoAlloc in include/tcb200.h)
// =================================================================================================
#include <mutex>

namespace tcb {

// Most elements one call may serialise. The look-back scan needs one 64-bit status word per tile
// of 1024 elements plus an 8-word header, so this bounds the scratch below at 131 144 bytes.
constexpr uint64_t kNoAllocMaxElements = 16ull << 20;
constexpr uint64_t kNoAllocScratchWords = WS_HEADER_WORDS + kNoAllocMaxElements / kSerTile + 1;

// The scan scratch of the no-alloc entry point. It is part of the library image: every GPU that
// loads libtcb200.so carries these 128 KiB from module load until the process exits, whether or
// not the entry point is ever called. In HOST code this name is only a stand-in for the symbol;
// the device address comes from cudaGetSymbolAddress below and that is what kernels receive.
static __device__ __align__(256) unsigned long long g_noalloc_scan_scratch[kNoAllocScratchWords];

// Device address of the array above, looked up once per device with that device current.
static int noalloc_scratch_for(int device_id, unsigned long long** out) {
  static std::mutex mu;
  static unsigned long long* addr[64] = {nullptr};
  TCB_CHECK_ARG(device_id >= 0 && device_id < 64, "TcbBytesSerializeNoAlloc: bad device %d", device_id);
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[device_id]) {
    TCB_DEVICE_GUARD(device_id);
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scan_scratch);
    if (e != cudaSuccess || p == nullptr)
      return set_error(TCB_ERR_CUDA, "TcbBytesSerializeNoAlloc: cudaGetSymbolAddress failed on device %d: %s",
                       device_id, cudaGetErrorString(e));
    addr[device_id] = static_cast<unsigned long long*>(p);
  }
  *out = addr[device_id];
  return TCB_OK;
}

}  // namespace tcb

extern "C" int TcbBytesSerializeNoAlloc(int device_id, const void* payload_dev, const void* lengths_dev,
                                        int length_size, uint64_t n_elements, void* dst_dev,
                                        uint64_t dst_capacity, uint64_t* total_host, void* stream) {
  if (n_elements > tcb::kNoAllocMaxElements)
    return tcb::set_error(TCB_ERR_OUT_OF_RANGE,
                          "TcbBytesSerializeNoAlloc: %llu elements exceed the built-in scratch, which covers at "
                          "most %llu elements per call; split the tensor or use TcbBytesSerialize with a "
                          "caller-provided workspace",
                          (unsigned long long)n_elements, (unsigned long long)tcb::kNoAllocMaxElements);
  unsigned long long* scratch = nullptr;
  const int rc = tcb::noalloc_scratch_for(device_id, &scratch);
  if (rc) return rc;
  // serialize_common zeroes the part of the scratch this call uses (header + one status word per
  // tile) on `stream` before the kernel, so whatever the previous call left behind is gone.
  return serialize_common(device_id, payload_dev, lengths_dev, length_size, false, 0, n_elements, dst_dev,
                          dst_capacity, scratch, tcb::kNoAllocScratchWords * 8, nullptr, total_host, stream);
}
