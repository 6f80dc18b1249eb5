// lookup16.cu — the pyramid window lookup on the 16-bit blocked pyramid (raftcorr_level16), sm_100a.
//
// Same arithmetic as lookup.cu / lookup_tma.cu (reference core/corr.py:29-50 + core/utils/utils.py:57-65 +
// ATen grid_sampler_2d, replayed op for op), different decomposition.  Measured on the fp32 kernel
// (profiles/r01_*): one warp per query spends ~635 warp-instructions per query, most of them per-query
// bookkeeping that only one or a few lanes really need (tap tables, window geometry, TMA issue, barrier
// traffic, the shared-memory transpose of the output), and the kernel is bound by issue slots, not by DRAM.
// Here a LANE owns a query:
//
//   unit      = 32 consecutive-x queries of one fmap row at ONE pyramid level (4 units per tile)
//   warp      = owns one unit buffer and walks every other tile of its level (8 warps: 4 levels x 2).  Lane q
//               derives its query's window ({2,3} x {3,4} whole 64-byte blocks; out-of-image blocks arrive as
//               zeros = grid_sample's zero padding).  Neighbouring queries look at nearly the same place of
//               neighbouring images, so the four lanes of a group take the union of their windows and the
//               group's first lane requests ONE box over the four images (cp.async.bulk.tensor.5d); a group
//               whose union does not fit a box shape (flow discontinuity, image border of the batch) falls back
//               to one box per lane.  All boxes of a unit arrive on one mbarrier.  Then lane q forms its
//               query's (2r+1)^2 samples out of its part of the buffer and stores them straight to the NCHW
//               output — for a fixed channel the 32 lanes are 32 consecutive x, so every store instruction is
//               one full 128-byte line and the output never passes through shared memory.  Then the warp refills
//               its buffer with its next unit; the other seven warps keep the SM busy meanwhile.
//   levels    = plain ([image][block row][block][64 B]) or interleaved ([group of 4 images][block row][block]
//               [image][64 B], raftcorr_level16.interleave_log2): with the latter a group's box reads pieces of
//               512-768 contiguous bytes instead of four pieces of 128-192 bytes 14 KB apart.  Measured with
//               tools/probe/tma_latency_probe.cu: the scattered form tops out at 3.4 TB/s and, at this kernel's
//               load, makes a warp wait ~5100 cycles for its unit; interleaved 4.8 TB/s and ~2400 cycles.
//
// Where the time goes (tools/trace16.py, cycles per unit and warp at BASELINE.json configs[1], smooth flow):
// prepare (taps, windows of this and the next unit) ~1600, wait for the boxes ~1750, sample ~3800, issue ~950; a
// warp issues a box only every ~70-110 cycles whatever the box holds (operands travel through uniform registers
// one elected lane at a time), which is why boxes are shared by groups.  Measured alternatives, all slower:
// dedicated producer warps; two warps per unit splitting the x taps (the per-unit work is paid twice: 35.7 us
// against 33.1); staging the patch in registers to refill early; L2 prefetch of the next unit (a prefetch costs
// the same issue time as the load); 12 warps with packed 18.5 KB buffers; groups of 2 or 8 (same time as 4).
//
// Shared memory: 8 unit buffers of 28 KB.  The 32 windows are walked in lock step, at the same in-window offset
// more often than not, and TMA wants 128-byte aligned destinations, so without a skew all lanes would hit one
// bank.  Boxes are therefore fetched with a TMA swizzle (64-byte swizzle: 16-byte chunk index bits 4-5 ^= address
// bits 7-8; measured layout: tools/probe/tma_swizzle_probe.cu — with 64-byte rows the 128-byte swizzle would pad
// every block to 128 bytes; interleaved group boxes have 128-byte rows and use the 128-byte swizzle, bits 4-6 ^=
// bits 7-9) into regions whose pitch is an odd multiple of 128 B, so that neighbouring groups start in
// different swizzle phases.
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <mutex>

#include "lookup_common.cuh"
#include "tma_util.cuh"

namespace raftcorr {

namespace {

using namespace lk;

constexpr int kL = RAFTCORR_MAX_LEVELS;           // the kernel is specialised for the 4-level pyramid
#ifndef RAFTCORR_L16_UNROLL
#define RAFTCORR_L16_UNROLL 2
#endif
constexpr int kUnroll16 = RAFTCORR_L16_UNROLL;     // unroll factor of the x-tap loop
constexpr int kWarps16 = 8;                        // per level: 2 warps, each with its own unit buffer
constexpr int kThreads16 = kWarps16 * 32;
constexpr int kDepth16 = 2;                        // unit buffers per level (each level runs its own ring)
constexpr int kBufs = kL * kDepth16;
constexpr int kBoxMax = 768;                       // the largest box of one query: 3 x 4 blocks
constexpr int kSlot = kBoxMax + 128;               // slot pitch of ungrouped queries: an odd multiple of 128 B
constexpr int kBufBytes = 32 * kSlot;              // 28 KB
// Queries whose images are neighbours in memory (consecutive x) and whose windows share one block-aligned box are
// fetched by ONE tensor box over 2^g images: a warp issues a box only every ~70 cycles (the operands travel
// through uniform registers, one elected lane at a time), and 32 boxes per unit were a quarter of a unit's time.
// Group size per level, log2 (neighbouring queries are 1, 1/2, 1/4, 1/8 level pixel apart):
#ifndef RAFTCORR_L16_GROUPS
#define RAFTCORR_L16_GROUPS 0x2222
#endif
constexpr int kGroups16 = RAFTCORR_L16_GROUPS;     // hex digit l = log2 of the group size at level l (0..3)
__host__ __device__ constexpr int group_log2(int level) { return (kGroups16 >> (4 * level)) & 15; }
// a group's region: its queries' boxes back to back (pitch = the box bytes) or, ungrouped, at kBoxMax; the region
// pitch is an odd multiple of 128 B so that neighbouring groups start in different swizzle phases
__host__ __device__ constexpr int group_pitch(int glog) { return (kBoxMax << glog) + 128; }
static_assert((32 >> group_log2(0)) * group_pitch(group_log2(0)) <= kBufBytes && group_log2(0) <= 3, "level 0 groups");
static_assert((32 >> group_log2(1)) * group_pitch(group_log2(1)) <= kBufBytes && group_log2(1) <= 3, "level 1 groups");
static_assert((32 >> group_log2(2)) * group_pitch(group_log2(2)) <= kBufBytes && group_log2(2) <= 3, "level 2 groups");
static_assert((32 >> group_log2(3)) * group_pitch(group_log2(3)) <= kBufBytes && group_log2(3) <= 3, "level 3 groups");
constexpr int kOffBars16 = kBufs * kBufBytes;
constexpr int kSmem16 = kOffBars16 + kBufs * 8;

struct Level16Dev {
  const unsigned char* data;
  int H, W, bxs, bys;
  int ilog;                                         // the blocks of 2^ilog neighbouring images are interleaved (0: not)
  long long istride;                                // bytes of one image's blocks
  float sm1x, invx, halfx, sm1y, invy, halfy, inv_pow2;
};

#ifdef RAFTCORR_L16_TRACE
// development build only (tools/trace16.py): cycles each warp spends per phase, summed over its units
__device__ unsigned long long g_trace16[148 * 8 * 32];   // [SM][warp]: 4 phase sums, units, whole warp, -, -, then 8 rounds x (wait, sample, issue)
#define L16_TICK(var) const long long var = clock64()
#else
#define L16_TICK(var)
#endif

struct Lookup16Params {
  Level16Dev lv[kL];
  const float* coords;
  float* out;
  int B, H, W, tiles_x;
  long long tiles;
  int debug;   // 0 in product builds; -DRAFTCORR_L16_DEV reads RAFTCORR_L16_DEBUG: bit 0 = skip the sampling (fill path only); bits 8.. = start skew
};

// box shapes: [level][one image / the level's group of images][wide: 2 or 3 blocks][tall: 3 or 4 block rows]
struct alignas(64) Lookup16Maps {
  CUtensorMap m[kL][2][2][2];
};

// slot of lane q in a unit buffer (a permutation of 0..31 per level; skew class = slot & 3): class mates are
// 8 lanes whose queries are 1, 2, 4 (4) fmap pixels apart at level 0, 1, 2 (3), i.e. ~one level pixel
__device__ __forceinline__ int slot_of(int lane, int level) {
  if (level == 0) return ((lane & 7) << 2) | (lane >> 3);
  if (level == 1) return (((lane >> 1) & 7) << 2) | (((lane >> 4) & 1) << 1) | (lane & 1);
  return lane;
}
// physical shared-memory address of logical byte `a` of a box fetched with CU_TENSOR_MAP_SWIZZLE_64B
// (mask 0x30); a group's box out of an interleaved level has 128-byte rows (the blocks of two neighbouring images) and
// is fetched with the 128-byte swizzle (chunk bits 4-6 ^= address bits 7-9, mask 0x70): eight phases for eight groups
__device__ __forceinline__ uint32_t swz(uint32_t a, uint32_t mask) { return a ^ ((a >> 3) & mask); }

// window of one query at one level, shared by producer and sampler (same code, same result)
struct Window {
  int bx0, by0;   // first block column / row of the box (may lie outside the image)
  int wide, tall; // 0/1: box is 2+wide blocks x 3+tall block rows
  int fast;       // all taps inside such a box (false only for |coords| so large that fp32 spacing breaks the pattern)
  int grouped;    // bx0/by0/wide/tall describe the box shared by the lane's whole group
};

__device__ __forceinline__ Window window_from(int x_lo, int x_hi, int y_lo, int y_hi) {
  Window w;
  w.bx0 = x_lo >> 3;
  w.by0 = y_lo >> 2;
  const int nbx = (x_hi >> 3) - w.bx0 + 1, nby = (y_hi >> 2) - w.by0 + 1;
  w.fast = (nbx >= 1) & (nbx <= 3) & (nby >= 1) & (nby <= 4);
  w.wide = nbx > 2;
  w.tall = nby > 3;
  w.grouped = 0;
  return w;
}
// the box that holds the windows of all 2^glog lanes of a group, if one of the four shapes does
__device__ __forceinline__ Window group_window(Window w, bool usable, int glog) {
  int x0 = w.bx0, y0 = w.by0, x1 = w.bx0 + 1 + w.wide, y1 = w.by0 + 2 + w.tall;
  bool ok = usable && w.fast;
#pragma unroll
  for (int s = 0; s < 3; ++s)
    if (s < glog) {
      const int d = 1 << s;
      x0 = min(x0, __shfl_xor_sync(0xffffffffu, x0, d));
      y0 = min(y0, __shfl_xor_sync(0xffffffffu, y0, d));
      x1 = max(x1, __shfl_xor_sync(0xffffffffu, x1, d));
      y1 = max(y1, __shfl_xor_sync(0xffffffffu, y1, d));
      ok = ok & (__shfl_xor_sync(0xffffffffu, (int)ok, d) != 0);
    }
  if (glog > 0 && ok && x1 - x0 <= 2 && y1 - y0 <= 3) {
    w.bx0 = x0;
    w.by0 = y0;
    w.wide = x1 - x0 > 1;
    w.tall = y1 - y0 > 2;
    w.grouped = 1;
  }
  return w;
}
// first and last tap decide the box (taps are monotone in their offset)
template <int R, int kMode>
__device__ __forceinline__ Window window_of(float cxl, float cyl, const Level16Dev& lv) {
  constexpr int rd = 2 * R + 1;
  const int x_lo = make_tap_fast(cxl, (float)(-R), lv.sm1x, lv.invx, lv.halfx, kMode).i0;
  const int x_hi = make_tap_fast(cxl, (float)(rd - 1 - R), lv.sm1x, lv.invx, lv.halfx, kMode).i0 + 1;
  const int y_lo = make_tap_fast(cyl, (float)(-R), lv.sm1y, lv.invy, lv.halfy, kMode).i0;
  const int y_hi = make_tap_fast(cyl, (float)(rd - 1 - R), lv.sm1y, lv.invy, lv.halfy, kMode).i0 + 1;
  return window_from(x_lo, x_hi, y_lo, y_hi);
}

__device__ __forceinline__ float bf16_bits_to_float(uint32_t h) { return __uint_as_float(h << 16); }

// rare path: straight from global memory, with bounds checks (zero padding)
__device__ __noinline__ float pixel16_global(const Level16Dev& lv, long long img, int y, int x) {
  if (x < 0 || y < 0 || x >= lv.W || y >= lv.H) return 0.0f;
  const long long member = img & ((1LL << lv.ilog) - 1), block = (long long)(y >> 2) * lv.bxs + (x >> 3);
  const unsigned char* p = lv.data + (img - member) * lv.istride + ((block << lv.ilog) + member) * 64 + (y & 3) * 16 +
                           (x & 7) * 2;
  return bf16_bits_to_float(*reinterpret_cast<const unsigned short*>(p));
}

// One query's samples at one level out of its part of the unit buffer.
//
// Tap j of an axis sits on pixel m + j + e_j with m = min_j(i0_j - j).  Normally every e_j is 0: the footprint is a
// dense (rd+1) x (rd+1) patch, each pixel is loaded once, and the right-hand column of x tap i is carried over as the
// left-hand column of tap i+1 (sample_unit_dense2 below).  On the integer grid (RAFT's first iteration) the
// reference's coordinate round trip moves floor() down by one on ~8 % of the taps: e_j is then 0 or 1 per lane and
// tap, the patch grows by one row and one column, it is still loaded once, and which of two neighbouring rows /
// columns a tap uses is a per-lane select on registers (sample_unit_flips).
//
// The x-tap loop is a real loop (the taps are recomputed from the loop counter, ~14 instructions): fully unrolled,
// the variants of these routines came to 10k SASS instructions and the warps of an SM — each in a different variant
// or phase — spent 40 % of their issue slots waiting for the instruction cache.
template <int R, int kMode>
__device__ __forceinline__ void sample_unit_flips(uint32_t slot_addr, uint32_t swz_mask, int cshift, int bw, int bh, int bx0,
                                                  int by0, int mx, int my, float cxl, const Level16Dev& lv,
                                                  const Tap (&ty)[2 * R + 1], float* optr, long long N, bool live) {
  constexpr int rd = 2 * R + 1, kRows = rd + 2;
  const int rowpitch = bw << cshift;   // bytes per block row of the box (a block column is 1 << cshift bytes wide)
  auto px = [&](uint32_t row, uint32_t col) -> float {
    unsigned short v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(swz(row + col, swz_mask)));
    return bf16_bits_to_float(v);
  };
  auto combine = [&](float vnw, float vne, float vsw, float vse, float w_, float e_, float n_, float s_) {
    // ATen's order (common.cuh bilinear_combine), with e = 1 - w and s = 1 - n hoisted
    const float nw = __fmul_rn(s_, e_), ne = __fmul_rn(s_, w_), sw = __fmul_rn(n_, e_), se = __fmul_rn(n_, w_);
    return __fmaf_rn(vse, se, __fmaf_rn(vsw, sw, __fmaf_rn(vne, ne, __fmul_rn(vnw, nw))));
  };
  // Patch rows / columns that no tap of this lane uses may fall outside the box: their addresses are clamped into
  // the lane's part of the buffer; the values are loaded but never selected.
  uint32_t rows[kRows];
#pragma unroll
  for (int r = 0; r < kRows; ++r) {
    const int yy = max(0, min(my + r - 4 * by0, 4 * bh - 1));
    rows[r] = slot_addr + (uint32_t)((yy >> 2) * rowpitch + (yy & 3) * 16);
  }
  auto col_off = [&](int c) {
    const int xx = max(0, min(mx + c - 8 * bx0, 8 * bw - 1));
    return (uint32_t)(((xx >> 3) << cshift) + (xx & 7) * 2);
  };
  float fy[rd], sy[rd];
  bool ey[rd];
#pragma unroll
  for (int j = 0; j < rd; ++j) {
    fy[j] = ty[j].frac;
    sy[j] = __fsub_rn(1.0f, fy[j]);
    ey[j] = (ty[j].i0 - j) != my;
  }
  float c0[kRows], c1[kRows], c2[kRows];
  {
    const uint32_t a = col_off(0), b = col_off(1);
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
      c0[r] = px(rows[r], a);
      c1[r] = px(rows[r], b);
    }
  }
  float* o = optr;
#pragma unroll kUnroll16
  for (int i = 0; i < rd; ++i) {
    const Tap t = make_tap_fast(cxl, (float)(i - R), lv.sm1x, lv.invx, lv.halfx, kMode);
    const float fx = t.frac, ex = __fsub_rn(1.0f, fx);
    const uint32_t c = col_off(i + 2);
#pragma unroll
    for (int r = 0; r < kRows; ++r) c2[r] = px(rows[r], c);
    const bool flip = (t.i0 - i) != mx;
    float lft[kRows], rgt[kRows];
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
      lft[r] = flip ? c1[r] : c0[r];
      rgt[r] = flip ? c2[r] : c1[r];
    }
    // outputs j and j + 1 share the packed fp32x2 instructions (the same IEEE operations, two at a time)
    const float2 e2 = make_float2(ex, ex), w2 = make_float2(fx, fx);
#pragma unroll
    for (int j = 0; j + 1 < rd; j += 2) {
      const float2 tl = make_float2(ey[j] ? lft[j + 1] : lft[j], ey[j + 1] ? lft[j + 2] : lft[j + 1]);
      const float2 bl = make_float2(ey[j] ? lft[j + 2] : lft[j + 1], ey[j + 1] ? lft[j + 3] : lft[j + 2]);
      const float2 tr = make_float2(ey[j] ? rgt[j + 1] : rgt[j], ey[j + 1] ? rgt[j + 2] : rgt[j + 1]);
      const float2 br = make_float2(ey[j] ? rgt[j + 2] : rgt[j + 1], ey[j + 1] ? rgt[j + 3] : rgt[j + 2]);
      const float2 n2 = make_float2(fy[j], fy[j + 1]), s2 = make_float2(sy[j], sy[j + 1]);
      float2 acc = __fmul2_rn(tl, __fmul2_rn(s2, e2));
      acc = __ffma2_rn(tr, __fmul2_rn(s2, w2), acc);
      acc = __ffma2_rn(bl, __fmul2_rn(n2, e2), acc);
      acc = __ffma2_rn(br, __fmul2_rn(n2, w2), acc);
      if (live) {
        __stcs(o, acc.x);
        __stcs(o + N, acc.y);
      }
      o += 2 * N;
    }
    {
      constexpr int j = rd - 1;
      const float tl = ey[j] ? lft[j + 1] : lft[j], bl = ey[j] ? lft[j + 2] : lft[j + 1];
      const float tr = ey[j] ? rgt[j + 1] : rgt[j], br = ey[j] ? rgt[j + 2] : rgt[j + 1];
      const float v = combine(tl, tr, bl, br, fx, ex, fy[j], sy[j]);
      if (live) __stcs(o, v);
      o += N;
    }
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
      c0[r] = c1[r];
      c1[r] = c2[r];
    }
  }
}

// The dense footprint, two outputs per floating-point instruction: sm_100's packed fp32
// operations (fma.rn.f32x2 & co.: each half is the IEEE operation, so the bits do not change) halve the issue slots
// of the weight products and the four-term sum, which are 45 % of this routine.  Outputs j and j + 2 of an x tap
// share an instruction: their upper-left pixels are patch rows (j, j + 2) and their lower-left ones (j + 1, j + 3),
// which is also the upper-left pair of outputs (j + 1, j + 3) — only rows 2, 4, 6 are needed in two pairs.
template <int R, int kMode, bool kAllLive>
__device__ __forceinline__ void sample_unit_dense2(uint32_t slot_addr, uint32_t swz_mask, int cshift, int bw, int bx0, int by0, int mx, int my,
                                                   float cxl, const Level16Dev& lv, const Tap (&ty)[2 * R + 1],
                                                   int i_begin, int i_end, float* optr, unsigned N, bool live) {
  constexpr int rd = 2 * R + 1, kRows = rd + 1;
  constexpr int kPairs = (rd - 1) / 2;                    // rd = 9: (0,2) (1,3) (4,6) (5,7) + 8;  rd = 7: (0,2) (1,3) (4,6) + 5
  constexpr int kSingle = rd == 9 ? 8 : 5;
  static_assert(rd == 9 || rd == 7, "pairing tables cover radius 3 and 4");
  auto lo_of = [](int p) { return 4 * (p >> 1) + (p & 1); };
  const int rowpitch = bw << cshift;
  auto px = [&](uint32_t row, uint32_t col) -> float {
    unsigned short v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(swz(row + col, swz_mask)));
    return bf16_bits_to_float(v);
  };
  uint32_t rows[kRows];
#pragma unroll
  for (int r = 0; r < kRows; ++r) {
    const int yy = my + r - 4 * by0;
    rows[r] = slot_addr + (uint32_t)((yy >> 2) * rowpitch + (yy & 3) * 16);
  }
  auto col_off = [&](int c) {
    const int xx = mx + c - 8 * bx0;
    return (uint32_t)(((xx >> 3) << cshift) + (xx & 7) * 2);
  };
  float2 wn[kPairs], ws[kPairs];   // (fy_j, fy_j+2) and (1 - fy_j, 1 - fy_j+2)
#pragma unroll
  for (int p = 0; p < kPairs; ++p) {
    const int lo = lo_of(p);
    wn[p] = make_float2(ty[lo].frac, ty[lo + 2].frac);
    ws[p] = make_float2(__fsub_rn(1.0f, ty[lo].frac), __fsub_rn(1.0f, ty[lo + 2].frac));
  }
  const float n1 = ty[kSingle].frac, s1 = __fsub_rn(1.0f, n1);
  struct Col {
    float2 tl[kPairs], bl[kPairs];
    float t1, b1;
  };
  auto load_col = [&](uint32_t coff) {
    float v[kRows];
#pragma unroll
    for (int r = 0; r < kRows; ++r) v[r] = px(rows[r], coff);
    Col c;
#pragma unroll
    for (int p = 0; p < kPairs; ++p) {
      const int lo = lo_of(p);
      c.tl[p] = make_float2(v[lo], v[lo + 2]);
      c.bl[p] = make_float2(v[lo + 1], v[lo + 3]);
    }
    c.t1 = v[kSingle];
    c.b1 = v[kSingle + 1];
    return c;
  };
  Col L = load_col(col_off(i_begin));
#pragma unroll kUnroll16
  for (int i = i_begin; i < i_end; ++i) {
    const Tap t = make_tap_fast(cxl, (float)(i - R), lv.sm1x, lv.invx, lv.halfx, kMode);
    const float fx = t.frac, ex = __fsub_rn(1.0f, fx);
    const float2 e2 = make_float2(ex, ex), w2 = make_float2(fx, fx);
    const Col Rt = load_col(col_off(i + 1));
    float* o = optr + (size_t)(unsigned)(i * rd) * N;
#pragma unroll
    for (int p = 0; p < kPairs; ++p) {
      const int lo = lo_of(p);
      // ATen's order (common.cuh bilinear_combine): nw = s*e, ne = s*w, sw = n*e, se = n*w; ((nw.v + ne.v) + sw.v) + se.v by fma
      float2 acc = __fmul2_rn(L.tl[p], __fmul2_rn(ws[p], e2));
      acc = __ffma2_rn(Rt.tl[p], __fmul2_rn(ws[p], w2), acc);
      acc = __ffma2_rn(L.bl[p], __fmul2_rn(wn[p], e2), acc);
      acc = __ffma2_rn(Rt.bl[p], __fmul2_rn(wn[p], w2), acc);
      if (kAllLive || live) {
        __stcs(o + (size_t)N * (unsigned)lo, acc.x);
        __stcs(o + (size_t)N * (unsigned)(lo + 2), acc.y);
      }
    }
    {
      float acc = __fmul_rn(L.t1, __fmul_rn(s1, ex));
      acc = __fmaf_rn(Rt.t1, __fmul_rn(s1, fx), acc);
      acc = __fmaf_rn(L.b1, __fmul_rn(n1, ex), acc);
      acc = __fmaf_rn(Rt.b1, __fmul_rn(n1, fx), acc);
      if (kAllLive || live) __stcs(o + (size_t)N * (unsigned)kSingle, acc);
    }
    L = Rt;
  }
}

// whole warp off the fast path (some lane's window does not fit a box): every lane samples from global memory
template <int R, int kMode>
__device__ __noinline__ void sample_unit_global(const Level16Dev& lv, long long img, float cxl, float cyl, int i_begin,
                                                int i_end, float* optr, long long N) {
  constexpr int rd = 2 * R + 1;
  for (int i = i_begin; i < i_end; ++i) {
    const Tap a = make_tap_fast(cxl, (float)(i - R), lv.sm1x, lv.invx, lv.halfx, kMode);
    for (int j = 0; j < rd; ++j) {
      const Tap c = make_tap_fast(cyl, (float)(j - R), lv.sm1y, lv.invy, lv.halfy, kMode);
      const float v = bilinear_combine(pixel16_global(lv, img, c.i0, a.i0), pixel16_global(lv, img, c.i0, a.i0 + 1),
                                       pixel16_global(lv, img, c.i0 + 1, a.i0), pixel16_global(lv, img, c.i0 + 1, a.i0 + 1),
                                       a.frac, c.frac);
      __stcs(optr + (long long)(i * rd + j) * N, v);
    }
  }
}

template <int R, int kMode>
__global__ void __launch_bounds__(kThreads16, 1)
lookup16_kernel(const __grid_constant__ Lookup16Params p, const __grid_constant__ Lookup16Maps maps) {
  constexpr int rd = 2 * R + 1, rd2 = rd * rd;
  constexpr int kChannels = kL * rd2;
  extern __shared__ __align__(1024) unsigned char smem[];
  const uint32_t smem_base = smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  auto full_bar = [&](int b) { return smem_base + kOffBars16 + b * 8; };

  if (threadIdx.x < kBufs) {
    mbar_init(full_bar(threadIdx.x), 1);   // one arrive.expect_tx per refill (+ the bytes of its boxes)
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  const long long N = (long long)p.H * p.W;
  const int n_tiles = (int)((p.tiles - blockIdx.x + gridDim.x - 1) / gridDim.x);   // tiles of this CTA

  // one warp = one unit buffer: warps 2l, 2l+1 belong to level l and take alternate tiles
  const int pair = warp;
  const int level = pair >> 1, first = pair & 1;
  const Level16Dev lv = p.lv[level];
  const int glog = group_log2(level), gl = lane & ((1 << glog) - 1);
  // region of the lane's group in the unit buffer
  const uint32_t group_addr =
      smem_base + pair * kBufBytes +
      (glog == 0 ? (uint32_t)slot_of(lane, level) * kSlot : (uint32_t)(lane >> glog) * (uint32_t)group_pitch(glog));
  // A group's box holds its images one after the other or, from an interleaved level, block by block
  // ([block row][block][image][64 B]); ungrouped queries keep kBoxMax bytes each.
  const int ilog = lv.ilog;
  auto slot_of_window = [&](const Window& w) {
    const uint32_t pitch = !w.grouped ? (uint32_t)kBoxMax : ilog ? 64u : (uint32_t)((2 + w.wide) * (3 + w.tall) * 64);
    return group_addr + (uint32_t)gl * pitch;
  };
  const uint32_t bar = full_bar(pair);

  struct Unit {
    float cxl, cyl;      // core/corr.py:41  centroid / 2**i
    long long img;
    float* optr;
    bool live;
  };
  auto group_usable = [&](const Unit& u) {   // interleaved: the lanes of a group must be one group in memory too
    return u.live && (ilog == 0 || ((int)u.img & ((1 << glog) - 1)) == gl);
  };
  // Tile cursors advance incrementally (a pair steps through tiles blockIdx + k * gridDim, k = first, first + 2, ...):
  // decoding a tile index costs two integer divisions, ~50 instructions each time, for every lane.
  struct Cursor {
    int tx, y, b;
  };
  const int tiles_x = p.tiles_x, fmH = p.H;
  auto cursor_at = [&](int k) {
    Cursor c;
    const unsigned tile = blockIdx.x + (unsigned)k * gridDim.x;
    const unsigned t2 = tile / (unsigned)tiles_x;
    c.tx = (int)(tile - t2 * (unsigned)tiles_x);
    c.b = (int)(t2 / (unsigned)fmH);
    c.y = (int)(t2 - (unsigned)c.b * (unsigned)fmH);
    return c;
  };
  Cursor step;   // kDepth16 * gridDim tiles, decomposed the same way
  {
    const unsigned st = (unsigned)kDepth16 * gridDim.x, t2 = st / (unsigned)tiles_x;
    step.tx = (int)(st - t2 * (unsigned)tiles_x);
    step.b = (int)(t2 / (unsigned)fmH);
    step.y = (int)(t2 - (unsigned)step.b * (unsigned)fmH);
  }
  auto advance = [&](Cursor& c) {
    c.tx += step.tx;
    if (c.tx >= tiles_x) {
      c.tx -= tiles_x;
      c.y += 1;
    }
    c.y += step.y;
    if (c.y >= fmH) {
      c.y -= fmH;
      c.b += 1;
    }
    c.b += step.b;
  };
  auto unit_at = [&](const Cursor& c, float cx, float cy) {
    Unit u;
    const int x = c.tx * kTileQ + lane;
    u.live = x < p.W;
    u.img = (long long)c.b * N + (long long)c.y * p.W + x;
    u.optr = p.out + ((long long)c.b * kChannels + level * rd2) * N + (long long)c.y * p.W + x;
    u.cxl = __fmul_rn(cx, lv.inv_pow2);
    u.cyl = __fmul_rn(cy, lv.inv_pow2);
    return u;
  };
  auto load_coords = [&](const Cursor& c, bool in_range, float& cx, float& cy) {
    cx = cy = 0.0f;
    const int x = c.tx * kTileQ + lane;
    if (in_range && x < p.W) {
      const float* ptr = p.coords + ((long long)c.b * 2 * fmH + c.y) * p.W + x;
      cx = __ldg(ptr);
      cy = __ldg(ptr + N);
    }
  };
  // request the boxes of a unit: one per group, or one per lane where the group has no common box
  auto issue = [&](const Unit& u, const Window& w) {
    const int bw = 2 + w.wide, bh = 3 + w.tall;
    uint32_t bytes = (u.live && w.fast) ? (uint32_t)(bw * bh * 64) : 0u;
    if (w.grouped) bytes = gl == 0 ? bytes << glog : 0u;   // the group's first lane fetches for all of them
    const uint32_t total = __reduce_add_sync(0xffffffffu, bytes);
    if (lane == 0) mbar_expect_tx(bar, total);
    __syncwarp();
    if (bytes) {
      // windows entirely outside the image are all zeros wherever the box sits: clamp far-away origins
      const int bx = max(-8, min(w.bx0, lv.bxs + 8)), by = max(-8, min(w.by0, lv.bys + 8));
      const int member = (ilog && !w.grouped) ? ((int)u.img & ((1 << ilog) - 1)) : 0;
      tma_load_5d(slot_of_window(w), &maps.m[level][w.grouped][w.wide][w.tall], bar, 0, member, bx, by, (int)u.img >> ilog);
    }
  };
  if (first >= n_tiles) return;
  float ncx, ncy;
  Cursor cu = cursor_at(first);            // the unit about to be fetched / sampled
  load_coords(cu, true, ncx, ncy);
  Unit cur = unit_at(cu, ncx, ncy);
  advance(cu);                             // -> first + 2
  Cursor cc = cu;                          // the tile whose coordinates are loaded next
  // All pairs take the same time per unit, so pairs that start together stay in lock step: every pair waits for
  // its boxes at the same moment and every pair samples at the same moment.  Half of them start one microsecond
  // late (measured best; RAFTCORR_L16_DEBUG bits 8.. override it in units of 256 ns, 255 = none).
  if (first) {
    const unsigned steps = (p.debug >> 8) ? (unsigned)(p.debug >> 8) : 4u;
    for (unsigned t = 0; t < steps && steps != 255u; ++t) __nanosleep(256);
  }
  Window w = group_window(window_of<R, kMode>(cur.cxl, cur.cyl, lv), group_usable(cur), glog);   // the box of `cur`
  issue(cur, w);
  load_coords(cc, first + kDepth16 < n_tiles, ncx, ncy);
  advance(cc);                             // -> first + 4
  uint32_t phase = 0;
#ifdef RAFTCORR_L16_TRACE
  long long tr_prep = 0, tr_wait = 0, tr_sample = 0, tr_issue = 0, tr_units = 0;
  const long long tr_begin = clock64();
#endif
  for (int k = first; k < n_tiles; k += kDepth16) {
    L16_TICK(t0);
    const bool more = k + kDepth16 < n_tiles;
    Unit nxt = cur;
    if (more) nxt = unit_at(cu, ncx, ncy);
    const Window wn = group_window(window_of<R, kMode>(nxt.cxl, nxt.cyl, lv), group_usable(nxt), glog);
    advance(cu);
    load_coords(cc, k + 2 * kDepth16 < n_tiles, ncx, ncy);   // two units ahead
    advance(cc);
    Tap tx[rd], ty[rd];
    bool chain = true;
#pragma unroll
    for (int j = 0; j < rd; ++j) {
      ty[j] = make_tap_fast(cur.cyl, (float)(j - R), lv.sm1y, lv.invy, lv.halfy, kMode);
      tx[j] = make_tap_fast(cur.cxl, (float)(j - R), lv.sm1x, lv.invx, lv.halfx, kMode);
      if (j) chain &= (ty[j].i0 == ty[j - 1].i0 + 1) & (tx[j].i0 == tx[j - 1].i0 + 1);
    }
    const int bw = 2 + w.wide, bh = 3 + w.tall;
    const uint32_t slot_addr = slot_of_window(w);
    const int cshift = 6 + (w.grouped ? ilog : 0);
    const uint32_t swz_mask = (w.grouped && ilog) ? 0x70u : 0x30u;
    // floor flips: is tap j on pixel m + j + e_j with e_j in {0,1}?
    int mx = tx[0].i0, my = ty[0].i0, hx = tx[0].i0, hy = ty[0].i0;
#pragma unroll
    for (int j = 1; j < rd; ++j) {
      mx = min(mx, tx[j].i0 - j);
      hx = max(hx, tx[j].i0 - j);
      my = min(my, ty[j].i0 - j);
      hy = max(hy, ty[j].i0 - j);
    }
    const bool flips_ok = (hx - mx <= 1) & (hy - my <= 1);
    const bool live = cur.live;
    const bool all_fast = __all_sync(0xffffffffu, w.fast || !live);
    const bool all_chain = __all_sync(0xffffffffu, chain || !live);
    const bool all_live = __all_sync(0xffffffffu, live);
    L16_TICK(t1);
    mbar_wait(bar, phase, 2);
    L16_TICK(t2);
    phase ^= 1u;
    const bool all_flips = __all_sync(0xffffffffu, flips_ok || !live);
    constexpr int ib = 0, ie = rd;
    if (p.debug & 1) {
    } else if (!all_fast || !all_flips) {
      // a window that does not fit a box (|coords| beyond ~2^20) or an irregular tap pattern somewhere in the warp:
      // every lane samples straight from global memory
      if (live) sample_unit_global<R, kMode>(lv, cur.img, cur.cxl, cur.cyl, ib, ie, cur.optr, N);
    } else if (all_chain && all_live) {
      sample_unit_dense2<R, kMode, true>(slot_addr, swz_mask, cshift, bw, w.bx0, w.by0, mx, my, cur.cxl, lv, ty, ib, ie, cur.optr, (unsigned)N, live);
    } else if (all_chain) {
      sample_unit_dense2<R, kMode, false>(slot_addr, swz_mask, cshift, bw, w.bx0, w.by0, mx, my, cur.cxl, lv, ty, ib, ie, cur.optr, (unsigned)N, live);
    } else {
      sample_unit_flips<R, kMode>(slot_addr, swz_mask, cshift, bw, bh, w.bx0, w.by0, mx, my, cur.cxl, lv, ty, cur.optr, N, live);
    }
    L16_TICK(t3);
    if (more) {
      // every lane is done reading the buffer before it is refilled
      __syncwarp();
      issue(nxt, wn);
    }
    cur = nxt;
    w = wn;
#ifdef RAFTCORR_L16_TRACE
    const long long t4 = clock64();
    if (lane == 0 && blockIdx.x < 148 && warp < 8 && tr_units < 8) {
      unsigned long long* t = g_trace16 + ((size_t)blockIdx.x * 8 + warp) * 32 + 8 + tr_units * 3;
      t[0] = t2 - t1, t[1] = t3 - t2, t[2] = t4 - t3;
    }
    tr_prep += t1 - t0, tr_wait += t2 - t1, tr_sample += t3 - t2, tr_issue += t4 - t3, tr_units += 1;
#endif
  }
#ifdef RAFTCORR_L16_TRACE
  if (lane == 0 && blockIdx.x < 148 && warp < 8) {
    unsigned long long* t = g_trace16 + ((size_t)blockIdx.x * 8 + warp) * 32;
    t[0] = tr_prep, t[1] = tr_wait, t[2] = tr_sample, t[3] = tr_issue, t[4] = tr_units, t[5] = clock64() - tr_begin;
  }
#endif
}

// Tensor maps depend only on (pointers, sizes, strides); RAFT calls the lookup once per refinement iteration on the
// same pyramid and encoding 16 maps costs more host time than the kernel runs: keep the last few sets behind a
// mutex (the library stays re-entrant under nn.DataParallel's one-thread-per-GPU calls).
struct MapKey16 {
  const void* data[kL];
  long long istride[kL], bn;
  int bxs[kL], bys[kL], ilog[kL];
};
constexpr int kMapCache16 = 8;
struct MapCache16 {
  std::mutex mu;
  MapKey16 keys[kMapCache16];
  Lookup16Maps maps[kMapCache16];
  unsigned long long stamp[kMapCache16] = {};
  unsigned long long clock = 0;
};
MapCache16& map_cache16() {
  static MapCache16 c;
  return c;
}

template <int R, int kMode>
int launch16(const Lookup16Params& p, const Lookup16Maps& maps, cudaStream_t stream) {
  int dev = 0;
  RC_CHECK_CUDA(cudaGetDevice(&dev));
  static std::atomic<int> sm_count[64];  // per device: 0 = kernel attribute not set yet
  int sms = (dev >= 0 && dev < 64) ? sm_count[dev].load(std::memory_order_acquire) : 0;
  if (sms == 0) {
    RC_CHECK_CUDA(cudaFuncSetAttribute(lookup16_kernel<R, kMode>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem16));
    RC_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (dev >= 0 && dev < 64) sm_count[dev].store(sms, std::memory_order_release);
  }
  const int grid = (int)(p.tiles < sms ? p.tiles : sms);
  lookup16_kernel<R, kMode><<<grid, kThreads16, kSmem16, stream>>>(p, maps);
  RC_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

int lookup16_group_log2(int level) { return level >= 0 && level < kL ? group_log2(level) : 0; }

int launch_lookup16_fwd(const raftcorr_level16* levels, int L, const float* coords, float* out, int B, int H, int W,
                        int r, int mode, cudaStream_t stream) {
  if (L != kL || (r != 3 && r != 4))
    return fail(RAFTCORR_E_UNSUPPORTED, "lookup16_fwd: needs num_levels == 4 and radius 3 or 4 (got %d, %d)", L, r);
  if (get_encode() == nullptr) return fail(RAFTCORR_E_DRIVER, "cuTensorMapEncodeTiled is not available from this driver");
  const long long BN = (long long)B * H * W;
  if (BN >= (1LL << 31)) return fail(RAFTCORR_E_UNSUPPORTED, "lookup16_fwd: B*H*W must be below 2^31");
  Lookup16Params p;
  MapKey16 key;
  memset(&key, 0, sizeof(key));
  key.bn = BN;
  for (int l = 0; l < kL; ++l) {
    const raftcorr_level16& s = levels[l];
    Level16Dev& d = p.lv[l];
    d.data = static_cast<const unsigned char*>(s.data);
    d.H = s.H;
    d.W = s.W;
    d.bxs = s.blocks_x;
    d.bys = s.blocks_y;
    d.istride = s.image_stride;
    d.ilog = s.interleave_log2;
    if (d.ilog != 0 && d.ilog != group_log2(l))
      return fail(RAFTCORR_E_UNSUPPORTED, "lookup16_fwd: level %d interleaves 2^%d images, this build fetches groups of 2^%d",
                  l, d.ilog, group_log2(l));
    d.sm1x = (float)(s.W - 1);
    d.invx = 1.0f / d.sm1x;
    d.halfx = d.sm1x / 2.0f;
    d.sm1y = (float)(s.H - 1);
    d.invy = 1.0f / d.sm1y;
    d.halfy = d.sm1y / 2.0f;
    d.inv_pow2 = 1.0f / (float)(1 << l);
    key.data[l] = s.data;
    key.istride[l] = s.image_stride;
    key.bxs[l] = s.blocks_x;
    key.bys[l] = s.blocks_y;
    key.ilog[l] = s.interleave_log2;
  }
  Lookup16Maps maps;
  bool hit = false;
  {
    MapCache16& c = map_cache16();
    std::lock_guard<std::mutex> g(c.mu);
    for (int i = 0; i < kMapCache16 && !hit; ++i)
      if (c.stamp[i] && memcmp(&c.keys[i], &key, sizeof(key)) == 0) {
        c.stamp[i] = ++c.clock;
        maps = c.maps[i];
        hit = true;
      }
  }
  if (!hit) {
    for (int l = 0; l < kL; ++l) {
      // (64-byte block, image within its interleaved group, block column, block row, image or group of images)
      const int ilog = levels[l].interleave_log2, G = 1 << group_log2(l);
      const cuuint64_t members = 1ull << ilog, bstride = 64ull << ilog;
      cuuint64_t dims[5] = {32, members, (cuuint64_t)levels[l].blocks_x, (cuuint64_t)levels[l].blocks_y,
                            (cuuint64_t)((BN + (long long)members - 1) >> ilog)};
      cuuint64_t strides[4] = {64, bstride, (cuuint64_t)levels[l].blocks_x * bstride,
                               (cuuint64_t)levels[l].image_stride << ilog};
      for (int grouped = 0; grouped < 2; ++grouped)
        for (int wide = 0; wide < 2; ++wide)
          for (int tall = 0; tall < 2; ++tall) {
            if (grouped && ilog) {
              // the blocks of two neighbouring images as one 128-byte row: (row, image pair, block column, block row, group)
              cuuint64_t d2[5] = {64, members / 2, dims[2], dims[3], dims[4]};
              cuuint64_t s2[4] = {128, strides[1], strides[2], strides[3]};
              cuuint32_t b2[5] = {64, (cuuint32_t)(G / 2), (cuuint32_t)(2 + wide), (cuuint32_t)(3 + tall), 1};
              if (int rc = encode(&maps.m[l][grouped][wide][tall], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, levels[l].data, d2, s2,
                                  b2, CU_TENSOR_MAP_SWIZZLE_128B, "lookup16 group window", CU_TENSOR_MAP_L2_PROMOTION_NONE))
                return rc;
              continue;
            }
            cuuint32_t box[5] = {32, 1, (cuuint32_t)(2 + wide), (cuuint32_t)(3 + tall), 1};
            if (grouped) box[4] = (cuuint32_t)G;
            if (box[4] > dims[4]) box[4] = (cuuint32_t)dims[4];   // (never used then: a group needs 2^g live queries)
            if (int rc = encode(&maps.m[l][grouped][wide][tall], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, levels[l].data, dims,
                                strides, box, CU_TENSOR_MAP_SWIZZLE_64B, "lookup16 window", CU_TENSOR_MAP_L2_PROMOTION_NONE))
              return rc;
          }
    }
    MapCache16& c = map_cache16();
    std::lock_guard<std::mutex> g(c.mu);
    int victim = 0;
    for (int i = 1; i < kMapCache16; ++i)
      if (c.stamp[i] < c.stamp[victim]) victim = i;
    c.keys[victim] = key;
    c.maps[victim] = maps;
    c.stamp[victim] = ++c.clock;
  }
  p.coords = coords;
  p.out = out;
  p.debug = 0;
#ifdef RAFTCORR_L16_DEV   // development builds only (tools/build_variant.py dev -DRAFTCORR_L16_DEV): switches from the environment
  if (const char* dbg = getenv("RAFTCORR_L16_DEBUG")) p.debug = atoi(dbg);
#endif
  p.B = B;
  p.H = H;
  p.W = W;
  p.tiles_x = (W + kTileQ - 1) / kTileQ;
  p.tiles = (long long)B * H * p.tiles_x;
  if (p.tiles == 0) return 0;
  switch (mode * 8 + r) {
    case RAFTCORR_COORD_ATEN_CPU * 8 + 4: return launch16<4, RAFTCORR_COORD_ATEN_CPU>(p, maps, stream);
    case RAFTCORR_COORD_ATEN_CPU * 8 + 3: return launch16<3, RAFTCORR_COORD_ATEN_CPU>(p, maps, stream);
    case RAFTCORR_COORD_ATEN_CUDA * 8 + 4: return launch16<4, RAFTCORR_COORD_ATEN_CUDA>(p, maps, stream);
    case RAFTCORR_COORD_ATEN_CUDA * 8 + 3: return launch16<3, RAFTCORR_COORD_ATEN_CUDA>(p, maps, stream);
    case RAFTCORR_COORD_DIRECT * 8 + 4: return launch16<4, RAFTCORR_COORD_DIRECT>(p, maps, stream);
    default: return launch16<3, RAFTCORR_COORD_DIRECT>(p, maps, stream);
  }
}

}  // namespace raftcorr

#ifdef RAFTCORR_L16_TRACE
extern "C" int raftcorr_debug_l16_trace(unsigned long long* host, int count) {
  return (int)cudaMemcpyFromSymbol(host, raftcorr::g_trace16, sizeof(unsigned long long) * count);
}
#endif
