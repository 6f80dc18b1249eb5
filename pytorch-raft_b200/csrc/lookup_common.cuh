// lookup_common.cuh — pieces shared by the generic-stride and the TMA lookup kernels.
#pragma once

#include "common.cuh"

namespace raftcorr {
namespace lk {

constexpr int kTileQ = 32;   // queries per CTA
constexpr int kWarps = 8;    // warps per CTA
constexpr int kFoot = 12;    // staged footprint edge: 2r+1 taps + 1 (bilinear) + 2 (floor flips)
constexpr int kFootN = kFoot * kFoot;
constexpr int kFootIters = (kFootN + 31) / 32;
constexpr int kMaxRd = 9;    // radius <= 4
constexpr int kOutPitch = kTileQ + 1;

struct LookupParams {
  LevelDev lv[RAFTCORR_MAX_LEVELS];
  const float* coords;
  float* out;
  int B, H, W, L, r, mode, tiles_x;
  int nhwc;   // 1: out is (B, H, W, C)
};

struct Tap {
  int i0;      // floor of the sampling index
  float frac;  // index - floor
};

__device__ __forceinline__ Tap make_tap(float centre, int tap_offset, float sm1, float inv,
                                        float half, int mode) {
  // core/corr.py:43  coords_lvl = centroid_lvl + delta_lvl  (one fp32 add), then the round trip
  const float p = __fadd_rn(centre, (float)tap_offset);
  const float idx = sample_index(p, sm1, inv, half, mode);
  const float fl = floorf(idx);
  Tap t;
  t.i0 = floor_index(fl);
  t.frac = __fsub_rn(idx, fl);
  return t;
}

// make_tap with the FMA division and a branch-free floor (NaN -> far outside, range clamped)
__device__ __forceinline__ Tap make_tap_fast(float centre, float tap_offset, float sm1, float inv,
                                             float half, int mode) {
  const float p = __fadd_rn(centre, tap_offset);
  const float idx = sample_index_fast(p, sm1, inv, half, mode);
  const float fl = floorf(idx);
  Tap t;
  const int i = __float2int_rd(fl);  // saturating conversion
  t.i0 = (fl == fl) ? max(-(1 << 29), min(i, 1 << 29)) : -(1 << 29);
  t.frac = __fsub_rn(idx, fl);
  return t;
}

__device__ __forceinline__ float load_pixel(const float* img, const LevelDev& lv, int gy, int gx) {
  const bool ok = (gy >= 0) & (gy < lv.H) & (gx >= 0) & (gx < lv.W);
  return ok ? __ldg(img + (long long)gy * lv.pitch + gx) : 0.0f;
}

}  // namespace lk
}  // namespace raftcorr
