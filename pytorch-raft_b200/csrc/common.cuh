// common.cuh — shared host/device helpers for libraftcorr_b200 (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/raftcorr.h"

namespace raftcorr {

// ---- host-side error plumbing (thread-local text, int codes across the C ABI) --------------
void set_error(const char* fmt, ...);
int fail(int code, const char* fmt, ...);
int cuda_fail(cudaError_t err, const char* what);

#define RC_CHECK_CUDA(expr)                                    \
  do {                                                         \
    cudaError_t _e = (expr);                                   \
    if (_e != cudaSuccess) return raftcorr::cuda_fail(_e, #expr); \
  } while (0)

#define RC_REQUIRE(cond, ...)                                              \
  do {                                                                     \
    if (!(cond)) return raftcorr::fail(RAFTCORR_E_INVALID, __VA_ARGS__);   \
  } while (0)

// ---- device-side description of one pyramid level -----------------------------------------
struct LevelDev {
  const float* data;
  int H, W;
  long long pitch, istride;
  // constants of the coordinate round trip, per axis (x uses W, y uses H)
  float sm1x, invx, halfx;  // (W-1), 1/(W-1), (W-1)/2
  float sm1y, invy, halfy;
  float inv_pow2;           // 2^-level  (core/corr.py:41  coords / 2**i, exact)
};

inline LevelDev make_level_dev(const float* data, int H, int W, long long pitch, long long istride,
                               int level) {
  LevelDev d;
  d.data = data;
  d.H = H;
  d.W = W;
  d.pitch = pitch;
  d.istride = istride;
  d.sm1x = (float)(W - 1);
  d.invx = 1.0f / d.sm1x;
  d.halfx = d.sm1x / 2.0f;
  d.sm1y = (float)(H - 1);
  d.invy = 1.0f / d.sm1y;
  d.halfy = d.sm1y / 2.0f;
  d.inv_pow2 = 1.0f / (float)(1 << level);
  return d;
}

#ifdef __CUDACC__
// Pixel coordinate -> sampling index, replaying the reference's fp32 op sequence with one
// rounding per op (no FMA contraction): core/utils/utils.py:61-62, then ATen grid_sampler's
// align_corners=True un-normalise (GridSampler.h:27-36).
__device__ __forceinline__ float sample_index(float p, float sm1, float inv, float half, int mode) {
  if (mode == RAFTCORR_COORD_DIRECT) return p;
  float t = __fmul_rn(2.0f, p);
  t = (mode == RAFTCORR_COORD_ATEN_CPU) ? __fdiv_rn(t, sm1) : __fmul_rn(t, inv);
  t = __fsub_rn(t, 1.0f);
  t = __fadd_rn(t, 1.0f);
  return __fmul_rn(t, half);
}

// Same value as sample_index with three FMAs instead of IEEE division: for b = size-1 (a small
// positive integer) and inv = RN(1/b), q' = fma(fma(-q, b, a), inv, q) with q = a*inv is the
// correctly rounded a/b (Markstein's correction step; exhaustively sampled against __fdiv_rn for
// every b <= 4100 on a B200 by tools/probe/fdiv_probe.cu: 0 mismatches in 2.9e11 quotients).
// Non-finite a yields NaN instead of +-inf, which samples identically (all taps out of range).
__device__ __forceinline__ float sample_index_fast(float p, float sm1, float inv, float half, int mode) {
  if (mode == RAFTCORR_COORD_DIRECT) return p;
  float t = __fmul_rn(2.0f, p);
  const float q = __fmul_rn(t, inv);
  if (mode == RAFTCORR_COORD_ATEN_CPU) {
    const float r = __fmaf_rn(-q, sm1, t);
    t = __fmaf_rn(r, inv, q);
  } else {
    t = q;
  }
  t = __fsub_rn(t, 1.0f);
  t = __fadd_rn(t, 1.0f);
  return __fmul_rn(t, half);
}

// floor index as int, clamped so that index arithmetic cannot overflow; NaN -> far out of range.
__device__ __forceinline__ int floor_index(float fl) {
  if (!(fl == fl)) return -(1 << 29);
  fl = fminf(fmaxf(fl, -536870912.0f), 536870912.0f);
  return __float2int_rd(fl);
}

// ATen (AVX2 build) evaluates nw_val*nw + ne_val*ne + sw_val*sw + se_val*se as this FMA chain;
// replaying it makes the lookup bit-exact against the CPU reference for identical pyramids.
__device__ __forceinline__ float bilinear_combine(float vnw, float vne, float vsw, float vse,
                                                  float w, float n) {
  const float e = __fsub_rn(1.0f, w), s = __fsub_rn(1.0f, n);
  const float nw = __fmul_rn(s, e), ne = __fmul_rn(s, w), sw = __fmul_rn(n, e), se = __fmul_rn(n, w);
  return __fmaf_rn(vse, se, __fmaf_rn(vsw, sw, __fmaf_rn(vne, ne, __fmul_rn(vnw, nw))));
}
#endif

}  // namespace raftcorr
