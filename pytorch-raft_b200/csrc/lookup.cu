// lookup.cu — pyramid window lookup (kernel family 2), forward and backward.
//
// Replaces CorrBlock.__call__ (reference core/corr.py:29-50) together with bilinear_sampler
// (core/utils/utils.py:57-65) and ATen grid_sampler_2d(bilinear, zeros, align_corners=True).
// Written from the published semantics; not derived from alt_cuda_corr.
//
// Work decomposition: one CTA owns 32 consecutive-x queries of one fmap row, one warp works on
// one query at a time.  Per query and level the warp (a) computes the 2*(2r+1) tap coordinates
// with the reference's exact fp32 op sequence, (b) stages the <=12x12 footprint of that query's
// own H_l x W_l image in shared memory (rows are 40-48 B runs: the only global reads),
// (c) forms the (2r+1)^2 bilinear samples from shared memory.  Results are transposed through
// shared memory so the NCHW output (core/corr.py:50) is written as full 128-byte rows.
#include "common.cuh"
#include "lookup_common.cuh"

namespace raftcorr {

namespace {

using namespace lk;

// kR > 0 fixes the radius at compile time (no run-time integer divisions); kR == 0 reads p.r
template <int kR>
__global__ void __launch_bounds__(kWarps * 32)
lookup_fwd_kernel(const __grid_constant__ LookupParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int rd = kR > 0 ? 2 * kR + 1 : 2 * p.r + 1;
  const int rd2 = rd * rd;
  const int C = p.L * rd2;
  float* out_s = reinterpret_cast<float*>(smem_raw);                    // [C][kOutPitch]
  float* foot_all = out_s + C * kOutPitch;                              // [warps][L][kFootN]
  Tap* tab_all = reinterpret_cast<Tap*>(foot_all + kWarps * RAFTCORR_MAX_LEVELS * kFootN);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* foot = foot_all + warp * RAFTCORR_MAX_LEVELS * kFootN;
  Tap* tab = tab_all + warp * RAFTCORR_MAX_LEVELS * 2 * kMaxRd;         // [L][2][kMaxRd]

  const int tile = blockIdx.x;
  const int xt = tile % p.tiles_x;
  const int y = (tile / p.tiles_x) % p.H;
  const int b = tile / (p.tiles_x * p.H);
  const int x0 = xt * kTileQ;
  const long long N = (long long)p.H * p.W;

  for (int qi = warp; qi < kTileQ; qi += kWarps) {
    const int x = x0 + qi;
    if (x >= p.W) break;
    const long long pix = (long long)y * p.W + x;
    const long long q = (long long)b * N + pix;
    const float cx = __ldg(p.coords + ((long long)b * 2 + 0) * N + pix);
    const float cy = __ldg(p.coords + ((long long)b * 2 + 1) * N + pix);

    // (a) tap tables
    const int taps_total = p.L * 2 * rd;
    for (int t = lane; t < taps_total; t += 32) {
      const int l = t / (2 * rd);
      const int rem = t - l * 2 * rd;
      const int axis = rem >= rd;
      const int tap = rem - axis * rd;
      const LevelDev& lv = p.lv[l];
      const float c = __fmul_rn(axis ? cy : cx, lv.inv_pow2);            // corr.py:41
      tab[(l * 2 + axis) * kMaxRd + tap] =
          axis ? make_tap(c, tap - p.r, lv.sm1y, lv.invy, lv.halfy, p.mode)
               : make_tap(c, tap - p.r, lv.sm1x, lv.invx, lv.halfx, p.mode);
    }
    __syncwarp();

    // (b) stage footprints; all loads of all levels are issued before any is stored
    float v[RAFTCORR_MAX_LEVELS][kFootIters];
    bool fast[RAFTCORR_MAX_LEVELS];
#pragma unroll
    for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l) {
      fast[l] = false;
      if (l < p.L) {
        const LevelDev& lv = p.lv[l];
        const float* img = lv.data + q * lv.istride;
        const int x_lo = tab[(l * 2 + 0) * kMaxRd].i0, y_lo = tab[(l * 2 + 1) * kMaxRd].i0;
        const int wdt = tab[(l * 2 + 0) * kMaxRd + rd - 1].i0 + 2 - x_lo;
        const int hgt = tab[(l * 2 + 1) * kMaxRd + rd - 1].i0 + 2 - y_lo;
        fast[l] = (wdt <= kFoot) & (hgt <= kFoot);
#pragma unroll
        for (int it = 0; it < kFootIters; ++it) {
          const int idx = lane + 32 * it;
          const int ry = idx / kFoot, rx = idx - ry * kFoot;
          v[l][it] = (fast[l] && ry < hgt && rx < wdt) ? load_pixel(img, lv, y_lo + ry, x_lo + rx)
                                                      : 0.0f;
        }
      }
    }
#pragma unroll
    for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l)
      if (l < p.L) {
#pragma unroll
        for (int it = 0; it < kFootIters; ++it) {
          const int idx = lane + 32 * it;
          if (idx < kFootN) foot[l * kFootN + idx] = v[l][it];
        }
      }
    __syncwarp();

    // (c) bilinear samples
#pragma unroll
    for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l)
      if (l < p.L) {
        const LevelDev& lv = p.lv[l];
        const Tap* tx = tab + (l * 2 + 0) * kMaxRd;
        const Tap* ty = tab + (l * 2 + 1) * kMaxRd;
        const int x_lo = tx[0].i0, y_lo = ty[0].i0;
        for (int k = lane; k < rd2; k += 32) {
          const int i = k / rd, j = k - i * rd;  // i moves x, j moves y (corr.py:37-43)
          const Tap a = tx[i], c = ty[j];
          float vnw, vne, vsw, vse;
          if (fast[l]) {
            const float* f = foot + l * kFootN + (c.i0 - y_lo) * kFoot + (a.i0 - x_lo);
            vnw = f[0];
            vne = f[1];
            vsw = f[kFoot];
            vse = f[kFoot + 1];
          } else {  // coordinates so large that fp32 spacing breaks the window: direct loads
            const float* img = lv.data + q * lv.istride;
            vnw = load_pixel(img, lv, c.i0, a.i0);
            vne = load_pixel(img, lv, c.i0, a.i0 + 1);
            vsw = load_pixel(img, lv, c.i0 + 1, a.i0);
            vse = load_pixel(img, lv, c.i0 + 1, a.i0 + 1);
          }
          out_s[(l * rd2 + k) * kOutPitch + qi] = bilinear_combine(vnw, vne, vsw, vse, a.frac, c.frac);
        }
      }
    __syncwarp();
  }
  __syncthreads();

  if (p.nhwc) {   // (B, H, W, C): each query's C samples are contiguous
    for (int qi = warp; qi < kTileQ; qi += kWarps) {
      if (x0 + qi >= p.W) break;
      float* dst = p.out + ((long long)b * N + (long long)y * p.W + x0 + qi) * C;
      for (int c = lane; c < C; c += 32) __stcs(dst + c, out_s[c * kOutPitch + qi]);
    }
    return;
  }
  // coalesced NCHW store: one 128-byte row (32 consecutive x) per warp instruction
  const int x = x0 + lane;
  if (x < p.W) {
    float* dst = p.out + (long long)b * C * N + (long long)y * p.W + x;
    for (int c = warp; c < C; c += kWarps) __stcs(dst + (long long)c * N, out_s[c * kOutPitch + lane]);
  }
}

}  // namespace

int launch_lookup_fwd_tma(const raftcorr_level* levels, int L, const float* coords, float* out, int B,
                          int H, int W, int r, int mode, int nhwc, cudaStream_t stream);

int launch_lookup_fwd(const raftcorr_level* levels, int L, const float* coords, float* out, int B,
                      int H, int W, int r, int mode, int nhwc, cudaStream_t stream) {
  // Rows on 16-byte boundaries (raftcorr_pyramid_layout) and r in {3,4}: TMA-gather pipeline.
  // Any other strides / radii: the generic-stride kernel below (same arithmetic).
  const int rc = launch_lookup_fwd_tma(levels, L, coords, out, B, H, W, r, mode, nhwc, stream);
  if (rc != 1) return rc;
  LookupParams p;
  for (int l = 0; l < L; ++l)
    p.lv[l] = make_level_dev(levels[l].data, levels[l].H, levels[l].W, levels[l].pitch,
                             levels[l].image_stride, l);
  for (int l = L; l < RAFTCORR_MAX_LEVELS; ++l) p.lv[l] = p.lv[0];
  p.coords = coords;
  p.out = out;
  p.nhwc = nhwc;
  p.B = B;
  p.H = H;
  p.W = W;
  p.L = L;
  p.r = r;
  p.mode = mode;
  p.tiles_x = (W + kTileQ - 1) / kTileQ;
  const int rd = 2 * r + 1;
  const size_t smem = (size_t)L * rd * rd * kOutPitch * sizeof(float) +
                      (size_t)kWarps * RAFTCORR_MAX_LEVELS * kFootN * sizeof(float) +
                      (size_t)kWarps * RAFTCORR_MAX_LEVELS * 2 * kMaxRd * sizeof(Tap);
  static_assert(sizeof(Tap) == 8, "Tap is an (int, float) pair");
  const long long tiles = (long long)B * H * p.tiles_x;
  if (tiles == 0) return 0;
  auto launch = [&](auto kernel) -> int {
    RC_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kernel<<<(unsigned)tiles, kWarps * 32, smem, stream>>>(p);
    return 0;
  };
  if (int rc = (r == 4 ? launch(lookup_fwd_kernel<4>) : r == 3 ? launch(lookup_fwd_kernel<3>) : launch(lookup_fwd_kernel<0>)))
    return rc;
  RC_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace raftcorr

// ---------------------------------------------------------------------------------------------
// Backward of the lookup.  The reference obtains it from autograd (GridSampler2DBackward0 per
// level per iteration): d(pyramid) receives grad_out * {nw,ne,sw,se} at the
// four in-image corners; d(coords) = sum over taps of grad_out * d(sample)/d(index) * 2^-level
// (the normalise/un-normalise pair has unit slope).  One accumulation buffer per level serves
// all iterations: corners are added with float atomics (order-dependent rounding, hence the
// tolerance-based parity test).
// ---------------------------------------------------------------------------------------------
namespace raftcorr {

namespace {

struct LookupBwdParams {
  LevelDev lv[RAFTCORR_MAX_LEVELS];   // values (needed for d coords only)
  LevelDev glv[RAFTCORR_MAX_LEVELS];  // gradient accumulators
  const float* coords;
  const float* gout;
  float* gcoords;
  int B, H, W, L, r, mode, tiles_x, want_levels, want_coords;
};

template <int kR>
__global__ void __launch_bounds__(kWarps * 32)
lookup_bwd_kernel(const __grid_constant__ LookupBwdParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int rd = kR > 0 ? 2 * kR + 1 : 2 * p.r + 1;
  const int rd2 = rd * rd;
  const int C = p.L * rd2;
  float* g_s = reinterpret_cast<float*>(smem_raw);  // [C][kOutPitch]
  Tap* tab_all = reinterpret_cast<Tap*>(g_s + C * kOutPitch);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  Tap* tab = tab_all + warp * RAFTCORR_MAX_LEVELS * 2 * kMaxRd;

  const int tile = blockIdx.x;
  const int xt = tile % p.tiles_x;
  const int y = (tile / p.tiles_x) % p.H;
  const int b = tile / (p.tiles_x * p.H);
  const int x0 = xt * kTileQ;
  const long long N = (long long)p.H * p.W;

  {
    const int x = x0 + lane;
    const float* src = p.gout + (long long)b * C * N + (long long)y * p.W + x;
    for (int c = warp; c < C; c += kWarps)
      g_s[c * kOutPitch + lane] = (x < p.W) ? __ldg(src + (long long)c * N) : 0.0f;
  }
  __syncthreads();

  for (int qi = warp; qi < kTileQ; qi += kWarps) {
    const int x = x0 + qi;
    if (x >= p.W) break;
    const long long pix = (long long)y * p.W + x;
    const long long q = (long long)b * N + pix;
    const float cx = __ldg(p.coords + ((long long)b * 2 + 0) * N + pix);
    const float cy = __ldg(p.coords + ((long long)b * 2 + 1) * N + pix);
    const int taps_total = p.L * 2 * rd;
    for (int t = lane; t < taps_total; t += 32) {
      const int l = t / (2 * rd);
      const int rem = t - l * 2 * rd;
      const int axis = rem >= rd;
      const int tap = rem - axis * rd;
      const LevelDev& lv = p.want_levels ? p.glv[l] : p.lv[l];
      const float c = __fmul_rn(axis ? cy : cx, lv.inv_pow2);
      tab[(l * 2 + axis) * kMaxRd + tap] =
          axis ? make_tap(c, tap - p.r, lv.sm1y, lv.invy, lv.halfy, p.mode)
               : make_tap(c, tap - p.r, lv.sm1x, lv.invx, lv.halfx, p.mode);
    }
    __syncwarp();

    float gx = 0.0f, gy = 0.0f;
    for (int l = 0; l < p.L; ++l) {
      const Tap* tx = tab + (l * 2 + 0) * kMaxRd;
      const Tap* ty = tab + (l * 2 + 1) * kMaxRd;
      for (int k = lane; k < rd2; k += 32) {
        const int i = k / rd, j = k - i * rd;
        const Tap a = tx[i], c = ty[j];
        const float g = g_s[(l * rd2 + k) * kOutPitch + qi];
        const float w = a.frac, e = 1.0f - w, n = c.frac, s = 1.0f - n;
        if (p.want_levels) {
          const LevelDev& gl = p.glv[l];
          float* img = const_cast<float*>(gl.data) + q * gl.istride;
          const bool x0ok = (a.i0 >= 0) & (a.i0 < gl.W), x1ok = (a.i0 + 1 >= 0) & (a.i0 + 1 < gl.W);
          const bool y0ok = (c.i0 >= 0) & (c.i0 < gl.H), y1ok = (c.i0 + 1 >= 0) & (c.i0 + 1 < gl.H);
          float* row0 = img + (long long)c.i0 * gl.pitch + a.i0;
          float* row1 = row0 + gl.pitch;
          if (y0ok && x0ok) atomicAdd(row0, g * (s * e));
          if (y0ok && x1ok) atomicAdd(row0 + 1, g * (s * w));
          if (y1ok && x0ok) atomicAdd(row1, g * (n * e));
          if (y1ok && x1ok) atomicAdd(row1 + 1, g * (n * w));
        }
        if (p.want_coords) {
          const LevelDev& lv = p.lv[l];
          const float* img = lv.data + q * lv.istride;
          const float vnw = load_pixel(img, lv, c.i0, a.i0), vne = load_pixel(img, lv, c.i0, a.i0 + 1);
          const float vsw = load_pixel(img, lv, c.i0 + 1, a.i0),
                      vse = load_pixel(img, lv, c.i0 + 1, a.i0 + 1);
          gx += g * ((vne - vnw) * s + (vse - vsw) * n) * lv.inv_pow2;
          gy += g * ((vsw - vnw) * e + (vse - vne) * w) * lv.inv_pow2;
        }
      }
    }
    if (p.want_coords) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        gx += __shfl_xor_sync(0xffffffffu, gx, o);
        gy += __shfl_xor_sync(0xffffffffu, gy, o);
      }
      if (lane == 0) {
        p.gcoords[((long long)b * 2 + 0) * N + pix] = gx;
        p.gcoords[((long long)b * 2 + 1) * N + pix] = gy;
      }
    }
    __syncwarp();
  }
}

}  // namespace

int launch_lookup_bwd_tma(const raftcorr_level* glevels, int L, const float* coords, const float* gout, int B, int H,
                          int W, int r, int mode, cudaStream_t stream);

int launch_lookup_bwd(const raftcorr_level* levels, const raftcorr_level* glevels, int L,
                      const float* coords, const float* gout, float* gcoords, int B, int H, int W,
                      int r, int mode, cudaStream_t stream) {
  if (glevels && !gcoords) {   // RAFT's case (coords are detached, core/raft.py:123): TMA reduce-add path
    const int rc = launch_lookup_bwd_tma(glevels, L, coords, gout, B, H, W, r, mode, stream);
    if (rc != 1) return rc;
  }
  LookupBwdParams p;
  for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l) {
    const int s = l < L ? l : 0;
    const raftcorr_level* a = (gcoords || !glevels) ? levels : glevels;
    const raftcorr_level* g = glevels ? glevels : levels;
    p.lv[l] = make_level_dev(a[s].data, a[s].H, a[s].W, a[s].pitch, a[s].image_stride, s);
    p.glv[l] = make_level_dev(g[s].data, g[s].H, g[s].W, g[s].pitch, g[s].image_stride, s);
  }
  p.coords = coords;
  p.gout = gout;
  p.gcoords = gcoords;
  p.B = B;
  p.H = H;
  p.W = W;
  p.L = L;
  p.r = r;
  p.mode = mode;
  p.tiles_x = (W + kTileQ - 1) / kTileQ;
  p.want_levels = glevels != nullptr;
  p.want_coords = gcoords != nullptr;
  const int rd = 2 * r + 1;
  const size_t smem = (size_t)L * rd * rd * kOutPitch * sizeof(float) +
                      (size_t)kWarps * RAFTCORR_MAX_LEVELS * 2 * kMaxRd * sizeof(Tap);
  const long long tiles = (long long)B * H * p.tiles_x;
  if (tiles == 0) return 0;
  auto launch = [&](auto kernel) -> int {
    RC_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kernel<<<(unsigned)tiles, kWarps * 32, smem, stream>>>(p);
    return 0;
  };
  if (int rc = (r == 4 ? launch(lookup_bwd_kernel<4>) : r == 3 ? launch(lookup_bwd_kernel<3>) : launch(lookup_bwd_kernel<0>)))
    return rc;
  RC_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace raftcorr
