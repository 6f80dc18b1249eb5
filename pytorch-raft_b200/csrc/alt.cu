// alt.cu — on-the-fly correlation (kernel family 3), forward and backward.
//
// Replaces alt_cuda_corr.forward/backward (reference alt_cuda_corr/correlation.cpp:23-48,
// correlation_kernel.cu:18-324) behind the same tensor layouts: fmap1 (B,H1,W1,C) NHWC, fmap2
// (B,H2,W2,C) NHWC, coords (B,1,H1,W1,2), corr (B,1,(2r+1)^2,H1,W1).  Not a port: the reference
// walks (2r+2)^2 taps for a 4x8 pixel block with one warp and read-modify-writes the output in
// global memory C/32*4 times; here a warp owns a query, takes each dot product of the <=12x12
// footprint exactly once, forms the (2r+1)^2 bilinear samples in shared memory and the CTA
// writes the output once as full 128-byte rows.
#include "common.cuh"

namespace raftcorr {

namespace {

constexpr int kTileQ = 32;
constexpr int kWarps = 8;
constexpr int kFoot = 12;
constexpr int kFootN = kFoot * kFoot;
constexpr int kMaxRd = 9;
constexpr int kOutPitch = kTileQ + 1;

struct Tap {
  int i0;
  float frac;
};

struct AltParams {
  const float* f1;
  const float* f2;
  const float* coords;
  float* corr;           // fwd output
  const float* gcorr;    // bwd input
  float* g1;
  float* g2;
  float* gcoords;
  int B, H1, W1, H2, W2, C, r, mode, tiles_x;
  float scale;
  float sm1x, invx, halfx, sm1y, invy, halfy;
};

__device__ __forceinline__ Tap make_tap(float centre, int off, float sm1, float inv, float half,
                                        int mode) {
  const float p = __fadd_rn(centre, (float)off);
  const float idx = sample_index(p, sm1, inv, half, mode);
  const float fl = floorf(idx);
  Tap t;
  t.i0 = floor_index(fl);
  t.frac = __fsub_rn(idx, fl);
  return t;
}

__device__ __forceinline__ void build_taps(const AltParams& p, Tap* tab, float cx, float cy, int lane) {
  const int rd = 2 * p.r + 1;
  for (int t = lane; t < 2 * rd; t += 32) {
    const int axis = t >= rd;
    const int tap = t - axis * rd;
    tab[axis * kMaxRd + tap] = axis ? make_tap(cy, tap - p.r, p.sm1y, p.invy, p.halfy, p.mode)
                                    : make_tap(cx, tap - p.r, p.sm1x, p.invx, p.halfx, p.mode);
  }
}

__device__ __forceinline__ bool in_image(const AltParams& p, int gy, int gx) {
  return (gy >= 0) & (gy < p.H2) & (gx >= 0) & (gx < p.W2);
}

// <f1s, fmap2[b, gy, gx, :]> by one thread (f1s in shared memory)
__device__ __forceinline__ float dot_thread(const AltParams& p, const float* f1s, int b, int gy, int gx) {
  const float* t = p.f2 + (((long long)b * p.H2 + gy) * p.W2 + gx) * p.C;
  float acc = 0.0f;
  if ((p.C & 3) == 0) {
    const float4* t4 = reinterpret_cast<const float4*>(t);
    const float4* a4 = reinterpret_cast<const float4*>(f1s);
    for (int c = 0; c < p.C / 4; ++c) {
      const float4 u = __ldg(t4 + c), a = a4[c];
      acc = fmaf(a.x, u.x, acc);
      acc = fmaf(a.y, u.y, acc);
      acc = fmaf(a.z, u.z, acc);
      acc = fmaf(a.w, u.w, acc);
    }
  } else {
    for (int c = 0; c < p.C; ++c) acc = fmaf(f1s[c], __ldg(t + c), acc);
  }
  return acc;
}

// shared-memory carve-up per CTA: [io tile][per-warp: f1 | g1 | S | G | taps]
__device__ __forceinline__ int warp_floats(int C) { return 2 * ((C + 3) & ~3) + 2 * kFootN + 2 * 2 * kMaxRd; }

__global__ void __launch_bounds__(kWarps * 32) alt_fwd_kernel(const __grid_constant__ AltParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int rd = 2 * p.r + 1, rd2 = rd * rd;
  const int Cp = (p.C + 3) & ~3;
  float* out_s = reinterpret_cast<float*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* wbase = out_s + ((rd2 * kOutPitch + 3) & ~3) + warp * warp_floats(p.C);
  float* f1s = wbase;
  float* S = wbase + 2 * Cp;
  Tap* tab = reinterpret_cast<Tap*>(wbase + 2 * Cp + 2 * kFootN);

  const int tile = blockIdx.x;
  const int xt = tile % p.tiles_x;
  const int y = (tile / p.tiles_x) % p.H1;
  const int b = tile / (p.tiles_x * p.H1);
  const int x0 = xt * kTileQ;
  const long long N1 = (long long)p.H1 * p.W1;

  for (int qi = warp; qi < kTileQ; qi += kWarps) {
    const int x = x0 + qi;
    if (x >= p.W1) break;
    const long long q = (long long)b * N1 + (long long)y * p.W1 + x;
    const float cx = __ldg(p.coords + q * 2), cy = __ldg(p.coords + q * 2 + 1);
    build_taps(p, tab, cx, cy, lane);
    for (int c = lane; c < p.C; c += 32) f1s[c] = __ldg(p.f1 + q * p.C + c);
    __syncwarp();
    const int x_lo = tab[0].i0, y_lo = tab[kMaxRd].i0;
    const int wdt = tab[rd - 1].i0 + 2 - x_lo, hgt = tab[kMaxRd + rd - 1].i0 + 2 - y_lo;
    const bool fast = (wdt <= kFoot) & (hgt <= kFoot);
    if (fast) {
      for (int idx = lane; idx < kFootN; idx += 32) {
        const int ry = idx / kFoot, rx = idx - ry * kFoot;
        const int gy = y_lo + ry, gx = x_lo + rx;
        S[idx] = (ry < hgt && rx < wdt && in_image(p, gy, gx)) ? dot_thread(p, f1s, b, gy, gx) : 0.0f;
      }
    }
    __syncwarp();
    for (int k = lane; k < rd2; k += 32) {
      const int i = k / rd, j = k - i * rd;
      const Tap a = tab[i], c = tab[kMaxRd + j];
      float vnw, vne, vsw, vse;
      if (fast) {
        const float* f = S + (c.i0 - y_lo) * kFoot + (a.i0 - x_lo);
        vnw = f[0];
        vne = f[1];
        vsw = f[kFoot];
        vse = f[kFoot + 1];
      } else {
        vnw = in_image(p, c.i0, a.i0) ? dot_thread(p, f1s, b, c.i0, a.i0) : 0.0f;
        vne = in_image(p, c.i0, a.i0 + 1) ? dot_thread(p, f1s, b, c.i0, a.i0 + 1) : 0.0f;
        vsw = in_image(p, c.i0 + 1, a.i0) ? dot_thread(p, f1s, b, c.i0 + 1, a.i0) : 0.0f;
        vse = in_image(p, c.i0 + 1, a.i0 + 1) ? dot_thread(p, f1s, b, c.i0 + 1, a.i0 + 1) : 0.0f;
      }
      out_s[k * kOutPitch + qi] = bilinear_combine(vnw, vne, vsw, vse, a.frac, c.frac) * p.scale;
    }
    __syncwarp();
  }
  __syncthreads();
  const int x = x0 + lane;
  if (x < p.W1) {
    float* dst = p.corr + (long long)b * rd2 * N1 + (long long)y * p.W1 + x;
    for (int k = warp; k < rd2; k += kWarps) __stcs(dst + (long long)k * N1, out_s[k * kOutPitch + lane]);
  }
}

// Backward.  d corr/d fmap1[q] = sum_cells G[cell] * fmap2[cell], d corr/d fmap2[cell] += G[cell] *
// fmap1[q], with G the footprint gradient (corr_grad scattered through the bilinear weights);
// d coords uses the recomputed dot products S.  The reference leaves coords_grad at zero
// (correlation_kernel.cu:130,307,323); filling it is a documented superset.
__global__ void __launch_bounds__(kWarps * 32) alt_bwd_kernel(const __grid_constant__ AltParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int rd = 2 * p.r + 1, rd2 = rd * rd;
  const int Cp = (p.C + 3) & ~3;
  float* g_s = reinterpret_cast<float*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* wbase = g_s + ((rd2 * kOutPitch + 3) & ~3) + warp * warp_floats(p.C);
  float* f1s = wbase;
  float* g1s = wbase + Cp;
  float* S = wbase + 2 * Cp;
  float* G = S + kFootN;
  Tap* tab = reinterpret_cast<Tap*>(wbase + 2 * Cp + 2 * kFootN);

  const int tile = blockIdx.x;
  const int xt = tile % p.tiles_x;
  const int y = (tile / p.tiles_x) % p.H1;
  const int b = tile / (p.tiles_x * p.H1);
  const int x0 = xt * kTileQ;
  const long long N1 = (long long)p.H1 * p.W1;
  {
    const int x = x0 + lane;
    const float* src = p.gcorr + (long long)b * rd2 * N1 + (long long)y * p.W1 + x;
    for (int k = warp; k < rd2; k += kWarps)
      g_s[k * kOutPitch + lane] = (x < p.W1) ? __ldg(src + (long long)k * N1) * p.scale : 0.0f;
  }
  __syncthreads();

  for (int qi = warp; qi < kTileQ; qi += kWarps) {
    const int x = x0 + qi;
    if (x >= p.W1) break;
    const long long q = (long long)b * N1 + (long long)y * p.W1 + x;
    const float cx = __ldg(p.coords + q * 2), cy = __ldg(p.coords + q * 2 + 1);
    build_taps(p, tab, cx, cy, lane);
    for (int c = lane; c < p.C; c += 32) {
      f1s[c] = __ldg(p.f1 + q * p.C + c);
      g1s[c] = 0.0f;
    }
    for (int idx = lane; idx < kFootN; idx += 32) G[idx] = 0.0f;
    __syncwarp();
    const int x_lo = tab[0].i0, y_lo = tab[kMaxRd].i0;
    const int wdt = tab[rd - 1].i0 + 2 - x_lo, hgt = tab[kMaxRd + rd - 1].i0 + 2 - y_lo;
    const bool fast = (wdt <= kFoot) & (hgt <= kFoot);
    if (!fast) continue;  // |coords| beyond 2^20: every tap is outside any real image
    if (p.gcoords) {
      for (int idx = lane; idx < kFootN; idx += 32) {
        const int ry = idx / kFoot, rx = idx - ry * kFoot;
        const int gy = y_lo + ry, gx = x_lo + rx;
        S[idx] = (ry < hgt && rx < wdt && in_image(p, gy, gx)) ? dot_thread(p, f1s, b, gy, gx) : 0.0f;
      }
      __syncwarp();
    }
    float gx_acc = 0.0f, gy_acc = 0.0f;
    for (int k = lane; k < rd2; k += 32) {
      const int i = k / rd, j = k - i * rd;
      const Tap a = tab[i], c = tab[kMaxRd + j];
      const float g = g_s[k * kOutPitch + qi];
      const float w = a.frac, e = 1.0f - w, n = c.frac, s = 1.0f - n;
      const int base = (c.i0 - y_lo) * kFoot + (a.i0 - x_lo);
      atomicAdd(G + base, g * (s * e));
      atomicAdd(G + base + 1, g * (s * w));
      atomicAdd(G + base + kFoot, g * (n * e));
      atomicAdd(G + base + kFoot + 1, g * (n * w));
      if (p.gcoords) {
        const float* f = S + base;
        gx_acc += g * ((f[1] - f[0]) * s + (f[kFoot + 1] - f[kFoot]) * n);
        gy_acc += g * ((f[kFoot] - f[0]) * e + (f[kFoot + 1] - f[1]) * w);
      }
    }
    __syncwarp();
    for (int idx = 0; idx < kFootN; ++idx) {
      const float gcell = G[idx];
      if (gcell == 0.0f) continue;
      const int ry = idx / kFoot, rx = idx - ry * kFoot;
      const int gy = y_lo + ry, gx = x_lo + rx;
      if (!in_image(p, gy, gx)) continue;
      const long long off = (((long long)b * p.H2 + gy) * p.W2 + gx) * p.C;
      for (int c = lane; c < p.C; c += 32) {
        g1s[c] = fmaf(gcell, __ldg(p.f2 + off + c), g1s[c]);
        atomicAdd(p.g2 + off + c, gcell * f1s[c]);
      }
    }
    for (int c = lane; c < p.C; c += 32) atomicAdd(p.g1 + q * p.C + c, g1s[c]);
    if (p.gcoords) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        gx_acc += __shfl_xor_sync(0xffffffffu, gx_acc, o);
        gy_acc += __shfl_xor_sync(0xffffffffu, gy_acc, o);
      }
      if (lane == 0) {
        p.gcoords[q * 2] = gx_acc;
        p.gcoords[q * 2 + 1] = gy_acc;
      }
    }
    __syncwarp();
  }
}

int fill_params(AltParams& p, const float* f1, const float* f2, const float* coords, int B, int H1,
                int W1, int H2, int W2, int C, int r, int mode, float scale, size_t* smem) {
  p.f1 = f1;
  p.f2 = f2;
  p.coords = coords;
  p.corr = nullptr;
  p.gcorr = nullptr;
  p.g1 = p.g2 = p.gcoords = nullptr;
  p.B = B;
  p.H1 = H1;
  p.W1 = W1;
  p.H2 = H2;
  p.W2 = W2;
  p.C = C;
  p.r = r;
  p.mode = mode;
  p.scale = scale;
  p.tiles_x = (W1 + kTileQ - 1) / kTileQ;
  p.sm1x = (float)(W2 - 1);
  p.invx = 1.0f / p.sm1x;
  p.halfx = p.sm1x / 2.0f;
  p.sm1y = (float)(H2 - 1);
  p.invy = 1.0f / p.sm1y;
  p.halfy = p.sm1y / 2.0f;
  const int rd = 2 * r + 1;
  const int Cp = (C + 3) & ~3;
  const size_t per_warp = (size_t)(2 * Cp + 2 * kFootN + 2 * 2 * kMaxRd);
  *smem = ((((size_t)rd * rd * kOutPitch + 3) & ~(size_t)3) + kWarps * per_warp) * sizeof(float);
  if (*smem > 200 * 1024) return fail(RAFTCORR_E_UNSUPPORTED, "alt: C=%d needs %zu B of shared memory", C, *smem);
  return 0;
}

}  // namespace

int launch_alt_fwd(const float* f1, const float* f2, const float* coords, float* corr, int B, int H1,
                   int W1, int H2, int W2, int C, int r, int mode, float scale, cudaStream_t stream) {
  AltParams p;
  size_t smem;
  if (int rc = fill_params(p, f1, f2, coords, B, H1, W1, H2, W2, C, r, mode, scale, &smem)) return rc;
  p.corr = corr;
  RC_CHECK_CUDA(cudaFuncSetAttribute(alt_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long tiles = (long long)B * H1 * p.tiles_x;
  if (tiles == 0) return 0;
  alt_fwd_kernel<<<(unsigned)tiles, kWarps * 32, smem, stream>>>(p);
  RC_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int launch_alt_bwd(const float* f1, const float* f2, const float* coords, const float* gcorr,
                   float* g1, float* g2, float* gcoords, int B, int H1, int W1, int H2, int W2, int C,
                   int r, int mode, float scale, cudaStream_t stream) {
  AltParams p;
  size_t smem;
  if (int rc = fill_params(p, f1, f2, coords, B, H1, W1, H2, W2, C, r, mode, scale, &smem)) return rc;
  p.gcorr = gcorr;
  p.g1 = g1;
  p.g2 = g2;
  p.gcoords = gcoords;
  if (gcoords)
    RC_CHECK_CUDA(cudaMemsetAsync(gcoords, 0, sizeof(float) * 2 * (size_t)B * H1 * W1, stream));
  RC_CHECK_CUDA(cudaFuncSetAttribute(alt_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long tiles = (long long)B * H1 * p.tiles_x;
  if (tiles == 0) return 0;
  alt_bwd_kernel<<<(unsigned)tiles, kWarps * 32, smem, stream>>>(p);
  RC_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace raftcorr
