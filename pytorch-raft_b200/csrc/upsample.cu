// upsample.cu — convex 8x flow upsampling: replaces RAFT.upsample_flow,
// core/raft.py:72-83, which the reference runs after EVERY refinement iteration (raft.py:134-139).
//
//   up[n,c,8y+dy,8x+dx] = sum_k softmax_k(mask[n, k*64+dy*8+dx, y, x]) * 8*flow[n,c,y+k/3-1,x+k%3-1]
//
// The reference spells this as view + softmax + unfold + mul + sum + permute + reshape: six passes
// over a (N,576,H,W)-sized tensor.  Here it is one pass: HBM-bound, 2304 B of mask read and 512 B
// of flow written per coarse pixel.  One CTA owns one coarse row segment of 32 pixels; a warp is
// 32 adjacent x for one dy, so every mask load is one full 128-byte line, and the 8x8 sub-pixel
// results go through shared memory so that the fine rows leave as contiguous float4 stores.
#include <algorithm>
#include <atomic>

#include "tma_util.cuh"

namespace raftcorr {
namespace {

constexpr int kTileX = 32;
constexpr int kThreads = 256;  // 32 x  *  8 dy
constexpr int kPitch = kTileX + 1;

__device__ __forceinline__ void load_flow_tile(float (&fl)[2][3][kTileX + 2], const float* __restrict__ flow,
                                               int n, int y, int tile_x, int H, int W) {
  for (int i = threadIdx.x; i < 2 * 3 * (kTileX + 2); i += kThreads) {
    const int c = i / (3 * (kTileX + 2)), r = (i / (kTileX + 2)) % 3, cx = i % (kTileX + 2);
    const int yy = y + r - 1, xx = tile_x + cx - 1;
    float v = 0.f;  // F.unfold(padding=1) pads with zeros
    if (yy >= 0 && yy < H && xx >= 0 && xx < W) v = 8.0f * __ldg(flow + ((size_t)(n * 2 + c) * H + yy) * W + xx);
    fl[c][r][cx] = v;
  }
}

// softmax over the 9 taps exactly as written in the reference: exp(v - max) / sum.
__device__ __forceinline__ void softmax9(float (&v)[9]) {
  float mx = v[0];
#pragma unroll
  for (int k = 1; k < 9; ++k) mx = fmaxf(mx, v[k]);
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    v[k] = expf(v[k] - mx);
    s += v[k];
  }
  const float inv = 1.0f / s;
#pragma unroll
  for (int k = 0; k < 9; ++k) v[k] *= inv;
}

__global__ void __launch_bounds__(kThreads) upsample_fwd_kernel(const float* __restrict__ flow,
                                                                const float* __restrict__ mask,
                                                                float* __restrict__ up, int H, int W) {
  __shared__ float fl[2][3][kTileX + 2];
  __shared__ float out_s[2][8][8][kPitch];  // [c][dy][dx][x]
  const int tile_x = blockIdx.x * kTileX, y = blockIdx.y, n = blockIdx.z;
  const int tx = threadIdx.x & 31, dy = threadIdx.x >> 5;
  load_flow_tile(fl, flow, n, y, tile_x, H, W);
  __syncthreads();
  const int x = tile_x + tx;
  if (x < W) {
    float f0[9], f1[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      f0[k] = fl[0][k / 3][tx + k % 3];
      f1[k] = fl[1][k / 3][tx + k % 3];
    }
    const size_t cs = (size_t)H * W;
    const float* m = mask + ((size_t)n * 576 + dy * 8) * cs + (size_t)y * W + x;
#pragma unroll 8
    for (int dx = 0; dx < 8; ++dx) {
      float v[9];
#pragma unroll
      for (int k = 0; k < 9; ++k) v[k] = __ldcs(m + (size_t)(k * 64 + dx) * cs);
      softmax9(v);
      float a0 = 0.f, a1 = 0.f;
#pragma unroll
      for (int k = 0; k < 9; ++k) {
        a0 = fmaf(v[k], f0[k], a0);
        a1 = fmaf(v[k], f1[k], a1);
      }
      out_s[0][dy][dx][tx] = a0;
      out_s[1][dy][dx][tx] = a1;
    }
  }
  __syncthreads();
  // 16 fine rows (2 channels x 8 dy) of up to 256 contiguous floats each
  const int W8 = 8 * W;
  for (int i = threadIdx.x; i < 16 * (kTileX * 2); i += kThreads) {
    const int row = i / (kTileX * 2), q = i % (kTileX * 2);
    const int c = row >> 3, r = row & 7, xx = q >> 1, dxb = (q & 1) * 4;
    if (tile_x + xx >= W) continue;
    const float4 o = make_float4(out_s[c][r][dxb][xx], out_s[c][r][dxb + 1][xx], out_s[c][r][dxb + 2][xx],
                                 out_s[c][r][dxb + 3][xx]);
    float* dst = up + ((size_t)(n * 2 + c) * (8 * H) + 8 * y + r) * W8 + 8 * (tile_x + xx) + dxb;
    __stcs(reinterpret_cast<float4*>(dst), o);
  }
}

// ---- TMA-fed kernels (H*W % 4 == 0) -----------------------------------------------------------------
// The mask is viewed as a 3-D tensor {H*W pixels, 64 sub-pixels, 9*B taps}; one box {32 pixels, 8 dx, 9 k}
// = 9216 B is everything one warp needs for one dy of one 32-pixel row segment.  (TMA wants the box to
// start on 16 bytes: when W is not a multiple of 4 the box is 36 pixels wide, starts at the aligned-down
// pixel and the lanes read at the remainder.)  Each warp of a persistent CTA owns a two-slot ring of such
// boxes and is its own producer: it issues the box of its next work item, then computes the current one,
// so ~12 x 9 KB are always in flight per SM and the loads cost one instruction each.  A lane ends up
// holding the 8 dx of one x for both flow channels = two 32-byte runs of the fine image, which move as
// 256-bit accesses; nothing is staged for the output.
template <int kBoxW>
struct UpCfg {
  static constexpr int kWarps = kBoxW == 32 ? 12 : 10;
  static constexpr int kBoxBytes = 9 * 8 * kBoxW * 4;
  static constexpr int kSmem = kWarps * 2 * kBoxBytes + kWarps * 2 * 8;
};

__device__ __forceinline__ void st_v8(float* dst, const float (&a)[8]) {
  asm volatile("st.global.L1::no_allocate.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst), "f"(a[0]),
               "f"(a[1]), "f"(a[2]), "f"(a[3]), "f"(a[4]), "f"(a[5]), "f"(a[6]), "f"(a[7])
               : "memory");
}
__device__ __forceinline__ void ld_v8(const float* src, float (&a)[8]) {
  asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=f"(a[0]), "=f"(a[1]), "=f"(a[2]), "=f"(a[3]), "=f"(a[4]), "=f"(a[5]), "=f"(a[6]), "=f"(a[7])
               : "l"(src));
}

struct UpItem {
  int n, y, xt, dy;
};
__device__ __forceinline__ UpItem decode_item(int it, int tiles_x, int H) {
  // item = ((n*H + y)*tiles_x + xt)*8 + dy: the warps of a CTA take the 8 dy of one row segment
  UpItem u;
  u.dy = it & 7;
  const int seg = it >> 3;
  u.xt = seg % tiles_x;
  const int row = seg / tiles_x;
  u.y = row % H;
  u.n = row / H;
  return u;
}

template <bool kBwd, int kBoxW>
__global__ void __launch_bounds__(UpCfg<kBoxW>::kWarps * 32, 1)
    upsample_tma_kernel(const __grid_constant__ CUtensorMap mmap, const float* __restrict__ flow,
                        float* __restrict__ up,          // fwd: output; bwd: grad_up (read)
                        float* __restrict__ gflow, float* __restrict__ gmask, int H, int W, int tiles_x,
                        int n_items) {
  using Cfg = UpCfg<kBoxW>;
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* ring = reinterpret_cast<float*>(smem + (size_t)warp * 2 * Cfg::kBoxBytes);
  const uint32_t bar0 = smem_u32(smem + (size_t)Cfg::kWarps * 2 * Cfg::kBoxBytes + warp * 16);
  if (lane == 0) {
    mbar_init(bar0, 1);
    mbar_init(bar0 + 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  const int stride = gridDim.x * Cfg::kWarps;
  int item = blockIdx.x * Cfg::kWarps + warp;
  auto issue = [&](int it, int slot) {
    const UpItem u = decode_item(it, tiles_x, H);
    const uint32_t bar = bar0 + slot * 8;
    mbar_expect_tx(bar, Cfg::kBoxBytes);
    tma_load_3d(smem_u32(ring) + slot * Cfg::kBoxBytes, &mmap, bar, (u.y * W + u.xt * kTileX) & ~3, u.dy * 8, u.n * 9);
  };
  const size_t cs = (size_t)H * W;
  // 8*flow of the 3x3 coarse neighbours of this lane's pixel (zero outside: F.unfold(padding=1)); fetched one
  // item ahead like the mask box, so that the L2 round trip is hidden behind the current item's arithmetic
  auto load_flow = [&](int it, float (&f0)[9], float (&f1)[9]) {
    const UpItem u = decode_item(it, tiles_x, H);
    const int x = u.xt * kTileX + lane;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const int yy = u.y + k / 3 - 1, xx = x + k % 3 - 1;
      const bool in = yy >= 0 && yy < H && xx >= 0 && xx < W;
      const float* p = flow + ((size_t)(u.n * 2) * H + yy) * W + xx;
      f0[k] = in ? __ldg(p) : 0.f;
      f1[k] = in ? __ldg(p + cs) : 0.f;
    }
  };
  const size_t fine_cs = (size_t)(8 * H) * (8 * W);
  auto fine_ptr = [&](const UpItem& u, int x) {
    return up + ((size_t)(u.n * 2) * (8 * H) + 8 * u.y + u.dy) * (8 * W) + 8 * x;
  };
  // backward only: the incoming gradient of this lane's 8 fine pixels per channel, also one item ahead
  auto load_grad = [&](int it, float (&g0)[8], float (&g1)[8]) {
    const UpItem u = decode_item(it, tiles_x, H);
    const int x = u.xt * kTileX + lane;
#pragma unroll
    for (int dx = 0; dx < 8; ++dx) g0[dx] = g1[dx] = 0.f;
    if (x < W) {
      ld_v8(fine_ptr(u, x), g0);
      ld_v8(fine_ptr(u, x) + fine_cs, g1);
    }
  };
  float f0[9], f1[9], a0[8], a1[8];  // a: forward results / backward incoming gradient
  if (item < n_items) {
    if (elect_one()) issue(item, 0);
    load_flow(item, f0, f1);
    if constexpr (kBwd) load_grad(item, a0, a1);
  }
  uint32_t phases = 0;
  for (int it = 0; item < n_items; item += stride, ++it) {
    const int slot = it & 1;
    float nf0[9], nf1[9], na0[8], na1[8];
    if (item + stride < n_items) {
      if (elect_one()) issue(item + stride, slot ^ 1);
      load_flow(item + stride, nf0, nf1);
      if constexpr (kBwd) load_grad(item + stride, na0, na1);
    }
    const UpItem u = decode_item(item, tiles_x, H);
    const int x = u.xt * kTileX + lane;
    const bool live = x < W;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      f0[k] *= 8.0f;
      f1[k] *= 8.0f;
    }
    float* fine = fine_ptr(u, x);
    mbar_wait(bar0 + slot * 8, (phases >> slot) & 1u);
    phases ^= 1u << slot;
    const int rem = kBoxW == 32 ? 0 : ((u.y * W + u.xt * kTileX) & 3);
    const float* box = ring + slot * (Cfg::kBoxBytes / 4) + rem + lane;
    float t0[9], t1[9];
    if constexpr (kBwd) {
#pragma unroll
      for (int k = 0; k < 9; ++k) t0[k] = t1[k] = 0.f;
    }
    float* gm = kBwd && gmask ? gmask + ((size_t)u.n * 576 + u.dy * 8) * cs + (size_t)u.y * W + x : nullptr;
#pragma unroll
    for (int dx = 0; dx < 8; ++dx) {
      float v[9];
#pragma unroll
      for (int k = 0; k < 9; ++k) v[k] = box[(k * 8 + dx) * kBoxW];
      float mx = v[0];
#pragma unroll
      for (int k = 1; k < 9; ++k) mx = fmaxf(mx, v[k]);
      constexpr float kLog2e = 1.4426950408889634f;
      float s = 0.f;
#pragma unroll
      for (int k = 0; k < 9; ++k) {
        v[k] = exp2f((v[k] - mx) * kLog2e);  // v - mx first: the largest tap is exactly 1
        s += v[k];
      }
      const float inv = __frcp_rn(s);
      if constexpr (!kBwd) {
        float b0 = 0.f, b1 = 0.f;
#pragma unroll
        for (int k = 0; k < 9; ++k) {
          b0 = fmaf(v[k], f0[k], b0);
          b1 = fmaf(v[k], f1[k], b1);
        }
        a0[dx] = b0 * inv;
        a1[dx] = b1 * inv;
      } else {
        const float g0 = a0[dx], g1 = a1[dx];
        float gk[9], dot = 0.f;
#pragma unroll
        for (int k = 0; k < 9; ++k) {
          v[k] *= inv;
          gk[k] = fmaf(g0, f0[k], g1 * f1[k]);
          dot = fmaf(v[k], gk[k], dot);
          t0[k] = fmaf(v[k], g0, t0[k]);
          t1[k] = fmaf(v[k], g1, t1[k]);
        }
        if (gm && live) {
#pragma unroll
          for (int k = 0; k < 9; ++k) __stcs(gm + (size_t)(k * 64 + dx) * cs, v[k] * (gk[k] - dot));
        }
      }
    }
    __syncwarp();  // every lane is done with this slot before the next iteration refills it
    if constexpr (!kBwd) {
      if (live) {
        st_v8(fine, a0);
        st_v8(fine + fine_cs, a1);
      }
    } else if (gflow) {
      // coarse pixel P gets tap kx=0 of lane P+1, kx=1 of lane P and kx=2 of lane P-1 (dead lanes hold 0)
#pragma unroll
      for (int ky = 0; ky < 3; ++ky) {
        const int yy = u.y + ky - 1;
        const bool row_ok = yy >= 0 && yy < H;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          const float l = c ? t1[ky * 3 + 0] : t0[ky * 3 + 0];
          const float m = c ? t1[ky * 3 + 1] : t0[ky * 3 + 1];
          const float r = c ? t1[ky * 3 + 2] : t0[ky * 3 + 2];
          const float from_right = __shfl_down_sync(0xffffffffu, l, 1);
          const float from_left = __shfl_up_sync(0xffffffffu, r, 1);
          float sum = m + (lane < 31 ? from_right : 0.f) + (lane > 0 ? from_left : 0.f);
          if (!row_ok) continue;
          float* dst = gflow + ((size_t)(u.n * 2 + c) * H + yy) * W;
          if (live) atomicAdd(dst + x, 8.0f * sum);
          if (lane == 0 && x > 0) atomicAdd(dst + x - 1, 8.0f * l);
          if (lane == 31 && x + 1 < W) atomicAdd(dst + x + 1, 8.0f * r);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      f0[k] = nf0[k];
      f1[k] = nf1[k];
    }
    if constexpr (kBwd) {
#pragma unroll
      for (int dx = 0; dx < 8; ++dx) {
        a0[dx] = na0[dx];
        a1[dx] = na1[dx];
      }
    }
  }
}

// Backward: d(mask) through the softmax, d(flow) by scattering to the 3x3 neighbours.
__global__ void __launch_bounds__(kThreads) upsample_bwd_kernel(const float* __restrict__ flow,
                                                                const float* __restrict__ mask,
                                                                const float* __restrict__ gup,
                                                                float* __restrict__ gflow,
                                                                float* __restrict__ gmask, int H, int W) {
  __shared__ float fl[2][3][kTileX + 2];
  __shared__ float acc[2][3][kTileX + 2];
  __shared__ float g_s[2][8][8][kPitch];
  const int tile_x = blockIdx.x * kTileX, y = blockIdx.y, n = blockIdx.z;
  const int tx = threadIdx.x & 31, dy = threadIdx.x >> 5;
  load_flow_tile(fl, flow, n, y, tile_x, H, W);
  for (int i = threadIdx.x; i < 2 * 3 * (kTileX + 2); i += kThreads) (&acc[0][0][0])[i] = 0.f;
  const int W8 = 8 * W;
  for (int i = threadIdx.x; i < 16 * (kTileX * 2); i += kThreads) {
    const int row = i / (kTileX * 2), q = i % (kTileX * 2);
    const int c = row >> 3, r = row & 7, xx = q >> 1, dxb = (q & 1) * 4;
    if (tile_x + xx >= W) continue;
    const float4 g = __ldcs(reinterpret_cast<const float4*>(
        gup + ((size_t)(n * 2 + c) * (8 * H) + 8 * y + r) * W8 + 8 * (tile_x + xx) + dxb));
    g_s[c][r][dxb][xx] = g.x;
    g_s[c][r][dxb + 1][xx] = g.y;
    g_s[c][r][dxb + 2][xx] = g.z;
    g_s[c][r][dxb + 3][xx] = g.w;
  }
  __syncthreads();
  const int x = tile_x + tx;
  if (x < W) {
    float f0[9], f1[9], t0[9], t1[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      f0[k] = fl[0][k / 3][tx + k % 3];
      f1[k] = fl[1][k / 3][tx + k % 3];
      t0[k] = 0.f;
      t1[k] = 0.f;
    }
    const size_t cs = (size_t)H * W;
    const size_t off = ((size_t)n * 576 + dy * 8) * cs + (size_t)y * W + x;
#pragma unroll 2
    for (int dx = 0; dx < 8; ++dx) {
      float v[9];
#pragma unroll
      for (int k = 0; k < 9; ++k) v[k] = __ldcs(mask + off + (size_t)(k * 64 + dx) * cs);
      softmax9(v);
      const float g0 = g_s[0][dy][dx][tx], g1 = g_s[1][dy][dx][tx];
      float gk[9], dot = 0.f;
#pragma unroll
      for (int k = 0; k < 9; ++k) {
        gk[k] = fmaf(g0, f0[k], g1 * f1[k]);
        dot = fmaf(v[k], gk[k], dot);
        t0[k] = fmaf(v[k], g0, t0[k]);
        t1[k] = fmaf(v[k], g1, t1[k]);
      }
      if (gmask) {
#pragma unroll
        for (int k = 0; k < 9; ++k) __stcs(gmask + off + (size_t)(k * 64 + dx) * cs, v[k] * (gk[k] - dot));
      }
    }
    if (gflow) {
#pragma unroll
      for (int k = 0; k < 9; ++k) {
        atomicAdd(&acc[0][k / 3][tx + k % 3], t0[k]);
        atomicAdd(&acc[1][k / 3][tx + k % 3], t1[k]);
      }
    }
  }
  if (!gflow) return;
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * 3 * (kTileX + 2); i += kThreads) {
    const int c = i / (3 * (kTileX + 2)), r = (i / (kTileX + 2)) % 3, cx = i % (kTileX + 2);
    const int yy = y + r - 1, xx = tile_x + cx - 1;
    if (yy >= 0 && yy < H && xx >= 0 && xx < W)
      atomicAdd(gflow + ((size_t)(n * 2 + c) * H + yy) * W + xx, 8.0f * acc[c][r][cx]);
  }
}

// 8 * F.interpolate(flow, scale 8, bilinear, align_corners=True): core/utils/utils.py:80-82, the
// upsampling of the small model (raft.py:134-135).  One thread per output float4.
__global__ void __launch_bounds__(256) upflow8_kernel(const float* __restrict__ flow, float* __restrict__ up,
                                                      int H, int W, long long total4) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total4) return;
  const int W8 = 8 * W, H8 = 8 * H;
  const int q = (int)(i % (W8 / 4));
  const long long rest = i / (W8 / 4);
  const int oy = (int)(rest % H8);
  const long long img = rest / H8;
  // ATen area_pixel_compute_source_index, align_corners: src = dst * (in-1)/(out-1)
  const float sy = H8 > 1 ? (float)(H - 1) / (float)(H8 - 1) : 0.f;
  const float sx = W8 > 1 ? (float)(W - 1) / (float)(W8 - 1) : 0.f;
  const float fy = sy * oy;
  const int y0 = (int)fy, y1 = y0 + (y0 < H - 1 ? 1 : 0);
  const float ly = fy - y0, hy = 1.f - ly;
  const float* img0 = flow + img * (long long)H * W;
  float o[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float fx = sx * (q * 4 + j);
    const int x0 = (int)fx, x1 = x0 + (x0 < W - 1 ? 1 : 0);
    const float lx = fx - x0, hx = 1.f - lx;
    const float v = hy * (hx * __ldg(img0 + y0 * W + x0) + lx * __ldg(img0 + y0 * W + x1)) +
                    ly * (hx * __ldg(img0 + y1 * W + x0) + lx * __ldg(img0 + y1 * W + x1));
    o[j] = 8.0f * v;
  }
  __stcs(reinterpret_cast<float4*>(up) + i, make_float4(o[0], o[1], o[2], o[3]));
}


template <bool kBwd, int kBoxW>
int launch_upsample_tma(const float* flow, const float* mask, float* fine, float* gflow, float* gmask, int B, int H,
                        int W, long long items, cudaStream_t stream) {
  using Cfg = UpCfg<kBoxW>;
  CUtensorMap m;
  const cuuint64_t dims[3] = {(cuuint64_t)H * W, 64, (cuuint64_t)9 * B};
  const cuuint64_t strides[2] = {(cuuint64_t)H * W * 4, (cuuint64_t)64 * H * W * 4};
  const cuuint32_t box[3] = {kBoxW, 8, 9};
  if (int rc = encode(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(mask), dims, strides, box,
                      CU_TENSOR_MAP_SWIZZLE_NONE, "upsample mask", CU_TENSOR_MAP_L2_PROMOTION_NONE))
    return rc;
  int dev = 0, sms = 0;
  RC_CHECK_CUDA(cudaGetDevice(&dev));
  RC_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  static std::atomic<unsigned long long> configured{0};  // one bit per device (per instantiation)
  if (!(configured.load() >> dev & 1ull)) {
    RC_CHECK_CUDA(cudaFuncSetAttribute(upsample_tma_kernel<kBwd, kBoxW>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       Cfg::kSmem));
    configured.fetch_or(1ull << dev);
  }
  const int grid = (int)std::min<long long>(sms, (items + Cfg::kWarps - 1) / Cfg::kWarps);
  upsample_tma_kernel<kBwd, kBoxW><<<grid, Cfg::kWarps * 32, Cfg::kSmem, stream>>>(
      m, flow, fine, gflow, gmask, H, W, (W + kTileX - 1) / kTileX, (int)items);
  RC_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// TMA path: 16-byte channel stride, 16-byte mask base, 32-byte fine rows; dev < 64 for the attribute bitmap
bool upsample_tma_ok(const float* mask, const float* fine, int H, int W, long long items) {
  int dev = 0;
  return ((long long)H * W) % 4 == 0 && !(reinterpret_cast<uintptr_t>(mask) & 15) &&
         !(reinterpret_cast<uintptr_t>(fine) & 31) && items < (1ll << 30) && (long long)H * W < (1ll << 31) &&
         cudaGetDevice(&dev) == cudaSuccess && dev < 64 && get_encode() != nullptr;
}

}  // namespace

int launch_upsample_fwd(const float* flow, const float* mask, float* up, int B, int H, int W, cudaStream_t stream) {
  if (B == 0) return 0;
  const long long items = (long long)B * H * ((W + kTileX - 1) / kTileX) * 8;
  if (upsample_tma_ok(mask, up, H, W, items)) {
    if (W % 4 == 0) return launch_upsample_tma<false, 32>(flow, mask, up, nullptr, nullptr, B, H, W, items, stream);
    return launch_upsample_tma<false, 36>(flow, mask, up, nullptr, nullptr, B, H, W, items, stream);
  }
  dim3 grid((W + kTileX - 1) / kTileX, H, B);
  upsample_fwd_kernel<<<grid, kThreads, 0, stream>>>(flow, mask, up, H, W);
  RC_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int launch_upsample_bwd(const float* flow, const float* mask, const float* gup, float* gflow, float* gmask, int B,
                        int H, int W, cudaStream_t stream) {
  if (B == 0) return 0;
  if (gflow) RC_CHECK_CUDA(cudaMemsetAsync(gflow, 0, sizeof(float) * 2 * (size_t)B * H * W, stream));
  const long long items = (long long)B * H * ((W + kTileX - 1) / kTileX) * 8;
  if (upsample_tma_ok(mask, gup, H, W, items) && (!gmask || !(reinterpret_cast<uintptr_t>(gmask) & 3))) {
    float* g = const_cast<float*>(gup);  // read-only in the backward instantiation
    if (W % 4 == 0) return launch_upsample_tma<true, 32>(flow, mask, g, gflow, gmask, B, H, W, items, stream);
    return launch_upsample_tma<true, 36>(flow, mask, g, gflow, gmask, B, H, W, items, stream);
  }
  dim3 grid((W + kTileX - 1) / kTileX, H, B);
  upsample_bwd_kernel<<<grid, kThreads, 0, stream>>>(flow, mask, gup, gflow, gmask, H, W);
  RC_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int launch_upflow8(const float* flow, float* up, int B, int H, int W, cudaStream_t stream) {
  const long long total4 = (long long)B * 2 * (8 * H) * (8 * W / 4);
  if (total4 == 0) return 0;
  upflow8_kernel<<<(unsigned)((total4 + 255) / 256), 256, 0, stream>>>(flow, up, H, W, total4);
  RC_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace raftcorr
