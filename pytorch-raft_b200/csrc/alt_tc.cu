// alt_tc.cu — on-the-fly correlation on the tensor cores (kernel family 3, fused over levels).
//
// Replaces AlternateCorrBlock.__call__ (reference core/corr.py:74-91: four launches of
// alt_cuda_corr.forward plus NHWC copies and a stack/divide per iteration) with one launch.
// The reference kernel takes, per query and per tap, a 256-long dot product on the CUDA cores
// and never shares a fmap2 row between queries.  Optical flow is smooth, so the windows of
// neighbouring queries overlap almost entirely: a tile of 8x16 = 128 queries needs, at level l,
// only the fmap2_l pixels inside the UNION of its windows (about (8/2^l+11) x (16/2^l+11) of
// them).  That is a small dense GEMM:
//
//     S[128 queries, union pixels] = F1_tile[128, C] . F2_l[union, C]^T      (bf16 -> fp32 in TMEM)
//
// after which each query picks its own (2r+2)^2 footprint out of its own TMEM lane and forms the
// (2r+1)^2 bilinear samples.  The union is covered by 8x32-pixel chunks (N = 256 per tcgen05.mma,
// TMA 4-D boxes with zero fill outside the image = the zero padding of the sampler), so rough
// flow only means more chunks, never a wrong result.
//
//   warp 0      TMA producer  (F1 tile once, then one 8x32xC chunk of F2_l per step)
//   warp 1      MMA issuer    (tcgen05.mma cta_group::1 kind::f16, M128 N256, 2 TMEM stages)
//   warps 2-13  samplers      (thread = query = TMEM lane, three warps per lane quarter sharing the
//                             x taps): tcgen05.ld a union row, park it in a private shared-memory
//                             row, accumulate the taps that use it
#include <cuda.h>
#include <cuda_bf16.h>

#include <atomic>

#include "lookup_common.cuh"
#include "tma_util.cuh"

namespace raftcorr {

namespace {

using namespace lk;

constexpr int kTileH = 8, kTileW = 16;          // query tile (fmap1 pixels)
constexpr int kChunkH = 8, kChunkW = 32;        // fmap2 pixels per MMA chunk
constexpr int kN = kChunkH * kChunkW;           // 256 accumulator columns
constexpr int kKBlockBytes = 128;               // bytes of K per smem row (64 bf16)
constexpr int kMaxKBlocks = 4;                  // C <= 256
constexpr int kABytes = 128 * kKBlockBytes;     // 16 KB per k-block of the F1 tile
constexpr int kBBytes = kN * kKBlockBytes;      // 32 KB per k-block of a chunk
constexpr int kBStages = 3;
constexpr int kSPitch = 36;                     // floats per query row of the parking buffer (16-byte rows)
constexpr int kXGroups = 3;                     // sampler warps per TMEM lane quarter: x taps split 3 ways
constexpr int kSamplerWarps = 4 * kXGroups;
constexpr int kAltThreads = (2 + kSamplerWarps) * 32;
constexpr int kOffB = kMaxKBlocks * kABytes;                    // 64 KB
constexpr int kOffS = kOffB + kBStages * kBBytes;               // +96 KB
constexpr int kOffMisc = kOffS + kSamplerWarps * 32 * kSPitch * 4;  // union box, barriers, tmem slot
constexpr int kAltSmem = kOffMisc + 256;

struct AltTcLevel {
  int H, W;
  float sm1x, invx, halfx, sm1y, invy, halfy, inv_pow2;
};

struct AltTcParams {
  AltTcLevel lv[RAFTCORR_MAX_LEVELS];
  const float* coords;   // (B, 2, H1, W1), ch0 = x, ch1 = y (full-resolution pixel units)
  float* out;            // (B, L*(2r+1)^2, H1, W1)
  int B, H1, W1, C, L, k_blocks, tiles_x, tiles_y;
  uint32_t idesc;
  float scale;
};

struct alignas(64) AltTcMaps {
  CUtensorMap f1;                          // (C, W1, H1, B) bf16, box 64 x 16 x 8 x 1
  CUtensorMap f2[RAFTCORR_MAX_LEVELS];     // (C, W_l, H_l, B) bf16, box 64 x 32 x 8 x 1
};

template <int R, int kMode>
__global__ void __launch_bounds__(kAltThreads, 1)
alt_tc_kernel(const __grid_constant__ AltTcParams p, const __grid_constant__ AltTcMaps maps) {
  constexpr int rd = 2 * R + 1;
  extern __shared__ __align__(1024) unsigned char smem[];
  const uint32_t smem_base = smem_u32(smem);
  float* S_s = reinterpret_cast<float*>(smem + kOffS);
  int* ubox = reinterpret_cast<int*>(smem + kOffMisc);          // per level: x0, y0, x1, y1 of the union
  const uint32_t bars = smem_base + kOffMisc + 64;
  auto full_bar = [&](int s) { return bars + 8 * s; };
  auto empty_bar = [&](int s) { return bars + 8 * (kBStages + s); };
  auto tfull_bar = [&](int a) { return bars + 8 * (2 * kBStages + a); };
  auto tempty_bar = [&](int a) { return bars + 8 * (2 * kBStages + 2 + a); };
  const uint32_t afull_bar = bars + 8 * (2 * kBStages + 4);
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kOffMisc + 64 + 8 * (2 * kBStages + 5));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kBStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), kSamplerWarps);
    }
    mbar_init(afull_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(
                     smem_u32(const_cast<uint32_t*>(tmem_slot)))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // this CTA's query tile
  const int tile = blockIdx.x;
  const int txi = tile % p.tiles_x;
  const int tyi = (tile / p.tiles_x) % p.tiles_y;
  const int b = tile / (p.tiles_x * p.tiles_y);
  const int qx0 = txi * kTileW, qy0 = tyi * kTileH;
  const long long N1 = (long long)p.H1 * p.W1;

  // sampler state (thread = query)
  const bool sampler = warp >= 2;
  const int xg = sampler ? (warp - 2) / 4 : 0;          // which third of the x taps this warp owns
  // TMEM lanes 32*(warp%4).. belong to this warp; query index inside the tile follows the lane
  const int mq = sampler ? ((warp & 3) * 32 + lane) : 0;
  const int qy = qy0 + mq / kTileW, qx = qx0 + mq % kTileW;
  const bool valid = sampler && qy < p.H1 && qx < p.W1;
  float cx = 0.0f, cy = 0.0f;
  if (valid) {
    cx = __ldg(p.coords + ((long long)b * 2 + 0) * N1 + (long long)qy * p.W1 + qx);
    cy = __ldg(p.coords + ((long long)b * 2 + 1) * N1 + (long long)qy * p.W1 + qx);
  }

  if (warp == 0 && lane == 0) {   // F1 tile: k_blocks boxes of 128 queries x 128 B, loaded once
    mbar_expect_tx(afull_bar, (uint32_t)(p.k_blocks * kABytes));
    for (int kb = 0; kb < p.k_blocks; ++kb)
      tma_load_4d(smem_base + kb * kABytes, &maps.f1, afull_bar, kb * 64, qx0, qy0, b);
  }

  uint32_t it = 0;      // B-stage counter (producer and MMA count the same sequence)
  uint32_t chunk_no = 0;  // accumulator-stage counter (MMA and samplers)

  // Union of the tile's windows at every level, computed once up front so that the TMA / MMA
  // pipeline can run across level boundaries without a CTA barrier per level.
  if (threadIdx.x < 4 * RAFTCORR_MAX_LEVELS) ubox[threadIdx.x] = (threadIdx.x & 2) ? -(1 << 30) : (1 << 30);
  __syncthreads();
  if (valid && warp < 6) {   // one sampler warp per TMEM lane quarter (x group 0) reports its queries
    for (int l = 0; l < p.L; ++l) {
      const AltTcLevel& lv = p.lv[l];
      const float lx = __fmul_rn(cx, lv.inv_pow2), ly = __fmul_rn(cy, lv.inv_pow2);
      const int xa = max(make_tap_fast(lx, (float)(-R), lv.sm1x, lv.invx, lv.halfx, kMode).i0, 0);
      const int xb = min(make_tap_fast(lx, (float)R, lv.sm1x, lv.invx, lv.halfx, kMode).i0 + 1, lv.W - 1);
      const int ya = max(make_tap_fast(ly, (float)(-R), lv.sm1y, lv.invy, lv.halfy, kMode).i0, 0);
      const int yb = min(make_tap_fast(ly, (float)R, lv.sm1y, lv.invy, lv.halfy, kMode).i0 + 1, lv.H - 1);
      if (xa <= xb && ya <= yb) {
        atomicMin(&ubox[4 * l + 0], xa);
        atomicMin(&ubox[4 * l + 1], ya);
        atomicMax(&ubox[4 * l + 2], xb);
        atomicMax(&ubox[4 * l + 3], yb);
      }
    }
  }
  __syncthreads();

  for (int l = 0; l < p.L; ++l) {
    const AltTcLevel& lv = p.lv[l];
    const int ux0 = ubox[4 * l + 0], uy0 = ubox[4 * l + 1], ux1 = ubox[4 * l + 2], uy1 = ubox[4 * l + 3];
    const int ncx = ux1 >= ux0 ? (ux1 - ux0 + kChunkW) / kChunkW : 0;
    const int ncy = uy1 >= uy0 ? (uy1 - uy0 + kChunkH) / kChunkH : 0;
    const int n_chunks = ncx * ncy;

    // ---- tap tables of this thread's query at this level (registers) ----
    constexpr int kXPer = (rd + kXGroups - 1) / kXGroups;   // x taps per warp group
    int x0[kXPer], y0[rd];
    float wx[kXPer], wy[rd];
    int my_y_lo = 0, my_y_hi = -1, my_x_lo = 0, my_x_hi = -1;
    if (valid) {
      const float lx = __fmul_rn(cx, lv.inv_pow2), ly = __fmul_rn(cy, lv.inv_pow2);
#pragma unroll
      for (int t = 0; t < rd; ++t) {
        const Tap c = make_tap_fast(ly, (float)(t - R), lv.sm1y, lv.invy, lv.halfy, kMode);
        y0[t] = c.i0;
        wy[t] = c.frac;
      }
#pragma unroll
      for (int t = 0; t < kXPer; ++t) {
        const int i = min(xg * kXPer + t, rd - 1);
        const Tap a = make_tap_fast(lx, (float)(i - R), lv.sm1x, lv.invx, lv.halfx, kMode);
        x0[t] = a.i0;
        wx[t] = a.frac;
      }
      // the window of the whole query (all x taps), clipped to the image; empty if it lies outside
      const int xa = make_tap_fast(lx, (float)(-R), lv.sm1x, lv.invx, lv.halfx, kMode).i0;
      const int xb = make_tap_fast(lx, (float)R, lv.sm1x, lv.invx, lv.halfx, kMode).i0 + 1;
      const int q_x_lo = max(xa, 0), q_x_hi = min(xb, lv.W - 1);
      my_y_lo = max(y0[0], 0);
      my_y_hi = min(y0[rd - 1] + 1, lv.H - 1);
      if (q_x_lo <= q_x_hi && my_y_lo <= my_y_hi) {
        // columns this warp's own taps touch
        my_x_lo = max(x0[0], 0);
        my_x_hi = min(x0[kXPer - 1] + 1, lv.W - 1);
        if (my_x_lo > my_x_hi) my_y_hi = my_y_lo - 1;
      } else {
        my_y_hi = my_y_lo - 1;   // nothing to read for this query at this level
      }
    } else {
#pragma unroll
      for (int t = 0; t < rd; ++t) {
        y0[t] = 0;
        wy[t] = 0.0f;
      }
#pragma unroll
      for (int t = 0; t < kXPer; ++t) {
        x0[t] = 0;
        wx[t] = 0.0f;
      }
    }
    if (warp == 0) {
      if (lane == 0) {
        for (int c = 0; c < n_chunks; ++c) {
          const int cyi = c / ncx, cxi = c - cyi * ncx;
          for (int kb = 0; kb < p.k_blocks; ++kb, ++it) {
            const int s = it % kBStages;
            mbar_wait(empty_bar(s), ((it / kBStages) & 1) ^ 1);
            mbar_expect_tx(full_bar(s), kBBytes);
            tma_load_4d(smem_base + kOffB + s * kBBytes, &maps.f2[l], full_bar(s), kb * 64,
                        ux0 + cxi * kChunkW, uy0 + cyi * kChunkH, b);
          }
        }
      }
    } else if (warp == 1) {
      if (lane == 0) {
        if (l == 0) {
          mbar_wait(afull_bar, 0);
          tc_fence_after();
        }
        for (int c = 0; c < n_chunks; ++c, ++chunk_no) {
          const int as = chunk_no & 1;
          mbar_wait(tempty_bar(as), ((chunk_no >> 1) & 1) ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + as * kN;
          for (int kb = 0; kb < p.k_blocks; ++kb, ++it) {
            const int s = it % kBStages;
            mbar_wait(full_bar(s), (it / kBStages) & 1);
            tc_fence_after();
            const uint64_t adesc = make_smem_desc(smem_base + kb * kABytes);
            const uint64_t bdesc = make_smem_desc(smem_base + kOffB + s * kBBytes);
#pragma unroll
            for (int k = 0; k < 4; ++k) tc_mma<false>(d_tmem, adesc + 2 * k, bdesc + 2 * k, p.idesc, (kb | k) != 0);
            tc_commit(empty_bar(s));
          }
          tc_commit(tfull_bar(as));
        }
      }
    } else {
      // ---- samplers ----
      float acc[kXPer][rd];   // [own x tap][y tap]
#pragma unroll
      for (int i = 0; i < kXPer; ++i)
#pragma unroll
        for (int j = 0; j < rd; ++j) acc[i][j] = 0.0f;
      float* srow = S_s + ((warp - 2) * 32 + lane) * kSPitch;
      const uint32_t lane_base = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
      for (int c = 0; c < n_chunks; ++c, ++chunk_no) {
        const int as = chunk_no & 1;
        const int cyi = c / ncx, cxi = c - cyi * ncx;
        const int gx0 = ux0 + cxi * kChunkW, gy0 = uy0 + cyi * kChunkH;
        mbar_wait(tfull_bar(as), (chunk_no >> 1) & 1);
        tc_fence_after();
        const bool cols_hit = (my_x_lo <= gx0 + kChunkW - 1) & (my_x_hi >= gx0);
        for (int r = 0; r < kChunkH; ++r) {
          const int gy = gy0 + r;
          const bool need = cols_hit & (gy >= my_y_lo) & (gy <= my_y_hi);
          if (!__any_sync(0xffffffffu, need)) continue;   // no query of this warp touches the row
          float f[kChunkW];
          tmem_ld_x32(lane_base + as * kN + r * kChunkW, f);
          tmem_ld_wait();
          if (need) {
#pragma unroll
            for (int k = 0; k < kChunkW; k += 4)
              *reinterpret_cast<float4*>(srow + k) = make_float4(f[k], f[k + 1], f[k + 2], f[k + 3]);
            // horizontal interpolation of this row at the warp's x taps (columns of a tap that fall
            // into a neighbouring chunk contribute there: the sum over chunks is the full value)
            float h[kXPer];
#pragma unroll
            for (int i = 0; i < kXPer; ++i) {
              const int c0 = x0[i] - gx0;
              const float v0 = (unsigned)c0 < (unsigned)kChunkW ? srow[c0] : 0.0f;
              const float v1 = (unsigned)(c0 + 1) < (unsigned)kChunkW ? srow[c0 + 1] : 0.0f;
              h[i] = fmaf(v1, wx[i], __fmul_rn(v0, __fsub_rn(1.0f, wx[i])));
            }
            // the row is the north row of taps with y0 == gy and the south row of taps with y0+1 == gy
#pragma unroll
            for (int j = 0; j < rd; ++j) {
              if (y0[j] == gy) {
                const float s_ = __fsub_rn(1.0f, wy[j]);
#pragma unroll
                for (int i = 0; i < kXPer; ++i) acc[i][j] = fmaf(h[i], s_, acc[i][j]);
              }
              if (y0[j] + 1 == gy) {
#pragma unroll
                for (int i = 0; i < kXPer; ++i) acc[i][j] = fmaf(h[i], wy[j], acc[i][j]);
              }
            }
          }
        }
        // all reads of this accumulator stage are done: hand it back to the MMA warp
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tempty_bar(as));
      }
      if (valid) {
        // a NaN weight (coordinates that are NaN/inf or overflow in the round trip) makes the
        // reference's sample NaN even though every corner is out of range (0 * NaN): keep that
        float* dst = p.out + ((long long)b * p.L * rd * rd + (long long)l * rd * rd) * N1 + (long long)qy * p.W1 + qx;
#pragma unroll
        for (int t = 0; t < kXPer; ++t) {
          const int i = xg * kXPer + t;
          if (i < rd) {
#pragma unroll
            for (int j = 0; j < rd; ++j) {
              const bool is_nan = (wx[t] != wx[t]) | (wy[j] != wy[j]);
              __stcs(dst + (long long)(i * rd + j) * N1, is_nan ? __int_as_float(0x7fc00000) : acc[t][j] * p.scale);
            }
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
  }
}

template <int R, int kMode>
int launch_one(const AltTcParams& p, const AltTcMaps& maps, cudaStream_t stream) {
  static std::atomic<int> ready[64];
  int dev = 0;
  RC_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !ready[dev].load(std::memory_order_acquire)) {
    RC_CHECK_CUDA(cudaFuncSetAttribute(alt_tc_kernel<R, kMode>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAltSmem));
    if (dev >= 0 && dev < 64) ready[dev].store(1, std::memory_order_release);
  }
  const long long tiles = (long long)p.B * p.tiles_x * p.tiles_y;
  if (tiles == 0) return 0;
  alt_tc_kernel<R, kMode><<<(unsigned)tiles, kAltThreads, kAltSmem, stream>>>(p, maps);
  RC_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

int launch_alt_tc(const void* f1_bf16, const void* const* f2_bf16, const float* coords, float* out, int B,
                  int H1, int W1, int C, int L, int r, int mode, float scale, cudaStream_t stream) {
  if (get_encode() == nullptr) return fail(RAFTCORR_E_DRIVER, "alt_fused_fwd: cuTensorMapEncodeTiled is not available");
  if (C % 8 != 0 || C > 64 * kMaxKBlocks)
    return fail(RAFTCORR_E_UNSUPPORTED, "alt_fused_fwd: needs C %% 8 == 0 and C <= %d (C=%d)", 64 * kMaxKBlocks, C);
  if (r != 4 && r != 3) return fail(RAFTCORR_E_UNSUPPORTED, "alt_fused_fwd: radius must be 3 or 4 (r=%d)", r);
  AltTcParams p;
  AltTcMaps maps;
  {
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W1, (cuuint64_t)H1, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W1 * C * 2, (cuuint64_t)H1 * W1 * C * 2};
    cuuint32_t box[4] = {64, kTileW, kTileH, 1};
    if (int rc = encode(&maps.f1, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(f1_bf16), dims, strides, box,
                        CU_TENSOR_MAP_SWIZZLE_128B, "alt f1"))
      return rc;
  }
  for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l) {
    const int s = l < L ? l : 0;
    const int Hl = H1 >> s, Wl = W1 >> s;
    if (Hl < 1 || Wl < 1) return fail(RAFTCORR_E_INVALID, "alt_fused_fwd: level %d of %dx%d is empty", s, H1, W1);
    if (reinterpret_cast<uintptr_t>(f2_bf16[s]) & 15)
      return fail(RAFTCORR_E_UNSUPPORTED, "alt_fused_fwd: fmap2 level %d must be 16-byte aligned", s);
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)Wl, (cuuint64_t)Hl, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)Wl * C * 2, (cuuint64_t)Hl * Wl * C * 2};
    cuuint32_t box[4] = {64, kChunkW, kChunkH, 1};
    if (int rc = encode(&maps.f2[l], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(f2_bf16[s]), dims, strides,
                        box, CU_TENSOR_MAP_SWIZZLE_128B, "alt f2 level"))
      return rc;
    AltTcLevel& lv = p.lv[l];
    lv.H = Hl;
    lv.W = Wl;
    lv.sm1x = (float)(Wl - 1);
    lv.invx = 1.0f / lv.sm1x;
    lv.halfx = lv.sm1x / 2.0f;
    lv.sm1y = (float)(Hl - 1);
    lv.invy = 1.0f / lv.sm1y;
    lv.halfy = lv.sm1y / 2.0f;
    lv.inv_pow2 = 1.0f / (float)(1 << s);
  }
  p.coords = coords;
  p.out = out;
  p.B = B;
  p.H1 = H1;
  p.W1 = W1;
  p.C = C;
  p.L = L;
  p.k_blocks = (C + 63) / 64;
  p.tiles_x = (W1 + kTileW - 1) / kTileW;
  p.tiles_y = (H1 + kTileH - 1) / kTileH;
  p.idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  p.scale = scale;
  switch (mode * 8 + r) {
    case RAFTCORR_COORD_ATEN_CPU * 8 + 4: return launch_one<4, RAFTCORR_COORD_ATEN_CPU>(p, maps, stream);
    case RAFTCORR_COORD_ATEN_CPU * 8 + 3: return launch_one<3, RAFTCORR_COORD_ATEN_CPU>(p, maps, stream);
    case RAFTCORR_COORD_ATEN_CUDA * 8 + 4: return launch_one<4, RAFTCORR_COORD_ATEN_CUDA>(p, maps, stream);
    case RAFTCORR_COORD_ATEN_CUDA * 8 + 3: return launch_one<3, RAFTCORR_COORD_ATEN_CUDA>(p, maps, stream);
    case RAFTCORR_COORD_DIRECT * 8 + 4: return launch_one<4, RAFTCORR_COORD_DIRECT>(p, maps, stream);
    default: return launch_one<3, RAFTCORR_COORD_DIRECT>(p, maps, stream);
  }
}

}  // namespace raftcorr
