// lookup_tma.cu — the pyramid window lookup as a persistent TMA-gather pipeline (sm_100a).
//
// Same arithmetic as lookup.cu (reference core/corr.py:29-50 + core/utils/utils.py:57-65 +
// ATen grid_sampler_2d), different data movement.  The lookup is a gather: each query reads a
// ~10x10 window out of its OWN H_l x W_l image at each of the 4 levels — 40-byte row pieces
// scattered over gigabytes.  Instead of having warps chase those rows through the LSU, one
// producer warp turns every (query, level) window into ONE TMA tensor-tile load
// (cp.async.bulk.tensor.3d, box 12 x {10,11,12} x 1 out of the (W_l, H_l, B*N) level tensor):
// the TMA engine generates the row requests, zero-fills everything outside the image (that IS
// grid_sample's zero padding) and signals an mbarrier.  Sixteen consumer warps form the
// (2r+1)^2 x 4 bilinear samples from shared memory while the next tile's windows are in flight
// (two stages of 32 queries x 4 levels), then write the NCHW tile as full 128-byte rows.
//
//   CTA = 1 SM, persistent over tiles of 32 consecutive-x queries
//   warp 16        producer: coords -> box origins -> 128 TMA loads per tile
//   warps 0..15    consumers: tap tables, bilinear samples, transposed store
#include "lookup_common.cuh"
#include "tma_util.cuh"

namespace raftcorr {

namespace {

using namespace lk;

constexpr int kConsumers = 16;                        // consumer warps = queries per stage
constexpr int kThreadsTma = (kConsumers + 1) * 32;
constexpr int kStageQ = 16;                           // queries per pipeline stage (half a tile)
constexpr int kStagesTma = 3;
constexpr int kSlotBytes = 768;                       // 16 x 12 fp32 box
constexpr int kStageBytes = kStageQ * RAFTCORR_MAX_LEVELS * kSlotBytes;  // 48 KB
constexpr int kBoxHeights = 3;                        // 10, 11, 12 rows
constexpr int kBoxWidths = 2;                         // 12 or 16 columns

// TMA tile loads need the innermost start coordinate on a 16-byte boundary (x % 4 == 0 for
// fp32), so a window whose first column is x_lo is fetched from x_lo & ~3: 12 columns suffice
// when (x_lo & 3) + width <= 12, else 16.  Heights 10..12 cover the floor flips of the
// coordinate round trip.  One tensor map per (level, width, height).
struct alignas(64) LookupTmaMaps {
  CUtensorMap m[RAFTCORR_MAX_LEVELS][kBoxWidths][kBoxHeights];
};

struct LookupTmaParams {
  LevelDev lv[RAFTCORR_MAX_LEVELS];
  const float* coords;
  float* out;
  int B, H, W, L, mode, tiles_x;
  long long tiles;
};

template <int R>
struct Cfg {
  static constexpr int kRd = 2 * R + 1;
  static constexpr int kRd2 = kRd * kRd;
  static constexpr int kJGroup = 3;                               // outputs per lane along y
  static constexpr int kJBlocks = (kRd + kJGroup - 1) / kJGroup;  // 3
  static constexpr int kActive = kRd * kJBlocks;                  // lanes with work (27 or 21)
  static constexpr int kOutBytes = ((RAFTCORR_MAX_LEVELS * kRd2 * kOutPitch * 4 + 127) / 128) * 128;
  static constexpr int kTabBytes = kConsumers * RAFTCORR_MAX_LEVELS * 2 * kMaxRd * 8;
  static constexpr int kOffOut = kStagesTma * kStageBytes;
  static constexpr int kOffTab = kOffOut + kOutBytes;
  static constexpr int kOffCoords = kOffTab + kTabBytes;
  static constexpr int kOffBars = kOffCoords + kStagesTma * kStageQ * 8;
  static constexpr int kSmem = kOffBars + 64;
};

__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// Window geometry of one (query, level): derived from the first and last tap of each axis with
// the same arithmetic on the producer and the consumer side, so both agree on the box.
struct Window {
  int x_lo, y_lo;   // floor index of tap -R
  int wdt, hgt;     // columns / rows the 2R+1 taps (+1 for bilinear) span
  int ox;           // x_lo - box start (0..3)
  int wsel, hsel;   // which tensor map
  int pitch;        // staged row pitch in floats (12 or 16)
  bool fast;        // window fits the staged box
};

__device__ __forceinline__ Window make_window(int x_lo, int x_last, int y_lo, int y_last) {
  Window w;
  w.x_lo = x_lo;
  w.y_lo = y_lo;
  w.wdt = x_last + 2 - x_lo;
  w.hgt = y_last + 2 - y_lo;
  w.fast = (w.wdt >= 1) & (w.wdt <= kFoot) & (w.hgt >= 1) & (w.hgt <= kFoot);
  w.ox = x_lo & 3;  // two's complement: also right for negative x_lo
  w.wsel = (w.ox + w.wdt > 12) ? 1 : 0;
  w.pitch = w.wsel ? 16 : 12;
  const int h = w.hgt < 10 ? 10 : (w.hgt > 12 ? 12 : w.hgt);
  w.hsel = h - 10;
  return w;
}

template <int R>
__global__ void __launch_bounds__(kThreadsTma, 1)
lookup_fwd_tma_kernel(const __grid_constant__ LookupTmaParams p, const __grid_constant__ LookupTmaMaps maps) {
  using C = Cfg<R>;
  constexpr int rd = C::kRd, rd2 = C::kRd2;
  extern __shared__ __align__(1024) unsigned char smem[];
  const uint32_t smem_base = smem_u32(smem);
  float* out_s = reinterpret_cast<float*>(smem + C::kOffOut);
  float2* coords_s = reinterpret_cast<float2*>(smem + C::kOffCoords);  // [stages][16]
  const uint32_t bars = smem_base + C::kOffBars;
  auto full_bar = [&](int s) { return bars + 8 * s; };
  auto empty_bar = [&](int s) { return bars + 8 * (kStagesTma + s); };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStagesTma; ++s) {
      mbar_init(full_bar(s), 32);             // one arrive (+ its bytes) per producer lane
      mbar_init(empty_bar(s), kConsumers);    // one arrive per consumer warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  const long long N = (long long)p.H * p.W;
  const int nC = p.L * rd2;

  if (warp == kConsumers) {
    // ------------------------------ producer ------------------------------
    // lane = (level pair, query): lanes 0-15 fetch levels 0,1 and lanes 16-31 levels 2,3
    const int pq = lane & 15, l0 = (lane >> 4) * 2;
    uint32_t u = 0;
    for (long long tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
      const int xt = (int)(tile % p.tiles_x);
      const int y = (int)((tile / p.tiles_x) % p.H);
      const int b = (int)(tile / ((long long)p.tiles_x * p.H));
      for (int half = 0; half < 2; ++half, ++u) {
        const int stage = u % kStagesTma;
        const uint32_t ph = (u / kStagesTma) & 1;
        const int x = xt * kTileQ + half * kStageQ + pq;
        const bool valid = x < p.W;
        const long long pix = (long long)y * p.W + (valid ? x : 0);
        const float cx = __ldg(p.coords + ((long long)b * 2 + 0) * N + pix);
        const float cy = __ldg(p.coords + ((long long)b * 2 + 1) * N + pix);
        mbar_wait(empty_bar(stage), ph ^ 1, 1);
        if (lane < kStageQ) coords_s[stage * kStageQ + pq] = make_float2(cx, cy);
        int bx[2], by[2], ws[2], hs[2];
        uint32_t bytes = 0;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
          const int l = l0 + k;
          bx[k] = by[k] = ws[k] = hs[k] = 0;
          if (valid && l < p.L) {
            const LevelDev& lv = p.lv[l];
            const float lx = __fmul_rn(cx, lv.inv_pow2), ly = __fmul_rn(cy, lv.inv_pow2);
            const Window w = make_window(make_tap(lx, -R, lv.sm1x, lv.invx, lv.halfx, p.mode).i0,
                                         make_tap(lx, R, lv.sm1x, lv.invx, lv.halfx, p.mode).i0,
                                         make_tap(ly, -R, lv.sm1y, lv.invy, lv.halfy, p.mode).i0,
                                         make_tap(ly, R, lv.sm1y, lv.invy, lv.halfy, p.mode).i0);
            // windows entirely outside the image are all zeros wherever the box sits
            bx[k] = max(-64, min(w.x_lo - w.ox, (lv.W + 64) & ~3));
            by[k] = max(-64, min(w.y_lo, lv.H + 64));
            ws[k] = w.wsel;
            hs[k] = w.hsel;
            bytes += (uint32_t)(w.pitch * 4 * (10 + w.hsel));
          }
        }
        if (bytes) mbar_expect_tx(full_bar(stage), bytes); else mbar_arrive(full_bar(stage));
        if (valid) {
          const int q = (int)((long long)b * N + pix);
          const uint32_t dst0 = smem_base + stage * kStageBytes + pq * (RAFTCORR_MAX_LEVELS * kSlotBytes);
#pragma unroll
          for (int k = 0; k < 2; ++k)
            if (l0 + k < p.L)
              tma_load_3d(dst0 + (l0 + k) * kSlotBytes, &maps.m[l0 + k][ws[k]][hs[k]], full_bar(stage),
                          bx[k], by[k], q);
        }
      }
    }
  } else {
    // ------------------------------ consumers ------------------------------
    Tap* tab = reinterpret_cast<Tap*>(smem + C::kOffTab) + warp * (RAFTCORR_MAX_LEVELS * 2 * kMaxRd);
    const int ti = lane / C::kJBlocks;            // x tap of this lane
    const int tj0 = (lane - ti * C::kJBlocks) * C::kJGroup;  // first y tap of this lane
    uint32_t u = 0;
    for (long long tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
      const int xt = (int)(tile % p.tiles_x);
      const int y = (int)((tile / p.tiles_x) % p.H);
      const int b = (int)(tile / ((long long)p.tiles_x * p.H));
      const int x0 = xt * kTileQ;
      for (int half = 0; half < 2; ++half, ++u) {
        const int stage = u % kStagesTma;
        const uint32_t ph = (u / kStagesTma) & 1;
        const int qi = half * kStageQ + warp;     // this warp's query within the tile
        mbar_wait(full_bar(stage), ph, 2);
        if (x0 + qi < p.W) {
          const float2 cxy = coords_s[stage * kStageQ + warp];
          // tap tables: L * 2 * rd coordinates, one per lane per round
#pragma unroll
          for (int t0 = 0; t0 < RAFTCORR_MAX_LEVELS * 2 * rd; t0 += 32) {
            const int t = t0 + lane;
            if (t < p.L * 2 * rd) {
              const int l = t / (2 * rd);
              const int rem = t - l * 2 * rd;
              const int axis = rem >= rd;
              const int tap = rem - axis * rd;
              const LevelDev& lv = p.lv[l];
              const float c = __fmul_rn(axis ? cxy.y : cxy.x, lv.inv_pow2);
              tab[(l * 2 + axis) * kMaxRd + tap] =
                  axis ? make_tap(c, tap - R, lv.sm1y, lv.invy, lv.halfy, p.mode)
                       : make_tap(c, tap - R, lv.sm1x, lv.invx, lv.halfx, p.mode);
            }
          }
          __syncwarp();
          const float* slot0 = reinterpret_cast<const float*>(smem + stage * kStageBytes +
                                                              warp * (RAFTCORR_MAX_LEVELS * kSlotBytes));
#pragma unroll
          for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l) {
            if (l < p.L) {
              const Tap* tx = tab + (l * 2 + 0) * kMaxRd;
              const Tap* ty = tab + (l * 2 + 1) * kMaxRd;
              const Window w = make_window(tx[0].i0, tx[rd - 1].i0, ty[0].i0, ty[rd - 1].i0);
              if (lane < C::kActive) {
                const Tap a = tx[ti];
                const float* f = slot0 + l * (kSlotBytes / 4) + w.ox;
#pragma unroll
                for (int jj = 0; jj < C::kJGroup; ++jj) {
                  const int j = tj0 + jj;
                  if (j < rd) {
                    const Tap c = ty[j];
                    float vnw, vne, vsw, vse;
                    if (w.fast) {
                      const float* g = f + (c.i0 - w.y_lo) * w.pitch + (a.i0 - w.x_lo);
                      vnw = g[0];
                      vne = g[1];
                      vsw = g[w.pitch];
                      vse = g[w.pitch + 1];
                    } else {  // |coords| so large that fp32 spacing breaks the window
                      const LevelDev& lv = p.lv[l];
                      const long long q = (long long)b * N + (long long)y * p.W + x0 + qi;
                      const float* img = lv.data + q * lv.istride;
                      vnw = load_pixel(img, lv, c.i0, a.i0);
                      vne = load_pixel(img, lv, c.i0, a.i0 + 1);
                      vsw = load_pixel(img, lv, c.i0 + 1, a.i0);
                      vse = load_pixel(img, lv, c.i0 + 1, a.i0 + 1);
                    }
                    out_s[(l * rd2 + ti * rd + j) * kOutPitch + qi] =
                        bilinear_combine(vnw, vne, vsw, vse, a.frac, c.frac);
                  }
                }
              }
            }
          }
        }
        // this warp is done with the stage's windows: hand the buffers back to the producer
        __syncwarp();
        if (lane == 0) mbar_arrive(empty_bar(stage));
      }
      named_bar_sync(1, kConsumers * 32);
      {
        const int x = x0 + lane;
        if (x < p.W) {
          float* dst = p.out + (long long)b * nC * N + (long long)y * p.W + x;
          for (int c = warp; c < nC; c += kConsumers)
            __stcs(dst + (long long)c * N, out_s[c * kOutPitch + lane]);
        }
      }
      named_bar_sync(1, kConsumers * 32);
    }
  }
}

bool tma_eligible(const raftcorr_level* levels, int L, int r) {
  if (r != 3 && r != 4) return false;
  for (int l = 0; l < L; ++l) {
    if ((levels[l].pitch & 3) || (levels[l].image_stride & 3)) return false;
    if (reinterpret_cast<uintptr_t>(levels[l].data) & 15) return false;
  }
  return get_encode() != nullptr;
}

template <int R>
int launch_tma(const LookupTmaParams& p, const LookupTmaMaps& maps, cudaStream_t stream) {
  using C = Cfg<R>;
  RC_CHECK_CUDA(cudaFuncSetAttribute(lookup_fwd_tma_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmem));
  int dev = 0, sms = 0;
  RC_CHECK_CUDA(cudaGetDevice(&dev));
  RC_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = (int)(p.tiles < sms ? p.tiles : sms);
  lookup_fwd_tma_kernel<R><<<grid, kThreadsTma, C::kSmem, stream>>>(p, maps);
  RC_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

#ifdef RAFTCORR_DEBUG_TIMEOUT
extern "C" int raftcorr_debug_wait_timeouts(unsigned long long* out8) {
  return (int)cudaMemcpyFromSymbol(out8, g_wait_timeout, sizeof(unsigned long long) * 8);
}
#endif

// returns 1 when the TMA path does not apply (caller uses the generic-stride kernel)
int launch_lookup_fwd_tma(const raftcorr_level* levels, int L, const float* coords, float* out, int B,
                          int H, int W, int r, int mode, cudaStream_t stream) {
  if (!tma_eligible(levels, L, r)) return 1;
  LookupTmaParams p;
  LookupTmaMaps maps;
  const long long BN = (long long)B * H * W;
  for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l) {
    const int s = l < L ? l : 0;
    p.lv[l] = make_level_dev(levels[s].data, levels[s].H, levels[s].W, levels[s].pitch,
                             levels[s].image_stride, s);
    cuuint64_t dims[3] = {(cuuint64_t)levels[s].W, (cuuint64_t)levels[s].H, (cuuint64_t)BN};
    cuuint64_t strides[2] = {(cuuint64_t)levels[s].pitch * 4, (cuuint64_t)levels[s].image_stride * 4};
    for (int wi = 0; wi < kBoxWidths; ++wi)
      for (int h = 0; h < kBoxHeights; ++h) {
        cuuint32_t box[3] = {(cuuint32_t)(wi ? 16 : 12), (cuuint32_t)(10 + h), 1};
        if (int rc = encode(&maps.m[l][wi][h], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, levels[s].data, dims,
                            strides, box, CU_TENSOR_MAP_SWIZZLE_NONE, "lookup level"))
          return rc;
      }
  }
  p.coords = coords;
  p.out = out;
  p.B = B;
  p.H = H;
  p.W = W;
  p.L = L;
  p.mode = mode;
  p.tiles_x = (W + kTileQ - 1) / kTileQ;
  p.tiles = (long long)B * H * p.tiles_x;
  if (p.tiles == 0) return 0;
  return r == 4 ? launch_tma<4>(p, maps, stream) : launch_tma<3>(p, maps, stream);
}

}  // namespace raftcorr
