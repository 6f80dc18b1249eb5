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
#include <string.h>

#include <atomic>
#include <mutex>

#include "lookup_common.cuh"
#include "tma_util.cuh"

namespace raftcorr {

namespace {

using namespace lk;

constexpr int kWarpsTma = 16;                         // one query per warp per pipeline unit
constexpr int kThreadsTma = kWarpsTma * 32;
constexpr int kUnitQ = 16;                            // queries per unit (half an output tile)
constexpr int kDepth = 3;                             // window buffers per warp (2 units in flight)
constexpr int kAuxOff = 160;                          // floats; aux box (4 x 12) starts 640 B in
constexpr int kSlotBytes = 896;                       // per (warp, level): main 576 B | aux 192 B
constexpr int kWarpBufBytes = RAFTCORR_MAX_LEVELS * kSlotBytes;
constexpr int kBoxHeights = 3;                        // 10, 11, 12 rows

// TMA tile loads need the innermost start coordinate on a 16-byte boundary (x % 4 == 0 for
// fp32).  A window whose first column is x_lo is therefore fetched from bx = x_lo & ~3 as a
// 12-column box, plus a 4-column box at bx + 12 when (x_lo & 3) + width > 12.  Heights 10..12
// cover the floor flips of the coordinate round trip.  Row pitch 12 keeps the 27 active lanes
// of the sampling phase on distinct banks.
struct alignas(64) LookupTmaMaps {
  CUtensorMap main[RAFTCORR_MAX_LEVELS][kBoxHeights];
  CUtensorMap aux[RAFTCORR_MAX_LEVELS][kBoxHeights];
};

struct LookupTmaParams {
  LevelDev lv[RAFTCORR_MAX_LEVELS];
  const float* coords;
  float* out;
  int B, H, W, mode, tiles_x;
  long long tiles;
};

template <int R, int L>
struct Cfg {
  static constexpr int kRd = 2 * R + 1;
  static constexpr int kRd2 = kRd * kRd;
  static constexpr int kChannels = L * kRd2;
  static constexpr int kTaps = L * 2 * kRd;                      // tap coordinates per query
  static constexpr int kTapRounds = (kTaps + 31) / 32;
  static constexpr int kIGroups = 3;                             // x taps per lane: i = ib + 3*ii
  static constexpr int kIPer = (kRd + kIGroups - 1) / kIGroups;
  static constexpr int kActive = kRd * kIGroups;                 // lanes with work (27 or 21)
  static constexpr int kOffOut = kDepth * kWarpsTma * kWarpBufBytes;
  static constexpr int kOutBytes = ((kChannels * kOutPitch * 4 + 127) / 128) * 128;
  static constexpr int kOffTab = kOffOut + kOutBytes;
  static constexpr int kTabBytes = kWarpsTma * RAFTCORR_MAX_LEVELS * 2 * kMaxRd * 8;
  static constexpr int kOffCoords = kOffTab + kTabBytes;         // [warp][depth] float2
  static constexpr int kOffBars = kOffCoords + kWarpsTma * kDepth * 8;
  static constexpr int kSmem = kOffBars + kWarpsTma * kDepth * 8;
};

__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// Window geometry of one (query, level), derived from the first and last tap of each axis.
struct Window {
  int x_lo, y_lo;   // floor index of tap -R per axis
  int ox;           // x_lo - box start (0..3)
  int hsel;         // box height - 10
  bool need_aux;    // columns 12..15 of the box are touched
  bool fast;        // window fits the staged boxes
};

__device__ __forceinline__ Window make_window(int x_lo, int x_last, int y_lo, int y_last) {
  Window w;
  const int wdt = x_last + 2 - x_lo, hgt = y_last + 2 - y_lo;
  w.x_lo = x_lo;
  w.y_lo = y_lo;
  w.fast = (wdt >= 1) & (wdt <= kFoot) & (hgt >= 1) & (hgt <= kFoot);
  w.ox = x_lo & 3;  // two's complement: also right for negative x_lo
  w.need_aux = w.ox + wdt > kFoot;
  const int h = hgt < 10 ? 10 : (hgt > 12 ? 12 : hgt);
  w.hsel = h - 10;
  return w;
}

// rare path: |coords| so large that fp32 spacing breaks the window; sample straight from global
__device__ __noinline__ float sample_direct(const LevelDev& lv, long long q, Tap a, Tap c) {
  const float* img = lv.data + q * lv.istride;
  return bilinear_combine(load_pixel(img, lv, c.i0, a.i0), load_pixel(img, lv, c.i0, a.i0 + 1),
                          load_pixel(img, lv, c.i0 + 1, a.i0), load_pixel(img, lv, c.i0 + 1, a.i0 + 1),
                          a.frac, c.frac);
}

template <int R, int L>
__global__ void __launch_bounds__(kThreadsTma, 1)
lookup_fwd_tma_kernel(const __grid_constant__ LookupTmaParams p, const __grid_constant__ LookupTmaMaps maps) {
  using C = Cfg<R, L>;
  constexpr int rd = C::kRd, rd2 = C::kRd2;
  extern __shared__ __align__(1024) unsigned char smem[];
  const uint32_t smem_base = smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* out_s = reinterpret_cast<float*>(smem + C::kOffOut);
  Tap* tab = reinterpret_cast<Tap*>(smem + C::kOffTab) + warp * (RAFTCORR_MAX_LEVELS * 2 * kMaxRd);
  float2* coords_w = reinterpret_cast<float2*>(smem + C::kOffCoords) + warp * kDepth;
  const uint32_t bar0 = smem_base + C::kOffBars + warp * kDepth * 8;   // this warp's mbarriers

  if (lane == 0) {
    for (int s = 0; s < kDepth; ++s) mbar_init(bar0 + 8 * s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  const long long N = (long long)p.H * p.W;
  const long long n_tiles = (p.tiles - blockIdx.x + gridDim.x - 1) / gridDim.x;  // tiles of this CTA
  const long long n_units = 2 * n_tiles;

  // unit -> (b, y, x) of this warp's query
  auto locate = [&](long long u, int& b, int& y, int& x) {
    const long long tile = blockIdx.x + (u >> 1) * gridDim.x;
    const int xt = (int)(tile % p.tiles_x);
    y = (int)((tile / p.tiles_x) % p.H);
    b = (int)(tile / ((long long)p.tiles_x * p.H));
    x = xt * kTileQ + (int)(u & 1) * kUnitQ + warp;
  };
  auto load_coords = [&](long long u, float& cx, float& cy) {
    int b, y, x;
    cx = cy = 0.0f;
    if (u < n_units) {
      locate(u, b, y, x);
      if (x < p.W) {
        const long long pix = (long long)y * p.W + x;
        cx = __ldg(p.coords + ((long long)b * 2 + 0) * N + pix);
        cy = __ldg(p.coords + ((long long)b * 2 + 1) * N + pix);
      }
    }
  };
  // issue the window loads of unit u into buffer u % kDepth
  auto issue = [&](long long u, float cx, float cy) {
    int b, y, x;
    locate(u, b, y, x);
    if (x >= p.W) return;
    const int s = (int)(u % kDepth);
    if (lane == 0) coords_w[s] = make_float2(cx, cy);
    // lanes 0..4L-1: (level, {x first, x last, y first, y last}) tap floors
    int fl = 0;
    {
      const int l = (lane >> 2) < L ? (lane >> 2) : 0;
      const LevelDev& lv = p.lv[l];
      const bool is_y = lane & 2;
      const float c = __fmul_rn(is_y ? cy : cx, lv.inv_pow2);
      fl = make_tap(c, (lane & 1) ? R : -R, is_y ? lv.sm1y : lv.sm1x, is_y ? lv.invy : lv.invx,
                    is_y ? lv.halfy : lv.halfx, p.mode).i0;
    }
    const int l = lane < L ? lane : 0;
    const Window w = make_window(__shfl_sync(0xffffffffu, fl, 4 * l), __shfl_sync(0xffffffffu, fl, 4 * l + 1),
                                 __shfl_sync(0xffffffffu, fl, 4 * l + 2), __shfl_sync(0xffffffffu, fl, 4 * l + 3));
    uint32_t bytes = lane < L ? (uint32_t)((10 + w.hsel) * (48 + (w.need_aux ? 16 : 0))) : 0u;
    bytes += __shfl_xor_sync(0xffffffffu, bytes, 1);
    bytes += __shfl_xor_sync(0xffffffffu, bytes, 2);
    const uint32_t bar = bar0 + 8 * s;
    if (lane == 0) mbar_expect_tx(bar, bytes);
    __syncwarp();
    if (lane < L) {
      const LevelDev& lv = p.lv[lane];
      // windows entirely outside the image are all zeros wherever the box sits
      const int bx = max(-64, min(w.x_lo - w.ox, (lv.W + 64) & ~3));
      const int by = max(-64, min(w.y_lo, lv.H + 64));
      const int q = (int)((long long)b * N + (long long)y * p.W + x);
      const uint32_t dst = smem_base + (s * kWarpsTma + warp) * kWarpBufBytes + lane * kSlotBytes;
      tma_load_3d(dst, &maps.main[lane][w.hsel], bar, bx, by, q);
      if (w.need_aux) tma_load_3d(dst + kAuxOff * 4, &maps.aux[lane][w.hsel], bar, bx + kFoot, by, q);
    }
  };

  // per-lane description of the tap this lane evaluates in each round of the tap-table phase
  float t_sm1[C::kTapRounds], t_inv[C::kTapRounds], t_half[C::kTapRounds], t_pow[C::kTapRounds];
  int t_off[C::kTapRounds], t_slot[C::kTapRounds];
  bool t_isy[C::kTapRounds];
#pragma unroll
  for (int r = 0; r < C::kTapRounds; ++r) {
    const int t = r * 32 + lane;
    const int tt = t < C::kTaps ? t : 0;
    const int l = tt / (2 * rd);
    const int rem = tt - l * 2 * rd;
    const int axis = rem >= rd;
    const int tap = rem - axis * rd;
    const LevelDev& lv = p.lv[l];
    t_sm1[r] = axis ? lv.sm1y : lv.sm1x;
    t_inv[r] = axis ? lv.invy : lv.invx;
    t_half[r] = axis ? lv.halfy : lv.halfx;
    t_pow[r] = lv.inv_pow2;
    t_off[r] = tap - R;
    t_isy[r] = axis;
    t_slot[r] = t < C::kTaps ? (l * 2 + axis) * kMaxRd + tap : -1;
  }
  // sampling phase: lane = j * 3 + ib handles y tap j and x taps i = ib + 3 * ii
  const int tj = lane / C::kIGroups;
  const int tib = lane - tj * C::kIGroups;

  {  // prologue: two units in flight
    float cx, cy;
    for (int u = 0; u < 2; ++u)
      if (u < n_units) {
        load_coords(u, cx, cy);
        issue(u, cx, cy);
      }
  }
  float ncx, ncy;
  load_coords(2, ncx, ncy);
  uint32_t phases = 0;  // bit s = parity of the next completion of this warp's buffer s

  for (long long u = 0; u < n_units; ++u) {
    if (u + 2 < n_units) issue(u + 2, ncx, ncy);
    load_coords(u + 3, ncx, ncy);

    int b, y, x;
    locate(u, b, y, x);
    const int s = (int)(u % kDepth);
    const int qi = (int)(u & 1) * kUnitQ + warp;
    if (x < p.W) {
      mbar_wait(bar0 + 8 * s, (phases >> s) & 1u, 2);
      phases ^= 1u << s;
      const float2 cxy = coords_w[s];
#pragma unroll
      for (int r = 0; r < C::kTapRounds; ++r) {
        if (t_slot[r] >= 0) {
          const float c = __fmul_rn(t_isy[r] ? cxy.y : cxy.x, t_pow[r]);
          tab[t_slot[r]] = make_tap(c, t_off[r], t_sm1[r], t_inv[r], t_half[r], p.mode);
        }
      }
      __syncwarp();
      const float* buf = reinterpret_cast<const float*>(smem + (s * kWarpsTma + warp) * kWarpBufBytes);
#pragma unroll
      for (int l = 0; l < L; ++l) {
        const Tap* tx = tab + (l * 2 + 0) * kMaxRd;
        const Tap* ty = tab + (l * 2 + 1) * kMaxRd;
        const Window w = make_window(tx[0].i0, tx[rd - 1].i0, ty[0].i0, ty[rd - 1].i0);
        if (lane < C::kActive) {
          const Tap c = ty[tj];
          const float* f = buf + l * (kSlotBytes / 4);
          const int r0 = (c.i0 - w.y_lo) * kFoot;        // main row offset; aux rows have pitch 4
          float* o = out_s + (l * rd2 + tj) * kOutPitch + qi;
#pragma unroll
          for (int ii = 0; ii < C::kIPer; ++ii) {
            const int i = tib + C::kIGroups * ii;
            if (i < rd) {
              const Tap a = tx[i];
              float v;
              if (w.fast) {
                const int c0 = w.ox + (a.i0 - w.x_lo), c1 = c0 + 1;   // box columns of the two corners
                const int o0 = c0 < kFoot ? r0 + c0 : kAuxOff + (r0 / 3) + (c0 - kFoot);
                const int o1 = c1 < kFoot ? r0 + c1 : kAuxOff + (r0 / 3) + (c1 - kFoot);
                const int d0 = c0 < kFoot ? kFoot : 4, d1 = c1 < kFoot ? kFoot : 4;
                v = bilinear_combine(f[o0], f[o1], f[o0 + d0], f[o1 + d1], a.frac, c.frac);
              } else {
                v = sample_direct(p.lv[l], (long long)b * N + (long long)y * p.W + x, a, c);
              }
              o[i * rd * kOutPitch] = v;
            }
          }
        }
      }
    }
    if (u & 1) {  // both halves of the output tile are in shared memory: write it out
      __syncwarp();
      named_bar_sync(1, kThreadsTma);
      const long long tile = blockIdx.x + (u >> 1) * gridDim.x;
      const int x0 = (int)(tile % p.tiles_x) * kTileQ;
      const int xx = x0 + lane;
      if (xx < p.W) {
        float* dst = p.out + (long long)b * C::kChannels * N + (long long)y * p.W + xx;
#pragma unroll 4
        for (int c = warp; c < C::kChannels; c += kWarpsTma)
          __stcs(dst + (long long)c * N, out_s[c * kOutPitch + lane]);
      }
      named_bar_sync(1, kThreadsTma);
    } else {
      __syncwarp();
    }
  }
}

struct MapKey {
  const void* data[RAFTCORR_MAX_LEVELS];
  long long pitch[RAFTCORR_MAX_LEVELS], istride[RAFTCORR_MAX_LEVELS], bn;
  int h[RAFTCORR_MAX_LEVELS], w[RAFTCORR_MAX_LEVELS];
};
constexpr int kMapCacheSize = 8;
struct MapCache {
  std::mutex mu;
  MapKey keys[kMapCacheSize];
  LookupTmaMaps maps[kMapCacheSize];
  unsigned long long stamp[kMapCacheSize] = {};
  unsigned long long clock = 0;
};
MapCache& map_cache() {
  static MapCache c;
  return c;
}
bool map_cache_get(const MapKey& k, LookupTmaMaps* out) {
  MapCache& c = map_cache();
  std::lock_guard<std::mutex> g(c.mu);
  for (int i = 0; i < kMapCacheSize; ++i)
    if (c.stamp[i] && memcmp(&c.keys[i], &k, sizeof(MapKey)) == 0) {
      c.stamp[i] = ++c.clock;
      *out = c.maps[i];
      return true;
    }
  return false;
}
void map_cache_put(const MapKey& k, const LookupTmaMaps& m) {
  MapCache& c = map_cache();
  std::lock_guard<std::mutex> g(c.mu);
  int victim = 0;
  for (int i = 1; i < kMapCacheSize; ++i)
    if (c.stamp[i] < c.stamp[victim]) victim = i;
  c.keys[victim] = k;
  c.maps[victim] = m;
  c.stamp[victim] = ++c.clock;
}

bool tma_eligible(const raftcorr_level* levels, int L, int r) {
  if (L != RAFTCORR_MAX_LEVELS || (r != 3 && r != 4)) return false;
  for (int l = 0; l < L; ++l) {
    if ((levels[l].pitch & 3) || (levels[l].image_stride & 3)) return false;
    if (reinterpret_cast<uintptr_t>(levels[l].data) & 15) return false;
  }
  return get_encode() != nullptr;
}

template <int R, int L>
int launch_tma(const LookupTmaParams& p, const LookupTmaMaps& maps, cudaStream_t stream) {
  using C = Cfg<R, L>;
  int dev = 0;
  RC_CHECK_CUDA(cudaGetDevice(&dev));
  static std::atomic<int> sm_count[64];  // per device: 0 = kernel attribute not set yet
  int sms = (dev >= 0 && dev < 64) ? sm_count[dev].load(std::memory_order_acquire) : 0;
  if (sms == 0) {
    RC_CHECK_CUDA(cudaFuncSetAttribute(lookup_fwd_tma_kernel<R, L>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmem));
    RC_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (dev >= 0 && dev < 64) sm_count[dev].store(sms, std::memory_order_release);
  }
  const int grid = (int)(p.tiles < sms ? p.tiles : sms);
  lookup_fwd_tma_kernel<R, L><<<grid, kThreadsTma, C::kSmem, stream>>>(p, maps);
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
  const long long BN = (long long)B * H * W;
  if (BN >= (1LL << 31)) return 1;
  for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l)
    p.lv[l] = make_level_dev(levels[l].data, levels[l].H, levels[l].W, levels[l].pitch,
                             levels[l].image_stride, l);
  // Tensor maps depend only on (pointers, sizes, strides); encoding 24 of them costs tens of
  // microseconds of host time, more than the kernel itself, and RAFT calls the lookup once per
  // refinement iteration on the same pyramid: keep the last few sets (mutex-protected, so the
  // library stays re-entrant under nn.DataParallel's one-thread-per-GPU calls).
  MapKey key;
  memset(&key, 0, sizeof(key));
  key.bn = BN;
  for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l) {
    key.data[l] = levels[l].data;
    key.h[l] = levels[l].H;
    key.w[l] = levels[l].W;
    key.pitch[l] = levels[l].pitch;
    key.istride[l] = levels[l].image_stride;
  }
  LookupTmaMaps maps;
  if (!map_cache_get(key, &maps)) {
    for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l) {
      cuuint64_t dims[3] = {(cuuint64_t)levels[l].W, (cuuint64_t)levels[l].H, (cuuint64_t)BN};
      cuuint64_t strides[2] = {(cuuint64_t)levels[l].pitch * 4, (cuuint64_t)levels[l].image_stride * 4};
      for (int h = 0; h < kBoxHeights; ++h) {
        cuuint32_t box[3] = {(cuuint32_t)kFoot, (cuuint32_t)(10 + h), 1};
        if (int rc = encode(&maps.main[l][h], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, levels[l].data, dims, strides,
                            box, CU_TENSOR_MAP_SWIZZLE_NONE, "lookup window", CU_TENSOR_MAP_L2_PROMOTION_NONE))
          return rc;
        box[0] = 4;
        if (int rc = encode(&maps.aux[l][h], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, levels[l].data, dims, strides,
                            box, CU_TENSOR_MAP_SWIZZLE_NONE, "lookup window tail", CU_TENSOR_MAP_L2_PROMOTION_NONE))
          return rc;
      }
    }
    map_cache_put(key, maps);
  }
  p.coords = coords;
  p.out = out;
  p.B = B;
  p.H = H;
  p.W = W;
  p.mode = mode;
  p.tiles_x = (W + kTileQ - 1) / kTileQ;
  p.tiles = (long long)B * H * p.tiles_x;
  if (p.tiles == 0) return 0;
  return r == 4 ? launch_tma<4, 4>(p, maps, stream) : launch_tma<3, 4>(p, maps, stream);
}

}  // namespace raftcorr
