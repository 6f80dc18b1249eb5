// lookup_tma.cu — the pyramid window lookup as a persistent TMA-gather pipeline (sm_100a).
//
// Same arithmetic as lookup.cu (reference core/corr.py:29-50 + core/utils/utils.py:57-65 +
// ATen grid_sampler_2d), different data movement.  The lookup is a gather: each query reads a
// ~10x10 window out of its OWN H_l x W_l image at each of the 4 levels — 40-byte row pieces
// scattered over gigabytes.  Instead of having warps chase those rows through the LSU, every
// (query, level) window becomes ONE TMA tensor-tile load (cp.async.bulk.tensor.3d, box
// 16 x {10,11,12} x 1 out of the (W_l, H_l, B*N) level tensor): the TMA engine generates the row
// requests, zero-fills everything outside the image (that IS grid_sample's zero padding) and
// signals an mbarrier.
//
//   CTA = 1 SM, persistent over tiles of 32 consecutive-x queries (two units of 16 queries)
//   warps 16..23   staging: per query the 72 tap coordinates (exact fp32 op sequence), the window
//                  of each level, 4 TMA loads into the ring slot of the sampling warp they feed
//   warps 0..15    sampling: one query per warp per unit out of a 3-slot ring (full/empty
//                  mbarriers), 81 x 4 bilinear samples from shared memory, then the CTA writes the
//                  NCHW tile as full 128-byte rows
#include <string.h>

#include <atomic>
#include <mutex>

#include "lookup_common.cuh"
#include "tma_util.cuh"

namespace raftcorr {

namespace {

using namespace lk;

constexpr int kWarpsTma = 16;                         // sampling warps: one query per warp per unit
constexpr int kStagers = 8;                           // staging warps: each feeds two sampling warps
constexpr int kSamplerThreads = kWarpsTma * 32;
constexpr int kThreadsTma = (kWarpsTma + kStagers) * 32;
constexpr int kUnitQ = 16;                            // queries per unit (half an output tile)
constexpr int kDepth = 3;                             // units buffered per warp (2 in flight)
constexpr int kBoxW = 16;                             // staged window: 16 columns ...
constexpr int kSlotBytes = kBoxW * kFoot * 4;         // ... x up to 12 rows of fp32 = 768 B
constexpr int kWarpBufBytes = RAFTCORR_MAX_LEVELS * kSlotBytes;
constexpr int kBoxHeights = 3;                        // 10, 11, 12 rows
constexpr int kTabEntries = RAFTCORR_MAX_LEVELS * 2 * kMaxRd;

// TMA tile loads need the innermost start coordinate on a 16-byte boundary (x % 4 == 0 for
// fp32): a window whose first column is x_lo is fetched from bx = x_lo & ~3 as a 16-column box
// (DRAM moves 64-byte granules anyway).  Heights 10..12 cover the floor flips of the coordinate
// round trip.  One tensor map per (level, height).
struct alignas(64) LookupTmaMaps {
  CUtensorMap m[RAFTCORR_MAX_LEVELS][kBoxHeights];
};

struct LookupTmaParams {
  LevelDev lv[RAFTCORR_MAX_LEVELS];
  const float* coords;
  float* out;
  int B, H, W, tiles_x;
  long long tiles;
  int nhwc;   // 1: out is (B, H, W, C) - each query's C samples contiguous, no transpose needed
};

template <int R, int L>
struct Cfg {
  static constexpr int kRd = 2 * R + 1;
  static constexpr int kRd2 = kRd * kRd;
  static constexpr int kChannels = L * kRd2;
  static constexpr int kTaps = L * 2 * kRd;                      // tap coordinates per query
  static constexpr int kTapRounds = (kTaps + 31) / 32;
  static constexpr int kJGroups = 3;                             // lane = i * 3 + jb, j = 3 * jb + jj
  static constexpr int kJPer = (kRd + kJGroups - 1) / kJGroups;
  static constexpr int kActive = kRd * kJGroups;                 // lanes with work (27 or 21)
  static constexpr int kOffOut = kDepth * kWarpsTma * kWarpBufBytes;
  static constexpr int kOutBytes = ((kChannels * kOutPitch * 4 + 127) / 128) * 128;
  static constexpr int kOffTab = kOffOut + kOutBytes;            // [warp][depth][kTabEntries] Tap
  static constexpr int kTabBytes = kWarpsTma * kDepth * kTabEntries * 8;
  static constexpr int kOffWin = kOffTab + kTabBytes;            // [warp][depth][L] int4
  static constexpr int kWinBytes = kWarpsTma * kDepth * RAFTCORR_MAX_LEVELS * 16;
  static constexpr int kOffBars = kOffWin + kWinBytes;
  static constexpr int kSmem = kOffBars + 2 * kWarpsTma * kDepth * 8;   // full[w][s], empty[w][s]
};

__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// rare path: |coords| so large that fp32 spacing breaks the window; sample straight from global
__device__ __noinline__ float sample_direct(const LevelDev& lv, long long q, Tap a, Tap c) {
  const float* img = lv.data + q * lv.istride;
  return bilinear_combine(load_pixel(img, lv, c.i0, a.i0), load_pixel(img, lv, c.i0, a.i0 + 1),
                          load_pixel(img, lv, c.i0 + 1, a.i0), load_pixel(img, lv, c.i0 + 1, a.i0 + 1),
                          a.frac, c.frac);
}

// A query with at least one window that does not fit the staged box (|coords| beyond ~2^20, where
// fp32 spacing breaks the tap pattern): per-sample choice between the staged box and global memory.
template <int R, int L>
__device__ __noinline__ void sample_query_slow(const LookupTmaParams& p, const Tap* tab, const int4* wv,
                                               const float* buf, float* out_s, long long q, int qi, int lane,
                                               int out_stride) {
  constexpr int rd = 2 * R + 1, rd2 = rd * rd;
  for (int l = 0; l < L; ++l) {
    const Tap* tx = tab + (l * 2 + 0) * kMaxRd;
    const Tap* ty = tab + (l * 2 + 1) * kMaxRd;
    for (int k = lane; k < rd2; k += 32) {
      const int i = k / rd, j = k - i * rd;
      const Tap a = tx[i], c = ty[j];
      float v;
      if (wv[l].w) {
        const float* gp = buf + l * (kSlotBytes / 4) + (c.i0 - wv[l].y) * kBoxW + (a.i0 + wv[l].x);
        v = bilinear_combine(gp[0], gp[1], gp[kBoxW], gp[kBoxW + 1], a.frac, c.frac);
      } else {
        v = sample_direct(p.lv[l], q, a, c);
      }
      out_s[(l * rd2 + k) * out_stride + qi] = v;
    }
  }
}

template <int R, int L, int kMode, bool kNhwc>
__global__ void __launch_bounds__(kThreadsTma, 1)
lookup_fwd_tma_kernel(const __grid_constant__ LookupTmaParams p, const __grid_constant__ LookupTmaMaps maps) {
  using C = Cfg<R, L>;
  constexpr int rd = C::kRd, rd2 = C::kRd2;
  extern __shared__ __align__(1024) unsigned char smem[];
  const uint32_t smem_base = smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* out_s = reinterpret_cast<float*>(smem + C::kOffOut);
  auto tab_of = [&](int w) { return reinterpret_cast<Tap*>(smem + C::kOffTab) + w * (kDepth * kTabEntries); };
  auto win_of = [&](int w) { return reinterpret_cast<int4*>(smem + C::kOffWin) + w * (kDepth * RAFTCORR_MAX_LEVELS); };
  auto full_bar = [&](int w, int s) { return smem_base + C::kOffBars + (w * kDepth + s) * 8; };
  auto empty_bar = [&](int w, int s) { return smem_base + C::kOffBars + ((kWarpsTma + w) * kDepth + s) * 8; };

  if (threadIdx.x < kWarpsTma) {
    for (int s = 0; s < kDepth; ++s) {
      mbar_init(full_bar(threadIdx.x, s), 1);    // stager's arrive.expect_tx (+ TMA bytes)
      mbar_init(empty_bar(threadIdx.x, s), 1);   // sampler's lane 0
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  const long long N = (long long)p.H * p.W;
  const int n_tiles = (int)((p.tiles - blockIdx.x + gridDim.x - 1) / gridDim.x);  // tiles of this CTA

  // 32-bit tile arithmetic only (64-bit div/mod costs ~40 instructions each on the SM)
  struct Cursor {
    int b, y, x0;
    bool live;
  };
  auto cursor_at = [&](int k) {
    Cursor c;
    c.live = k < n_tiles;
    const unsigned tile = blockIdx.x + (unsigned)(c.live ? k : 0) * gridDim.x;
    const unsigned t2 = tile / (unsigned)p.tiles_x;
    c.x0 = (int)(tile - t2 * (unsigned)p.tiles_x) * kTileQ;
    c.b = (int)(t2 / (unsigned)p.H);
    c.y = (int)(t2 - (unsigned)c.b * (unsigned)p.H);
    return c;
  };

  if (warp >= kWarpsTma) {
    // ================================ staging warps ================================
    // Warp g feeds sampling warps 2g and 2g+1: tap table -> window geometry -> one TMA box per
    // level, into ring slot (unit % kDepth) of that sampling warp.
    const int g = warp - kWarpsTma;
    float t_sm1[C::kTapRounds], t_inv[C::kTapRounds], t_half[C::kTapRounds], t_pow[C::kTapRounds], t_off[C::kTapRounds];
    int t_slot[C::kTapRounds];
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
      t_off[r] = (float)(tap - R);
      t_isy[r] = axis;
      t_slot[r] = t < C::kTaps ? (l * 2 + axis) * kMaxRd + tap : -1;
    }
    // lanes 0..3 fetch (x, y) of the two queries this warp stages in a unit
    auto load_coords = [&](const Cursor& c, int half) -> float {
      float v = 0.0f;
      const int x = c.x0 + half * kUnitQ + 2 * g + (lane >> 1);
      if (c.live && lane < 4 && x < p.W)
        v = __ldg(p.coords + ((long long)c.b * 2 + (lane & 1)) * N + (long long)c.y * p.W + x);
      return v;
    };
    uint32_t ephase = 0x3fu;  // bit (cc*kDepth+s): parity to wait for on empty[2g+cc][s]; fresh = passes
    int slot = 0;
    Cursor cur = cursor_at(0), nxt = cursor_at(1);
    float cnow = load_coords(cur, 0);
    for (int k = 0; k < n_tiles; ++k) {
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const float cnext = half == 0 ? load_coords(cur, 1) : load_coords(nxt, 0);   // one unit ahead
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int w = 2 * g + cc;
          const int x = cur.x0 + half * kUnitQ + w;
          if (x < p.W) {
            const float cx = __shfl_sync(0xffffffffu, cnow, 2 * cc), cy = __shfl_sync(0xffffffffu, cnow, 2 * cc + 1);
            const int bit = cc * kDepth + slot;
            mbar_wait_relaxed(empty_bar(w, slot), (ephase >> bit) & 1u, 1);
            ephase ^= 1u << bit;
            Tap* tab = tab_of(w) + slot * kTabEntries;
            int4* win = win_of(w) + slot * RAFTCORR_MAX_LEVELS;
#pragma unroll
            for (int r = 0; r < C::kTapRounds; ++r)
              if (t_slot[r] >= 0)
                tab[t_slot[r]] = make_tap_fast(__fmul_rn(t_isy[r] ? cy : cx, t_pow[r]), t_off[r], t_sm1[r],
                                               t_inv[r], t_half[r], kMode);
            __syncwarp();
            if (lane < L) {
              const Tap* tx = tab + (lane * 2 + 0) * kMaxRd;
              const Tap* ty = tab + (lane * 2 + 1) * kMaxRd;
              const int x_lo = tx[0].i0, y_lo = ty[0].i0;
              const int wdt = tx[rd - 1].i0 + 2 - x_lo, hgt = ty[rd - 1].i0 + 2 - y_lo;
              const int fast = (wdt >= 1) & (wdt <= kFoot) & (hgt >= 1) & (hgt <= kFoot);
              const int h = hgt < 10 ? 10 : (hgt > 12 ? 12 : hgt);
              // .x: box column of a tap with floor index i0 is (i0 + .x); .y: first row; .z: box height; .w: fast
              win[lane] = make_int4((x_lo & 3) - x_lo, y_lo, h, fast);
            }
            __syncwarp();
            if (elect_one()) {
              const uint32_t bar = full_bar(w, slot);
              const int q = (int)((long long)cur.b * N + (long long)cur.y * p.W + x);
              const uint32_t dst = smem_base + (slot * kWarpsTma + w) * kWarpBufBytes;
              int4 wv[L];
              uint32_t bytes = 0;
#pragma unroll
              for (int l = 0; l < L; ++l) {
                wv[l] = win[l];
                bytes += (uint32_t)(wv[l].z * kBoxW * 4);
              }
              mbar_expect_tx(bar, bytes);   // release: the tap table and windows above are visible to the waiter
#pragma unroll
              for (int l = 0; l < L; ++l) {
                // box start = x_lo & ~3 = -(w.x); windows entirely outside the image are all zeros
                // wherever the box sits, so far-away origins are clamped to keep the coordinates sane
                const int bx = max(-64, min(-wv[l].x, (p.lv[l].W + 64) & ~3));
                const int by = max(-64, min(wv[l].y, p.lv[l].H + 64));
                const CUtensorMap* m = &maps.m[l][0] + (wv[l].z - 10);
                tma_load_3d(dst + l * kSlotBytes, m, bar, bx, by, q);
              }
            }
            __syncwarp();
          }
        }
        cnow = cnext;
        slot = slot + 1 == kDepth ? 0 : slot + 1;
      }
      cur = nxt;
      nxt = cursor_at(k + 2);
    }
  } else {
    // ================================ sampling warps ================================
    // lane = i * 3 + jb handles x tap i and y taps j = 3 * jb + jj of this warp's query
    const int ti = lane / C::kJGroups;
    const int tj0 = (lane - ti * C::kJGroups) * C::kJPer;
    const Tap* tab_w = tab_of(warp);
    const int4* win_w = win_of(warp);
    uint32_t phases = 0;   // bit s = parity of the next completion of full[warp][s]
    int slot = 0;
    Cursor cur = cursor_at(0);
    for (int k = 0; k < n_tiles; ++k) {
      const Cursor nxt = cursor_at(k + 1);
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int x = cur.x0 + half * kUnitQ + warp;
        const int qi = half * kUnitQ + warp;
        if (x < p.W) {
          mbar_wait(full_bar(warp, slot), (phases >> slot) & 1u, 2);
          phases ^= 1u << slot;
          const Tap* tab = tab_w + slot * kTabEntries;
          const float* buf = reinterpret_cast<const float*>(smem + (slot * kWarpsTma + warp) * kWarpBufBytes);
          const long long qglob = (long long)cur.b * N + (long long)cur.y * p.W + x;
          // NCHW: park in the transposing tile; NHWC: the query's C samples are contiguous in global memory
          float* obase = kNhwc ? p.out + qglob * C::kChannels : out_s + qi;
          constexpr int ostride = kNhwc ? 1 : kOutPitch;
          int4 wv[L];
          int all_fast = 1;
#pragma unroll
          for (int l = 0; l < L; ++l) {
            wv[l] = win_w[slot * RAFTCORR_MAX_LEVELS + l];
            all_fast &= wv[l].w;
          }
          if (all_fast) {
            // common case, branch-free so the 12 samples of a lane overlap their shared-memory latency
            if (lane < C::kActive) {
#pragma unroll
              for (int l = 0; l < L; ++l) {
                const Tap a = tab[(l * 2 + 0) * kMaxRd + ti];
                const Tap* ty = tab + (l * 2 + 1) * kMaxRd;
                const float* f = buf + l * (kSlotBytes / 4) + (a.i0 + wv[l].x) - wv[l].y * kBoxW;
                float* o = obase + (l * rd2 + ti * rd) * ostride;
                // this lane's y taps; when their rows are consecutive (no floor flip inside the
                // group, the usual case) adjacent samples share a row: 8 loads instead of 12
                Tap cj[C::kJPer];
#pragma unroll
                for (int jj = 0; jj < C::kJPer; ++jj) cj[jj] = ty[(tj0 + jj < rd) ? tj0 + jj : rd - 1];
                bool chain = true;
#pragma unroll
                for (int jj = 1; jj < C::kJPer; ++jj) chain &= cj[jj].i0 == cj[jj - 1].i0 + 1;
                if (__all_sync(__activemask(), chain)) {
                  const float* gp = f + cj[0].i0 * kBoxW;
                  float r0[C::kJPer + 1], r1[C::kJPer + 1];
#pragma unroll
                  for (int t = 0; t <= C::kJPer; ++t) {
                    r0[t] = gp[t * kBoxW];
                    r1[t] = gp[t * kBoxW + 1];
                  }
#pragma unroll
                  for (int jj = 0; jj < C::kJPer; ++jj) {
                    const float v = bilinear_combine(r0[jj], r1[jj], r0[jj + 1], r1[jj + 1], a.frac, cj[jj].frac);
                    if (tj0 + jj < rd) o[(tj0 + jj) * ostride] = v;
                  }
                } else {
#pragma unroll
                  for (int jj = 0; jj < C::kJPer; ++jj) {
                    const float* gp = f + cj[jj].i0 * kBoxW;
                    const float v = bilinear_combine(gp[0], gp[1], gp[kBoxW], gp[kBoxW + 1], a.frac, cj[jj].frac);
                    if (tj0 + jj < rd) o[(tj0 + jj) * ostride] = v;
                  }
                }
              }
            }
          } else {
            sample_query_slow<R, L>(p, tab, win_w + slot * RAFTCORR_MAX_LEVELS, buf, kNhwc ? obase : out_s, qglob,
                                    kNhwc ? 0 : qi, lane, ostride);
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(empty_bar(warp, slot));   // ring slot may be refilled
        }
        slot = slot + 1 == kDepth ? 0 : slot + 1;
      }
      // both halves of the output tile are in shared memory: write it out as 128-byte rows
      if (kNhwc) {
        cur = nxt;
        continue;
      }
      named_bar_sync(1, kSamplerThreads);
      {
        const int xx = cur.x0 + lane;
        if (xx < p.W) {
          float* dst = p.out + (long long)cur.b * C::kChannels * N + (long long)cur.y * p.W + xx;
#pragma unroll 4
          for (int c = warp; c < C::kChannels; c += kWarpsTma)
            __stcs(dst + (long long)c * N, out_s[c * kOutPitch + lane]);
        }
      }
      named_bar_sync(1, kSamplerThreads);
      cur = nxt;
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

template <int R, int L, int kMode, bool kNhwc>
int launch_tma_k(const LookupTmaParams& p, const LookupTmaMaps& maps, cudaStream_t stream) {
  using C = Cfg<R, L>;
  int dev = 0;
  RC_CHECK_CUDA(cudaGetDevice(&dev));
  static std::atomic<int> sm_count[64];  // per device: 0 = kernel attribute not set yet
  int sms = (dev >= 0 && dev < 64) ? sm_count[dev].load(std::memory_order_acquire) : 0;
  if (sms == 0) {
    RC_CHECK_CUDA(cudaFuncSetAttribute(lookup_fwd_tma_kernel<R, L, kMode, kNhwc>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmem));
    RC_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (dev >= 0 && dev < 64) sm_count[dev].store(sms, std::memory_order_release);
  }
  const int grid = (int)(p.tiles < sms ? p.tiles : sms);
  lookup_fwd_tma_kernel<R, L, kMode, kNhwc><<<grid, kThreadsTma, C::kSmem, stream>>>(p, maps);
  RC_CHECK_CUDA(cudaGetLastError());
  return 0;
}

template <int R, int L, int kMode>
int launch_tma(const LookupTmaParams& p, const LookupTmaMaps& maps, cudaStream_t stream) {
  return p.nhwc ? launch_tma_k<R, L, kMode, true>(p, maps, stream) : launch_tma_k<R, L, kMode, false>(p, maps, stream);
}

}  // namespace

#ifdef RAFTCORR_DEBUG_TIMEOUT
extern "C" int raftcorr_debug_wait_timeouts(unsigned long long* out8) {
  return (int)cudaMemcpyFromSymbol(out8, g_wait_timeout, sizeof(unsigned long long) * 8);
}
#endif

// returns 1 when the TMA path does not apply (caller uses the generic-stride kernel)
int launch_lookup_fwd_tma(const raftcorr_level* levels, int L, const float* coords, float* out, int B,
                          int H, int W, int r, int mode, int nhwc, cudaStream_t stream) {
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
        cuuint32_t box[3] = {(cuuint32_t)kBoxW, (cuuint32_t)(10 + h), 1};
        if (int rc = encode(&maps.m[l][h], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, levels[l].data, dims, strides,
                            box, CU_TENSOR_MAP_SWIZZLE_NONE, "lookup window", CU_TENSOR_MAP_L2_PROMOTION_NONE))
          return rc;
      }
    }
    map_cache_put(key, maps);
  }
  p.coords = coords;
  p.out = out;
  p.nhwc = nhwc;
  p.B = B;
  p.H = H;
  p.W = W;
  p.tiles_x = (W + kTileQ - 1) / kTileQ;
  p.tiles = (long long)B * H * p.tiles_x;
  if (p.tiles == 0) return 0;
  switch (mode * 8 + r) {
    case RAFTCORR_COORD_ATEN_CPU * 8 + 4: return launch_tma<4, 4, RAFTCORR_COORD_ATEN_CPU>(p, maps, stream);
    case RAFTCORR_COORD_ATEN_CPU * 8 + 3: return launch_tma<3, 4, RAFTCORR_COORD_ATEN_CPU>(p, maps, stream);
    case RAFTCORR_COORD_ATEN_CUDA * 8 + 4: return launch_tma<4, 4, RAFTCORR_COORD_ATEN_CUDA>(p, maps, stream);
    case RAFTCORR_COORD_ATEN_CUDA * 8 + 3: return launch_tma<3, 4, RAFTCORR_COORD_ATEN_CUDA>(p, maps, stream);
    case RAFTCORR_COORD_DIRECT * 8 + 4: return launch_tma<4, 4, RAFTCORR_COORD_DIRECT>(p, maps, stream);
    default: return launch_tma<3, 4, RAFTCORR_COORD_DIRECT>(p, maps, stream);
  }
}

}  // namespace raftcorr
