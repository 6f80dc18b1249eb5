// lookup_bwd_tma.cu — backward of the window lookup w.r.t. the pyramid as TMA reduce-add stores (sm_100a).
//
// The reference gets this from autograd (GridSampler2DBackward0 per level per iteration):
// every output gradient is scattered with its four bilinear weights into a zero-filled, volume-sized
// tensor.  Here the scatter of one (query, level) is first assembled on chip: the (2r+1)^2 gradients of
// the window, spread through the 2x2 bilinear stencil, are a (2r+2)^2 footprint of the query's own image,
// written into a 16 x 10 shared-memory box.  One cp.reduce.async.bulk.tensor (.add) per box then adds it
// into the accumulated d(pyramid): the read-modify-write happens in L2, no atomic instructions are issued,
// clipping at the image border (= the zero padding of the forward) is done by the TMA unit, and because
// each query owns its image the sums are deterministic.
//
// The assembly needs the taps of a window to sit on consecutive pixels (true except where the fp32
// coordinate round trip flips a floor, DESIGN.md §4, and for absurd coordinates); a query where they do
// not takes the atomic scatter of lookup.cu instead, so both paths add exactly the same terms.
//
//   CTA = 1 SM, persistent over tiles of 32 consecutive-x queries; 16 warps, two queries per warp per tile;
//   the (C x 32) tile of output gradients is staged through shared memory (double-buffered, one CTA barrier
//   per tile) so that global reads are full 128-byte rows.
#include <atomic>

#include "lookup_common.cuh"
#include "tma_util.cuh"

namespace raftcorr {

namespace {

using namespace lk;

constexpr int kBWarps = 16;
constexpr int kBThreads = kBWarps * 32;
constexpr int kBoxW = 16;                      // box columns (the window starts at column x_lo & 3)
constexpr int kBoxH = 10;                      // 2r+2 rows for r = 4
constexpr int kBoxFloats = kBoxW * kBoxH;
constexpr int kBoxBytes = kBoxFloats * 4;      // 640
constexpr int kSlots = 2;                      // boxes sets in flight per warp
constexpr int kTabEntries = RAFTCORR_MAX_LEVELS * 2 * kMaxRd;

struct alignas(64) BwdMaps {
  CUtensorMap m[RAFTCORR_MAX_LEVELS];
};

struct BwdTmaParams {
  LevelDev glv[RAFTCORR_MAX_LEVELS];   // the accumulated d(pyramid)
  const float* coords;
  const float* gout;
  int B, H, W, tiles_x;
  long long tiles;
};

template <int R>
struct BCfg {
  static constexpr int kRd = 2 * R + 1;
  static constexpr int kRd2 = kRd * kRd;
  static constexpr int kChannels = RAFTCORR_MAX_LEVELS * kRd2;
  static constexpr int kTaps = RAFTCORR_MAX_LEVELS * 2 * kRd;
  static constexpr int kTapRounds = (kTaps + 31) / 32;
  static constexpr int kGBytes = ((kChannels * kOutPitch * 4 + 127) / 128) * 128;   // one gradient tile
  static constexpr int kOffSlots = 2 * kGBytes;
  static constexpr int kSlotBytesAll = kBWarps * kSlots * RAFTCORR_MAX_LEVELS * kBoxBytes;
  static constexpr int kOffTab = kOffSlots + kSlotBytesAll;
  static constexpr int kOffWin = kOffTab + kBWarps * kTabEntries * 8;
  static constexpr int kSmem = kOffWin + kBWarps * RAFTCORR_MAX_LEVELS * 16;
};

__device__ __forceinline__ void tma_reduce_add_3d(const CUtensorMap* map, uint32_t src, int c0, int c1, int c2) {
  asm volatile("cp.reduce.async.bulk.tensor.3d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4}], [%1];"
               ::"l"(map), "r"(src), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}

template <int R, int kMode>
__global__ void __launch_bounds__(kBThreads, 1)
lookup_bwd_tma_kernel(const __grid_constant__ BwdTmaParams p, const __grid_constant__ BwdMaps maps) {
  using C = BCfg<R>;
  constexpr int rd = C::kRd, rd2 = C::kRd2;
  extern __shared__ __align__(1024) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* slots = reinterpret_cast<float*>(smem + C::kOffSlots) + warp * (kSlots * RAFTCORR_MAX_LEVELS * kBoxFloats);
  Tap* tab = reinterpret_cast<Tap*>(smem + C::kOffTab) + warp * kTabEntries;
  int4* win = reinterpret_cast<int4*>(smem + C::kOffWin) + warp * RAFTCORR_MAX_LEVELS;
  const long long N = (long long)p.H * p.W;

  // which tap coordinate each lane computes in each round (as in the forward's staging warps)
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
    const LevelDev& lv = p.glv[l];
    t_sm1[r] = axis ? lv.sm1y : lv.sm1x;
    t_inv[r] = axis ? lv.invy : lv.invx;
    t_half[r] = axis ? lv.halfy : lv.halfx;
    t_pow[r] = lv.inv_pow2;
    t_off[r] = (float)(tap - R);
    t_isy[r] = axis;
    t_slot[r] = t < C::kTaps ? (l * 2 + axis) * kMaxRd + tap : -1;
  }
  // box cell ownership: lane -> column (lane & 15) and rows 5*(lane >> 4) .. +4
  const int col = lane & 15, half = lane >> 4;

  // (C x 32) tile of output gradients -> shared memory with cp.async (zero-filled beyond the row end); the
  // copy of tile k+1 is in flight while tile k is processed
  auto decode = [&](long long tile, int& b, int& y, int& x0) {
    const unsigned t32 = (unsigned)tile;
    const unsigned t2 = t32 / (unsigned)p.tiles_x;
    x0 = (int)(t32 - t2 * (unsigned)p.tiles_x) * kTileQ;
    b = (int)(t2 / (unsigned)p.H);
    y = (int)(t2 - (unsigned)b * (unsigned)p.H);
  };
  auto prefetch_tile = [&](long long tile, unsigned buf) {
    int b, y, x0;
    decode(tile, b, y, x0);
    const int x = x0 + lane;
    const bool in = x < p.W;
    const float* src = p.gout + (long long)b * C::kChannels * N + (long long)y * p.W + (in ? x : 0);
    const uint32_t dst = smem_u32(smem + buf * C::kGBytes) + lane * 4;
    const uint32_t bytes = in ? 4u : 0u;
    for (int c = warp; c < C::kChannels; c += kBWarps)
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst + c * (kOutPitch * 4)),
                   "l"(src + (long long)c * N), "r"(bytes)
                   : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  unsigned use = 0;   // boxes sets this warp has sent
  unsigned k = 0;
  if ((long long)blockIdx.x < p.tiles) prefetch_tile(blockIdx.x, 0);
  for (long long tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++k) {
    int b, y, x0;
    decode(tile, b, y, x0);
    const float* g_s = reinterpret_cast<const float*>(smem + (k & 1u) * C::kGBytes);
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncthreads();   // tile k is complete; every warp is done with tile k-1, whose buffer is refilled next
    if (tile + gridDim.x < p.tiles) prefetch_tile(tile + gridDim.x, (k + 1) & 1u);

#pragma unroll 1
    for (int h2 = 0; h2 < 2; ++h2) {
      const int qi = warp + kBWarps * h2;
      const int x = x0 + qi;
      if (x >= p.W) continue;
      const long long pix = (long long)y * p.W + x;
      const int q = (int)((long long)b * N + pix);
      const float cx = __ldg(p.coords + ((long long)b * 2 + 0) * N + pix);
      const float cy = __ldg(p.coords + ((long long)b * 2 + 1) * N + pix);
      int my_i0[C::kTapRounds];
#pragma unroll
      for (int r = 0; r < C::kTapRounds; ++r) {
        my_i0[r] = 0;
        if (t_slot[r] >= 0) {
          const Tap t = make_tap_fast(__fmul_rn(t_isy[r] ? cy : cx, t_pow[r]), t_off[r], t_sm1[r], t_inv[r],
                                      t_half[r], kMode);
          tab[t_slot[r]] = t;
          my_i0[r] = t.i0;
        }
      }
      __syncwarp();
      // regular = every tap sits exactly (tap index) pixels after the first tap of its level and axis
      bool regular = true;
#pragma unroll
      for (int r = 0; r < C::kTapRounds; ++r)
        if (t_slot[r] >= 0) {
          const int tapidx = (int)t_off[r] + R;
          regular &= my_i0[r] == tab[t_slot[r] - tapidx].i0 + tapidx;
        }
      if (lane < RAFTCORR_MAX_LEVELS) {
        const int x_lo = tab[(lane * 2 + 0) * kMaxRd].i0, y_lo = tab[(lane * 2 + 1) * kMaxRd].i0;
        regular &= (x_lo > -(1 << 28)) & (x_lo < (1 << 28)) & (y_lo > -(1 << 28)) & (y_lo < (1 << 28));
        // TMA stores/reductions take no negative coordinates (tools/probe/tma_reduce_probe.cu) and, like loads,
        // need the first column on 16 bytes: the box starts at the aligned-down window origin clamped into the
        // image, and the taps are shifted inside it.  Origins far outside only ever add zeros: keep them sane.
        const int bx = min(max(x_lo & ~3, 0), (p.glv[lane].W + 64) & ~3);
        const int by = min(max(y_lo, 0), p.glv[lane].H + 64);
        win[lane] = make_int4(bx, by, bx - x_lo, by - y_lo);   // .z/.w: tap index of box column / row 0
      }
      const bool all_regular = __all_sync(0xffffffffu, regular);
      __syncwarp();
      if (all_regular) {
        const unsigned slot = use & 1u;
        ++use;
        if (lane == 0) bulk_wait_read<1>();   // the boxes sent two queries ago have been read out of this slot
        __syncwarp();
        float* sl = slots + slot * (RAFTCORR_MAX_LEVELS * kBoxFloats);
#pragma unroll
        for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l) {
          const Tap* tx = tab + (l * 2 + 0) * kMaxRd;
          const Tap* ty = tab + (l * 2 + 1) * kMaxRd;
          const int4 wv = win[l];
          const int cxr = col + wv.z;                     // column relative to the first tap's pixel
          // column cxr receives tap cxr with weight (1 - frac) and tap cxr-1 with weight frac
          // the TMA unit clips whole 16-byte chunks: a chunk straddling column W would spill into the row
          // padding, so columns at or beyond W carry zeros (their gradient is dropped: zero padding)
          const bool col_in = wv.x + col < p.glv[l].W;
          const bool has_a = col_in & (cxr >= 0) & (cxr < rd), has_b = col_in & (cxr >= 1) & (cxr <= rd);
          const int ia = min(max(cxr, 0), rd - 1), ib = min(max(cxr - 1, 0), rd - 1);
          const float wa = has_a ? __fsub_rn(1.0f, tx[ia].frac) : 0.0f;
          const float wb = has_b ? tx[ib].frac : 0.0f;
          const float* ga = g_s + (l * rd2 + ia * rd) * kOutPitch + qi;
          const float* gb = g_s + (l * rd2 + ib * rd) * kOutPitch + qi;
          // t[rr] = horizontal pass of y tap (5*half + rr - 1); rows 5*half .. 5*half+4 need taps one lower too
          float t[6], wy[6];
#pragma unroll
          for (int rr = 0; rr < 6; ++rr) {
            const int j = 5 * half + rr - 1 + wv.w;
            const bool ok = (j >= 0) & (j < rd);
            const int jc = min(max(j, 0), rd - 1);
            t[rr] = ok ? fmaf(wa, ga[jc * kOutPitch], wb * gb[jc * kOutPitch]) : 0.0f;
            wy[rr] = ty[jc].frac;
          }
#pragma unroll
          for (int rr = 0; rr < 5; ++rr) {
            // box row 5*half + rr = tap row ry: tap ry contributes (1 - frac), tap ry-1 contributes frac
            const float f = fmaf(t[rr + 1], __fsub_rn(1.0f, wy[rr + 1]), t[rr] * wy[rr]);
            sl[l * kBoxFloats + (5 * half + rr) * kBoxW + col] = f;
          }
        }
        fence_async_smem();
        __syncwarp();
        if (lane == 0) {
#pragma unroll
          for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l) {
            tma_reduce_add_3d(&maps.m[l], smem_u32(sl + l * kBoxFloats), win[l].x, win[l].y, q);
          }
          bulk_commit();
        }
      } else {
        // irregular window: the same terms through atomics (lookup.cu's scatter)
        for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l) {
          const Tap* tx = tab + (l * 2 + 0) * kMaxRd;
          const Tap* ty = tab + (l * 2 + 1) * kMaxRd;
          const LevelDev& gl = p.glv[l];
          float* img = const_cast<float*>(gl.data) + (long long)q * gl.istride;
          for (int kk = lane; kk < rd2; kk += 32) {
            const int i = kk / rd, j = kk - i * rd;
            const Tap a = tx[i], c = ty[j];
            const float g = g_s[(l * rd2 + kk) * kOutPitch + qi];
            const float w = a.frac, e = 1.0f - w, n = c.frac, s = 1.0f - n;
            const bool x0ok = (a.i0 >= 0) & (a.i0 < gl.W), x1ok = (a.i0 + 1 >= 0) & (a.i0 + 1 < gl.W);
            const bool y0ok = (c.i0 >= 0) & (c.i0 < gl.H), y1ok = (c.i0 + 1 >= 0) & (c.i0 + 1 < gl.H);
            float* row0 = img + (long long)c.i0 * gl.pitch + a.i0;
            float* row1 = row0 + gl.pitch;
            if (y0ok && x0ok) atomicAdd(row0, g * (s * e));
            if (y0ok && x1ok) atomicAdd(row0 + 1, g * (s * w));
            if (y1ok && x0ok) atomicAdd(row1, g * (n * e));
            if (y1ok && x1ok) atomicAdd(row1 + 1, g * (n * w));
          }
        }
      }
      __syncwarp();
    }
  }
  if (lane == 0) bulk_wait_all();   // shared memory must outlive the last reduce; results complete before exit
}

template <int R, int kMode>
int launch_k(const BwdTmaParams& p, const BwdMaps& maps, cudaStream_t stream) {
  using C = BCfg<R>;
  int dev = 0;
  RC_CHECK_CUDA(cudaGetDevice(&dev));
  static std::atomic<int> sm_count[64];
  int sms = (dev >= 0 && dev < 64) ? sm_count[dev].load(std::memory_order_acquire) : 0;
  if (sms == 0) {
    RC_CHECK_CUDA(cudaFuncSetAttribute(lookup_bwd_tma_kernel<R, kMode>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmem));
    RC_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (dev >= 0 && dev < 64) sm_count[dev].store(sms, std::memory_order_release);
  }
  const int grid = (int)(p.tiles < sms ? p.tiles : sms);
  lookup_bwd_tma_kernel<R, kMode><<<grid, kBThreads, C::kSmem, stream>>>(p, maps);
  RC_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

// returns 1 when this path does not apply (caller uses lookup.cu's atomic scatter)
int launch_lookup_bwd_tma(const raftcorr_level* glevels, int L, const float* coords, const float* gout, int B, int H,
                          int W, int r, int mode, cudaStream_t stream) {
  if (L != RAFTCORR_MAX_LEVELS || (r != 3 && r != 4) || get_encode() == nullptr) return 1;
  const long long BN = (long long)B * H * W;
  if (BN >= (1LL << 31)) return 1;
  for (int l = 0; l < L; ++l) {
    if ((glevels[l].pitch & 3) || (glevels[l].image_stride & 3)) return 1;
    if (reinterpret_cast<uintptr_t>(glevels[l].data) & 15) return 1;
  }
  BwdTmaParams p;
  BwdMaps maps;
  for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l) {
    p.glv[l] = make_level_dev(glevels[l].data, glevels[l].H, glevels[l].W, glevels[l].pitch, glevels[l].image_stride, l);
    cuuint64_t dims[3] = {(cuuint64_t)glevels[l].W, (cuuint64_t)glevels[l].H, (cuuint64_t)BN};
    cuuint64_t strides[2] = {(cuuint64_t)glevels[l].pitch * 4, (cuuint64_t)glevels[l].image_stride * 4};
    cuuint32_t box[3] = {(cuuint32_t)kBoxW, (cuuint32_t)kBoxH, 1};
    if (int rc = encode(&maps.m[l], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, glevels[l].data, dims, strides, box,
                        CU_TENSOR_MAP_SWIZZLE_NONE, "lookup_bwd window", CU_TENSOR_MAP_L2_PROMOTION_NONE))
      return rc;
  }
  p.coords = coords;
  p.gout = gout;
  p.B = B;
  p.H = H;
  p.W = W;
  p.tiles_x = (W + kTileQ - 1) / kTileQ;
  p.tiles = (long long)B * H * p.tiles_x;
  if (p.tiles == 0) return 0;
  if (p.tiles >= (1LL << 32)) return 1;
  switch (mode * 8 + r) {
    case RAFTCORR_COORD_ATEN_CPU * 8 + 4: return launch_k<4, RAFTCORR_COORD_ATEN_CPU>(p, maps, stream);
    case RAFTCORR_COORD_ATEN_CPU * 8 + 3: return launch_k<3, RAFTCORR_COORD_ATEN_CPU>(p, maps, stream);
    case RAFTCORR_COORD_ATEN_CUDA * 8 + 4: return launch_k<4, RAFTCORR_COORD_ATEN_CUDA>(p, maps, stream);
    case RAFTCORR_COORD_ATEN_CUDA * 8 + 3: return launch_k<3, RAFTCORR_COORD_ATEN_CUDA>(p, maps, stream);
    case RAFTCORR_COORD_DIRECT * 8 + 4: return launch_k<4, RAFTCORR_COORD_DIRECT>(p, maps, stream);
    default: return launch_k<3, RAFTCORR_COORD_DIRECT>(p, maps, stream);
  }
}

}  // namespace raftcorr
