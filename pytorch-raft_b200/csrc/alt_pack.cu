// alt_pack.cu — the operands of the on-the-fly path in one pass, sm_100a.
//
// core/corr.py:64-72 pools both fmaps with avg_pool2d four times and core/corr.py:82-83 re-permutes fmap1 and the
// pooled fmap2 to NHWC on every call (8 copies per refinement iteration).  alt_tc.cu wants bfloat16 NHWC copies of
// fmap1 and of the pooled fmap2 pyramid; as ATen ops that is 18 launches and ~1.5 GB of traffic at BASELINE.json
// configs[3] (0.71 ms measured).  Here one launch per fmap reads the fp32 NCHW tensor once and writes every level:
//
//   CTA  = one image, a tile of 8 rows x 32 columns, a slice of 64 channels
//   load = 128-byte row segments per channel (coalesced along W) into shared memory, fp32
//   pool = 2x2 means in fp32 in ATen's order ((a+b)+c)+d, * 0.25, level by level (4x16, 2x8, 1x4 pixels); floor on odd
//          sizes falls out of storing only pixels below (H >> l, W >> l)
//   store= per pixel the slice's 64 channels as 128 contiguous bytes of bfloat16 (lane = channel pair)
//
// HBM-bound: 4 bytes read + 2 * (1 + 1/4 + 1/16 + 1/64) bytes written per element.
#include <cuda_bf16.h>

#include <atomic>

#include "common.cuh"
#include "tma_util.cuh"

namespace raftcorr {

namespace {

constexpr int kPTileY = 8, kPTileX = 32, kPCh = 64;
constexpr int kPPitch = kPTileY * kPTileX + 1;                    // floats per channel (odd: neighbouring channels on neighbouring banks)
constexpr int kPSmem0 = kPCh * kPPitch;                           // level 0
constexpr int kPSmem1 = kPCh * (4 * 16 + 1), kPSmem2 = kPCh * (2 * 8 + 1), kPSmem3 = kPCh * (1 * 4 + 1);
constexpr int kPSmemBytes = (kPSmem0 + kPSmem1 + kPSmem2 + kPSmem3) * 4;

struct AltPackParams {
  const float* in;
  __nv_bfloat16* out[RAFTCORR_MAX_LEVELS];
  int B, D, H, W, L;
};

// pixels of level l covered by the tile (TW x TH of them), stored for channels c0 .. c0+63: lane = channel pair, one
// pixel per warp and step; pixel p = warp + 8 i sits at row i / (TW / 8), column warp + 8 (i % (TW / 8)) of the tile
template <int TW, int TH>
__device__ __forceinline__ void store_level(const float* src, int pitch, __nv_bfloat16* out, int b, int Hl, int Wl, int D,
                                            int y0, int x0, int c0, int warp, int lane) {
  const int c = c0 + 2 * lane;
  if (c >= D || y0 >= Hl || x0 >= Wl) return;
  __nv_bfloat16* base = out + (((long long)b * Hl + y0) * Wl + x0) * D + c;
  const float* s0 = src + (2 * lane) * pitch;
  constexpr int kPer = TW >= 8 ? TW / 8 : 1;             // steps per tile row
  constexpr int kSteps = (TW * TH + 7) / 8;
#pragma unroll
  for (int i = 0; i < kSteps; ++i) {
    const int py = i / kPer, px = warp + 8 * (i % kPer);
    if (TW < 8 && warp >= TW) continue;
    if (y0 + py >= Hl || x0 + px >= Wl) continue;
    const int q = py * TW + px;
    *reinterpret_cast<uint32_t*>(base + ((long long)py * Wl + px) * D) = pack_bf16x2(s0[q], s0[pitch + q]);
  }
}

template <int SW, int DW, int DH>
__device__ __forceinline__ void pool_level(const float* src, int spitch, float* dst, int dpitch, int tid) {
  constexpr int kPix = DW * DH;
#pragma unroll
  for (int k0 = 0; k0 < kPCh * kPix; k0 += 256) {
    const int k = k0 + tid, c = k / kPix, q = k % kPix;
    const int py = q / DW, px = q % DW;
    const float* s = src + c * spitch + (2 * py) * SW + 2 * px;
    dst[c * dpitch + q] = (((s[0] + s[1]) + s[SW]) + s[SW + 1]) * 0.25f;   // ATen avg_pool2d: row-major sum, then / 4
  }
}

__global__ void __launch_bounds__(256) alt_pack_kernel(const AltPackParams p) {
  extern __shared__ float sm[];
  float* t0 = sm;
  float* t1 = t0 + kPSmem0;
  float* t2 = t1 + kPSmem1;
  float* t3 = t2 + kPSmem2;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int slices = (p.D + kPCh - 1) / kPCh;
  const int b = blockIdx.z / slices, c0 = (blockIdx.z - b * slices) * kPCh;
  const int y0 = blockIdx.y * kPTileY, x0 = blockIdx.x * kPTileX;
  // warp w loads row w of the tile for every channel of the slice (lane = column); all 64 row segments are requested
  // before the first is used
  {
    const bool ok = (y0 + warp < p.H) & (x0 + lane < p.W);
    const int nch = min(kPCh, p.D - c0);
    const long long cstride = (long long)p.H * p.W;
    const float* src = p.in + (((long long)b * p.D + c0) * p.H + y0 + warp) * p.W + x0 + lane;
    float v[kPCh];
#pragma unroll
    for (int k = 0; k < kPCh; ++k) v[k] = (ok && k < nch) ? __ldcs(src + k * cstride) : 0.0f;
    float* d = t0 + warp * kPTileX + lane;
#pragma unroll
    for (int k = 0; k < kPCh; ++k) d[k * kPPitch] = v[k];
  }
  __syncthreads();
  store_level<32, 8>(t0, kPPitch, p.out[0], b, p.H, p.W, p.D, y0, x0, c0, warp, lane);
  if (p.L > 1) {
    pool_level<32, 16, 4>(t0, kPPitch, t1, 4 * 16 + 1, tid);
    __syncthreads();
    store_level<16, 4>(t1, 4 * 16 + 1, p.out[1], b, p.H >> 1, p.W >> 1, p.D, y0 >> 1, x0 >> 1, c0, warp, lane);
  }
  if (p.L > 2) {
    pool_level<16, 8, 2>(t1, 4 * 16 + 1, t2, 2 * 8 + 1, tid);
    __syncthreads();
    store_level<8, 2>(t2, 2 * 8 + 1, p.out[2], b, p.H >> 2, p.W >> 2, p.D, y0 >> 2, x0 >> 2, c0, warp, lane);
  }
  if (p.L > 3) {
    pool_level<8, 4, 1>(t2, 2 * 8 + 1, t3, 1 * 4 + 1, tid);
    __syncthreads();
    store_level<4, 1>(t3, 1 * 4 + 1, p.out[3], b, p.H >> 3, p.W >> 3, p.D, y0 >> 3, x0 >> 3, c0, warp, lane);
  }
}

}  // namespace

int launch_alt_pack(const float* fmap, int B, int D, int H, int W, int L, void* const* out_levels, cudaStream_t stream) {
  if (B == 0) return 0;
  if (D % 2 != 0) return fail(RAFTCORR_E_UNSUPPORTED, "alt_pack: needs an even channel count (D=%d)", D);
  AltPackParams p;
  p.in = fmap;
  for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l) p.out[l] = l < L ? static_cast<__nv_bfloat16*>(out_levels[l]) : nullptr;
  p.B = B;
  p.D = D;
  p.H = H;
  p.W = W;
  p.L = L;
  const long long gz = (long long)B * ((D + kPCh - 1) / kPCh);
  if (gz > 65535) return fail(RAFTCORR_E_UNSUPPORTED, "alt_pack: B * ceil(D/64) = %lld exceeds the grid limit", gz);
  static std::atomic<unsigned long long> configured{0};   // bit per device: shared-memory attribute set
  int dev = 0;
  RC_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !((configured.load(std::memory_order_acquire) >> dev) & 1ull)) {
    RC_CHECK_CUDA(cudaFuncSetAttribute(alt_pack_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kPSmemBytes));
    if (dev >= 0 && dev < 64) configured.fetch_or(1ull << dev, std::memory_order_release);
  }
  dim3 grid((W + kPTileX - 1) / kPTileX, (H + kPTileY - 1) / kPTileY, (unsigned)gz);
  alt_pack_kernel<<<grid, 256, kPSmemBytes, stream>>>(p);
  RC_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace raftcorr
