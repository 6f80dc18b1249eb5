// conv1x1_tc.cu — the lookup's consumer as one tensor-core GEMM.
//
// BasicMotionEncoder.convc1 + bias + ReLU (reference core/update.py:83,90) over a channels-last lookup:
//
//     out[b][o][pix] = act( sum_c W[o][c] * S[b][pix][c] + bias[o] )        o < 256, c = 324 (or 196)
//
// Output channels are the M operand so that the accumulator rows are NCHW rows: a tile is ALL output
// channels (two M128 halves, each its own TMEM accumulator) x 128 pixels, so the samples are read once
// and 8 x 55 tiles at cfg2 fill 148 SMs in 2.97 waves.  K-major fp32 operands, kind::tf32, fp32 accumulate;
// bias + ReLU in the TMEM epilogue; TMA stores straight into NCHW.
//
// The samples may also come as the plain NCHW lookup (S[b][c][pix], what CorrBlock returns by default and what the
// 16-bit lookup writes): the pixel axis is then contiguous, i.e. the B operand is MN-major — tcgen05 takes that for
// tf32 (instruction-descriptor bit 16) in one layout only, the 128-byte swizzle with 32-byte atoms (TMA
// SWIZZLE_128B_ATOM_32B, descriptor layout type 1; atom = 32 pixels x 4 channels): four TMA boxes of {32 pixels,
// 32 channels} per stage, leading byte offset 4096 (next 32 pixels), stride byte offset 512 (next 4 channels), and
// an MMA step (K = 8) advances the descriptor by 1024 bytes.
//
//   warp 0      TMA producer (3 stages of {W half 0, W half 1, 128 pixels} x 32 channels)
//   warp 1      MMA issuer   (tcgen05.mma cta_group::1 kind::tf32, M128 N128, 2 x 2 accumulators)
//   warps 2-9   epilogue     (warps 2-5 the first 128 output channels, 6-9 the second)
#include <atomic>

#include "tma_util.cuh"

namespace raftcorr {

namespace {

constexpr int kCM = 128;                      // rows per MMA (output channels per half)
constexpr int kCN = 128;                      // pixels per tile
constexpr int kCStages = 3;
constexpr int kCBytesA = kCM * 128;           // one half of W, 32 channels deep
constexpr int kCBytesB = kCN * 128;
constexpr int kCStageBytes = 2 * kCBytesA + kCBytesB;   // 48 KB
constexpr int kCEpiWarps = 8;
constexpr int kCEpiBuf = 4096;                // 32 rows x 128 B
constexpr int kCOffEpi = kCStages * kCStageBytes;
constexpr int kCOffBars = kCOffEpi + kCEpiWarps * 2 * kCEpiBuf;
constexpr int kCSmem = kCOffBars + 128;
constexpr int kCThreads = 32 * (2 + kCEpiWarps);

struct ConvParams {
  int B, M, Nn, k_blocks, halves, tiles_n;
  int nchw;        // 1: activations are (B, K, pixels), the B operand is MN-major
  long long tiles;
  const float* bias;
  int relu;
  uint32_t idesc;
};

__global__ void __launch_bounds__(kCThreads, 1)
conv1x1_tc_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_s,
                  const __grid_constant__ CUtensorMap map_o, const __grid_constant__ ConvParams p) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t bars = smem_base + kCOffBars;
  auto full_bar = [&](int s) { return bars + 8 * s; };
  auto empty_bar = [&](int s) { return bars + 8 * (kCStages + s); };
  auto tfull_bar = [&](int a) { return bars + 8 * (2 * kCStages + a); };
  auto tempty_bar = [&](int a) { return bars + 8 * (2 * kCStages + 2 + a); };
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kCOffBars + 8 * (2 * kCStages + 4));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < kCStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 4 * p.halves);
    }
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

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (long long t = blockIdx.x; t < p.tiles; t += gridDim.x) {
        const int nt = (int)(t % p.tiles_n), b = (int)(t / p.tiles_n);
        for (int kb = 0; kb < p.k_blocks; ++kb, ++it) {
          const int s = it % kCStages;
          mbar_wait(empty_bar(s), ((it / kCStages) & 1) ^ 1);
          mbar_expect_tx(full_bar(s), (uint32_t)(p.halves * kCBytesA + kCBytesB));
          const uint32_t dst = smem_base + s * kCStageBytes;
          for (int h = 0; h < p.halves; ++h) tma_load_3d(dst + h * kCBytesA, &map_w, full_bar(s), kb * 32, h * kCM, 0);
          if (p.nchw) {
            for (int j = 0; j < kCN / 32; ++j)
              tma_load_3d(dst + 2 * kCBytesA + j * 4096, &map_s, full_bar(s), nt * kCN + j * 32, kb * 32, b);
          } else {
            tma_load_3d(dst + 2 * kCBytesA, &map_s, full_bar(s), kb * 32, nt * kCN, b);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      uint32_t it = 0, i = 0;
      for (long long t = blockIdx.x; t < p.tiles; t += gridDim.x, ++i) {
        const int as = i & 1;
        mbar_wait(tempty_bar(as), ((i >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * (2 * kCN);
        for (int kb = 0; kb < p.k_blocks; ++kb, ++it) {
          const int s = it % kCStages;
          mbar_wait(full_bar(s), (it / kCStages) & 1);
          tc_fence_after();
          const uint32_t base = smem_base + s * kCStageBytes;
          // K-major: 32 bytes of K per step (+2 in the 16-byte address field); MN-major: one 1024-byte atom per step
          const uint64_t bdesc = p.nchw ? make_smem_desc_mn32(base + 2 * kCBytesA, 4096, 512) : make_smem_desc(base + 2 * kCBytesA);
          const uint64_t bstep = p.nchw ? (1024 >> 4) : 2;
          for (int h = 0; h < p.halves; ++h) {
            const uint64_t adesc = make_smem_desc(base + h * kCBytesA);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              tc_mma<true>(d_tmem + h * kCN, adesc + 2 * k, bdesc + bstep * k, p.idesc, (kb | k) != 0);
          }
          tc_commit(empty_bar(s));
        }
        tc_commit(tfull_bar(as));
      }
    }
  } else {
    const int ew = warp - 2;
    const int half = ew >> 2;         // which 128 output channels
    const int quarter = warp & 3;     // TMEM lanes 32*quarter .. +31 are visible to this warp
    if (half < p.halves) {
      const uint32_t buf_base = smem_base + kCOffEpi + ew * 2 * kCEpiBuf;
      const int row0 = half * kCM + quarter * 32;          // first output channel of this warp
      const bool rows_live = row0 < p.M;
      const float bias = (p.bias && row0 + lane < p.M) ? __ldg(p.bias + row0 + lane) : 0.0f;
      uint32_t nstores = 0, i = 0;
      for (long long t = blockIdx.x; t < p.tiles; t += gridDim.x, ++i) {
        const int nt = (int)(t % p.tiles_n), b = (int)(t / p.tiles_n);
        const int as = i & 1;
        mbar_wait(tfull_bar(as), (i >> 1) & 1);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + as * (2 * kCN) + half * kCN;
#pragma unroll 1
        for (int cc = 0; cc < kCN / 32; ++cc) {
          float f[32];
          tmem_ld_x32(taddr + cc * 32, f);
          tmem_ld_wait();
          if (cc == kCN / 32 - 1) {   // accumulator fully read: hand the TMEM stage back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(as));
          }
          const int n0 = nt * kCN + cc * 32;
          if (!rows_live || n0 >= p.Nn) continue;
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            f[c] += bias;
            if (p.relu) f[c] = f[c] < 0.0f ? 0.0f : f[c];   // NaN stays NaN, as in torch.relu
          }
          // 32 channels (lanes) x 32 pixels -> 128-byte-swizzled staging buffer -> TMA store (clipped at the edges)
          const uint32_t buf = buf_base + (nstores & 1) * kCEpiBuf;
          ++nstores;
          if (lane == 0) bulk_wait_read<1>();
          __syncwarp();
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const uint32_t addr = buf + lane * 128 + ((c ^ (lane & 7)) * 16);
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(f[4 * c]), "f"(f[4 * c + 1]),
                         "f"(f[4 * c + 2]), "f"(f[4 * c + 3])
                         : "memory");
          }
          fence_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_4d(&map_o, buf, n0, 0, row0, b);
            bulk_commit();
          }
        }
      }
      if (lane == 0) bulk_wait_all();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
  }
}

}  // namespace

// returns 1 when this kernel does not apply (caller uses the generic plain-GEMM mode of build_tc.cu)
int launch_conv1x1_tc(const float* weight, const float* act, float* out, int B, int M, int Nn, int K,
                      const float* bias, int relu, cudaStream_t stream, int nchw) {
  if (M > 2 * kCM || (K & 3) || (Nn & 3) || get_encode() == nullptr) return 1;
  if (B == 0) return 0;
  CUtensorMap map_w, map_s, map_o;
  {
    cuuint64_t dims[3] = {(cuuint64_t)K, (cuuint64_t)M, 1};
    cuuint64_t strides[2] = {(cuuint64_t)K * 4, (cuuint64_t)M * K * 4};
    cuuint32_t box[3] = {32, kCM, 1};
    if (int rc = encode(&map_w, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(weight), dims, strides, box,
                        CU_TENSOR_MAP_SWIZZLE_128B, "conv1x1 weight"))
      return rc;
  }
  if (nchw) {
    cuuint64_t dims[3] = {(cuuint64_t)Nn, (cuuint64_t)K, (cuuint64_t)B};
    cuuint64_t strides[2] = {(cuuint64_t)Nn * 4, (cuuint64_t)Nn * K * 4};
    cuuint32_t box[3] = {32, 32, 1};
    if (int rc = encode(&map_s, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(act), dims, strides, box,
                        CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, "conv1x1 activations (NCHW)"))
      return rc;
  } else {
    cuuint64_t dims[3] = {(cuuint64_t)K, (cuuint64_t)Nn, (cuuint64_t)B};
    cuuint64_t strides[2] = {(cuuint64_t)K * 4, (cuuint64_t)Nn * K * 4};
    cuuint32_t box[3] = {32, kCN, 1};
    if (int rc = encode(&map_s, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(act), dims, strides, box,
                        CU_TENSOR_MAP_SWIZZLE_128B, "conv1x1 activations"))
      return rc;
  }
  {
    cuuint64_t dims[4] = {(cuuint64_t)Nn, 1, (cuuint64_t)M, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)Nn * 4, (cuuint64_t)Nn * 4, (cuuint64_t)M * Nn * 4};
    cuuint32_t box[4] = {32, 1, 32, 1};
    if (int rc = encode(&map_o, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, out, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B,
                        "conv1x1 out"))
      return rc;
  }
  ConvParams p;
  p.B = B;
  p.M = M;
  p.Nn = Nn;
  p.k_blocks = (K + 31) / 32;
  p.halves = (M + kCM - 1) / kCM;
  p.tiles_n = (Nn + kCN - 1) / kCN;
  p.tiles = (long long)B * p.tiles_n;
  p.bias = bias;
  p.relu = relu;
  p.nchw = nchw;
  p.idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kCN >> 3) << 17) | ((uint32_t)(kCM >> 4) << 24) |
            (nchw ? 1u << 16 : 0u);   // bit 16: B is MN-major
  int dev = 0;
  RC_CHECK_CUDA(cudaGetDevice(&dev));
  static std::atomic<int> sm_count[64];
  int sms = (dev >= 0 && dev < 64) ? sm_count[dev].load(std::memory_order_acquire) : 0;
  if (sms == 0) {
    RC_CHECK_CUDA(cudaFuncSetAttribute(conv1x1_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kCSmem));
    RC_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (dev >= 0 && dev < 64) sm_count[dev].store(sms, std::memory_order_release);
  }
  const int grid = (int)(p.tiles < sms ? p.tiles : sms);
  conv1x1_tc_kernel<<<grid, kCThreads, kCSmem, stream>>>(map_w, map_s, map_o, p);
  RC_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace raftcorr
