// build_tc.cu — all-pairs volume + 4-level pyramid on the 5th-gen tensor cores (kernel family 1).
//
// Replaces CorrBlock.corr + the avg_pool2d chain (reference core/corr.py:21-27, 52-60):
//   C[b, i, j] = <fmap1[b,:,i], fmap2[b,:,j]> / sqrt(D), then three 2x2 means.
// The reference materialises C with cuBLAS, re-reads it for the division and re-reads it again
// for every pooling level.  Here a persistent, warp-specialised tcgen05 GEMM keeps the 128 x 256
// accumulator tile in TMEM; the N side of a tile is a 2-D block of 8 x 32 target pixels, so one
// TMEM lane (= one query) holds a whole 8x32 patch and the epilogue scales it, forms the 4x16,
// 2x8 and 1x4 pooled patches in registers and writes all four levels once, through swizzled
// shared-memory staging and TMA tensor stores (which also clip ragged edges).
//
//   warp 0      TMA producer   (A: 128 queries x 128 B of K;  B: 8x32 pixels x 128 B of K)
//   warp 1      MMA issuer     (tcgen05.mma cta_group::1, kind::f16 / kind::tf32, M128 N256)
//   warps 2-5   epilogue group 0 (accumulator stage 0)     warps 6-9   epilogue group 1
//
// Operands come from pack_kernel: NCHW fp32 -> (B, N, D) K-major bf16 (or tf32-rounded fp32).
#include <cuda.h>
#include <cuda_bf16.h>

#include <mutex>

#include "common.cuh"
#include "tma_util.cuh"

namespace raftcorr {

namespace {

constexpr int kBlockM = 128;             // queries per tile  (TMEM lanes)
constexpr int kTileY = 8, kTileX = 32;   // target-pixel block per tile
constexpr int kBlockN = kTileY * kTileX; // 256 TMEM columns
constexpr int kStages = 3;
constexpr int kBytesA = kBlockM * 128;   // 16 KB per k-block (128 B of K per row)
constexpr int kBytesB = kBlockN * 128;   // 32 KB
constexpr int kStageBytes = kBytesA + kBytesB;
constexpr int kEpiWarps = 8;
constexpr int kEpiBufBytes = 4096;       // one staged row: 32 queries x 128 B
constexpr int kSmemStaging = kStages * kStageBytes;
constexpr int kSmemBars = kSmemStaging + kEpiWarps * 2 * kEpiBufBytes;
constexpr int kSmemTotal = kSmemBars + 128;
constexpr int kThreads = 32 * (2 + kEpiWarps);

struct BuildTcParams {
  int B, N, H, W, L;
  int tiles_m, tiles_y, tiles_x;
  int k_blocks;       // ceil(K * elem_size / 128) per pass
  int passes;         // 1, or 3 for the split-TF32 product (hi*hi + hi*lo + lo*hi)
  int plain;          // 0: pyramid epilogue (8x32 pixel tiles); 1: plain C[M][N] epilogue (1x256 tiles)
  int k_elems;        // K elements per 128 B
  uint32_t idesc;
  float inv_scale;    // 1 / sqrt(D)
  int Hl[RAFTCORR_MAX_LEVELS], Wl[RAFTCORR_MAX_LEVELS];
  // plain epilogue extras (1x1-convolution use): per-row bias, ReLU, A shared by all batches
  const float* bias;
  int relu;
  int a_batch_mul;    // 1: A has a batch dimension; 0: one A for every batch
  // 16-bit blocked pyramid (raftcorr_level16): block grid per level, and level 3 for plain stores
  int bxs[RAFTCORR_MAX_LEVELS], bys[RAFTCORR_MAX_LEVELS];
  unsigned char* l3_data;
  long long l3_istride;   // bytes
  int ilog[2];            // levels 0, 1: log2 of the number of neighbouring images whose blocks are interleaved (0: none)
};

// Stage one row segment (kNF floats per query, 32 queries) in swizzled shared memory and hand it
// to the TMA store engine.  slot alternates between the warp's two 4 KB buffers.
template <int kNF>
__device__ __forceinline__ void stage_and_store(const float* vals, const CUtensorMap* map, uint32_t buf_base,
                                                uint32_t& nstores, int lane, int X0, int Y, int q0, int b) {
  const uint32_t buf = buf_base + (nstores & 1) * kEpiBufBytes;
  ++nstores;
  if (lane == 0) bulk_wait_read<1>();  // the store issued two rows ago has drained this buffer
  __syncwarp();
  constexpr int kChunks = kNF / 4;
#pragma unroll
  for (int c = 0; c < kChunks; ++c) {
    int phys = c;
    if (kNF == 32) phys = c ^ (lane & 7);          // SWIZZLE_128B
    if (kNF == 16) phys = c ^ ((lane >> 1) & 3);   // SWIZZLE_64B
    if (kNF == 8) phys = c ^ ((lane >> 2) & 1);    // SWIZZLE_32B
    const uint32_t addr = buf + lane * (kNF * 4) + phys * 16;
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(vals[4 * c]),
                 "f"(vals[4 * c + 1]), "f"(vals[4 * c + 2]), "f"(vals[4 * c + 3])
                 : "memory");
  }
  fence_async_smem();
  __syncwarp();
  if (lane == 0) {
    tma_store_4d(map, buf, X0, Y, q0, b);
    bulk_commit();
  }
}

// Operand maps per pass: C += A[pass] . B[pass]^T.  One pass for bf16 / tf32; three passes with
// (A_hi,B_hi), (A_hi,B_lo), (A_lo,B_hi) give an fp32-accurate product on the tf32 tensor pipe.
struct alignas(64) TcOperandMaps {
  CUtensorMap a[3];
  CUtensorMap b[3];
};

// ---- epilogue of the 16-bit blocked pyramid (include/raftcorr.h, raftcorr_level16) ---------------------------
// A staged box is 32 queries x kRowBytes; lane = query.  Chunks of 16 bytes are placed with the TMA swizzle of
// the box's tensor map (128-byte rows: SWIZZLE_128B, 32-byte rows: SWIZZLE_32B) so a quarter warp fills whole
// 128-byte wavefronts.
__device__ __forceinline__ uint32_t stage16_begin(uint32_t buf_base, uint32_t& nstores, int lane) {
  const uint32_t buf = buf_base + (nstores & 1) * kEpiBufBytes;
  ++nstores;
  if (lane == 0) bulk_wait_read<1>();  // the store issued two boxes ago has drained this buffer
  __syncwarp();
  return buf;
}
// Interleaved levels (raftcorr_level16.interleave_log2 = g >= 1): memory is [group of 2^g images][block row][block]
// [image][64 B].  The blocks of two neighbouring images are one 128-byte row of the tensor map (64-byte rows halve
// the store engine's throughput), so a staged box is {128 B, 2^(g-1) image pairs, 2 blocks, 1 block row, 32 >> g
// groups} with SWIZZLE_128B (16-byte chunk index ^= 128-byte row index & 7; the staging buffers are 1 KB aligned).
// Chunk r of block bl of the lane's query:
__device__ __forceinline__ void stage16_chunk_il(uint32_t buf, int lane, int ilog, int bl, int r, uint32_t a, uint32_t b,
                                                 uint32_t cc, uint32_t d) {
  const int m = lane & ((1 << ilog) - 1);   // image within its group
  const uint32_t row = ((uint32_t)(lane >> ilog) << ilog) + ((uint32_t)bl << (ilog - 1)) + (uint32_t)(m >> 1);
  const uint32_t phys = (uint32_t)((m & 1) * 4 + r) ^ (row & 7u);
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(buf + row * 128 + phys * 16), "r"(a), "r"(b), "r"(cc), "r"(d)
               : "memory");
}
// one box for the warp's 32 queries or, in a tile that ends before them (N % 32 != 0), one per live group
__device__ __forceinline__ void store16_il(const CUtensorMap* map, const CUtensorMap* map_tail, uint32_t buf, int ilog,
                                           int bx, int by, long long image0, int live) {
  const int g0 = (int)(image0 >> ilog);
  if (live >= 32) {
    tma_store_5d(map, buf, 0, 0, bx, by, g0);
  } else {
    for (int g = 0; (g << ilog) < live; ++g) tma_store_5d(map_tail, buf + ((uint32_t)g << (7 + ilog)), 0, 0, bx, by, g0 + g);
  }
}
template <int kRowBytes>
__device__ __forceinline__ void stage16_chunk(uint32_t buf, int lane, int c, uint32_t a, uint32_t b, uint32_t cc,
                                              uint32_t d) {
  int phys = c;
  if (kRowBytes == 128) phys = c ^ (lane & 7);
  if (kRowBytes == 32) phys = c ^ ((lane >> 2) & 1);
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(buf + lane * kRowBytes + phys * 16), "r"(a), "r"(b),
               "r"(cc), "r"(d)
               : "memory");
}

// kOut16: 0 = fp32 rows, 1 = 16-bit blocked levels, 2 = the same with levels 0 and 1 interleaved by four images
template <bool kTf32, int kOut16>
__global__ void __launch_bounds__(kThreads, 1)
build_tc_kernel(const __grid_constant__ TcOperandMaps ops,
                const __grid_constant__ CUtensorMap mapL0, const __grid_constant__ CUtensorMap mapL1,
                const __grid_constant__ CUtensorMap mapL2, const __grid_constant__ CUtensorMap mapL3,
                const __grid_constant__ CUtensorMap mapL0t, const __grid_constant__ CUtensorMap mapL1t,
                const __grid_constant__ BuildTcParams p) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t bars = smem_base + kSmemBars;
  // barrier slots (8 B each): full[3] empty[3] tmem_full[2] tmem_empty[2]; then the TMEM address
  auto full_bar = [&](int s) { return bars + 8 * s; };
  auto empty_bar = [&](int s) { return bars + 8 * (kStages + s); };
  auto tfull_bar = [&](int a) { return bars + 8 * (2 * kStages + a); };
  auto tempty_bar = [&](int a) { return bars + 8 * (2 * kStages + 2 + a); };
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kSmemBars + 8 * (2 * kStages + 4));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&ops.a[0]) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&ops.b[0]) : "memory");
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

  const long long tiles_n = (long long)p.tiles_y * p.tiles_x;
  const long long total = (long long)p.B * p.tiles_m * tiles_n;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (long long t = blockIdx.x; t < total; t += gridDim.x) {
        const int nt = (int)(t % tiles_n);
        const int mt = (int)((t / tiles_n) % p.tiles_m);
        const int b = (int)(t / (tiles_n * p.tiles_m));
        const int ty = nt / p.tiles_x, tx = nt - ty * p.tiles_x;
        const int bx = p.plain ? kBlockN : kTileX, by = p.plain ? 1 : kTileY;   // N-tile origin step
        for (int pass = 0; pass < p.passes; ++pass)
          for (int kb = 0; kb < p.k_blocks; ++kb, ++it) {
            const int s = it % kStages;
            const uint32_t ph = (it / kStages) & 1;
            mbar_wait(empty_bar(s), ph ^ 1);
            mbar_expect_tx(full_bar(s), kStageBytes);
            const uint32_t a_dst = smem_base + s * kStageBytes;
            tma_load_3d(a_dst, &ops.a[pass], full_bar(s), kb * p.k_elems, mt * kBlockM, b * p.a_batch_mul);
            tma_load_4d(a_dst + kBytesA, &ops.b[pass], full_bar(s), kb * p.k_elems, tx * bx, ty * by, b);
          }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      uint32_t it = 0, i = 0;
      for (long long t = blockIdx.x; t < total; t += gridDim.x, ++i) {
        const int as = i & 1;
        const uint32_t aph = (i >> 1) & 1;
        mbar_wait(tempty_bar(as), aph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * kBlockN;
        for (int kb = 0; kb < p.k_blocks * p.passes; ++kb, ++it) {
          const int s = it % kStages;
          const uint32_t ph = (it / kStages) & 1;
          mbar_wait(full_bar(s), ph);
          tc_fence_after();
          const uint32_t a_addr = smem_base + s * kStageBytes;
          const uint64_t adesc = make_smem_desc(a_addr);
          const uint64_t bdesc = make_smem_desc(a_addr + kBytesA);
#pragma unroll
          for (int k = 0; k < 4; ++k)  // 4 x 32 B of K per 128 B row; +2 in the 16-byte address field
            tc_mma<kTf32>(d_tmem, adesc + 2 * k, bdesc + 2 * k, p.idesc, (kb | k) != 0);
          tc_commit(empty_bar(s));
        }
        tc_commit(tfull_bar(as));
      }
    }
  } else {
    const int ew = warp - 2;          // 0..7
    const int group = ew >> 2;        // accumulator stage this warp drains
    const int quarter = warp & 3;     // TMEM lanes 32*quarter .. +31 are visible to this warp
    const uint32_t buf_base = smem_base + kSmemStaging + ew * 2 * kEpiBufBytes;
    uint32_t nstores = 0, i = 0;
    for (long long t = blockIdx.x; t < total; t += gridDim.x, ++i) {
      if ((int)(i & 1) != group) continue;
      const uint32_t aph = (i >> 1) & 1;
      const int nt = (int)(t % tiles_n);
      const int mt = (int)((t / tiles_n) % p.tiles_m);
      const int b = (int)(t / (tiles_n * p.tiles_m));
      const int ty = nt / p.tiles_x, tx = nt - ty * p.tiles_x;
      const int y0 = ty * kTileY, x0 = tx * kTileX;
      const int q0 = mt * kBlockM + quarter * 32;
      const bool rows_live = q0 < p.N;
      mbar_wait(tfull_bar(group), aph);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + group * kBlockN;
      if (p.plain) {
        // C[b][m][n]: this warp's 32 rows x 256 columns, eight 32x32 boxes through the staging buffers
#pragma unroll 1
        for (int cc = 0; cc < kBlockN / 32; ++cc) {
          float f[32];
          tmem_ld_x32(taddr + cc * 32, f);
          tmem_ld_wait();
          if (cc == kBlockN / 32 - 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(group));
          }
          if (!rows_live) continue;
          const float bias = p.bias ? __ldg(p.bias + min(q0 + lane, p.N - 1)) : 0.0f;
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            f[c] = fmaf(f[c], p.inv_scale, bias);
            if (p.relu) f[c] = f[c] < 0.0f ? 0.0f : f[c];   // NaN stays NaN, as in torch.relu
          }
          const int n0 = tx * kBlockN + cc * 32;
          if (n0 < p.W) stage_and_store<32>(f, &mapL0, buf_base, nstores, lane, n0, 0, q0, b);
        }
        continue;
      }
      if constexpr (kOut16 != 0) {
        constexpr int kIlog = kOut16 == 2 ? 2 : 0;
        // bf16 pixels in 4x8 blocks.  An 8x32 tile is 2 block rows x 4 blocks at level 0, one block row x 2
        // blocks at level 1, half a block at level 2 and a quarter of a block row at level 3.
        uint32_t prev[32], l1pk[32];
        float q1[2][16], q2[2][8];
        const bool edge = (x0 + kTileX > p.W) | (y0 + kTileY > p.H);   // pooled patches reach padding pixels
#pragma unroll
        for (int rp = 0; rp < 4; ++rp) {
          float f[64];
          tmem_ld_x64(taddr + rp * 64, f);
          tmem_ld_wait();
          if (rp == 3) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(group));
          }
          if (!rows_live) continue;
#pragma unroll
          for (int c = 0; c < 64; ++c) f[c] *= p.inv_scale;  // corr.py:60 (pixels beyond H x W are exact zeros:
                                                             // TMA zero-filled their fmap2 rows)
          uint32_t cur[32];
#pragma unroll
          for (int k = 0; k < 32; ++k) cur[k] = pack_bf16x2(f[2 * k], f[2 * k + 1]);
          // 2x2 means in fp32, ATen's order ((a+b)+c)+d then /4 (corr.py:26); padding pixels of the pooled levels
          // (x >= W_l or y >= H_l, reachable only from edge tiles) must be stored as zeros
#pragma unroll
          for (int k = 0; k < 16; ++k)
            q1[rp & 1][k] = (((f[2 * k] + f[2 * k + 1]) + f[32 + 2 * k]) + f[33 + 2 * k]) * 0.25f;
          if (edge) {
            const bool row_ok = (y0 >> 1) + rp < p.Hl[1];
#pragma unroll
            for (int k = 0; k < 16; ++k)
              if (!row_ok || (x0 >> 1) + k >= p.Wl[1]) q1[rp & 1][k] = 0.0f;
          }
#pragma unroll
          for (int k = 0; k < 8; ++k) l1pk[rp * 8 + k] = pack_bf16x2(q1[rp & 1][2 * k], q1[rp & 1][2 * k + 1]);
          if (rp & 1) {
            const int by = (y0 >> 2) + (rp >> 1);
            if (by < p.bys[0]) {
#pragma unroll
              for (int h = 0; h < 2; ++h) {
                const int bx = (x0 >> 3) + 2 * h;
                if (bx >= p.bxs[0]) continue;
                const uint32_t buf = stage16_begin(buf_base, nstores, lane);
#pragma unroll
                for (int bl = 0; bl < 2; ++bl)
#pragma unroll
                  for (int r = 0; r < 4; ++r) {
                    const uint32_t* src = (r < 2 ? prev : cur) + (r & 1) * 16 + (2 * h + bl) * 4;
                    if constexpr (kIlog) stage16_chunk_il(buf, lane, kIlog, bl, r, src[0], src[1], src[2], src[3]);
                    else stage16_chunk<128>(buf, lane, bl * 4 + r, src[0], src[1], src[2], src[3]);
                  }
                fence_async_smem();
                __syncwarp();
                if (lane == 0) {
                  if constexpr (kIlog) store16_il(&mapL0, &mapL0t, buf, kIlog, bx, by, (long long)b * p.N + q0, p.N - q0);
                  else tma_store_4d(&mapL0, buf, bx * 32, by, q0, b);
                  bulk_commit();
                }
              }
            }
#pragma unroll
            for (int k = 0; k < 8; ++k)
              q2[rp >> 1][k] = (((q1[0][2 * k] + q1[0][2 * k + 1]) + q1[1][2 * k]) + q1[1][2 * k + 1]) * 0.25f;
            if (edge) {
              const bool row_ok = (y0 >> 2) + (rp >> 1) < p.Hl[2];
#pragma unroll
              for (int k = 0; k < 8; ++k)
                if (!row_ok || (x0 >> 2) + k >= p.Wl[2]) q2[rp >> 1][k] = 0.0f;
            }
          } else {
#pragma unroll
            for (int k = 0; k < 32; ++k) prev[k] = cur[k];
          }
        }
        if (!rows_live) continue;
        if (p.L > 1 && ty < p.bys[1] && 2 * tx < p.bxs[1]) {     // level 1: block row ty, blocks 2tx, 2tx+1
          const uint32_t buf = stage16_begin(buf_base, nstores, lane);
#pragma unroll
          for (int bl = 0; bl < 2; ++bl)
#pragma unroll
            for (int r = 0; r < 4; ++r) {
              const uint32_t* src = l1pk + r * 8 + bl * 4;
              if constexpr (kIlog) stage16_chunk_il(buf, lane, kIlog, bl, r, src[0], src[1], src[2], src[3]);
              else stage16_chunk<128>(buf, lane, bl * 4 + r, src[0], src[1], src[2], src[3]);
            }
          fence_async_smem();
          __syncwarp();
          if (lane == 0) {
            if constexpr (kIlog) store16_il(&mapL1, &mapL1t, buf, kIlog, 2 * tx, ty, (long long)b * p.N + q0, p.N - q0);
            else tma_store_4d(&mapL1, buf, 2 * tx * 32, ty, q0, b);
            bulk_commit();
          }
        }
        if (p.L > 2 && (ty >> 1) < p.bys[2] && tx < p.bxs[2]) {  // level 2: rows 2(ty&1), +1 of block (ty>>1, tx)
          // the level is addressed in 32-byte half blocks: unit = 2 * block + half
          const int unit = 2 * ((ty >> 1) * p.bxs[2] + tx) + (ty & 1);
          const bool tail = (ty == p.tiles_y - 1) && !(ty & 1);  // nobody else writes the block's lower half
#pragma unroll
          for (int part = 0; part < 2; ++part) {
            if (part == 1 && !tail) break;
            const uint32_t buf = stage16_begin(buf_base, nstores, lane);
#pragma unroll
            for (int r = 0; r < 2; ++r) {
              uint32_t w[4];
#pragma unroll
              for (int k = 0; k < 4; ++k) w[k] = part ? 0u : pack_bf16x2(q2[r][2 * k], q2[r][2 * k + 1]);
              stage16_chunk<32>(buf, lane, r, w[0], w[1], w[2], w[3]);
            }
            fence_async_smem();
            __syncwarp();
            if (lane == 0) {
              tma_store_4d(&mapL2, buf, 0, unit + part, q0, b);
              bulk_commit();
            }
          }
        }
        if (p.L > 3 && (ty >> 2) < p.bys[3] && (tx >> 1) < p.bxs[3] && q0 + lane < p.N) {
          // level 3: row ty, columns 4tx..4tx+3 = 8 bytes per query, plain stores.  The last tile row / column
          // also zero what is left of the block (rows below, the block's other four columns).
          float q3[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            q3[k] = (((q2[0][2 * k] + q2[0][2 * k + 1]) + q2[1][2 * k]) + q2[1][2 * k + 1]) * 0.25f;
            if (ty >= p.Hl[3] || 4 * tx + k >= p.Wl[3]) q3[k] = 0.0f;
          }
          unsigned char* blk = p.l3_data + ((long long)b * p.N + q0 + lane) * p.l3_istride +
                               ((long long)(ty >> 2) * p.bxs[3] + (tx >> 1)) * 64;
          const uint2 v = make_uint2(pack_bf16x2(q3[0], q3[1]), pack_bf16x2(q3[2], q3[3]));
          const int colb = (tx & 1) * 8;
          const bool zero_cols = (tx == p.tiles_x - 1) && !(tx & 1);
          const int row_end = (ty == p.tiles_y - 1) ? 4 : (ty & 3) + 1;
          for (int r = ty & 3; r < row_end; ++r) {
            const uint2 w = r == (ty & 3) ? v : make_uint2(0u, 0u);
            *reinterpret_cast<uint2*>(blk + r * 16 + colb) = w;
            if (zero_cols) *reinterpret_cast<uint2*>(blk + r * 16 + 8) = make_uint2(0u, 0u);
          }
        }
        continue;
      }
      float p1[2][16];
      float p2[2][8];
#pragma unroll
      for (int rp = 0; rp < 4; ++rp) {
        float f[64];
        tmem_ld_x64(taddr + rp * 64, f);
        tmem_ld_wait();
        if (rp == 3) {  // accumulator fully read: hand the TMEM stage back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(tempty_bar(group));
        }
        if (!rows_live) continue;
#pragma unroll
        for (int c = 0; c < 64; ++c) f[c] *= p.inv_scale;  // corr.py:60
        if (y0 + 2 * rp < p.H) stage_and_store<32>(f, &mapL0, buf_base, nstores, lane, x0, y0 + 2 * rp, q0, b);
        if (y0 + 2 * rp + 1 < p.H)
          stage_and_store<32>(f + 32, &mapL0, buf_base, nstores, lane, x0, y0 + 2 * rp + 1, q0, b);
        // 2x2 means in ATen's order ((a+b)+c)+d, then /4 (corr.py:26)
#pragma unroll
        for (int k = 0; k < 16; ++k)
          p1[rp & 1][k] = (((f[2 * k] + f[2 * k + 1]) + f[32 + 2 * k]) + f[33 + 2 * k]) * 0.25f;
        if (p.L > 1 && (y0 >> 1) + rp < p.Hl[1] && (x0 >> 1) < p.Wl[1])
          stage_and_store<16>(p1[rp & 1], &mapL1, buf_base, nstores, lane, x0 >> 1, (y0 >> 1) + rp, q0, b);
        if (rp & 1) {
#pragma unroll
          for (int k = 0; k < 8; ++k)
            p2[rp >> 1][k] =
                (((p1[0][2 * k] + p1[0][2 * k + 1]) + p1[1][2 * k]) + p1[1][2 * k + 1]) * 0.25f;
          if (p.L > 2 && (y0 >> 2) + (rp >> 1) < p.Hl[2] && (x0 >> 2) < p.Wl[2])
            stage_and_store<8>(p2[rp >> 1], &mapL2, buf_base, nstores, lane, x0 >> 2, (y0 >> 2) + (rp >> 1), q0, b);
        }
      }
      if (rows_live && p.L > 3 && (y0 >> 3) < p.Hl[3] && (x0 >> 3) < p.Wl[3]) {
        float p3[4];
#pragma unroll
        for (int k = 0; k < 4; ++k)
          p3[k] = (((p2[0][2 * k] + p2[0][2 * k + 1]) + p2[1][2 * k]) + p2[1][2 * k + 1]) * 0.25f;
        stage_and_store<4>(p3, &mapL3, buf_base, nstores, lane, x0 >> 3, y0 >> 3, q0, b);
      }
    }
    if (lane == 0) bulk_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
  }
}

// NCHW fp32 (B, D, N) -> K-major (B, N, D): bf16, or fp32 rounded to tf32 (cvt.rna)
template <typename T>
__global__ void __launch_bounds__(256)
pack_kernel(const float* __restrict__ src, T* __restrict__ dst, int D, int N) {
  __shared__ float tile[64][33];
  const int b = blockIdx.z;
  const int n0 = blockIdx.x * 32, d0 = blockIdx.y * 64;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const float* s = src + (long long)b * D * N;
  for (int c = w; c < 64; c += 8) {
    const int d = d0 + c, n = n0 + lane;
    tile[c][lane] = (d < D && n < N) ? __ldg(s + (long long)d * N + n) : 0.0f;
  }
  __syncthreads();
  T* o = dst + (long long)b * N * D;
  for (int px = w; px < 32; px += 8) {
    const int n = n0 + px;
    const int d = d0 + 2 * lane;
    if (n >= N || d >= D) continue;  // D is even (checked by the launcher)
    const float v0 = tile[2 * lane][px], v1 = tile[2 * lane + 1][px];
    if constexpr (sizeof(T) == 2) {
      *reinterpret_cast<__nv_bfloat162*>(o + (long long)n * D + d) = __floats2bfloat162_rn(v0, v1);
    } else {
      uint32_t r0, r1;
      asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r0) : "f"(v0));
      asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r1) : "f"(v1));
      *reinterpret_cast<float2*>(o + (long long)n * D + d) = make_float2(__uint_as_float(r0), __uint_as_float(r1));
    }
  }
}

// Both feature maps in one launch, bf16: NCHW fp32 (B, D, N) -> K-major (B, N, D).  64 channels x 64
// pixels per block, 16-byte loads along pixels and 16-byte stores along channels.
__global__ void __launch_bounds__(256)
pack2_bf16_kernel(const float* __restrict__ src1, const float* __restrict__ src2, __nv_bfloat16* __restrict__ dst1,
                  __nv_bfloat16* __restrict__ dst2, int D, int N) {
  __shared__ float tile[64][65];
  const int b = blockIdx.z >> 1;
  const float* s = ((blockIdx.z & 1) ? src2 : src1) + (long long)b * D * N;
  __nv_bfloat16* o = ((blockIdx.z & 1) ? dst2 : dst1) + (long long)b * N * D;
  const int n0 = blockIdx.x * 64, d0 = blockIdx.y * 64;
  const int tid = threadIdx.x;
  if ((N & 3) == 0) {
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const int e = tid + 256 * it;          // 1024 float4 per tile
      const int c = e >> 4, p4 = (e & 15) * 4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (d0 + c < D && n0 + p4 < N) v = __ldg(reinterpret_cast<const float4*>(s + (long long)(d0 + c) * N + n0 + p4));
      tile[c][p4] = v.x;
      tile[c][p4 + 1] = v.y;
      tile[c][p4 + 2] = v.z;
      tile[c][p4 + 3] = v.w;
    }
  } else {
    for (int e = tid; e < 64 * 64; e += 256) {
      const int c = e >> 6, px = e & 63;
      tile[c][px] = (d0 + c < D && n0 + px < N) ? __ldg(s + (long long)(d0 + c) * N + n0 + px) : 0.0f;
    }
  }
  __syncthreads();
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    const int px = (tid >> 3) + 32 * it, q = tid & 7;   // 8 channels (16 B) per thread
    const int n = n0 + px, d = d0 + 8 * q;
    if (n < N && d < D) {                              // D % 8 == 0 (checked by the launcher)
      __nv_bfloat162 h[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) h[k] = __floats2bfloat162_rn(tile[8 * q + 2 * k][px], tile[8 * q + 2 * k + 1][px]);
      *reinterpret_cast<uint4*>(o + (long long)n * D + d) = *reinterpret_cast<uint4*>(h);
    }
  }
}

// NCHW fp32 (B, D, N) -> two K-major (B, N, D) fp32 arrays holding tf32 values: hi = rna(x),
// lo = rna(x - hi); hi + lo reproduces x to ~2^-24, so hi*hi + hi*lo + lo*hi is an fp32-accurate
// product on the tf32 tensor pipe.
__global__ void __launch_bounds__(256)
pack_split_kernel(const float* __restrict__ src, float* __restrict__ hi, float* __restrict__ lo, int D, int N) {
  __shared__ float tile[64][33];
  const int b = blockIdx.z;
  const int n0 = blockIdx.x * 32, d0 = blockIdx.y * 64;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const float* s = src + (long long)b * D * N;
  for (int c = w; c < 64; c += 8) {
    const int d = d0 + c, n = n0 + lane;
    tile[c][lane] = (d < D && n < N) ? __ldg(s + (long long)d * N + n) : 0.0f;
  }
  __syncthreads();
  for (int px = w; px < 32; px += 8) {
    const int n = n0 + px;
    const int d = d0 + 2 * lane;
    if (n >= N || d >= D) continue;
    float h[2], l[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const float v = tile[2 * lane + k][px];
      uint32_t r;
      asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
      h[k] = __uint_as_float(r);
      asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v - h[k]));
      l[k] = __uint_as_float(r);
    }
    const long long o = (long long)b * N * D + (long long)n * D + d;
    *reinterpret_cast<float2*>(hi + o) = make_float2(h[0], h[1]);
    *reinterpret_cast<float2*>(lo + o) = make_float2(l[0], l[1]);
  }
}

size_t elem_size(int precision) { return precision == RAFTCORR_PREC_BF16 ? 2 : 4; }  // tf32 and split-tf32 are fp32-sized

int64_t align256(int64_t v) { return (v + 255) & ~(int64_t)255; }

}  // namespace

int64_t build_tc_workspace_bytes(int B, int D, int H, int W, int precision) {
  const int64_t one = align256((int64_t)B * H * W * D * (int64_t)elem_size(precision));
  return (precision == RAFTCORR_PREC_FP32 ? 4 : 2) * one;   // fp32: hi and lo copies of both fmaps
}

// Shared launcher of build_tc_kernel: operand maps for `passes` (A, B) pairs are already encoded.
static int launch_tc_kernel(bool tf32, const TcOperandMaps& ops, const CUtensorMap* out_maps, const BuildTcParams& p,
                            cudaStream_t stream, bool out16 = false, const CUtensorMap* tail_maps = nullptr) {
  if (tail_maps == nullptr) tail_maps = out_maps;   // (only the interleaved 16-bit epilogue uses them)
  int dev = 0, sms = 0;
  RC_CHECK_CUDA(cudaGetDevice(&dev));
  RC_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const long long total = (long long)p.B * p.tiles_m * p.tiles_y * p.tiles_x;
  if (total == 0) return 0;
  const int grid = (int)(total < sms ? total : sms);
  if (out16 && p.ilog[0]) {
    RC_CHECK_CUDA(cudaFuncSetAttribute(build_tc_kernel<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTotal));
    build_tc_kernel<false, 2><<<grid, kThreads, kSmemTotal, stream>>>(ops, out_maps[0], out_maps[1], out_maps[2], out_maps[3], tail_maps[0], tail_maps[1], p);
  } else if (out16) {
    RC_CHECK_CUDA(cudaFuncSetAttribute(build_tc_kernel<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTotal));
    build_tc_kernel<false, 1><<<grid, kThreads, kSmemTotal, stream>>>(ops, out_maps[0], out_maps[1], out_maps[2], out_maps[3], tail_maps[0], tail_maps[1], p);
  } else if (!tf32) {
    RC_CHECK_CUDA(cudaFuncSetAttribute(build_tc_kernel<false, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTotal));
    build_tc_kernel<false, 0><<<grid, kThreads, kSmemTotal, stream>>>(ops, out_maps[0], out_maps[1], out_maps[2], out_maps[3], tail_maps[0], tail_maps[1], p);
  } else {
    RC_CHECK_CUDA(cudaFuncSetAttribute(build_tc_kernel<true, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTotal));
    build_tc_kernel<true, 0><<<grid, kThreads, kSmemTotal, stream>>>(ops, out_maps[0], out_maps[1], out_maps[2], out_maps[3], tail_maps[0], tail_maps[1], p);
  }
  RC_CHECK_CUDA(cudaGetLastError());
  return 0;
}

static uint32_t make_idesc(bool tf32) {
  const uint32_t fmt = tf32 ? 2u : 1u;
  return (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(kBlockN >> 3) << 17) | ((uint32_t)(kBlockM >> 4) << 24);
}

int launch_build_tc(const float* f1, const float* f2, int B, int D, int H, int W, int L,
                    const raftcorr_level* levels, int precision, void* ws, int64_t ws_bytes,
                    cudaStream_t stream) {
  const int N = H * W;
  if (B == 0) return 0;
  const size_t es = elem_size(precision);
  const bool split = precision == RAFTCORR_PREC_FP32;   // three tf32 passes over hi/lo operands
  const bool tf32 = precision != RAFTCORR_PREC_BF16;
  if (D % 8 != 0) return fail(RAFTCORR_E_UNSUPPORTED, "build_pyramid: tensor-core path needs D %% 8 == 0 (D=%d)", D);
  if (ws == nullptr || ws_bytes < build_tc_workspace_bytes(B, D, H, W, precision) ||
      (reinterpret_cast<uintptr_t>(ws) & 255))
    return fail(RAFTCORR_E_WORKSPACE, "build_pyramid: workspace must be 256-byte aligned and >= %lld B",
                (long long)build_tc_workspace_bytes(B, D, H, W, precision));
  for (int l = 0; l < L; ++l)
    if ((levels[l].pitch & 3) || (levels[l].image_stride & 3) ||
        (reinterpret_cast<uintptr_t>(levels[l].data) & 15))
      return fail(RAFTCORR_E_UNSUPPORTED,
                  "build_pyramid: level %d needs 16-byte aligned rows (pitch %% 4 == 0; use raftcorr_pyramid_layout)", l);

  char* wsb = static_cast<char*>(ws);
  const int64_t one = align256((int64_t)B * N * D * (int64_t)es);
  void* packA = wsb;            // fmap1: bf16 / tf32 / tf32-hi
  void* packB = wsb + one;      // fmap2
  void* packAlo = wsb + 2 * one;  // split mode only
  void* packBlo = wsb + 3 * one;
  dim3 pgrid((N + 31) / 32, (D + 63) / 64, B);
  if (precision == RAFTCORR_PREC_BF16) {
    dim3 g2((N + 63) / 64, (D + 63) / 64, 2 * B);
    pack2_bf16_kernel<<<g2, 256, 0, stream>>>(f1, f2, static_cast<__nv_bfloat16*>(packA),
                                              static_cast<__nv_bfloat16*>(packB), D, N);
  } else if (!split) {
    pack_kernel<float><<<pgrid, 256, 0, stream>>>(f1, static_cast<float*>(packA), D, N);
    pack_kernel<float><<<pgrid, 256, 0, stream>>>(f2, static_cast<float*>(packB), D, N);
  } else {
    pack_split_kernel<<<pgrid, 256, 0, stream>>>(f1, static_cast<float*>(packA), static_cast<float*>(packAlo), D, N);
    pack_split_kernel<<<pgrid, 256, 0, stream>>>(f2, static_cast<float*>(packB), static_cast<float*>(packBlo), D, N);
  }
  RC_CHECK_CUDA(cudaGetLastError());

  const CUtensorMapDataType dt =
      precision == RAFTCORR_PREC_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  const cuuint32_t k_elems = (cuuint32_t)(128 / es);
  TcOperandMaps ops;
  CUtensorMap mapL[RAFTCORR_MAX_LEVELS];
  // pass p multiplies A[a_of[p]] by B[b_of[p]]: hi*hi, hi*lo, lo*hi
  void* a_src[3] = {packA, packA, packAlo};
  void* b_src[3] = {packB, packBlo, packB};
  const int passes = split ? 3 : 1;
  for (int ps = 0; ps < 3; ++ps) {
    const int q = ps < passes ? ps : 0;
    {
      cuuint64_t dims[3] = {(cuuint64_t)D, (cuuint64_t)N, (cuuint64_t)B};
      cuuint64_t strides[2] = {(cuuint64_t)D * es, (cuuint64_t)N * D * es};
      cuuint32_t box[3] = {k_elems, kBlockM, 1};
      if (int rc = encode(&ops.a[ps], dt, 3, a_src[q], dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B, "A")) return rc;
    }
    {
      cuuint64_t dims[4] = {(cuuint64_t)D, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
      cuuint64_t strides[3] = {(cuuint64_t)D * es, (cuuint64_t)W * D * es, (cuuint64_t)N * D * es};
      cuuint32_t box[4] = {k_elems, kTileX, kTileY, 1};
      if (int rc = encode(&ops.b[ps], dt, 4, b_src[q], dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B, "B")) return rc;
    }
  }
  const CUtensorMapSwizzle sw[4] = {CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_SWIZZLE_64B,
                                    CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_SWIZZLE_NONE};
  for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l) {
    const raftcorr_level& lv = levels[l < L ? l : 0];
    cuuint64_t dims[4] = {(cuuint64_t)lv.W, (cuuint64_t)lv.H, (cuuint64_t)N, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)lv.pitch * 4, (cuuint64_t)lv.image_stride * 4,
                             (cuuint64_t)lv.image_stride * 4 * N};
    const int sl = l < L ? l : 0;
    cuuint32_t box[4] = {(cuuint32_t)(kTileX >> sl), 1, 32, 1};
    if (int rc = encode(&mapL[l], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, lv.data, dims, strides, box, sw[sl], "level"))
      return rc;
  }

  BuildTcParams p;
  p.ilog[0] = p.ilog[1] = 0;
  p.B = B;
  p.N = N;
  p.H = H;
  p.W = W;
  p.L = L;
  p.tiles_m = (N + kBlockM - 1) / kBlockM;
  p.tiles_y = (H + kTileY - 1) / kTileY;
  p.tiles_x = (W + kTileX - 1) / kTileX;
  p.k_elems = (int)k_elems;
  p.k_blocks = (int)((D + k_elems - 1) / k_elems);
  p.passes = passes;
  p.plain = 0;
  p.bias = nullptr;
  p.relu = 0;
  p.a_batch_mul = 1;
  p.l3_data = nullptr;
  p.l3_istride = 0;
  for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l) p.bxs[l] = p.bys[l] = 0;
  p.idesc = make_idesc(tf32);
  p.inv_scale = 1.0f / sqrtf((float)D);
  for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l) {
    p.Hl[l] = H >> l;
    p.Wl[l] = W >> l;
  }
  return launch_tc_kernel(tf32, ops, mapL, p, stream);
}

// The same build into the 16-bit blocked pyramid (raftcorr_level16): bf16 operands, fp32 accumulation and
// pooling in TMEM / registers, one rounding to bf16 per stored pixel.  Halves the bytes the epilogue writes.
int launch_build_tc16(const float* f1, const float* f2, int B, int D, int H, int W, int L,
                      const raftcorr_level16* levels, void* ws, int64_t ws_bytes, cudaStream_t stream) {
  const int N = H * W;
  if (B == 0) return 0;
  if (D % 8 != 0)
    return fail(RAFTCORR_E_UNSUPPORTED, "build_pyramid16: needs D %% 8 == 0 (D=%d); use the fp32 pyramid", D);
  const int64_t need = build_tc_workspace_bytes(B, D, H, W, RAFTCORR_PREC_BF16);
  if (ws == nullptr || ws_bytes < need || (reinterpret_cast<uintptr_t>(ws) & 255))
    return fail(RAFTCORR_E_WORKSPACE, "build_pyramid16: workspace must be 256-byte aligned and >= %lld B", (long long)need);
  char* wsb = static_cast<char*>(ws);
  const int64_t one = align256((int64_t)B * N * D * 2);
  void* packA = wsb;
  void* packB = wsb + one;
  dim3 g2((N + 63) / 64, (D + 63) / 64, 2 * B);
  pack2_bf16_kernel<<<g2, 256, 0, stream>>>(f1, f2, static_cast<__nv_bfloat16*>(packA), static_cast<__nv_bfloat16*>(packB), D, N);
  RC_CHECK_CUDA(cudaGetLastError());

  TcOperandMaps ops;
  for (int ps = 0; ps < 3; ++ps) {
    {
      cuuint64_t dims[3] = {(cuuint64_t)D, (cuuint64_t)N, (cuuint64_t)B};
      cuuint64_t strides[2] = {(cuuint64_t)D * 2, (cuuint64_t)N * D * 2};
      cuuint32_t box[3] = {64, kBlockM, 1};
      if (int rc = encode(&ops.a[ps], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, packA, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B, "A")) return rc;
    }
    {
      cuuint64_t dims[4] = {(cuuint64_t)D, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
      cuuint64_t strides[3] = {(cuuint64_t)D * 2, (cuuint64_t)W * D * 2, (cuuint64_t)N * D * 2};
      cuuint32_t box[4] = {64, kTileX, kTileY, 1};
      if (int rc = encode(&ops.b[ps], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, packB, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B, "B")) return rc;
    }
  }
  CUtensorMap mapL[RAFTCORR_MAX_LEVELS], mapT[2];
  for (int l = 0; l < 2; ++l) {   // levels 0, 1: two whole blocks (128 B) per query per box
    if (l < L && levels[l].interleave_log2 > 0) {
      // [group of 2^g images][block row][block][image][64 B]: one box = 2 blocks of the warp's 32 queries
      const raftcorr_level16& lv = levels[l];
      const int g = lv.interleave_log2;
      if (N % (1 << g) != 0 || g > 3 || g < 1)
        return fail(RAFTCORR_E_UNSUPPORTED, "build_pyramid16: level %d interleaves 2^%d images, H*W=%d is not a multiple", l, g, N);
      cuuint64_t dims[5] = {64, 1ull << (g - 1), (cuuint64_t)lv.blocks_x, (cuuint64_t)lv.blocks_y, ((cuuint64_t)B * N) >> g};
      cuuint64_t strides[4] = {128, 64ull << g, (cuuint64_t)lv.blocks_x * (64ull << g), (cuuint64_t)lv.image_stride << g};
      cuuint32_t box[5] = {64, 1u << (g - 1), 2, 1, 32u >> g};
      if (int rc = encode(&mapL[l], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, lv.data, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B, "level16 interleaved"))
        return rc;
      box[4] = 1;
      if (int rc = encode(&mapT[l], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, lv.data, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B, "level16 interleaved, one group"))
        return rc;
      continue;
    }
    // A block row of an image is contiguous (blocks_x * 64 B), so the two innermost axes (pixel in block, block
    // column) are one axis of blocks_x * 32 elements: the box's inner row is then exactly the 128-byte swizzle span
    // (a 64-byte inner axis would be padded to 128-byte rows in shared memory: tools/probe/tma_swizzle_probe.cu),
    // and TMA clips the second block of a pair at the end of the block row.
    const raftcorr_level16& lv = levels[l < L ? l : 0];
    cuuint64_t dims[4] = {(cuuint64_t)lv.blocks_x * 32, (cuuint64_t)lv.blocks_y, (cuuint64_t)N, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)lv.blocks_x * 64, (cuuint64_t)lv.image_stride, (cuuint64_t)lv.image_stride * N};
    cuuint32_t box[4] = {64, 1, 32, 1};
    if (int rc = encode(&mapL[l], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, lv.data, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B, "level16"))
      return rc;
    mapT[l] = mapL[l];
  }
  for (int l = 2; l < L; ++l)
    if (levels[l].interleave_log2 != 0)
      return fail(RAFTCORR_E_UNSUPPORTED, "build_pyramid16: only levels 0 and 1 may be interleaved (level %d)", l);
  {                               // level 2: 32-byte half blocks (rows 0-1 / 2-3 of a block), one per query per box
    const raftcorr_level16& lv = levels[2 < L ? 2 : 0];
    cuuint64_t dims[4] = {16, (cuuint64_t)lv.blocks_x * lv.blocks_y * 2, (cuuint64_t)N, (cuuint64_t)B};
    cuuint64_t strides[3] = {32, (cuuint64_t)lv.image_stride, (cuuint64_t)lv.image_stride * N};
    cuuint32_t box[4] = {16, 1, 32, 1};
    if (int rc = encode(&mapL[2], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, lv.data, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_32B, "level16 half blocks"))
      return rc;
    mapL[3] = mapL[2];            // level 3 is written with plain stores
  }

  BuildTcParams p;
  p.ilog[0] = p.ilog[1] = 0;
  p.B = B;
  p.N = N;
  p.H = H;
  p.W = W;
  p.L = L;
  p.tiles_m = (N + kBlockM - 1) / kBlockM;
  p.tiles_y = (H + kTileY - 1) / kTileY;
  p.tiles_x = (W + kTileX - 1) / kTileX;
  p.k_elems = 64;
  p.k_blocks = (D + 63) / 64;
  p.passes = 1;
  p.plain = 0;
  p.bias = nullptr;
  p.relu = 0;
  p.a_batch_mul = 1;
  p.idesc = make_idesc(false);
  p.inv_scale = 1.0f / sqrtf((float)D);
  for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l) {
    p.Hl[l] = H >> l;
    p.Wl[l] = W >> l;
    p.bxs[l] = l < L ? levels[l].blocks_x : 0;
    p.bys[l] = l < L ? levels[l].blocks_y : 0;
  }
  p.l3_data = L > 3 ? static_cast<unsigned char*>(levels[3].data) : nullptr;
  p.l3_istride = L > 3 ? levels[3].image_stride : 0;
  p.ilog[0] = levels[0].interleave_log2;
  p.ilog[1] = L > 1 ? levels[1].interleave_log2 : p.ilog[0];
  if (p.ilog[0] != p.ilog[1] || (p.ilog[0] != 0 && p.ilog[0] != 2))
    return fail(RAFTCORR_E_UNSUPPORTED, "build_pyramid16: levels 0 and 1 must both be plain or both interleave 4 images (got 2^%d, 2^%d)",
                p.ilog[0], p.ilog[1]);
  return launch_tc_kernel(false, ops, mapL, p, stream, true, mapT);
}

// Batched fp32-accurate GEMM on the tf32 tensor pipe (plain epilogue), used by build_bwd.cu:
//   C[b][m][n] = sum_k A[b][m][k] * B[b][n][k]      A = a_hi + a_lo, B = b_hi + b_lo (tf32 values)
// All operands are K-major fp32 with row pitches / batch strides in elements; pitches and o_pitch
// must be multiples of 4 (16-byte TMA strides).  K and N tails are zero-filled / clipped by TMA.
namespace {
// passes = 3: a_src/b_src = {hi, hi, lo} / {hi, lo, hi}; passes = 1: one tf32 pass over a_src[0] . b_src[0]^T.
// a_bstride == 0 means one A for all batches.  bias (per row of C) and relu are applied in the epilogue.
int launch_tc_gemm_any(const float* const* a_src, const float* const* b_src, int passes, long long a_pitch,
                       long long a_bstride, long long b_pitch, long long b_bstride, float* out, long long o_pitch,
                       long long o_bstride, int B, int M, int Nn, int K, const float* bias, int relu,
                       cudaStream_t stream) {
  if (B == 0 || M == 0 || Nn == 0) return 0;
  if ((a_pitch | b_pitch | o_pitch | a_bstride | b_bstride | o_bstride) & 3)
    return fail(RAFTCORR_E_UNSUPPORTED, "tc_gemm: pitches must be multiples of 4 elements");
  TcOperandMaps ops;
  const bool a_shared = a_bstride == 0;
  for (int ps = 0; ps < 3; ++ps) {
    const int src = ps < passes ? ps : 0;
    {
      cuuint64_t dims[3] = {(cuuint64_t)K, (cuuint64_t)M, (cuuint64_t)(a_shared ? 1 : B)};
      cuuint64_t strides[2] = {(cuuint64_t)a_pitch * 4, (cuuint64_t)(a_shared ? (long long)M * a_pitch : a_bstride) * 4};
      cuuint32_t box[3] = {32, kBlockM, 1};
      if (int rc = encode(&ops.a[ps], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(a_src[src]), dims, strides,
                          box, CU_TENSOR_MAP_SWIZZLE_128B, "gemm A"))
        return rc;
    }
    {
      cuuint64_t dims[4] = {(cuuint64_t)K, (cuuint64_t)Nn, 1, (cuuint64_t)B};
      cuuint64_t strides[3] = {(cuuint64_t)b_pitch * 4, (cuuint64_t)b_bstride * 4, (cuuint64_t)b_bstride * 4};
      cuuint32_t box[4] = {32, kBlockN, 1, 1};
      if (int rc = encode(&ops.b[ps], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(b_src[src]), dims, strides,
                          box, CU_TENSOR_MAP_SWIZZLE_128B, "gemm B"))
        return rc;
    }
  }
  CUtensorMap outm[RAFTCORR_MAX_LEVELS];
  {
    cuuint64_t dims[4] = {(cuuint64_t)Nn, 1, (cuuint64_t)M, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)o_pitch * 4, (cuuint64_t)o_pitch * 4, (cuuint64_t)o_bstride * 4};
    cuuint32_t box[4] = {32, 1, 32, 1};
    for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l)
      if (int rc = encode(&outm[l], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, out, dims, strides, box,
                          CU_TENSOR_MAP_SWIZZLE_128B, "gemm C"))
        return rc;
  }
  BuildTcParams p;
  p.ilog[0] = p.ilog[1] = 0;
  p.B = B;
  p.N = M;     // rows of C (TMEM lanes)
  p.H = 1;
  p.W = Nn;    // columns of C
  p.L = 1;
  p.tiles_m = (M + kBlockM - 1) / kBlockM;
  p.tiles_y = 1;
  p.tiles_x = (Nn + kBlockN - 1) / kBlockN;
  p.k_elems = 32;
  p.k_blocks = (K + 31) / 32;
  p.passes = passes;
  p.plain = 1;
  p.idesc = make_idesc(true);
  p.inv_scale = 1.0f;
  p.bias = bias;
  p.relu = relu;
  p.a_batch_mul = a_shared ? 0 : 1;
  p.l3_data = nullptr;
  p.l3_istride = 0;
  for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l) p.bxs[l] = p.bys[l] = 0;
  for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l) {
    p.Hl[l] = 1;
    p.Wl[l] = Nn;
  }
  return launch_tc_kernel(true, ops, outm, p, stream);
}
}  // namespace

int launch_tc_gemm_split(const float* a_hi, const float* a_lo, long long a_pitch, long long a_bstride,
                         const float* b_hi, const float* b_lo, long long b_pitch, long long b_bstride,
                         float* out, long long o_pitch, long long o_bstride, int B, int M, int Nn, int K,
                         cudaStream_t stream) {
  const float* a_src[3] = {a_hi, a_hi, a_lo};
  const float* b_src[3] = {b_hi, b_lo, b_hi};
  return launch_tc_gemm_any(a_src, b_src, 3, a_pitch, a_bstride, b_pitch, b_bstride, out, o_pitch, o_bstride, B, M, Nn,
                            K, nullptr, 0, stream);
}

// One tf32 pass, C[b][m][n] = act(sum_k A[m][k] * B[b][n][k] + bias[m]) with A shared by all batches: a 1x1
// convolution over NHWC activations whose result lands in NCHW (m = output channel, n = pixel).
int launch_tc_conv1x1(const float* weight, long long w_pitch, const float* act, long long act_pitch,
                      long long act_bstride, float* out, long long o_pitch, long long o_bstride, int B, int M, int Nn,
                      int K, const float* bias, int relu, cudaStream_t stream) {
  const float* a_src[3] = {weight, weight, weight};
  const float* b_src[3] = {act, act, act};
  return launch_tc_gemm_any(a_src, b_src, 1, w_pitch, 0, act_pitch, act_bstride, out, o_pitch, o_bstride, B, M, Nn, K,
                            bias, relu, stream);
}

bool tc_gemm_available() { return get_encode() != nullptr; }

// ---------------------------------------------------------------------------------------------
// Allocation-free variant of the tensor-core build.
//
// launch_build_tc() packs both feature maps into a caller-supplied workspace.  Some integrations
// cannot supply one: a host application that gives the whole device to its own memory pool at
// start-up and hands a plug-in nothing but raw input/output pointers (a ctypes / JNI / cgo
// binding has no handle on the host's caching allocator), or a latency-critical loop whose
// rules forbid cudaMalloc/cudaFree because those calls synchronise the device.  For them the
// packed operands live in the fixed array below, which is part of the loaded library image:
// every GPU that loads libraftcorr_b200.so holds these RAFTCORR_NOALLOC_SCRATCH_BYTES until the
// process exits, whether or not the entry point is ever called.
//
// g_noalloc_scratch is a device symbol.  In host code its name is only a host-side stand-in, not
// a device address, so the address is looked up once per device with cudaGetSymbolAddress (with
// that device current) and that pointer is what the kernels and tensor maps receive.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_noalloc_scratch[RAFTCORR_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;
void* g_noalloc_ptr[kMaxDevices] = {};   // device address of g_noalloc_scratch, per device

// First use on a device performs the symbol lookup (which may initialise the module on that
// device): make that first call outside any stream capture.
int noalloc_scratch_for_current_device(void** out) {
  int dev = -1;
  RC_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail(RAFTCORR_E_UNSUPPORTED, "build_pyramid_noalloc: device %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (g_noalloc_ptr[dev] == nullptr) {
    void* ptr = nullptr;
    const cudaError_t err = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
    if (err != cudaSuccess) return cuda_fail(err, "cudaGetSymbolAddress(g_noalloc_scratch)");
    if (ptr == nullptr) return fail(RAFTCORR_E_DRIVER, "build_pyramid_noalloc: scratch symbol has no device address");
    g_noalloc_ptr[dev] = ptr;
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}

}  // namespace

int launch_build_tc_noalloc(const float* f1, const float* f2, int B, int D, int H, int W, int L,
                            const raftcorr_level* levels, int precision, cudaStream_t stream) {
  const int64_t need = build_tc_workspace_bytes(B, D, H, W, precision);
  if (need > (int64_t)RAFTCORR_NOALLOC_SCRATCH_BYTES)
    return fail(RAFTCORR_E_WORKSPACE,
                "build_pyramid_noalloc: B=%d D=%d H=%d W=%d needs %lld B of packed operands, the built-in "
                "scratch holds %lld B; use raftcorr_build_pyramid with a caller workspace",
                B, D, H, W, (long long)need, (long long)RAFTCORR_NOALLOC_SCRATCH_BYTES);
  void* scratch = nullptr;
  if (int rc = noalloc_scratch_for_current_device(&scratch)) return rc;
  // The pack kernels overwrite every byte of the scratch that the GEMM reads (rows beyond D or
  // N are produced by TMA zero fill, not read from memory), so nothing has to be cleared between
  // calls even though the array still holds the previous call's operands.
  return launch_build_tc(f1, f2, B, D, H, W, L, levels, precision, scratch, RAFTCORR_NOALLOC_SCRATCH_BYTES, stream);
}

}  // namespace raftcorr
