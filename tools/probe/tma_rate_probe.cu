// Probe: how fast can one SM's TMA unit fetch many small boxes?  What does a box cost: per box, per row, per byte?
//   usage: probe <inner_bytes 64|128|192|256> <rows> <swizzle 0|64|128> [issuers=8] [boxes_per_issuer=512] [in_flight=4] [log2_row_groups=19]
//   (a small log2_row_groups keeps the source L2-resident: the box rate of the TMA unit itself, DRAM out of the picture)
// Global tensor: 2-D rows of `inner_bytes` (u16 elements), 2^19 x rows rows (>= 256 MB at 64 B: no L2 reuse); every box
// reads `rows` consecutive rows at a pseudo-random position.  148 CTAs x `issuers` warps; each warp keeps 4 boxes in
// flight on its own mbarriers.  Reports boxes/us per SM, rows/us per SM and GB/s over the chip.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
__global__ void __launch_bounds__(512, 1) k(const __grid_constant__ CUtensorMap m, int box_bytes, int rows, int nboxes,
                                            unsigned total_row_groups, unsigned long long* sink, int depth) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const int slot_bytes = (box_bytes + 127) & ~127;
  unsigned bar0 = (unsigned)__cvta_generic_to_shared(smem) + warp * 256;   // 16 warps x 32 barriers = 4 KB
  unsigned dst0 = (unsigned)__cvta_generic_to_shared(smem) + 4096 + warp * depth * slot_bytes;
  if (lane == 0)
    for (int i = 0; i < depth; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8 * i));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();
  if (lane != 0) return;
  unsigned seed = (blockIdx.x * nw + warp) * 2654435761u + 12345u;
  unsigned long long acc = 0;
  int s = 0, round = 0;
  for (int i = 0; i < nboxes + depth; ++i, ++s) {
    if (s == depth) s = 0, ++round;
    if (i >= depth) {
      const unsigned parity = (round - 1) & 1;
      unsigned done = 0;
      while (!done)
        asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.b32 %0,1,0,p;}"
                     : "=r"(done) : "r"(bar0 + 8 * s), "r"(parity) : "memory");
    }
    if (i < nboxes) {
      seed = seed * 1664525u + 1013904223u;
      const int row = (int)((seed >> 8) & (total_row_groups - 1)) * rows;   // total_row_groups is a power of two
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar0 + 8 * s), "r"(box_bytes) : "memory");
      asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                   ::"r"(dst0 + s * slot_bytes), "l"(&m), "r"(bar0 + 8 * s), "r"(0), "r"(row) : "memory");
    }
  }
  if (acc == 0x123456789ULL) *sink = acc;
}
int main(int argc, char** argv) {
  if (argc < 4) return 1;
  const int inner = atoi(argv[1]), rows = atoi(argv[2]), sw = atoi(argv[3]);
  const int issuers = argc > 4 ? atoi(argv[4]) : 8, nboxes = argc > 5 ? atoi(argv[5]) : 512;
  const int depth = argc > 6 ? atoi(argv[6]) : 4;
  const unsigned groups = 1u << (argc > 7 ? atoi(argv[7]) : 19);
  if (depth > 32 || (size_t)issuers * depth * ((inner * rows + 127) & ~127) > 196 * 1024) { printf("too much shared memory\n"); return 0; }
  const size_t bytes = (size_t)groups * rows * inner;
  unsigned char* d; cudaMalloc(&d, bytes); cudaMemset(d, 1, bytes);
  unsigned long long* sink; cudaMalloc(&sink, 8);
  void* fn = nullptr; cudaDriverEntryPointQueryResult qr;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr);
  typedef CUresult (*E)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  CUtensorMap m; cuuint64_t dims[2] = {(cuuint64_t)inner / 2, (cuuint64_t)groups * rows}; cuuint64_t st[1] = {(cuuint64_t)inner};
  cuuint32_t box[2] = {(cuuint32_t)inner / 2, (cuuint32_t)rows}, ones[2] = {1, 1};
  CUtensorMapSwizzle s = sw == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : sw == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUresult r = ((E)fn)(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, d, dims, st, box, ones, CU_TENSOR_MAP_INTERLEAVE_NONE, s, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r) { printf("inner %d rows %d sw %d: encode=%d\n", inner, rows, sw, (int)r); return 0; }
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 201 * 1024 + 512);
  float best = 1e9f;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    k<<<148, issuers * 32, 201 * 1024 + 512>>>(m, inner * rows, rows, nboxes, groups, sink, depth);
    cudaEventRecord(b);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("inner %d rows %d sw %d: %s\n", inner, rows, sw, cudaGetErrorString(e)); return 0; }
    float ms; cudaEventElapsedTime(&ms, a, b);
    if (rep && ms < best) best = ms;
  }
  const double boxes_per_sm = (double)issuers * nboxes, us = best * 1e3;
  printf("inner %3d B x %2d rows, swizzle %3d, %2d issuers x %2d in flight, %4zu MB source: %7.1f us  %6.2f boxes/us/SM  %6.2f rows/us/SM  %7.1f GB/s chip  (%.0f ns per box per SM)\n",
         inner, rows, sw, issuers, depth, bytes >> 20, us, boxes_per_sm / us, boxes_per_sm * rows / us, 148.0 * boxes_per_sm * inner * rows / us / 1e3,
         us * 1e3 / boxes_per_sm);
  return 0;
}
