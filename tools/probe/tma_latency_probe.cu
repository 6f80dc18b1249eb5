// Probe: how long does a warp wait for one "unit" of lookup16 boxes (K boxes of 2 x 4 blocks of 64 B out of K
// neighbouring images), as a function of how many warps per SM fetch at the same time, of a compute pause between
// fetches, and of an L2 prefetch issued some rounds ahead?
//   usage: probe <warps_per_sm> <pause_cycles> <prefetch_rounds_ahead 0=off> [images_log2=14] [boxes_per_unit=32] [group=1] [interleaved=0]
//   interleaved=1: the 64-byte blocks of the `group` images of a group are neighbours in memory ([group][by][bx][image][64 B]),
//   so that a group's box reads pieces of group x 128 B instead of `group` pieces of 128 B 14 KB apart
// Tensor = lookup16's level 0 at 55x128: {64 B block, 16 blocks, 14 block rows, images}; warp w of SM s fetches, in
// round r, images ((r * 148 + s) * W + w) * K ... + K-1 (a stream through memory, like the kernel's tile order).
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
constexpr int kRounds = 24;
__global__ void __launch_bounds__(512, 1) k(const __grid_constant__ CUtensorMap m, int K, int group, int pause, int ahead,
                                            unsigned nimg_mask, unsigned long long* out, int inter) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, W = blockDim.x >> 5;
  const unsigned bar = (unsigned)__cvta_generic_to_shared(smem) + warp * 8;
  const unsigned dst = (unsigned)__cvta_generic_to_shared(smem) + 1024 + warp * (32 * 512);
  if (lane == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();
  if (lane != 0) return;
  long long t_issue = 0, t_wait = 0;
  const int box_bytes = 512 * group, nbox = K / group;
  auto first_image = [&](int r) { return (unsigned)(((r * 148 + blockIdx.x) * W + warp) * K) & nimg_mask; };
  const long long t_begin = clock64();
  for (int r = 0; r < kRounds; ++r) {
    const long long t0 = clock64();
    if (ahead && r + ahead < kRounds) {
      const unsigned img = first_image(r + ahead);
      for (int i = 0; i < nbox; ++i)
        asm volatile("cp.async.bulk.prefetch.tensor.4d.L2.global.tile [%0, {%1, %2, %3, %4}];"
                     ::"l"(&m), "r"(0), "r"((r + ahead) % 14), "r"((r + ahead) % 10), "r"((int)(img + i * group)) : "memory");
    }
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(box_bytes * nbox) : "memory");
    const unsigned img = first_image(r);
    if (inter) {
      for (int i = 0; i < nbox; ++i)
        asm volatile("cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
                     ::"r"(dst + i * box_bytes), "l"(&m), "r"(bar), "r"(0), "r"(0), "r"(r % 14), "r"(r % 10), "r"((int)(img / group + i)) : "memory");
    } else
    for (int i = 0; i < nbox; ++i)
      asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
                   ::"r"(dst + i * box_bytes), "l"(&m), "r"(bar), "r"(0), "r"(r % 14), "r"(r % 10), "r"((int)(img + i * group)) : "memory");
    const long long t1 = clock64();
    unsigned done = 0;
    while (!done)
      asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.b32 %0,1,0,p;}"
                   : "=r"(done) : "r"(bar), "r"(r & 1) : "memory");
    const long long t2 = clock64();
    t_issue += t1 - t0, t_wait += t2 - t1;
    while (clock64() - t2 < pause) {}
  }
  unsigned long long* o = out + (blockIdx.x * W + warp) * 4;
  o[0] = t_issue, o[1] = t_wait, o[2] = clock64() - t_begin;
}
int main(int argc, char** argv) {
  if (argc < 4) return 1;
  const int W = atoi(argv[1]), pause = atoi(argv[2]), ahead = atoi(argv[3]);
  const int lg = argc > 4 ? atoi(argv[4]) : 14, K = argc > 5 ? atoi(argv[5]) : 32, group = argc > 6 ? atoi(argv[6]) : 1;
  const int inter = argc > 7 ? atoi(argv[7]) : 0;
  const size_t nimg = (size_t)1 << lg, bytes = nimg * 14336;
  unsigned char* d; cudaMalloc(&d, bytes); cudaMemset(d, 1, bytes);
  unsigned long long* out; cudaMalloc(&out, 148 * 16 * 4 * 8);
  void* fn = nullptr; cudaDriverEntryPointQueryResult qr;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr);
  typedef CUresult (*E)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  CUtensorMap m; cuuint64_t dims[4] = {32, 16, 14, nimg}; cuuint64_t st[3] = {64, 1024, 14336};
  cuuint32_t box[4] = {32, 2, 4, (cuuint32_t)group}, ones[4] = {1, 1, 1, 1};
  CUresult r = ((E)fn)(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, d, dims, st, box, ones, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (inter) {
    cuuint64_t d5[5] = {32, (cuuint64_t)group, 16, 14, nimg / group};
    cuuint64_t s5[4] = {64, (cuuint64_t)group * 64, (cuuint64_t)group * 1024, (cuuint64_t)group * 14336};
    cuuint32_t b5[5] = {32, (cuuint32_t)group, 2, 4, 1}, o5[5] = {1, 1, 1, 1, 1};
    r = ((E)fn)(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, d, d5, s5, b5, o5, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  if (r) { printf("encode=%d\n", (int)r); return 0; }
  const int smem = 1024 + W * 32 * 512;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  unsigned char* flush; cudaMalloc(&flush, 256 << 20);
  float ms = 0;
  for (int rep = 0; rep < 3; ++rep) {
    cudaMemset(flush, rep, 256 << 20);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    k<<<148, W * 32, smem>>>(m, K, group, pause, ahead, (unsigned)(nimg - 1), out, inter);
    cudaEventRecord(b);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s\n", cudaGetErrorString(e)); return 0; }
    cudaEventElapsedTime(&ms, a, b);
  }
  static unsigned long long h[148 * 16 * 4];
  cudaMemcpy(h, out, sizeof(unsigned long long) * 148 * W * 4, cudaMemcpyDeviceToHost);
  double ti = 0, tw = 0, tt = 0;
  for (int i = 0; i < 148 * W; ++i) ti += h[i * 4], tw += h[i * 4 + 1], tt += h[i * 4 + 2];
  const double n = 148.0 * W * kRounds;
  printf("%2d warps/SM, pause %5d cyc, prefetch %d ahead, %5zu MB, %2d boxes x %d images%s: issue %6.0f  wait %6.0f cycles per unit; "
         "%7.1f us, %6.0f GB/s\n", W, pause, ahead, bytes >> 20, K / group, group, inter ? " interleaved" : "", ti / n, tw / n, ms * 1e3,
         148.0 * W * kRounds * K * 512 / (ms * 1e-3) / 1e9);
  return 0;
}
