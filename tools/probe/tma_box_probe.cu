// Probe: which TMA box shapes load correctly with SWIZZLE_NONE?  usage: probe bw bh W H pitch x0 y0
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
__global__ void k(const __grid_constant__ CUtensorMap m, float* out, int bw, int bh, int x0, int y0, int q) {
  extern __shared__ __align__(1024) unsigned char smem[];
  unsigned bar = (unsigned)__cvta_generic_to_shared(smem + 8192);
  unsigned dst = (unsigned)__cvta_generic_to_shared(smem);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bw * bh * 4) : "memory");
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(dst), "l"(&m), "r"(bar), "r"(x0), "r"(y0), "r"(q) : "memory");
  }
  unsigned done = 0; long long t0 = clock64();
  while (!done && clock64() - t0 < 200000000LL)
    asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.b32 %0,1,0,p;}" : "=r"(done) : "r"(bar) : "memory");
  if (threadIdx.x == 0) out[1023] = done ? 1.f : -1.f;
  for (int i = threadIdx.x; i < bw * bh; i += blockDim.x) out[i] = reinterpret_cast<float*>(smem)[i];
}
int main(int argc, char** argv) {
  int bw = atoi(argv[1]), bh = atoi(argv[2]), W = atoi(argv[3]), H = atoi(argv[4]), pitch = atoi(argv[5]), x0 = atoi(argv[6]), y0 = atoi(argv[7]);
  int Q = 5;
  std::vector<float> h((size_t)Q * H * pitch);
  for (int q = 0; q < Q; ++q) for (int y = 0; y < H; ++y) for (int x = 0; x < pitch; ++x) h[((size_t)q * H + y) * pitch + x] = q * 10000 + y * 100 + x;
  float *d, *o; cudaMalloc(&d, h.size() * 4); cudaMalloc(&o, 4096); cudaMemset(o, 0, 4096);
  cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
  void* fn = nullptr; cudaDriverEntryPointQueryResult qr;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr);
  typedef CUresult (*E)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  CUtensorMap m; cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)Q}; cuuint64_t st[2] = {(cuuint64_t)pitch * 4, (cuuint64_t)pitch * H * 4};
  cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, 1}, ones[3] = {1, 1, 1};
  CUresult r = ((E)fn)(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, d, dims, st, box, ones, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("box %dx%d W=%d H=%d pitch=%d at (%d,%d): encode=%d ", bw, bh, W, H, pitch, x0, y0, (int)r);
  if (r) { printf("\n"); return 0; }
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384);
  k<<<1, 128, 16384>>>(m, o, bw, bh, x0, y0, 2);
  cudaError_t e = cudaDeviceSynchronize();
  std::vector<float> res(1024); cudaMemcpy(res.data(), o, 4096, cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int y = 0; y < bh; ++y) for (int x = 0; x < bw; ++x) {
    int gx = x0 + x, gy = y0 + y; float exp = (gx >= 0 && gx < W && gy >= 0 && gy < H) ? 2 * 10000 + gy * 100 + gx : 0.f;
    if (res[y * bw + x] != exp) ++bad;
  }
  printf("sync=%s done=%g mismatches=%d first row: %g %g %g %g\n", cudaGetErrorString(e), res[1023], bad, res[0], res[1], res[2], res[3]);
  return 0;
}
