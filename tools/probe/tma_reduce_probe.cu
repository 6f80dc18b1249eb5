// Probe: cp.reduce.async.bulk.tensor .add of a shared-memory box into a 3-D fp32 tensor.
// usage: probe bw bh W H pitch x0 y0   -> adds a box of ones at (x0, y0) of image 2, prints the sum / mismatches
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
__global__ void k(const __grid_constant__ CUtensorMap m, int bw, int bh, int x0, int y0, int q) {
  extern __shared__ __align__(1024) unsigned char smem[];
  float* box = reinterpret_cast<float*>(smem);
  for (int i = threadIdx.x; i < bw * bh; i += blockDim.x) box[i] = 1.0f + i;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned src = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.reduce.async.bulk.tensor.3d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4}], [%1];"
                 ::"l"(&m), "r"(src), "r"(x0), "r"(y0), "r"(q) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}
int main(int argc, char** argv) {
  int bw = atoi(argv[1]), bh = atoi(argv[2]), W = atoi(argv[3]), H = atoi(argv[4]), pitch = atoi(argv[5]), x0 = atoi(argv[6]), y0 = atoi(argv[7]);
  int Q = 5;
  std::vector<float> h((size_t)Q * H * pitch, 100.0f);
  float* d; cudaMalloc(&d, h.size() * 4);
  cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
  void* fn = nullptr; cudaDriverEntryPointQueryResult qr;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr);
  typedef CUresult (*E)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  CUtensorMap m; cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)Q}; cuuint64_t st[2] = {(cuuint64_t)pitch * 4, (cuuint64_t)pitch * H * 4};
  cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, 1}, ones[3] = {1, 1, 1};
  CUresult r = ((E)fn)(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, d, dims, st, box, ones, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("reduce box %dx%d W=%d H=%d pitch=%d at (%d,%d): encode=%d ", bw, bh, W, H, pitch, x0, y0, (int)r);
  if (r) { printf("\n"); return 0; }
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384);
  k<<<1, 128, 16384>>>(m, bw, bh, x0, y0, 2);
  cudaError_t e = cudaDeviceSynchronize();
  std::vector<float> res(h.size()); cudaMemcpy(res.data(), d, h.size() * 4, cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int q = 0; q < Q; ++q) for (int y = 0; y < H; ++y) for (int x = 0; x < pitch; ++x) {
    float exp = 100.0f;
    int bx = x - x0, by = y - y0;
    if (q == 2 && x < W && bx >= 0 && bx < bw && by >= 0 && by < bh) exp += 1.0f + by * bw + bx;
    if (res[((size_t)q * H + y) * pitch + x] != exp) ++bad;
  }
  printf("sync=%s mismatches=%d\n", cudaGetErrorString(e), bad);
  return 0;
}
