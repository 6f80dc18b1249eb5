// Probe: where does a 4-D TMA box of 64-byte blocks land in shared memory?
//   usage: probe <swizzle 0|32|64|128> <dst_offset_bytes> <nbx> <nby> <bx0> <by0>
// Global tensor: (32 u16 [one 64-byte block], BX blocks, BY block rows, IMG images), element value =
// (by*BX + bx) * 32 + e of image 2 (+ 0x8000 marker in other images).  The kernel loads box (32, nbx, nby, 1)
// at (0, bx0, by0, 2) to smem + dst_offset and dumps 2 KB of shared memory around it; the host reports, for every
// 16-byte chunk of the box, the byte offset (relative to dst) at which it was found.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
__global__ void k(const __grid_constant__ CUtensorMap m, unsigned short* out, int bytes, int off, int bx0, int by0) {
  extern __shared__ __align__(1024) unsigned char smem[];
  unsigned bar = (unsigned)__cvta_generic_to_shared(smem + 8192);
  unsigned dst = (unsigned)__cvta_generic_to_shared(smem + off);
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) reinterpret_cast<unsigned short*>(smem)[i] = 0xEEEE;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
                 ::"r"(dst), "l"(&m), "r"(bar), "r"(0), "r"(bx0), "r"(by0), "r"(2) : "memory");
  }
  unsigned done = 0; long long t0 = clock64();
  while (!done && clock64() - t0 < 200000000LL)
    asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.b32 %0,1,0,p;}" : "=r"(done) : "r"(bar) : "memory");
  __syncthreads();
  if (threadIdx.x == 0) out[4096] = done ? 1 : 0;
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) out[i] = reinterpret_cast<unsigned short*>(smem)[i];
}
int main(int argc, char** argv) {
  if (argc < 7) return 1;
  int sw = atoi(argv[1]), off = atoi(argv[2]), nbx = atoi(argv[3]), nby = atoi(argv[4]), bx0 = atoi(argv[5]), by0 = atoi(argv[6]);
  const int BX = 6, BY = 5, IMG = 4;
  std::vector<unsigned short> h((size_t)IMG * BY * BX * 32);
  for (int q = 0; q < IMG; ++q) for (int b = 0; b < BY * BX; ++b) for (int e = 0; e < 32; ++e)
    h[((size_t)q * BY * BX + b) * 32 + e] = (unsigned short)((b * 32 + e) | (q == 2 ? 0 : 0x8000));
  unsigned short *d, *o; cudaMalloc(&d, h.size() * 2); cudaMalloc(&o, 8194 + 64); cudaMemset(o, 0, 8194);
  cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice);
  void* fn = nullptr; cudaDriverEntryPointQueryResult qr;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr);
  typedef CUresult (*E)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  CUtensorMap m; cuuint64_t dims[4] = {32, BX, BY, IMG}; cuuint64_t st[3] = {64, 64 * BX, 64 * BX * BY};
  cuuint32_t box[4] = {32, (cuuint32_t)nbx, (cuuint32_t)nby, 1}, ones[4] = {1, 1, 1, 1};
  CUtensorMapSwizzle s = sw == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : sw == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : sw == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUresult r = ((E)fn)(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, d, dims, st, box, ones, CU_TENSOR_MAP_INTERLEAVE_NONE, s, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("swizzle %d dst+%d box (32,%d,%d) at block (%d,%d): encode=%d\n", sw, off, nbx, nby, bx0, by0, (int)r);
  if (r) return 0;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384);
  k<<<1, 128, 16384>>>(m, o, nbx * nby * 64, off, bx0, by0);
  cudaError_t e = cudaDeviceSynchronize();
  std::vector<unsigned short> res(4097); cudaMemcpy(res.data(), o, 8194, cudaMemcpyDeviceToHost);
  printf("sync=%s done=%d\n", cudaGetErrorString(e), (int)res[4096]);
  if (e != cudaSuccess) return 0;
  // for each logical 16-byte chunk (block jy*nbx+jx, chunk c) find where it sits
  int linear_ok = 1, sw128_ok = 1, sw64_ok = 1;
  for (int jy = 0; jy < nby; ++jy) for (int jx = 0; jx < nbx; ++jx) for (int c = 0; c < 4; ++c) {
    int gx = bx0 + jx, gy = by0 + jy;
    bool inb = gx >= 0 && gx < BX && gy >= 0 && gy < BY;
    unsigned short first = inb ? (unsigned short)((gy * BX + gx) * 32 + c * 8) : 0;
    int logical = ((jy * nbx + jx) * 4 + c) * 16;
    int found = -1;
    for (int p = 0; p < 8192 - 16; p += 16) {
      bool okc = true;
      for (int e2 = 0; e2 < 8 && okc; ++e2) okc = res[p / 2 + e2] == (unsigned short)(inb ? first + e2 : 0);
      if (okc && inb) { found = p - off; break; }
    }
    int abs_addr = off + logical;                       // smem is 1024-aligned, so offsets are address bits
    int exp128 = (abs_addr ^ (((abs_addr >> 7) & 7) << 4)) - off;
    int exp64 = (abs_addr ^ (((abs_addr >> 7) & 3) << 4)) - off;
    if (inb) {
      if (found != logical) linear_ok = 0;
      if (found != exp128) sw128_ok = 0;
      if (found != exp64) sw64_ok = 0;
      if (jy < 2) printf("  block(%d,%d) chunk %d: logical +%d found +%d (sw128 model +%d)\n", jy, jx, c, logical, found, exp128);
    }
  }
  printf("RESULT sw=%d off=%d: linear=%d sw128model=%d sw64model=%d\n", sw, off, linear_ok, sw128_ok, sw64_ok);
  return 0;
}
