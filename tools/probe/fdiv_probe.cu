// Is q' = fma(fma(-q, b, a), inv, q), q = a*inv, inv = RN(1/b) equal to RN(a/b) for small integer b?
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(unsigned long long* bad, unsigned* first, int bmax) {
  const unsigned stride = gridDim.x * blockDim.x;
  for (int b = 1; b <= bmax; ++b) {
    const float fb = (float)b, inv = __frcp_rn(fb);
    for (unsigned bits = blockIdx.x * blockDim.x + threadIdx.x; bits < 0xFFFFFFFFu - stride; bits += stride * 61u) {
      const float a = __uint_as_float(bits);
      if (!(fabsf(a) < 3e38f) || (fabsf(a) < 1e-30f)) continue;  // finite, away from underflow
      const float ref = __fdiv_rn(a, fb);
      const float q = __fmul_rn(a, inv);
      const float r = __fmaf_rn(-q, fb, a);
      const float q2 = __fmaf_rn(r, inv, q);
      if (q2 != ref) { if (atomicAdd(bad, 1ULL) == 0) { first[0] = bits; first[1] = b; } }
    }
  }
}
int main() {
  unsigned long long* bad; unsigned* first;
  cudaMallocManaged(&bad, 8); cudaMallocManaged(&first, 8); *bad = 0;
  k<<<148 * 8, 256>>>(bad, first, 4100);
  cudaError_t e = cudaDeviceSynchronize();
  printf("%s mismatches=%llu first bits=%08x b=%u\n", cudaGetErrorString(e), *bad, first[0], first[1]);
}
