// Probe: what does HBM take from a kernel that only writes?  (The roof of the 16-bit pyramid build, which writes 1.02 GB
// and reads 58 MB.)  Variants: plain 16-byte stores, streaming (.cs) stores, 32-byte stores, TMA bulk stores from shared
// memory; 2 GiB per launch, best of 5.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
__global__ void k_st(uint4* p, size_t n, int mode) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const uint4 v = make_uint4(1, 2, 3, 4);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    if (mode == 0) p[i] = v;
    else if (mode == 1) __stcs(p + i, v);
    else asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p + i), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
  }
}
// each CTA streams 16 KB chunks out of shared memory with cp.async.bulk (1-D TMA store)
__global__ void k_bulk(unsigned char* p, size_t bytes) {
  extern __shared__ __align__(128) unsigned char sm[];
  for (int i = threadIdx.x; i < 16384 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sm)[i] = i;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t src = (uint32_t)__cvta_generic_to_shared(sm);
    int inflight = 0;
    for (size_t off = (size_t)blockIdx.x * 16384; off + 16384 <= bytes; off += (size_t)gridDim.x * 16384) {
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(p + off), "r"(src), "r"(16384) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      if (++inflight >= 8) asm volatile("cp.async.bulk.wait_group.read 7;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}
int main() {
  const size_t bytes = (size_t)2 << 30;
  unsigned char* d; cudaMalloc(&d, bytes);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  auto run = [&](const char* name, auto launch) {
    float best = 1e9f;
    for (int r = 0; r < 6; ++r) {
      cudaEventRecord(a); launch(); cudaEventRecord(b);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
      float ms; cudaEventElapsedTime(&ms, a, b);
      if (r && ms < best) best = ms;
    }
    printf("%-44s %7.1f us  %6.0f GB/s\n", name, best * 1e3, bytes / (best * 1e-3) / 1e9);
  };
  run("cudaMemsetAsync", [&] { cudaMemsetAsync(d, 1, bytes); });
  for (int mode = 0; mode < 3; ++mode)
    for (int ctas : {148 * 4, 148 * 8, 148 * 16}) {
      char name[96]; snprintf(name, 96, "%s 16-byte stores, %d CTAs x 512", mode == 0 ? "plain" : mode == 1 ? "st.cs" : "no_allocate", ctas);
      run(name, [&] { k_st<<<ctas, 512>>>((uint4*)d, bytes / 16, mode); });
    }
  cudaFuncSetAttribute(k_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384);
  for (int ctas : {148, 148 * 2, 148 * 4, 148 * 8})  {
    char name[96]; snprintf(name, 96, "TMA bulk stores of 16 KB, %d CTAs", ctas);
    run(name, [&] { k_bulk<<<ctas, 128, 16384>>>(d, bytes); });
  }
  return 0;
}
