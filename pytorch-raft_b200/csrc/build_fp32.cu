// build_fp32.cu — fp32-accurate all-pairs volume + pyramid (RAFTCORR_PREC_FP32).
//
// Replaces CorrBlock.corr + the avg_pool2d chain (reference core/corr.py:21-27, 52-60) for the
// "fp32 mode" of north_star (<=1e-5 of max|ref|): plain fp32 FMA on the CUDA cores, fp32
// accumulation in k order, true division by sqrt(D), then the 2x2 mean chain level by level in
// the reference's summation order ((a+b)+c)+d)/4.  The tensor-core path is build_tc.cu.
#include "common.cuh"

namespace raftcorr {

namespace {

constexpr int kBM = 64, kBN = 64, kBK = 16;

// C[b,i,j] = sum_d f1[b,d,i] * f2[b,d,j] / sqrtD ; level-0 image i, pixel j.
__global__ void __launch_bounds__(256)
volume_fp32_kernel(const float* __restrict__ f1, const float* __restrict__ f2, float* __restrict__ lvl0,
                   int D, int N, int W, long long pitch, long long istride, float sqrtD) {
  __shared__ float As[kBK][kBM];
  __shared__ float Bs[kBK][kBN];
  const int b = blockIdx.z;
  const int i0 = blockIdx.y * kBM, j0 = blockIdx.x * kBN;
  const float* A = f1 + (long long)b * D * N;
  const float* Bm = f2 + (long long)b * D * N;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < D; k0 += kBK) {
    for (int e = threadIdx.x; e < kBK * kBM; e += 256) {
      const int k = e / kBM, m = e - k * kBM;
      const bool kin = (k0 + k) < D;
      As[k][m] = (kin && i0 + m < N) ? A[(long long)(k0 + k) * N + i0 + m] : 0.0f;
      Bs[k][m] = (kin && j0 + m < N) ? Bm[(long long)(k0 + k) * N + j0 + m] : 0.0f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kBK; ++k) {
      float a[4], c[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        a[u] = As[k][ty * 4 + u];
        c[u] = Bs[k][tx * 4 + u];
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int w = 0; w < 4; ++w) acc[u][w] = fmaf(a[u], c[w], acc[u][w]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int i = i0 + ty * 4 + u;
    if (i >= N) continue;
    float* img = lvl0 + ((long long)b * N + i) * istride;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const int j = j0 + tx * 4 + w;
      if (j >= N) continue;
      const int jy = j / W, jx = j - jy * W;
      img[(long long)jy * pitch + jx] = __fdiv_rn(acc[u][w], sqrtD);  // corr.py:60
    }
  }
}

// one 2x2 mean level (core/corr.py:26), floor on odd sizes
__global__ void __launch_bounds__(256)
pool2_kernel(const float* __restrict__ src, int Ws_pitch, long long s_istride, float* __restrict__ dst,
             int Ho, int Wo, long long d_pitch, long long d_istride, long long images) {
  const long long per = (long long)Ho * Wo;
  const long long total = images * per;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
       e += (long long)gridDim.x * blockDim.x) {
    const long long q = e / per;
    const int rem = (int)(e - q * per);
    const int y = rem / Wo, x = rem - y * Wo;
    const float* s = src + q * s_istride + (long long)(2 * y) * Ws_pitch + 2 * x;
    float sum = __fadd_rn(s[0], s[1]);
    sum = __fadd_rn(sum, s[Ws_pitch]);
    sum = __fadd_rn(sum, s[Ws_pitch + 1]);
    dst[q * d_istride + (long long)y * d_pitch + x] = __fmul_rn(sum, 0.25f);
  }
}

}  // namespace

int launch_pool_chain(const raftcorr_level* levels, int L, long long images, cudaStream_t stream) {
  for (int l = 1; l < L; ++l) {
    const long long total = images * levels[l].H * levels[l].W;
    if (total == 0) continue;
    const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
    pool2_kernel<<<blocks, 256, 0, stream>>>(levels[l - 1].data, (int)levels[l - 1].pitch,
                                             levels[l - 1].image_stride, levels[l].data, levels[l].H,
                                             levels[l].W, levels[l].pitch, levels[l].image_stride,
                                             images);
    RC_CHECK_CUDA(cudaGetLastError());
  }
  return 0;
}

int launch_build_fp32(const float* f1, const float* f2, int B, int D, int H, int W, int L,
                      const raftcorr_level* levels, cudaStream_t stream) {
  const int N = H * W;
  if (B == 0 || N == 0) return 0;
  dim3 grid((N + kBN - 1) / kBN, (N + kBM - 1) / kBM, B);
  volume_fp32_kernel<<<grid, 256, 0, stream>>>(f1, f2, levels[0].data, D, N, W, levels[0].pitch,
                                               levels[0].image_stride, sqrtf((float)D));
  RC_CHECK_CUDA(cudaGetLastError());
  return launch_pool_chain(levels, L, (long long)B * N, stream);
}

}  // namespace raftcorr
