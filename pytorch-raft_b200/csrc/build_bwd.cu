// build_bwd.cu — backward of the volume build: d(pyramid) -> d(fmap1), d(fmap2).
//
// The reference gets this from autograd: AvgPool2DBackward0 x3 spreads each
// pooled gradient over its 2x2 parents (/4), DivBackward0 applies 1/sqrt(D), BmmBackward0 forms
//   d f1[b,d,i] = sum_j dC[b,i,j] f2[b,d,j]      d f2[b,d,j] = sum_i dC[b,i,j] f1[b,d,i].
// Here the un-pool + scale is one pass that materialises dC in the caller's workspace as tf32 hi/lo
// splits (plain and transposed), followed by two split-tf32 tcgen05 GEMMs (hi*hi + hi*lo + lo*hi:
// fp32-accurate, gradients are checked at rtol 1e-3); shapes TMA cannot address (H*W % 4 != 0) take
// the fp32 CUDA-core GEMMs further down.
#include "common.cuh"

namespace raftcorr {

namespace {

struct GradLevels {
  const float* data[RAFTCORR_MAX_LEVELS];
  int H[RAFTCORR_MAX_LEVELS], W[RAFTCORR_MAX_LEVELS];
  long long pitch[RAFTCORR_MAX_LEVELS], istride[RAFTCORR_MAX_LEVELS];
  int L;
};

// dC[q, j] = (1/sqrtD) * sum_l 4^-l * dP_l[q, jy>>l, jx>>l]   (levels that cover pixel j)
__global__ void __launch_bounds__(256)
unpool_kernel(const __grid_constant__ GradLevels g, float* __restrict__ dC, long long images, int H,
              int W, float inv_sqrtD) {
  const long long N = (long long)H * W;
  const long long total = images * N;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
       e += (long long)gridDim.x * blockDim.x) {
    const long long q = e / N;
    const int j = (int)(e - q * N);
    const int jy = j / W, jx = j - jy * W;
    float acc = 0.0f, wgt = 1.0f;
    for (int l = 0; l < g.L; ++l) {
      const int yy = jy >> l, xx = jx >> l;
      if (yy < g.H[l] && xx < g.W[l])
        acc += wgt * __ldg(g.data[l] + q * g.istride[l] + (long long)yy * g.pitch[l] + xx);
      wgt *= 0.25f;
    }
    dC[e] = acc * inv_sqrtD;
  }
}

__device__ __forceinline__ void split_tf32(float v, float& hi, float& lo) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
  hi = __uint_as_float(r);
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v - hi));
  lo = __uint_as_float(r);
}

// Tensor-core variant of the un-pool: a 32 (i) x 128 (j) tile of dC, written as tf32 hi/lo splits both
// as dC[i][j] (pitch Np) and transposed as dCT[j][i], so that both backward GEMMs see K-major operands.
// Both copies leave as 16-byte stores (N is a multiple of 4 on this path, so is Np).
constexpr int kUI = 32, kUJ = 128, kUPitch = kUJ + 1;   // odd pitch: conflict-free transposed reads
__global__ void __launch_bounds__(256)
unpool_split_kernel(const __grid_constant__ GradLevels g, float* __restrict__ c_hi, float* __restrict__ c_lo,
                    float* __restrict__ t_hi, float* __restrict__ t_lo, int N, int Np, int W, float inv_sqrtD) {
  __shared__ float tile[kUI][kUPitch];
  const int b = blockIdx.z;
  const int i0 = blockIdx.y * kUI, j0 = blockIdx.x * kUJ;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const long long base = (long long)b * N * Np;
  // phase 1: lane owns 4 consecutive j of one row i; pixel coordinates of the 4 are decoded once
  const int j = j0 + 4 * lane;
  int jy[4], jx[4];
  {
    const int y = j / W, x = j - y * W;
#pragma unroll
    for (int k = 0; k < 4; ++k) {   // the 4 columns may run over the end of an image row
      const int over = (x + k) / W;
      jy[k] = y + over;
      jx[k] = (x + k) - over * W;
    }
  }
  for (int r = w; r < kUI; r += 8) {
    const int i = i0 + r;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (i < N && j < N) {     // N % 4 == 0 and j % 4 == 0: all four columns are in range together
      const long long q = (long long)b * N + i;
      float wgt = inv_sqrtD;
      for (int l = 0; l < g.L; ++l) {
        const float* img = g.data[l] + q * g.istride[l];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int yy = jy[k] >> l, xx = jx[k] >> l;
          if (yy < g.H[l] && xx < g.W[l]) v[k] = fmaf(wgt, __ldg(img + (long long)yy * g.pitch[l] + xx), v[k]);
        }
        wgt *= 0.25f;
      }
      float hi[4], lo[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) split_tf32(v[k], hi[k], lo[k]);
      const long long o = base + (long long)i * Np + j;
      __stcs(reinterpret_cast<float4*>(c_hi + o), make_float4(hi[0], hi[1], hi[2], hi[3]));
      __stcs(reinterpret_cast<float4*>(c_lo + o), make_float4(lo[0], lo[1], lo[2], lo[3]));
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) tile[r][4 * lane + k] = v[k];
  }
  __syncthreads();
  // phase 2: transposed rows dCT[j][i0 .. i0+31]; a lane writes 4 consecutive i of one j, a warp 4 rows
  const int i4 = (lane & 7) * 4, jsub = lane >> 3;
  for (int r = w * 4 + jsub; r < kUJ; r += 32) {
    const int jj = j0 + r, i = i0 + i4;
    if (jj < N && i < N) {     // i % 4 == 0 and N % 4 == 0: the four rows are in range together
      float hi[4], lo[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) split_tf32(tile[i4 + k][r], hi[k], lo[k]);
      const long long o = base + (long long)jj * Np + i;
      __stcs(reinterpret_cast<float4*>(t_hi + o), make_float4(hi[0], hi[1], hi[2], hi[3]));
      __stcs(reinterpret_cast<float4*>(t_lo + o), make_float4(lo[0], lo[1], lo[2], lo[3]));
    }
  }
}

// rows x N fp32 -> tf32 hi/lo with row pitch Np
__global__ void __launch_bounds__(256)
split_rows_kernel(const float* __restrict__ src, float* __restrict__ hi, float* __restrict__ lo, long long rows,
                  int N, int Np) {
  const long long total = rows * N;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const long long r = e / N;
    const int c = (int)(e - r * N);
    float h, l;
    split_tf32(__ldg(src + e), h, l);
    hi[r * Np + c] = h;
    lo[r * Np + c] = l;
  }
}

constexpr int kT = 64, kK = 16;

// out[m, n] = sum_k A[m*a_m + k*a_k] * Bm[k*b_k + n*b_n]   (per batch), 64x64 tiles, 4x4 per thread
__global__ void __launch_bounds__(256)
sgemm_strided_kernel(const float* __restrict__ A, long long a_m, long long a_k, long long a_b,
                     const float* __restrict__ Bm, long long b_k, long long b_n, long long b_b,
                     float* __restrict__ out, long long o_m, long long o_b, int M, int Nn, int K) {
  __shared__ float As[kK][kT + 1];
  __shared__ float Bs[kK][kT + 1];
  const int b = blockIdx.z;
  A += b * a_b;
  Bm += b * b_b;
  out += b * o_b;
  const int m0 = blockIdx.y * kT, n0 = blockIdx.x * kT;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += kK) {
    for (int e = threadIdx.x; e < kK * kT; e += 256) {
      // pick the faster-varying index so that unit-stride operands are read coalesced
      int ka, ma, kb, nb;
      if (a_k == 1) { ka = e % kK; ma = e / kK; } else { ma = e % kT; ka = e / kT; }
      if (b_k == 1) { kb = e % kK; nb = e / kK; } else { nb = e % kT; kb = e / kT; }
      As[ka][ma] = (k0 + ka < K && m0 + ma < M) ? __ldg(A + (m0 + ma) * a_m + (k0 + ka) * a_k) : 0.0f;
      Bs[kb][nb] = (k0 + kb < K && n0 + nb < Nn) ? __ldg(Bm + (k0 + kb) * b_k + (n0 + nb) * b_n) : 0.0f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kK; ++k) {
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
    const int m = m0 + ty * 4 + u;
    if (m >= M) continue;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const int n = n0 + tx * 4 + w;
      if (n < Nn) out[m * o_m + n] = acc[u][w];
    }
  }
}

}  // namespace

int launch_tc_gemm_split(const float* a_hi, const float* a_lo, long long a_pitch, long long a_bstride,
                         const float* b_hi, const float* b_lo, long long b_pitch, long long b_bstride,
                         float* out, long long o_pitch, long long o_bstride, int B, int M, int Nn, int K,
                         cudaStream_t stream);
bool tc_gemm_available();

namespace {
int64_t align256b(int64_t v) { return (v + 255) & ~(int64_t)255; }
int64_t simt_ws_bytes(int B, int N) { return (int64_t)B * N * N * (int64_t)sizeof(float); }
int64_t tc_ws_bytes(int B, int D, int N) {
  const int64_t Np = (N + 3) & ~3;
  return 4 * align256b((int64_t)B * N * Np * 4) + 4 * align256b((int64_t)B * D * Np * 4);
}
}  // namespace

int64_t build_bwd_workspace_bytes(int B, int D, int H, int W) {
  const int N = H * W;
  // tensor-core path (N % 4 == 0: TMA needs 16-byte row strides in d fmap) or the CUDA-core fallback
  return (N % 4 == 0) ? tc_ws_bytes(B, D, N) : simt_ws_bytes(B, N);
}

int launch_build_bwd(const raftcorr_level* glevels, int L, const float* f1, const float* f2,
                     float* g1, float* g2, int B, int D, int H, int W, void* ws, int64_t ws_bytes,
                     cudaStream_t stream) {
  const long long N = (long long)H * W;
  if (B == 0) return 0;
  GradLevels g;
  g.L = L;
  for (int l = 0; l < RAFTCORR_MAX_LEVELS; ++l) {
    const int s = l < L ? l : 0;
    g.data[l] = glevels[s].data;
    g.H[l] = glevels[s].H;
    g.W[l] = glevels[s].W;
    g.pitch[l] = glevels[s].pitch;
    g.istride[l] = glevels[s].image_stride;
  }
  const float inv_sqrtD = 1.0f / sqrtf((float)D);
  const bool tc = (N % 4 == 0) && tc_gemm_available() && ws != nullptr && ws_bytes >= tc_ws_bytes(B, D, (int)N) &&
                  !(reinterpret_cast<uintptr_t>(ws) & 255) && !(reinterpret_cast<uintptr_t>(g1) & 15) &&
                  !(reinterpret_cast<uintptr_t>(g2) & 15);
  if (tc) {
    // fp32-accurate tensor-core path: d f1 = dC . f2^T and d f2 = dC^T . f1^T as two split-tf32 GEMMs
    const long long Np = (N + 3) & ~3LL;
    char* w = static_cast<char*>(ws);
    const int64_t cn = align256b((int64_t)B * N * Np * 4), fn = align256b((int64_t)B * D * Np * 4);
    float* c_hi = reinterpret_cast<float*>(w);
    float* c_lo = reinterpret_cast<float*>(w + cn);
    float* t_hi = reinterpret_cast<float*>(w + 2 * cn);
    float* t_lo = reinterpret_cast<float*>(w + 3 * cn);
    float* f1_hi = reinterpret_cast<float*>(w + 4 * cn);
    float* f1_lo = reinterpret_cast<float*>(w + 4 * cn + fn);
    float* f2_hi = reinterpret_cast<float*>(w + 4 * cn + 2 * fn);
    float* f2_lo = reinterpret_cast<float*>(w + 4 * cn + 3 * fn);
    dim3 ugrid((unsigned)((N + kUJ - 1) / kUJ), (unsigned)((N + kUI - 1) / kUI), (unsigned)B);
    unpool_split_kernel<<<ugrid, 256, 0, stream>>>(g, c_hi, c_lo, t_hi, t_lo, (int)N, (int)Np, W, inv_sqrtD);
    RC_CHECK_CUDA(cudaGetLastError());
    split_rows_kernel<<<148 * 8, 256, 0, stream>>>(f1, f1_hi, f1_lo, (long long)B * D, (int)N, (int)Np);
    split_rows_kernel<<<148 * 8, 256, 0, stream>>>(f2, f2_hi, f2_lo, (long long)B * D, (int)N, (int)Np);
    RC_CHECK_CUDA(cudaGetLastError());
    // d f1[b][d][i] = sum_j f2[b][d][j] * dC[b][i][j]
    if (int rc = launch_tc_gemm_split(f2_hi, f2_lo, Np, (long long)D * Np, c_hi, c_lo, Np, N * Np, g1, N,
                                      (long long)D * N, B, D, (int)N, (int)N, stream))
      return rc;
    // d f2[b][d][j] = sum_i f1[b][d][i] * dC[b][i][j]
    return launch_tc_gemm_split(f1_hi, f1_lo, Np, (long long)D * Np, t_hi, t_lo, Np, N * Np, g2, N,
                                (long long)D * N, B, D, (int)N, (int)N, stream);
  }
  if (ws == nullptr || ws_bytes < simt_ws_bytes(B, (int)N))
    return fail(RAFTCORR_E_WORKSPACE, "build_bwd: workspace %lld B < %lld B", (long long)ws_bytes,
                (long long)simt_ws_bytes(B, (int)N));
  float* dC = static_cast<float*>(ws);
  unpool_kernel<<<148 * 8, 256, 0, stream>>>(g, dC, (long long)B * N, H, W, inv_sqrtD);
  RC_CHECK_CUDA(cudaGetLastError());
  dim3 grid((unsigned)((N + kT - 1) / kT), (unsigned)((D + kT - 1) / kT), (unsigned)B);
  // d f1[d, i] = sum_j f2[d, j] * dC[i, j]
  sgemm_strided_kernel<<<grid, 256, 0, stream>>>(f2, N, 1, (long long)D * N, dC, 1, N, N * N, g1, N,
                                                 (long long)D * N, D, (int)N, (int)N);
  RC_CHECK_CUDA(cudaGetLastError());
  // d f2[d, j] = sum_i f1[d, i] * dC[i, j]
  sgemm_strided_kernel<<<grid, 256, 0, stream>>>(f1, N, 1, (long long)D * N, dC, N, 1, N * N, g2, N,
                                                 (long long)D * N, D, (int)N, (int)N);
  RC_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace raftcorr
