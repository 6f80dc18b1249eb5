/*
 * corr_oracle.c — CPU restatement of RAFT's correlation hot path (TEST INFRASTRUCTURE ONLY).
 *
 * This file is the checker, never the product: only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it.  The product path
 * (pytorch-raft_b200/) never links, imports or calls anything under oracle/.
 *
 * Each function restates the algorithm of the reference (zhaoyuzhi/PyTorch-RAFT) and cites the
 * file:line it follows.  The arithmetic that lives in PyTorch (ATen bmm / avg_pool2d /
 * grid_sampler_2d, torch 2.11) is restated from its published semantics and pinned against the
 * reference itself run in the build container (oracle/make_golden.py -> tests/golden/).
 *
 * Pin status: oracle_corr_volume / oracle_avg_pool2 / oracle_lookup are pinned against vectors the
 * unmodified reference produced (tests/golden/, tests/test_oracle.py: pooling and lookup bit-exact,
 * volume <= 1e-5).  oracle_alt_forward restates alt_cuda_corr's CUDA kernel, which cannot be executed
 * without a GPU and may not be recompiled: for that function PARITY IS UNPINNED by execution; it is
 * only cross-checked against the reference CorrBlock's output (pooling commutes with the dot product).
 *
 * Build: gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC (see oracle/Makefile).
 * -ffp-contract=off matters: the coordinate arithmetic must round after every operation
 * exactly as the separate ATen kernels do.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_COORD_ATEN_CPU 0  /* true division  (CPU div kernel)                       */
#define ORACLE_COORD_ATEN_CUDA 1 /* multiply by 1/(size-1) (CUDA div-by-scalar kernel)    */
#define ORACLE_COORD_DIRECT 2    /* no normalise round trip (alt_cuda_corr kernel)        */

int oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

void oracle_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

/* -------------------------------------------------------------------------------------------
 * All-pairs volume.  Reference: core/corr.py:52-60 (CorrBlock.corr):
 *   corr = matmul(fmap1.view(B,D,N).transpose(1,2), fmap2.view(B,D,N)) / sqrt(tensor(D).float())
 * fmaps are NCHW fp32; vol is (B, N, N) == (B,H,W,1,H,W).
 * The dot product accumulates in double and rounds once to float (ATen's sgemm accumulates in
 * float in an unspecified order; both are within 1e-6 relative of the exact value), then the
 * division by sqrtf(D) is a float division like the reference's.
 * ----------------------------------------------------------------------------------------- */
void oracle_corr_volume(const float* f1, const float* f2, float* vol, int B, int D, int H, int W) {
  const int64_t N = (int64_t)H * W;
  const float s = sqrtf((float)D);
  for (int b = 0; b < B; ++b) {
    const float* a = f1 + (int64_t)b * D * N;
    const float* c = f2 + (int64_t)b * D * N;
    float* v = vol + (int64_t)b * N * N;
#pragma omp parallel
    {
      double* acc = (double*)malloc(sizeof(double) * (size_t)N);
#pragma omp for schedule(static)
      for (int64_t i = 0; i < N; ++i) {
        for (int64_t j = 0; j < N; ++j) acc[j] = 0.0;
        for (int d = 0; d < D; ++d) {
          const double ai = (double)a[(int64_t)d * N + i];
          const float* row = c + (int64_t)d * N;
          for (int64_t j = 0; j < N; ++j) acc[j] += ai * (double)row[j];
        }
        for (int64_t j = 0; j < N; ++j) v[i * N + j] = (float)acc[j] / s;
      }
      free(acc);
    }
  }
}

/* -------------------------------------------------------------------------------------------
 * 2x2 stride-2 average pool, floor on odd sizes, no padding.
 * Reference: core/corr.py:25-27 (F.avg_pool2d(corr, 2, stride=2)); ATen cpu_avg_pool2d sums
 * the window in (kh, kw) order in float and divides by 4.
 * in: (n, H, W) contiguous, out: (n, H/2, W/2) contiguous.
 * ----------------------------------------------------------------------------------------- */
void oracle_avg_pool2(const float* in, float* out, int64_t n, int H, int W) {
  const int Ho = H / 2, Wo = W / 2;
#pragma omp parallel for schedule(static)
  for (int64_t q = 0; q < n; ++q) {
    const float* src = in + q * H * W;
    float* dst = out + q * Ho * Wo;
    for (int y = 0; y < Ho; ++y)
      for (int x = 0; x < Wo; ++x) {
        float sum = src[(2 * y) * W + 2 * x];
        sum += src[(2 * y) * W + 2 * x + 1];
        sum += src[(2 * y + 1) * W + 2 * x];
        sum += src[(2 * y + 1) * W + 2 * x + 1];
        dst[y * Wo + x] = sum / 4.0f;
      }
  }
}

/* -------------------------------------------------------------------------------------------
 * Pixel coordinate -> sampling index, exactly as the reference evaluates it in fp32.
 * Reference: core/utils/utils.py:61-62 (xgrid = 2*xgrid/(W-1) - 1) followed by ATen
 * grid_sampler_2d with align_corners=True (GridSampler.h:27-36: ((coord+1)/2)*(size-1); the
 * vectorised CPU kernel evaluates (coord+1)*((size-1)/2), which rounds identically because the
 * halving is exact).
 * ----------------------------------------------------------------------------------------- */
static inline float sample_index(float p, int size, int mode) {
  if (mode == ORACLE_COORD_DIRECT) return p;
  const float sm1 = (float)(size - 1);
  float t = 2.0f * p;
  if (mode == ORACLE_COORD_ATEN_CPU) {
    t = t / sm1;
  } else {
    const float inv = 1.0f / sm1;
    t = t * inv;
  }
  t = t - 1.0f;
  t = t + 1.0f;
  t = t * (sm1 / 2.0f);
  return t;
}

static inline int floor_to_int(float f) {
  /* floorf then saturating conversion; NaN -> INT32_MIN (out of bounds) */
  if (!(f == f)) return INT32_MIN;
  if (f >= 2147483520.0f) return INT32_MAX - 1;
  if (f <= -2147483648.0f) return INT32_MIN;
  return (int)f;
}

/* One bilinear sample with zero padding.  ATen GridSamplerKernel.cpp (ApplyGridSample,
 * bilinear, zeros): x_w=floor(x), w=x-x_w, e=1-w, n=y-y_n, s=1-n, nw=s*e, ne=s*w, sw=n*e,
 * se=n*w; out = nw_val*nw + ne_val*ne + sw_val*sw + se_val*se, out-of-image corners read 0. */
static inline float bilinear_zero(const float* img, int H, int W, int64_t pitch, float ix, float iy) {
  const float xw = floorf(ix), yn = floorf(iy);
  const float w = ix - xw, e = 1.0f - w, n = iy - yn, s = 1.0f - n;
  const float nw = s * e, ne = s * w, sw = n * e, se = n * w;
  const int x0 = floor_to_int(xw), y0 = floor_to_int(yn);
  const int x1 = x0 + 1, y1 = y0 + 1;
  const int mx0 = (x0 > -1) && (x0 < W), mx1 = (x1 > -1) && (x1 < W);
  const int my0 = (y0 > -1) && (y0 < H), my1 = (y1 > -1) && (y1 < H);
  const float vnw = (mx0 && my0) ? img[(int64_t)y0 * pitch + x0] : 0.0f;
  const float vne = (mx1 && my0) ? img[(int64_t)y0 * pitch + x1] : 0.0f;
  const float vsw = (mx0 && my1) ? img[(int64_t)y1 * pitch + x0] : 0.0f;
  const float vse = (mx1 && my1) ? img[(int64_t)y1 * pitch + x1] : 0.0f;
  /* ATen's AVX2 build contracts the sum into this FMA chain; bit-exact vs torch 2.11 CPU */
  return fmaf(vse, se, fmaf(vsw, sw, fmaf(vne, ne, vnw * nw)));
}

/* -------------------------------------------------------------------------------------------
 * Pyramid window lookup.  Reference: core/corr.py:29-50 (CorrBlock.__call__) +
 * core/utils/utils.py:57-65 (bilinear_sampler).
 *   levels[l]: (B*N, Hl[l], Wl[l]) fp32 with row pitch pitch[l] (elements) and image stride
 *              istride[l] (elements)
 *   coords:    (B, 2, H, W) fp32, channel 0 = x, channel 1 = y          (utils.py:74-77)
 *   out:       (B, L*(2r+1)^2, H, W) fp32; channel = l*rd*rd + i*rd + j where i moves x and j
 *              moves y (corr.py:37-43: delta=stack(meshgrid(dy,dx)) added to (x,y))
 * ----------------------------------------------------------------------------------------- */
void oracle_lookup(const float* const* levels, const int* Hl, const int* Wl, const int64_t* pitch,
                   const int64_t* istride, const float* coords, float* out, int B, int H, int W,
                   int L, int r, int mode) {
  const int rd = 2 * r + 1;
  const int64_t N = (int64_t)H * W;
  const int C = L * rd * rd;
#pragma omp parallel for schedule(static)
  for (int64_t q = 0; q < (int64_t)B * N; ++q) {
    const int b = (int)(q / N);
    const int64_t p = q % N;
    const float cx = coords[((int64_t)b * 2 + 0) * N + p];
    const float cy = coords[((int64_t)b * 2 + 1) * N + p];
    for (int l = 0; l < L; ++l) {
      const float* img = levels[l] + q * istride[l];
      const float lx = cx / (float)(1 << l); /* corr.py:41  coords / 2**i */
      const float ly = cy / (float)(1 << l);
      for (int i = 0; i < rd; ++i) {
        const float px = lx + (float)(i - r); /* corr.py:43 centroid + delta */
        const float ix = sample_index(px, Wl[l], mode);
        for (int j = 0; j < rd; ++j) {
          const float py = ly + (float)(j - r);
          const float iy = sample_index(py, Hl[l], mode);
          const int ch = l * rd * rd + i * rd + j;
          out[((int64_t)b * C + ch) * N + p] = bilinear_zero(img, Hl[l], Wl[l], pitch[l], ix, iy);
        }
      }
    }
  }
}

/* -------------------------------------------------------------------------------------------
 * On-the-fly correlation, one level.  Reference: alt_cuda_corr/correlation_kernel.cu:18-119
 * (corr_forward_kernel) as called from core/corr.py:80-87.
 *   fmap1: (B,H1,W1,C) NHWC, fmap2: (B,H2,W2,C) NHWC, coords: (B,1,H1,W1,2) last dim (x,y)
 *   out:   (B,1,rd*rd,H1,W1), UNSCALED (corr.py:91 divides by sqrt(C) afterwards)
 * The reference kernel takes the dot product at the (2r+2)^2 integer taps around floor(coords)
 * (.cu:71-90) and scatters each into <=4 outputs with weights dy*dx, dy*(1-dx), (1-dy)*dx,
 * (1-dy)*(1-dx) (.cu:92-114); output index iy + rd*ix is x-major.  Gathering instead of
 * scattering gives, for output (ix, iy):
 *   out = (1-dy)(1-dx)*S[iy][ix] + (1-dy)dx*S[iy][ix+1] + dy(1-dx)*S[iy+1][ix] + dy*dx*S[iy+1][ix+1]
 * with S[iy][ix] = <fmap1[q], fmap2[floor(y)-r+iy, floor(x)-r+ix]> (zero outside the image).
 * mode selects how the tap position is derived from coords (ORACLE_COORD_DIRECT = reference
 * kernel; the ATEN modes replay CorrBlock's normalise round trip per tap, as oracle_lookup).
 * ----------------------------------------------------------------------------------------- */
void oracle_alt_forward(const float* fmap1, const float* fmap2, const float* coords, float* out,
                        int B, int H1, int W1, int H2, int W2, int C, int r, int mode) {
  const int rd = 2 * r + 1;
  const int64_t N1 = (int64_t)H1 * W1;
#pragma omp parallel for schedule(static)
  for (int64_t q = 0; q < (int64_t)B * N1; ++q) {
    const int b = (int)(q / N1);
    const int64_t p = q % N1;
    const float* a = fmap1 + q * C;
    const float cx = coords[q * 2 + 0], cy = coords[q * 2 + 1];
    for (int i = 0; i < rd; ++i) {
      const float ix = sample_index(cx + (float)(i - r), W2, mode);
      const float xw = floorf(ix);
      const float w = ix - xw, e = 1.0f - w;
      const int x0 = floor_to_int(xw);
      for (int j = 0; j < rd; ++j) {
        const float iy = sample_index(cy + (float)(j - r), H2, mode);
        const float yn = floorf(iy);
        const float n = iy - yn, s = 1.0f - n;
        const int y0 = floor_to_int(yn);
        float acc = 0.0f;
        for (int cy2 = 0; cy2 < 2; ++cy2)
          for (int cx2 = 0; cx2 < 2; ++cx2) {
            const int yy = y0 + cy2, xx = x0 + cx2;
            if (yy < 0 || yy >= H2 || xx < 0 || xx >= W2) continue;
            const float* t = fmap2 + (((int64_t)b * H2 + yy) * W2 + xx) * C;
            double dot = 0.0;
            for (int c = 0; c < C; ++c) dot += (double)a[c] * (double)t[c];
            const float wgt = (cy2 ? n : s) * (cx2 ? w : e);
            acc += (float)dot * wgt;
          }
        out[((int64_t)b * rd * rd + (i * rd + j)) * N1 + p] = acc;
      }
    }
  }
}

/* ---- the step after the path: convex 8x upsampling (core/raft.py:72-83) --------------------
 * mask (B,576,H,W) viewed (B,1,9,8,8,H,W) and soft-maxed over the 9 (raft.py:75-76); 8*flow
 * unfolded 3x3 with zero padding (raft.py:78-79); weighted sum over the 9 (raft.py:81); the
 * permute/reshape of raft.py:82-83 puts sub-pixel (dy,dx) of coarse (y,x) at (8y+dy, 8x+dx). */
void oracle_upsample_flow(const float* flow, const float* mask, float* up, int B, int H, int W) {
  const int64_t cs = (int64_t)H * W;
#pragma omp parallel for schedule(static)
  for (int64_t row = 0; row < (int64_t)B * H; ++row) {
    const int b = (int)(row / H), y = (int)(row % H);
    for (int x = 0; x < W; ++x)
      for (int dy = 0; dy < 8; ++dy)
        for (int dx = 0; dx < 8; ++dx) {
          float v[9], mx = -INFINITY, sum = 0.0f;
          for (int k = 0; k < 9; ++k) {
            v[k] = mask[((int64_t)b * 576 + k * 64 + dy * 8 + dx) * cs + (int64_t)y * W + x];
            if (v[k] > mx) mx = v[k];
          }
          for (int k = 0; k < 9; ++k) {
            v[k] = expf(v[k] - mx);
            sum += v[k];
          }
          for (int c = 0; c < 2; ++c) {
            float acc = 0.0f;
            for (int k = 0; k < 9; ++k) {
              const int yy = y + k / 3 - 1, xx = x + k % 3 - 1;
              const float f = (yy < 0 || yy >= H || xx < 0 || xx >= W)
                                  ? 0.0f
                                  : 8.0f * flow[((int64_t)b * 2 + c) * cs + (int64_t)yy * W + xx];
              acc += (v[k] / sum) * f;
            }
            up[(((int64_t)b * 2 + c) * 8 * H + 8 * y + dy) * 8 * W + 8 * x + dx] = acc;
          }
        }
  }
}

/* 8 * F.interpolate(flow, 8x, bilinear, align_corners=True) (core/utils/utils.py:80-82); source index
 * = dst * (in-1)/(out-1) as ATen's area_pixel_compute_source_index does for align_corners. */
void oracle_upflow8(const float* flow, float* up, int B, int H, int W) {
  const int H8 = 8 * H, W8 = 8 * W;
  const float sy = H8 > 1 ? (float)(H - 1) / (float)(H8 - 1) : 0.0f;
  const float sx = W8 > 1 ? (float)(W - 1) / (float)(W8 - 1) : 0.0f;
#pragma omp parallel for schedule(static)
  for (int64_t row = 0; row < (int64_t)B * 2 * H8; ++row) {
    const int64_t img = row / H8;
    const int oy = (int)(row % H8);
    const float fy = sy * (float)oy;
    const int y0 = (int)fy, y1 = y0 + (y0 < H - 1 ? 1 : 0);
    const float ly = fy - (float)y0, hy = 1.0f - ly;
    const float* src = flow + img * (int64_t)H * W;
    for (int ox = 0; ox < W8; ++ox) {
      const float fx = sx * (float)ox;
      const int x0 = (int)fx, x1 = x0 + (x0 < W - 1 ? 1 : 0);
      const float lx = fx - (float)x0, hx = 1.0f - lx;
      const float v = hy * (hx * src[y0 * W + x0] + lx * src[y0 * W + x1]) +
                      ly * (hx * src[y1 * W + x0] + lx * src[y1 * W + x1]);
      up[row * W8 + ox] = 8.0f * v;
    }
  }
}
