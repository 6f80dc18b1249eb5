// api.cu — the extern "C" surface declared in include/raftcorr.h: argument checks + launches.
#include <stdarg.h>
#include <stdio.h>

#include "common.cuh"

namespace raftcorr {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

int cuda_fail(cudaError_t err, const char* what) {
  snprintf(g_err, sizeof(g_err), "CUDA error %d (%s) at %s", (int)err, cudaGetErrorString(err), what);
  return (int)err;
}

// launchers implemented in the kernel files
int launch_lookup_fwd(const raftcorr_level* levels, int L, const float* coords, float* out, int B,
                      int H, int W, int r, int mode, int nhwc, cudaStream_t stream);
int launch_lookup_bwd(const raftcorr_level* levels, const raftcorr_level* glevels, int L,
                      const float* coords, const float* gout, float* gcoords, int B, int H, int W,
                      int r, int mode, cudaStream_t stream);
int launch_build_fp32(const float* f1, const float* f2, int B, int D, int H, int W, int L,
                      const raftcorr_level* levels, cudaStream_t stream);
int64_t build_tc_workspace_bytes(int B, int D, int H, int W, int precision);
int launch_build_tc(const float* f1, const float* f2, int B, int D, int H, int W, int L,
                    const raftcorr_level* levels, int precision, void* ws, int64_t ws_bytes,
                    cudaStream_t stream);
int launch_build_tc_noalloc(const float* f1, const float* f2, int B, int D, int H, int W, int L,
                            const raftcorr_level* levels, int precision, cudaStream_t stream);
int launch_build_tc16(const float* f1, const float* f2, int B, int D, int H, int W, int L,
                      const raftcorr_level16* levels, void* ws, int64_t ws_bytes, cudaStream_t stream);
int launch_alt_pack(const float* fmap, int B, int D, int H, int W, int L, void* const* out_levels, cudaStream_t stream);
int lookup16_group_log2(int level);   // lookup16.cu: the lookup fetches the windows of 2^g neighbouring queries as one box
int launch_lookup16_fwd(const raftcorr_level16* levels, int L, const float* coords, float* out, int B, int H, int W,
                        int r, int mode, cudaStream_t stream);
int64_t build_bwd_workspace_bytes(int B, int D, int H, int W);
int launch_build_bwd(const raftcorr_level* glevels, int L, const float* f1, const float* f2,
                     float* g1, float* g2, int B, int D, int H, int W, void* ws, int64_t ws_bytes,
                     cudaStream_t stream);
int launch_alt_fwd(const float* f1, const float* f2, const float* coords, float* corr, int B, int H1,
                   int W1, int H2, int W2, int C, int r, int mode, float scale, cudaStream_t stream);
int launch_alt_tc(const void* f1_bf16, const void* const* f2_bf16, const float* coords, float* out, int B,
                  int H1, int W1, int C, int L, int r, int mode, float scale, cudaStream_t stream);
int launch_upsample_fwd(const float* flow, const float* mask, float* up, int B, int H, int W, cudaStream_t stream);
int launch_upsample_bwd(const float* flow, const float* mask, const float* gup, float* gflow, float* gmask, int B,
                        int H, int W, cudaStream_t stream);
int launch_upflow8(const float* flow, float* up, int B, int H, int W, cudaStream_t stream);
int launch_tc_conv1x1(const float* weight, long long w_pitch, const float* act, long long act_pitch,
                      long long act_bstride, float* out, long long o_pitch, long long o_bstride, int B, int M, int Nn,
                      int K, const float* bias, int relu, cudaStream_t stream);
bool tc_gemm_available();
int launch_conv1x1_tc(const float* weight, const float* act, float* out, int B, int M, int Nn, int K,
                      const float* bias, int relu, cudaStream_t stream, int nchw = 0);
int launch_alt_bwd(const float* f1, const float* f2, const float* coords, const float* gcorr,
                   float* g1, float* g2, float* gcoords, int B, int H1, int W1, int H2, int W2, int C,
                   int r, int mode, float scale, cudaStream_t stream);

static int check_levels(const raftcorr_level* levels, int L, int H, int W, const char* what) {
  RC_REQUIRE(levels != nullptr, "%s: levels is NULL", what);
  RC_REQUIRE(L >= 1 && L <= RAFTCORR_MAX_LEVELS, "%s: num_levels=%d not in [1,%d]", what, L,
             RAFTCORR_MAX_LEVELS);
  for (int l = 0; l < L; ++l) {
    RC_REQUIRE(levels[l].data != nullptr, "%s: levels[%d].data is NULL", what, l);
    RC_REQUIRE(levels[l].H == (H >> l) && levels[l].W == (W >> l),
               "%s: levels[%d] is %dx%d, expected %dx%d (avg_pool2d floors)", what, l, levels[l].H,
               levels[l].W, H >> l, W >> l);
    RC_REQUIRE(levels[l].H >= 1 && levels[l].W >= 1, "%s: level %d is empty", what, l);
    RC_REQUIRE(levels[l].pitch >= levels[l].W &&
                   levels[l].image_stride >= levels[l].pitch * levels[l].H,
               "%s: levels[%d] pitch/image_stride too small", what, l);
  }
  return 0;
}

static int check_levels16(const raftcorr_level16* levels, int L, int H, int W, const char* what, long long images = 0) {
  RC_REQUIRE(levels != nullptr, "%s: levels is NULL", what);
  RC_REQUIRE(L >= 1 && L <= RAFTCORR_MAX_LEVELS, "%s: num_levels=%d not in [1,%d]", what, L, RAFTCORR_MAX_LEVELS);
  for (int l = 0; l < L; ++l) {
    const raftcorr_level16& v = levels[l];
    RC_REQUIRE(v.data != nullptr, "%s: levels[%d].data is NULL", what, l);
    RC_REQUIRE(!(reinterpret_cast<uintptr_t>(v.data) & 63), "%s: levels[%d].data must be 64-byte aligned", what, l);
    RC_REQUIRE(v.H == (H >> l) && v.W == (W >> l) && v.H >= 1 && v.W >= 1,
               "%s: levels[%d] is %dx%d, expected %dx%d (avg_pool2d floors)", what, l, v.H, v.W, H >> l, W >> l);
    RC_REQUIRE(v.blocks_x == (v.W + 7) / 8 && v.blocks_y == (v.H + 3) / 4,
               "%s: levels[%d] block grid %dx%d does not match %dx%d pixels (use raftcorr_pyramid16_layout)", what, l,
               v.blocks_x, v.blocks_y, v.W, v.H);
    RC_REQUIRE(v.image_stride >= (int64_t)v.blocks_x * v.blocks_y * 64 && !(v.image_stride & 63),
               "%s: levels[%d].image_stride must be a multiple of 64 and >= blocks_x*blocks_y*64", what, l);
    RC_REQUIRE(v.interleave_log2 >= 0 && v.interleave_log2 <= 3, "%s: levels[%d].interleave_log2=%d not in [0,3]", what, l,
               v.interleave_log2);
    RC_REQUIRE(v.interleave_log2 == 0 || images % (1LL << v.interleave_log2) == 0,
               "%s: levels[%d] interleaves 2^%d images, B*H*W=%lld is not a multiple", what, l, v.interleave_log2, images);
    RC_REQUIRE(v.interleave_log2 == 0 || v.image_stride == (int64_t)v.blocks_x * v.blocks_y * 64,
               "%s: levels[%d]: an interleaved level needs image_stride == blocks_x*blocks_y*64", what, l);
  }
  return 0;
}

static bool valid_mode(int m) { return m >= RAFTCORR_COORD_ATEN_CPU && m <= RAFTCORR_COORD_DIRECT; }

}  // namespace raftcorr

using namespace raftcorr;

extern "C" {

int raftcorr_version(void) { return 100; }

const char* raftcorr_last_error(void) { return g_err; }

int raftcorr_pyramid_layout(int B, int H, int W, int num_levels, raftcorr_level* levels,
                            int64_t* floats_per_level) {
  RC_REQUIRE(levels && floats_per_level, "pyramid_layout: NULL output");
  RC_REQUIRE(B >= 0 && H >= 1 && W >= 1, "pyramid_layout: bad sizes B=%d H=%d W=%d", B, H, W);
  RC_REQUIRE(num_levels >= 1 && num_levels <= RAFTCORR_MAX_LEVELS, "pyramid_layout: num_levels=%d",
             num_levels);
  for (int l = 0; l < num_levels; ++l) {
    const int Hl = H >> l, Wl = W >> l;
    RC_REQUIRE(Hl >= 1 && Wl >= 1, "pyramid_layout: level %d of %dx%d is empty", l, H, W);
    levels[l].H = Hl;
    levels[l].W = Wl;
    levels[l].pitch = (Wl + 3) & ~3;  // rows start on 16 B so TMA can address them
    levels[l].image_stride = levels[l].pitch * Hl;
    floats_per_level[l] = (int64_t)B * H * W * levels[l].image_stride;
  }
  return 0;
}

int64_t raftcorr_build_workspace_bytes(int B, int D, int H, int W, int precision) {
  if (D % 8 != 0) return 0;   // CUDA-core build (any precision request), no workspace
  return build_tc_workspace_bytes(B, D, H, W, precision);
}

int raftcorr_build_pyramid(const float* fmap1, const float* fmap2, int B, int D, int H, int W,
                           int num_levels, const raftcorr_level* levels, int precision,
                           void* workspace, int64_t workspace_bytes, void* stream) {
  RC_REQUIRE(fmap1 && fmap2, "build_pyramid: fmap pointer is NULL");
  RC_REQUIRE(B >= 0 && D >= 1 && H >= 1 && W >= 1, "build_pyramid: bad sizes B=%d D=%d H=%d W=%d", B,
             D, H, W);
  RC_REQUIRE((long long)H * W <= (1 << 24), "build_pyramid: fmap too large");
  if (int rc = check_levels(levels, num_levels, H, W, "build_pyramid")) return rc;
  RC_REQUIRE(precision == RAFTCORR_PREC_FP32 || precision == RAFTCORR_PREC_BF16 || precision == RAFTCORR_PREC_TF32,
             "build_pyramid: unknown precision %d", precision);
  if (precision == RAFTCORR_PREC_FP32) {
    // fp32-accurate: three tf32 tensor-core passes over hi/lo splits when the layout allows it and
    // the caller brought the workspace for them; otherwise plain fp32 FMA on the CUDA cores
    bool tc_ok = D % 8 == 0 && workspace != nullptr && !(reinterpret_cast<uintptr_t>(workspace) & 255) &&
                 workspace_bytes >= build_tc_workspace_bytes(B, D, H, W, precision);
    for (int l = 0; l < num_levels && tc_ok; ++l)
      tc_ok = !(levels[l].pitch & 3) && !(levels[l].image_stride & 3) &&
              !(reinterpret_cast<uintptr_t>(levels[l].data) & 15);
    if (!tc_ok) return launch_build_fp32(fmap1, fmap2, B, D, H, W, num_levels, levels, (cudaStream_t)stream);
  } else if (D % 8 != 0) {
    // the tensor-core operand layout needs 16-byte rows of K; odd channel counts take the (more
    // accurate) fp32 CUDA-core build instead of failing - the reference accepts any D
    return launch_build_fp32(fmap1, fmap2, B, D, H, W, num_levels, levels, (cudaStream_t)stream);
  }
  return launch_build_tc(fmap1, fmap2, B, D, H, W, num_levels, levels, precision, workspace,
                         workspace_bytes, (cudaStream_t)stream);
}

int raftcorr_build_pyramid_noalloc(const float* fmap1, const float* fmap2, int B, int D, int H, int W,
                                   int num_levels, const raftcorr_level* levels, int precision,
                                   void* stream) {
  RC_REQUIRE(fmap1 && fmap2, "build_pyramid_noalloc: fmap pointer is NULL");
  RC_REQUIRE(B >= 0 && D >= 1 && H >= 1 && W >= 1, "build_pyramid_noalloc: bad sizes B=%d D=%d H=%d W=%d",
             B, D, H, W);
  RC_REQUIRE((long long)H * W <= (1 << 24), "build_pyramid_noalloc: fmap too large");
  RC_REQUIRE(precision == RAFTCORR_PREC_BF16 || precision == RAFTCORR_PREC_TF32,
             "build_pyramid_noalloc: precision must be RAFTCORR_PREC_BF16 or RAFTCORR_PREC_TF32 "
             "(the fp32 build needs no workspace: call raftcorr_build_pyramid)");
  if (int rc = check_levels(levels, num_levels, H, W, "build_pyramid_noalloc")) return rc;
  // channel counts the tensor-core operand layout cannot take go to the fp32 CUDA-core build, exactly as in
  // raftcorr_build_pyramid; that build uses no scratch memory at all
  if (D % 8 != 0) return launch_build_fp32(fmap1, fmap2, B, D, H, W, num_levels, levels, (cudaStream_t)stream);
  return launch_build_tc_noalloc(fmap1, fmap2, B, D, H, W, num_levels, levels, precision,
                                 (cudaStream_t)stream);
}

int raftcorr_pyramid16_layout(int B, int H, int W, int num_levels, raftcorr_level16* levels,
                              int64_t* bytes_per_level) {
  RC_REQUIRE(levels && bytes_per_level, "pyramid16_layout: NULL output");
  RC_REQUIRE(B >= 0 && H >= 1 && W >= 1, "pyramid16_layout: bad sizes B=%d H=%d W=%d", B, H, W);
  RC_REQUIRE(num_levels >= 1 && num_levels <= RAFTCORR_MAX_LEVELS, "pyramid16_layout: num_levels=%d", num_levels);
  for (int l = 0; l < num_levels; ++l) {
    const int Hl = H >> l, Wl = W >> l;
    RC_REQUIRE(Hl >= 1 && Wl >= 1, "pyramid16_layout: level %d of %dx%d is empty", l, H, W);
    levels[l].H = Hl;
    levels[l].W = Wl;
    levels[l].blocks_x = (Wl + 7) / 8;
    levels[l].blocks_y = (Hl + 3) / 4;
    levels[l].image_stride = (int64_t)levels[l].blocks_x * levels[l].blocks_y * 64;
    // levels 0 and 1 (images of several KB, windows = a few 128..192-byte pieces of them) keep the blocks of the
    // 2^g queries the lookup fetches together side by side: one DRAM piece of 2^g * 128..192 bytes instead of 2^g
    // (W a multiple of the group size: every row of queries then starts a group, so the lookup's lane groups are the
    // groups in memory; otherwise the plain layout, where any four neighbouring lanes can share a box)
    levels[l].interleave_log2 = (l < 2 && W % (1 << lookup16_group_log2(l)) == 0) ? lookup16_group_log2(l) : 0;
    levels[l].reserved = 0;
    bytes_per_level[l] = (int64_t)B * H * W * levels[l].image_stride;
  }
  return 0;
}

int raftcorr_build_pyramid16(const float* fmap1, const float* fmap2, int B, int D, int H, int W,
                             int num_levels, const raftcorr_level16* levels, void* workspace,
                             int64_t workspace_bytes, void* stream) {
  RC_REQUIRE(fmap1 && fmap2, "build_pyramid16: fmap pointer is NULL");
  RC_REQUIRE(B >= 0 && D >= 1 && H >= 1 && W >= 1, "build_pyramid16: bad sizes B=%d D=%d H=%d W=%d", B, D, H, W);
  RC_REQUIRE((long long)H * W <= (1 << 24), "build_pyramid16: fmap too large");
  if (int rc = check_levels16(levels, num_levels, H, W, "build_pyramid16", (long long)B * H * W)) return rc;
  return launch_build_tc16(fmap1, fmap2, B, D, H, W, num_levels, levels, workspace, workspace_bytes,
                           (cudaStream_t)stream);
}

int raftcorr_lookup16_fwd(const raftcorr_level16* levels, int num_levels, const float* coords, float* out,
                          int B, int H, int W, int radius, int coord_mode, void* stream) {
  RC_REQUIRE(coords && out, "lookup16_fwd: coords/out is NULL");
  RC_REQUIRE(B >= 0 && H >= 1 && W >= 1, "lookup16_fwd: bad sizes B=%d H=%d W=%d", B, H, W);
  RC_REQUIRE(valid_mode(coord_mode), "lookup16_fwd: coord_mode=%d", coord_mode);
  if (int rc = check_levels16(levels, num_levels, H, W, "lookup16_fwd", (long long)B * H * W)) return rc;
  return launch_lookup16_fwd(levels, num_levels, coords, out, B, H, W, radius, coord_mode, (cudaStream_t)stream);
}

int raftcorr_alt_pack(const float* fmap, int B, int D, int H, int W, int num_levels, void* const* out_levels,
                      void* stream) {
  RC_REQUIRE(fmap && out_levels, "alt_pack: NULL pointer");
  RC_REQUIRE(B >= 0 && D >= 1 && H >= 1 && W >= 1, "alt_pack: bad sizes B=%d D=%d H=%d W=%d", B, D, H, W);
  RC_REQUIRE(num_levels >= 1 && num_levels <= RAFTCORR_MAX_LEVELS, "alt_pack: num_levels=%d", num_levels);
  for (int l = 0; l < num_levels; ++l) {
    RC_REQUIRE((H >> l) >= 1 && (W >> l) >= 1, "alt_pack: level %d of %dx%d is empty", l, H, W);
    RC_REQUIRE(out_levels[l] != nullptr && !(reinterpret_cast<uintptr_t>(out_levels[l]) & 3),
               "alt_pack: out_levels[%d] is NULL or not 4-byte aligned", l);
  }
  return launch_alt_pack(fmap, B, D, H, W, num_levels, out_levels, (cudaStream_t)stream);
}

int raftcorr_lookup_fwd(const raftcorr_level* levels, int num_levels, const float* coords,
                        float* out, int B, int H, int W, int radius, int coord_mode, void* stream) {
  RC_REQUIRE(coords && out, "lookup_fwd: coords/out is NULL");
  RC_REQUIRE(B >= 0 && H >= 1 && W >= 1, "lookup_fwd: bad sizes B=%d H=%d W=%d", B, H, W);
  RC_REQUIRE(radius >= 0 && radius <= 4, "lookup_fwd: radius=%d not in [0,4]", radius);
  RC_REQUIRE(valid_mode(coord_mode), "lookup_fwd: coord_mode=%d", coord_mode);
  if (int rc = check_levels(levels, num_levels, H, W, "lookup_fwd")) return rc;
  return launch_lookup_fwd(levels, num_levels, coords, out, B, H, W, radius, coord_mode, 0,
                           (cudaStream_t)stream);
}

int raftcorr_lookup_fwd_nhwc(const raftcorr_level* levels, int num_levels, const float* coords,
                             float* out, int B, int H, int W, int radius, int coord_mode, void* stream) {
  RC_REQUIRE(coords && out, "lookup_fwd_nhwc: coords/out is NULL");
  RC_REQUIRE(B >= 0 && H >= 1 && W >= 1, "lookup_fwd_nhwc: bad sizes B=%d H=%d W=%d", B, H, W);
  RC_REQUIRE(radius >= 0 && radius <= 4, "lookup_fwd_nhwc: radius=%d not in [0,4]", radius);
  RC_REQUIRE(valid_mode(coord_mode), "lookup_fwd_nhwc: coord_mode=%d", coord_mode);
  if (int rc = check_levels(levels, num_levels, H, W, "lookup_fwd_nhwc")) return rc;
  return launch_lookup_fwd(levels, num_levels, coords, out, B, H, W, radius, coord_mode, 1,
                           (cudaStream_t)stream);
}

int raftcorr_lookup_bwd(const raftcorr_level* levels, const raftcorr_level* grad_levels,
                        int num_levels, const float* coords, const float* grad_out,
                        float* grad_coords, int B, int H, int W, int radius, int coord_mode,
                        void* stream) {
  RC_REQUIRE(coords && grad_out, "lookup_bwd: coords/grad_out is NULL");
  RC_REQUIRE(B >= 0 && H >= 1 && W >= 1, "lookup_bwd: bad sizes B=%d H=%d W=%d", B, H, W);
  RC_REQUIRE(radius >= 0 && radius <= 4, "lookup_bwd: radius=%d not in [0,4]", radius);
  RC_REQUIRE(valid_mode(coord_mode), "lookup_bwd: coord_mode=%d", coord_mode);
  RC_REQUIRE(grad_levels || grad_coords, "lookup_bwd: nothing to compute");
  if (grad_coords || !grad_levels)
    if (int rc = check_levels(levels, num_levels, H, W, "lookup_bwd(levels)")) return rc;
  if (grad_levels)
    if (int rc = check_levels(grad_levels, num_levels, H, W, "lookup_bwd(grad_levels)")) return rc;
  return launch_lookup_bwd(levels, grad_levels, num_levels, coords, grad_out, grad_coords, B, H, W,
                           radius, coord_mode, (cudaStream_t)stream);
}

int64_t raftcorr_build_bwd_workspace_bytes(int B, int D, int H, int W) {
  return build_bwd_workspace_bytes(B, D, H, W);
}

int raftcorr_build_bwd(const raftcorr_level* grad_levels, int num_levels, const float* fmap1,
                       const float* fmap2, float* grad_fmap1, float* grad_fmap2, int B, int D,
                       int H, int W, void* workspace, int64_t workspace_bytes, void* stream) {
  RC_REQUIRE(fmap1 && fmap2 && grad_fmap1 && grad_fmap2, "build_bwd: NULL pointer");
  RC_REQUIRE(B >= 0 && D >= 1 && H >= 1 && W >= 1, "build_bwd: bad sizes");
  if (int rc = check_levels(grad_levels, num_levels, H, W, "build_bwd")) return rc;
  return launch_build_bwd(grad_levels, num_levels, fmap1, fmap2, grad_fmap1, grad_fmap2, B, D, H, W,
                          workspace, workspace_bytes, (cudaStream_t)stream);
}

int raftcorr_alt_fwd(const float* fmap1, const float* fmap2, const float* coords, float* corr,
                     int B, int H1, int W1, int H2, int W2, int C, int radius, int coord_mode,
                     float scale, void* stream) {
  RC_REQUIRE(fmap1 && fmap2 && coords && corr, "alt_fwd: NULL pointer");
  RC_REQUIRE(B >= 0 && H1 >= 1 && W1 >= 1 && H2 >= 1 && W2 >= 1 && C >= 1, "alt_fwd: bad sizes");
  RC_REQUIRE(radius >= 0 && radius <= 4, "alt_fwd: radius=%d not in [0,4]", radius);
  RC_REQUIRE(valid_mode(coord_mode), "alt_fwd: coord_mode=%d", coord_mode);
  return launch_alt_fwd(fmap1, fmap2, coords, corr, B, H1, W1, H2, W2, C, radius, coord_mode, scale,
                        (cudaStream_t)stream);
}

int raftcorr_alt_fused_fwd(const void* fmap1_bf16, const void* const* fmap2_bf16, const float* coords,
                           float* out, int B, int H1, int W1, int C, int num_levels, int radius,
                           int coord_mode, float scale, void* stream) {
  RC_REQUIRE(fmap1_bf16 && fmap2_bf16 && coords && out, "alt_fused_fwd: NULL pointer");
  RC_REQUIRE(B >= 0 && H1 >= 1 && W1 >= 1 && C >= 1, "alt_fused_fwd: bad sizes");
  RC_REQUIRE(num_levels >= 1 && num_levels <= RAFTCORR_MAX_LEVELS, "alt_fused_fwd: num_levels=%d", num_levels);
  RC_REQUIRE(valid_mode(coord_mode), "alt_fused_fwd: coord_mode=%d", coord_mode);
  for (int l = 0; l < num_levels; ++l) RC_REQUIRE(fmap2_bf16[l] != nullptr, "alt_fused_fwd: fmap2 level %d is NULL", l);
  RC_REQUIRE(!(reinterpret_cast<uintptr_t>(fmap1_bf16) & 15), "alt_fused_fwd: fmap1 must be 16-byte aligned");
  return launch_alt_tc(fmap1_bf16, fmap2_bf16, coords, out, B, H1, W1, C, num_levels, radius, coord_mode, scale,
                       (cudaStream_t)stream);
}

int raftcorr_alt_bwd(const float* fmap1, const float* fmap2, const float* coords,
                     const float* corr_grad, float* fmap1_grad, float* fmap2_grad,
                     float* coords_grad, int B, int H1, int W1, int H2, int W2, int C, int radius,
                     int coord_mode, float scale, void* stream) {
  RC_REQUIRE(fmap1 && fmap2 && coords && corr_grad && fmap1_grad && fmap2_grad,
             "alt_bwd: NULL pointer");
  RC_REQUIRE(B >= 0 && H1 >= 1 && W1 >= 1 && H2 >= 1 && W2 >= 1 && C >= 1, "alt_bwd: bad sizes");
  RC_REQUIRE(radius >= 0 && radius <= 4, "alt_bwd: radius=%d not in [0,4]", radius);
  RC_REQUIRE(valid_mode(coord_mode), "alt_bwd: coord_mode=%d", coord_mode);
  return launch_alt_bwd(fmap1, fmap2, coords, corr_grad, fmap1_grad, fmap2_grad, coords_grad, B, H1,
                        W1, H2, W2, C, radius, coord_mode, scale, (cudaStream_t)stream);
}

int raftcorr_upsample_flow_fwd(const float* flow, const float* mask, float* up, int B, int H, int W,
                               void* stream) {
  RC_REQUIRE(flow && mask && up, "upsample_flow_fwd: NULL pointer");
  RC_REQUIRE(B >= 0 && H >= 1 && W >= 1, "upsample_flow_fwd: bad sizes B=%d H=%d W=%d", B, H, W);
  RC_REQUIRE(B <= 65535 && H <= 65535, "upsample_flow_fwd: B and H must fit a CUDA grid dimension");
  RC_REQUIRE(!(reinterpret_cast<uintptr_t>(up) & 15), "upsample_flow_fwd: up must be 16-byte aligned");
  return launch_upsample_fwd(flow, mask, up, B, H, W, (cudaStream_t)stream);
}

int raftcorr_upsample_flow_bwd(const float* flow, const float* mask, const float* grad_up, float* grad_flow,
                               float* grad_mask, int B, int H, int W, void* stream) {
  RC_REQUIRE(flow && mask && grad_up, "upsample_flow_bwd: NULL pointer");
  RC_REQUIRE(grad_flow || grad_mask, "upsample_flow_bwd: nothing to compute");
  RC_REQUIRE(B >= 0 && H >= 1 && W >= 1, "upsample_flow_bwd: bad sizes B=%d H=%d W=%d", B, H, W);
  RC_REQUIRE(B <= 65535 && H <= 65535, "upsample_flow_bwd: B and H must fit a CUDA grid dimension");
  RC_REQUIRE(!(reinterpret_cast<uintptr_t>(grad_up) & 15), "upsample_flow_bwd: grad_up must be 16-byte aligned");
  return launch_upsample_bwd(flow, mask, grad_up, grad_flow, grad_mask, B, H, W, (cudaStream_t)stream);
}

int raftcorr_upflow8(const float* flow, float* up, int B, int H, int W, void* stream) {
  RC_REQUIRE(flow && up, "upflow8: NULL pointer");
  RC_REQUIRE(B >= 0 && H >= 1 && W >= 1, "upflow8: bad sizes B=%d H=%d W=%d", B, H, W);
  RC_REQUIRE(!(reinterpret_cast<uintptr_t>(up) & 15), "upflow8: up must be 16-byte aligned");
  return launch_upflow8(flow, up, B, H, W, (cudaStream_t)stream);
}

int64_t raftcorr_lookup_conv1x1_workspace_bytes(int B, int H, int W, int num_levels, int radius) {
  if (B < 0 || H < 1 || W < 1 || num_levels < 1 || num_levels > RAFTCORR_MAX_LEVELS || radius < 0 || radius > 4)
    return RAFTCORR_E_INVALID;
  const int64_t rd = 2 * radius + 1;
  return (int64_t)B * H * W * num_levels * rd * rd * (int64_t)sizeof(float);
}

int raftcorr_conv1x1_nhwc_fwd(const float* act, const float* weight, const float* bias, float* out, int B,
                              int pixels, int in_channels, int out_channels, int relu, void* stream) {
  RC_REQUIRE(act && weight && out, "conv1x1_nhwc_fwd: act/weight/out is NULL");
  RC_REQUIRE(B >= 0 && pixels >= 1 && in_channels >= 1 && out_channels >= 1, "conv1x1_nhwc_fwd: bad sizes");
  if ((in_channels & 3) || (pixels & 3) || !tc_gemm_available() || (reinterpret_cast<uintptr_t>(weight) & 15) ||
      (reinterpret_cast<uintptr_t>(act) & 15) || (reinterpret_cast<uintptr_t>(out) & 15))
    return raftcorr::fail(RAFTCORR_E_UNSUPPORTED,
                          "conv1x1_nhwc_fwd: needs in_channels %% 4 == 0 and pixels %% 4 == 0 (%d, %d) and 16-byte "
                          "aligned pointers", in_channels, pixels);
  // all output channels x 128 pixels per tile when there are at most 256 output channels (csrc/conv1x1_tc.cu)
  const int rc = launch_conv1x1_tc(weight, act, out, B, out_channels, pixels, in_channels, bias, relu,
                                   (cudaStream_t)stream);
  if (rc != 1) return rc;
  return launch_tc_conv1x1(weight, in_channels, act, in_channels, (long long)pixels * in_channels, out, pixels,
                           (long long)out_channels * pixels, B, out_channels, pixels, in_channels, bias, relu,
                           (cudaStream_t)stream);
}

int raftcorr_conv1x1_nchw_fwd(const float* act, const float* weight, const float* bias, float* out, int B,
                              int pixels, int in_channels, int out_channels, int relu, void* stream) {
  RC_REQUIRE(act && weight && out, "conv1x1_nchw_fwd: act/weight/out is NULL");
  RC_REQUIRE(B >= 0 && pixels >= 1 && in_channels >= 1 && out_channels >= 1, "conv1x1_nchw_fwd: bad sizes");
  if ((in_channels & 3) || (pixels & 3) || out_channels > 256 || !tc_gemm_available() ||
      (reinterpret_cast<uintptr_t>(weight) & 15) || (reinterpret_cast<uintptr_t>(act) & 15) ||
      (reinterpret_cast<uintptr_t>(out) & 15))
    return raftcorr::fail(RAFTCORR_E_UNSUPPORTED,
                          "conv1x1_nchw_fwd: needs in_channels %% 4 == 0, pixels %% 4 == 0, out_channels <= 256 (%d, %d, %d) "
                          "and 16-byte aligned pointers", in_channels, pixels, out_channels);
  const int rc = launch_conv1x1_tc(weight, act, out, B, out_channels, pixels, in_channels, bias, relu,
                                   (cudaStream_t)stream, 1);
  if (rc == 1) return raftcorr::fail(RAFTCORR_E_UNSUPPORTED, "conv1x1_nchw_fwd: shape not covered");
  return rc;
}

int raftcorr_lookup_conv1x1_fwd(const raftcorr_level* levels, int num_levels, const float* coords,
                                const float* weight, const float* bias, float* out, int B, int H, int W,
                                int radius, int coord_mode, int out_channels, int relu, void* workspace,
                                int64_t workspace_bytes, void* stream) {
  RC_REQUIRE(coords && weight && out, "lookup_conv1x1_fwd: coords/weight/out is NULL");
  RC_REQUIRE(B >= 0 && H >= 1 && W >= 1, "lookup_conv1x1_fwd: bad sizes B=%d H=%d W=%d", B, H, W);
  RC_REQUIRE(radius >= 0 && radius <= 4, "lookup_conv1x1_fwd: radius=%d not in [0,4]", radius);
  RC_REQUIRE(valid_mode(coord_mode), "lookup_conv1x1_fwd: coord_mode=%d", coord_mode);
  RC_REQUIRE(out_channels >= 1, "lookup_conv1x1_fwd: out_channels=%d", out_channels);
  if (int rc = check_levels(levels, num_levels, H, W, "lookup_conv1x1_fwd")) return rc;
  const int rd = 2 * radius + 1, C = num_levels * rd * rd;
  const long long N = (long long)H * W;
  if ((C & 3) || (N & 3) || !tc_gemm_available() || (reinterpret_cast<uintptr_t>(weight) & 15) ||
      (reinterpret_cast<uintptr_t>(out) & 15))
    return raftcorr::fail(RAFTCORR_E_UNSUPPORTED,
                          "lookup_conv1x1_fwd: needs L*(2r+1)^2 %% 4 == 0 and H*W %% 4 == 0 (C=%d, H*W=%lld) and "
                          "16-byte aligned weight/out", C, N);
  const int64_t need = raftcorr_lookup_conv1x1_workspace_bytes(B, H, W, num_levels, radius);
  if (!workspace || workspace_bytes < need || (reinterpret_cast<uintptr_t>(workspace) & 15))
    return raftcorr::fail(RAFTCORR_E_WORKSPACE, "lookup_conv1x1_fwd: workspace %lld B < %lld B (or not 16-byte aligned)",
                          (long long)workspace_bytes, (long long)need);
  float* samples = static_cast<float*>(workspace);   // (B, H, W, C): every query's C samples contiguous
  if (int rc = launch_lookup_fwd(levels, num_levels, coords, samples, B, H, W, radius, coord_mode, 1,
                                 (cudaStream_t)stream))
    return rc;
  return raftcorr_conv1x1_nhwc_fwd(samples, weight, bias, out, B, (int)N, C, out_channels, relu, stream);
}

}  // extern "C"
