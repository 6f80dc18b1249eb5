/*
 * raftcorr.h — C ABI of libraftcorr_b200.so: RAFT's correlation hot path on B200 (sm_100a).
 *
 * The reference (zhaoyuzhi/PyTorch-RAFT) reaches this path through Python only
 * (core/corr.py) and through one pybind11 extension (alt_cuda_corr/correlation.cpp:51-54).
 * This header is the boundary a binding would use instead: plain pointers, sizes and a
 * cudaStream_t, no torch types.  pytorch-raft_b200/raftcorr_b200/_capi.py binds it with ctypes;
 * INTEGRATION.md shows the binding a maintainer of the reference would add.
 *
 * Conventions
 *  - every function returns 0 on success, a negative RAFTCORR_E_* for a rejected argument,
 *    or a positive cudaError_t; raftcorr_last_error() gives the text (thread-local).
 *  - no function allocates device memory, synchronises the device or changes the current
 *    device; kernels are enqueued on `stream` (a cudaStream_t passed as void*; NULL = the
 *    legacy default stream).  All pointers are device pointers on the current device unless
 *    the name says host.
 *  - re-entrant (nn.DataParallel calls it from one thread per GPU): no per-call state is kept.  What the
 *    library does keep is caches behind mutexes - encoded TMA tensor maps (a small LRU keyed on pointers,
 *    sizes and strides), per-device SM counts / kernel attributes - and, for the one entry point that says so
 *    (raftcorr_build_pyramid_noalloc), a fixed scratch array shared by all its calls on a device.
 */
#ifndef RAFTCORR_H_
#define RAFTCORR_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RAFTCORR_MAX_LEVELS 4

#define RAFTCORR_E_INVALID (-1)     /* bad pointer / size / enum                     */
#define RAFTCORR_E_UNSUPPORTED (-2) /* valid request this build cannot serve         */
#define RAFTCORR_E_WORKSPACE (-3)   /* workspace too small                           */
#define RAFTCORR_E_DRIVER (-4)      /* CUDA driver entry point (tensor maps) missing */

/* How a pixel coordinate becomes a sampling index (DESIGN.md "coordinate semantics").
 * core/utils/utils.py:61-62 normalises to [-1,1] and ATen's grid_sampler un-normalises again;
 * that fp32 round trip moves floor() on ~8 % of integer-grid taps, so it is replayed exactly. */
#define RAFTCORR_COORD_ATEN_CPU 0  /* 2x/(W-1): true division, as ATen's CPU kernels do it      */
#define RAFTCORR_COORD_ATEN_CUDA 1 /* 2x*(1/(W-1)): as ATen's CUDA div-by-scalar kernel does it */
#define RAFTCORR_COORD_DIRECT 2    /* floor(x) directly: alt_cuda_corr/correlation_kernel.cu:66-76 */

/* Arithmetic of the all-pairs contraction (core/corr.py:58 is an fp32 cuBLAS SGEMM). */
#define RAFTCORR_PREC_FP32 0 /* fp32-accurate: <=1e-5 of max|ref| (north_star "fp32 mode")   */
#define RAFTCORR_PREC_BF16 1 /* one bf16 tcgen05 pass, fp32 accumulate: <=1e-2 of max|ref|   */
#define RAFTCORR_PREC_TF32 2 /* one tf32 tcgen05 pass, fp32 accumulate: <=1e-2 of max|ref|   */

/* One pyramid level: `images` 2-D fp32 images of H x W, row pitch and image stride in elements.
 * This is core/corr.py:21-27's corr_pyramid[l] (shape (B*H1*W1, 1, H, W)) with explicit
 * strides, so that rows can be padded to 16 bytes for TMA. */
typedef struct raftcorr_level {
  float* data;
  int32_t H, W;
  int64_t pitch;        /* elements between rows   (>= W)        */
  int64_t image_stride; /* elements between images (>= H*pitch)  */
} raftcorr_level;

int raftcorr_version(void);
const char* raftcorr_last_error(void);

/* Level sizes of core/corr.py:25-27 (avg_pool2d floors) and the padded layout this library
 * prefers: pitch rounded up to 4 floats.  Fills H/W/pitch/image_stride of levels[0..L) (data
 * untouched) and returns the number of floats each level needs in floats_per_level[]. */
int raftcorr_pyramid_layout(int B, int H, int W, int num_levels, raftcorr_level* levels,
                            int64_t* floats_per_level);

/* ---- (1) all-pairs volume + pyramid: replaces CorrBlock.__init__ / CorrBlock.corr
 *      (core/corr.py:13-27, 52-60).  fmap1/fmap2: (B,D,H,W) fp32 NCHW contiguous.
 *      levels[l] for l<L receive mean-pooled fmap1^T.fmap2/sqrt(D); image index = b*H*W + y*W + x
 *      of the fmap1 pixel.  workspace: raftcorr_build_workspace_bytes() bytes, 256-byte aligned. */
int64_t raftcorr_build_workspace_bytes(int B, int D, int H, int W, int precision);
int raftcorr_build_pyramid(const float* fmap1, const float* fmap2, int B, int D, int H, int W,
                           int num_levels, const raftcorr_level* levels, int precision,
                           void* workspace, int64_t workspace_bytes, void* stream);

/* Same result as raftcorr_build_pyramid at RAFTCORR_PREC_BF16 / RAFTCORR_PREC_TF32 (including its fp32
 * CUDA-core build, which needs no scratch, when D % 8 != 0), for callers that
 * cannot hand over a workspace.  This entry point has no counterpart in the reference.
 * Who needs it: a host that claims the whole device for its own pool at start-up and passes a
 * plug-in only input/output pointers (the binding has no way to ask that pool for more), or a
 * real-time loop that bans cudaMalloc/cudaFree because they stall the device.  The packed
 * operands (2 * B*H*W*D elements of 2 or 4 bytes, each half rounded up to 256 B) go into a fixed
 * RAFTCORR_NOALLOC_SCRATCH_BYTES array inside the library image: every GPU that loads the library
 * holds that memory until the process ends.  Larger problems are refused with
 * RAFTCORR_E_WORKSPACE.  It calls no cudaMalloc/cudaFree and no allocating library routine.
 *  - The first call on each device looks the array's device address up; make that first call
 *    before starting any stream capture.
 *  - All calls on a device share the one array.  Each call fully rewrites the part it reads, so
 *    stale contents are harmless, but two calls whose kernels overlap on the GPU (different
 *    streams, or different host threads on the same device) race on it: serialise them. */
#define RAFTCORR_NOALLOC_SCRATCH_BYTES (8u << 20)
int raftcorr_build_pyramid_noalloc(const float* fmap1, const float* fmap2, int B, int D, int H, int W,
                                   int num_levels, const raftcorr_level* levels, int precision,
                                   void* stream);

/* ---- (1b) the same pyramid in 16-bit storage: core/corr.py:21-27 stores fp32, which makes
 *      the build write 5.3 B per 512 flop (2 GB at configs[1]) and the lookup read 4-byte pixels out of 64-byte
 *      DRAM granules.  Here every level holds bfloat16 pixels in blocks of 4 rows x 8 columns = 64 bytes (one
 *      DRAM granule; a 10x10 window touches ~7 of them instead of ~13 row pieces):
 *        pixel (y, x) of image n lives at byte  n*image_stride + ((y>>2)*blocks_x + (x>>3))*64 + (y&3)*16 + (x&7)*2
 *      with blocks_x = ceil(W/8), blocks_y = ceil(H/4), image_stride = blocks_x*blocks_y*64.
 *      A level may keep the blocks of G = 2^interleave_log2 neighbouring images side by side (image n = b*H*W +
 *      y*W + x, so neighbours are neighbouring queries of a row, whose windows nearly coincide: the lookup fetches
 *      them as one box, and DRAM then sees pieces of G*128..192 bytes instead of G pieces 14 KB apart):
 *        byte  (n / G) * G*image_stride + (((y>>2)*blocks_x + (x>>3)) * G + n % G) * 64 + (y&3)*16 + (x&7)*2
 *      raftcorr_pyramid16_layout chooses it (levels 0 and 1, when W is a multiple of G); the lookup takes
 *      either form per level, the build takes levels 0 and 1 both plain or both with interleave_log2 = 2 and
 *      the other levels plain.  Pixels of a block that lie outside the H x W image are stored as +0.0 (the lookup relies on it: they are grid_sample's zero
 *      padding).  Values: bf16 tcgen05 products, fp32 accumulation and pooling, one rounding to bf16 per stored
 *      pixel: <= 1e-2 of max|ref| like RAFTCORR_PREC_BF16.  Inference only (gradients use the fp32 pyramid). */
typedef struct raftcorr_level16 {
  void* data;           /* bfloat16, 64-byte aligned                    */
  int32_t H, W;         /* pixels                                       */
  int32_t blocks_x, blocks_y;
  int32_t interleave_log2; /* g: the blocks of 2^g neighbouring images are interleaved (see above); 0 = not */
  int32_t reserved;        /* 0 */
  int64_t image_stride; /* bytes of one image = blocks_x*blocks_y*64   */
} raftcorr_level16;
/* Fills H/W/blocks_x/blocks_y/image_stride of levels[0..L) (data untouched); bytes_per_level[l] = B*H*W images. */
int raftcorr_pyramid16_layout(int B, int H, int W, int num_levels, raftcorr_level16* levels,
                              int64_t* bytes_per_level);
/* Build (replaces core/corr.py:13-27, 52-60).  Needs D % 8 == 0 (else RAFTCORR_E_UNSUPPORTED: use the fp32
 * pyramid).  workspace: raftcorr_build_workspace_bytes(B, D, H, W, RAFTCORR_PREC_BF16) bytes, 256-byte aligned. */
int raftcorr_build_pyramid16(const float* fmap1, const float* fmap2, int B, int D, int H, int W,
                             int num_levels, const raftcorr_level16* levels, void* workspace,
                             int64_t workspace_bytes, void* stream);
/* Lookup on it (replaces core/corr.py:29-50; same semantics, layouts and channel order as raftcorr_lookup_fwd;
 * bit-identical to the reference's arithmetic applied to the stored bf16 values).  num_levels == 4, radius 3 or 4. */
int raftcorr_lookup16_fwd(const raftcorr_level16* levels, int num_levels, const float* coords, float* out,
                          int B, int H, int W, int radius, int coord_mode, void* stream);

/* ---- (2) 4-level (2r+1)^2 bilinear window lookup: replaces CorrBlock.__call__
 *      (core/corr.py:29-50) including bilinear_sampler (core/utils/utils.py:57-65) and
 *      grid_sample(align_corners=True, zero padding).
 *      coords: (B,2,H,W) fp32, ch0=x ch1=y.  out: (B, L*(2r+1)^2, H, W) fp32 contiguous,
 *      channel = l*(2r+1)^2 + i*(2r+1) + j, i moving x and j moving y. */
int raftcorr_lookup_fwd(const raftcorr_level* levels, int num_levels, const float* coords,
                        float* out, int B, int H, int W, int radius, int coord_mode, void* stream);

/* Same samples, channels-last: out is (B, H, W, L*(2r+1)^2), i.e. a (B, C, H, W) tensor in PyTorch's
 * channels_last memory format.  The consumer of the lookup, BasicMotionEncoder.convc1 (a 1x1 convolution,
 * core/update.py:82-90), is run by cuDNN in NHWC anyway; writing a query's C samples contiguously also
 * spares the kernel its shared-memory transpose.  Values are identical to
 * raftcorr_lookup_fwd. */
int raftcorr_lookup_fwd_nhwc(const raftcorr_level* levels, int num_levels, const float* coords,
                             float* out, int B, int H, int W, int radius, int coord_mode, void* stream);

/* Backward of (2) (the reference gets it from autograd: GridSampler2DBackward0, SURVEY.md §8
 * a-10).  grad_levels[l] must be zero-filled by the caller and are accumulated into with
 * atomics (several lookups may accumulate into the same buffers).  grad_coords (B,2,H,W) may be
 * NULL (RAFT detaches coords, core/raft.py:123); when given it is overwritten. */
int raftcorr_lookup_bwd(const raftcorr_level* levels, const raftcorr_level* grad_levels,
                        int num_levels, const float* coords, const float* grad_out,
                        float* grad_coords, int B, int H, int W, int radius, int coord_mode,
                        void* stream);

/* Backward of (1): grad_fmap1/2 (B,D,H,W) from the accumulated d(pyramid)
 * (AvgPool2DBackward0 x3, DivBackward0, BmmBackward0 of core/corr.py:21-27,52-60).
 * workspace: raftcorr_build_bwd_workspace_bytes() bytes. */
int64_t raftcorr_build_bwd_workspace_bytes(int B, int D, int H, int W);
int raftcorr_build_bwd(const raftcorr_level* grad_levels, int num_levels, const float* fmap1,
                       const float* fmap2, float* grad_fmap1, float* grad_fmap2, int B, int D,
                       int H, int W, void* workspace, int64_t workspace_bytes, void* stream);

/* ---- (3) on-the-fly correlation: replaces alt_cuda_corr.forward / .backward
 *      (alt_cuda_corr/correlation.cpp:23-48, correlation_kernel.cu:18-324).
 *      fmap1: (B,H1,W1,C) NHWC, fmap2: (B,H2,W2,C) NHWC, coords: (B,1,H1,W1,2) last dim (x,y),
 *      corr: (B,1,(2r+1)^2,H1,W1), multiplied by `scale` (the reference kernel is unscaled:
 *      scale=1; core/corr.py:91 divides by sqrt(C) afterwards). */
int raftcorr_alt_fwd(const float* fmap1, const float* fmap2, const float* coords, float* corr,
                     int B, int H1, int W1, int H2, int W2, int C, int radius, int coord_mode,
                     float scale, void* stream);
/* The same path fused over all levels on the tensor cores: replaces one whole
 * AlternateCorrBlock.__call__ (core/corr.py:74-91: num_levels launches + NHWC copies + stack + divide).
 *      fmap1_bf16: (B,H1,W1,C) NHWC bfloat16; fmap2_bf16[l]: (B,H1>>l,W1>>l,C) NHWC bfloat16, the 2x2
 *      mean-pooled fmap2 pyramid of core/corr.py:68-72; coords: (B,2,H1,W1) fp32 in level-0 pixels
 *      (the kernel applies the /2^l of corr.py:85); out: (B, L*(2r+1)^2, H1, W1) fp32, multiplied by
 *      `scale` (1/sqrt(C) reproduces corr.py:91).  bf16 products, fp32 accumulation: <=1e-2 of max|ref|. */
int raftcorr_alt_fused_fwd(const void* fmap1_bf16, const void* const* fmap2_bf16, const float* coords,
                           float* out, int B, int H1, int W1, int C, int num_levels, int radius,
                           int coord_mode, float scale, void* stream);

/* The operands of raftcorr_alt_fused_fwd in one pass: replaces the avg_pool2d chain of AlternateCorrBlock.__init__
 * (core/corr.py:64-72) and the per-call NHWC copies of core/corr.py:82-83.
 *      fmap: (B,D,H,W) fp32 NCHW contiguous; out_levels[l], l < num_levels: (B, H>>l, W>>l, D) NHWC bfloat16, level
 *      l+1 = the 2x2 mean of level l in fp32 (ATen's avg_pool2d arithmetic, floor on odd sizes), each value rounded
 *      to bfloat16 once.  num_levels = 1 for fmap1, 4 for fmap2.  D must be even. */
int raftcorr_alt_pack(const float* fmap, int B, int D, int H, int W, int num_levels, void* const* out_levels,
                      void* stream);

/* fmap1_grad/fmap2_grad must be zero-filled by the caller (accumulated with atomics);
 * coords_grad may be NULL.  The reference returns coords_grad as all zeros
 * (correlation_kernel.cu:130,307,323); this library fills it (documented superset). */
int raftcorr_alt_bwd(const float* fmap1, const float* fmap2, const float* coords,
                     const float* corr_grad, float* fmap1_grad, float* fmap2_grad,
                     float* coords_grad, int B, int H1, int W1, int H2, int W2, int C, int radius,
                     int coord_mode, float scale, void* stream);

/* ---- (4) the step after the path: convex 8x upsampling, replaces RAFT.upsample_flow
 *      (core/raft.py:72-83: view + softmax + unfold + mul + sum + permute + reshape) in one pass.
 *      flow: (B,2,H,W) fp32; mask: (B,576,H,W) fp32, channel = k*64 + dy*8 + dx with k = 3*ky+kx the 3x3
 *      neighbour (zero padded) and (dy,dx) the sub-pixel; up: (B,2,8H,8W) fp32.
 *      up[b,c,8y+dy,8x+dx] = sum_k softmax_k(mask[b,k,dy,dx,y,x]) * 8*flow[b,c,y+ky-1,x+kx-1]. */
int raftcorr_upsample_flow_fwd(const float* flow, const float* mask, float* up, int B, int H, int W,
                               void* stream);
/* Backward of the above (the training loss reads every iteration's upsampled flow, train.py:47-72).
 * grad_flow (B,2,H,W) is overwritten (zeroed on `stream`, then accumulated with atomics); grad_mask
 * (B,576,H,W) is overwritten; either may be NULL. */
int raftcorr_upsample_flow_bwd(const float* flow, const float* mask, const float* grad_up, float* grad_flow,
                               float* grad_mask, int B, int H, int W, void* stream);
/* 8 * bilinear x8 interpolation with align_corners=True: replaces upflow8, core/utils/utils.py:80-82
 * (the small model's upsampling, raft.py:134-135).  flow (B,2,H,W) -> up (B,2,8H,8W). */
int raftcorr_upflow8(const float* flow, float* up, int B, int H, int W, void* stream);

/* ---- (5) lookup + the 1x1 convolution that consumes it: replaces
 *      CorrBlock.__call__ (core/corr.py:29-50) followed by BasicMotionEncoder.convc1 + ReLU
 *      (core/update.py:83,90; SmallMotionEncoder: update.py:66,72).  The samples are written once,
 *      query-major, into `workspace` and consumed by a tf32 tcgen05 GEMM whose epilogue adds the bias,
 *      applies the ReLU and stores NCHW — no 324-channel NCHW tensor, no layout shuffles, no separate
 *      bias / ReLU passes.  weight: (out_channels, L*(2r+1)^2) row-major = the convolution's
 *      (out_channels, C, 1, 1) weight; bias: (out_channels) or NULL; out: (B, out_channels, H, W).
 *      TF32 products (what cuDNN uses for this convolution by default), fp32 accumulation.
 *      Needs C % 4 == 0 and H*W % 4 == 0, else RAFTCORR_E_UNSUPPORTED. */
int64_t raftcorr_lookup_conv1x1_workspace_bytes(int B, int H, int W, int num_levels, int radius);
/* The convolution half on its own, for a caller that already holds the channels-last lookup
 * (raftcorr_lookup_fwd_nhwc): act (B, pixels, in_channels) -> out (B, out_channels, pixels),
 * out = act . weight^T + bias, optional ReLU. */
int raftcorr_conv1x1_nhwc_fwd(const float* act, const float* weight, const float* bias, float* out, int B,
                              int pixels, int in_channels, int out_channels, int relu, void* stream);
/* The same product on a plain NCHW activation tensor (B, in_channels, pixels) — what raftcorr_lookup_fwd and
 * raftcorr_lookup16_fwd write: the pixel axis is the contiguous one, the tensor cores take it as an MN-major operand.
 * out_channels <= 256. */
int raftcorr_conv1x1_nchw_fwd(const float* act, const float* weight, const float* bias, float* out, int B,
                              int pixels, int in_channels, int out_channels, int relu, void* stream);
int raftcorr_lookup_conv1x1_fwd(const raftcorr_level* levels, int num_levels, const float* coords,
                                const float* weight, const float* bias, float* out, int B, int H, int W,
                                int radius, int coord_mode, int out_channels, int relu, void* workspace,
                                int64_t workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RAFTCORR_H_ */
