"""Drop-in replacement for the reference's ``core/corr.py`` (zhaoyuzhi/PyTorch-RAFT).

Put this directory ahead of the reference's ``core/`` on ``sys.path`` and ``core/raft.py`` runs
unchanged: it does ``from corr import CorrBlock, AlternateCorrBlock`` (raft.py:8), constructs
``CorrBlock(fmap1, fmap2, radius=r)`` once per frame pair (raft.py:104-107) and calls
``corr_fn(coords1)`` once per refinement iteration (raft.py:124).

Same names, arguments, shapes and error behaviour as the reference; all arithmetic runs in
libraftcorr_b200.so (hand-written sm_100a kernels).  There is no CPU or ATen fallback: tensors
must live on a CUDA device and the extension must be built, otherwise the call raises.

Extras (keyword-only, all optional): ``precision`` ('bf16' default | 'tf32' | 'fp32') selects the
arithmetic of the all-pairs contraction, ``coord_mode`` ('aten_cpu' default | 'aten_cuda' |
'direct') selects which evaluation of the reference's coordinate round trip is replayed
(DESIGN.md).  Environment variables RAFTCORR_PRECISION / RAFTCORR_COORD_MODE set the defaults.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from raftcorr_b200 import ops


class CorrBlock:
    """core/corr.py:12-60.  ``corr_pyramid`` keeps the reference's shapes
    ``(B*H*W, 1, H_l, W_l)`` (strided views: rows are padded to 16 bytes for TMA)."""

    def __init__(self, fmap1, fmap2, num_levels=4, radius=4, *, precision=None, coord_mode=None,
                 channels_last=False, storage=None):
        self.num_levels = num_levels
        self.radius = radius
        self._coord_mode = ops.resolve_coord_mode(coord_mode)
        # channels_last=True returns the lookup in torch.channels_last memory format (same shape and
        # values; cuDNN runs the consuming 1x1 convolution in NHWC anyway).  Inference path only.
        self._channels_last = bool(channels_last)
        self._holder = {}
        self._token = None
        self._views = None
        if torch.is_grad_enabled() and (fmap1.requires_grad or fmap2.requires_grad):
            # gradients flow through the fp32 pyramid (the format core/corr.py itself stores)
            self._token, *views = ops.BuildFn.apply(fmap1, fmap2, num_levels,
                                                    ops.resolve_precision(precision), self._holder)
            self._views = list(views)
        else:
            # inference: with the (default) bf16 arithmetic the pyramid is also STORED in bf16, in 64-byte blocks
            # (storage='auto'; 'fp32' keeps the reference's format) - half the bytes to write and to look up
            if self._channels_last and storage in (None, "auto"):
                storage = "fp32"
            self._holder["pyr"] = ops.build_pyramid(fmap1, fmap2, num_levels, precision, storage=storage,
                                                    radius=radius)

    @property
    def corr_pyramid(self):
        """core/corr.py:21-27's public list of (B*H*W, 1, H_l, W_l) float32 levels.  Views of the fp32 pyramid;
        for the 16-bit pyramid a de-blocked, widened copy made on first access (RAFT itself never reads it)."""
        if self._views is None:
            self._views = self._holder["pyr"].views()
        return self._views

    def __call__(self, coords):
        needs_graph = torch.is_grad_enabled() and (
            coords.requires_grad or (self._token is not None and self._token.requires_grad))
        if needs_graph:
            token = self._token
            if token is None:
                token = torch.zeros(1, dtype=torch.float32, device=coords.device)
            return ops.LookupFn.apply(coords, token, self._holder, self.radius, self._coord_mode)
        return ops.lookup(self._holder["pyr"], coords, self.radius, self._coord_mode, self._channels_last)

    def lookup_conv1x1(self, coords, weight, bias=None, relu=True):
        """Extra: ``relu(conv1x1(self(coords)))`` — the lookup followed by the first layer of the
        motion encoder (``convc1``, core/update.py:83,90) behind one call.  Two launches: the samples are written
        once, query-major ((B, H, W, 324) in a scratch buffer), and a tcgen05 GEMM with bias + ReLU in its epilogue
        reads them back; the NCHW (B, 324, H, W) tensor and the separate bias / ReLU passes are what is saved.
        ``weight`` is the convolution's (out_channels, 324, 1, 1) weight.  TF32 products; inference only."""
        if torch.is_grad_enabled() and (coords.requires_grad or weight.requires_grad or
                                        (self._token is not None and self._token.requires_grad)):
            raise RuntimeError("lookup_conv1x1 is an inference path: call the block and the convolution separately "
                               "when gradients are needed")
        if not isinstance(self._holder["pyr"], ops.Pyramid):
            raise RuntimeError("lookup_conv1x1 reads the fp32 pyramid: construct the block with storage='fp32'")
        return ops.lookup_conv1x1(self._holder["pyr"], coords, self.radius, weight, bias, relu, self._coord_mode)

    @staticmethod
    def corr(fmap1, fmap2, *, precision=None):
        """core/corr.py:52-60: the all-pairs volume (B,H,W,1,H,W), scaled by 1/sqrt(D)."""
        B, D, H, W = fmap1.shape
        pyr = ops.build_pyramid(fmap1, fmap2, 1, precision, storage="fp32")
        return pyr.views()[0].reshape(B, H, W, 1, H, W)


def conv1x1_relu(lookup, weight, bias=None):
    """``relu(conv2d(lookup, weight, bias))`` for the 1x1 convolution that consumes the lookup
    (BasicMotionEncoder.convc1, core/update.py:83,90) on the tensor cores, reading a lookup emitted with
    ``CorrBlock(..., channels_last=True)`` in place (any other layout is converted first).  Inference only."""
    if torch.is_grad_enabled() and (lookup.requires_grad or weight.requires_grad):
        raise RuntimeError("conv1x1_relu is an inference path")
    return ops.conv1x1_channels_last(lookup, weight, bias, relu=True)


class AlternateCorrBlock:
    """core/corr.py:63-91: memory-lean path that never materialises the volume.

    ``pyramid`` keeps the reference's list of pooled (fmap1, fmap2) pairs (num_levels+1 entries,
    corr.py:68-72).  The NHWC copies the reference re-creates on every call (corr.py:82-83) are
    made once here.

    ``precision='bf16'`` (default without autograd): one fused tensor-core launch per call — the
    windows of a tile of 128 neighbouring queries overlap, so their union is a small dense GEMM
    against bf16 NHWC copies of fmap1 and the pooled fmap2 pyramid (csrc/alt_tc.cu); the 1/sqrt(D) of
    corr.py:91 and the /2^l of corr.py:85 are folded in; with autograd the backward runs the fp32
    per-level kernels.  ``precision='fp32'``: one fp32 CUDA-core launch per level (csrc/alt.cu),
    differentiable.

    ``coord_mode`` defaults to CorrBlock's coordinate semantics so both blocks produce identical
    zero-padding masks; pass 'direct' for the reference CUDA kernel's floor(coords) behaviour
    (correlation_kernel.cu:66-76)."""

    def __init__(self, fmap1, fmap2, num_levels=4, radius=4, *, coord_mode=None, precision=None):
        self.num_levels = num_levels
        self.radius = radius
        self._coord_mode = ops.resolve_coord_mode(coord_mode)
        if precision not in (None, "bf16", "fp32"):
            raise ValueError("AlternateCorrBlock precision must be 'bf16' or 'fp32'")
        self._precision = precision or "bf16"
        self._fmaps = (fmap1, fmap2)
        self._pyramid = None          # the reference's list of pooled NCHW pairs, made when somebody reads it
        self._nhwc = None             # fp32 NHWC copies (fp32 kernels, autograd)
        self._f1_bf16 = self._f2_bf16 = None
        dim = fmap1.shape[1]
        if self._precision == "bf16" and dim % 8 == 0 and dim <= 256 and radius in (3, 4):
            # one launch per fmap: pooled, channels-last, bfloat16 (csrc/alt_pack.cu)
            self._f1_bf16 = ops.alt_pack(fmap1.detach(), 1)[0]
            self._f2_bf16 = ops.alt_pack(fmap2.detach(), num_levels)

    @property
    def pyramid(self):
        """core/corr.py:68-72: [(fmap1, fmap2)] followed by num_levels 2x2 mean-poolings of both."""
        if self._pyramid is None:
            fmap1, fmap2 = self._fmaps
            self._pyramid = [(fmap1, fmap2)]
            for _ in range(self.num_levels):
                fmap1 = F.avg_pool2d(fmap1, 2, stride=2)
                fmap2 = F.avg_pool2d(fmap2, 2, stride=2)
                self._pyramid.append((fmap1, fmap2))
        return self._pyramid

    def _fp32_nhwc(self):
        """fmap1 and the pooled fmap2 levels channels-last in fp32 (the copies core/corr.py:82-83 makes per call)."""
        if self._nhwc is None:
            pyr = self.pyramid
            self._nhwc = (pyr[0][0].permute(0, 2, 3, 1).contiguous(),
                          [pyr[i][1].permute(0, 2, 3, 1).contiguous() for i in range(self.num_levels)])
        return self._nhwc

    def __call__(self, coords):
        dim = self._fmaps[0].shape[1]
        scale = 1.0 / math.sqrt(dim)
        graph = torch.is_grad_enabled() and (coords.requires_grad or self._fmaps[0].requires_grad
                                             or self._fmaps[1].requires_grad)
        if self._f1_bf16 is not None and not graph:
            return ops.alt_fused_forward(self._f1_bf16, self._f2_bf16, coords, self.radius, self._coord_mode, scale)
        f1, f2 = self._fp32_nhwc()
        if self._f1_bf16 is not None:
            holder = {"f1_bf16": self._f1_bf16, "f2_bf16": self._f2_bf16}
            return ops.AltFusedFn.apply(coords, f1, holder, self.radius, self._coord_mode, scale, *f2)
        coords = coords.permute(0, 2, 3, 1)
        B, H, W, _ = coords.shape
        corr_list = []
        for i in range(self.num_levels):
            coords_i = (coords / 2 ** i).reshape(B, 1, H, W, 2).contiguous()
            if graph:
                corr = ops.AltFn.apply(f1, f2[i], coords_i, self.radius, self._coord_mode, scale)
            else:
                corr = ops.alt_forward(f1, f2[i], coords_i, self.radius, self._coord_mode, scale)
            corr_list.append(corr.squeeze(1))
        corr = torch.stack(corr_list, dim=1)
        return corr.reshape(B, -1, H, W)
