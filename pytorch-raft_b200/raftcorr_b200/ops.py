"""torch-facing wrappers over the C ABI: tensors in, tensors out, current device + stream.

PyTorch is plumbing here (device memory, streams, autograd bookkeeping); all arithmetic of the
correlation path runs in libraftcorr_b200.so.  Nothing in this file computes on the CPU or
falls back to ATen ops.
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional, Sequence

import torch

from . import _capi
from ._capi import (COORD_ATEN_CPU, COORD_ATEN_CUDA, COORD_DIRECT, PREC_BF16, PREC_FP32, PREC_TF32,
                    Level, Level16Array, LevelArray)

_PRECISIONS = {"fp32": PREC_FP32, "bf16": PREC_BF16, "tf32": PREC_TF32}
_COORD_MODES = {"aten_cpu": COORD_ATEN_CPU, "aten_cuda": COORD_ATEN_CUDA, "direct": COORD_DIRECT}


def resolve_precision(precision) -> int:
    """'bf16' (default; one tcgen05 pass, 5e-4 px EPE drift on RAFT), 'tf32' or 'fp32'."""
    if precision is None:
        precision = os.environ.get("RAFTCORR_PRECISION", "bf16")
    if isinstance(precision, str):
        if precision not in _PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(_PRECISIONS)}, got {precision!r}")
        return _PRECISIONS[precision]
    return int(precision)


def resolve_coord_mode(mode) -> int:
    if mode is None:
        mode = os.environ.get("RAFTCORR_COORD_MODE", "aten_cpu")
    if isinstance(mode, str):
        if mode not in _COORD_MODES:
            raise ValueError(f"coord_mode must be one of {sorted(_COORD_MODES)}, got {mode!r}")
        return _COORD_MODES[mode]
    return int(mode)


def _stream(t: torch.Tensor) -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


class _on_device:
    """`with torch.cuda.device(d)` only when d is not already current (the context manager
    costs more host time than a lookup kernel runs)."""

    def __init__(self, device):
        self.ctx = None if device.index == torch.cuda.current_device() else torch.cuda.device(device)

    def __enter__(self):
        if self.ctx is not None:
            self.ctx.__enter__()

    def __exit__(self, *exc):
        if self.ctx is not None:
            self.ctx.__exit__(*exc)


def _require_cuda_f32(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32, got {t.dtype}")


class Pyramid:
    """Device storage of a correlation pyramid (core/corr.py:21-27's ``corr_pyramid``) with rows
    padded to 16 bytes, plus the raftcorr_level descriptors that describe it to the kernels."""

    def __init__(self, B: int, H: int, W: int, num_levels: int, device, zero: bool = False):
        self.B, self.H, self.W, self.num_levels = B, H, W, num_levels
        self.desc = LevelArray()
        floats = (ctypes.c_int64 * _capi.MAX_LEVELS)()
        _capi.check(_capi.lib().raftcorr_pyramid_layout(B, H, W, num_levels, self.desc, floats),
                    "raftcorr_pyramid_layout")
        alloc = torch.zeros if zero else torch.empty
        self.storage: List[torch.Tensor] = []
        for l in range(num_levels):
            d = self.desc[l]
            buf = alloc((B * H * W, d.H, d.pitch), dtype=torch.float32, device=device)
            d.data = buf.data_ptr()
            self.storage.append(buf)

    def views(self) -> List[torch.Tensor]:
        """(B*H*W, 1, H_l, W_l) strided views, the shape core/corr.py:22-27 exposes."""
        return [self.storage[l][:, :, : self.desc[l].W].unsqueeze(1) for l in range(self.num_levels)]


class Pyramid16:
    """Device storage of a correlation pyramid in 16 bits (include/raftcorr.h, raftcorr_level16): bfloat16
    pixels in blocks of 4 rows x 8 columns, ``storage[l]`` indexed
    [group of G images][block row][block col][image in group][row][col] with G = 2**interleave_log2 (1 = plain).
    Half the bytes of ``Pyramid`` for the build to write and the lookup to read; inference only."""

    def __init__(self, B: int, H: int, W: int, num_levels: int, device, zero: bool = False, interleave=None):
        """interleave: None = the library's choice (raftcorr_pyramid16_layout), False = plain blocked levels,
        or interleave_log2 per level."""
        self.B, self.H, self.W, self.num_levels = B, H, W, num_levels
        self.desc = Level16Array()
        nbytes = (ctypes.c_int64 * _capi.MAX_LEVELS)()
        _capi.check(_capi.lib().raftcorr_pyramid16_layout(B, H, W, num_levels, self.desc, nbytes),
                    "raftcorr_pyramid16_layout")
        if interleave is False:
            interleave = (0,) * num_levels
        if interleave is not None:                      # explicit log2 per level (tests, measurements)
            for l in range(num_levels):
                self.desc[l].interleave_log2 = int(interleave[l])
        alloc = torch.zeros if zero else torch.empty
        self.storage: List[torch.Tensor] = []
        for l in range(num_levels):
            d = self.desc[l]
            g = 1 << d.interleave_log2
            assert (B * H * W) % g == 0
            buf = alloc((B * H * W // g, d.blocks_y, d.blocks_x, g, 4, 8), dtype=torch.bfloat16, device=device)
            d.data = buf.data_ptr()
            self.storage.append(buf)

    def padded(self, l: int) -> torch.Tensor:
        """Level l de-blocked, padding pixels included: (B*H*W, 4*blocks_y, 8*blocks_x) bfloat16 (a copy)."""
        d, s = self.desc[l], self.storage[l]                    # [group][by][bx][g][4][8]
        return s.permute(0, 3, 1, 4, 2, 5).reshape(s.shape[0] * s.shape[3], 4 * d.blocks_y, 8 * d.blocks_x)

    def dense(self, l: int) -> torch.Tensor:
        """Level l de-blocked and widened: (B*H*W, H_l, W_l) float32 (a copy; layout conversion only)."""
        d = self.desc[l]
        return self.padded(l)[:, : d.H, : d.W].float()

    def views(self) -> List[torch.Tensor]:
        """(B*H*W, 1, H_l, W_l) float32 tensors, the shape core/corr.py:22-27 exposes (copies)."""
        return [self.dense(l).unsqueeze(1) for l in range(self.num_levels)]

    @classmethod
    def from_dense(cls, levels: Sequence[torch.Tensor], B: int, H: int, W: int, device="cuda",
                   interleave=None) -> "Pyramid16":
        """Blocked bf16 copy of reference-shaped levels (B*H*W, 1, H_l, W_l) (rounded to bf16, padding zeroed)."""
        pyr = cls(B, H, W, len(levels), device, zero=True, interleave=interleave)
        for l, lv in enumerate(levels):
            d = pyr.desc[l]
            t = torch.as_tensor(lv, device=device).reshape(B * H * W, d.H, d.W).to(torch.bfloat16)
            full = torch.zeros((B * H * W, 4 * d.blocks_y, 8 * d.blocks_x), dtype=torch.bfloat16, device=device)
            full[:, : d.H, : d.W] = t
            g = 1 << d.interleave_log2
            pyr.storage[l].copy_(full.reshape(B * H * W // g, g, d.blocks_y, 4, d.blocks_x, 8).permute(0, 2, 4, 1, 3, 5))
        return pyr


def resolve_storage(storage, precision: int, D: int, num_levels: int, radius) -> str:
    """'fp32' (row-major fp32, core/corr.py's own format, needed for gradients) or 'bf16' (blocked bfloat16).
    'auto': 'bf16' whenever the bf16 arithmetic was asked for and the 16-bit kernels cover the case."""
    if storage is None:
        storage = os.environ.get("RAFTCORR_PYRAMID", "auto")
    if storage not in ("auto", "fp32", "bf16"):
        raise ValueError(f"pyramid storage must be 'auto', 'fp32' or 'bf16', got {storage!r}")
    covered = precision == PREC_BF16 and D % 8 == 0 and num_levels == 4 and radius in (None, 3, 4)
    if storage == "bf16" and not covered:
        raise RuntimeError("16-bit pyramid storage needs precision='bf16', D % 8 == 0, num_levels == 4 and radius 3 or 4")
    if storage == "auto":
        storage = "bf16" if covered else "fp32"
    return storage


def build_pyramid(fmap1: torch.Tensor, fmap2: torch.Tensor, num_levels: int = 4,
                  precision=None, storage="fp32", radius=None, interleave=None):
    """core/corr.py:13-27 + 52-60 in one launch sequence (pack, tcgen05 GEMM + pooled epilogue).
    Returns a ``Pyramid`` (fp32 rows) or, for storage 'bf16' / 'auto', a ``Pyramid16`` (blocked bfloat16)."""
    _require_cuda_f32(fmap1, "fmap1")
    _require_cuda_f32(fmap2, "fmap2")
    if fmap1.shape != fmap2.shape or fmap1.dim() != 4:
        raise RuntimeError(f"fmap1/fmap2 must both be (B,D,H,W); got {tuple(fmap1.shape)} and {tuple(fmap2.shape)}")
    fmap1, fmap2 = fmap1.contiguous(), fmap2.contiguous()
    B, D, H, W = fmap1.shape
    prec = resolve_precision(precision)
    lib = _capi.lib()
    if resolve_storage(storage, prec, D, num_levels, radius) == "bf16":
        with torch.cuda.device(fmap1.device):
            pyr16 = Pyramid16(B, H, W, num_levels, fmap1.device, interleave=interleave)
            ws_bytes = int(lib.raftcorr_build_workspace_bytes(B, D, H, W, PREC_BF16))
            ws = torch.empty(max(ws_bytes, 1), dtype=torch.uint8, device=fmap1.device)
            _capi.check(lib.raftcorr_build_pyramid16(fmap1.data_ptr(), fmap2.data_ptr(), B, D, H, W, num_levels,
                                                     pyr16.desc, ws.data_ptr(), ws_bytes, _stream(fmap1)),
                        "raftcorr_build_pyramid16")
            ws.record_stream(torch.cuda.current_stream(fmap1.device))
        return pyr16
    with torch.cuda.device(fmap1.device):
        pyr = Pyramid(B, H, W, num_levels, fmap1.device)
        ws_bytes = int(lib.raftcorr_build_workspace_bytes(B, D, H, W, prec))
        ws = torch.empty(max(ws_bytes, 1), dtype=torch.uint8, device=fmap1.device)
        _capi.check(lib.raftcorr_build_pyramid(fmap1.data_ptr(), fmap2.data_ptr(), B, D, H, W, num_levels,
                                               pyr.desc, prec, ws.data_ptr(), ws_bytes, _stream(fmap1)),
                    "raftcorr_build_pyramid")
        ws.record_stream(torch.cuda.current_stream(fmap1.device))
    return pyr


def lookup(pyr: Pyramid, coords: torch.Tensor, radius: int, coord_mode=None, channels_last: bool = False) -> torch.Tensor:
    """core/corr.py:29-50: coords (B,2,H,W) -> (B, L*(2r+1)^2, H, W) fp32, contiguous (NCHW) or, with
    channels_last=True, the same logical tensor in torch.channels_last memory format."""
    _require_cuda_f32(coords, "coords")
    B, two, H, W = coords.shape
    if two != 2 or (B, H, W) != (pyr.B, pyr.H, pyr.W):
        raise RuntimeError(f"coords must be ({pyr.B},2,{pyr.H},{pyr.W}); got {tuple(coords.shape)}")
    coords = coords.contiguous()
    rd = 2 * radius + 1
    C = pyr.num_levels * rd * rd
    if isinstance(pyr, Pyramid16):
        if channels_last:
            raise RuntimeError("the channels-last lookup reads the fp32 pyramid: build it with storage='fp32'")
        out = torch.empty((B, C, H, W), dtype=torch.float32, device=coords.device)
        with _on_device(coords.device):
            _capi.check(_capi.lib().raftcorr_lookup16_fwd(pyr.desc, pyr.num_levels, coords.data_ptr(), out.data_ptr(),
                                                          B, H, W, radius, resolve_coord_mode(coord_mode),
                                                          _stream(coords)), "raftcorr_lookup16_fwd")
        return out
    if channels_last:
        out = torch.empty((B, H, W, C), dtype=torch.float32, device=coords.device)
        fn, name = _capi.lib().raftcorr_lookup_fwd_nhwc, "raftcorr_lookup_fwd_nhwc"
    else:
        out = torch.empty((B, C, H, W), dtype=torch.float32, device=coords.device)
        fn, name = _capi.lib().raftcorr_lookup_fwd, "raftcorr_lookup_fwd"
    with _on_device(coords.device):
        _capi.check(fn(pyr.desc, pyr.num_levels, coords.data_ptr(), out.data_ptr(), B, H, W, radius,
                       resolve_coord_mode(coord_mode), _stream(coords)), name)
    return out.permute(0, 3, 1, 2) if channels_last else out


def lookup_backward(pyr: Optional[Pyramid], grad_pyr: Optional[Pyramid], coords: torch.Tensor,
                    grad_out: torch.Tensor, radius: int, want_coords: bool, coord_mode=None):
    """Scatter grad_out into grad_pyr (accumulating) and/or return d coords."""
    coords = coords.contiguous()
    grad_out = grad_out.contiguous()
    B, _, H, W = coords.shape
    gcoords = torch.empty_like(coords) if want_coords else None
    ref = grad_pyr if grad_pyr is not None else pyr
    with torch.cuda.device(coords.device):
        _capi.check(_capi.lib().raftcorr_lookup_bwd(
            pyr.desc if pyr is not None else None, grad_pyr.desc if grad_pyr is not None else None,
            ref.num_levels, coords.data_ptr(), grad_out.data_ptr(),
            gcoords.data_ptr() if want_coords else None, B, H, W, radius,
            resolve_coord_mode(coord_mode), _stream(coords)), "raftcorr_lookup_bwd")
    return gcoords


def build_backward(grad_pyr: Pyramid, fmap1: torch.Tensor, fmap2: torch.Tensor):
    fmap1, fmap2 = fmap1.contiguous(), fmap2.contiguous()
    B, D, H, W = fmap1.shape
    lib = _capi.lib()
    g1, g2 = torch.empty_like(fmap1), torch.empty_like(fmap2)
    with torch.cuda.device(fmap1.device):
        ws_bytes = int(lib.raftcorr_build_bwd_workspace_bytes(B, D, H, W))
        ws = torch.empty(max(ws_bytes, 1), dtype=torch.uint8, device=fmap1.device)
        _capi.check(lib.raftcorr_build_bwd(grad_pyr.desc, grad_pyr.num_levels, fmap1.data_ptr(), fmap2.data_ptr(),
                                           g1.data_ptr(), g2.data_ptr(), B, D, H, W, ws.data_ptr(), ws_bytes,
                                           _stream(fmap1)), "raftcorr_build_bwd")
        ws.record_stream(torch.cuda.current_stream(fmap1.device))
    return g1, g2


def alt_forward(fmap1: torch.Tensor, fmap2: torch.Tensor, coords: torch.Tensor, radius: int,
                coord_mode=COORD_DIRECT, scale: float = 1.0) -> torch.Tensor:
    """alt_cuda_corr.forward layouts: fmap1 (B,H1,W1,C), fmap2 (B,H2,W2,C), coords (B,1,H1,W1,2)."""
    B, H1, W1, C = fmap1.shape
    _, H2, W2, _ = fmap2.shape
    rd = 2 * radius + 1
    corr = torch.empty((B, 1, rd * rd, H1, W1), dtype=torch.float32, device=fmap1.device)
    with torch.cuda.device(fmap1.device):
        _capi.check(_capi.lib().raftcorr_alt_fwd(fmap1.data_ptr(), fmap2.data_ptr(), coords.data_ptr(),
                                                 corr.data_ptr(), B, H1, W1, H2, W2, C, radius,
                                                 resolve_coord_mode(coord_mode), scale, _stream(fmap1)),
                    "raftcorr_alt_fwd")
    return corr


def alt_fused_forward(fmap1_bf16: torch.Tensor, fmap2_bf16: Sequence[torch.Tensor], coords: torch.Tensor,
                      radius: int, coord_mode=None, scale: float = 1.0) -> torch.Tensor:
    """All levels of the on-the-fly path in one tensor-core launch (replaces corr.py:74-91).
    fmap1_bf16 (B,H,W,C) and fmap2_bf16[l] (B,H>>l,W>>l,C) are NHWC bfloat16; coords (B,2,H,W) fp32."""
    B, H, W, C = fmap1_bf16.shape
    L = len(fmap2_bf16)
    coords = coords.contiguous()
    _require_cuda_f32(coords, "coords")
    rd = 2 * radius + 1
    out = torch.empty((B, L * rd * rd, H, W), dtype=torch.float32, device=coords.device)
    ptrs = (ctypes.c_void_p * L)(*[t.data_ptr() for t in fmap2_bf16])
    with _on_device(coords.device):
        _capi.check(_capi.lib().raftcorr_alt_fused_fwd(fmap1_bf16.data_ptr(), ptrs, coords.data_ptr(), out.data_ptr(),
                                                       B, H, W, C, L, radius, resolve_coord_mode(coord_mode), scale,
                                                       _stream(coords)), "raftcorr_alt_fused_fwd")
    return out


def alt_pack(fmap: torch.Tensor, num_levels: int) -> List[torch.Tensor]:
    """(B,D,H,W) fp32 NCHW -> [(B, H>>l, W>>l, D) bfloat16 NHWC for l < num_levels], level l+1 = the 2x2 mean of level l
    in fp32 (core/corr.py:64-72's avg_pool2d chain + the NHWC copies of corr.py:82-83 + the bf16 cast, one launch)."""
    _require_cuda_f32(fmap, "fmap")
    fmap = fmap.contiguous()
    B, D, H, W = fmap.shape
    outs = [torch.empty((B, H >> l, W >> l, D), dtype=torch.bfloat16, device=fmap.device) for l in range(num_levels)]
    ptrs = (ctypes.c_void_p * num_levels)(*[t.data_ptr() for t in outs])
    with _on_device(fmap.device):
        _capi.check(_capi.lib().raftcorr_alt_pack(fmap.data_ptr(), B, D, H, W, num_levels, ptrs, _stream(fmap)),
                    "raftcorr_alt_pack")
    return outs


def alt_backward(fmap1, fmap2, coords, corr_grad, radius: int, coord_mode=COORD_DIRECT,
                 scale: float = 1.0, want_coords: bool = True):
    B, H1, W1, C = fmap1.shape
    _, H2, W2, _ = fmap2.shape
    g1, g2 = torch.zeros_like(fmap1), torch.zeros_like(fmap2)
    gc = torch.empty_like(coords) if want_coords else None
    with torch.cuda.device(fmap1.device):
        _capi.check(_capi.lib().raftcorr_alt_bwd(fmap1.data_ptr(), fmap2.data_ptr(), coords.data_ptr(),
                                                 corr_grad.data_ptr(), g1.data_ptr(), g2.data_ptr(),
                                                 gc.data_ptr() if want_coords else None, B, H1, W1, H2, W2,
                                                 C, radius, resolve_coord_mode(coord_mode), scale,
                                                 _stream(fmap1)), "raftcorr_alt_bwd")
    return g1, g2, gc


# ---- autograd -------------------------------------------------------------------------------
class _GradSink:
    """Shared d(pyramid) accumulator: every lookup of one CorrBlock scatters into the same
    buffers instead of materialising a volume-sized gradient per iteration.

    The accumulated gradient belongs to ONE backward pass: BuildFn.backward consumes and drops it.  A pass that
    scatters but never reaches BuildFn (``torch.autograd.grad(out, coords)`` prunes that node; an exception can
    abort the pass) must not leave its partial sums behind for the next pass to inherit, so the first scatter of
    a pass queues an end-of-pass callback with the autograd engine that drops whatever is still here."""

    def __init__(self):
        self.grad_pyr: Optional[Pyramid] = None
        self._pass_open = False

    def buffers(self, like: Pyramid, device) -> Pyramid:
        if self.grad_pyr is None:
            self.grad_pyr = Pyramid(like.B, like.H, like.W, like.num_levels, device, zero=True)
        if not self._pass_open:
            self._pass_open = True
            try:   # only valid while the engine is running a backward pass - which is when this is called
                torch.autograd.Variable._execution_engine.queue_callback(self.end_of_pass)
            except RuntimeError:
                self._pass_open = False
        return self.grad_pyr

    def end_of_pass(self):
        self.grad_pyr = None
        self._pass_open = False


class BuildFn(torch.autograd.Function):
    """fmaps -> (token, level views).  The 1-element token is fed to every LookupFn so that
    autograd runs this node's backward only after all lookups have scattered their gradients."""

    @staticmethod
    def forward(ctx, fmap1, fmap2, num_levels, precision, holder):
        pyr = build_pyramid(fmap1, fmap2, num_levels, precision)
        holder["pyr"] = pyr
        holder["sink"] = _GradSink()
        ctx.holder = holder
        ctx.save_for_backward(fmap1, fmap2)
        token = torch.zeros(1, dtype=torch.float32, device=fmap1.device)
        return (token, *pyr.views())

    @staticmethod
    def backward(ctx, gtoken, *glevels):
        fmap1, fmap2 = ctx.saved_tensors
        pyr: Pyramid = ctx.holder["pyr"]
        sink: _GradSink = ctx.holder["sink"]
        gp = sink.buffers(pyr, fmap1.device)
        for l, g in enumerate(glevels):  # gradients that reached corr_pyramid[l] directly
            if g is not None:
                gp.storage[l][:, :, : gp.desc[l].W].add_(g.squeeze(1))
        g1, g2 = build_backward(gp, fmap1, fmap2)
        sink.grad_pyr = None     # consumed (the end-of-pass callback then finds nothing to drop)
        return g1, g2, None, None, None


class LookupFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, coords, token, holder, radius, coord_mode):
        ctx.holder, ctx.radius, ctx.coord_mode = holder, radius, coord_mode
        ctx.save_for_backward(coords)
        return lookup(holder["pyr"], coords, radius, coord_mode)

    @staticmethod
    def backward(ctx, grad_out):
        (coords,) = ctx.saved_tensors
        pyr: Pyramid = ctx.holder["pyr"]
        want_levels = ctx.needs_input_grad[1]
        want_coords = ctx.needs_input_grad[0]
        gp = ctx.holder["sink"].buffers(pyr, coords.device) if want_levels else None
        gcoords = None
        if want_levels or want_coords:
            gcoords = lookup_backward(pyr if want_coords else None, gp, coords, grad_out, ctx.radius,
                                      want_coords, ctx.coord_mode)
        gtoken = torch.zeros(1, dtype=torch.float32, device=coords.device) if want_levels else None
        return gcoords, gtoken, None, None, None


class AltFn(torch.autograd.Function):
    """One level of the on-the-fly path with alt_cuda_corr's layouts, differentiable."""

    @staticmethod
    def forward(ctx, fmap1, fmap2, coords, radius, coord_mode, scale):
        ctx.radius, ctx.coord_mode, ctx.scale = radius, coord_mode, scale
        ctx.save_for_backward(fmap1, fmap2, coords)
        return alt_forward(fmap1, fmap2, coords, radius, coord_mode, scale)

    @staticmethod
    def backward(ctx, grad):
        fmap1, fmap2, coords = ctx.saved_tensors
        g1, g2, gc = alt_backward(fmap1, fmap2, coords, grad.contiguous(), ctx.radius, ctx.coord_mode,
                                  ctx.scale, want_coords=ctx.needs_input_grad[2])
        return g1, g2, gc, None, None, None


class AltFusedFn(torch.autograd.Function):
    """All levels of the on-the-fly path in one tensor-core launch (bf16 products), differentiable: the
    backward runs the fp32 per-level kernels on the fp32 feature maps (it does not need the forward's output,
    the op is bilinear in the feature maps)."""

    @staticmethod
    def forward(ctx, coords, fmap1, holder, radius, coord_mode, scale, *fmap2_levels):
        ctx.radius, ctx.coord_mode, ctx.scale = radius, coord_mode, scale
        ctx.save_for_backward(coords, fmap1, *fmap2_levels)
        return alt_fused_forward(holder["f1_bf16"], holder["f2_bf16"], coords, radius, coord_mode, scale)

    @staticmethod
    def backward(ctx, grad):
        coords, fmap1, *levels = ctx.saved_tensors
        B, _, H, W = coords.shape
        win = (2 * ctx.radius + 1) ** 2
        want_coords = ctx.needs_input_grad[0]
        base = coords.permute(0, 2, 3, 1)
        g1_total, gc_total, g2 = None, None, []
        for l, f2 in enumerate(levels):
            coords_l = (base / 2 ** l).reshape(B, 1, H, W, 2).contiguous()
            g_l = grad[:, l * win:(l + 1) * win].reshape(B, 1, win, H, W).contiguous()
            g1, g2_l, gc = alt_backward(fmap1, f2, coords_l, g_l, ctx.radius, ctx.coord_mode, ctx.scale,
                                        want_coords=want_coords)
            g1_total = g1 if g1_total is None else g1_total.add_(g1)
            g2.append(g2_l)
            if want_coords:
                gc = gc.reshape(B, H, W, 2).permute(0, 3, 1, 2) / 2 ** l
                gc_total = gc if gc_total is None else gc_total + gc
        return (gc_total.contiguous() if want_coords else None, g1_total, None, None, None, None, *g2)


def lookup_conv1x1(pyr: Pyramid, coords: torch.Tensor, radius: int, weight: torch.Tensor,
                   bias: Optional[torch.Tensor] = None, relu: bool = True, coord_mode=None) -> torch.Tensor:
    """Window lookup followed by the 1x1 convolution (+ bias, + ReLU) that consumes it (update.py:83,90):
    (B,2,H,W) coords -> (B, out_channels, H, W), TF32 products.  Inference only."""
    _require_cuda_f32(coords, "coords")
    _require_cuda_f32(weight, "weight")
    coords = coords.contiguous()
    B, _, H, W = coords.shape
    rd = 2 * radius + 1
    C = pyr.num_levels * rd * rd
    cout = weight.shape[0]
    if weight.numel() != cout * C:
        raise RuntimeError(f"weight must hold (out_channels, {C}) values, got {tuple(weight.shape)}")
    weight = weight.reshape(cout, C).contiguous()
    if bias is not None:
        _require_cuda_f32(bias, "bias")
        bias = bias.contiguous()
    out = torch.empty((B, cout, H, W), dtype=torch.float32, device=coords.device)
    need = _capi.lib().raftcorr_lookup_conv1x1_workspace_bytes(B, H, W, pyr.num_levels, radius)
    ws = torch.empty(max(int(need), 16) // 4, dtype=torch.float32, device=coords.device)
    with _on_device(coords.device):
        _capi.check(_capi.lib().raftcorr_lookup_conv1x1_fwd(
            pyr.desc, pyr.num_levels, coords.data_ptr(), weight.data_ptr(),
            bias.data_ptr() if bias is not None else None, out.data_ptr(), B, H, W, radius,
            resolve_coord_mode(coord_mode), cout, 1 if relu else 0, ws.data_ptr(), ws.numel() * 4, _stream(coords)),
            "raftcorr_lookup_conv1x1_fwd")
    return out


def conv1x1_channels_last(act: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None,
                          relu: bool = True) -> torch.Tensor:
    """relu(conv1x1(act)) for a (B, C, H, W) tensor held in channels_last memory (what ``lookup(...,
    channels_last=True)`` returns) or plain NCHW (the default lookup) -> (B, out_channels, H, W) contiguous.
    tcgen05 TF32 GEMM with the bias and the ReLU in its epilogue.  Inference only."""
    _require_cuda_f32(act, "act")
    _require_cuda_f32(weight, "weight")
    B, C, H, W = act.shape
    cout = weight.shape[0]
    # a plain NCHW tensor (the default lookup, the 16-bit lookup) goes to the MN-major variant of the same GEMM
    nchw = act.is_contiguous() and not act.is_contiguous(memory_format=torch.channels_last) and cout <= 256 \
        and C % 4 == 0 and (H * W) % 4 == 0
    if not nchw and not act.is_contiguous(memory_format=torch.channels_last):
        act = act.contiguous(memory_format=torch.channels_last)
    if weight.numel() != cout * C:
        raise RuntimeError(f"weight must hold (out_channels, {C}) values, got {tuple(weight.shape)}")
    weight = weight.reshape(cout, C).contiguous()
    if bias is not None:
        _require_cuda_f32(bias, "bias")
        bias = bias.contiguous()
    out = torch.empty((B, cout, H, W), dtype=torch.float32, device=act.device)
    with _on_device(act.device):
        fn, name = ((_capi.lib().raftcorr_conv1x1_nchw_fwd, "raftcorr_conv1x1_nchw_fwd") if nchw else
                    (_capi.lib().raftcorr_conv1x1_nhwc_fwd, "raftcorr_conv1x1_nhwc_fwd"))
        _capi.check(fn(act.data_ptr(), weight.data_ptr(), bias.data_ptr() if bias is not None else None, out.data_ptr(), B,
                       H * W, C, cout, 1 if relu else 0, _stream(act)), name)
    return out


# ---- the step after the path: flow upsampling ------------------------------------

def _check_flow_mask(flow: torch.Tensor, mask: Optional[torch.Tensor]):
    _require_cuda_f32(flow, "flow")
    if flow.dim() != 4 or flow.shape[1] != 2:
        raise RuntimeError(f"flow must be (B,2,H,W), got {tuple(flow.shape)}")
    if mask is not None:
        _require_cuda_f32(mask, "mask")
        B, _, H, W = flow.shape
        if tuple(mask.shape) != (B, 576, H, W):
            raise RuntimeError(f"mask must be (B,576,H,W) = {(B, 576, H, W)}, got {tuple(mask.shape)}")
        if mask.device != flow.device:
            raise RuntimeError("flow and mask must be on the same device")


def upsample_flow_forward(flow: torch.Tensor, mask: torch.Tensor) -> torch.Tensor:
    _check_flow_mask(flow, mask)
    flow, mask = flow.contiguous(), mask.contiguous()
    B, _, H, W = flow.shape
    up = torch.empty((B, 2, 8 * H, 8 * W), dtype=torch.float32, device=flow.device)
    with _on_device(flow.device):
        _capi.check(_capi.lib().raftcorr_upsample_flow_fwd(flow.data_ptr(), mask.data_ptr(), up.data_ptr(), B, H, W,
                                                           _stream(flow)), "raftcorr_upsample_flow_fwd")
    return up


def upsample_flow_backward(flow: torch.Tensor, mask: torch.Tensor, grad_up: torch.Tensor, want_flow: bool = True,
                           want_mask: bool = True):
    _check_flow_mask(flow, mask)
    flow, mask, grad_up = flow.contiguous(), mask.contiguous(), grad_up.contiguous()
    B, _, H, W = flow.shape
    gflow = torch.empty_like(flow) if want_flow else None
    gmask = torch.empty_like(mask) if want_mask else None
    with _on_device(flow.device):
        _capi.check(_capi.lib().raftcorr_upsample_flow_bwd(
            flow.data_ptr(), mask.data_ptr(), grad_up.data_ptr(), gflow.data_ptr() if want_flow else None,
            gmask.data_ptr() if want_mask else None, B, H, W, _stream(flow)), "raftcorr_upsample_flow_bwd")
    return gflow, gmask


def upflow8_forward(flow: torch.Tensor) -> torch.Tensor:
    _check_flow_mask(flow, None)
    flow = flow.contiguous()
    B, _, H, W = flow.shape
    up = torch.empty((B, 2, 8 * H, 8 * W), dtype=torch.float32, device=flow.device)
    with _on_device(flow.device):
        _capi.check(_capi.lib().raftcorr_upflow8(flow.data_ptr(), up.data_ptr(), B, H, W, _stream(flow)),
                    "raftcorr_upflow8")
    return up


class UpsampleFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, flow, mask):
        ctx.save_for_backward(flow, mask)
        return upsample_flow_forward(flow, mask)

    @staticmethod
    def backward(ctx, grad_up):
        flow, mask = ctx.saved_tensors
        if not (ctx.needs_input_grad[0] or ctx.needs_input_grad[1]):
            return None, None
        return upsample_flow_backward(flow, mask, grad_up, ctx.needs_input_grad[0], ctx.needs_input_grad[1])
