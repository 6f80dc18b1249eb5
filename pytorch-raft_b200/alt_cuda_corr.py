"""Drop-in for the reference's compiled ``alt_cuda_corr`` extension module.

Same surface as alt_cuda_corr/correlation.cpp:51-54: ``forward(fmap1, fmap2, coords, radius)
-> [corr]`` and ``backward(fmap1, fmap2, coords, corr_grad, radius) -> [fmap1_grad,
fmap2_grad, coords_grad]`` with the reference's layouts (fmaps NHWC, coords (B,1,H1,W1,2),
corr (B,1,(2r+1)^2,H1,W1), unscaled) and its checks (CUDA + contiguous, else RuntimeError;
correlation.cpp:19-21).  Coordinates follow the reference kernel (floor(coords) directly).
One documented superset: ``coords_grad`` is filled, the reference returns zeros
(correlation_kernel.cu:130,307,323).
"""
from __future__ import annotations

from raftcorr_b200 import ops
from raftcorr_b200._capi import COORD_DIRECT


def _check_input(t, name):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    if t.dtype != ops.torch.float32:
        raise RuntimeError(f"{name} must be float32")


def _check_shapes(fmap1, fmap2, coords):
    if fmap1.dim() != 4 or fmap2.dim() != 4 or coords.dim() != 5:
        raise RuntimeError("expected fmap1/fmap2 (B,H,W,C) and coords (B,N,H1,W1,2)")
    if coords.shape[1] != 1:
        raise RuntimeError("only one coordinate set per pixel (N == 1) is supported, as core/corr.py:85 uses")


def forward(fmap1, fmap2, coords, radius):
    for t, n in ((fmap1, "fmap1"), (fmap2, "fmap2"), (coords, "coords")):
        _check_input(t, n)
    _check_shapes(fmap1, fmap2, coords)
    return [ops.alt_forward(fmap1, fmap2, coords, int(radius), COORD_DIRECT, 1.0)]


def backward(fmap1, fmap2, coords, corr_grad, radius):
    for t, n in ((fmap1, "fmap1"), (fmap2, "fmap2"), (coords, "coords"), (corr_grad, "corr_grad")):
        _check_input(t, n)
    _check_shapes(fmap1, fmap2, coords)
    g1, g2, gc = ops.alt_backward(fmap1, fmap2, coords, corr_grad, int(radius), COORD_DIRECT, 1.0, True)
    return [g1, g2, gc]
