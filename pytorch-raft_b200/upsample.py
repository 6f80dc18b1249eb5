"""Flow upsampling on the B200: the step right after the correlation path.

``upsample_flow(flow, mask)`` replaces the method ``RAFT.upsample_flow`` (reference core/raft.py:72-83), which
the reference calls after every refinement iteration (raft.py:134-139); ``upflow8(flow)`` replaces
core/utils/utils.py:80-82 (the small model, which predicts no mask).  To use them without editing the
reference::

    import raft, upsample
    raft.RAFT.upsample_flow = lambda self, flow, mask: upsample.upsample_flow(flow, mask)
    raft.upflow8 = upsample.upflow8

Same shapes and meaning as the reference: flow (N,2,H,W), mask (N,576,H,W) -> (N,2,8H,8W), fp32, differentiable
w.r.t. both inputs (upflow8: forward only).  CUDA tensors only; there is no CPU fallback.
"""
from __future__ import annotations

import torch

from raftcorr_b200 import ops


def upsample_flow(flow: torch.Tensor, mask: torch.Tensor) -> torch.Tensor:
    """[H/8, W/8, 2] -> [H, W, 2] by the learned convex combination of the 3x3 coarse neighbours."""
    if mask.dtype != torch.float32:      # autocast hands the mask over in half precision (update.py:134)
        mask = mask.float()
    if flow.dtype != torch.float32:
        flow = flow.float()
    if torch.is_grad_enabled() and (flow.requires_grad or mask.requires_grad):
        return ops.UpsampleFn.apply(flow, mask)
    return ops.upsample_flow_forward(flow, mask)


def upflow8(flow: torch.Tensor, mode: str = "bilinear") -> torch.Tensor:
    if mode != "bilinear":
        raise NotImplementedError("upflow8: only mode='bilinear' (the reference's default and only use)")
    if torch.is_grad_enabled() and flow.requires_grad:
        raise NotImplementedError("upflow8 is forward-only in this library")
    return ops.upflow8_forward(flow.float())
