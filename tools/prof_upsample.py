"""Tiny driver for ncu: a few convex-upsampling launches (forward and backward) at configs[1] shape."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200")]
from raftcorr_b200 import ops
B, H, W = 8, 55, 128
masks = [torch.randn(B, 576, H, W, device="cuda") for _ in range(3)]
flow = torch.randn(B, 2, H, W, device="cuda")
gup = torch.randn(B, 2, 8 * H, 8 * W, device="cuda")
for m in masks:
    ops.upsample_flow_forward(flow, m)
for m in masks:
    ops.upsample_flow_backward(flow, m, gup)
torch.cuda.synchronize()
print("done")
