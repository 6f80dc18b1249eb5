"""Tiny driver for ncu: lookup backward (TMA reduce path) and build backward at the training shape (configs[4])."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200")]
from raftcorr_b200 import ops
B, D, H, W = 12, 256, 46, 62
ys, xs = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
c = torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1) + 4 * torch.randn(B, 2, H, W, device="cuda")
gp = ops.Pyramid(B, H, W, 4, "cuda", zero=True)
go = torch.randn(B, 324, H, W, device="cuda")
for _ in range(4):
    ops.lookup_backward(None, gp, c, go, 4, False)
torch.cuda.synchronize()
import time
t0 = time.perf_counter()
for _ in range(200):
    ops.lookup_backward(None, gp, c, go, 4, False)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"host issue {1e6*(t1-t0)/200:.1f} us/call, with drain {1e6*(t2-t0)/200:.1f} us/call")
f1 = torch.randn(B, D, H, W, device="cuda"); f2 = torch.randn(B, D, H, W, device="cuda")
for _ in range(2):
    ops.build_backward(gp, f1, f2)
torch.cuda.synchronize()
