"""Tiny driver for ncu: runs the build and a few lookups at BASELINE.json configs[1] shape."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200")]
from raftcorr_b200 import ops
B, D, H, W = 8, 256, 55, 128
what = sys.argv[1] if len(sys.argv) > 1 else "both"
f1 = torch.randn(B, D, H, W, device="cuda"); f2 = torch.randn(B, D, H, W, device="cuda")
pyr = ops.build_pyramid(f1, f2, 4, "bf16")
if what in ("both", "build"):
    for _ in range(2):
        pyr = ops.build_pyramid(f1, f2, 4, "bf16")
ys, xs = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
c = torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1) + 4 * torch.randn(B, 2, H, W, device="cuda")
if what in ("both", "lookup"):
    for _ in range(4):
        out = ops.lookup(pyr, c, 4)
torch.cuda.synchronize()
print("done")
