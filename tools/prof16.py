"""Tiny driver for ncu: the 16-bit build and a few lookup16 launches at BASELINE.json configs[1] shape.
usage: prof16.py [build|lookup|both] [grid|smooth|sigma4]"""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200")]
from raftcorr_b200 import ops
B, D, H, W = 8, 256, 55, 128
what = sys.argv[1] if len(sys.argv) > 1 else "both"
kind = sys.argv[2] if len(sys.argv) > 2 else "smooth"
torch.manual_seed(0)
f1 = torch.randn(B, D, H, W, device="cuda"); f2 = torch.randn(B, D, H, W, device="cuda")
pyr = ops.build_pyramid(f1, f2, 4, "bf16", storage="bf16")
if what in ("both", "build"):
    for _ in range(2):
        pyr = ops.build_pyramid(f1, f2, 4, "bf16", storage="bf16")
ys, xs = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
grid = torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1)
if kind == "smooth":
    low = 8 * torch.randn(B, 2, H // 8, W // 8, device="cuda")
    c = grid + torch.nn.functional.interpolate(low, size=(H, W), mode="bilinear", align_corners=True)
elif kind == "sigma4":
    c = grid + 4 * torch.randn_like(grid)
else:
    c = grid
if what in ("both", "lookup"):
    for _ in range(4):
        out = ops.lookup(pyr, c, 4)
torch.cuda.synchronize()
print("done")
