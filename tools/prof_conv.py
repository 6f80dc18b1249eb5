"""Tiny driver for ncu: channels-last lookup + convc1 GEMM at configs[1] shape."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200")]
import corr
B, D, H, W = 8, 256, 55, 128
f1 = torch.randn(B, D, H, W, device="cuda"); f2 = torch.randn(B, D, H, W, device="cuda")
ys, xs = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
c = torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1) + 4 * torch.randn(B, 2, H, W, device="cuda")
w = torch.randn(256, 324, 1, 1, device="cuda") / 18; b = torch.randn(256, device="cuda")
with torch.no_grad():
    blk = corr.CorrBlock(f1, f2, channels_last=True)
    for _ in range(4):
        out = corr.conv1x1_relu(blk(c), w, b)
torch.cuda.synchronize()
print("done")
