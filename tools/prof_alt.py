import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200")]
import corr
B, D, H, W = 4, 256, 136, 240
f1 = torch.randn(B, D, H, W, device="cuda"); f2 = torch.randn(B, D, H, W, device="cuda")
ys, xs = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
grid = torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1)
low = 8 * torch.randn(B, 2, H // 8, W // 8, device="cuda")
c = grid + torch.nn.functional.interpolate(low, size=(H, W), mode="bilinear", align_corners=True)
with torch.no_grad():
    blk = corr.AlternateCorrBlock(f1, f2)
    for _ in range(3):
        out = blk(c)
torch.cuda.synchronize(); print("done")
