import os, sys, torch
sys.path[:0] = ["/root/repo", "/root/repo/pytorch-raft_b200", "/root/repo/tools"]
import corr
from microbench import timeit
B, D, H, W = 4, 256, 136, 240
f1 = torch.randn(B, D, H, W, device="cuda"); f2 = torch.randn(B, D, H, W, device="cuda")
with torch.no_grad():
    ms = timeit(lambda: corr.AlternateCorrBlock(f1, f2), reps=5, warm=2)
    print(f"AlternateCorrBlock.__init__ cfg4: {ms:.3f} ms")
