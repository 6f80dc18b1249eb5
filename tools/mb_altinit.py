"""AlternateCorrBlock.__init__ at BASELINE.json configs[3] (B=4, D=256, 136x240) and its alt_pack launches alone."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200"), os.path.join(ROOT, "tools")]
import corr
from raftcorr_b200 import ops
from microbench import timeit
B, D, H, W = 4, 256, 136, 240
f1 = torch.randn(B, D, H, W, device="cuda"); f2 = torch.randn(B, D, H, W, device="cuda")
with torch.no_grad():
    ms = timeit(lambda: corr.AlternateCorrBlock(f1, f2), reps=5, warm=2)
    print(f"AlternateCorrBlock.__init__ cfg4: {ms:.3f} ms")
    for L in (1, 4):
        ms = timeit(lambda: ops.alt_pack(f2, L), reps=10, warm=2)
        byts = f2.numel() * 4 + sum(B * (H >> l) * (W >> l) * D * 2 for l in range(L))
        print(f"alt_pack {L} level(s): {ms*1e3:.1f} us  {byts/ms/1e6:.0f} GB/s algorithmic")
