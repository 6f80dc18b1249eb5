"""lookup16 + the tensor-core convc1 on its NCHW output at BASELINE.json configs[1] (B=8, 55x128), against stock torch."""
import os, sys, torch, torch.nn.functional as F
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200"), os.path.join(ROOT, "tools")]
import corr
from microbench import timeit
B, D, H, W = 8, 256, 55, 128
torch.manual_seed(0)
f1 = torch.randn(B, D, H, W, device="cuda"); f2 = torch.randn(B, D, H, W, device="cuda")
ys, xs = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
c = torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1) + 2 * torch.randn(B, 2, H, W, device="cuda")
w = torch.randn(256, 324, 1, 1, device="cuda") / 18; b = torch.randn(256, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
with torch.no_grad():
    blk = corr.CorrBlock(f1, f2)
    lk = blk(c)
    t_l = timeit(lambda: blk(c), reps=10, warm=3, flush=flush) if "flush" in timeit.__code__.co_varnames else timeit(lambda: blk(c), reps=10, warm=3)
    t_c = timeit(lambda: corr.conv1x1_relu(lk, w, b), reps=10, warm=3)
    t_s = timeit(lambda: F.relu(F.conv2d(lk, w, b)), reps=10, warm=3)
    t_b = timeit(lambda: corr.conv1x1_relu(blk(c), w, b), reps=10, warm=3)
    print(f"lookup16 {t_l*1e3:.1f} us; conv1x1 on NCHW (tcgen05 tf32, MN-major) {t_c*1e3:.1f} us; stock conv2d + relu {t_s*1e3:.1f} us; "
          f"lookup16 + conv1x1 back to back {t_b*1e3:.1f} us")
