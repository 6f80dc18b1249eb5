"""Timings of the on-the-fly path (cfg4) and the backward kernels (cfg5) on the B200."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200")]
from raftcorr_b200 import ops
import corr
sys.path.insert(0, os.path.join(ROOT, "tools"))
from microbench import timeit

def grid(B, H, W):
    ys, xs = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
    return torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1)

which = sys.argv[1] if len(sys.argv) > 1 else "all"
if which in ("all", "alt"):
    B, D, H, W = 4, 256, 136, 240   # BASELINE.json configs[3]
    f1 = torch.randn(B, D, H, W, device="cuda"); f2 = torch.randn(B, D, H, W, device="cuda")
    with torch.no_grad():
        blk = corr.AlternateCorrBlock(f1, f2)
        # smooth flow: N(0, 8^2) at 1/8 resolution, bilinearly upsampled
        low = 8 * torch.randn(B, 2, H // 8, W // 8, device="cuda")
        smooth = grid(B, H, W) + torch.nn.functional.interpolate(low, size=(H, W), mode="bilinear", align_corners=True)
        rough = grid(B, H, W) + 4 * torch.randn(B, 2, H, W, device="cuda")
        for name, c in (("smooth", smooth), ("iid sigma4", rough)):
            ms = timeit(lambda: blk(c), reps=5, warm=2)
            N = H * W
            flops = B * N * 4 * 100 * 2 * D
            print(f"alt cfg4 {name}: {ms:.3f} ms/iter  {flops/ms/1e9:.1f} TFLOP/s (useful, (2r+2)^2 taps)  compulsory 120.4 MB/pair -> {B*120.4e6/ms/1e6:.0f} GB/s")
if which in ("all", "bwd"):
    B, D, H, W = 12, 256, 46, 62    # BASELINE.json configs[4]
    f1 = torch.randn(B, D, H, W, device="cuda", requires_grad=True); f2 = torch.randn(B, D, H, W, device="cuda", requires_grad=True)
    c = grid(B, H, W) + 4 * torch.randn(B, 2, H, W, device="cuda")
    def step():
        blk = corr.CorrBlock(f1, f2)
        loss = 0
        for _ in range(12):
            loss = loss + blk(c).sum()
        loss.backward()
    ms = timeit(step, reps=3, warm=1)
    print(f"cfg5 CorrBlock fwd+bwd (build + 12 lookups + 12 lookup_bwd + build_bwd), B=12 46x62: {ms:.2f} ms")
    pyr = ops.build_pyramid(f1.detach(), f2.detach(), 4)
    gp = ops.Pyramid(B, H, W, 4, "cuda", zero=True)
    go = torch.randn(B, 324, H, W, device="cuda")
    ms = timeit(lambda: ops.lookup_backward(None, gp, c, go, 4, False), reps=5)
    print(f"  lookup_bwd (TMA reduce-add boxes, incl. host call): {ms*1e3:.1f} us")
    ms = timeit(lambda: ops.build_backward(gp, f1.detach(), f2.detach()), reps=3)
    print(f"  build_bwd (unpool + 2 fp32 GEMMs): {ms:.2f} ms  {4*B*(H*W)**2*D/ms/1e9:.1f} TFLOP/s")
if which in ("all", "altbwd"):
    B, D, H, W = 12, 256, 46, 62    # training shape, on-the-fly block with autograd (fp32 per-level kernels)
    f1 = torch.randn(B, D, H, W, device="cuda", requires_grad=True); f2 = torch.randn(B, D, H, W, device="cuda", requires_grad=True)
    c = grid(B, H, W) + 4 * torch.randn(B, 2, H, W, device="cuda")
    blk = corr.AlternateCorrBlock(f1, f2)
    out = blk(c)
    go = torch.randn_like(out)
    ms_f = timeit(lambda: blk(c), reps=3, warm=1)
    def fb():
        f1.grad = None; f2.grad = None
        corr.AlternateCorrBlock(f1, f2)(c).backward(go)
    ms_fb = timeit(fb, reps=3, warm=1)
    print(f"alt (autograd: fused TC forward, fp32 per-level backward) B=12 46x62: fwd {ms_f:.2f} ms, fwd+bwd {ms_fb:.2f} ms per iteration")
    with torch.no_grad():
        ms = timeit(lambda: corr.AlternateCorrBlock(f1.detach(), f2.detach())(c), reps=3, warm=1)
    print(f"alt (no grad, fused bf16 TC incl. constructor) same shape: {ms:.3f} ms")
