"""Prototype check: lookup16 on a pyramid whose levels 0..k keep the blocks of 2^g neighbouring images interleaved
(made here by a torch permutation of the plain blocked pyramid) — same values?  how fast?"""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200")]
from raftcorr_b200 import ops

LEVELS = int(sys.argv[1]) if len(sys.argv) > 1 else 2      # how many levels to interleave
G = 4
B, D, H, W = 8, 256, 55, 128
torch.manual_seed(0)
f1 = torch.randn(B, D, H, W, device="cuda"); f2 = torch.randn(B, D, H, W, device="cuda")
pyr = ops.build_pyramid(f1, f2, 4, "bf16", storage="bf16")
pyi = ops.build_pyramid(f1, f2, 4, "bf16", storage="bf16")
keep = []
for l in range(LEVELS):
    s = pyr.storage[l]                                   # (BN, by, bx, 4, 8)
    n, by, bx = s.shape[:3]
    assert n % G == 0
    t = s.reshape(n // G, G, by, bx, 32).permute(0, 2, 3, 1, 4).contiguous()
    keep.append(t)
    pyi.desc[l].data = t.data_ptr()
    pyi.desc[l].interleave_log2 = 2
ys, xs = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
grid = torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1)
low = 8 * torch.randn(B, 2, H // 8, W // 8, device="cuda")
smooth = grid + torch.nn.functional.interpolate(low, size=(H, W), mode="bilinear", align_corners=True)
rough = grid + 30 * torch.randn_like(grid)
for name, c in (("smooth", smooth), ("grid", grid), ("sigma30", rough)):
    a, b = ops.lookup(pyr, c, 4), ops.lookup(pyi, c, 4)
    print(f"{name}: interleaved == plain: {torch.equal(a, b)}  max|d| {(a - b).abs().max().item():.3g}", flush=True)
cs = {"smooth": [smooth + 0.3 * torch.randn_like(grid) for _ in range(12)], "grid": [grid.clone() for _ in range(12)],
      "sigma30": [grid + 30 * torch.randn_like(grid) for _ in range(12)]}
for tag, P in (("plain", pyr), (f"interleaved x{G} (levels 0..{LEVELS - 1})", pyi)):
    for name, cl in cs.items():
        ops.lookup(P, cl[0], 4)
        g = torch.cuda.CUDAGraph(); s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            with torch.cuda.graph(g, stream=s):
                outs = [ops.lookup(P, c, 4) for c in cl]
        torch.cuda.current_stream().wait_stream(s)
        for _ in range(3): g.replay()
        ts = []
        for _ in range(10):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); g.replay(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b) / 12)
        ts.sort()
        print(f"{tag:36s} {name:8s}: {ts[5]*1e3:.1f} us per launch ({B*H*W*2104/ts[5]/1e6/6453.4:.3f} of HBM at 2104 B/query)", flush=True)
        del g, outs
