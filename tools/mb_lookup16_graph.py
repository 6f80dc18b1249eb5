import os, sys, torch
sys.path[:0] = ["/root/repo", "/root/repo/pytorch-raft_b200"]
from raftcorr_b200 import ops
B, D, H, W = 8, 256, 55, 128
torch.manual_seed(0)
f1 = torch.randn(B, D, H, W, device="cuda"); f2 = torch.randn(B, D, H, W, device="cuda")
pyr = ops.build_pyramid(f1, f2, 4, "bf16", storage="bf16")
ys, xs = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
grid = torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1)
low = 8 * torch.randn(B, 2, H // 8, W // 8, device="cuda")
smooth = grid + torch.nn.functional.interpolate(low, size=(H, W), mode="bilinear", align_corners=True)
cs = {"smooth": [smooth + 0.3 * torch.randn_like(grid) for _ in range(12)], "grid": [grid.clone() for _ in range(12)]}
for name, cl in cs.items():
    ops.lookup(pyr, cl[0], 4)
    g = torch.cuda.CUDAGraph(); s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            outs = [ops.lookup(pyr, c, 4) for c in cl]
    torch.cuda.current_stream().wait_stream(s)
    for _ in range(3): g.replay()
    ts = []
    for _ in range(10):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); g.replay(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b) / 12)
    ts.sort()
    print(f"RAFTCORR_L16_DEBUG={os.environ.get('RAFTCORR_L16_DEBUG','0')} {name}: {ts[5]*1e3:.1f} us per launch ({B*H*W*2104/ts[5]/1e6/6453.4:.3f} of HBM at 2104 B/query)", flush=True)
