import os, sys, torch
sys.path[:0] = ["/root/repo", "/root/repo/pytorch-raft_b200"]
from raftcorr_b200 import ops
B, D, H, W = 8, 256, 55, 128
torch.manual_seed(0)
fs = [(torch.randn(B, D, H, W, device="cuda"), torch.randn(B, D, H, W, device="cuda")) for _ in range(2)]
fl = 2.0 * B * (H * W) ** 2 * D
for il in (None, False):
    for i in range(3): ops.build_pyramid(*fs[i & 1], 4, "bf16", storage="bf16", interleave=il)
    ts = []
    for i in range(10):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); p = ops.build_pyramid(*fs[i & 1], 4, "bf16", storage="bf16", interleave=il); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b)); del p
    ts.sort()
    print(f"interleave={il}: build16 (pack + gemm) {ts[5]*1e3:.1f} us  {fl/ts[5]/1e9:.0f} TFLOP/s", flush=True)
# as bench.py times it: six builds back to back between two events
for il in (None, False):
    for rep in range(2):
        for i in range(2): ops.build_pyramid(*fs[i & 1], 4, "bf16", storage="bf16", interleave=il)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(6): ops.build_pyramid(*fs[i & 1], 4, "bf16", storage="bf16", interleave=il)
        b.record(); torch.cuda.synchronize()
    print(f"interleave={il}: six builds back to back: {a.elapsed_time(b)/6*1e3:.1f} us each", flush=True)
