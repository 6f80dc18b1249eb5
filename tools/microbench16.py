"""Kernel-level timings of the 16-bit pyramid path on the B200 (CUDA events, L2 flushed between reps).
Development aid; bench.py is the contract.   usage: microbench16.py [B D H W]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200")]
from raftcorr_b200 import ops  # noqa: E402

PK = os.path.join(ROOT, "MEASURED_PEAKS.json")
PEAKS = json.load(open(PK)) if os.path.exists(PK) else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}


def timeit(fn, reps=10, warm=3, flush=None):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        if flush is not None:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    B, D, H, W = (int(v) for v in (sys.argv[1:5] if len(sys.argv) >= 5 else (8, 256, 55, 128)))
    N = H * W
    hbm = PEAKS["hbm_gbs"]
    f1 = torch.randn(B, D, H, W, device="cuda")
    f2 = torch.randn(B, D, H, W, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    T = sum((H >> l) * (W >> l) for l in range(4))
    flops = 2.0 * B * N * N * D
    for storage, es in (("fp32", 4), ("bf16", 2)):
        ms = timeit(lambda: ops.build_pyramid(f1, f2, 4, "bf16", storage=storage), reps=10, flush=flush)
        byts = B * N * T * es + 2 * B * N * D * 4
        print(f"build bf16 mma, {storage} pyramid: {ms:.3f} ms  {flops/ms/1e9:.0f} TFLOP/s "
              f"({flops/ms/1e9/PEAKS['bf16_tflops']:.2f} of burst, {flops/ms/1e9/PEAKS['bf16_tflops_sustained']:.2f} of sustained)  "
              f"algorithmic {byts/ms/1e6:.0f} GB/s ({byts/ms/1e6/hbm:.2f} of HBM)", flush=True)
    pyr32 = ops.build_pyramid(f1, f2, 4, "bf16", storage="fp32")
    pyr16 = ops.build_pyramid(f1, f2, 4, "bf16", storage="bf16")
    ys, xs = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
    grid = torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1)
    low = 8 * torch.randn(B, 2, max(H // 8, 2), max(W // 8, 2), device="cuda")
    smooth = grid + torch.nn.functional.interpolate(low, size=(H, W), mode="bilinear", align_corners=True)
    cases = (("grid", grid), ("smooth8", smooth), ("sigma4", grid + 4 * torch.randn_like(grid)),
             ("sigma30", grid + 30 * torch.randn_like(grid)))
    for name, c in cases:
        for tag, pyr, bpq in (("fp32 pyramid (lookup_tma)", pyr32, 2904), ("bf16 pyramid (lookup16)", pyr16, 2104)):
            ms = timeit(lambda: ops.lookup(pyr, c, 4), reps=20, flush=flush)
            byts = B * N * bpq
            print(f"lookup {name:8s} {tag:28s}: {ms*1e3:6.1f} us  {byts/ms/1e6:5.0f} GB/s algorithmic at {bpq} B/query "
                  f"({byts/ms/1e6/hbm:.3f} of measured HBM)", flush=True)
    # kernel-only: 12 lookups with different coords in one CUDA graph (no host time between launches)
    cs = [smooth + 0.3 * torch.randn_like(grid) for _ in range(12)]
    for tag, pyr, bpq in (("fp32", pyr32, 2904), ("bf16", pyr16, 2104)):
        ops.lookup(pyr, cs[0], 4)
        g = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            with torch.cuda.graph(g, stream=s):
                outs = [ops.lookup(pyr, c, 4) for c in cs]
        torch.cuda.current_stream().wait_stream(s)
        ms = timeit(g.replay, reps=10) / len(cs)
        byts = B * N * bpq
        print(f"lookup graph x12 smooth {tag}: {ms*1e3:6.1f} us per launch  {byts/ms/1e6:5.0f} GB/s ({byts/ms/1e6/hbm:.3f})", flush=True)
        del g, outs


if __name__ == "__main__":
    main()
