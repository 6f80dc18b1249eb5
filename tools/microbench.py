"""Kernel-level timings on the B200 (CUDA events, L2 flushed between reps).  Development aid;
bench.py is the contract."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200")]
from raftcorr_b200 import ops  # noqa: E402

PEAKS = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0}


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
    f1 = torch.randn(B, D, H, W, device="cuda")
    f2 = torch.randn(B, D, H, W, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    T = sum((H >> l) * (W >> l) for l in range(4))
    big = torch.empty(2 << 30, dtype=torch.uint8, device="cuda")
    ms = timeit(lambda: big.zero_(), reps=10)
    print(f"write-only roof (torch zero_ of 2 GiB): {ms:.3f} ms  {big.numel()/ms/1e6:.0f} GB/s ({big.numel()/ms/1e6/PEAKS['hbm_gbs']:.2f} of measured copy peak)")
    src = torch.empty(1 << 30, dtype=torch.uint8, device="cuda"); dst = torch.empty_like(src)
    ms = timeit(lambda: dst.copy_(src), reps=10)
    print(f"copy roof (1 GiB -> 1 GiB): {ms:.3f} ms  {2*src.numel()/ms/1e6:.0f} GB/s")
    del big, src, dst
    for prec in ("bf16", "tf32", "fp32"):
        ms = timeit(lambda: ops.build_pyramid(f1, f2, 4, prec), reps=5 if prec == "fp32" else 10, flush=flush)
        byts = B * N * T * 4 + 2 * B * N * D * 4
        print(f"build {prec}: {ms:.3f} ms  {2*B*N*N*D/ms/1e9:.1f} TFLOP/s  write-roof {byts/ms/1e6:.0f} GB/s ({byts/ms/1e6/PEAKS['hbm_gbs']:.2f} of measured HBM)")
    pyr = ops.build_pyramid(f1, f2, 4, "bf16")
    ys, xs = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
    grid = torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1)
    for name, c in (("grid", grid), ("sigma4", grid + 4 * torch.randn_like(grid)), ("sigma30", grid + 30 * torch.randn_like(grid))):
        ms = timeit(lambda: ops.lookup(pyr, c, 4), reps=20, flush=flush)
        byts = B * N * 2904
        print(f"lookup {name}: {ms*1e3:.1f} us  {byts/ms/1e6:.0f} GB/s algorithmic ({byts/ms/1e6/PEAKS['hbm_gbs']:.2f} of measured HBM)")
    # kernel-only: 10 lookups with different coords captured in one CUDA graph (no host time)
    cs = [grid + 4 * torch.randn_like(grid) for _ in range(10)]
    ops.lookup(pyr, cs[0], 4)
    g = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            outs = [ops.lookup(pyr, c, 4) for c in cs]
    torch.cuda.current_stream().wait_stream(side)
    ms = timeit(g.replay, reps=10) / len(cs)
    byts = B * N * 2904
    print(f"lookup sigma4 (graph of 10, kernel only): {ms*1e3:.1f} us  {byts/ms/1e6:.0f} GB/s algorithmic ({byts/ms/1e6/PEAKS['hbm_gbs']:.2f} of measured HBM)")
    g2 = torch.cuda.CUDAGraph()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        with torch.cuda.graph(g2, stream=side):
            outs2 = [ops.lookup(pyr, c, 4, channels_last=True) for c in cs]
    torch.cuda.current_stream().wait_stream(side)
    ms = timeit(g2.replay, reps=10) / len(cs)
    print(f"lookup sigma4 channels_last (graph of 10, kernel only): {ms*1e3:.1f} us  {byts/ms/1e6:.0f} GB/s algorithmic ({byts/ms/1e6/PEAKS['hbm_gbs']:.2f} of measured HBM)")
    import time
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(200): ops.lookup(pyr, cs[0], 4)
    t1 = time.perf_counter(); torch.cuda.synchronize()
    print(f"lookup host time per call: {(t1-t0)/200*1e6:.1f} us")


if __name__ == "__main__":
    main()
