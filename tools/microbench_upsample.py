"""Time the convex upsampling kernel (and its backward) against the reference's op sequence on the same GPU.
Inputs rotate over 4 buffers (4 x 130 MB > L2) so no launch finds its mask in L2."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "pytorch-raft_b200"))
import upsample  # noqa: E402
from oracle import torch_restatement as tr  # noqa: E402
from raftcorr_b200 import ops  # noqa: E402


def timeit(fn, n=40, warm=5):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(n):
        fn(i)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3


def main():
    B, H, W = 8, 55, 128
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6453.4) if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6453.4
    masks = [torch.randn(B, 576, H, W, device="cuda") for _ in range(4)]
    flows = [torch.randn(B, 2, H, W, device="cuda") for _ in range(4)]
    gup = torch.randn(B, 2, 8 * H, 8 * W, device="cuda")
    px = B * H * W
    fwd_bytes = px * (2304 + 8 + 512)
    bwd_bytes = px * (2304 + 8 + 512 + 2304 + 8)
    with torch.no_grad():
        t = timeit(lambda i: upsample.upsample_flow(flows[i % 4], masks[i % 4]))
        print(f"upsample fwd ours   {t:8.1f} us  {fwd_bytes / t / 1e3:7.0f} GB/s  frac {fwd_bytes / t / 1e3 / peak:.2f}")
        t = timeit(lambda i: tr.upsample_flow(flows[i % 4], masks[i % 4]))
        print(f"upsample fwd torch  {t:8.1f} us")
        t = timeit(lambda i: ops.upsample_flow_backward(flows[i % 4], masks[i % 4], gup))
        print(f"upsample bwd ours   {t:8.1f} us  {bwd_bytes / t / 1e3:7.0f} GB/s  frac {bwd_bytes / t / 1e3 / peak:.2f}")
        t = timeit(lambda i: upsample.upflow8(flows[i % 4]))
        print(f"upflow8 ours        {t:8.1f} us")
        t = timeit(lambda i: tr.upflow8(flows[i % 4]))
        print(f"upflow8 torch       {t:8.1f} us")

    def torch_bwd(i):
        a, b = flows[i % 4].clone().requires_grad_(True), masks[i % 4].clone().requires_grad_(True)
        tr.upsample_flow(a, b).backward(gup)
    t = timeit(torch_bwd, n=10, warm=2)
    print(f"upsample fwd+bwd torch {t:8.1f} us")


if __name__ == "__main__":
    main()
