"""Does cudaLimitMaxL2FetchGranularity (default 64 B) change the lookup's time? (GPU box only)"""
import os, sys, ctypes, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200"), os.path.join(ROOT, "tools")]
from raftcorr_b200 import ops
from microbench import timeit
rt = ctypes.CDLL("libcudart.so.12")
def get():
    v = ctypes.c_size_t(0); rc = rt.cudaDeviceGetLimit(ctypes.byref(v), 5); return rc, v.value
B, D, H, W = 8, 256, 55, 128
f1 = torch.randn(B, D, H, W, device="cuda"); f2 = torch.randn(B, D, H, W, device="cuda")
pyr = ops.build_pyramid(f1, f2, 4, "bf16")
ys, xs = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
grid = torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1)
cs = [grid + 4 * torch.randn_like(grid) for _ in range(10)]
def run(tag):
    ops.lookup(pyr, cs[0], 4)
    g = torch.cuda.CUDAGraph(); side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            outs = [ops.lookup(pyr, c, 4) for c in cs]
    torch.cuda.current_stream().wait_stream(side)
    ms = timeit(g.replay, reps=10) / len(cs)
    mb = timeit(lambda: ops.build_pyramid(f1, f2, 4, "bf16"), reps=5)
    a = torch.randn(8192, 8192, device="cuda", dtype=torch.bfloat16)
    mm = timeit(lambda: a @ a, reps=5)
    print(f"{tag}: limit={get()} lookup {ms*1e3:.1f} us  build {mb:.3f} ms  bf16 matmul 8192^3 {mm:.3f} ms")
run("default")
for gran in (32, 128, 64):
    torch.cuda.synchronize()
    rc = rt.cudaDeviceSetLimit(5, ctypes.c_size_t(gran))
    run(f"set {gran} (rc={rc})")
