"""bench.py's inference step at batch 8: eager launches vs one CUDA-graph replay (raftcorr_b200.graphed)."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200")]
import bench
import corr as raftcorr_dropin
import upsample as raftcorr_upsample
import raftcorr_b200
dev = torch.device("cuda", 0)
model = bench.make_model(dev)
h1, h2 = bench.synthetic_frames(bench.BATCH_PER_GPU, seed=100, pin=True)
d1, d2 = h1.to(dev), h2.to(dev)
factory = lambda f1, f2, r: raftcorr_dropin.CorrBlock(f1, f2, num_levels=4, radius=r)
def fwd(a, b):
    with torch.no_grad():
        return model(a, b, factory, iters=bench.ITERS, upsampler=raftcorr_upsample.upsample_flow)[1]
def timeit(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
eager = timeit(lambda: fwd(d1, d2))
g = raftcorr_b200.graphed(fwd, (d1, d2))
out_g = g(d1, d2).clone()
out_e = fwd(d1, d2)
print("graph == eager:", torch.equal(out_g, out_e), float((out_g - out_e).abs().max()))
rep = timeit(lambda: g(d1, d2))
print(f"batch 8: eager {eager:.2f} ms per step, graph replay {rep:.2f} ms per step")
