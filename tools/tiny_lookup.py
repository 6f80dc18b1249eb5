import ctypes, os, sys, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200"), os.path.join(ROOT, "tests")]
from raftcorr_b200 import ops, _capi
from helpers import pyramid_from_numpy
g = np.load(os.path.join(ROOT, "tests/golden/corr_tiny.npz"))
pyr = pyramid_from_numpy([g[f"pyr{l}"] for l in range(4)], 1, 16, 18)
for kind in ("sigma4", "grid"):
    out = ops.lookup(pyr, torch.from_numpy(g[f"coords_{kind}"]).cuda(), 4, "aten_cpu")
    torch.cuda.synchronize()
    o = out.cpu().numpy(); ref = g[f"out_{kind}"]
    print(kind, "bit exact:", np.array_equal(o, ref), "n mismatch", int((o != ref).sum()), "of", o.size)
    if hasattr(_capi.lib(), "raftcorr_debug_wait_timeouts"):
        buf = (ctypes.c_uint64 * 8)()
        _capi.lib().raftcorr_debug_wait_timeouts(buf)
        print("  timeouts:", buf[0], [(hex(v)) for v in buf[1:5]])
    bad = np.argwhere(o != ref)
    if len(bad):
        print("  first mismatches (b,ch,y,x):", bad[:6].tolist())
        chans = np.unique(bad[:, 1]); print("  channels affected:", len(chans), "levels", np.unique(chans // 81), "xs", np.unique(bad[:, 3])[:20], "ys", np.unique(bad[:,2])[:20])
