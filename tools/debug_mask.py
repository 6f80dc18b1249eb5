import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200"), os.path.join(ROOT, "tests")]
from raftcorr_b200 import ops
from oracle import torch_restatement as tr
from helpers import coords_for
B, H, W = 8, 55, 128
gen = torch.Generator().manual_seed(7)
pyr = ops.Pyramid(B, H, W, 4, "cuda", zero=True)
for l in range(4):
    pyr.storage[l][:, :, : pyr.desc[l].W].normal_()
levels = [v.contiguous() for v in pyr.views()]
for kind in ("grid", "sigma4", "sigma30"):
    c = coords_for(kind, B, H, W, gen).cuda()
    out = ops.lookup(pyr, c, 4, "aten_cuda")
    ref = tr.lookup(levels, c, 4)
    bad = (out == 0) != (ref == 0)
    print(kind, "mask mismatches", int(bad.sum()), "of", bad.numel(), "maxdiff", float((out - ref).abs().max()))
    idx = bad.nonzero()[:8]
    for b, ch, y, x in idx.tolist():
        l, k = divmod(ch, 81); i, j = divmod(k, 9)
        print("  b,y,x", b, y, x, "level", l, "i,j", i, j, "coords", c[b, :, y, x].tolist(), "ours", float(out[b, ch, y, x]), "ref", float(ref[b, ch, y, x]))
    # per-batch check: does the reference itself change when run per batch element?
    ref1 = tr.lookup([lv[:H * W] for lv in levels], c[:1], 4)
    print("   ref batch0 alone equals batched:", bool(torch.equal(ref1, ref[:1])))
