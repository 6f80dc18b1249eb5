"""Randomised stress of the 16-bit path on the GPU: many shapes, flows and layouts; lookup16 on the blocked pyramid
against lookup_fwd (fp32 pyramid kernels, themselves pinned to the oracle) on the same stored values, bit for bit,
and build16 against the bf16 rounding of the fp32-pyramid build.   usage: stress16.py [cases=200] [seed=0]"""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200")]
from raftcorr_b200 import ops

n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 200
rng = np.random.RandomState(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
bad = 0
for case in range(n_cases):
    B = int(rng.randint(1, 5)); H = int(rng.randint(8, 80)); W = int(rng.randint(8, 200)); r = int(rng.choice([3, 4]))
    D = int(rng.choice([8, 16, 64])); kind = str(rng.choice(["grid", "smooth", "sigma1", "sigma8", "sigma40", "jumps"]))
    il = rng.choice(["auto", "plain", "forced"])
    gen = torch.Generator(device="cuda").manual_seed(case)
    f1 = torch.randn(B, D, H, W, device="cuda", generator=gen); f2 = torch.randn(B, D, H, W, device="cuda", generator=gen)
    interleave = None if il == "auto" else False if il == "plain" else ((2, 2, 0, 0) if (B * H * W) % 4 == 0 else None)
    pyr16 = ops.build_pyramid(f1, f2, 4, "bf16", storage="bf16")
    if interleave is not None:          # same values in another layout
        pyr16 = ops.Pyramid16.from_dense([pyr16.dense(l).unsqueeze(1) for l in range(4)], B, H, W, interleave=interleave)
    wide = ops.build_pyramid(f1, f2, 4, "bf16", storage="fp32")
    ok_build = all(torch.equal(pyr16.dense(l), wide.views()[l][:, 0].to(torch.bfloat16).float()) for l in range(4))
    for l in range(4):                  # the fp32 pyramid now holds exactly the stored values
        w_ = wide.desc[l].W
        wide.storage[l][:, :, :w_].copy_(pyr16.dense(l))
    ys, xs = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
    grid = torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1)
    if kind == "grid":
        c = grid
    elif kind == "smooth":
        low = 6 * torch.randn(B, 2, max(H // 8, 2), max(W // 8, 2), device="cuda", generator=gen)
        c = grid + torch.nn.functional.interpolate(low, size=(H, W), mode="bilinear", align_corners=True)
    elif kind == "jumps":               # piecewise constant flow: groups that straddle a motion boundary
        low = 12 * torch.randn(B, 2, max(H // 6, 2), max(W // 6, 2), device="cuda", generator=gen)
        c = grid + torch.nn.functional.interpolate(low, size=(H, W), mode="nearest")
    else:
        c = grid + float(kind[5:]) * torch.randn(B, 2, H, W, device="cuda", generator=gen)
    for mode in ("aten_cpu", "aten_cuda", "direct"):
        a = ops.lookup(pyr16, c, r, mode); b = ops.lookup(wide, c, r, mode)
        same = torch.equal(torch.nan_to_num(a), torch.nan_to_num(b)) and torch.equal(torch.isnan(a), torch.isnan(b))
        if not (same and ok_build):
            bad += 1
            print(f"MISMATCH case {case}: B={B} D={D} H={H} W={W} r={r} {kind} {il} {mode} build_ok={ok_build} "
                  f"max|d|={float((a - b).abs().max()):.3g} n={int((a != b).sum())}", flush=True)
print(f"{n_cases} cases, {bad} mismatches")
