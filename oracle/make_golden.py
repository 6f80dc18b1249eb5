"""Generate tests/golden/*.npz by running the UNMODIFIED reference in the build container.

    python oracle/make_golden.py            # needs /root/reference (read-only) and torch (CPU)

The reference (zhaoyuzhi/PyTorch-RAFT) ships no tests or golden vectors for its correlation path
(SURVEY.md §4), so the vectors written here — outputs of the reference's own ``CorrBlock``
(core/corr.py:12-60) imported from /root/reference on seeded inputs — are what pins the oracle.
/root/reference does not exist on the GPU box; the committed .npz files travel instead.

Cases
  tiny      B=1 D=8   16x18  r=4 L=4  inputs + full pyramid + outputs for 4 coord distributions,
                                      plus autograd grads (d fmap1, d fmap2, d coords)
  odd_r3    B=2 D=8   17x23  r=3 L=4  inputs + levels 1..3 + outputs (odd sizes, sqrt(8) scale, r=3)
  upsample  N=2 7x11: RAFT.upsample_flow (raft.py:72-83) output and autograd grads, upflow8 output
  cfg1      B=1 D=256 46x62  r=4 L=4  BASELINE.json configs[0]; inputs regenerated from the seed
                                      (torch CPU generator), outputs stored as strided samples +
                                      float64 checksums + zero counts
"""
from __future__ import annotations

import os
import sys
import warnings

import numpy as np
import torch

REF = "/root/reference/core"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")

warnings.filterwarnings("ignore")


def coords_for(kind: str, B: int, H: int, W: int, gen: torch.Generator) -> torch.Tensor:
    """SURVEY.md §8(d) coordinate distributions. ch0 = x, ch1 = y (utils.py:74-77)."""
    ys, xs = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    grid = torch.stack([xs, ys], dim=0).float()[None].repeat(B, 1, 1, 1)
    if kind == "grid":      # iteration-0 state: exact integers (floor-flip semantics)
        return grid
    if kind == "half":      # exact half-integers and quarter steps
        return grid + 0.5 - 0.25 * (torch.arange(W) % 3).float().view(1, 1, 1, W)
    if kind == "sigma4":
        return grid + 4.0 * torch.randn(B, 2, H, W, generator=gen)
    if kind == "sigma30":   # many windows partly / wholly outside the image
        return grid + 30.0 * torch.randn(B, 2, H, W, generator=gen)
    raise ValueError(kind)


def seeded_inputs(seed: int, B: int, D: int, H: int, W: int):
    gen = torch.Generator().manual_seed(seed)
    f1 = torch.randn(B, D, H, W, generator=gen)
    f2 = torch.randn(B, D, H, W, generator=gen)
    return f1, f2, gen


def checksum(a: np.ndarray):
    a64 = a.astype(np.float64)
    return np.array([a64.sum(), np.abs(a64).sum(), float((a == 0).sum())], dtype=np.float64)


def main() -> None:
    sys.path.insert(0, REF)
    from corr import CorrBlock  # the reference, unmodified

    os.makedirs(OUT, exist_ok=True)

    # ---- tiny: everything stored -------------------------------------------------------------
    B, D, H, W, r, L = 1, 8, 16, 18, 4, 4
    f1, f2, gen = seeded_inputs(1234, B, D, H, W)
    blk = CorrBlock(f1, f2, num_levels=L, radius=r)
    rec = {"fmap1": f1.numpy(), "fmap2": f2.numpy(), "radius": np.int32(r), "levels": np.int32(L)}
    for l, p in enumerate(blk.corr_pyramid):
        rec[f"pyr{l}"] = p.numpy()
    for kind in ("grid", "half", "sigma4", "sigma30"):
        c = coords_for(kind, B, H, W, gen)
        rec[f"coords_{kind}"] = c.numpy()
        rec[f"out_{kind}"] = blk(c).numpy()
    # autograd through the reference: fmaps and coords all require grad
    g1 = f1.clone().requires_grad_(True)
    g2 = f2.clone().requires_grad_(True)
    gc = torch.from_numpy(rec["coords_sigma4"]).clone().requires_grad_(True)
    out = CorrBlock(g1, g2, num_levels=L, radius=r)(gc)
    up = torch.randn(out.shape, generator=gen)
    out.backward(up)
    rec.update(grad_out=up.numpy(), grad_fmap1=g1.grad.numpy(), grad_fmap2=g2.grad.numpy(),
               grad_coords=gc.grad.numpy())
    np.savez(os.path.join(OUT, "corr_tiny.npz"), **rec)

    # ---- odd sizes, r=3, inexact sqrt(D) ------------------------------------------------------
    B, D, H, W, r, L = 2, 8, 17, 23, 3, 4
    f1, f2, gen = seeded_inputs(1235, B, D, H, W)
    blk = CorrBlock(f1, f2, num_levels=L, radius=r)
    c = coords_for("sigma4", B, H, W, gen)
    rec = {"fmap1": f1.numpy(), "fmap2": f2.numpy(), "radius": np.int32(r), "levels": np.int32(L),
           "coords_sigma4": c.numpy(), "out_sigma4": blk(c).numpy()}
    for l in range(1, L):
        rec[f"pyr{l}"] = blk.corr_pyramid[l].numpy()
    rec["pyr0_sample"] = blk.corr_pyramid[0].numpy()[::7]
    np.savez(os.path.join(OUT, "corr_odd_r3.npz"), **rec)

    # ---- BASELINE.json configs[0]: 46x62, D=256 (inputs regenerated from the seed) -----------
    B, D, H, W, r, L = 1, 256, 46, 62, 4, 4
    f1, f2, gen = seeded_inputs(1236, B, D, H, W)
    blk = CorrBlock(f1, f2, num_levels=L, radius=r)
    rec = {"seed": np.int64(1236), "shape": np.array([B, D, H, W, r, L], dtype=np.int64),
           "fmap1_probe": f1.numpy()[0, :4, :2, :3], "fmap2_probe": f2.numpy()[0, :4, :2, :3]}
    for l, p in enumerate(blk.corr_pyramid):
        a = p.numpy()
        rec[f"pyr{l}_sum"] = checksum(a)
        rec[f"pyr{l}_sample"] = a[::97, 0, ::3, ::5]
    for kind in ("grid", "sigma4", "sigma30"):
        c = coords_for(kind, B, H, W, gen)
        o = blk(c).numpy()
        rec[f"coords_{kind}"] = c.numpy()
        rec[f"out_{kind}_sum"] = checksum(o)
        rec[f"out_{kind}_sample"] = o[:, :, ::5, ::7]
        rec[f"out_{kind}_zero_rows"] = (o == 0).sum(axis=(0, 2, 3)).astype(np.int32)
    np.savez(os.path.join(OUT, "corr_cfg1.npz"), **rec)
    # ---- the step after the path: RAFT.upsample_flow / upflow8 (raft.py:72-83, utils.py:80-82) ----
    import argparse
    from raft import RAFT  # the reference, unmodified
    from utils.utils import upflow8
    model = RAFT(argparse.Namespace(small=False, mixed_precision=False, alternate_corr=False))
    gen = torch.Generator().manual_seed(1237)
    N, H, W = 2, 7, 11
    flow = (4 * torch.randn(N, 2, H, W, generator=gen)).requires_grad_(True)
    mask = (3 * torch.randn(N, 576, H, W, generator=gen)).requires_grad_(True)
    up = model.upsample_flow(flow, mask)
    gup = torch.randn(up.shape, generator=gen)
    up.backward(gup)
    np.savez(os.path.join(OUT, "upsample_tiny.npz"), flow=flow.detach().numpy(), mask=mask.detach().numpy(),
             up=up.detach().numpy(), grad_up=gup.numpy(), grad_flow=flow.grad.numpy(),
             grad_mask=mask.grad.numpy(), up8=upflow8(flow.detach()).numpy())
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
