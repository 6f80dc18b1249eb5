"""Shared helpers for the GPU parity tests (the oracle is only ever the checker)."""
import numpy as np
import torch

from raftcorr_b200 import ops


def pyramid_from_numpy(levels, B, H, W, device="cuda"):
    """Upload reference-shaped levels (B*N,1,H_l,W_l) into the library's padded storage."""
    pyr = ops.Pyramid(B, H, W, len(levels), device, zero=True)
    for l, lv in enumerate(levels):
        t = torch.as_tensor(np.ascontiguousarray(lv), device=device).reshape(B * H * W, lv.shape[-2], lv.shape[-1])
        pyr.storage[l][:, :, : lv.shape[-1]].copy_(t)
    return pyr


def pyramid_to_numpy(pyr):
    return [v.contiguous().cpu().numpy() for v in pyr.views()]


def coords_for(kind, B, H, W, gen):
    ys, xs = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    grid = torch.stack([xs, ys], dim=0).float()[None].repeat(B, 1, 1, 1)
    if kind == "grid":
        return grid
    if kind == "half":
        return grid + 0.5 - 0.25 * (torch.arange(W) % 3).float().view(1, 1, 1, W)
    sigma = {"sigma4": 4.0, "sigma30": 30.0}[kind]
    return grid + sigma * torch.randn(B, 2, H, W, generator=gen)


def rel_max(a, ref):
    """max|a-ref| / max|ref| — the normalisation SURVEY.md §8(d) prescribes."""
    a, ref = np.asarray(a, dtype=np.float64), np.asarray(ref, dtype=np.float64)
    return float(np.abs(a - ref).max() / max(np.abs(ref).max(), 1e-30))


def expected_zero_mask(block, coords, radius, ref):
    """Which outputs must be exactly zero.  For the fp32 pyramid that is the reference's own zero pattern (all of
    it zero padding: fp32 sums do not cancel by accident).  With 16-bit storage two neighbouring pixels can round
    to exact opposites, so a tap at a half-integer position may cancel to 0.0 where the fp32 reference keeps a tiny
    residue; the padding pattern is then pinned by the oracle applied to the values the block actually stores."""
    import oracle
    if isinstance(block._holder["pyr"], ops.Pyramid16):
        stored = [v.cpu().numpy() for v in block.corr_pyramid]
        c = coords.detach().cpu().numpy() if torch.is_tensor(coords) else np.asarray(coords)
        return oracle.lookup(stored, c, radius) == 0
    return np.asarray(ref) == 0
