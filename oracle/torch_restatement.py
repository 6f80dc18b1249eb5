"""torch-op restatement of the reference CorrBlock — TEST INFRASTRUCTURE ONLY.

The reference's correlation path (``core/corr.py:12-60``, ``core/utils/utils.py:57-65``) is
nothing but a sequence of ATen ops; this module issues the same ops in the same order so that

* ``bench.py`` can time "the reference's CPU implementation" on the GPU box's host cores
  (``/root/reference`` does not exist there), and
* GPU tests can use ATen's own CUDA kernels as a second oracle (``coord_mode=1`` of the
  product kernels follows ATen-CUDA's reciprocal-multiply division).

It is pinned against the unmodified reference by ``tests/test_oracle.py`` via the vectors
``oracle/make_golden.py`` wrote.  It also provides autograd (d fmaps, d coords) for the
backward-parity tests, exactly as the reference obtains its gradients.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F


def all_pairs(fmap1: torch.Tensor, fmap2: torch.Tensor) -> torch.Tensor:
    """corr.py:52-60: (B,D,H,W)x2 -> (B,H,W,1,H,W), divided by sqrt(D) as a float tensor."""
    B, D, H, W = fmap1.shape
    a = fmap1.reshape(B, D, H * W)
    b = fmap2.reshape(B, D, H * W)
    vol = torch.matmul(a.transpose(1, 2), b).view(B, H, W, 1, H, W)
    return vol / torch.sqrt(torch.tensor(D).float())


def build_pyramid(fmap1: torch.Tensor, fmap2: torch.Tensor, num_levels: int = 4):
    """corr.py:13-27: level 0 is the volume folded to (B*N,1,H,W); each next level is a 2x2 mean."""
    vol = all_pairs(fmap1, fmap2)
    B, H, W = vol.shape[:3]
    level = vol.reshape(B * H * W, 1, H, W)
    pyramid = [level]
    while len(pyramid) < num_levels:
        level = F.avg_pool2d(level, 2, stride=2)
        pyramid.append(level)
    return pyramid


def _sample_pixels(img: torch.Tensor, xy: torch.Tensor) -> torch.Tensor:
    """utils.py:57-65: pixel coordinates -> [-1,1] -> grid_sample(align_corners=True)."""
    Hl, Wl = img.shape[-2:]
    gx, gy = xy.split([1, 1], dim=-1)
    gx = 2 * gx / (Wl - 1) - 1
    gy = 2 * gy / (Hl - 1) - 1
    return F.grid_sample(img, torch.cat([gx, gy], dim=-1), align_corners=True)


def lookup(pyramid, coords: torch.Tensor, radius: int = 4) -> torch.Tensor:
    """corr.py:29-50: (B,2,H,W) coords -> (B, L*(2r+1)^2, H, W)."""
    B, _, H, W = coords.shape
    centre = coords.permute(0, 2, 3, 1).reshape(B * H * W, 1, 1, 2)
    taps = torch.linspace(-radius, radius, 2 * radius + 1, device=coords.device)
    # first window axis moves x, second moves y (corr.py:37-43)
    window = torch.stack(torch.meshgrid(taps, taps, indexing="ij"), dim=-1)
    window = window.view(1, 2 * radius + 1, 2 * radius + 1, 2)
    per_level = []
    for l, level in enumerate(pyramid):
        sampled = _sample_pixels(level, centre / 2 ** l + window)
        per_level.append(sampled.view(B, H, W, -1))
    return torch.cat(per_level, dim=-1).permute(0, 3, 1, 2).contiguous().float()


class TorchCorrBlock:
    """Same constructor / call signature as the reference CorrBlock (corr.py:12-50)."""

    def __init__(self, fmap1, fmap2, num_levels: int = 4, radius: int = 4):
        self.num_levels = num_levels
        self.radius = radius
        self.corr_pyramid = build_pyramid(fmap1, fmap2, num_levels)

    def __call__(self, coords):
        return lookup(self.corr_pyramid, coords, self.radius)


def alt_corr_block_math(fmap1, fmap2, coords, num_levels: int = 4, radius: int = 4):
    """What AlternateCorrBlock (corr.py:63-91) computes, expressed without the CUDA extension:
    correlate fmap1 with the 2x2-pooled fmap2 pyramid and sample bilinearly at floor(coords/2^l)
    directly (no normalise round trip; correlation_kernel.cu:66-76).  Differentiable."""
    B, D, H, W = fmap1.shape
    r = radius
    rd = 2 * r + 1
    q = fmap1.reshape(B, D, H * W).transpose(1, 2)  # (B,N,D)
    taps = torch.arange(-r, r + 1, device=coords.device, dtype=coords.dtype)
    outs = []
    f2 = fmap2
    for l in range(num_levels):
        Hl, Wl = f2.shape[-2:]
        vol = torch.matmul(q, f2.reshape(B, D, Hl * Wl)).view(B, H * W, Hl, Wl)  # unscaled
        cx = (coords[:, 0] / 2 ** l).reshape(B, H * W, 1)
        cy = (coords[:, 1] / 2 ** l).reshape(B, H * W, 1)
        x0 = torch.floor(cx)
        y0 = torch.floor(cy)
        dx = (cx - x0)[..., None]
        dy = (cy - y0)[..., None]
        xs = x0.long() + taps.long().view(1, 1, rd)  # (B,N,rd) moves with i
        ys = y0.long() + taps.long().view(1, 1, rd)  # (B,N,rd) moves with j

        def gather(yy, xx):
            ok = ((yy >= 0) & (yy < Hl))[:, :, None, :] & ((xx >= 0) & (xx < Wl))[:, :, :, None]
            idx = yy.clamp(0, Hl - 1)[:, :, None, :] * Wl + xx.clamp(0, Wl - 1)[:, :, :, None]
            v = torch.gather(vol.reshape(B, H * W, Hl * Wl), 2, idx.reshape(B, H * W, rd * rd))
            return v.view(B, H * W, rd, rd) * ok

        out = ((1 - dy) * (1 - dx) * gather(ys, xs) + (1 - dy) * dx * gather(ys, xs + 1)
               + dy * (1 - dx) * gather(ys + 1, xs) + dy * dx * gather(ys + 1, xs + 1))
        outs.append(out.reshape(B, H, W, rd * rd))
        f2 = F.avg_pool2d(f2, 2, stride=2)
    corr = torch.cat(outs, dim=-1).permute(0, 3, 1, 2).contiguous()
    return corr / math.sqrt(D)


def upsample_flow(flow: torch.Tensor, mask: torch.Tensor) -> torch.Tensor:
    """RAFT.upsample_flow restated op for op (core/raft.py:72-83): the op sequence the reference runs after
    every iteration; used to time "what the reference does" and as the autograd oracle for the backward."""
    N, _, H, W = flow.shape
    weights = torch.softmax(mask.view(N, 1, 9, 8, 8, H, W), dim=2)                      # raft.py:75-76
    patches = F.unfold(8 * flow, [3, 3], padding=1).view(N, 2, 9, 1, 1, H, W)           # raft.py:78-79
    fine = torch.sum(weights * patches, dim=2)                                          # raft.py:81
    return fine.permute(0, 1, 4, 2, 5, 3).reshape(N, 2, 8 * H, 8 * W)                   # raft.py:82-83


def upflow8(flow: torch.Tensor) -> torch.Tensor:
    """core/utils/utils.py:80-82."""
    size = (8 * flow.shape[2], 8 * flow.shape[3])
    return 8 * F.interpolate(flow, size=size, mode="bilinear", align_corners=True)
