"""Randomised shapes through the whole path (build -> lookup, and the on-the-fly block) against the
oracle: ragged widths, odd heights, both radii, channel counts that do and do not fit the
tensor-core layouts.  Small sizes so the CPU oracle stays fast."""
import warnings

import numpy as np
import pytest
import torch

import oracle
from helpers import expected_zero_mask, coords_for, rel_max

warnings.filterwarnings("ignore")
pytestmark = pytest.mark.gpu

CASES = []
_rng = np.random.RandomState(2024)
for _ in range(10):
    CASES.append((int(_rng.randint(1, 4)), int(_rng.choice([8, 16, 24, 40, 64, 12])), int(_rng.randint(16, 41)),
                  int(_rng.randint(16, 71)), int(_rng.choice([3, 4])), str(_rng.choice(["grid", "half", "sigma4", "sigma30"]))))


@pytest.mark.parametrize("B,D,H,W,r,kind", CASES)
def test_random_shape_end_to_end(B, D, H, W, r, kind):
    import corr
    gen = torch.Generator().manual_seed(B * 1000 + D * 100 + H * 10 + W)
    f1 = torch.randn(B, D, H, W, generator=gen)
    f2 = torch.randn(B, D, H, W, generator=gen)
    c = coords_for(kind, B, H, W, gen)
    ref_pyr = oracle.build_pyramid(f1, f2, 4)
    ref = oracle.lookup(ref_pyr, c.numpy(), r)
    with torch.no_grad():
        for precision, tol in (("fp32", 1e-5), ("bf16", 1e-2), ("tf32", 1e-2)):
            blk = corr.CorrBlock(f1.cuda(), f2.cuda(), radius=r, precision=precision)
            out = blk(c.cuda()).cpu().numpy()
            assert np.array_equal(out == 0, expected_zero_mask(blk, c, r, ref)), (precision,)
            assert rel_max(out, ref) <= tol, (precision, rel_max(out, ref))
            # the channels-last lookup reads the fp32 pyramid: identical to the NCHW lookup on that pyramid
            wide = corr.CorrBlock(f1.cuda(), f2.cuda(), radius=r, precision=precision, storage="fp32")(c.cuda())
            nhwc = corr.CorrBlock(f1.cuda(), f2.cuda(), radius=r, precision=precision, channels_last=True)(c.cuda())
            assert torch.equal(nhwc.contiguous(), wide)
            assert rel_max(wide.cpu().numpy(), ref) <= tol
        for precision, tol in (("fp32", 2e-5), ("bf16", 1e-2)):
            alt = corr.AlternateCorrBlock(f1.cuda(), f2.cuda(), radius=r, precision=precision)(c.cuda()).cpu().numpy()
            assert np.array_equal(alt == 0, ref == 0), ("alt", precision)
            assert rel_max(alt, ref) <= tol, ("alt", precision, rel_max(alt, ref))


@pytest.mark.parametrize("levels,r", [(1, 4), (2, 3), (3, 2), (4, 1), (2, 0), (1, 2)])
def test_other_level_counts_and_radii(levels, r):
    """num_levels and radius are runtime arguments of the reference API (corr.py:13); anything but (4, 3|4)
    goes through the generic-stride kernels."""
    import corr
    B, D, H, W = 2, 24, 20, 36
    gen = torch.Generator().manual_seed(levels * 10 + r)
    f1 = torch.randn(B, D, H, W, generator=gen)
    f2 = torch.randn(B, D, H, W, generator=gen)
    c = coords_for("sigma4", B, H, W, gen)
    ref = oracle.lookup(oracle.build_pyramid(f1, f2, levels), c.numpy(), r)
    with torch.no_grad():
        blk = corr.CorrBlock(f1.cuda(), f2.cuda(), num_levels=levels, radius=r, precision="fp32")
        assert len(blk.corr_pyramid) == levels
        out = blk(c.cuda()).cpu().numpy()
        assert out.shape == (B, levels * (2 * r + 1) ** 2, H, W)
        assert np.array_equal(out == 0, ref == 0) and rel_max(out, ref) <= 1e-5
        alt = corr.AlternateCorrBlock(f1.cuda(), f2.cuda(), num_levels=levels, radius=r)(c.cuda()).cpu().numpy()
        assert np.array_equal(alt == 0, ref == 0) and rel_max(alt, ref) <= 1e-2
    # and with gradients
    a, b = f1.cuda().requires_grad_(True), f2.cuda().requires_grad_(True)
    out = corr.CorrBlock(a, b, num_levels=levels, radius=r, precision="fp32")(c.cuda())
    out.sum().backward()
    ra, rb = f1.clone().requires_grad_(True), f2.clone().requires_grad_(True)
    from oracle import torch_restatement as tr
    tr.TorchCorrBlock(ra, rb, levels, r)(c).sum().backward()
    for got, want in ((a.grad, ra.grad), (b.grad, rb.grad)):
        want = want.numpy()
        assert np.allclose(got.cpu().numpy(), want, rtol=1e-3, atol=1e-4 * np.abs(want).max())
