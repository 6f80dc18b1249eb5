"""Randomised shapes through the whole path (build -> lookup, and the on-the-fly block) against the
oracle: ragged widths, odd heights, both radii, channel counts that do and do not fit the
tensor-core layouts.  Small sizes so the CPU oracle stays fast."""
import warnings

import numpy as np
import pytest
import torch

import oracle
from helpers import coords_for, rel_max

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
            assert np.array_equal(out == 0, ref == 0), (precision,)
            assert rel_max(out, ref) <= tol, (precision, rel_max(out, ref))
            nhwc = corr.CorrBlock(f1.cuda(), f2.cuda(), radius=r, precision=precision, channels_last=True)(c.cuda())
            assert np.array_equal(nhwc.cpu().numpy(), out)
        for precision, tol in (("fp32", 2e-5), ("bf16", 1e-2)):
            alt = corr.AlternateCorrBlock(f1.cuda(), f2.cuda(), radius=r, precision=precision)(c.cuda()).cpu().numpy()
            assert np.array_equal(alt == 0, ref == 0), ("alt", precision)
            assert rel_max(alt, ref) <= tol, ("alt", precision, rel_max(alt, ref))
