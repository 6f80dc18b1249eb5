"""Lookup + the 1x1 convolution that consumes it (BasicMotionEncoder.convc1, update.py:83,90).
The GEMM runs TF32 products with fp32 accumulation, which is what cuDNN uses for this convolution by default:
the bar is the bf16/TF32 one, 1e-2 of max|ref| (measured ~1e-3), against an exact-fp32 torch convolution of
the oracle's lookup."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

import oracle
from raftcorr_b200 import ops
from helpers import coords_for, rel_max

pytestmark = pytest.mark.gpu


def _exact_conv(x, w, b):
    with torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
        old = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = False
        try:
            return F.relu(F.conv2d(x, w, b))
        finally:
            torch.backends.cuda.matmul.allow_tf32 = old


@pytest.mark.parametrize("B,H,W,r,cout", [(2, 24, 40, 4, 256), (1, 46, 62, 4, 256), (2, 23, 44, 3, 96), (1, 16, 20, 4, 40), (1, 16, 20, 4, 320)])
def test_conv1x1_of_the_lookup_matches_torch(B, H, W, r, cout):
    import corr
    gen = torch.Generator().manual_seed(B * 100 + H)
    D = 64
    f1 = torch.randn(B, D, H, W, generator=gen)
    f2 = torch.randn(B, D, H, W, generator=gen)
    c = coords_for("sigma4", B, H, W, gen)
    C = 4 * (2 * r + 1) ** 2
    w = (torch.randn(cout, C, 1, 1, generator=gen) / C ** 0.5).cuda()
    b = torch.randn(cout, generator=gen).cuda()
    ref_lookup = torch.from_numpy(oracle.lookup(oracle.build_pyramid(f1, f2, 4), c.numpy(), r)).cuda()
    ref = _exact_conv(ref_lookup, w, b).cpu().numpy()
    with torch.no_grad():
        blk = corr.CorrBlock(f1.cuda(), f2.cuda(), radius=r, precision="fp32", channels_last=True)
        lk = blk(c.cuda())
        assert lk.is_contiguous(memory_format=torch.channels_last)
        two_calls = corr.conv1x1_relu(lk, w, b)
        one_call = blk.lookup_conv1x1(c.cuda(), w, b)
        nchw_in = corr.conv1x1_relu(lk.contiguous(), w, b)     # any layout is accepted (converted first)
    assert two_calls.shape == (B, cout, H, W) and two_calls.is_contiguous()
    assert torch.equal(two_calls, one_call) and torch.equal(two_calls, nchw_in)
    assert rel_max(two_calls.cpu().numpy(), ref) <= 1e-2
    assert rel_max(two_calls.cpu().numpy(), ref) <= 3e-3        # TF32 is far inside the bar
    assert (two_calls >= 0).all()
    no_act = ops.conv1x1_channels_last(lk, w, None, relu=False).cpu().numpy()
    with torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
        want = F.conv2d(ref_lookup, w).cpu().numpy()
    assert rel_max(no_act, want) <= 3e-3


def test_nan_samples_stay_nan_through_the_relu():
    """A degenerate level (W_l == 1) makes the reference's lookup NaN; torch.relu keeps NaN."""
    import corr
    gen = torch.Generator().manual_seed(3)
    f1, f2 = torch.randn(1, 16, 8, 12, generator=gen).cuda(), torch.randn(1, 16, 8, 12, generator=gen).cuda()
    c = coords_for("sigma4", 1, 8, 12, gen).cuda()
    w = torch.randn(8, 324, 1, 1, generator=gen).cuda()
    with torch.no_grad():
        lk = corr.CorrBlock(f1, f2, channels_last=True)(c)
        got = corr.conv1x1_relu(lk, w)
    assert torch.isnan(lk).any()
    assert torch.equal(torch.isnan(got), torch.isnan(F.relu(F.conv2d(lk, w))))


def test_unsupported_shapes_say_so():
    import corr
    f = torch.randn(1, 16, 9, 9).cuda()          # H*W = 81 is not a multiple of 4
    blk = corr.CorrBlock(f, f, channels_last=True)
    c = coords_for("grid", 1, 9, 9, torch.Generator()).cuda()
    with pytest.raises(RuntimeError, match="% 4"):
        corr.conv1x1_relu(blk(c), torch.randn(8, 324, 1, 1).cuda())
    with pytest.raises(RuntimeError, match="inference"):
        corr.conv1x1_relu(blk(c), torch.randn(8, 324, 1, 1).cuda().requires_grad_(True))


@pytest.mark.parametrize("B,H,W,r,cout", [(2, 24, 40, 4, 256), (1, 55, 128, 4, 256), (2, 23, 44, 3, 96)])
def test_conv1x1_on_the_default_nchw_lookup(B, H, W, r, cout):
    """The default inference lookup (16-bit pyramid, plain NCHW output) feeds the same GEMM as an MN-major operand
    (raftcorr_conv1x1_nchw_fwd): identical to the channels-last variant on the same values, and within the TF32 bar
    of an exact fp32 convolution of that lookup."""
    import corr
    gen = torch.Generator().manual_seed(B * 100 + W)
    D = 64
    f1 = torch.randn(B, D, H, W, generator=gen).cuda()
    f2 = torch.randn(B, D, H, W, generator=gen).cuda()
    c = coords_for("sigma4", B, H, W, gen).cuda()
    C = 4 * (2 * r + 1) ** 2
    w = (torch.randn(cout, C, 1, 1, generator=gen) / C ** 0.5).cuda()
    b = torch.randn(cout, generator=gen).cuda()
    with torch.no_grad():
        lk = corr.CorrBlock(f1, f2, radius=r)(c)                      # bf16 blocked pyramid, lookup16
        assert lk.is_contiguous()
        got = corr.conv1x1_relu(lk, w, b)
        same = corr.conv1x1_relu(lk.contiguous(memory_format=torch.channels_last), w, b)
    assert got.shape == (B, cout, H, W) and got.is_contiguous()
    assert torch.equal(got, same)
    ref = _exact_conv(lk, w, b)
    assert rel_max(got.cpu().numpy(), ref.cpu().numpy()) <= 3e-3
