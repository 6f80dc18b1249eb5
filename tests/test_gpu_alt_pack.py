"""``raftcorr_alt_pack``: the operands of the fused on-the-fly path in one launch — the avg_pool2d chain of
``AlternateCorrBlock.__init__`` (reference core/corr.py:64-72), the NHWC copies of corr.py:82-83 and the bfloat16
cast — against the same chain made of ATen ops on the same GPU.  Same fp32 arithmetic, one rounding: bit-exact."""
import ctypes
import os
import sys

import pytest
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200")]
from raftcorr_b200 import _capi, ops  # noqa: E402
import corr  # noqa: E402

pytestmark = pytest.mark.gpu


def aten_chain(fmap, levels):
    outs = []
    for l in range(levels):
        outs.append(fmap.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16))
        if l + 1 < levels:
            fmap = F.avg_pool2d(fmap, 2, stride=2)
    return outs


@pytest.mark.parametrize("shape", [(1, 256, 46, 62), (2, 128, 17, 33), (1, 24, 9, 70), (3, 64, 8, 8), (1, 200, 55, 128),
                                   (2, 256, 16, 16), (1, 8, 136, 240)])
def test_alt_pack_equals_the_aten_chain(shape):
    B, D, H, W = shape
    gen = torch.Generator().manual_seed(D + H * W)
    fmap = torch.randn(B, D, H, W, generator=gen).cuda()
    got = ops.alt_pack(fmap, 4)
    want = aten_chain(fmap, 4)
    for l in range(4):
        assert got[l].shape == want[l].shape == (B, H >> l, W >> l, D), (shape, l)
        assert torch.equal(got[l].view(torch.int16), want[l].view(torch.int16)), (shape, l)
    one = ops.alt_pack(fmap, 1)
    assert len(one) == 1 and torch.equal(one[0].view(torch.int16), want[0].view(torch.int16))


def test_alt_pack_rejections():
    lib = _capi.lib()
    fmap = torch.randn(1, 7, 16, 16, device="cuda")
    out = torch.empty(1, 16, 16, 7, dtype=torch.bfloat16, device="cuda")
    ptrs = (ctypes.c_void_p * 1)(out.data_ptr())
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert lib.raftcorr_alt_pack(fmap.data_ptr(), 1, 7, 16, 16, 1, ptrs, stream) == -2     # RAFTCORR_E_UNSUPPORTED: odd channel count
    assert lib.raftcorr_alt_pack(fmap.data_ptr(), 1, 8, 4, 4, 4, ptrs, stream) == -1           # RAFTCORR_E_INVALID: level 3 of 4x4 is empty
    with pytest.raises(RuntimeError):
        ops.alt_pack(torch.randn(1, 8, 16, 16), 1)                                                          # host tensor: no CPU path


def test_alternate_block_keeps_the_reference_attributes():
    """``pyramid`` stays core/corr.py:68-72's list of num_levels+1 pooled NCHW pairs (made on first access)."""
    f1 = torch.randn(1, 64, 24, 40, device="cuda")
    f2 = torch.randn(1, 64, 24, 40, device="cuda")
    blk = corr.AlternateCorrBlock(f1, f2)
    assert blk._pyramid is None                       # inference never needs the fp32 copies
    pyr = blk.pyramid
    assert len(pyr) == 5 and pyr[0][0] is f1 and pyr[0][1] is f2
    assert torch.equal(pyr[2][1], F.avg_pool2d(F.avg_pool2d(f2, 2, stride=2), 2, stride=2))
    assert [tuple(t.shape) for t in blk._f2_bf16] == [(1, 24 >> l, 40 >> l, 64) for l in range(4)]
