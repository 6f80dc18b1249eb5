"""End-to-end bar of north_star: after 12 refinement iterations with random-init weights the flow
computed with this repo's correlation path must stay within 0.01 px mean end-point error of the
flow computed with the reference's correlation path (here: its torch restatement, i.e. the same
ATen ops as core/corr.py, on the same device and with the same stock-PyTorch model)."""
import warnings

import pytest
import torch

from harness.raft_stock import StockRAFT
from oracle import torch_restatement as tr

warnings.filterwarnings("ignore")
pytestmark = pytest.mark.gpu


def _flows(model, im1, im2, factory, iters=12):
    with torch.no_grad():
        return model(im1, im2, factory, iters=iters)


@pytest.mark.parametrize("shape", [(1, 368, 496), (2, 440, 1024)])
def test_twelve_iteration_flow_epe(shape):
    import corr
    B, Hi, Wi = shape
    torch.manual_seed(1234)
    model = StockRAFT().eval().cuda()
    gen = torch.Generator().manual_seed(17)
    im1 = (torch.rand(B, 3, Hi, Wi, generator=gen) * 255).cuda()
    im2 = (torch.rand(B, 3, Hi, Wi, generator=gen) * 255).cuda()
    lo_ref, up_ref = _flows(model, im1, im2, lambda a, b, r: tr.TorchCorrBlock(a, b, 4, r))
    mag = up_ref.pow(2).sum(1).sqrt().mean().item()
    assert mag > 0.1   # random-init RAFT does move pixels: the comparison is not vacuous
    for precision, bound in (("fp32", 0.01), ("tf32", 0.01), ("bf16", 0.01)):
        lo, up = _flows(model, im1, im2, lambda a, b, r: corr.CorrBlock(a, b, radius=r, precision=precision))
        assert up.shape == up_ref.shape == (B, 2, Hi, Wi)
        epe = (up - up_ref).pow(2).sum(1).sqrt().mean().item()
        assert epe <= bound, (precision, epe, mag)
    # ... and with the library's convex-upsampling kernel in place of raft.py:72-83 as well
    import upsample
    with torch.no_grad():
        lo, up = model(im1, im2, lambda a, b, r: corr.CorrBlock(a, b, radius=r), iters=12,
                       upsampler=upsample.upsample_flow)
    epe = (up - up_ref).pow(2).sum(1).sqrt().mean().item()
    assert epe <= 0.01, ("upsample", epe)
    # ... and with convc1 (+ bias + ReLU) as the library's tensor-core GEMM over a channels-last lookup
    with torch.no_grad():
        lo, up = model(im1, im2, lambda a, b, r: corr.CorrBlock(a, b, radius=r, channels_last=True), iters=12,
                       upsampler=upsample.upsample_flow, convc1_fn=corr.conv1x1_relu)
    epe = (up - up_ref).pow(2).sum(1).sqrt().mean().item()
    assert epe <= 0.01, ("conv1x1", epe)
    # the memory-lean block computes the same flow
    lo, up = _flows(model, im1[:1], im2[:1], lambda a, b, r: corr.AlternateCorrBlock(a, b, radius=r))
    epe = (up - up_ref[:1]).pow(2).sum(1).sqrt().mean().item()
    assert epe <= 0.01, ("alternate", epe)


def test_small_model_flow_epe():
    """The --small configuration (raft.py:29-33,49-51: radius 3, D=128, 196 lookup channels, sqrt(128) inexact,
    bilinear upflow8): same 0.01 px bar, with the library's upflow8 kernel in place of utils.py:80-82."""
    import corr
    import upsample
    torch.manual_seed(4321)
    model = StockRAFT(small=True).eval().cuda()
    gen = torch.Generator().manual_seed(18)
    im1 = (torch.rand(2, 3, 368, 496, generator=gen) * 255).cuda()
    im2 = (torch.rand(2, 3, 368, 496, generator=gen) * 255).cuda()
    lo_ref, up_ref = _flows(model, im1, im2, lambda a, b, r: tr.TorchCorrBlock(a, b, 4, r))
    assert up_ref.pow(2).sum(1).sqrt().mean().item() > 0.05
    for precision in ("fp32", "bf16"):
        with torch.no_grad():
            lo, up = model(im1, im2, lambda a, b, r: corr.CorrBlock(a, b, radius=r, precision=precision), iters=12,
                           upsampler8=upsample.upflow8)
        assert lo.shape == (2, 2, 46, 62) and up.shape == (2, 2, 368, 496)
        epe = (up - up_ref).pow(2).sum(1).sqrt().mean().item()
        assert epe <= 0.01, (precision, epe)
    lo, up = _flows(model, im1, im2, lambda a, b, r: corr.AlternateCorrBlock(a, b, radius=r))
    assert (up - up_ref).pow(2).sum(1).sqrt().mean().item() <= 0.01
