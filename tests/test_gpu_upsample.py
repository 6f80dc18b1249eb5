"""Convex 8x upsampling and upflow8 on the GPU against the reference's golden outputs
(RAFT.upsample_flow, raft.py:72-83) and the oracle.  Floating point: softmax uses expf, so the bar is
1e-5 of max|ref| forward, allclose(rtol 1e-3, atol 1e-4*max) for the atomically accumulated grads."""
import numpy as np
import pytest
import torch

import oracle
from oracle import torch_restatement as tr
from helpers import rel_max

pytestmark = pytest.mark.gpu


def test_forward_matches_reference_golden(golden):
    import upsample
    g = golden("upsample_tiny")
    up = upsample.upsample_flow(torch.from_numpy(g["flow"]).cuda(), torch.from_numpy(g["mask"]).cuda())
    assert up.shape == g["up"].shape and up.dtype == torch.float32 and up.is_contiguous()
    assert rel_max(up.cpu().numpy(), g["up"]) <= 1e-5
    up8 = upsample.upflow8(torch.from_numpy(g["flow"]).cuda())
    assert rel_max(up8.cpu().numpy(), g["up8"]) <= 1e-5


def test_backward_matches_reference_golden(golden):
    import upsample
    g = golden("upsample_tiny")
    flow = torch.from_numpy(g["flow"]).cuda().requires_grad_(True)
    mask = torch.from_numpy(g["mask"]).cuda().requires_grad_(True)
    upsample.upsample_flow(flow, mask).backward(torch.from_numpy(g["grad_up"]).cuda())
    for got, ref in ((flow.grad, g["grad_flow"]), (mask.grad, g["grad_mask"])):
        assert np.allclose(got.cpu().numpy(), ref, rtol=1e-3, atol=1e-4 * np.abs(ref).max())


@pytest.mark.parametrize("B,H,W", [(1, 1, 1), (1, 3, 33), (2, 5, 64), (1, 46, 62), (2, 55, 128), (1, 6, 34), (1, 47, 156)])
def test_forward_matches_oracle(B, H, W):
    import upsample
    gen = torch.Generator().manual_seed(H * 1000 + W)
    flow = 6 * torch.randn(B, 2, H, W, generator=gen)
    mask = 4 * torch.randn(B, 576, H, W, generator=gen)
    ref = oracle.upsample_flow(flow, mask)
    got = upsample.upsample_flow(flow.cuda(), mask.cuda()).cpu().numpy()
    assert rel_max(got, ref) <= 1e-5
    assert rel_max(upsample.upflow8(flow.cuda()).cpu().numpy(), oracle.upflow8(flow)) <= 1e-5


@pytest.mark.parametrize("B,H,W", [(3, 46, 62), (2, 17, 128), (1, 5, 7), (2, 6, 34)])
def test_backward_matches_autograd_of_restatement(B, H, W):
    """cfg5 shape (46x62, the 36-wide boxes), a 32-wide-box shape, and two that take the plain-load kernels:
    grads against torch autograd through the reference's op sequence on the GPU."""
    import upsample
    gen = torch.Generator().manual_seed(77)
    flow = (5 * torch.randn(B, 2, H, W, generator=gen)).cuda()
    mask = (3 * torch.randn(B, 576, H, W, generator=gen)).cuda()
    gup = torch.randn(B, 2, 8 * H, 8 * W, generator=gen).cuda()
    a, b = flow.clone().requires_grad_(True), mask.clone().requires_grad_(True)
    tr.upsample_flow(a, b).backward(gup)
    c, d = flow.clone().requires_grad_(True), mask.clone().requires_grad_(True)
    upsample.upsample_flow(c, d).backward(gup)
    for got, ref in ((c.grad, a.grad), (d.grad, b.grad)):
        assert torch.allclose(got, ref, rtol=1e-3, atol=1e-4 * ref.abs().max().item())
    # only one of the two gradients wanted
    e = mask.clone().requires_grad_(True)
    upsample.upsample_flow(flow, e).backward(gup)
    assert torch.allclose(e.grad, b.grad, rtol=1e-3, atol=1e-4 * b.grad.abs().max().item())


def test_properties_at_full_size():
    """cfg2 size: a constant flow stays constant away from the border (weights sum to 1), and a huge
    logit selects exactly one neighbour."""
    import upsample
    B, H, W = 8, 55, 128
    gen = torch.Generator().manual_seed(5)
    mask = torch.randn(B, 576, H, W, generator=gen).cuda()
    flow = torch.full((B, 2, H, W), 1.25).cuda()
    up = upsample.upsample_flow(flow, mask)
    assert torch.allclose(up[:, :, 8:-8, 8:-8], torch.full_like(up[:, :, 8:-8, 8:-8], 10.0), rtol=0, atol=1e-5)
    flow = torch.randn(B, 2, H, W, generator=gen).cuda()
    mask = torch.zeros(B, 576, H, W).cuda()
    mask[:, 5 * 64:6 * 64] = 200.0   # k=5: neighbour (y, x+1)
    up = upsample.upsample_flow(flow, mask)
    want = torch.zeros_like(flow)
    want[..., :-1] = 8 * flow[..., 1:]
    want = want.repeat_interleave(8, dim=2).repeat_interleave(8, dim=3)
    assert torch.equal(up, want)


def test_host_tensor_is_rejected():
    import upsample
    with pytest.raises(RuntimeError):
        upsample.upsample_flow(torch.zeros(1, 2, 4, 4), torch.zeros(1, 576, 4, 4))
    with pytest.raises(RuntimeError):
        upsample.upsample_flow(torch.zeros(1, 2, 4, 4).cuda(), torch.zeros(1, 575, 4, 4).cuda())
