"""Values of the ``alt_cuda_corr`` drop-in's backward (reference alt_cuda_corr/correlation.cpp:36-48,
correlation_kernel.cu:122-256: dead code there — never called from Python, coords_grad left at zero) in the
reference's own layouts, against autograd through a float64 restatement of what ``alt_cuda_corr.forward`` computes:

    corr[b, 0, ix*rd + iy, h, w] = < fmap1[b, h, w, :],  bilinear(fmap2[b], x + ix - r, y + iy - r) >

with (x, y) = coords[b, 0, h, w, :], floor() taken directly (correlation_kernel.cu:66-76) and zero padding.
fmap1_grad and fmap2_grad are what the reference kernel computes; coords_grad is this library's documented superset
(the reference returns zeros), checked against the same autograd.  Bar: allclose(rtol 1e-3, atol 1e-4 * max)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def alt_forward_f64(f1, f2, coords, r):
    """f1 (B,H1,W1,C), f2 (B,H2,W2,C), coords (B,1,H1,W1,2) -> (B,1,rd*rd,H1,W1); differentiable."""
    B, H1, W1, C = f1.shape
    H2, W2 = f2.shape[1:3]
    rd = 2 * r + 1
    x, y = coords[:, 0, :, :, 0], coords[:, 0, :, :, 1]
    bidx = torch.arange(B, device=f1.device).view(B, 1, 1).expand(B, H1, W1)

    def pixel(yy, xx):
        ok = (yy >= 0) & (yy < H2) & (xx >= 0) & (xx < W2)
        v = f2[bidx, yy.clamp(0, H2 - 1), xx.clamp(0, W2 - 1)]            # (B,H1,W1,C)
        return v * ok.unsqueeze(-1)

    outs = []
    for ix in range(rd):
        for iy in range(rd):
            px, py = x + (ix - r), y + (iy - r)
            x0, y0 = torch.floor(px), torch.floor(py)
            wx, wy = (px - x0).unsqueeze(-1), (py - y0).unsqueeze(-1)
            x0, y0 = x0.long(), y0.long()
            s = (pixel(y0, x0) * (1 - wx) * (1 - wy) + pixel(y0, x0 + 1) * wx * (1 - wy) +
                 pixel(y0 + 1, x0) * (1 - wx) * wy + pixel(y0 + 1, x0 + 1) * wx * wy)
            outs.append((s * f1).sum(-1))
    return torch.stack(outs, dim=1).unsqueeze(1)


CASES = [  # B, C, H1, W1, H2, W2, r, sigma
    (2, 32, 12, 20, 12, 20, 4, 3.0),
    (1, 24, 17, 23, 8, 11, 3, 2.0),      # a pooled level (the caller divides coords by 2**l), radius 3, odd sizes
    (2, 16, 9, 33, 9, 33, 4, 12.0),      # windows far across the border
    (1, 256, 46, 62, 46, 62, 4, 4.0),    # BASELINE.json configs[4] shape (368x496), D = 256
    (1, 128, 46, 62, 23, 31, 3, 4.0),    # the small model at level 1
]


@pytest.mark.parametrize("case", CASES)
def test_alt_cuda_corr_backward_values(case):
    import alt_cuda_corr
    B, C, H1, W1, H2, W2, r, sigma = case
    gen = torch.Generator().manual_seed(B * 100 + C + H1)
    f1 = torch.randn(B, H1, W1, C, generator=gen).cuda()
    f2 = torch.randn(B, H2, W2, C, generator=gen).cuda()
    ys, xs = torch.meshgrid(torch.arange(H1), torch.arange(W1), indexing="ij")
    scale = W2 / W1
    base = torch.stack([xs * scale, ys * scale], dim=-1).float()                       # (H1,W1,2), last = (x, y)
    coords = (base[None, None] + sigma * torch.randn(B, 1, H1, W1, 2, generator=gen) + 0.013).contiguous().cuda()
    rd = 2 * r + 1
    up = torch.randn(B, 1, rd * rd, H1, W1, generator=gen).cuda()

    fwd = alt_cuda_corr.forward(f1, f2, coords, r)[0]
    a, b, c = (t.double().requires_grad_(True) for t in (f1, f2, coords))
    ref = alt_forward_f64(a, b, c, r)
    assert fwd.shape == ref.shape
    assert (fwd.double() - ref).abs().max().item() <= 1e-5 * ref.abs().max().item()
    ref.backward(up.double())

    g1, g2, gc = alt_cuda_corr.backward(f1, f2, coords, up, r)
    assert (g1.shape, g2.shape, gc.shape) == (f1.shape, f2.shape, coords.shape)
    for name, got, want in (("fmap1_grad", g1, a.grad), ("fmap2_grad", g2, b.grad), ("coords_grad", gc, c.grad)):
        want = want.float()
        assert torch.allclose(got, want, rtol=1e-3, atol=1e-4 * want.abs().max().item()), \
            (case, name, (got - want).abs().max().item(), want.abs().max().item())
