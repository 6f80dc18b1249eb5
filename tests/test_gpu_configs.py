"""BASELINE.json's remaining configurations as parity cases at (or near) their full sizes.

configs[1] and configs[3] are covered in test_gpu_build / test_gpu_lookup / test_gpu_corrblock;
here: configs[2] (KITTI shape, batch 16, many iterations) and configs[4] (training step:
CorrBlock forward + backward of 12 lookups at batch 12), each against the torch restatement of the
reference running ATen's native CUDA kernels on the same device."""
import warnings

import pytest
import torch

from oracle import torch_restatement as tr

warnings.filterwarnings("ignore")
pytestmark = pytest.mark.gpu


def _grid(B, H, W):
    ys, xs = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
    return torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1)


def test_cfg3_kitti_shape_batch16():
    """376x1248 -> 47x156 fmaps (odd sizes at every level), batch 16 per GPU, 24 lookups."""
    import corr
    B, D, H, W = 16, 256, 47, 156
    gen = torch.Generator(device="cuda").manual_seed(41)
    f1 = torch.randn(B, D, H, W, device="cuda", generator=gen)
    f2 = torch.randn(B, D, H, W, device="cuda", generator=gen)
    with torch.no_grad():
        blk = corr.CorrBlock(f1, f2, radius=4, coord_mode="aten_cuda")
        assert [tuple(p.shape[-2:]) for p in blk.corr_pyramid] == [(47, 156), (23, 78), (11, 39), (5, 19)]
        # reference pyramid for 2 of the 16 pairs (the fp32 volume of all 16 is 4.5 GB; two are enough
        # to pin batch indexing at both ends)
        for b in (0, B - 1):
            ref = tr.TorchCorrBlock(f1[b:b + 1], f2[b:b + 1], 4, 4)
            N = H * W
            for l in range(4):
                got = blk.corr_pyramid[l][b * N:(b + 1) * N]
                err = (got - ref.corr_pyramid[l]).abs().max().item() / ref.corr_pyramid[l].abs().max().item()
                assert err <= 1e-2, (b, l, err)
            coords = _grid(1, H, W)
            for it in range(24):     # 24 refinement iterations, coordinates drifting like a flow would
                coords = coords + 0.7 * torch.randn(1, 2, H, W, device="cuda", generator=gen)
                full = _grid(B, H, W)
                full[b] = coords[0]
                out = blk(full)[b:b + 1]
                with torch.backends.cudnn.flags(enabled=False):
                    want = tr.lookup([blk.corr_pyramid[l][b * N:(b + 1) * N].contiguous() for l in range(4)], coords, 4)
                assert torch.equal(out == 0, want == 0), (b, it)
                assert (out - want).abs().max().item() <= 1e-5 * want.abs().max().item(), (b, it)


def test_cfg5_training_step_gradients():
    """368x496 -> 46x62 fmaps, batch 12, 12 lookups with detached coordinates (raft.py:122-124):
    d fmap1 / d fmap2 against autograd through the reference's op sequence."""
    import corr
    B, D, H, W, iters = 12, 256, 46, 62, 12
    gen = torch.Generator(device="cuda").manual_seed(43)
    f1 = torch.randn(B, D, H, W, device="cuda", generator=gen)
    f2 = torch.randn(B, D, H, W, device="cuda", generator=gen)
    coords = [_grid(B, H, W) + (0.5 + 0.4 * k) * torch.randn(B, 2, H, W, device="cuda", generator=gen) for k in range(iters)]
    ups = [torch.randn(B, 324, H, W, device="cuda", generator=gen) for _ in range(iters)]

    a, b_ = f1.clone().requires_grad_(True), f2.clone().requires_grad_(True)
    blk = corr.CorrBlock(a, b_, precision="fp32", coord_mode="aten_cuda")
    loss = sum((blk(c) * u).sum() for c, u in zip(coords, ups))
    loss.backward()

    ra, rb = f1.clone().requires_grad_(True), f2.clone().requires_grad_(True)
    with torch.backends.cudnn.flags(enabled=False):
        ref = tr.TorchCorrBlock(ra, rb, 4, 4)
        rloss = sum((ref(c) * u).sum() for c, u in zip(coords, ups))
        rloss.backward()
    assert abs(loss.item() - rloss.item()) <= 1e-4 * abs(rloss.item()) + 1e-2
    for got, want in ((a.grad, ra.grad), (b_.grad, rb.grad)):
        assert torch.allclose(got, want, rtol=1e-3, atol=1e-4 * want.abs().max().item())


def test_cfg4_volume_path_equals_on_the_fly_path_at_1088x1920():
    """BASELINE.json configs[3] shape, one pair: the all-pairs volume at 136x240 is 5.7 GB per pair
    (the reason the reference offers AlternateCorrBlock); on a 180 GB part both blocks fit, so the
    largest configuration doubles as a cross-check of the two independent code paths: tensor-core
    volume + TMA lookup versus fp32 on-the-fly dot products.  Same zero masks, values within the
    bf16 bar."""
    import corr
    B, D, H, W = 1, 256, 136, 240
    gen = torch.Generator(device="cuda").manual_seed(47)
    f1 = torch.randn(B, D, H, W, device="cuda", generator=gen)
    f2 = torch.randn(B, D, H, W, device="cuda", generator=gen)
    low = 8 * torch.randn(B, 2, H // 8, W // 8, device="cuda", generator=gen)
    c = _grid(B, H, W) + torch.nn.functional.interpolate(low, size=(H, W), mode="bilinear", align_corners=True)
    with torch.no_grad():
        vol = corr.CorrBlock(f1, f2)
        assert [tuple(p.shape) for p in vol.corr_pyramid] == [(32640, 1, 136, 240), (32640, 1, 68, 120),
                                                              (32640, 1, 34, 60), (32640, 1, 17, 30)]
        a = vol(c)
        b = corr.AlternateCorrBlock(f1, f2, precision="fp32")(c)
        del vol
    assert a.shape == b.shape == (1, 324, 136, 240)
    assert torch.equal(a == 0, b == 0)
    assert (a - b).abs().max().item() <= 1e-2 * b.abs().max().item()
