"""Volume + pyramid build (family 1) through the C ABI against the oracle / golden vectors.
Tolerances are north_star's: fp32 mode 1e-5, bf16/TF32 modes 1e-2, relative to max|ref|."""
import warnings

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import oracle
from raftcorr_b200 import ops
from helpers import pyramid_to_numpy, rel_max

warnings.filterwarnings("ignore")
pytestmark = pytest.mark.gpu

TOL = {"fp32": 1e-5, "tf32": 1e-2, "bf16": 1e-2}


@pytest.mark.parametrize("precision", ["fp32", "bf16", "tf32"])
def test_tiny_golden(golden, precision):
    g = golden("corr_tiny")
    pyr = ops.build_pyramid(torch.from_numpy(g["fmap1"]).cuda(), torch.from_numpy(g["fmap2"]).cuda(), 4, precision)
    got = pyramid_to_numpy(pyr)
    for l in range(4):
        assert got[l].shape == g[f"pyr{l}"].shape
        assert rel_max(got[l], g[f"pyr{l}"]) <= TOL[precision], (precision, l)


@pytest.mark.parametrize("precision", ["fp32", "bf16", "tf32"])
def test_odd_sizes_golden(golden, precision):
    """17x23: every level has odd sizes (floor pooling), D=8 (sqrt(8) inexact, K padded by TMA)."""
    g = golden("corr_odd_r3")
    pyr = ops.build_pyramid(torch.from_numpy(g["fmap1"]).cuda(), torch.from_numpy(g["fmap2"]).cuda(), 4, precision)
    got = pyramid_to_numpy(pyr)
    assert [p.shape[-2:] for p in got] == [(17, 23), (8, 11), (4, 5), (2, 2)]
    assert rel_max(got[0][::7], g["pyr0_sample"]) <= TOL[precision]
    for l in range(1, 4):
        assert rel_max(got[l], g[f"pyr{l}"]) <= TOL[precision], (precision, l)


@pytest.mark.parametrize("precision", ["fp32", "bf16", "tf32"])
def test_cfg1_baseline_config(golden, precision):
    g = golden("corr_cfg1")
    B, D, H, W, r, L = (int(v) for v in g["shape"])
    gen = torch.Generator().manual_seed(int(g["seed"]))
    f1 = torch.randn(B, D, H, W, generator=gen)
    f2 = torch.randn(B, D, H, W, generator=gen)
    if not np.array_equal(f1.numpy()[0, :4, :2, :3], g["fmap1_probe"]):
        pytest.skip("torch CPU generator differs from the one that made the fixture")
    got = pyramid_to_numpy(ops.build_pyramid(f1.cuda(), f2.cuda(), L, precision))
    for l in range(L):
        assert rel_max(got[l][::97, 0, ::3, ::5], g[f"pyr{l}_sample"]) <= TOL[precision], (precision, l)
        assert abs(got[l].astype(np.float64).sum() - g[f"pyr{l}_sum"][0]) <= TOL[precision] * g[f"pyr{l}_sum"][1]


@pytest.mark.parametrize("shape", [(8, 256, 55, 128), (2, 256, 47, 156), (2, 128, 46, 62)])
@pytest.mark.parametrize("precision", ["bf16", "tf32"])
def test_full_size_tensor_core_vs_fp32_path(shape, precision):
    """BASELINE.json configs[1]/[2] shapes (+ the --small model's D=128).  The tensor-core build
    is compared with the fp32 CUDA-core build on the device (both from this library; the fp32
    path is pinned to the golden vectors above), and each pooled level must equal the 2x2 mean
    of the level below it (size-independent property)."""
    B, D, H, W = shape
    gen = torch.Generator(device="cuda").manual_seed(21)
    f1 = torch.randn(B, D, H, W, device="cuda", generator=gen)
    f2 = torch.randn(B, D, H, W, device="cuda", generator=gen)
    ref = ops.build_pyramid(f1, f2, 4, "fp32").views()
    got = ops.build_pyramid(f1, f2, 4, precision).views()
    for l in range(4):
        assert got[l].shape == ref[l].shape
        err = (got[l] - ref[l]).abs().max().item() / ref[l].abs().max().item()
        assert err <= 1e-2, (precision, l, err)
    for l in range(1, 4):
        pooled = F.avg_pool2d(got[l - 1].contiguous(), 2, stride=2)
        assert (pooled - got[l]).abs().max().item() <= 2e-6 * got[l].abs().max().item()


def test_fp32_mode_full_size_is_fp32_accurate():
    """'fp32' precision = three tf32 tensor-core passes over hi/lo splits (csrc/build_tc.cu).  At the
    BASELINE.json configs[1] shape it must stay within 1e-5 of max|ref| of an fp64 contraction, and the
    CUDA-core fallback (D not a multiple of 8) must meet the same bar."""
    gen = torch.Generator(device="cuda").manual_seed(5)
    for (B, D, H, W) in [(1, 256, 55, 128), (1, 100, 24, 40)]:
        f1 = torch.randn(B, D, H, W, device="cuda", generator=gen)
        f2 = torch.randn(B, D, H, W, device="cuda", generator=gen)
        lv0 = ops.build_pyramid(f1, f2, 4, "fp32").views()[0].reshape(B, H * W, H * W)
        rows = torch.arange(0, H * W, 37, device="cuda")
        ref = torch.einsum("bdi,bdj->bij", f1.double().flatten(2)[:, :, rows], f2.double().flatten(2)) / (D ** 0.5)
        err = (lv0[:, rows].double() - ref).abs().max().item() / ref.abs().max().item()
        assert err <= 1e-5, (D, err)


def test_odd_channel_count_is_served_by_the_fp32_kernel():
    """D % 8 != 0 cannot use the tensor-core operand layout; every precision request then runs the
    fp32 CUDA-core build (the reference accepts any D)."""
    gen = torch.Generator(device="cuda").manual_seed(6)
    f1 = torch.randn(1, 12, 16, 20, device="cuda", generator=gen)
    f2 = torch.randn(1, 12, 16, 20, device="cuda", generator=gen)
    ref = torch.einsum("bdi,bdj->bij", f1.double().flatten(2), f2.double().flatten(2)) / (12 ** 0.5)
    for precision in ("bf16", "tf32", "fp32"):
        lv0 = ops.build_pyramid(f1, f2, 4, precision).views()[0].reshape(1, 320, 320)
        assert (lv0.double() - ref).abs().max().item() <= 1e-5 * ref.abs().max().item(), precision


def test_static_corr_and_idempotence():
    """CorrBlock.corr (corr.py:52-60) shape, and two builds of the same input agree bit for bit."""
    import corr
    gen = torch.Generator(device="cuda").manual_seed(4)
    f1 = torch.randn(2, 64, 24, 40, device="cuda", generator=gen)
    f2 = torch.randn(2, 64, 24, 40, device="cuda", generator=gen)
    v1 = corr.CorrBlock.corr(f1, f2)
    v2 = corr.CorrBlock.corr(f1, f2)
    assert v1.shape == (2, 24, 40, 1, 24, 40)
    assert torch.equal(v1, v2)
    ref = torch.einsum("bdi,bdj->bij", f1.double().flatten(2), f2.double().flatten(2)) / 8.0
    assert (v1.reshape(2, 960, 960).double() - ref).abs().max().item() <= 1e-2 * ref.abs().max().item()
    v3 = corr.CorrBlock.corr(f1, f2, precision="fp32")
    assert (v3.reshape(2, 960, 960).double() - ref).abs().max().item() <= 1e-5 * ref.abs().max().item()
