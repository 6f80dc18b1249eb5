"""The oracle (oracle/) against the vectors the unmodified reference produced
(oracle/make_golden.py -> tests/golden/).  CPU only."""
import warnings

import numpy as np
import pytest
import torch

import oracle
from oracle import torch_restatement as tr

warnings.filterwarnings("ignore")

KINDS = ("grid", "half", "sigma4", "sigma30")


def _pyr(g, n=4):
    return [g[f"pyr{l}"] for l in range(n)]


def test_volume_and_pyramid_tiny(golden):
    g = golden("corr_tiny")
    pyr = oracle.build_pyramid(g["fmap1"], g["fmap2"], 4)
    for l in range(4):
        ref = g[f"pyr{l}"]
        assert pyr[l].shape == ref.shape
        # fp32 mode bar of north_star: 1e-5 of max|ref|
        assert np.abs(pyr[l] - ref).max() <= 1e-5 * np.abs(ref).max()


def test_pool_is_bit_exact(golden):
    g = golden("corr_tiny")
    p = g["pyr0"]
    for l in range(1, 4):
        p = oracle.avg_pool2(p)
        assert np.array_equal(p, g[f"pyr{l}"])


@pytest.mark.parametrize("kind", KINDS)
def test_lookup_bit_exact_on_reference_pyramid(golden, kind):
    g = golden("corr_tiny")
    out = oracle.lookup(_pyr(g), g[f"coords_{kind}"], int(g["radius"]))
    ref = g[f"out_{kind}"]
    assert out.shape == ref.shape == (1, 324, 16, 18)
    assert np.array_equal(out == 0, ref == 0)      # zero-padding masks
    assert np.array_equal(out, ref)                # and every value, bit for bit


@pytest.mark.parametrize("kind", ("grid", "half"))
def test_direct_floor_has_different_masks(golden, kind):
    """The normalise round trip moves floor() on integer grids: a kernel that
    samples at floor(x) directly (alt_cuda_corr) cannot reproduce CorrBlock's zero masks."""
    g = golden("corr_tiny")
    out = oracle.lookup(_pyr(g), g[f"coords_{kind}"], 4, mode=oracle.COORD_DIRECT)
    ref = g[f"out_{kind}"]
    assert not np.array_equal(out == 0, ref == 0)
    assert np.abs(out - ref).max() <= 1e-5 * np.abs(ref).max()


def test_odd_sizes_radius3(golden):
    g = golden("corr_odd_r3")
    r = int(g["radius"])
    pyr = oracle.build_pyramid(g["fmap1"], g["fmap2"], 4)
    assert [p.shape[-2:] for p in pyr] == [(17, 23), (8, 11), (4, 5), (2, 2)]
    for l in range(1, 4):
        assert np.abs(pyr[l] - g[f"pyr{l}"]).max() <= 1e-5 * np.abs(g[f"pyr{l}"]).max()
    assert np.abs(pyr[0][::7] - g["pyr0_sample"]).max() <= 1e-5 * np.abs(g["pyr0_sample"]).max()
    out = oracle.lookup(pyr, g["coords_sigma4"], r)
    ref = g["out_sigma4"]
    assert out.shape == ref.shape == (2, 4 * 49, 17, 23)
    assert np.array_equal(out == 0, ref == 0)
    assert np.abs(out - ref).max() <= 1e-5 * np.abs(ref).max()


def _cfg1_inputs(g):
    B, D, H, W, r, L = (int(v) for v in g["shape"])
    gen = torch.Generator().manual_seed(int(g["seed"]))
    f1 = torch.randn(B, D, H, W, generator=gen)
    f2 = torch.randn(B, D, H, W, generator=gen)
    if not (np.array_equal(f1.numpy()[0, :4, :2, :3], g["fmap1_probe"])
            and np.array_equal(f2.numpy()[0, :4, :2, :3], g["fmap2_probe"])):
        pytest.skip("torch CPU generator differs from the one that made the fixture")
    return f1, f2, r, L


def test_cfg1_baseline_config(golden):
    """BASELINE.json configs[0]: B=1, D=256, 46x62, 4 levels, r=4."""
    g = golden("corr_cfg1")
    f1, f2, r, L = _cfg1_inputs(g)
    pyr = oracle.build_pyramid(f1, f2, L)
    for l in range(L):
        s = pyr[l][::97, 0, ::3, ::5]
        ref = g[f"pyr{l}_sample"]
        assert np.abs(s - ref).max() <= 1e-5 * np.abs(ref).max()
    tpyr = tr.build_pyramid(f1, f2, L)
    for kind in ("grid", "sigma4", "sigma30"):
        c = g[f"coords_{kind}"]
        out = oracle.lookup(pyr, c, r)
        ref_s = g[f"out_{kind}_sample"]
        assert np.abs(out[:, :, ::5, ::7] - ref_s).max() <= 1e-5 * np.abs(ref_s).max()
        # zero counts per channel pin the masks on the full tensor
        assert np.array_equal((out == 0).sum(axis=(0, 2, 3)), g[f"out_{kind}_zero_rows"])
        tout = tr.lookup(tpyr, torch.from_numpy(c), r).numpy()
        assert np.array_equal(tout[:, :, ::5, ::7], ref_s)
        assert np.isclose(tout.astype(np.float64).sum(), g[f"out_{kind}_sum"][0])


@pytest.mark.parametrize("kind", KINDS)
def test_torch_restatement_bit_exact(golden, kind):
    g = golden("corr_tiny")
    f1, f2 = torch.from_numpy(g["fmap1"]), torch.from_numpy(g["fmap2"])
    blk = tr.TorchCorrBlock(f1, f2, 4, 4)
    for l in range(4):
        assert np.array_equal(blk.corr_pyramid[l].numpy(), g[f"pyr{l}"])
    out = blk(torch.from_numpy(g[f"coords_{kind}"])).numpy()
    assert np.array_equal(out, g[f"out_{kind}"])


def test_torch_restatement_grads(golden):
    g = golden("corr_tiny")
    f1 = torch.from_numpy(g["fmap1"]).requires_grad_(True)
    f2 = torch.from_numpy(g["fmap2"]).requires_grad_(True)
    c = torch.from_numpy(g["coords_sigma4"]).requires_grad_(True)
    out = tr.TorchCorrBlock(f1, f2, 4, 4)(c)
    out.backward(torch.from_numpy(g["grad_out"]))
    for got, key in ((f1.grad, "grad_fmap1"), (f2.grad, "grad_fmap2"), (c.grad, "grad_coords")):
        ref = g[key]
        assert np.abs(got.numpy() - ref).max() <= 1e-5 * np.abs(ref).max()


def test_alt_path_matches_corrblock_math(golden):
    """AlternateCorrBlock == CorrBlock up to the coordinate round trip:
    pooling commutes with the dot product.  The reference's CUDA extension cannot run without
    a GPU, so its restatement (oracle.alt_forward, 'parity unpinned') is cross-checked against
    the reference CorrBlock's golden output at non-integer coordinates."""
    g = golden("corr_tiny")
    ref = g["out_sigma4"]
    out = oracle.alt_corr_block(g["fmap1"], g["fmap2"], g["coords_sigma4"], 4, 4)
    assert out.shape == ref.shape
    assert np.abs(out - ref).max() <= 2e-5 * np.abs(ref).max()
    # with the ATen round trip replayed per tap, masks match CorrBlock exactly too
    out0 = oracle.alt_corr_block(g["fmap1"], g["fmap2"], g["coords_grid"], 4, 4,
                                 mode=oracle.COORD_ATEN_CPU)
    assert np.array_equal(out0 == 0, g["out_grid"] == 0)
    t = tr.alt_corr_block_math(torch.from_numpy(g["fmap1"]), torch.from_numpy(g["fmap2"]),
                               torch.from_numpy(g["coords_sigma4"]), 4, 4).numpy()
    assert np.abs(t - out).max() <= 2e-5 * np.abs(ref).max()


def test_upsample_flow_restatements_match_reference(golden):
    """RAFT.upsample_flow / upflow8 (raft.py:72-83, utils.py:80-82): C loops and torch restatement vs the
    reference's own outputs and autograd grads."""
    from helpers import rel_max
    g = golden("upsample_tiny")
    up = oracle.upsample_flow(g["flow"], g["mask"])
    assert rel_max(up, g["up"]) <= 1e-6
    assert rel_max(oracle.upflow8(g["flow"]), g["up8"]) <= 1e-6
    flow = torch.from_numpy(g["flow"]).requires_grad_(True)
    mask = torch.from_numpy(g["mask"]).requires_grad_(True)
    out = tr.upsample_flow(flow, mask)
    assert np.array_equal(out.detach().numpy(), g["up"])
    out.backward(torch.from_numpy(g["grad_up"]))
    assert np.array_equal(flow.grad.numpy(), g["grad_flow"])
    assert np.array_equal(mask.grad.numpy(), g["grad_mask"])
    assert np.array_equal(tr.upflow8(flow.detach()).numpy(), g["up8"])
