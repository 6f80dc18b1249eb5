"""Lookup kernel (family 2) through the C ABI against the oracle and the reference's golden
vectors.  Bar: zero-padding masks identical, values bit-exact for an identical fp32 pyramid."""
import warnings

import numpy as np
import pytest
import torch

import oracle
from oracle import torch_restatement as tr
from raftcorr_b200 import ops
from helpers import coords_for, pyramid_from_numpy, rel_max

warnings.filterwarnings("ignore")
pytestmark = pytest.mark.gpu

KINDS = ("grid", "half", "sigma4", "sigma30")


@pytest.mark.parametrize("kind", KINDS)
def test_tiny_golden_bit_exact(golden, kind):
    g = golden("corr_tiny")
    pyr = pyramid_from_numpy([g[f"pyr{l}"] for l in range(4)], 1, 16, 18)
    out = ops.lookup(pyr, torch.from_numpy(g[f"coords_{kind}"]).cuda(), 4, "aten_cpu").cpu().numpy()
    ref = g[f"out_{kind}"]
    assert out.shape == ref.shape
    assert np.array_equal(out == 0, ref == 0)
    assert np.array_equal(out, ref)


def test_odd_sizes_radius3_golden(golden):
    g = golden("corr_odd_r3")
    levels = oracle.build_pyramid(g["fmap1"], g["fmap2"], 4)
    pyr = pyramid_from_numpy(levels, 2, 17, 23)
    out = ops.lookup(pyr, torch.from_numpy(g["coords_sigma4"]).cuda(), 3, "aten_cpu").cpu().numpy()
    assert np.array_equal(out, oracle.lookup(levels, g["coords_sigma4"], 3))
    ref = g["out_sigma4"]
    assert np.array_equal(out == 0, ref == 0)
    assert rel_max(out, ref) <= 1e-5


def test_cfg1_baseline_config(golden):
    """BASELINE.json configs[0] (46x62, D=256): oracle pyramid -> CUDA lookup vs oracle + golden."""
    g = golden("corr_cfg1")
    B, D, H, W, r, L = (int(v) for v in g["shape"])
    gen = torch.Generator().manual_seed(int(g["seed"]))
    f1 = torch.randn(B, D, H, W, generator=gen)
    f2 = torch.randn(B, D, H, W, generator=gen)
    levels = oracle.build_pyramid(f1, f2, L)
    pyr = pyramid_from_numpy(levels, B, H, W)
    same_inputs = np.array_equal(f1.numpy()[0, :4, :2, :3], g["fmap1_probe"])
    for kind in ("grid", "sigma4", "sigma30"):
        c = g[f"coords_{kind}"]
        out = ops.lookup(pyr, torch.from_numpy(c).cuda(), r, "aten_cpu").cpu().numpy()
        assert np.array_equal(out, oracle.lookup(levels, c, r))
        if same_inputs:
            assert rel_max(out[:, :, ::5, ::7], g[f"out_{kind}_sample"]) <= 1e-5
            assert np.array_equal((out == 0).sum(axis=(0, 2, 3)), g[f"out_{kind}_zero_rows"])


@pytest.mark.parametrize("shape", [(8, 55, 128), (2, 47, 156)])
def test_full_size_against_aten_cuda(shape):
    """BASELINE.json configs[1]/[2] shapes.  Second oracle: the torch restatement of the
    reference run on the GPU with ATen's native CUDA kernels; coord_mode 'aten_cuda' replays
    their reciprocal-multiply division, so the zero masks must agree exactly.  (cuDNN is
    switched off for the reference call: with it on, F.grid_sample dispatches to cuDNN's
    spatial-transformer sampler, whose undocumented arithmetic differs in the last ulp at
    integer taps - see DESIGN.md "coordinate semantics".)"""
    B, H, W = shape
    gen = torch.Generator().manual_seed(7)
    pyr = ops.Pyramid(B, H, W, 4, "cuda", zero=True)
    for l in range(4):
        d = pyr.desc[l]
        pyr.storage[l][:, :, : d.W].normal_(generator=None)
    levels = [v.contiguous() for v in pyr.views()]
    for kind in ("grid", "sigma4", "sigma30"):
        c = coords_for(kind, B, H, W, gen).cuda()
        out = ops.lookup(pyr, c, 4, "aten_cuda")
        with torch.backends.cudnn.flags(enabled=False):
            ref = tr.lookup(levels, c, 4)
        assert out.shape == ref.shape == (B, 324, H, W)
        assert out.is_contiguous()
        assert torch.equal(out == 0, ref == 0)
        assert (out - ref).abs().max().item() <= 1e-5 * ref.abs().max().item()


def test_linearity_and_one_hot():
    """Size-independent properties at full size: the lookup is linear in the pyramid, and a
    one-hot level-0 image lights up exactly the taps whose footprint covers that pixel."""
    B, H, W = 1, 55, 128
    gen = torch.Generator().manual_seed(11)
    c = coords_for("sigma4", B, H, W, gen).cuda()
    p1 = ops.Pyramid(B, H, W, 4, "cuda", zero=True)
    p2 = ops.Pyramid(B, H, W, 4, "cuda", zero=True)
    p3 = ops.Pyramid(B, H, W, 4, "cuda", zero=True)
    for l in range(4):
        w = p1.desc[l].W
        p1.storage[l][:, :, :w].normal_()
        p2.storage[l][:, :, :w].normal_()
        p3.storage[l].copy_(2.0 * p1.storage[l] + p2.storage[l])
    o1, o2, o3 = (ops.lookup(p, c, 4) for p in (p1, p2, p3))
    assert (o3 - (2.0 * o1 + o2)).abs().max().item() <= 1e-5 * o3.abs().max().item()

    hot = ops.Pyramid(B, H, W, 4, "cuda", zero=True)
    hot.storage[0][:, 20, 40] = 1.0   # pixel (x=40, y=20) of every query's level-0 image
    out = ops.lookup(hot, c, 4)[0, :81].reshape(9, 9, H, W)   # [i (x tap), j (y tap)]
    cx, cy = c[0, 0], c[0, 1]
    for i in (0, 4, 8):
        for j in (0, 3, 8):
            px, py = cx + (i - 4), cy + (j - 4)
            covers = ((px - 40).abs() < 1 - 1e-4) & ((py - 20).abs() < 1 - 1e-4)
            outside = ((px - 40).abs() > 1 + 1e-4) | ((py - 20).abs() > 1 + 1e-4)
            assert bool((out[i, j][covers] > 0).all())
            assert bool((out[i, j][outside] == 0).all())


def test_wild_coordinates_match_oracle():
    """NaN / inf / 1e9 / -1e9 coordinates: no fault, same result as the oracle (NaN == NaN)."""
    B, H, W = 1, 16, 40
    gen = torch.Generator().manual_seed(3)
    levels = [torch.randn(B * H * W, 1, H >> l, W >> l, generator=gen).numpy() for l in range(4)]
    c = coords_for("sigma4", B, H, W, gen)
    c[0, 0, 0, 0] = float("nan")
    c[0, 1, 0, 1] = float("inf")
    c[0, 0, 0, 2] = -float("inf")
    c[0, 0, 1, 0] = 1e9
    c[0, 1, 1, 1] = -1e9
    c[0, 0, 1, 2] = 3e38
    c[0, 0, 2, 0] = 20000.7
    c[0, 1, 2, 1] = -70000.2
    pyr = pyramid_from_numpy(levels, B, H, W)
    out = ops.lookup(pyr, c.cuda(), 4, "aten_cpu").cpu().numpy()
    ref = oracle.lookup(levels, c.numpy(), 4)
    assert np.array_equal(np.isnan(out), np.isnan(ref))
    assert np.array_equal(np.nan_to_num(out, nan=0.0), np.nan_to_num(ref, nan=0.0))


def test_ragged_and_small_shapes():
    """W not a multiple of the 32-query tile, single-row maps, fewer levels, radius 0..4."""
    gen = torch.Generator().manual_seed(5)
    for (B, H, W, L, r) in [(1, 8, 8, 3, 4), (3, 9, 33, 2, 2), (1, 16, 95, 4, 1), (2, 20, 31, 1, 0), (1, 17, 64, 4, 3)]:
        levels = [torch.randn(B * H * W, 1, H >> l, W >> l, generator=gen).numpy() for l in range(L)]
        c = coords_for("sigma4", B, H, W, gen)
        out = ops.lookup(pyramid_from_numpy(levels, B, H, W), c.cuda(), r, "aten_cpu").cpu().numpy()
        assert np.array_equal(out, oracle.lookup(levels, c.numpy(), r)), (B, H, W, L, r)


def test_unpadded_contiguous_levels_are_accepted():
    """The C ABI takes any pitch >= W (e.g. a plain contiguous reference-style pyramid)."""
    from raftcorr_b200 import _capi
    B, H, W, r = 1, 17, 23, 4
    gen = torch.Generator().manual_seed(9)
    levels = [torch.randn(B * H * W, 1, H >> l, W >> l, generator=gen) for l in range(4)]
    dev = [lv.cuda() for lv in levels]
    desc = _capi.LevelArray()
    for l, t in enumerate(dev):
        desc[l].data, desc[l].H, desc[l].W = t.data_ptr(), t.shape[-2], t.shape[-1]
        desc[l].pitch, desc[l].image_stride = t.shape[-1], t.shape[-1] * t.shape[-2]
    c = coords_for("sigma4", B, H, W, gen)
    cd = c.cuda()
    out = torch.empty(B, 324, H, W, device="cuda")
    rc = _capi.lib().raftcorr_lookup_fwd(desc, 4, cd.data_ptr(), out.data_ptr(), B, H, W, r, 0,
                                         torch.cuda.current_stream().cuda_stream)
    assert rc == 0, _capi.lib().raftcorr_last_error()
    assert np.array_equal(out.cpu().numpy(), oracle.lookup([l.numpy() for l in levels], c.numpy(), r))


def test_channels_last_output_is_the_same_tensor():
    """raftcorr_lookup_fwd_nhwc: same values as the NCHW entry point, laid out (B,H,W,C) — through both
    the TMA kernel (4 levels, r=4/3) and the generic-stride kernel (other level counts / radii)."""
    gen = torch.Generator().manual_seed(23)
    for (B, H, W, L, r) in [(2, 55, 128, 4, 4), (1, 46, 62, 4, 4), (2, 17, 23, 4, 3), (1, 16, 95, 3, 2), (1, 9, 33, 2, 1)]:
        pyr = ops.Pyramid(B, H, W, L, "cuda", zero=True)
        for l in range(L):
            pyr.storage[l][:, :, : pyr.desc[l].W].normal_()
        c = coords_for("sigma4", B, H, W, gen).cuda()
        a = ops.lookup(pyr, c, r)
        b = ops.lookup(pyr, c, r, channels_last=True)
        assert b.shape == a.shape and b.is_contiguous(memory_format=torch.channels_last)
        assert torch.equal(a, b), (B, H, W, L, r)


@pytest.mark.parametrize("r", [4, 3])
@pytest.mark.parametrize("kind", ["grid", "half", "sigma4", "sigma30", "wild"])
def test_backward_tma_reduce_path_equals_atomic_scatter(kind, r):
    """d(pyramid) through TMA reduce-add boxes (the path RAFT takes: coords detached) against the atomic
    scatter kernel (forced by also asking for d coords).  Same terms, so only the summation order differs;
    'grid' has floor flips on ~8 % of its taps (irregular windows fall back per query), 'sigma30' pushes
    windows across the border, 'wild' has NaN / inf / huge coordinates.  Padding columns stay untouched."""
    B, H, W = 2, 30, 44
    gen = torch.Generator().manual_seed(100 + r)
    if kind == "wild":
        c = coords_for("sigma4", B, H, W, gen)
        c[0, 0, 3, 5], c[0, 1, 4, 6], c[1, 0, 7, 7], c[1, 1, 0, 0] = float("nan"), float("inf"), 3.0e38, -1.0e9
        c[1, :, 10, 10] = torch.tensor([1.7e7, -2.5e7])
    else:
        c = coords_for(kind, B, H, W, gen)
    c = c.cuda()
    gout = torch.randn(B, 4 * (2 * r + 1) ** 2, H, W, generator=gen).cuda()
    pyr = ops.Pyramid(B, H, W, 4, "cuda", zero=True)
    ga = ops.Pyramid(B, H, W, 4, "cuda", zero=True)
    gb = ops.Pyramid(B, H, W, 4, "cuda", zero=True)
    for _ in range(2):      # two calls accumulate, like two refinement iterations
        ops.lookup_backward(None, ga, c, gout, r, False)
        ops.lookup_backward(pyr, gb, c, gout, r, True)
    torch.cuda.synchronize()
    for l in range(4):
        a, b = ga.storage[l], gb.storage[l]
        assert torch.isfinite(a).all()
        scale = b.abs().max().item()
        assert torch.allclose(a, b, rtol=1e-5, atol=1e-6 * max(scale, 1.0)), (l, (a - b).abs().max().item())
        assert torch.equal(a[:, :, ga.desc[l].W:], torch.zeros_like(a[:, :, ga.desc[l].W:]))
