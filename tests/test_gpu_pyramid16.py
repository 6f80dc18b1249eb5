"""The 16-bit blocked pyramid (include/raftcorr.h: raftcorr_level16, raftcorr_build_pyramid16,
raftcorr_lookup16_fwd) through the C ABI.

Bars:  build   — <= 1e-2 of max|ref| against the oracle (north_star's bf16 bar), exactly the bf16 rounding of the
                 fp32-pyramid build, every padding pixel of every block +0.0;
       lookup  — zero masks identical and values BIT-EXACT against the oracle applied to the stored (bf16)
                 pixel values, on every coordinate family of tests/test_gpu_lookup.py;
       both    — at BASELINE.json's full sizes, against the fp32-pyramid kernels on the same values."""
import ctypes
import warnings

import numpy as np
import pytest
import torch

import oracle
from raftcorr_b200 import _capi, ops
from helpers import coords_for, pyramid_from_numpy, rel_max

warnings.filterwarnings("ignore")
pytestmark = pytest.mark.gpu

KINDS = ("grid", "half", "sigma4", "sigma30")


def _levels_numpy(pyr16):
    return [pyr16.dense(l).unsqueeze(1).cpu().numpy() for l in range(pyr16.num_levels)]


def _padding_is_zero(pyr16):
    for l in range(pyr16.num_levels):
        d = pyr16.desc[l]
        bits = pyr16.padded(l).contiguous().view(torch.int16)
        if bits[:, d.H:, :].ne(0).any() or bits[:, :, d.W:].ne(0).any():
            return False
    return True


@pytest.mark.parametrize("size,interleave", [((1, 9, 21), None), ((2, 8, 24), None), ((2, 8, 24), False), ((2, 8, 22), (2, 2, 0))])
def test_layout_matches_the_header_formula(size, interleave):
    """pixel (y, x) of image n at byte (n // G) * G * image_stride + (((y>>2)*blocks_x + (x>>3)) * G + n % G) * 64
    + (y&3)*16 + (x&7)*2, G = 2**interleave_log2 (include/raftcorr.h)."""
    B, H, W = size
    gen = torch.Generator().manual_seed(1)
    levels = [torch.randn(B * H * W, 1, max(H >> l, 1), max(W >> l, 1), generator=gen) for l in range(3)]
    pyr = ops.Pyramid16.from_dense(levels, B, H, W, interleave=interleave)
    want_g = list(interleave) if isinstance(interleave, tuple) else [2, 2, 0] if W % 4 == 0 and interleave is None else [0, 0, 0]
    for l in range(3):
        d = pyr.desc[l]
        assert (d.blocks_x, d.blocks_y, d.image_stride) == ((d.W + 7) // 8, (d.H + 3) // 4, ((d.W + 7) // 8) * ((d.H + 3) // 4) * 64)
        assert d.interleave_log2 == want_g[l]
        G = 1 << d.interleave_log2
        raw = pyr.storage[l].reshape(-1).view(torch.int16).cpu()
        want = levels[l].to(torch.bfloat16).view(torch.int16)
        for (n, y, x) in [(0, 0, 0), (5, d.H - 1, d.W - 1), (B * H * W - 1, d.H // 2, d.W // 3), (6, 0, d.W - 1)]:
            byte = (n // G) * G * d.image_stride + (((y >> 2) * d.blocks_x + (x >> 3)) * G + n % G) * 64 + (y & 3) * 16 + (x & 7) * 2
            assert raw[byte // 2].item() == want[n, 0, y, x].item(), (l, n, y, x)
        assert torch.equal(pyr.dense(l).cpu(), levels[l][:, 0].to(torch.bfloat16).float())
    assert _padding_is_zero(pyr)


SHAPES = [(1, 64, 16, 18), (2, 32, 17, 23), (1, 128, 46, 62), (1, 8, 55, 96), (2, 16, 24, 40), (1, 24, 9, 33),
          (1, 16, 8, 8), (1, 40, 33, 65),
          (2, 32, 10, 12), (3, 16, 9, 20)]   # W % 4 == 0 but H*W % 32 != 0: interleaved levels, tiles that end mid-warp


@pytest.mark.parametrize("interleave", [None, False])
@pytest.mark.parametrize("shape", SHAPES)
def test_build16_against_oracle_and_fp32_pyramid(shape, interleave):
    B, D, H, W = shape
    gen = torch.Generator().manual_seed(B * 1000 + H * W)
    f1 = torch.randn(B, D, H, W, generator=gen)
    f2 = torch.randn(B, D, H, W, generator=gen)
    pyr16 = ops.build_pyramid(f1.cuda(), f2.cuda(), 4, "bf16", storage="bf16", interleave=interleave)
    assert isinstance(pyr16, ops.Pyramid16)
    assert pyr16.desc[0].interleave_log2 == (2 if interleave is None and W % 4 == 0 else 0)
    want = oracle.build_pyramid(f1, f2, 4)
    wide = ops.build_pyramid(f1.cuda(), f2.cuda(), 4, "bf16", storage="fp32")
    for l in range(4):
        got = pyr16.dense(l).cpu().numpy()
        assert got.shape == want[l][:, 0].shape
        assert rel_max(got, want[l][:, 0]) <= 1e-2, (shape, l)
        # same accumulators, same scale, same pooling chain; only the final rounding differs
        assert torch.equal(pyr16.dense(l), wide.views()[l][:, 0].to(torch.bfloat16).float()), (shape, l)
    assert _padding_is_zero(pyr16), shape


def test_build16_overwrites_stale_padding():
    """The build must leave +0.0 in every padding pixel even when the buffers held garbage (the lookup relies on it)."""
    B, D, H, W = 1, 16, 55, 96          # level 3 is 6 x 12: 4 padding columns, 2 padding rows nobody's tile covers
    gen = torch.Generator().manual_seed(8)
    f1 = torch.randn(B, D, H, W, generator=gen).cuda()
    f2 = torch.randn(B, D, H, W, generator=gen).cuda()
    lib = _capi.lib()
    pyr = ops.Pyramid16(B, H, W, 4, "cuda")
    for s in pyr.storage:
        s.view(torch.int16).fill_(0x7FC1)        # NaNs everywhere
    ws_bytes = int(lib.raftcorr_build_workspace_bytes(B, D, H, W, _capi.PREC_BF16))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    rc = lib.raftcorr_build_pyramid16(f1.data_ptr(), f2.data_ptr(), B, D, H, W, 4, pyr.desc, ws.data_ptr(), ws_bytes,
                                      ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.raftcorr_last_error()
    torch.cuda.synchronize()
    assert _padding_is_zero(pyr)
    for l in range(4):
        assert torch.isfinite(pyr.dense(l)).all()


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("case", [(1, 16, 18, 4), (2, 17, 23, 3), (1, 46, 62, 4), (3, 9, 33, 4), (1, 16, 95, 3), (1, 16, 16, 4),
                                  (6, 40, 70, 4)])   # the last one: several tiles per CTA, the rings wrap
@pytest.mark.parametrize("interleave", [None, False, (2, 2, 0, 0)])
def test_lookup16_bit_exact_against_oracle(case, kind, interleave):
    B, H, W, r = case
    if isinstance(interleave, tuple) and (B * H * W) % 4 != 0:
        pytest.skip("an interleaved level needs a multiple of four images")
    gen = torch.Generator().manual_seed(H * 100 + W + r)
    levels = [torch.randn(B * H * W, 1, H >> l, W >> l, generator=gen) for l in range(4)]
    pyr = ops.Pyramid16.from_dense(levels, B, H, W, interleave=interleave)
    stored = _levels_numpy(pyr)                     # what the kernel reads: bf16 values, widened
    c = coords_for(kind, B, H, W, gen)
    out = ops.lookup(pyr, c.cuda(), r, "aten_cpu").cpu().numpy()
    ref = oracle.lookup(stored, c.numpy(), r)
    assert out.shape == ref.shape
    # (a level with H_l == 1 or W_l == 1 divides by zero in the reference, utils.py:61-62: NaN there, in both)
    assert np.array_equal(np.isnan(out), np.isnan(ref)), (case, kind)
    assert np.array_equal(out == 0, ref == 0), (case, kind)
    assert np.array_equal(np.nan_to_num(out, nan=0.0), np.nan_to_num(ref, nan=0.0)), (case, kind)


@pytest.mark.parametrize("mode", ["aten_cuda", "direct"])
def test_lookup16_other_coordinate_modes(mode):
    B, H, W, r = 2, 24, 40, 4
    gen = torch.Generator().manual_seed(12)
    levels = [torch.randn(B * H * W, 1, H >> l, W >> l, generator=gen) for l in range(4)]
    pyr = ops.Pyramid16.from_dense(levels, B, H, W)
    wide = pyramid_from_numpy(_levels_numpy(pyr), B, H, W)      # same values in the fp32 pyramid format
    for kind in KINDS:
        c = coords_for(kind, B, H, W, gen).cuda()
        assert torch.equal(ops.lookup(pyr, c, r, mode), ops.lookup(wide, c, r, mode)), (mode, kind)


def test_lookup16_wild_coordinates():
    """NaN / inf / 1e9 / -1e9 coordinates: no fault, same result as the oracle (NaN == NaN)."""
    B, H, W = 1, 16, 40
    gen = torch.Generator().manual_seed(3)
    levels = [torch.randn(B * H * W, 1, H >> l, W >> l, generator=gen) for l in range(4)]
    c = coords_for("sigma4", B, H, W, gen)
    c[0, 0, 0, 0] = float("nan")
    c[0, 1, 0, 1] = float("inf")
    c[0, 0, 0, 2] = -float("inf")
    c[0, 0, 1, 0] = 1e9
    c[0, 1, 1, 1] = -1e9
    c[0, 0, 1, 2] = 3e38
    c[0, 0, 2, 0] = 20000.7
    c[0, 1, 2, 1] = -70000.2
    c[0, :, 3, 3] = torch.tensor([1.7e7, -2.5e7])
    pyr = ops.Pyramid16.from_dense(levels, B, H, W)
    out = ops.lookup(pyr, c.cuda(), 4, "aten_cpu").cpu().numpy()
    ref = oracle.lookup(_levels_numpy(pyr), c.numpy(), 4)
    assert np.array_equal(np.isnan(out), np.isnan(ref))
    assert np.array_equal(np.nan_to_num(out, nan=0.0), np.nan_to_num(ref, nan=0.0))


def test_lookup16_every_window_needs_the_largest_box():
    """Every query looking at the same spot, placed so that every level-0 window needs the largest box (3 x 4
    blocks = a whole 768-byte slot, for all 32 lanes of a unit at once).  Same bit-exact bar."""
    B, H, W = 2, 24, 64
    gen = torch.Generator().manual_seed(77)
    levels = [torch.randn(B * H * W, 1, H >> l, W >> l, generator=gen) for l in range(4)]
    pyr = ops.Pyramid16.from_dense(levels, B, H, W)
    c = torch.empty(B, 2, H, W)
    c[:, 0] = 11.5      # x_lo = 7: the window straddles three block columns
    c[:, 1] = 7.5       # y_lo = 3: ... and four block rows
    c[1] += 0.25 * torch.randn(2, H, W, generator=gen).clamp(-1, 1) * 0.1
    out = ops.lookup(pyr, c.cuda(), 4, "aten_cpu").cpu().numpy()
    ref = oracle.lookup(_levels_numpy(pyr), c.numpy(), 4)
    assert np.array_equal(out, ref)


@pytest.mark.parametrize("shape", [(8, 55, 128), (4, 47, 156)])
def test_full_size_build16_and_lookup16(shape):
    """BASELINE.json configs[1] / configs[2] shapes, D=256: the 16-bit build against the fp32-pyramid build
    (rounded), then the 16-bit lookup against the fp32-pyramid lookup kernel on the same stored values: identical."""
    B, H, W = shape
    gen = torch.Generator(device="cuda").manual_seed(21)
    f1 = torch.randn(B, 256, H, W, device="cuda", generator=gen)
    f2 = torch.randn(B, 256, H, W, device="cuda", generator=gen)
    pyr16 = ops.build_pyramid(f1, f2, 4, "bf16", storage="bf16")
    wide = ops.build_pyramid(f1, f2, 4, "bf16", storage="fp32")
    for l in range(4):
        w = wide.desc[l].W
        rounded = wide.storage[l][:, :, :w].to(torch.bfloat16).float()
        assert torch.equal(pyr16.dense(l), rounded), l
        wide.storage[l][:, :, :w].copy_(rounded)          # now both pyramids hold the same values
    assert _padding_is_zero(pyr16)
    cgen = torch.Generator().manual_seed(5)
    for kind in ("grid", "sigma4", "sigma30"):
        c = coords_for(kind, B, H, W, cgen).cuda()
        a = ops.lookup(pyr16, c, 4)
        b = ops.lookup(wide, c, 4)
        assert a.shape == (B, 324, H, W) and a.is_contiguous()
        assert torch.equal(a, b), kind


def test_corrblock_defaults_to_the_16bit_pyramid():
    """The drop-in class: bf16 arithmetic (default) now also stores bf16; the public corr_pyramid attribute keeps
    core/corr.py's shapes and dtype; gradients still run through the fp32 pyramid."""
    import corr
    gen = torch.Generator().manual_seed(4)
    f1 = torch.randn(1, 64, 24, 40, generator=gen)
    f2 = torch.randn(1, 64, 24, 40, generator=gen)
    c = coords_for("sigma4", 1, 24, 40, gen)
    blk = corr.CorrBlock(f1.cuda(), f2.cuda(), radius=4)
    assert isinstance(blk._holder["pyr"], ops.Pyramid16)
    shapes = [tuple(t.shape) for t in blk.corr_pyramid]
    assert shapes == [(960, 1, 24, 40), (960, 1, 12, 20), (960, 1, 6, 10), (960, 1, 3, 5)]
    assert all(t.dtype == torch.float32 for t in blk.corr_pyramid)
    ref_pyr = oracle.build_pyramid(f1, f2, 4)
    ref = oracle.lookup(ref_pyr, c.numpy(), 4)
    out = blk(c.cuda()).cpu().numpy()
    assert np.array_equal(out == 0, ref == 0)
    assert rel_max(out, ref) <= 1e-2
    assert isinstance(corr.CorrBlock(f1.cuda(), f2.cuda(), precision="fp32")._holder["pyr"], ops.Pyramid)
    assert isinstance(corr.CorrBlock(f1.cuda(), f2.cuda(), storage="fp32")._holder["pyr"], ops.Pyramid)
    g = corr.CorrBlock(f1.cuda().requires_grad_(True), f2.cuda())
    assert isinstance(g._holder["pyr"], ops.Pyramid)


def test_lookup16_rejects_what_it_does_not_cover():
    lib = _capi.lib()
    pyr = ops.Pyramid16(1, 16, 16, 3, "cuda", zero=True)
    c = torch.zeros(1, 2, 16, 16, device="cuda")
    out = torch.empty(1, 3 * 81, 16, 16, device="cuda")
    rc = lib.raftcorr_lookup16_fwd(pyr.desc, 3, c.data_ptr(), out.data_ptr(), 1, 16, 16, 4, 0, None)
    assert rc == -2 and b"num_levels" in lib.raftcorr_last_error()
    pyr = ops.Pyramid16(1, 16, 16, 4, "cuda", zero=True)
    rc = lib.raftcorr_lookup16_fwd(pyr.desc, 4, c.data_ptr(), out.data_ptr(), 1, 16, 16, 2, 0, None)
    assert rc == -2
    pyr.desc[1].blocks_x += 1
    rc = lib.raftcorr_lookup16_fwd(pyr.desc, 4, c.data_ptr(), out.data_ptr(), 1, 16, 16, 4, 0, None)
    assert rc == -1 and b"block grid" in lib.raftcorr_last_error()
    f = torch.zeros(1, 12, 16, 16, device="cuda")
    pyr = ops.Pyramid16(1, 16, 16, 4, "cuda", zero=True)
    rc = lib.raftcorr_build_pyramid16(f.data_ptr(), f.data_ptr(), 1, 12, 16, 16, 4, pyr.desc, None, 0, None)
    assert rc == -2 and b"D % 8" in lib.raftcorr_last_error()


def test_interleaved_levels_are_checked():
    """interleave_log2 outside [0, 3], an image count that is no multiple of the group, a build asked for a layout
    it does not write: refused with RAFTCORR_E_INVALID / RAFTCORR_E_UNSUPPORTED before any launch."""
    lib = _capi.lib()
    c = torch.zeros(1, 2, 9, 21, device="cuda")
    out = torch.empty(1, 4 * 81, 9, 21, device="cuda")
    pyr = ops.Pyramid16(1, 9, 21, 4, "cuda", zero=True)        # 189 images: plain layout
    pyr.desc[0].interleave_log2 = 2
    rc = lib.raftcorr_lookup16_fwd(pyr.desc, 4, c.data_ptr(), out.data_ptr(), 1, 9, 21, 4, 0, None)
    assert rc == -1 and b"not a multiple" in lib.raftcorr_last_error()
    pyr.desc[0].interleave_log2 = 5
    rc = lib.raftcorr_lookup16_fwd(pyr.desc, 4, c.data_ptr(), out.data_ptr(), 1, 9, 21, 4, 0, None)
    assert rc == -1 and b"interleave_log2" in lib.raftcorr_last_error()
    f = torch.zeros(1, 16, 16, 16, device="cuda")
    pyr = ops.Pyramid16(1, 16, 16, 4, "cuda", zero=True, interleave=(2, 0, 0, 0))
    ws = torch.empty(1 << 20, dtype=torch.uint8, device="cuda")
    rc = lib.raftcorr_build_pyramid16(f.data_ptr(), f.data_ptr(), 1, 16, 16, 16, 4, pyr.desc, ws.data_ptr(), ws.numel(), None)
    assert rc == -2 and b"both" in lib.raftcorr_last_error()
    pyr = ops.Pyramid16(1, 16, 16, 4, "cuda", zero=True, interleave=(2, 2, 2, 0))
    rc = lib.raftcorr_build_pyramid16(f.data_ptr(), f.data_ptr(), 1, 16, 16, 16, 4, pyr.desc, ws.data_ptr(), ws.numel(), None)
    assert rc == -2 and b"only levels 0 and 1" in lib.raftcorr_last_error()
