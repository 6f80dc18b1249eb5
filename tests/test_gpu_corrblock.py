"""The drop-in classes end to end (build + lookup), forward and backward, against the golden
vectors produced by the reference CorrBlock and its autograd."""
import warnings

import numpy as np
import pytest
import torch

import oracle
from oracle import torch_restatement as tr
from helpers import expected_zero_mask, rel_max

warnings.filterwarnings("ignore")
pytestmark = pytest.mark.gpu


def _t(a, grad=False):
    return torch.from_numpy(np.array(a)).cuda().requires_grad_(grad)


@pytest.mark.parametrize("kind", ("grid", "half", "sigma4", "sigma30"))
@pytest.mark.parametrize("precision,tol", [("fp32", 1e-5), ("tf32", 1e-2), ("bf16", 1e-2)])
def test_corrblock_forward_golden(golden, kind, precision, tol):
    import corr
    g = golden("corr_tiny")
    with torch.no_grad():
        blk = corr.CorrBlock(_t(g["fmap1"]), _t(g["fmap2"]), num_levels=4, radius=4, precision=precision)
        out = blk(_t(g[f"coords_{kind}"]))
    assert blk.num_levels == 4 and blk.radius == 4
    assert [tuple(p.shape) for p in blk.corr_pyramid] == [(288, 1, 16, 18), (288, 1, 8, 9), (288, 1, 4, 4), (288, 1, 2, 2)]
    ref = g[f"out_{kind}"]
    assert out.shape == ref.shape and out.is_contiguous() and out.dtype == torch.float32
    o = out.cpu().numpy()
    assert np.array_equal(o == 0, expected_zero_mask(blk, g[f"coords_{kind}"], 4, ref))   # zero-padding masks: exact
    assert rel_max(o, ref) <= tol


def test_corrblock_backward_golden(golden):
    """d fmap1, d fmap2, d coords vs the reference's autograd.
    Atomics make the sums order-dependent: allclose(rtol 1e-3, atol 1e-4 max)."""
    import corr
    g = golden("corr_tiny")
    f1, f2, c = _t(g["fmap1"], True), _t(g["fmap2"], True), _t(g["coords_sigma4"], True)
    blk = corr.CorrBlock(f1, f2, precision="fp32")
    out = blk(c)
    assert rel_max(out.detach().cpu().numpy(), g["out_sigma4"]) <= 1e-5
    out.backward(_t(g["grad_out"]))
    for got, key in ((f1.grad, "grad_fmap1"), (f2.grad, "grad_fmap2"), (c.grad, "grad_coords")):
        ref = g[key]
        assert got is not None, key
        assert np.allclose(got.cpu().numpy(), ref, rtol=1e-3, atol=1e-4 * np.abs(ref).max()), key


def test_twelve_lookups_share_one_gradient_buffer(golden):
    """RAFT's training graph: one build, many lookups with detached coords (raft.py:122-124).
    The gradient of a sum over lookups must equal the reference autograd's."""
    import corr
    g = golden("corr_tiny")
    gen = torch.Generator().manual_seed(2)
    coords = [torch.from_numpy(g["coords_sigma4"]) + 0.37 * k + 0.1 * torch.randn(1, 2, 16, 18, generator=gen)
              for k in range(3)]
    ups = [torch.randn(1, 324, 16, 18, generator=gen) for _ in range(3)]
    f1r = torch.from_numpy(g["fmap1"]).requires_grad_(True)
    f2r = torch.from_numpy(g["fmap2"]).requires_grad_(True)
    rb = tr.TorchCorrBlock(f1r, f2r, 4, 4)
    sum((rb(c) * u).sum() for c, u in zip(coords, ups)).backward()
    f1, f2 = _t(g["fmap1"], True), _t(g["fmap2"], True)
    blk = corr.CorrBlock(f1, f2, precision="fp32")
    sum((blk(c.cuda()) * u.cuda()).sum() for c, u in zip(coords, ups)).backward()
    for got, ref in ((f1.grad, f1r.grad), (f2.grad, f2r.grad)):
        ref = ref.numpy()
        assert np.allclose(got.cpu().numpy(), ref, rtol=1e-3, atol=1e-4 * np.abs(ref).max())


def test_partial_backward_leaves_nothing_behind():
    """A backward pass that scatters into the shared d(pyramid) buffers but never reaches the build node
    (torch.autograd.grad w.r.t. coords prunes it) must not leak its sums into the next, full pass."""
    import corr
    gen = torch.Generator(device="cuda").manual_seed(3)
    f1 = torch.randn(1, 32, 16, 24, device="cuda", generator=gen, requires_grad=True)
    f2 = torch.randn(1, 32, 16, 24, device="cuda", generator=gen, requires_grad=True)
    ys, xs = torch.meshgrid(torch.arange(16.0), torch.arange(24.0), indexing="ij")
    c = (torch.stack([xs, ys])[None] + 0.3).cuda().requires_grad_(True)
    up = torch.randn(1, 324, 16, 24, device="cuda", generator=gen)

    blk = corr.CorrBlock(f1, f2)
    out = blk(c)
    (gc,) = torch.autograd.grad(out, c, up, retain_graph=True)      # scatters, BuildFn.backward never runs
    assert torch.isfinite(gc).all()
    out.backward(up)                                                 # the full pass
    got1, got2 = f1.grad.clone(), f2.grad.clone()
    f1.grad = f2.grad = c.grad = None
    corr.CorrBlock(f1, f2)(c).backward(up)                           # a fresh block, one pass
    assert torch.equal(got1, f1.grad) and torch.equal(got2, f2.grad)


def test_no_grad_builds_no_graph(golden):
    import corr
    g = golden("corr_tiny")
    f1, f2 = _t(g["fmap1"], True), _t(g["fmap2"], True)
    with torch.no_grad():
        out = corr.CorrBlock(f1, f2)(_t(g["coords_grid"]))
    assert not out.requires_grad and out.grad_fn is None


@pytest.mark.parametrize("kind", ("grid", "half", "sigma4", "sigma30"))
@pytest.mark.parametrize("precision,tol", [("fp32", 2e-5), ("bf16", 1e-2)])
def test_alternate_corr_block_matches_corrblock_golden(golden, kind, precision, tol):
    """AlternateCorrBlock computes the same numbers as CorrBlock (pooling commutes with the dot
    product); with CorrBlock's coordinate semantics the zero masks agree exactly as well.
    'bf16' is the fused tensor-core kernel (csrc/alt_tc.cu), 'fp32' the per-level CUDA-core one."""
    import corr
    g = golden("corr_tiny")
    with torch.no_grad():
        blk = corr.AlternateCorrBlock(_t(g["fmap1"]), _t(g["fmap2"]), num_levels=4, radius=4, precision=precision)
        out = blk(_t(g[f"coords_{kind}"]))
    assert len(blk.pyramid) == 5 and blk.pyramid[1][1].shape == (1, 8, 8, 9)
    ref = g[f"out_{kind}"]
    o = out.cpu().numpy()
    assert o.shape == ref.shape and out.is_contiguous()
    assert np.array_equal(o == 0, ref == 0)
    assert rel_max(o, ref) <= tol


def test_alternate_corr_block_odd_sizes_radius3(golden):
    import corr
    g = golden("corr_odd_r3")
    with torch.no_grad():
        out = corr.AlternateCorrBlock(_t(g["fmap1"]), _t(g["fmap2"]), radius=3)(_t(g["coords_sigma4"]))
    ref = g["out_sigma4"]
    o = out.cpu().numpy()
    assert np.array_equal(o == 0, ref == 0)
    assert rel_max(o, ref) <= 1e-2


@pytest.mark.parametrize("flow", ("smooth", "iid", "wild"))
def test_alternate_tensor_core_vs_fp32_path_full_size(flow):
    """BASELINE.json configs[3] shape (136x240 fmaps, D=256; B=1 here): the fused tensor-core kernel
    against this library's fp32 per-level kernel, which is pinned to the oracle / golden vectors
    above.  'wild' adds NaN / inf / 1e9 coordinates: same NaN pattern, no fault."""
    import corr
    B, D, H, W = 1, 256, 136, 240
    gen = torch.Generator(device="cuda").manual_seed(31)
    f1 = torch.randn(B, D, H, W, device="cuda", generator=gen)
    f2 = torch.randn(B, D, H, W, device="cuda", generator=gen)
    ys, xs = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
    grid = torch.stack([xs, ys]).float()[None]
    if flow == "smooth":
        low = 8 * torch.randn(B, 2, H // 8, W // 8, device="cuda", generator=gen)
        c = grid + torch.nn.functional.interpolate(low, size=(H, W), mode="bilinear", align_corners=True)
    else:
        c = grid + 6 * torch.randn(B, 2, H, W, device="cuda", generator=gen)
    if flow == "wild":
        c[0, 0, 0, 0] = float("nan")
        c[0, 1, 3, 5] = float("inf")
        c[0, 0, 7, 9] = 1e9
        c[0, 1, 9, 200] = -1e9
        c[0, 0, 100, 100] = 3e38
    with torch.no_grad():
        a = corr.AlternateCorrBlock(f1, f2, precision="bf16")(c)
        b = corr.AlternateCorrBlock(f1, f2, precision="fp32")(c)
    assert a.shape == b.shape == (B, 324, H, W)
    assert torch.equal(torch.isnan(a), torch.isnan(b))
    a, b = torch.nan_to_num(a), torch.nan_to_num(b)
    assert torch.equal(a == 0, b == 0)
    assert (a - b).abs().max().item() <= 1e-2 * b.abs().max().item()


def test_alt_cuda_corr_module_surface(golden):
    """alt_cuda_corr.forward/backward: list returns, reference layouts, unscaled, CUDA+contiguous
    checks raising RuntimeError (correlation.cpp:19-21), floor(coords) semantics."""
    import alt_cuda_corr
    g = golden("corr_tiny")
    f1 = _t(g["fmap1"]).permute(0, 2, 3, 1).contiguous()
    f2 = _t(g["fmap2"]).permute(0, 2, 3, 1).contiguous()
    c = _t(g["coords_sigma4"]).permute(0, 2, 3, 1).reshape(1, 1, 16, 18, 2).contiguous()
    res = alt_cuda_corr.forward(f1, f2, c, 4)
    assert isinstance(res, list) and len(res) == 1 and res[0].shape == (1, 1, 81, 16, 18)
    ref = oracle.alt_forward(f1.cpu().numpy(), f2.cpu().numpy(), c.cpu().numpy(), 4, oracle.COORD_DIRECT)
    assert rel_max(res[0].cpu().numpy(), ref) <= 1e-5
    assert np.array_equal(res[0].cpu().numpy() == 0, ref == 0)
    with pytest.raises(RuntimeError, match="contiguous"):
        alt_cuda_corr.forward(f1.transpose(1, 2), f2, c, 4)
    with pytest.raises(RuntimeError, match="CUDA"):
        alt_cuda_corr.forward(f1.cpu(), f2, c, 4)
    up = torch.randn(res[0].shape, device="cuda")
    grads = alt_cuda_corr.backward(f1, f2, c, up, 4)
    assert isinstance(grads, list) and [tuple(t.shape) for t in grads] == [tuple(f1.shape), tuple(f2.shape), tuple(c.shape)]


@pytest.mark.parametrize("precision", [None, "fp32"])
def test_alternate_corr_block_backward(precision):
    """Gradients of the on-the-fly path vs autograd through its torch restatement (CPU): the default block
    (fused tensor-core forward, fp32 per-level backward) and the all-fp32 one."""
    import corr
    gen = torch.Generator().manual_seed(8)
    B, D, H, W = 1, 16, 16, 24
    f1 = torch.randn(B, D, H, W, generator=gen)
    f2 = torch.randn(B, D, H, W, generator=gen)
    ys, xs = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    c = torch.stack([xs, ys]).float()[None] + 3.0 * torch.randn(B, 2, H, W, generator=gen)
    up = torch.randn(B, 324, H, W, generator=gen)
    a, b_, cc = f1.clone().requires_grad_(True), f2.clone().requires_grad_(True), c.clone().requires_grad_(True)
    tr.alt_corr_block_math(a, b_, cc, 4, 4).backward(up)
    x, y, z = f1.cuda().requires_grad_(True), f2.cuda().requires_grad_(True), c.cuda().requires_grad_(True)
    out = corr.AlternateCorrBlock(x, y, coord_mode="direct", precision=precision)(z)
    with torch.no_grad():
        plain = corr.AlternateCorrBlock(x, y, coord_mode="direct", precision=precision)(z)
    assert torch.equal(out, plain)   # recording a graph does not change the forward values
    out.backward(up.cuda())
    for got, ref in ((x.grad, a.grad), (y.grad, b_.grad), (z.grad, cc.grad)):
        ref = ref.numpy()
        assert np.allclose(got.cpu().numpy(), ref, rtol=1e-3, atol=1e-4 * np.abs(ref).max())
