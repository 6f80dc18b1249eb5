"""raftcorr_build_pyramid_noalloc: the build for callers that cannot provide a workspace.

Some users of this library receive device pointers for inputs and outputs but have no allocator
to ask for scratch memory (a plug-in inside a host that owns the whole device pool, or a loop in
which cudaMalloc is banned because it stalls the GPU).  For them the library carries a fixed 8 MiB
scratch array of its own.  This test drives that entry point through the C ABI exactly as such a
caller would - outputs allocated up front, nothing allocated around the calls - and checks it
against raftcorr_build_pyramid (caller workspace) on two back-to-back calls: a large problem
first, then a smaller one with different data, so that operands left in the shared array by the
first call would show up as errors in the second.  Collected last on purpose (file name)."""
import ctypes

import pytest
import torch

from raftcorr_b200 import _capi, ops

pytestmark = pytest.mark.gpu

TOL = 1e-2  # the bf16/TF32 bar used by tests/test_gpu_build.py, relative to max|ref|


def _build_both(fmap1, fmap2, precision):
    B, D, H, W = fmap1.shape
    lib = _capi.lib()
    prec = ops.resolve_precision(precision)
    ref = ops.build_pyramid(fmap1, fmap2, 4, precision)          # existing entry point
    out = ops.Pyramid(B, H, W, 4, fmap1.device, zero=True)       # caller-owned outputs
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    free_before, _ = torch.cuda.mem_get_info()    # cudaMemGetInfo: sees raw cudaMalloc too, not only torch's pool
    rc = lib.raftcorr_build_pyramid_noalloc(fmap1.data_ptr(), fmap2.data_ptr(), B, D, H, W, 4, out.desc, prec,
                                            ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.raftcorr_last_error()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before   # nothing went through torch's allocator either
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after >= free_before, (free_before, free_after)   # ... and nothing was taken from the device
    return ref, out


@pytest.mark.parametrize("precision", ["bf16", "tf32"])
def test_noalloc_matches_workspace_build_twice_in_a_row(precision):
    gen = torch.Generator(device="cuda").manual_seed(77)
    shapes = [(1, 256, 46, 62), (2, 64, 17, 23)] if precision == "bf16" else [(1, 128, 46, 62), (1, 64, 24, 40)]
    for shape in shapes:
        f1 = torch.randn(*shape, device="cuda", generator=gen)
        f2 = torch.randn(*shape, device="cuda", generator=gen)
        ref, out = _build_both(f1, f2, precision)
        for l in range(4):
            a, b = out.views()[l], ref.views()[l]
            assert a.shape == b.shape
            err = (a - b).abs().max().item() / b.abs().max().item()
            assert err <= TOL, (precision, shape, l, err)
            assert torch.equal(a, b)   # same kernels, same operands: in fact bit-identical
        if shape == shapes[-1]:    # ... and once against the checker itself, not only against the sibling entry point
            import oracle
            want = oracle.build_pyramid(f1.cpu(), f2.cpu(), 4)
            for l in range(4):
                got = out.views()[l].cpu().numpy()
                assert got.shape == want[l].shape
                assert abs(got - want[l]).max() <= TOL * abs(want[l]).max(), (precision, l)


def test_noalloc_rejects_what_does_not_fit():
    lib = _capi.lib()
    B, D, H, W = 8, 256, 55, 128     # 57.7 MB of packed bf16 operands > 8 MiB
    f = torch.zeros(1, device="cuda")
    pyr = ops.Pyramid(1, 8, 8, 4, "cuda")   # descriptors only matter after the size check
    lv = _capi.LevelArray()
    n = (ctypes.c_int64 * 4)()
    assert lib.raftcorr_pyramid_layout(B, H, W, 4, lv, n) == 0
    for l in range(4):
        lv[l].data = pyr.storage[0].data_ptr()
    rc = lib.raftcorr_build_pyramid_noalloc(f.data_ptr(), f.data_ptr(), B, D, H, W, 4, lv, _capi.PREC_BF16, None)
    assert rc == -3
    msg = lib.raftcorr_last_error().decode()
    assert "scratch" in msg and str(_capi.NOALLOC_SCRATCH_BYTES) in msg
    rc = lib.raftcorr_build_pyramid_noalloc(f.data_ptr(), f.data_ptr(), 1, 8, 16, 16, 4, lv, _capi.PREC_FP32, None)
    assert rc == -1 and b"precision" in lib.raftcorr_last_error()


def test_noalloc_odd_channel_count_takes_the_scratch_free_build():
    """D % 8 != 0 cannot use the tensor-core operand layout; like raftcorr_build_pyramid the entry point then
    runs the fp32 CUDA-core build (which needs no scratch) instead of refusing."""
    import oracle
    gen = torch.Generator(device="cuda").manual_seed(5)
    f1 = torch.randn(1, 12, 16, 24, device="cuda", generator=gen)
    f2 = torch.randn(1, 12, 16, 24, device="cuda", generator=gen)
    _, out = _build_both(f1, f2, "bf16")
    want = oracle.build_pyramid(f1.cpu(), f2.cpu(), 4)
    for l in range(4):
        got = out.views()[l].cpu().numpy()
        assert abs(got - want[l]).max() <= 1e-5 * abs(want[l]).max(), l
