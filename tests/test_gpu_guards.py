"""Out-of-bounds writes would land in guard bands: every output buffer of the round-2 kernels (16-bit pyramid build,
lookup16, alt_pack) is carved out of a larger allocation whose surroundings hold a sentinel pattern, at ragged sizes
(tiles that end inside a warp, odd level sizes, groups that end at the batch border), and the pattern must survive.
(compute-sanitizer is not available on the GPU pool; this is the bounds check of our own.)"""
import ctypes
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-raft_b200")]
from raftcorr_b200 import _capi, ops  # noqa: E402

pytestmark = pytest.mark.gpu
GUARD = 1 << 16           # bytes on each side
SENTINEL = 0x5A


class Guarded:
    """nbytes of device memory (256-byte aligned) between two guard bands."""

    def __init__(self, nbytes):
        self.n = (nbytes + 255) // 256 * 256
        self.buf = torch.full((GUARD + self.n + GUARD,), SENTINEL, dtype=torch.uint8, device="cuda")
        assert self.buf.data_ptr() % 256 == 0
        self.ptr = self.buf.data_ptr() + GUARD

    def payload(self):
        return self.buf[GUARD:GUARD + self.n]

    def intact(self):
        return bool((self.buf[:GUARD] == SENTINEL).all() and (self.buf[GUARD + self.n:] == SENTINEL).all())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize("shape", [(2, 16, 10, 10), (1, 24, 9, 33), (3, 16, 12, 14), (1, 64, 46, 62), (2, 8, 17, 23)])
def test_build16_and_lookup16_stay_inside_their_buffers(shape):
    B, D, H, W = shape
    lib = _capi.lib()
    gen = torch.Generator().manual_seed(H * W)
    f1 = torch.randn(B, D, H, W, generator=gen).cuda()
    f2 = torch.randn(B, D, H, W, generator=gen).cuda()
    desc = _capi.Level16Array()
    nbytes = (ctypes.c_int64 * _capi.MAX_LEVELS)()
    assert lib.raftcorr_pyramid16_layout(B, H, W, 4, desc, nbytes) == 0
    levels = [Guarded(int(nbytes[l])) for l in range(4)]
    for l in range(4):
        desc[l].data = levels[l].ptr
    ws_bytes = int(lib.raftcorr_build_workspace_bytes(B, D, H, W, _capi.PREC_BF16))
    ws = Guarded(ws_bytes)
    rc = lib.raftcorr_build_pyramid16(f1.data_ptr(), f2.data_ptr(), B, D, H, W, 4, desc, ctypes.c_void_p(ws.ptr), ws_bytes,
                                      _stream())
    assert rc == 0, lib.raftcorr_last_error()
    torch.cuda.synchronize()
    assert ws.intact() and all(g.intact() for g in levels), shape
    # the guarded pyramid holds what the ordinary build produces
    ref = ops.build_pyramid(f1, f2, 4, "bf16", storage="bf16")
    for l in range(4):
        n = ref.storage[l].numel() * 2
        assert torch.equal(levels[l].payload()[:n], ref.storage[l].reshape(-1).view(torch.uint8)), (shape, l)
    out = Guarded(B * 324 * H * W * 4)
    for sigma in (0.0, 4.0, 40.0):
        coords = torch.stack(torch.meshgrid(torch.arange(W), torch.arange(H), indexing="xy")).float()[None].repeat(B, 1, 1, 1)
        coords = (coords + sigma * torch.randn(B, 2, H, W, generator=gen)).cuda().contiguous()
        rc = lib.raftcorr_lookup16_fwd(desc, 4, coords.data_ptr(), ctypes.c_void_p(out.ptr), B, H, W, 4, 0, _stream())
        assert rc == 0, lib.raftcorr_last_error()
        torch.cuda.synchronize()
        assert out.intact() and all(g.intact() for g in levels), (shape, sigma)
        want = ops.lookup(ref, coords, 4, "aten_cpu")
        got = out.payload()[: B * 324 * H * W * 4].view(torch.float32).reshape(B, 324, H, W)
        assert torch.equal(torch.nan_to_num(got), torch.nan_to_num(want)), (shape, sigma)


@pytest.mark.parametrize("shape", [(1, 24, 9, 70), (2, 128, 17, 33), (1, 200, 8, 8)])
def test_alt_pack_stays_inside_its_buffers(shape):
    B, D, H, W = shape
    lib = _capi.lib()
    fmap = torch.randn(B, D, H, W, device="cuda")
    outs = [Guarded(B * (H >> l) * (W >> l) * D * 2) for l in range(4)]
    ptrs = (ctypes.c_void_p * 4)(*[g.ptr for g in outs])
    rc = lib.raftcorr_alt_pack(fmap.data_ptr(), B, D, H, W, 4, ptrs, _stream())
    assert rc == 0, lib.raftcorr_last_error()
    torch.cuda.synchronize()
    assert all(g.intact() for g in outs), shape
    want = ops.alt_pack(fmap, 4)
    for l in range(4):
        n = want[l].numel() * 2
        assert torch.equal(outs[l].payload()[:n], want[l].reshape(-1).view(torch.uint8)), (shape, l)
