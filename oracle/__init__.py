"""CPU oracle for RAFT's correlation hot path — TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package, and only as the checker.  The product
(``pytorch-raft_b200/``) never imports it and has no CPU fallback.

Two restatements of the reference (zhaoyuzhi/PyTorch-RAFT, ``core/corr.py``) live here:

* ``corr_oracle.c`` (loaded through ctypes by this module): explicit fp32 loops that spell out
  the arithmetic hidden inside ATen (``bmm``, ``avg_pool2d``, ``grid_sampler_2d``).
* ``torch_restatement.py``: the same op sequence as the reference, written against torch ops,
  used to time "what the reference does on host cores" and, on a GPU box, as a second oracle
  running ATen's CUDA kernels.

Parity pin: ``oracle/make_golden.py`` imports the unmodified reference from ``/root/reference``
in the build container and stores its outputs under ``tests/golden/``; ``tests/test_oracle.py``
checks both restatements against those vectors.  The reference has no tests or golden vectors
of its own for this path, so these generated vectors are the pin.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")

COORD_ATEN_CPU = 0   # utils.py:61-62 evaluated by ATen's CPU kernels (true division)
COORD_ATEN_CUDA = 1  # same expression evaluated by ATen's CUDA kernels (x * (1/(W-1)))
COORD_DIRECT = 2     # alt_cuda_corr kernel: floor(coords) directly, no normalise round trip


def build(force: bool = False) -> str:
    """Compile corr_oracle.c with the committed Makefile (gcc, seconds)."""
    src = os.path.join(_HERE, "corr_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"],
                              stdout=subprocess.DEVNULL)
    return _LIB_PATH


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        _lib.oracle_num_threads.restype = ctypes.c_int
    return _lib


def _fp(a: np.ndarray):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def _f32(a) -> np.ndarray:
    if hasattr(a, "detach"):
        a = a.detach().cpu().numpy()
    return np.ascontiguousarray(a, dtype=np.float32)


def num_threads() -> int:
    return int(lib().oracle_num_threads())


def set_num_threads(n: int) -> None:
    lib().oracle_set_num_threads(ctypes.c_int(int(n)))


def level_dims(H: int, W: int, num_levels: int):
    """corr.py:25-27: avg_pool2d(2, stride=2) floors odd sizes."""
    return [(H >> l, W >> l) for l in range(num_levels)]


def corr_volume(fmap1, fmap2) -> np.ndarray:
    """corr.py:52-60.  fmaps (B,D,H,W) -> (B,H,W,1,H,W) fp32."""
    f1, f2 = _f32(fmap1), _f32(fmap2)
    B, D, H, W = f1.shape
    vol = np.empty((B, H * W, H * W), dtype=np.float32)
    lib().oracle_corr_volume(_fp(f1), _fp(f2), _fp(vol), B, D, H, W)
    return vol.reshape(B, H, W, 1, H, W)


def avg_pool2(x) -> np.ndarray:
    """corr.py:26.  (n,1,H,W) -> (n,1,H//2,W//2)."""
    x = _f32(x)
    n, c, H, W = x.shape
    out = np.empty((n, c, H // 2, W // 2), dtype=np.float32)
    lib().oracle_avg_pool2(_fp(x), _fp(out), ctypes.c_int64(n * c), H, W)
    return out


def build_pyramid(fmap1, fmap2, num_levels: int = 4):
    """corr.py:13-27 (CorrBlock.__init__): list of (B*N,1,H_l,W_l) fp32 arrays."""
    vol = corr_volume(fmap1, fmap2)
    B, H, W = vol.shape[:3]
    pyr = [vol.reshape(B * H * W, 1, H, W)]
    for _ in range(num_levels - 1):
        pyr.append(avg_pool2(pyr[-1]))
    return pyr


def lookup(pyramid, coords, radius: int = 4, mode: int = COORD_ATEN_CPU) -> np.ndarray:
    """corr.py:29-50 (CorrBlock.__call__).  coords (B,2,H,W) -> (B, L*(2r+1)^2, H, W)."""
    levels = [_f32(p) for p in pyramid]
    c = _f32(coords)
    B, _, H, W = c.shape
    L = len(levels)
    rd = 2 * radius + 1
    out = np.empty((B, L * rd * rd, H, W), dtype=np.float32)
    ptrs = (ctypes.POINTER(ctypes.c_float) * L)(*[_fp(p) for p in levels])
    Hl = (ctypes.c_int * L)(*[p.shape[-2] for p in levels])
    Wl = (ctypes.c_int * L)(*[p.shape[-1] for p in levels])
    pitch = (ctypes.c_int64 * L)(*[p.shape[-1] for p in levels])
    istr = (ctypes.c_int64 * L)(*[p.shape[-1] * p.shape[-2] for p in levels])
    lib().oracle_lookup(ptrs, Hl, Wl, pitch, istr, _fp(c), _fp(out), B, H, W, L, radius, mode)
    return out


def alt_forward(fmap1_nhwc, fmap2_nhwc, coords, radius: int, mode: int = COORD_DIRECT) -> np.ndarray:
    """alt_cuda_corr.forward (correlation_kernel.cu:18-119, :260-286): one level, UNSCALED.

    fmap1 (B,H1,W1,C), fmap2 (B,H2,W2,C), coords (B,1,H1,W1,2) -> (B,1,rd*rd,H1,W1).
    """
    f1, f2, c = _f32(fmap1_nhwc), _f32(fmap2_nhwc), _f32(coords)
    B, H1, W1, C = f1.shape
    _, H2, W2, _ = f2.shape
    assert c.shape == (B, 1, H1, W1, 2)
    rd = 2 * radius + 1
    out = np.empty((B, 1, rd * rd, H1, W1), dtype=np.float32)
    lib().oracle_alt_forward(_fp(f1), _fp(f2), _fp(c), _fp(out), B, H1, W1, H2, W2, C, radius, mode)
    return out


def alt_corr_block(fmap1, fmap2, coords, num_levels: int = 4, radius: int = 4,
                   mode: int = COORD_DIRECT) -> np.ndarray:
    """corr.py:63-91 (AlternateCorrBlock): pooled fmap2 pyramid, per-level alt_forward, /sqrt(D)."""
    f1, f2, c = _f32(fmap1), _f32(fmap2), _f32(coords)
    B, D, H, W = f1.shape
    f1_nhwc = np.ascontiguousarray(f1.transpose(0, 2, 3, 1))
    cc = c.transpose(0, 2, 3, 1)
    outs = []
    f2_l = f2
    for l in range(num_levels):
        f2_nhwc = np.ascontiguousarray(f2_l.transpose(0, 2, 3, 1))
        coords_l = np.ascontiguousarray((cc / np.float32(2 ** l)).reshape(B, 1, H, W, 2))
        outs.append(alt_forward(f1_nhwc, f2_nhwc, coords_l, radius, mode)[:, 0])
        f2_l = avg_pool2(f2_l)
    corr = np.stack(outs, axis=1).reshape(B, -1, H, W)
    return corr / np.sqrt(np.float32(D))


def upsample_flow(flow, mask) -> np.ndarray:
    """RAFT.upsample_flow (core/raft.py:72-83).  flow (B,2,H,W), mask (B,576,H,W) -> (B,2,8H,8W)."""
    f, m = _f32(flow), _f32(mask)
    B, _, H, W = f.shape
    assert m.shape == (B, 576, H, W)
    up = np.empty((B, 2, 8 * H, 8 * W), dtype=np.float32)
    lib().oracle_upsample_flow(_fp(f), _fp(m), _fp(up), B, H, W)
    return up


def upflow8(flow) -> np.ndarray:
    """upflow8 (core/utils/utils.py:80-82).  flow (B,2,H,W) -> (B,2,8H,8W)."""
    f = _f32(flow)
    B, _, H, W = f.shape
    up = np.empty((B, 2, 8 * H, 8 * W), dtype=np.float32)
    lib().oracle_upflow8(_fp(f), _fp(up), B, H, W)
    return up
