"""Where does a lookup16 warp spend its time?  Builds a copy of the library with -DRAFTCORR_L16_TRACE (clock64 around
the phases of every unit: prepare / wait for the boxes / sample / issue the next boxes) and prints the per-warp sums.
Development aid.   usage: trace16.py build   (here, no GPU)   |   trace16.py run [smooth|grid]   (on the B200)"""
import ctypes
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "pytorch-raft_b200")
LIB = os.path.join(ROOT, "tools", "probe", "libraftcorr_trace.so")


def build(extra=()):
    sys.path[:0] = [PKG]
    from raftcorr_b200 import build as b
    cmd = ["/usr/local/cuda/bin/nvcc"] + b.NVCC_FLAGS + ["-DRAFTCORR_L16_TRACE", *extra, "-o", LIB] + b.sources()
    subprocess.check_call(cmd)
    print(LIB)


def run(kind):
    os.environ["RAFTCORR_LIB"] = LIB
    sys.path[:0] = [ROOT, PKG]
    import numpy as np
    import torch
    from raftcorr_b200 import _capi, ops
    B, D, H, W = 8, 256, 55, 128
    torch.manual_seed(0)
    f1 = torch.randn(B, D, H, W, device="cuda")
    f2 = torch.randn(B, D, H, W, device="cuda")
    pyr = ops.build_pyramid(f1, f2, 4, "bf16", storage="bf16")
    keep = []
    for l in range(int(os.environ.get("TRACE16_INTERLEAVE", "0"))):   # levels 0..n-1 with 4 images interleaved
        s = pyr.storage[l]
        n, by, bx = s.shape[:3]
        keep.append(s.reshape(n // 4, 4, by, bx, 32).permute(0, 2, 3, 1, 4).contiguous())
        pyr.desc[l].data = keep[-1].data_ptr()
        pyr.desc[l].interleave_log2 = 2
    ys, xs = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
    grid = torch.stack([xs, ys]).float()[None].repeat(B, 1, 1, 1)
    low = 8 * torch.randn(B, 2, H // 8, W // 8, device="cuda")
    smooth = grid + torch.nn.functional.interpolate(low, size=(H, W), mode="bilinear", align_corners=True)
    c = smooth if kind == "smooth" else grid
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        ops.lookup(pyr, c, 4)
        b.record()
        torch.cuda.synchronize()
    us = a.elapsed_time(b) * 1e3
    n = 148 * 8 * 32
    buf = (ctypes.c_ulonglong * n)()
    fn = _capi.lib().raftcorr_debug_l16_trace
    fn.argtypes = [ctypes.c_void_p, ctypes.c_int]
    rc = fn(buf, n)
    assert rc == 0, rc
    t = np.frombuffer(buf, dtype=np.uint64).reshape(148, 8, 32).astype(np.float64)
    units = t[..., 4]
    print(f"{kind}: {us:.1f} us per launch (traced build); cycles per unit, mean over SMs, by warp (level = warp // 2):")
    print("  warp  units   prepare     wait   sample    issue   | whole warp (cycles)")
    for w in range(8):
        u = units[:, w].mean()
        print(f"  {w:4d} {u:6.1f} " + " ".join(f"{(t[:, w, k] / units[:, w]).mean():8.0f}" for k in range(4)) +
              f"   | {t[:, w, 5].mean():9.0f}")
    print("  by round (mean over SMs and warps): wait / sample / issue")
    for r in range(8):
        sel = units > r
        if sel.any():
            print(f"   round {r}: " + " ".join(f"{t[..., 8 + 3 * r + k][sel].mean():8.0f}" for k in range(3)) +
                  f"   ({int(sel.sum())} warps)")
    tot = t[..., :4].sum(axis=(0, 1)) / units.sum()
    print("  all          " + " ".join(f"{v:8.0f}" for v in tot) + f"   sum {tot.sum():.0f} cycles per unit")


if __name__ == "__main__":
    if sys.argv[1] == "build":
        build(sys.argv[2:])
    else:
        run(sys.argv[2] if len(sys.argv) > 2 else "smooth")
