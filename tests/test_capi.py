"""CPU-side checks of the boundary: the C-ABI library loads, exports every symbol that
include/raftcorr.h declares, and rejects bad arguments without touching a GPU."""
import ctypes
import os
import re

import pytest

from raftcorr_b200 import _capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "raftcorr.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(raftcorr_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = _capi.lib()
    declared = _declared_symbols()
    assert len(declared) >= 11
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/raftcorr.h but not exported"
        assert name in _capi.SIGNATURES, f"{name} has no ctypes signature"
    assert sorted(_capi.SIGNATURES) == declared
    assert lib.raftcorr_version() >= 100


def test_pyramid_layout_matches_avg_pool_floor():
    lib = _capi.lib()
    lv = _capi.LevelArray()
    n = (ctypes.c_int64 * 4)()
    assert lib.raftcorr_pyramid_layout(8, 55, 128, 4, lv, n) == 0
    assert [(lv[i].H, lv[i].W) for i in range(4)] == [(55, 128), (27, 64), (13, 32), (6, 16)]
    assert lib.raftcorr_pyramid_layout(16, 47, 156, 4, lv, n) == 0
    assert [(lv[i].H, lv[i].W) for i in range(4)] == [(47, 156), (23, 78), (11, 39), (5, 19)]
    assert all(lv[i].pitch % 4 == 0 and lv[i].pitch >= lv[i].W for i in range(4))
    assert n[1] == 16 * 47 * 156 * 23 * 80
    # a level that would be empty is rejected, not silently produced
    assert lib.raftcorr_pyramid_layout(1, 4, 4, 4, lv, n) == -1
    assert b"empty" in lib.raftcorr_last_error()


def test_bad_arguments_are_rejected_before_any_launch():
    lib = _capi.lib()
    lv = _capi.LevelArray()
    assert lib.raftcorr_lookup_fwd(lv, 4, None, None, 1, 8, 8, 4, 0, None) == -1
    assert b"NULL" in lib.raftcorr_last_error()
    assert lib.raftcorr_lookup_fwd(lv, 4, 16, 16, 1, 8, 8, 9, 0, None) == -1   # radius
    assert b"radius" in lib.raftcorr_last_error()
    assert lib.raftcorr_lookup_fwd(lv, 4, 16, 16, 1, 8, 8, 4, 7, None) == -1   # coord_mode
    assert lib.raftcorr_lookup_fwd(lv, 9, 16, 16, 1, 8, 8, 4, 0, None) == -1   # num_levels
    assert lib.raftcorr_build_pyramid(None, None, 1, 8, 8, 8, 4, lv, 1, None, 0, None) == -1
    assert lib.raftcorr_alt_fwd(None, None, None, None, 1, 8, 8, 8, 8, 8, 4, 2, 1.0, None) == -1
    assert lib.raftcorr_upsample_flow_fwd(None, None, None, 1, 8, 8, None) == -1
    assert b"NULL" in lib.raftcorr_last_error()
    assert lib.raftcorr_upsample_flow_fwd(16, 16, 16, 1, 0, 8, None) == -1
    assert lib.raftcorr_upsample_flow_bwd(16, 16, 16, None, None, 1, 8, 8, None) == -1
    assert b"nothing to compute" in lib.raftcorr_last_error()
    assert lib.raftcorr_upflow8(16, 8, 1, 8, 8, None) == -1    # output not 16-byte aligned
    assert lib.raftcorr_build_workspace_bytes(8, 256, 55, 128, _capi.PREC_BF16) == 2 * 8 * 7040 * 256 * 2
    # fp32-accurate mode: hi/lo tf32 splits of both fmaps for the tensor-core path; the CUDA-core
    # fallback (D not a multiple of 8) needs none
    assert lib.raftcorr_build_workspace_bytes(8, 256, 55, 128, _capi.PREC_FP32) == 4 * 8 * 7040 * 256 * 4
    assert lib.raftcorr_build_workspace_bytes(8, 100, 55, 128, _capi.PREC_FP32) == 0


def test_python_boundary_refuses_cpu_tensors():
    """No CPU fallback: the drop-in classes raise on host tensors instead of computing."""
    import torch
    import corr
    import alt_cuda_corr

    f = torch.randn(1, 8, 16, 16)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        corr.CorrBlock(f, f)
    import upsample
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        upsample.upsample_flow(torch.zeros(1, 2, 4, 4), torch.zeros(1, 576, 4, 4))
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        upsample.upflow8(torch.zeros(1, 2, 4, 4))
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        alt_cuda_corr.forward(f.permute(0, 2, 3, 1).contiguous(), f.permute(0, 2, 3, 1).contiguous(),
                              torch.zeros(1, 1, 16, 16, 2), 4)


def test_header_is_plain_c(tmp_path):
    """The boundary is a C ABI: include/raftcorr.h must compile as C99 on its own and a C program must link
    against the library using nothing but that header."""
    import shutil
    import subprocess
    gcc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else shutil.which("gcc")
    if gcc is None:
        pytest.skip("no C compiler")
    header = os.path.join(ROOT, "include", "raftcorr.h")
    subprocess.check_call([gcc, "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-x", "c", header])
    src = tmp_path / "use.c"
    src.write_text('#include "raftcorr.h"\n#include <stdio.h>\n'
                   'int main(void) {\n'
                   '  raftcorr_level lv[RAFTCORR_MAX_LEVELS]; int64_t n[RAFTCORR_MAX_LEVELS];\n'
                   '  if (raftcorr_pyramid_layout(1, 46, 62, 4, lv, n) != 0) return 1;\n'
                   '  if (raftcorr_lookup_fwd(lv, 4, 0, 0, 1, 46, 62, 4, 0, 0) != RAFTCORR_E_INVALID) return 2;\n'
                   '  printf("%d %d %d %s\\n", raftcorr_version(), lv[3].H, lv[3].W, raftcorr_last_error());\n'
                   '  return 0;\n}\n')
    exe = tmp_path / "use"
    libdir = os.path.dirname(_capi.LIB_PATH)
    subprocess.check_call([gcc, "-std=c99", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe),
                           "-L", libdir, "-l:" + os.path.basename(_capi.LIB_PATH), "-Wl,-rpath," + libdir])
    out = subprocess.check_output([str(exe)], text=True)
    assert out.split()[1:3] == ["5", "7"] and "NULL" in out
