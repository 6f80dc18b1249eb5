"""Build a copy of the library with extra nvcc flags (kernel variants for A/B measurements):
   build_variant.py <name> [-DFLAG=VALUE ...]  ->  tools/probe/libraftcorr_<name>.so   (use with RAFTCORR_LIB=...)"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "pytorch-raft_b200")]
from raftcorr_b200 import build as b  # noqa: E402

out = os.path.join(ROOT, "tools", "probe", f"libraftcorr_{sys.argv[1]}.so")
subprocess.check_call(["/usr/local/cuda/bin/nvcc"] + b.NVCC_FLAGS + sys.argv[2:] + ["-o", out] + b.sources(),
                      stderr=subprocess.DEVNULL)
print(out)
