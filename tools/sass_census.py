"""Instruction census per kernel of the built library (cuobjdump -sass): the SASS mnemonics that prove tcgen05 / TMA.
   usage: sass_census.py [round tag] > profiles/<tag>_sass_census.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "pytorch-raft_b200", "libraftcorr_b200.so")
COLS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UTMAPF", "UBLKCP", "SYNCS",
        "HMMA", "FFMA2", "LDGSTS"]
tag = sys.argv[1] if len(sys.argv) > 1 else "r02"
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
counts, order, cur = collections.defaultdict(collections.Counter), [], None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        order.append(cur)
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and cur:
        op = m.group(1)
        counts[cur]["total"] += 1
        for c in COLS:
            if op == c or (c == "HMMA" and op.startswith("HMMA")):
                counts[cur][c] += 1
names = subprocess.run(["c++filt"], input="\n".join(order), capture_output=True, text=True).stdout.splitlines()
print(f"# cuobjdump -sass pytorch-raft_b200/libraftcorr_b200.so (sm_100a), {tag}: instruction census per kernel")
print("# UTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG / UTMASTG / UTMAREDG = TMA tensor load / store / reduce-add, "
      "SYNCS = mbarrier ops, FFMA2 = packed fp32x2 fma; HMMA (legacy mma.sync) must be 0")
print("#" + "".join(f"{c:>9s}" for c in ["total"] + COLS) + "  kernel")
tot = collections.Counter()
for f, n in zip(order, names):
    short = re.sub(r"raftcorr::|\(anonymous namespace\)::|<unnamed>::", "", n)
    short = re.sub(r"\(.*", "", short)
    print(" " + "".join(f"{counts[f][c]:9d}" for c in ["total"] + COLS) + "  " + short[:70])
    tot.update(counts[f])
print("#" + "".join(f"{tot[c]:9d}" for c in ["total"] + COLS) + "  all kernels")
