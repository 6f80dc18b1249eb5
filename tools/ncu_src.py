"""Per-kernel hot spots of an .ncu-rep source page: usage ncu_src.py rep launch_index [top]"""
import csv, io, subprocess, sys
rep, which = sys.argv[1], int(sys.argv[2])
top_n = int(sys.argv[3]) if len(sys.argv) > 3 else 30
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
starts = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
s = starts[which]
e = starts[which + 1] - 1 if which + 1 < len(starts) else len(rows)
hdr = rows[s]
data = [r for r in rows[s + 1:e] if len(r) == len(hdr)]
iS, iN, iM = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
iW, iWi = hdr.index("L1 Wavefronts Shared"), hdr.index("L1 Wavefronts Shared Ideal")
tot = sum(int(r[iN]) for r in data) or 1
tots = sum(int(r[iM]) for r in data) or 1
print(f"== launch {which}: {len(data)} SASS instructions, {tot} executed (warp-level), {tots} samples")
agg = {}
for r in data:
    for i in stall:
        try:
            agg[hdr[i]] = agg.get(hdr[i], 0) + int(r[i])
        except ValueError:
            pass
print("   stall totals:", ", ".join(f"{k[6:]}={v/tots*100:.1f}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v / tots > 0.01))
wf = sum(int(r[iW] or 0) for r in data); wfi = sum(int(r[iWi] or 0) for r in data)
print(f"   shared wavefronts {wf} (ideal {wfi})")
ops = {}
for r in data:
    op = r[iS].strip().split()[0] if r[iS].strip() else "?"
    if op.startswith("@"):
        op = r[iS].strip().split()[1]
    op = op.split(".")[0]
    ops[op] = ops.get(op, 0) + int(r[iN])
print("   executed by opcode:", ", ".join(f"{k}={v/tot*100:.1f}%" for k, v in sorted(ops.items(), key=lambda kv: -kv[1])[:18]))
top = sorted(range(len(data)), key=lambda k: -int(data[k][iM]))[:top_n]
for k in sorted(top):
    r = data[k]
    why = max(stall, key=lambda i: int(r[i] or 0))
    print(f"   [{k:4d}] exec={int(r[iN]):9d} samples={int(r[iM])/tots*100:5.1f}% {hdr[why][6:]:>10s} wf={r[iW]}/{r[iWi]}  {r[iS].strip()[:90]}")
