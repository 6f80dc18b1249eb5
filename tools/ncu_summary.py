"""Summarise an .ncu-rep (read on the CPU box): key raw metrics + hottest SASS regions."""
import csv, io, subprocess, sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__cycles_elapsed.max", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor",
        "smsp__average_warps_issue_stalled", "launch__grid_size", "launch__block_size", "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum"]
for k, vals in enumerate(rows[2:]):
    print(f"== launch {k}: {vals[hdr.index('Kernel Name')][:90] if 'Kernel Name' in hdr else ''}")
    for i, h in enumerate(hdr):
        if any(h == w or (w.endswith("stalled") and h.startswith(w)) or (w in ("sm__inst_executed_pipe_tensor",) and h.startswith(w)) for w in WANT):
            if h.startswith("smsp__average_warps_issue_stalled") and not h.endswith("per_issue_active.ratio"):
                continue
            try:
                if h.startswith("smsp__average_warps_issue_stalled") and float(vals[i]) < 0.15:
                    continue
            except ValueError:
                pass
            print(f"   {h} [{units[i]}] = {vals[i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
start = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
if start:
    hdr = rows[start[0]]
    end = start[1] - 1 if len(start) > 1 else len(rows)
    data = [r for r in rows[start[0] + 1:end] if len(r) == len(hdr)]
    iS, iN, iM = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
    tot = sum(int(r[iN]) for r in data) or 1
    tots = sum(int(r[iM]) for r in data) or 1
    print(f"== SASS: {len(data)} instructions, {tot} executed (warp-level), {tots} samples")
    top = sorted(range(len(data)), key=lambda k: -int(data[k][iM]))[:25]
    for k in sorted(top):
        print(f"   [{k:4d}] exec={int(data[k][iN]):9d} samples={int(data[k][iM])/tots*100:5.1f}%  {data[k][iS].strip()[:80]}")
