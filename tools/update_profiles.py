"""Turn one gpurun capture set into the tracked files under profiles/ (run here, on the CPU box):
   update_profiles.py <round tag, e.g. r02> <bench.json> <launches.csv> <full .ncu-rep> "<ncu command of the .ncu-rep>"
Writes  profiles/<tag>_bench_1gpu.json, <tag>_bench_launches.csv (+ _summary.txt), <tag>_ncu_full_bench.txt and
refreshes the lookup16_kernel entry of profiles/lookup_traffic.json (keyed on the blob hash of csrc/lookup16.cu)."""
import collections
import csv
import io
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PROF = os.path.join(ROOT, "profiles")
tag, bench, launches, rep, ncu_cmd = sys.argv[1:6]

shutil.copy(bench, os.path.join(PROF, f"{tag}_bench_1gpu.json"))
shutil.copy(launches, os.path.join(PROF, f"{tag}_bench_launches.csv"))
rows = [r for r in csv.reader(open(launches)) if len(r) > 14 and r[0].isdigit()]
agg = collections.OrderedDict()
for r in rows:
    name, ns = r[4], float(r[14])
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += ns / 1e3
total = sum(v[1] for v in agg.values())
with open(os.path.join(PROF, f"{tag}_bench_launches_summary.txt"), "w") as f:
    f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none -c {len(rows)} --csv  python bench.py --steps 2 --warmup 3 "
            f"--skip-cpu-baseline  ({tag}, B200)\n")
    f.write(f"# {len(rows)} launches captured (first {len(rows)} of the run: warm-up and timed steps), {total / 1e3:.1f} ms of kernel time; "
            "per-launch times are cold-cache and serialised: compare SHARES\n")
    f.write("# launches     total us   share    avg us  kernel\n")
    for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"{n:10d} {us:12.1f} {us / total * 100:6.2f}% {us / n:9.2f}  {name[:110]}\n")
summary = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), rep], capture_output=True, text=True).stdout
with open(os.path.join(PROF, f"{tag}_ncu_full_bench.txt"), "w") as f:
    f.write(f"# {ncu_cmd}\n# ({tag}, B200; read with tools/ncu_summary.py)\n")
    f.write(summary)

raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(io.StringIO(raw)))
hdr, units = rr[0], rr[1]
look = [v for v in rr[2:] if "lookup16_kernel" in v[hdr.index("Kernel Name")]]
if len(look) >= 2:
    v = look[1]                                    # refinement iteration 1 (iteration 0 = integer grid)

    def val(name):
        i = hdr.index(name)
        x = float(v[i])
        return x * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "us": 1.0, "ns": 1e-3, "ms": 1e3}.get(units[i], 1.0)
    blob = subprocess.run(["git", "hash-object", os.path.join(ROOT, "pytorch-raft_b200", "csrc", "lookup16.cu")],
                          capture_output=True, text=True).stdout.strip()
    path = os.path.join(PROF, "lookup_traffic.json")
    t = json.load(open(path))
    rd, wr = val("dram__bytes_read.sum"), val("dram__bytes_write.sum")
    t["kernels"]["lookup16_kernel"] = {
        "workload": "cfg2_b8_55x128", "source_file": "lookup16.cu", "source_blob": blob,
        "capture": f"{ncu_cmd} (profiles/{tag}_ncu_full_bench.txt, second lookup16 launch = refinement iteration 1)",
        "dram_bytes_read": int(rd), "dram_bytes_write": int(wr), "dram_bytes_per_launch": int(rd + wr),
        "algorithmic_bytes_per_launch": 8 * 55 * 128 * 2104, "duration_us": round(val("gpu__time_duration.sum"), 2)}
    json.dump(t, open(path, "w"), indent=1)
    print("lookup16 traffic:", int(rd + wr), "bytes per launch, blob", blob)
print("wrote", tag)
