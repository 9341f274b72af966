"""Summarise gpurun_out/launches.csv (+ an optional --set full report) into profiles/<tag>_*.{md,csv}."""
import collections
import csv
import re
import subprocess
import sys

tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
rep = sys.argv[2] if len(sys.argv) > 2 else None

rows = list(csv.reader(open("gpurun_out/launches.csv")))
hdr, agg = None, collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    if "Kernel Name" in r:
        hdr = r
        continue
    if hdr is None or len(r) != len(hdr):
        continue
    d = dict(zip(hdr, r))
    if d.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*", "", d["Kernel Name"])
    scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0}.get(d["Metric Unit"][:2], 1e-6)
    agg[name][0] += 1
    agg[name][1] += float(d["Metric Value"].replace(",", "")) * scale
tot = sum(v[1] for v in agg.values())
with open(f"profiles/{tag}_launches.md", "w") as f:
    f.write(f"# ncu launch list ({tag}): `ncu --metrics gpu__time_duration.sum --clock-control none` on "
            "`python bench.py --steps 1 --warmup 3 --no-cpu-baseline`\n\n"
            f"{sum(v[0] for v in agg.values())} launches captured, {tot:.1f} ms of kernel time "
            "(cold-cache, serialised: compare shares, not absolutes).\n\n| kernel | launches | total ms | share |\n|---|---|---|---|\n")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:30]:
        f.write(f"| `{k[:100]}` | {v[0]} | {v[1]:.2f} | {100 * v[1] / tot:.1f} % |\n")
if rep:
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rr = list(csv.reader(raw.splitlines()))
    h, u = rr[0], rr[1]
    want = ["Kernel Name", "launch__grid_size", "launch__registers_per_thread", "gpu__time_duration.sum",
            "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers"]
    idx = [h.index(w) for w in want if w in h]
    with open(f"profiles/{tag}_full_capture.csv", "w", newline="") as f:
        w = csv.writer(f)
        w.writerow([h[i] + (f" [{u[i]}]" if u[i] else "") for i in idx])
        for r in rr[2:]:
            w.writerow([re.sub(r"\(.*", "", r[i]) if h[i] == "Kernel Name" else r[i] for i in idx])
    print("wrote full capture summary")
print("wrote", f"profiles/{tag}_launches.md")
