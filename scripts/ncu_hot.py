"""Top stall lines of an ncu report's source page:  python scripts/ncu_hot.py <report.ncu-rep> [kernel-index] [n]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks, cur = [], []
for line in out.splitlines():
    if line.startswith('"Kernel Name"'):
        if cur:
            blocks.append(cur)
        cur = [line]
    elif cur:
        cur.append(line)
if cur:
    blocks.append(cur)
blk = blocks[which]
print(blk[0][:150])
rows = list(csv.reader(io.StringIO("\n".join(blk[1:]))))
hdr = rows[0]
col = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
data = []
total = 0
for r in rows[1:]:
    try:
        n = int(r[col["# Samples"]])
    except (ValueError, IndexError):
        continue
    total += n
    data.append((n, r))
data_sorted = sorted(enumerate(data), key=lambda t: -t[1][0])[:topn]
print("total samples", total)
agg = {}
for n, r in data:
    for h in stall_cols:
        try:
            agg[h] = agg.get(h, 0) + int(r[col[h]])
        except ValueError:
            pass
print("stall totals:", sorted(((v, k) for k, v in agg.items() if v), reverse=True)[:8])
for idx, (n, r) in data_sorted:
    st = sorted(((int(r[col[h]]), h[6:]) for h in stall_cols if r[col[h]].isdigit() and int(r[col[h]]) > 0), reverse=True)[:3]
    print(f"{idx:5d} {n:6d} {100.0*n/total:5.1f}%  {r[col['Source']].strip()[:70]:70s} {st}")
