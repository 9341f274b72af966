"""Summarise an `ncu --page source --csv --print-source sass` export: total stall samples by reason, the top
instructions by samples, and per-opcode executed counts.  usage: python scripts/ncu_source_stalls.py file.csv [top]"""
import csv
import sys
from collections import Counter

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path)))
hdr = rows[1]
col = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
data = rows[2:]
tot = Counter()
ops = Counter()
samples = []
for k, r in enumerate(data):
    if len(r) < len(hdr):
        continue
    n = int(r[col["# Samples"]] or 0)
    ex = int(r[col["Instructions Executed"]] or 0)
    src = r[col["Source"]].strip()
    op = src.split()[0] if not src.startswith("@") else src.split()[1]
    ops[op] += ex
    for h in stall_cols:
        tot[h] += int(r[col[h]] or 0)
    samples.append((n, k, src, ex, {h: int(r[col[h]] or 0) for h in stall_cols if int(r[col[h]] or 0)}))
all_s = sum(tot.values())
print("total samples", all_s)
for h, v in tot.most_common():
    print(f"  {h:28s} {v:8d} {100.0 * v / all_s:5.1f}%")
print("\ntop instructions by samples (index, samples %, executed, source, reasons)")
for n, k, src, ex, why in sorted(samples, reverse=True)[:top]:
    w = " ".join(f"{h[6:]}={v}" for h, v in sorted(why.items(), key=lambda kv: -kv[1])[:3])
    print(f"  {k:5d} {100.0 * n / all_s:5.2f}% ex={ex:8d}  {src[:70]:70s} {w}")
print("\nexecuted warp instructions by opcode")
te = sum(ops.values())
for op, v in ops.most_common(25):
    print(f"  {op:34s} {v:10d} {100.0 * v / te:5.1f}%")
