#!/bin/bash
# bench lines of the other BASELINE configurations (builder-run): adaptive, superglue, sweep
# usage: scripts/gpu_workloads.sh <tag> [sweep pairs]
mkdir -p gpurun_out
TAG="${1:-r02}"; SW="${2:-8192}"
for w in adaptive superglue; do
  timeout 600 python bench.py --workload $w --steps 5 --warmup 3 > gpurun_out/bench_${w}_$TAG.json 2> gpurun_out/bench_${w}_$TAG.err; echo "$w exit=$?"; tail -n 3 gpurun_out/bench_${w}_$TAG.err
done
timeout 900 python bench.py --workload sweep --sweep-pairs $SW > gpurun_out/bench_sweep_$TAG.json 2> gpurun_out/bench_sweep_$TAG.err; echo "sweep exit=$?"; tail -n 3 gpurun_out/bench_sweep_$TAG.err
python - <<PY
import json
for w in ("adaptive","superglue","sweep"):
    try:
        d=json.load(open("gpurun_out/bench_%s_$TAG.json" % w))
        print(w, "pairs/s", round(d["value"],1), "ms/step", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["value"],1), "roofline", d["roofline"]["kernel"], round(d["roofline"]["achieved"],1), d["roofline"]["unit"], round(d["roofline"]["frac"],3), "cpu", d.get("cpu_baseline",{}).get("value"))
        for k in d["kernels"][:6]: print("   %-24s %7.3f ms x%d" % (k["tag"], k["ms_per_step"], k["launches_per_step"]))
        print("   ", json.dumps(d.get("adaptive") or d.get("sweep") or {}))
    except Exception as e: print(w, "failed", e)
PY
