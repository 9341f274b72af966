#!/bin/bash
# one GPU call: the no-alloc test alone (fresh process), the whole GPU suite, then a bench line
# usage: scripts/gpu_round.sh <tag> [bench args...]
mkdir -p gpurun_out
TAG="${1:-r02}"; shift
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/pytest_zz_$TAG.log 2>&1; echo "zz alone exit=$?"
tail -n 3 gpurun_out/pytest_zz_$TAG.log
timeout 1500 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/pytest_full_$TAG.log 2>&1; echo "full suite exit=$?"
grep -E "passed|failed|error" gpurun_out/pytest_full_$TAG.log | tail -n 5
grep -E "^(FAILED|ERROR)" gpurun_out/pytest_full_$TAG.log | head -n 20
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -n 2
timeout 900 python bench.py "$@" > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit=$?"; tail -n 5 gpurun_out/bench_$TAG.err
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_$TAG.json"))
    print("pairs/s", round(d["value"],1), "ms/step", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["value"],1), "launches", d["gpu_launches"], "roofline", d["roofline"]["frac"])
    for k in d["kernels"]: print("  %-24s %7.3f ms  x%-3d rate %.1f" % (k["tag"], k["ms_per_step"], k["launches_per_step"], k["rate"] or 0))
    print("gpu_library_baseline", json.dumps(d.get("gpu_library_baseline",{}).get("modes")))
    print("cpu_baseline", json.dumps(d.get("cpu_baseline")))
    print("parity", json.dumps(d.get("parity")))
except Exception as e: print("bench parse failed", e)
PY
