#!/bin/bash
# quick GPU check: selected tests (args: pytest -k expression, file) then a short bench with the per-kernel table
mkdir -p gpurun_out
K="$1"; F="${2:-tests/test_gpu_dense.py}"; TAG="${3:-q}"
if [ -n "$K" ]; then timeout 600 python -m pytest $F -m gpu -q -p no:cacheprovider -k "$K" 2>&1 | tail -8; fi
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit=$?"; tail -n 3 gpurun_out/bench_$TAG.err
python - <<PY
import json
d=json.load(open("gpurun_out/bench_$TAG.json"))
print("pairs/s", round(d["value"],1), "ms/step", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["value"],1), "launches", d["gpu_launches"])
for k in d["kernels"]: print("  %-24s %7.3f ms  x%-3d rate %.1f" % (k["tag"], k["ms_per_step"], k["launches_per_step"], k["rate"] or 0))
PY
