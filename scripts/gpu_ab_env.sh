#!/bin/bash
# A/B of an environment switch inside the bench step: scripts/gpu_ab_env.sh VAR value_a value_b [kernel tag substring]
mkdir -p gpurun_out
VAR=$1; A=$2; B=$3; TAG=${4:-}
for v in $A $B $A $B; do
  echo "== $VAR=$v"
  env $VAR=$v timeout 300 python bench.py --steps 10 --warmup 5 --no-cpu-baseline --no-gpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('pairs/s', round(d['value'],1), 'ms/step', round(d['ms_per_step'],3))
for k in d['kernels']:
    if '$TAG' in k['tag']: print('  %-24s %7.3f ms  x%d' % (k['tag'], k['ms_per_step'], k['launches_per_step']))
"
done
