#!/bin/bash
# FFN-1 (Linear + LayerNorm + GELU) experiments: accumulator part width 128 / 256
mkdir -p gpurun_out
TAG="${1:-ffn}"
LOG=gpurun_out/ffn_exp_$TAG.log
: > $LOG
for part in 256 128; do
  echo "== LN part $part: tests" >> $LOG
  B200GLUE_LN_PART=$part timeout 900 python -m pytest tests/test_gpu_dense.py tests/test_gpu_e2e.py -m gpu -q -p no:cacheprovider -k "ffn or lightglue_vs_oracle or attention" 2>&1 | tail -4 >> $LOG
  echo "== LN part $part: bench" >> $LOG
  B200GLUE_LN_PART=$part timeout 300 python bench.py --steps 10 --warmup 5 --no-cpu-baseline --no-gpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('pairs/s', round(d['value'],1), 'ms/step', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1))
for k in d['kernels'][:9]: print('  %-24s %7.3f ms  x%d' % (k['tag'], k['ms_per_step'], k['launches_per_step']))
" >> $LOG 2>&1
done
cat $LOG
