#!/bin/bash
# attention experiments of round 2: stream layouts (narrow / wide), issuer placement, softmax instruction mix
mkdir -p gpurun_out
TAG="${1:-attn}"
LOG=gpurun_out/attn_exp_$TAG.log
: > $LOG
timeout 120 gpurun_out/ubench_mufu >> $LOG 2>&1
timeout 900 python -m pytest tests/test_gpu_dense.py -m gpu -q -p no:cacheprovider -k "attention" 2>&1 | tail -8 | tee -a $LOG
for lib in "" $(ls glue_factory_b200/build/variants/libb200glue_*.so 2>/dev/null); do
  for layout in narrow wide; do
    for shape in "64 2048" "16 4096" "64 1024" "2 2048" "128 512"; do
      echo "== lib=${lib:-default} layout=$layout shape=$shape" >> $LOG
      B200GLUE_LIB=$lib B200GLUE_ATT_LAYOUT=$layout timeout 120 python scripts/bench_attention.py $shape 30 2>&1 | grep "^attention" >> $LOG
    done
  done
done
for layout in narrow wide; do
  echo "== bench layout=$layout" >> $LOG
  B200GLUE_ATT_LAYOUT=$layout timeout 300 python bench.py --steps 10 --warmup 5 --no-cpu-baseline --no-gpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('pairs/s', round(d['value'],1), 'ms/step', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1))
for k in d['kernels'][:3]: print('  %-24s %7.3f ms' % (k['tag'], k['ms_per_step']))
" >> $LOG 2>&1
done
cat $LOG
