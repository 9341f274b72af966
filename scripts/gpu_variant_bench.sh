#!/bin/bash
# bench the default library and every variant under glue_factory_b200/build/variants (kernel table only)
mkdir -p gpurun_out
LOG=gpurun_out/variant_bench_${1:-v}.log
: > $LOG
for lib in "" $(ls glue_factory_b200/build/variants/libb200glue_*.so 2>/dev/null); do
  echo "== lib=${lib:-default}" >> $LOG
  B200GLUE_LIB=$lib timeout 300 python bench.py --steps 10 --warmup 5 --no-cpu-baseline --no-gpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('pairs/s', round(d['value'],1), 'ms/step', round(d['ms_per_step'],3))
for k in d['kernels'][:10]: print('  %-24s %7.3f ms  x%d' % (k['tag'], k['ms_per_step'], k['launches_per_step']))
" >> $LOG 2>&1
done
cat $LOG
