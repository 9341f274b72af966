#!/bin/bash
# attention alone, default library and every variant, both layouts at the bench shape
mkdir -p gpurun_out
LOG=gpurun_out/attn_variants_${1:-v}.log
: > $LOG
for lib in "" $(ls glue_factory_b200/build/variants/libb200glue_*.so 2>/dev/null); do
  for layout in wide narrow; do
    echo "== lib=${lib:-default} layout=$layout" >> $LOG
    for rep in 1 2; do
      B200GLUE_LIB=$lib B200GLUE_ATT_LAYOUT=$layout timeout 120 python scripts/bench_attention.py 64 2048 40 2>&1 | grep "^attention" >> $LOG
    done
  done
done
cat $LOG
