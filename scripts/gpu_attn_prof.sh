#!/bin/bash
# source-level capture of the attention kernel alone (stall samples per SASS instruction); args: tag layout
mkdir -p gpurun_out
TAG="${1:-attn}"; LAYOUT="${2:-wide}"
export B200GLUE_ATT_LAYOUT=$LAYOUT
python scripts/bench_attention.py 64 2048 5 || exit 1
ncu --set full --import-source on --clock-control none -k regex:attention_ -s 4 -c 1 -f -o gpurun_out/att_$TAG python scripts/bench_attention.py 64 2048 3 > gpurun_out/att_ncu_$TAG.log 2>&1; echo "capture exit=$?"
ncu -i gpurun_out/att_$TAG.ncu-rep --page source --csv --print-source sass > gpurun_out/att_source_$TAG.csv 2>gpurun_out/att_source_$TAG.err
ncu -i gpurun_out/att_$TAG.ncu-rep --page raw --csv > gpurun_out/att_raw_$TAG.csv 2>/dev/null
ncu -i gpurun_out/att_$TAG.ncu-rep --page details > gpurun_out/att_details_$TAG.txt 2>/dev/null
sz=$(stat -c %s gpurun_out/att_$TAG.ncu-rep); echo "report bytes $sz"; if [ $sz -gt 20000000 ]; then rm -f gpurun_out/att_$TAG.ncu-rep; fi
ls -la gpurun_out | tail -8
if [ -x scripts/microbench/softmax_mix ]; then timeout 120 scripts/microbench/softmax_mix | tee gpurun_out/softmax_mix.log; fi
