#!/bin/bash
# source-level capture (stall samples per SASS instruction) of one launch of a kernel inside the bench step
# usage: scripts/gpu_kernel_prof.sh <tag> <kernel regex> [launches to skip]
mkdir -p gpurun_out
TAG="$1"; REGEX="$2"; SKIP="${3:-20}"
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-gpu-baseline"
ncu --set full --import-source on --clock-control none --kernel-name-base demangled -k regex:"$REGEX" -s $SKIP -c 1 -f -o gpurun_out/k_$TAG $CMD > gpurun_out/k_ncu_$TAG.log 2>&1; echo "capture exit=$?"
ncu -i gpurun_out/k_$TAG.ncu-rep --page source --csv --print-source sass > gpurun_out/k_source_$TAG.csv 2>/dev/null
ncu -i gpurun_out/k_$TAG.ncu-rep --page raw --csv > gpurun_out/k_raw_$TAG.csv 2>/dev/null
sz=$(stat -c %s gpurun_out/k_$TAG.ncu-rep); echo "report bytes $sz"; if [ $sz -gt 20000000 ]; then rm -f gpurun_out/k_$TAG.ncu-rep; fi
