#!/bin/bash
# bench (plain) && ncu launch list of the same command, then --set full captures of the top kernels
# usage: scripts/gpu_profile.sh <tag> [what: list att gemm conv]
mkdir -p gpurun_out
TAG="${1:-r01}"; shift
WHAT="${*:-list att gemm conv}"
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/bench_small.json 2> gpurun_out/bench_small.err || { echo "plain run failed"; tail -5 gpurun_out/bench_small.err; exit 1; }
cap() {  # name, regex, skip, count
  ncu --set full --clock-control none --import-source on -k regex:"$2" -s $3 -c $4 -f -o gpurun_out/prof_${TAG}_$1 $CMD > gpurun_out/ncu_$1.log 2>&1
  echo "capture $1 exit=$?"
}
for w in $WHAT; do case $w in
  list) ncu --metrics gpu__time_duration.sum --clock-control none -c 100000 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list.log 2>&1; echo "launch list exit=$?";;
  att) cap attention "attention_tcgen05" 60 1;;
  gemm) cap gemm "gemm_tcgen05" 14 8;;
  conv) cap conv "conv_stem_kernel|conv_patch_kernel|nms_list_kernel" 6 5;;
esac; done
ls -la gpurun_out | tail -8
