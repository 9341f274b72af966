#!/bin/bash
# bench (plain) && ncu launch list of the same command, then --set full captures of the top kernels
# usage: scripts/gpu_profile.sh <tag>
mkdir -p gpurun_out
TAG="${1:-r01}"
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/bench_small.json 2> gpurun_out/bench_small.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 1206 -c 402 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "launch list exit=$?"
cap() {  # name, regex, skip, count
  ncu --set full --clock-control none --import-source on -k regex:"$2" -s $3 -c $4 -f -o gpurun_out/prof_${TAG}_$1 $CMD > gpurun_out/ncu_$1.log 2>&1
  echo "capture $1 exit=$?"
}
cap attention "attention_tcgen05" 60 2
cap gemm "gemm_tcgen05_kernel<256, 3|gemm_tcgen05_kernel<512, 2|gemm_tcgen05_kernel<256, 1" 150 6
cap conv "conv_patch_kernel|conv1a_kernel|nms_list_kernel" 12 5
ls -la gpurun_out | tail -12
