#!/bin/bash
# bench (plain) && ncu launch list of the same command, then one --set full capture of the top kernels
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/bench_small.json 2> gpurun_out/bench_small.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "launch list exit=$?"
$CMD > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"attention_tcgen05|gemm_tcgen05" -s 60 -c 14 -o gpurun_out/prof_r01 $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture exit=$?"
ls -la gpurun_out | tail -8
