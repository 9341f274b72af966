#!/bin/bash
# final records of a round: tests + smoke + bench (20 steps), the other workloads, the reference arm, the ncu launch list
mkdir -p gpurun_out
TAG="${1:-final}"
scripts/gpu_round.sh $TAG --steps 20 --warmup 5 > gpurun_out/round_$TAG.log 2>&1; grep -E "passed|failed|smoke|pairs/s" gpurun_out/round_$TAG.log | cut -c1-200
scripts/gpu_workloads.sh $TAG 8192 > gpurun_out/workloads_$TAG.log 2>&1; grep -E "pairs/s|exit" gpurun_out/workloads_$TAG.log
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_reference_$TAG.json 2>/dev/null; cut -c1-120 gpurun_out/bench_reference_$TAG.json
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-gpu-baseline"
$CMD > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 100000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1; echo "launch list exit=$?"
du -sh gpurun_out
