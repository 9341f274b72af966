#!/bin/bash
# round-2 evidence: plain bench, ncu launch list of the same command, --set full captures
# usage: scripts/gpu_profile_r02.sh <tag>
mkdir -p gpurun_out
TAG="${1:-r02}"
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-gpu-baseline"
$CMD > gpurun_out/bench_small_$TAG.json 2> gpurun_out/bench_small_$TAG.err || { echo "plain run failed"; tail -5 gpurun_out/bench_small_$TAG.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 100000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1; echo "launch list exit=$?"
# one step's attention + bandwidth kernels (31 matching launches per step; skip the 3 warm-up steps + packing pass)
ncu --set full --clock-control none -k regex:"attention_tcgen05|nms_list|select_list|sample_descriptors|ingest_kernel|posenc_kernel|rowdot_kernel|assign_stats|assign_finalize" -s 93 -c 31 -f -o gpurun_out/prof_${TAG}_step $CMD > gpurun_out/ncu_step.log 2>&1; echo "step capture exit=$?"
ncu --set full --clock-control none -k regex:"gemm_tcgen05|conv_patch|conv_stem" -s 150 -c 30 -f -o gpurun_out/prof_${TAG}_dense $CMD > gpurun_out/ncu_dense.log 2>&1; echo "dense capture exit=$?"
python scripts/prof_io.py > gpurun_out/prof_io.log 2>&1 || { echo "prof_io failed"; tail -3 gpurun_out/prof_io.log; }
ncu --set full --clock-control none -k regex:"prep_|export_pack|eval_" -s 12 -c 6 -f -o gpurun_out/prof_${TAG}_io python scripts/prof_io.py > gpurun_out/ncu_io.log 2>&1; echo "io capture exit=$?"
# gpurun copies back at most 64 MiB: keep the raw-metric CSV of every report, and the report itself only when small
for r in gpurun_out/prof_${TAG}_*.ncu-rep; do
  ncu -i $r --page raw --csv > ${r%.ncu-rep}.csv 2>/dev/null
  python scripts/ncu_summary.py $r > ${r%.ncu-rep}.txt 2>/dev/null
  sz=$(stat -c %s $r); if [ $sz -gt 12000000 ]; then rm -f $r; fi
done
ls -la gpurun_out | tail -12; du -sh gpurun_out
