#!/bin/bash
# bench + launch list (ncu) on one GPU; logs into gpurun_out/
mkdir -p gpurun_out
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit=$?"; tail -n 2 gpurun_out/smoke.log
python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit=$?"
tail -n 5 gpurun_out/bench.err; cat gpurun_out/bench.json
