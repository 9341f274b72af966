#!/bin/bash
# Run the GPU test files one pytest process each (a hung kernel then only costs that file), logs to gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
run() {  # name, timeout, pytest args...
  local name=$1; local to=$2; shift 2
  timeout -k 10 "$to" python -m pytest "$@" -m gpu -q -s -p no:cacheprovider > "gpurun_out/$name.log" 2>&1
  echo "$name exit=$?" | tee -a gpurun_out/summary.txt
  tail -n 3 "gpurun_out/$name.log"
}
: > gpurun_out/summary.txt
run bw 600 tests/test_gpu_bandwidth.py
for t in test_linear_bias test_linear_strided test_qkv_rotary test_ffn_layernorm test_similarity test_conv3x3 test_heads test_attention; do
  run dense_$t 300 tests/test_gpu_dense.py -k $t
done
run e2e 900 tests/test_gpu_e2e.py
run nn 300 tests/test_gpu_nn.py
run superglue 600 tests/test_gpu_superglue.py
run extractor_paths 300 tests/test_gpu_extractor_paths.py
run parity_config 600 tests/test_gpu_parity_config.py
run zz_no_alloc 300 tests/test_zz_no_alloc.py
cat gpurun_out/summary.txt
