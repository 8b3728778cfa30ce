#!/bin/bash
# One `ncu --set full` capture of one kernel, the way profiles/*.md were produced (1 GPU, after the same command
# exited 0 without ncu; never report a time measured under ncu).
#   tools/ncu_kernel.sh gemm <case>              case: qkv|ffn1|ffn2|wgrad|dgelu|square  (stand-alone harness)
#   tools/ncu_kernel.sh step <kernel-name-regex>   e.g. attn_bwd_kernel, ln_bwd_kernel (one bench step)
# Writes gpurun_out/<name>.ncu-rep; read it in the dev container with
#   ncu -i gpurun_out/<name>.ncu-rep --page raw --csv      /     --page source --csv
set -e
mkdir -p gpurun_out
MODE=${1:-gemm}
ARG=${2:-qkv}
if [ "$MODE" = gemm ]; then
  build/gemm_test --case "$ARG" > gpurun_out/plain_$ARG.log
  ncu --set full --clock-control none --import-source on -f -o gpurun_out/gemm_$ARG -c 14 \
      build/gemm_test --case "$ARG" > gpurun_out/ncu_$ARG.log 2>&1
  tail -2 gpurun_out/ncu_$ARG.log
else
  python bench.py --steps 2 --warmup 3 --no-e2e --no-cuda-graph > gpurun_out/plain_step.log 2>&1
  ncu --set full --clock-control none --import-source on -f -o "gpurun_out/step_$ARG" \
      --kernel-name "regex:$ARG" --launch-skip 30 --launch-count 2 \
      python bench.py --steps 1 --warmup 3 --no-e2e --no-cuda-graph > "gpurun_out/ncu_step_$ARG.log" 2>&1
  tail -2 "gpurun_out/ncu_step_$ARG.log"
fi
