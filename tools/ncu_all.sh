#!/bin/bash
# One `ncu --set full` capture per hand-written kernel family of a BERT-large update (1 GPU), the way the
# per-kernel summaries in profiles/ncu_kernels.md are produced. Run on the GPU box:
#     tools/ncu_all.sh            -> gpurun_out/ncu_<family>.ncu-rep
# then, in the dev container:  python tools/ncu_summary.py gpurun_out/ncu_*.ncu-rep > profiles/ncu_kernels.md
# The plain command runs first and must exit 0 (never profile a command that has not passed unprofiled);
# times measured under ncu are never reported as bench values.
set -e
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cuda-graph"
$CMD > gpurun_out/ncu_plain.log 2>&1
# kernel-name regex : launches of that kernel to skip (warm-up steps) : tag
FAMILIES=(
  "attn_fwd_kernel:24:attn_fwd"
  "attn_bwd_kernel:24:attn_bwd"
  "ln_fwd_kernel:60:ln_fwd"
  "ln_bwd_kernel:60:ln_bwd"
  "xent_fwd_bwd_kernel:4:xent"
  "embed_gather_sum_kernel:2:embed_gather"
  "embed_scatter_sorted_kernel:2:embed_scatter"
  "adam_flat_kernel:2:adam"
  "sumsq_partial_kernel:2:sumsq"
  "gemm_tcgen05_2cta_kernel:300:gemm_2cta"
)
for f in "${FAMILIES[@]}"; do
  IFS=: read -r regex skip tag <<< "$f"
  ncu --set full --clock-control none --import-source on -f -o "gpurun_out/ncu_$tag" \
      --kernel-name "regex:$regex" --launch-skip "$skip" --launch-count 2 \
      $CMD > "gpurun_out/ncu_$tag.log" 2>&1 || echo "ncu failed for $tag (see gpurun_out/ncu_$tag.log)"
done
ls -la gpurun_out/ncu_*.ncu-rep
