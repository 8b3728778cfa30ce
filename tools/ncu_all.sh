#!/bin/bash
# `ncu --set full` captures of every hand-written kernel family of a BERT-large update (1 GPU), the way the
# per-kernel summaries in profiles/ncu_kernels.md are produced. Run on the GPU box:
#     tools/ncu_all.sh       -> gpurun_out/ncu_{fwd,bwd,misc,gemm}.ncu-rep and gpurun_out/ncu_kernels.md
# The plain command runs first and must exit 0 (never profile a command that has not passed unprofiled);
# times measured under ncu are never reported as bench values.
set -e
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cuda-graph"
$CMD > gpurun_out/ncu_plain.log 2>&1
cap() {   # tag, kernel regex, launches to skip, launches to capture
  ncu --set full --clock-control none --import-source on -f -o "gpurun_out/ncu_$1" \
      --kernel-name "regex:$2" --launch-skip "$3" --launch-count "$4" $CMD > "gpurun_out/ncu_$1.log" 2>&1 \
      || echo "ncu failed for $1 (see gpurun_out/ncu_$1.log)"
}
# attention + LayerNorm launch 148 times per update (24 layers x (1+2) forward, x (2+1) backward, + embeddings/heads):
# a window in the forward part and one in the backward part of the 4th update
cap fwd  'attn_|ln_fwd|ln_bwd' 460 6
cap bwd  'attn_|ln_fwd|ln_bwd' 560 6
# everything else that is not a GEMM: ~15 launches per update
cap misc 'xent_|embed_|adam_flat|sumsq_|gather_rows|scatter_add_rows|colsum|sum_over_batch' 45 18
cap gemm 'gemm_tcgen05' 1200 14
python tools/ncu_summary.py gpurun_out/ncu_fwd.ncu-rep gpurun_out/ncu_bwd.ncu-rep gpurun_out/ncu_misc.ncu-rep \
    gpurun_out/ncu_gemm.ncu-rep > gpurun_out/ncu_kernels.md || true
# gpurun brings back at most 64 MiB: the summary is what gets committed; keep only the two small reports
ls -la gpurun_out/ncu_*
rm -f gpurun_out/ncu_gemm.ncu-rep gpurun_out/ncu_misc.ncu-rep
