#!/bin/bash
# bench.py at N GPUs under the gradient-path switches (environment variables of parallel/fused_dp.py), one JSON
# line each -> gpurun_out/variants_n$N.log.   usage: tools/bench_variants.sh N [extra bench args]
N=${1:-2}; shift
OUT=gpurun_out/variants_n$N.log
mkdir -p gpurun_out; : > $OUT
run() {   # label, env assignments..., then "--" and extra args
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  echo "### $label" >> $OUT
  env "${envs[@]}" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port $((29600 + RANDOM % 300)) bench.py --gpus $N --steps 20 --warmup 6 --no-e2e --no-dp-check "$@" 2>&1 \
      | grep -E '^\{"metric"|Error|error' | tail -2 >> $OUT
}
# VARIANTS picks a subset (default: all)
V=${VARIANTS:-"default p2p multimem fusedupd norotate rs74 rs296 nccl"}
for v in $V; do
  case $v in
    default)  run "default (split update, gather overlapped, transport auto)" X=1 -- "$@" ;;
    p2p)      run "peer loads/stores" HETSEQ_DP_TRANSPORT=p2p -- "$@" ;;
    multimem) run "multimem (NVLS) loads/stores" HETSEQ_DP_TRANSPORT=multimem -- "$@" ;;
    fusedupd) run "fused update+gather kernel rotated into the forward" HETSEQ_DP_SPLIT=0 -- "$@" ;;
    norotate) run "gather not overlapped (applied at the end of the step)" HETSEQ_DP_ROTATE=0 -- "$@" ;;
    rs74)     run "reduce-scatter with 74 CTAs" HETSEQ_RS_BLOCKS=74 -- "$@" ;;
    rs37)     run "reduce-scatter with 37 CTAs" HETSEQ_RS_BLOCKS=37 -- "$@" ;;
    rs74g37)  run "reduce-scatter with 74 CTAs, gather with 37" HETSEQ_RS_BLOCKS=74 HETSEQ_GATHER_BLOCKS=37 -- "$@" ;;
    rs74g74)  run "reduce-scatter with 74 CTAs, gather with 74" HETSEQ_RS_BLOCKS=74 HETSEQ_GATHER_BLOCKS=74 -- "$@" ;;
    env:*)    IFS=, read -ra kv <<< "${v#env:}"; run "$v" "${kv[@]}" -- "$@" ;;
    rs296)    run "reduce-scatter with 296 CTAs" HETSEQ_RS_BLOCKS=296 -- "$@" ;;
    nccl)     run "NCCL all-reduce backend (eager)" X=1 -- --dp-backend nccl "$@" ;;
  esac
done
python - <<'PY' $OUT
import json, sys
for line in open(sys.argv[1]):
    if line.startswith('###'):
        label = line[4:].strip()
    elif line.startswith('{'):
        d = json.loads(line)
        print('%-62s %8.3f ms/step %9.1f seq/s  [%s]' % (label, d['ms_per_step'], d['value'], d['config']['dp_backend']))
    else:
        print(label, 'FAILED:', line.strip()[:160])
PY
