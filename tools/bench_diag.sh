#!/bin/bash
# Diagnosis runs of bench.py at N GPUs (phase timing needs eager mode). usage: tools/bench_diag.sh N
N=${1:-2}
OUT=gpurun_out/diag_n$N.log
mkdir -p gpurun_out; : > $OUT
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  echo "### $label" >> $OUT
  env "${envs[@]}" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port $((29600 + RANDOM % 300)) bench.py --gpus $N --steps 20 --warmup 6 --no-e2e --no-dp-check "$@" 2>&1 \
      | grep -E '^\{"metric"|phase timing|Error|error' | tail -4 >> $OUT
}
run "eager p2p phases" HETSEQ_PHASE_TIMING=1 HETSEQ_DP_TRANSPORT=p2p -- --no-cuda-graph
run "eager p2p ipc-alloc phases" HETSEQ_PHASE_TIMING=1 HETSEQ_DP_TRANSPORT=p2p HETSEQ_DP_ALLOC=ipc -- --no-cuda-graph
run "eager multimem phases" HETSEQ_PHASE_TIMING=1 -- --no-cuda-graph
run "graph p2p ipc-alloc" HETSEQ_DP_TRANSPORT=p2p HETSEQ_DP_ALLOC=ipc --
run "graph p2p no-overlap of reduce-scatter" HETSEQ_DP_TRANSPORT=p2p HETSEQ_DP_NO_OVERLAP=1 --
run "graph p2p no PDL" HETSEQ_DP_TRANSPORT=p2p HETSEQ_PDL=0 --
python - <<'PY' $OUT
import json, sys
for line in open(sys.argv[1]):
    if line.startswith('###'):
        print(line.strip())
    elif line.startswith('{'):
        d = json.loads(line)
        print('   %8.3f ms/step %9.1f seq/s' % (d['ms_per_step'], d['value']))
    else:
        print('   ' + line.strip()[:200])
PY
