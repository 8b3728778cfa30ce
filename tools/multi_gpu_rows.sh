#!/bin/bash
# The BASELINE.json configurations beyond the headline one, at N GPUs: BERT-base seq 128, BERT-large seq 512, and the
# flattened-gradient all-reduce + Adam sweep. usage: tools/multi_gpu_rows.sh N   -> gpurun_out/rows_n$N.log, sweep_n$N.json
N=${1:-8}
OUT=gpurun_out/rows_n$N.log
mkdir -p gpurun_out; : > $OUT
tr() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) "$@"; }
if [ -z "$SWEEP_ONLY" ]; then
  echo "### bert-base seq128 batch 32/GPU" >> $OUT
  tr bench.py --gpus $N --steps 20 --warmup 6 --no-e2e --no-dp-check --model bert-base 2>&1 | grep '^{"metric"' >> $OUT
  echo "### bert-large seq512 batch 8/GPU" >> $OUT
  tr bench.py --gpus $N --steps 20 --warmup 6 --no-e2e --no-dp-check --seq-len 512 --batch 8 2>&1 | grep '^{"metric"' >> $OUT
fi
tr tools/allreduce_adam_sweep.py --sizes-mb ${SIZES:-1,16,256,1345} --iters 5 --out gpurun_out/sweep_n$N.json 2>&1 | grep '^{' >> $OUT
python - <<'PY' $OUT
import json, sys
for line in open(sys.argv[1]):
    if line.startswith('###'):
        print(line.strip())
    elif line.startswith('{"metric'):
        d = json.loads(line); print('   %.3f ms/step  %.1f seq/s' % (d['ms_per_step'], d['value']))
    elif line.startswith('{'):
        d = json.loads(line)
        print('   %7.1f MB: ' % d['grad_mb'] + '  '.join('%s=%s' % (k, d[k]) for k in sorted(d) if k.endswith('_ms') or k.endswith('_parts_ms')))
PY
