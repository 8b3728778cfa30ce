#!/bin/bash
# Every kernel launch of one bench step with its device time (ncu, single metric). Usage (GPU box):
#   tools/launch_list.sh [bench args]   -> gpurun_out/launches.csv
set -e
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 2 --no-e2e "$@" > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 1 --warmup 1 --no-e2e "$@" > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/plain.log
