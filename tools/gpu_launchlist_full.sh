#!/bin/bash
# ncu launch list (gpu__time_duration only, no replay) of the bench at the HEADLINE size, so that each kernel's share
# of the step can be set beside the CUDA-event shares of the bench line. Runs only after the same command exits 0.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; O=gpurun_out
ARGS="--steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-secondary --no-qmap"
timeout 600 python bench.py $ARGS > $O/plain_full.log 2>&1 || { echo "plain bench failed"; tail -5 $O/plain_full.log; exit 1; }
grep -h '^{' $O/plain_full.log > $O/plain_full.json
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches_full.csv python bench.py $ARGS > $O/ncu_launches_full.log 2>&1
echo "exit=$? ncu launch list (full size)"; wc -l $O/launches_full.csv
