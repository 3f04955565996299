#!/bin/bash
cd "$(dirname "$0")/.."
bash tools/gpu_final.sh "$@"
O=gpurun_out
ARGS="--steps 1 --warmup 1 --global-batch 4096 --no-cpu-baseline --no-e2e --no-secondary --no-qmap"
timeout 600 python bench.py $ARGS > $O/plain_small.log 2>&1 || { echo "plain bench failed"; tail -5 $O/plain_small.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches_final.csv python bench.py $ARGS > $O/ncu_launches.log 2>&1
echo "exit=$? ncu launch list"; wc -l $O/launches_final.csv
timeout 600 python tools/time_qmap.py > $O/time_qmap.log 2>&1; echo "exit=$? time_qmap"
