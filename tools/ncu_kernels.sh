#!/bin/bash
# ncu --set full of the backward kernels at global batch ${GB:-4096}: usage ncu_kernels.sh <regex> <out> [skip] [count]
cd "$(dirname "$0")/.."
ARGS="--steps 1 --warmup 1 --global-batch ${GB:-4096} --no-cpu-baseline --no-e2e"
python bench.py $ARGS > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k "regex:$1" -s ${3:-0} -c ${4:-4} -f -o gpurun_out/$2 python bench.py $ARGS > gpurun_out/ncu_$2.log 2>&1
tail -2 gpurun_out/ncu_$2.log
