#!/bin/bash
# ncu --set full of one nce_bwd_kernel launch at global batch ${GB:-32768}. usage: ncu_nce.sh <out>
cd "$(dirname "$0")/.."
ARGS="--steps 1 --warmup 1 --global-batch ${GB:-32768} --no-cpu-baseline --no-e2e"
python bench.py $ARGS > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k 'regex:nce_bwd' -c 1 -f -o gpurun_out/${1:-prof_nce} python bench.py $ARGS > gpurun_out/ncu_nce.log 2>&1
tail -2 gpurun_out/ncu_nce.log
