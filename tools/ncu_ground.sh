#!/bin/bash
# one ncu --set full capture of the fused grounding kernel (ground2_kernel<49>, image launch) at global batch ${GB:-4096}
cd "$(dirname "$0")/.."
ARGS="--steps 1 --warmup 1 --global-batch ${GB:-4096} --no-cpu-baseline --no-e2e"
python bench.py $ARGS > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k 'regex:ground2_kernel<.int.49' -c 2 -f -o gpurun_out/${OUT:-prof_ground} python bench.py $ARGS > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
