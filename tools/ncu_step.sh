#!/bin/bash
# one ncu --set full capture of the two tensor-core kernels of the step at global batch ${GB:-32768}:
# ground2_kernel<49> (image grounding) and nce_bwd_kernel (InfoNCE backward). usage: ncu_step.sh <out> [count]
cd "$(dirname "$0")/.."
ARGS="--steps 1 --warmup 1 --global-batch ${GB:-32768} --no-cpu-baseline --no-e2e"
python bench.py $ARGS > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k 'regex:ground2_kernel<.int.49|nce_bwd_kernel' -c ${2:-4} -f -o gpurun_out/${1:-prof_step} python bench.py $ARGS > gpurun_out/ncu_step.log 2>&1
tail -2 gpurun_out/ncu_step.log
