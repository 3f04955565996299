#!/bin/bash
# ncu --set full of the dq scatter kernel at the headline size, default against ILFDT_GBWD_KEEP=1 (evict-last codebook rows)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; O=gpurun_out
ARGS="--steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-secondary --no-qmap"
for tag in default keep; do
  if [ $tag = keep ]; then export ILFDT_GBWD_KEEP=1; fi
  timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
      -k 'regex:ground_bwd_dq_v3' -s 2 -c 1 -f -o $O/prof_dq_$tag python bench.py $ARGS > $O/ncu_dq_$tag.log 2>&1
  echo "exit=$? ncu dq $tag"; tail -1 $O/ncu_dq_$tag.log
done
