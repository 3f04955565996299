#!/bin/bash
# ncu --set full of the dsd scatter kernel at the headline size, with and without the ranged sweep (ILFDT_DSD_SPLIT=0),
# plus the launch list of the bench at the headline size. Each only after the plain command exited 0.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; O=gpurun_out
bash tools/gpu_launchlist_full.sh
ARGS="--steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-secondary --no-qmap"
for tag in ranged whole; do
  if [ $tag = whole ]; then export ILFDT_DSD_SPLIT=0 ILFDT_DSD_STREAM=0; fi
  timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
      -k 'regex:ground_bwd_dsd_v3' -s 2 -c 1 -f -o $O/prof_dsd_$tag python bench.py $ARGS > $O/ncu_dsd_$tag.log 2>&1
  echo "exit=$? ncu dsd $tag"; tail -1 $O/ncu_dsd_$tag.log
done
ls -la $O/prof_dsd_*.ncu-rep
