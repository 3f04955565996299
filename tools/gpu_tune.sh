#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; O=gpurun_out/tune.log; : > $O
ILFDT_GBWD_SKIP=2 timeout 300 python tools/tune_gbwd.py 8192 >> $O 2>&1
for c in ${CFGS:-0 12 20 21 22 23}; do ILFDT_DSD_CFG=$c timeout 300 python tools/tune_gbwd.py 8192 >> $O 2>&1; done
cat $O
