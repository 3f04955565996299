#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; O=gpurun_out
python bench.py --steps 5 --warmup 3 > $O/bench.log 2>&1; echo "exit=$? bench"; grep -h '^{' $O/bench.log > $O/bench.json
GB=32768 OUT=prof_ground_v2_b32768 bash tools/ncu_ground.sh
