#!/bin/bash
# quick loop: selected GPU tests + one bench line. usage: gpu_quick.sh "<pytest -k expr>" [bench args]
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py -x -q -k "$1" -p no:cacheprovider > $O/t_quick.log 2>&1; echo "exit=$? tests"; tail -3 $O/t_quick.log
shift
timeout 900 python bench.py --steps 3 --warmup 2 --no-cpu-baseline "$@" > $O/bench_quick.log 2>&1; echo "exit=$? bench"
grep -h '^{' $O/bench_quick.log > $O/bench_quick.json; python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/bench_quick.json'))
    print('ms_per_step',d['ms_per_step'],'value',d['value'])
    print({k:round(v,2) for k,v in d['kernel_ms_per_step'].items()})
    if d.get('e2e'): print('e2e',d['e2e']['value'],d['e2e']['ms_per_step'])
except Exception as e: print('no bench json',e)
PY
tail -3 $O/bench_quick.log | cut -c1-300
