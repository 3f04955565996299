#!/bin/bash
# round-end style validation in one call: the whole -m gpu suite, smoke(), the reference arm and the bench line.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; O=gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total,power.limit --format=csv > $O/gpu.txt 2>&1
timeout 900 python -m pytest tests -x -q -m gpu --timeout 300 -p no:cacheprovider > $O/t_gpu.log 2>&1; echo "exit=$? pytest -m gpu"; tail -4 $O/t_gpu.log
timeout 300 python __graft_entry__.py smoke > $O/smoke.log 2>&1; echo "exit=$? smoke"; tail -3 $O/smoke.log
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.log 2>&1; echo "exit=$? bench ref"
grep -h '^{' $O/bench_ref.log > $O/bench_ref.json
timeout 900 python bench.py "$@" > $O/bench.log 2>&1; echo "exit=$? bench"
grep -h '^{' $O/bench.log > $O/bench.json
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/bench.json'))
    print('ms_per_step',d['ms_per_step'],'value',d['value'],'frac',d['roofline']['frac'])
    print({k:round(v,2) for k,v in d['kernel_ms_per_step'].items()})
    if d.get('e2e'): print('e2e',d['e2e']['value'],d['e2e']['ms_per_step'])
except Exception as e: print('no bench json',e)
PY
