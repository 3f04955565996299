#!/bin/bash
# prints headline + trained-like kernel times of one bench run (A/B comparisons inside one gpurun call: box-to-box
# variation is ~5 %). usage: gpu_ab.sh [ENV=val ...]
cd "$(dirname "$0")/.."
for v in "${@:-X=0}"; do
  env $v python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-e2e 2>/dev/null | grep '^{' | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['kernel_ms_per_step']; s=d['secondary']
for kk in (k, s['kernel_ms_per_step']):
    if 'ground_bwd_w16' in kk: kk['ground_bwd']=kk.pop('ground_bwd_w16')
print('$v', 'step', round(d['ms_per_step'],2), 'gbwd', round(k['ground_bwd'],2), 'spx', round(k['sparsemax_fwd'],2), '| trained-like step', round(s['ms_per_step'],2), {a:round(b,2) for a,b in s['kernel_ms_per_step'].items() if a in ('ground_bwd','sparsemax_fwd','gemm_nt')})"
done
