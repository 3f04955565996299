#!/bin/bash
# timing experiments: runs the bench under several env settings and prints nce_bwd / step times
cd "$(dirname "$0")/.."
for v in "$@"; do
  env $v python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-e2e 2>&1 | grep '^{' | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['kernel_ms_per_step']
print('$v', 'step', round(d['ms_per_step'],2), {a:round(b,2) for a,b in k.items() if a in ('nce_bwd_gathered','gemm_nt','sparsemax_fwd','nce_fwd','ground_bwd')})"
done
