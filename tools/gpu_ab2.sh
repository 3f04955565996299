#!/bin/bash
# A/B of environment knobs inside ONE gpurun call (box-to-box variation is ~5 %): headline step + scatter backward time.
cd "$(dirname "$0")/.."
for v in "$@"; do
  env $v python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-e2e --no-secondary --no-qmap 2>/dev/null | grep '^{' | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['kernel_ms_per_step']
print('$v', 'step', round(d['ms_per_step'],2), 'gbwd', round(k.get('ground_bwd_w16',k.get('ground_bwd',0)),2), 'loss', d['loss'])"
done
