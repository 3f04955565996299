#!/bin/bash
# One gpurun call: GPU test suite, smoke, and one bench line per BASELINE.json config at N=1.
# usage: gpurun --timeout 2400 -- 'bash tools/gpu_suite.sh [tests|bench|all]'
what=${1:-all}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/gpu.txt 2>&1
if [ "$what" = tests ] || [ "$what" = all ]; then
  timeout 1500 python -m pytest tests -m gpu -q --maxfail=25 -s > gpurun_out/pytest_gpu.log 2>&1
  echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
  tail -5 gpurun_out/pytest_gpu.log
  timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke.log
fi
if [ "$what" = bench ] || [ "$what" = all ]; then
  timeout 600 python bench.py > gpurun_out/bench_headline.json 2> gpurun_out/bench_headline.err; echo "headline exit $?"
  for cfg in 1 2 3 3b 4 5; do
    timeout 400 python bench.py --config $cfg --steps 20 --warmup 5 > gpurun_out/bench_cfg$cfg.json 2> gpurun_out/bench_cfg$cfg.err; echo "cfg $cfg exit $?"
  done
  timeout 600 python bench.py --mode fp32 --steps 3 --warmup 3 --no-secondary --no-qmap --no-cpu-baseline > gpurun_out/bench_headline_fp32.json 2> gpurun_out/bench_headline_fp32.err; echo "fp32 exit $?"
  timeout 300 python bench.py --impl reference > gpurun_out/bench_ref.json 2>/dev/null; echo "ref exit $?"
  for f in gpurun_out/bench_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d.get("roofline") or {}
    e=d.get("e2e") or {}
    print(sys.argv[1], d.get("dtype"), "value %.4g" % d["value"], "ms %.4g" % d["ms_per_step"], "roof %.3f" % (r.get("frac") or 0), "e2e %.4g" % (e.get("value") or 0), d.get("cuda_graph","")[:12])
except Exception as ex:
    print(sys.argv[1], "unreadable", ex)
PY
  done
fi
