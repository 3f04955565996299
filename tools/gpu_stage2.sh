#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
O=gpurun_out
run() { local name=$1 to=$2; shift 2; echo "=== $name" | tee -a $O/stage.log; timeout $to "$@" > $O/$name.log 2>&1; echo "exit=$? $name" | tee -a $O/stage.log; tail -n 4 $O/$name.log | cut -c1-400 | tee -a $O/stage.log; }
rm -f $O/stage.log
run t_all 1500 python -m pytest tests -x -q -m gpu --timeout 600 -p no:cacheprovider
run smoke 600 python __graft_entry__.py smoke
run bench_small 900 python bench.py --steps 2 --warmup 1 --global-batch 4096 --no-cpu-baseline
run bench 1500 python bench.py --steps 3 --warmup 2
run bench_ref 600 python bench.py --impl reference --steps 2 --warmup 1
grep -h '^{' $O/bench.log > $O/bench.json 2>/dev/null
if grep -q '^{' $O/bench_small.log; then
  echo "=== ncu launches" | tee -a $O/stage.log
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_small.csv python bench.py --steps 1 --warmup 1 --global-batch 4096 --no-cpu-baseline --no-e2e > $O/ncu_small.log 2>&1
  echo "exit=$? ncu_launches" | tee -a $O/stage.log
fi
grep exit= $O/stage.log
