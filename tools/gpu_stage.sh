#!/bin/bash
# Staged GPU validation: each group runs in its own process under its own timeout so that a faulting
# tcgen05 kernel cannot take the safe groups down with it. Logs land in gpurun_out/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > $O/gpu.txt 2>&1
run() { # name timeout cmd...
  local name=$1 to=$2; shift 2
  echo "=== $name" | tee -a $O/stage.log
  timeout $to "$@" > $O/$name.log 2>&1
  echo "exit=$? $name" | tee -a $O/stage.log
  tail -n 6 $O/$name.log | tee -a $O/stage.log
}
PY="python -m pytest tests/test_parity_gpu.py -q -x --timeout 300 -p no:cacheprovider"
run t_simt    600 $PY -k "gemm_f32 or sparsemax or l2norm or randn or sum_mean"
run t_umma    300 $PY -k "gemm_nt_tcgen05"
run t_ground  600 $PY -k "test_ground_fwd"
run t_gbwd    600 $PY -k "grounding_fwd_bwd"
run t_nce_f32 600 $PY -k "test_infonce and fp32"
run t_nce_bf  600 $PY -k "test_infonce and bf16"
run t_golden  600 $PY -k "golden and not sparsemax"
run t_full    900 $PY -k "full_step or full_size"
run smoke     600 python __graft_entry__.py smoke
run bench_small 900 python bench.py --steps 2 --warmup 1 --global-batch 4096 --no-cpu-baseline
run bench     1500 python bench.py --steps 3 --warmup 2
grep -h '^{' $O/bench.log > $O/bench.json 2>/dev/null
cat $O/stage.log | grep exit=
