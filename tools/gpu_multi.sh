#!/bin/bash
# usage: gpu_multi.sh N   (inside gpurun --gpus N)
cd "$(dirname "$0")/.."
N=${1:-2}; O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611"
[ -z "$SKIP_CHECK" ] && timeout 200 $TR tools/check_multi_gpu.py > $O/multi_check_$N.log 2>&1; echo "exit=$? check"; grep -E "rank|Error" $O/multi_check_$N.log | head -20
timeout 240 $TR bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_$N.log 2>&1; echo "exit=$? bench$N"; grep -h '^{' $O/bench_$N.log | cut -c1-600
[ -z "$SKIP_ONE" ] && timeout 200 python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_1.log 2>&1; echo "exit=$? bench1"; grep -h '^{' $O/bench_1.log | cut -c1-300
