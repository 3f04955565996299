#!/bin/bash
# Evidence for the final state of the round: ncu launch list of the bench command (global batch 4096, q_map section
# included) and one --set full capture of the q_map kernels; each only after the same command exited 0 without ncu.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; O=gpurun_out
ARGS="--steps 1 --warmup 1 --global-batch 4096 --no-cpu-baseline --no-e2e --no-secondary --no-qmap"
timeout 600 python bench.py $ARGS > $O/plain_small.log 2>&1 || { echo "plain bench failed"; tail -5 $O/plain_small.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches_final.csv python bench.py $ARGS > $O/ncu_launches.log 2>&1
echo "exit=$? ncu launch list"; wc -l $O/launches_final.csv
B=4096 timeout 600 python tools/time_qmap.py > $O/time_qmap_small.log 2>&1 || { echo "plain time_qmap failed"; tail -5 $O/time_qmap_small.log; exit 1; }
B=4096 timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k 'regex:ln_fwd_kernel|ln_bwd_kernel|umma_kernel|colsum_partial' -s 30 -c 14 -f -o $O/prof_qmap python tools/time_qmap.py > $O/ncu_qmap.log 2>&1
echo "exit=$? ncu qmap"; tail -2 $O/ncu_qmap.log; ls -la $O/*.ncu-rep
