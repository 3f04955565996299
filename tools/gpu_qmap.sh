#!/bin/bash
# q_map producer (N2): its parity tests, then a timing of the producer at the headline batch.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; O=gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py -x -q -k "qmap or q_map or gemm_tn or linear_bf16 or gemm_nt" --timeout 300 -p no:cacheprovider > $O/t_qmap.log 2>&1; echo "exit=$? tests"; tail -15 $O/t_qmap.log
timeout 600 python tools/time_qmap.py > $O/time_qmap.log 2>&1; echo "exit=$? time_qmap"; tail -20 $O/time_qmap.log
