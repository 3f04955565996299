#!/bin/bash
# ncu --set full capture of the hot kernels inside one headline-size step (second step of tools/step_once.py).
#   gpurun --timeout 1800 -- 'bash tools/ncu_step.sh'   -> gpurun_out/r02_step_raw.csv (+ the .ncu-rep when it is small enough
#   to travel back: gpurun returns at most 64 MiB)
mkdir -p gpurun_out
python tools/step_once.py --steps 2 > /dev/null 2>&1 || exit 1
ncu --set full --clock-control none \
    -k regex:"sparsemax_fwd_reg_kernel|nce_bwd2_kernel|ground_bwd_dq_v3_kernel|ground_bwd_dsd_v4_kernel|ground2_kernel" \
    --launch-skip 14 --launch-count 14 -f -o /tmp/r02_step python tools/step_once.py --steps 2 > gpurun_out/ncu_step.log 2>&1
ncu -i /tmp/r02_step.ncu-rep --page raw --csv > gpurun_out/r02_step_raw.csv 2>/dev/null
ncu -i /tmp/r02_step.ncu-rep --page details > gpurun_out/r02_step_details.txt 2>/dev/null
sz=$(stat -c %s /tmp/r02_step.ncu-rep)
echo "rep size $sz"
if [ "$sz" -lt 40000000 ]; then cp /tmp/r02_step.ncu-rep gpurun_out/; fi
ls -la gpurun_out/
