#!/bin/bash
# builds tools/l2_probe.cu for sm_100a and runs it; output (one JSON line per configuration) -> gpurun_out/l2_probe.jsonl
set -e
mkdir -p gpurun_out
nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o gpurun_out/l2_probe tools/l2_probe.cu
./gpurun_out/l2_probe | tee gpurun_out/l2_probe.jsonl
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm --format=csv,noheader >> gpurun_out/l2_probe.jsonl
