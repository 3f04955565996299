#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_parity_gpu.py -x -q -m gpu -k "without_fp32_weights or grounding" --timeout 200 -p no:cacheprovider 2>&1 | tail -5
bash tools/gpu_ab.sh ILFDT_LEAN_WEIGHTS=0 ILFDT_LEAN_WEIGHTS=1 ILFDT_LEAN_WEIGHTS=0 ILFDT_LEAN_WEIGHTS=1
