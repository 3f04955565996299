#!/bin/bash
# compute-sanitizer pass over the small-shape unit tests of every kernel family.
#   gpurun --timeout 3000 -- 'bash tools/sanitize.sh'        logs -> gpurun_out/sanitize_<tool>.log (+ a summary)
# memcheck runs the whole small-shape parity file; racecheck / synccheck / initcheck run the subset that reaches every
# kernel family, above all the CTA-pair kernels whose remote mbarrier arrives are relaxed (csrc/umma_pair.cuh).
mkdir -p gpurun_out
SEL='test_ground_fwd or test_grounding_fwd_bwd or test_ground_fwd_text_buckets or test_infonce or test_gemm_nt_tcgen05 or test_gemm_tn_tcgen05 or test_sparsemax_golden or test_fused_topk or test_linear_bf16 or test_qmap_layernorm_rows or test_l2norm or test_randn_fill'
run() {  # tool, time limit, pytest selection
  local tool=$1 limit=$2 sel=$3
  local log=gpurun_out/sanitize_$tool.log
  timeout $limit compute-sanitizer --tool $tool --error-exitcode 86 --print-limit 20 \
      python -m pytest tests/test_parity_gpu.py -m gpu -q -x -p no:cacheprovider -k "$sel" > $log 2>&1
  echo "$tool exit $? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY|passed|failed' $log | tr '\n' ';' | cut -c1-300)"
}
run memcheck 1500 "$SEL"
run racecheck 1500 "$SEL"
run synccheck 1200 "$SEL"
grep -hE "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|passed|failed" gpurun_out/sanitize_*.log | sort | uniq -c | sort -rn | head -40 > gpurun_out/sanitize_summary.txt
cat gpurun_out/sanitize_summary.txt
