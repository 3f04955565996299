#!/bin/bash
# usage: gpurun --gpus N --timeout 1800 -- 'bash tools/gpu_multi.sh N "headline 3 3b 4 5"'
# One bench line per named config at N ranks (torchrun), the copy-only H2D probe, and (N >= 2) the W-rank oracle check.
N=$1; CFGS=${2:-headline}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
port=29600
for cfg in $CFGS; do
  port=$((port+1))
  timeout 500 $TR --master-port $port bench.py --gpus $N --config $cfg --steps 10 --warmup 3 > gpurun_out/bench_${cfg}_n$N.json 2> gpurun_out/bench_${cfg}_n$N.err
  echo "cfg $cfg N=$N exit $?"
  python - gpurun_out/bench_${cfg}_n$N.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    e=d.get("e2e") or {}
    print("   value %.5g pairs/s  %.4g ms/step  step_frac %.3f  e2e %.5g (fp32 host %s)  h2d probe %.1f GB/s/GPU  graph %s  kernel_sum %.3f ms" % (
        d["value"], d["ms_per_step"], d["step_roofline"]["frac_of_sustained_peak"], e.get("value",0),
        round(e["fp32_host"]["value"]) if "fp32_host" in e else None, e.get("h2d_copy_only",{}).get("gb_per_s_per_gpu",0), d["cuda_graph"][:3], sum(d["kernel_ms_per_step"].values())))
except Exception as ex:
    print("   unreadable:", ex)
PY
done
for dt in "" "--fp32"; do port=$((port+1)); timeout 200 $TR --master-port $port tools/h2d_probe.py $dt >> gpurun_out/h2d_probe_n$N.jsonl 2>/dev/null; done
cat gpurun_out/h2d_probe_n$N.jsonl
if [ "$N" -ge 2 ]; then
  port=$((port+1))
  timeout 400 $TR --master-port $port tools/check_multi_gpu.py > gpurun_out/multi_check_$N.log 2>&1; echo "check exit $?"
  grep -E "^rank 0|destroy|MISMATCH|Error" gpurun_out/multi_check_$N.log | head -12
fi
