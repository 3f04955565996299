"""python tools/time_sparsemax.py [b]  -- CUDA-event timing of ilfdt_sparsemax_fwd on pooled scores of the headline statistics
(random init: support ~3000 of 16384) and of the trained-like case (q x 200: support of tens), plus the value check against
an fp64 sort-based evaluation of prototype/model/sparsemax.py:40-63 (support identical where the margin is not rounding-level)."""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "iterated-learning-for-vlm_b200"))
from ilfdt import ops
dev = torch.device("cuda:0")
b = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
K, D, temp = 16384, 512, 1000.0
g_ = torch.Generator(device="cuda").manual_seed(0)
sd_k = ops.cast_bf16(torch.randn(K, D, device=dev, generator=g_))
out = {}
for name, scale in (("random_init", 1.0), ("trained_like", 200.0)):
    q = torch.randn(b, 49, D, device=dev, generator=g_) * scale
    pooled, _ = ops.ground_fwd(ops.cast_bf16_rows(q, None), sd_k, None, temp, "max")
    del q
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for it in range(6):
        e0.record()
        w, w_b, tau, nnz = ops.sparsemax_fwd(pooled, want_bf16=True, want_f32=(it == 0))
        e1.record()
        torch.cuda.synchronize()
        if it >= 2:
            ts.append(e0.elapsed_time(e1))
        if it == 0:
            w0 = w
    # fp64 reference on a slice
    n = min(b, 512)
    z = pooled[:n].double()
    z = z - z.max(1, keepdim=True)[0]
    zs, _ = z.sort(1, descending=True)
    cs = zs.cumsum(1) - 1.0
    rng = torch.arange(1, K + 1, device=dev, dtype=torch.float64)
    supp = (rng * zs > cs).sum(1, keepdim=True)
    tau_ref = cs.gather(1, supp - 1) / supp.double()
    w_ref = (z - tau_ref).clamp_min(0)
    mis = ((w0[:n] > 0) != (w_ref > 0))
    margin = (z - tau_ref).abs()[mis].max().item() if mis.any() else 0.0
    out[name] = dict(ms=round(sum(ts) / len(ts), 4), gb_per_s=round((b * K * 6) / (sum(ts) / len(ts)) / 1e6, 1),
                     mean_support=float(nnz.float().mean()), max_abs_err=float((w0[:n].double() - w_ref).abs().max()),
                     support_mismatches=int(mis.sum()), worst_mismatch_margin=margin)
    del pooled, w0, w, w_b
print(json.dumps({"b": b, "sparsemax_fwd": out}, indent=1))
