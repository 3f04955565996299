"""Times ilfdt_ground_bwd (dq scatter + dsd scatter) on realistic supports: ILFDT_DSD_CFG / ILFDT_GBWD_SKIP select
the kernel variant (read once per process). usage: python tools/tune_gbwd.py [b]"""
import os, sys, math, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "iterated-learning-for-vlm_b200"))
from ilfdt import ops

b = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
K, D = 16384, 512
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(0)
sd = torch.randn(K, D, device=dev, generator=g)
sd_b = ops.cast_bf16(sd)
out = []
for n_tok in (49, 77):
    q = torch.randn(b, n_tok, D, device=dev, generator=g)
    q_b = ops.cast_bf16(q)
    pooled, arg = ops.ground_fwd(q_b, sd_b, None, 1000.0, "max")
    w, _, _, nnz = ops.sparsemax_fwd(pooled)
    g0 = torch.randn(b, K, device=dev, generator=g)
    dsd = torch.zeros(K, D, device=dev)
    ts = []
    for it in range(4):
        gg = g0.clone()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.ground_bwd(gg, w, arg, q_b, sd_b, None, 1000.0, dsd)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    dsd.zero_()
    gg = g0.clone()
    dq = ops.ground_bwd(gg, w, arg, q_b, sd_b, None, 1000.0, dsd)
    out.append("n_tok=%d nnz/row=%.0f ms=%.3f chk=%.6e/%.6e" % (n_tok, nnz.float().mean().item(), min(ts[1:]),
                                                             dsd.double().abs().sum().item(), dq.double().abs().sum().item()))
print("cfg=%s skip=%s b=%d  %s" % (os.environ.get("ILFDT_DSD_CFG", "0"), os.environ.get("ILFDT_GBWD_SKIP", "0"), b, "  ".join(out)), flush=True)
