"""Diagnostic (GPU): where do our argmax tokens differ from the fp64 oracle's, and how close were the two best tokens?"""
import math, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "iterated-learning-for-vlm_b200"))
from ilfdt import ops
from oracle import fdt_oracle as O
dev = torch.device("cuda:0")

def run(tag, b, P, K, mode, seed):
    D = 512
    g = torch.Generator(device="cuda").manual_seed(seed)
    q = torch.randn(b, P, D, device=dev, generator=g)
    sd = torch.randn(K, D, device=dev, generator=g)
    if mode == "bf16":
        q, sd = q.to(torch.bfloat16).float(), sd.to(torch.bfloat16).float()
        pooled, arg = ops.ground_fwd(ops.cast_bf16(q), ops.cast_bf16(sd), None, 1000.0, "max")
    else:
        pooled, arg = ops.ground_fwd(q, sd, None, 1000.0, "max")
    w = ops.sparsemax_fwd(pooled)[0]
    n_mis = 0; gaps = []; n_sup = 0
    for b0 in range(0, b, 32):
        s = torch.matmul(q[b0:b0+32].double(), sd.double().t())
        top, idx = s.topk(2, dim=1)
        mis = (idx[:, 0] != arg[b0:b0+32].long())
        sup = w[b0:b0+32] > 0
        n_sup += int(sup.sum())
        m2 = mis & sup
        n_mis += int(m2.sum())
        if m2.any():
            gap = (top[:, 0] - top[:, 1])[m2]
            second = (idx[:, 1] == arg[b0:b0+32].long())[m2]
            gaps += [(float(x), bool(y)) for x, y in zip(gap.tolist(), second.tolist())]
        smax = float(s.abs().max())
    gaps.sort(reverse=True)
    print(tag, "b=%d P=%d K=%d %s: %d argmax mismatches inside the support of %d entries; max|S|=%.1f; largest gaps (gap, ours==2nd best): %s" % (
        b, P, K, mode, n_mis, n_sup, smax, gaps[:8]), flush=True)

run("img", 256, 49, 4096, "fp32", 1)
run("img", 256, 49, 16384, "fp32", 2)
run("img", 256, 49, 16384, "bf16", 3)
run("vitb16", 256, 196, 16384, "bf16", 4)
run("vitb16", 256, 196, 16384, "fp32", 5)
run("vitb16", 1024, 196, 16384, "bf16", 6)
