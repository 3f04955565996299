"""Diagnostic (GPU): which samples / rows of dq disagree with the fp64 oracle, and does the disagreement start upstream?"""
import math, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "iterated-learning-for-vlm_b200"))
from ilfdt import ops, functional as Fn
from oracle import fdt_oracle as O
dev = torch.device("cuda:0")

def per_sample(a, ref):
    a = a.detach().double(); ref = ref.detach().double()
    return ((a - ref).abs().flatten(1).amax(1) / ref.abs().max()).cpu()

def grounding_case(b, P, K, mode, seed, chunk=64):
    D, temp = 512, 1000.0
    g_ = torch.Generator(device="cuda").manual_seed(seed)
    q = torch.randn(b, P, D, device=dev, generator=g_)
    sd = torch.randn(K, D, device=dev, generator=g_)
    if mode == "bf16":
        q, sd = q.to(torch.bfloat16).float(), sd.to(torch.bfloat16).float()
    up = torch.randn(b, D, device=dev, generator=g_)
    qd = q.clone().requires_grad_(True); sdd = sd.clone().requires_grad_(True)
    w, f = Fn.GroundingFunction.apply(qd, sdd, None, temp, mode, True)
    (f * up).sum().backward()
    errs = []; werr = []; ferr = []
    for b0 in range(0, b, chunk):
        qo = q[b0:b0+chunk].double().requires_grad_(True); so = sd.double().requires_grad_(True)
        w_ref, f_ref, pooled = O.grounding(qo, so, None, temp)
        (f_ref * up[b0:b0+chunk].double()).sum().backward()
        errs.append(per_sample(qd.grad[b0:b0+chunk], qo.grad))
        werr.append(per_sample(w[b0:b0+chunk].float(), w_ref)); ferr.append(per_sample(f[b0:b0+chunk], f_ref))
    e = torch.cat(errs); we = torch.cat(werr); fe = torch.cat(ferr)
    bad = (e > 1e-3).nonzero().flatten().tolist()
    print("grounding b=%d P=%d K=%d %s: dq per-sample err max %.2e median %.2e; w max %.2e f max %.2e; %d samples > 1e-3: %s" % (
        b, P, K, mode, e.max(), e.median(), we.max(), fe.max(), len(bad), bad[:20]), flush=True)
    if bad:
        s0 = bad[0]
        qo = q[s0:s0+1].double().requires_grad_(True); so = sd.double().requires_grad_(True)
        w_ref, f_ref, pooled = O.grounding(qo, so, None, temp)
        (f_ref * up[s0:s0+1].double()).sum().backward()
        rowerr = (qd.grad[s0].double() - qo.grad[0]).abs().amax(1)
        rel = rowerr / qo.grad.abs().max()
        print("  sample %d: tokens with error > 1e-4:" % s0, [(int(t), round(float(rel[t]), 4)) for t in (rel > 1e-4).nonzero().flatten().tolist()[:10]])
        s = (qo.detach()[0] @ so.detach().t())           # [P,K]
        top, idx = s.topk(2, 0)
        nz = w_ref[0] > 0
        gap = (top[0] - top[1])
        gg = (up[s0].double() @ so.detach().t())
        mean = gg[nz].mean()
        coef = (gg - mean) * nz
        for t in (rel > 1e-4).nonzero().flatten().tolist()[:4]:
            ks = ((idx[0] == t) | (idx[1] == t)) & nz
            kk = ks.nonzero().flatten()
            small = kk[gap[kk].argsort()[:3]]
            print("   token %d: smallest top-2 gaps among its support codes: %s (max|S| %.1f) coef %s w %s tau-margin %s" % (
                t, [float("%.3g" % x) for x in gap[small].tolist()], s.abs().max(), [float("%.3g" % x) for x in coef[small].tolist()],
                [float("%.3g" % x) for x in w_ref[0][small].tolist()], ""))
        # support comparison for this sample
        mism = ((w[s0] > 0) != (w_ref[0] > 0)).nonzero().flatten().tolist()
        print("   support mismatches in this sample:", mism[:10], "weights there (ours, ref):",
              [(float("%.3g" % w[s0][k]), float("%.3g" % w_ref[0][k])) for k in mism[:5]],
              "their coef:", [float("%.3g" % ((gg[k] - mean))) for k in mism[:5]], "typical |coef| %.3g" % coef[nz].abs().mean())

grounding_case(256, 49, 4096, "fp32", 11)
grounding_case(256, 49, 16384, "fp32", 12)
grounding_case(256, 49, 16384, "bf16", 13)
grounding_case(256, 196, 16384, "bf16", 14)
grounding_case(1024, 196, 16384, "bf16", 196)
