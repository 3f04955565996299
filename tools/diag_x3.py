"""Feasibility of the 3 x bf16 fp32-grade score contraction on tcgen05: accuracy vs fp64 and time vs the SIMT fp32 kernel."""
import math, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "iterated-learning-for-vlm_b200"))
from ilfdt import ops
dev = torch.device("cuda:0")
def t(fn, n=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
for b, P, K in ((256, 49, 16384), (2048, 49, 16384), (1024, 77, 16384)):
    D = 512
    g = torch.Generator(device="cuda").manual_seed(b)
    q = torch.randn(b, P, D, device=dev, generator=g); sd = torch.randn(K, D, device=dev, generator=g)
    def x3():
        return ops.ground_fwd(ops.split3_bf16(q, 1), ops.split3_bf16(sd, 0), None, 1000.0, "max", sqrt_d=math.sqrt(D))
    p3, a3 = x3()
    p32, a32 = ops.ground_fwd(q, sd, None, 1000.0, "max")
    n = min(b, 128)
    s64 = (q[:n].double() @ sd.double().t()) / math.sqrt(D) / 1000.0
    ref, aref = s64.max(1)
    e3 = ((p3[:n].double() - ref).abs().max() / ref.abs().max()).item()
    e32 = ((p32[:n].double() - ref).abs().max() / ref.abs().max()).item()
    print("b=%d P=%d: x3 rel err %.2e (argmax mismatches %d of %d), SIMT fp32 rel err %.2e (%d); time x3 %.2f ms, SIMT %.2f ms" % (
        b, P, e3, int((a3[:n].long() != aref).sum()), n * K, e32, int((a32[:n].long() != aref).sum()), t(x3), t(lambda: ops.ground_fwd(q, sd, None, 1000.0, "max"))), flush=True)
