"""Times ilfdt_ground_fwd_ragged against the full-width kernel for fixed caption lengths (per-box TMA cost)."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "iterated-learning-for-vlm_b200"))
from ilfdt import ops
b = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
K, D, T = 16384, 512, 77
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(0)
sd_b = ops.cast_bf16(torch.randn(K, D, device=dev, generator=g))
q = torch.randn(b, T, D, device=dev, generator=g)
def t_of(fn):
    ts = []
    for _ in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return min(ts[1:])
for L in (77, 76, 60, 40, 20, 8, -1):
    if L > 0:
        lens = torch.full((b,), L, device=dev)
    else:
        lens = torch.randint(5, T + 1, (b,), device=dev, generator=g)
    mask = torch.zeros(b, T, device=dev)
    mask[torch.arange(T, device=dev).view(1, -1) >= lens.view(-1, 1)] = float("-inf")
    q_k = ops.cast_bf16_rows(q, mask)
    tf = t_of(lambda: ops.ground_fwd(q_k, sd_b, None, 1000.0, "max"))
    tr = t_of(lambda: ops.ground_fwd_ragged(q_k, sd_b, mask, 1000.0))
    groups = ((torch.clamp(lens + 1, max=T) + 15) // 16).sum().item()
    print("L=%3d full=%.3f ms ragged=%.3f ms groups/sample=%.2f  ragged us per 1k groups=%.2f" % (L, tf, tr, groups / b, tr * 1e3 / (groups / 1e3)), flush=True)
