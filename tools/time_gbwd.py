"""python tools/time_gbwd.py [b] [split]  -- CUDA-event timing of ilfdt_ground_bwd_w16 (dq + dsd kernels) at the headline statistics
for the shapes of the codebook-gradient kernel (ILFDT_TUNE_DSD_PIPE) and the dq kernel variants; image and text."""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "iterated-learning-for-vlm_b200"))
from ilfdt import ops
dev = torch.device("cuda:0")
b = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
K, D, temp = 16384, 512, 1000.0
g_ = torch.Generator(device="cuda").manual_seed(0)
sd_k = ops.cast_bf16(torch.randn(K, D, device=dev, generator=g_))
out = {}
for name, n_tok, masked in (("image", 49, False), ("text", 77, True)):
    q = torch.randn(b, n_tok, D, device=dev, generator=g_)
    mask = row_base = None
    if masked:
        lens = torch.randint(5, n_tok + 1, (b,), device=dev, generator=g_)
        mask = torch.zeros(b, n_tok, device=dev)
        mask[torch.arange(n_tok, device=dev).view(1, -1) >= lens.view(-1, 1)] = float("-inf")
        q_k, plan = ops.text_compact(q, mask)
        row_base = plan.row_base
        pooled, arg = ops.ground_fwd_text(q_k, sd_k, plan, n_tok, temp)
    else:
        q_k = ops.cast_bf16_rows(q, None)
        pooled, arg = ops.ground_fwd(q_k, sd_k, None, temp, "max")
    del q
    _, w_b, _, nnz = ops.sparsemax_fwd(pooled, want_bf16=True, want_f32=False)
    del pooled
    g = torch.randn(b, K, device=dev, generator=g_)
    gg = torch.empty_like(g)
    dsd = torch.zeros(K, D, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    res = {}
    variants = [dict(dsd_pipe=p) for p in (0, 1, 2, 3)] + [dict(dq_pipe=p) for p in (1, 2)]
    if "split" in sys.argv:      # python tools/time_gbwd.py 4096 split: samples per range of the codebook-gradient sweep
        variants = [dict(dsd_split_rows=r) for r in (512, 1024, 2048, 4096, 0)]
    for tune in variants:
        try:
            with ops.tune(**tune):
                ts = []
                for it in range(4):
                    gg.copy_(g)
                    e0.record()
                    ops.ground_bwd(gg, w_b, arg, q_k, sd_k, mask, temp, dsd, torch.bfloat16, row_base, n_tok, nnz)
                    e1.record()
                    torch.cuda.synchronize()
                    if it:
                        ts.append(e0.elapsed_time(e1))
            res[json.dumps(tune)] = round(sum(ts) / len(ts), 3)
        except Exception as ex:  # noqa: BLE001
            res[json.dumps(tune)] = str(ex)[:80]
    out[name] = res
    del g, gg, w_b, arg, q_k
    torch.cuda.empty_cache()
print(json.dumps({"b": b, "ms": out}, indent=1))
