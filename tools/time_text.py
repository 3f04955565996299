"""python tools/time_text.py  -- CUDA-event timing of ilfdt_ground_fwd_text (padded text, lengths U{5..77}, K = 16384, D = 512):
all length buckets in one launch (ground2_text_kernel) against one launch per bucket (ground2_kernel<R>), per batch size.
Decides ILFDT_TUNE_TEXT_MERGED_MAX_B. Both routes produce the same bits (asserted)."""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "iterated-learning-for-vlm_b200"))
from ilfdt import ops
dev = torch.device("cuda:0")
T, D = 77, 512
g_ = torch.Generator(device="cuda").manual_seed(0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
out = []
for K in (16384, 4096):
    sd_k = ops.cast_bf16(torch.randn(K, D, device=dev, generator=g_))
    for b in (256, 1024, 2048, 4096, 8192, 16384, 32768):
        q = torch.randn(b, T, D, device=dev, generator=g_)
        lens = torch.randint(5, T + 1, (b,), device=dev, generator=g_)
        mask = torch.zeros(b, T, device=dev)
        mask[torch.arange(T, device=dev).view(1, -1) >= lens.view(-1, 1)] = float("-inf")
        q_c, plan = ops.text_compact(q, mask)
        res, ref = {}, None
        for name, mb in (("merged", 1 << 30), ("per_bucket", 0)):
            with ops.tune(text_merged_max_b=mb):
                ts = []
                for it in range(8):
                    flush.zero_()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    pooled, arg = ops.ground_fwd_text(q_c, sd_k, plan, T, 1000.0)
                    e1.record()
                    torch.cuda.synchronize()
                    if it >= 3:
                        ts.append(e0.elapsed_time(e1))
            if ref is None:
                ref = (pooled.clone(), arg.clone())
            else:
                assert torch.equal(ref[0], pooled) and torch.equal(ref[1], arg)
            res[name] = round(sorted(ts)[len(ts) // 2] * 1e3, 1)
        out.append(dict(K=K, b=b, us=res))
        print(json.dumps(out[-1]), flush=True)
        del q, mask, q_c, plan, pooled, arg, ref
