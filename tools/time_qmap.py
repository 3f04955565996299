"""Times the q_map producer at the headline batch: libilfdt route (QMapFunction, bf16 operands) against
the reference's fp32 nn.Sequential in PyTorch eager on the same GPU. CUDA events, 3 warm-up + 5 timed passes."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "iterated-learning-for-vlm_b200"))
from ilfdt import functional as Fn, ops  # noqa: E402
from ilfdt.model.clip_fdt import Query_model  # noqa: E402

torch.backends.cuda.matmul.allow_tf32 = False   # the reference runs true fp32
dev = torch.device("cuda:0")
b = int(os.environ.get("B", 32768))
out = {}
for name, n_tok, F in (("image", 49, 768), ("text", 77, 512)):
    torch.manual_seed(0)
    qm = Query_model(F, 512, 1000.0, "sparsemax", "max").to(dev)
    x = torch.randn(b, n_tok, F, device=dev, requires_grad=True)
    up_b = torch.randn(b, n_tok, 512, device=dev, dtype=torch.bfloat16)

    def fused():
        q = Fn.q_map_apply(qm.q_map, x)
        q.backward(up_b)

    def eager(chunk=4096):   # chunked: the fp32 autograd graph of the whole batch would not be the point here
        for i in range(0, b, chunk):
            xi = x[i:i + chunk].detach().requires_grad_(True)
            q = qm.q_map(xi)
            q.backward(up_b[i:i + chunk].float())

    for tag, fn, reps in (("ilfdt_bf16", fused, 5), ("torch_fp32_eager", eager, 2)):
        for _ in range(3 if tag == "ilfdt_bf16" else 1):
            fn()
        ops.PROFILE.clear()
        ops.PROFILE_ON = tag == "ilfdt_bf16"
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ops.PROFILE_ON = False
        ms = e0.elapsed_time(e1) / reps
        rows = b * n_tok
        flops = 3 * 2 * rows * (F * 512 + 512 * 512)
        out[name + "/" + tag] = dict(ms=ms, tflops=flops / ms / 1e9)
        if tag == "ilfdt_bf16":
            out[name + "/kernels_ms"] = {k: round(v[1] / reps, 3) for k, v in ops.profile_summary().items()}
        x.grad = None
        for p in qm.parameters():
            p.grad = None
print(json.dumps(out, indent=1))
