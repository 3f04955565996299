"""Times the two gemm_nt shapes of the step (pooling f = w.sd, g = df.sd^T) at b = 32768 (tuning only)."""
import os, sys, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "iterated-learning-for-vlm_b200"))
from ilfdt import ops
dev = torch.device("cuda")
b, K, D = 32768, 16384, 512
sd = torch.randn(K, D, device=dev)
sd_k = ops.cast_bf16(sd); sd_kt = ops.transpose_bf16(sd_k)
w = torch.rand(b, K, device=dev); w_b = ops.cast_bf16(w); del w
df_b = ops.cast_bf16(torch.randn(b, D, device=dev))
def t(fn, n=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
print("route wide=%s" % os.environ.get("ILFDT_GEMM_WIDE", "1"))
print("pool  f = w.sd     [b,K]x[K,D]: %.3f ms  (%.0f TFLOP/s)" % ((tp := t(lambda: ops.gemm_nt(sd_kt, w_b))), 2 * b * K * D / tp / 1e9))
print("g = df.sd^T        [b,D]x[D,K]: %.3f ms  (writes %.1f GB)" % (t(lambda: ops.gemm_nt(sd_k, df_b)), b * K * 4 / 1e9))
print("dsd += w^T.df      gemm_tn    : %.3f ms" % t(lambda: ops.gemm_tn(df_b, w_b)))
buf = torch.empty(b, K, device=dev)
tf = t(lambda: buf.fill_(1.0))
print("torch fill of [b,K] fp32 (2.1 GB): %.3f ms = %.2f TB/s" % (tf, b * K * 4 / tf / 1e9))
src = torch.empty(b, K // 2, device=dev)
tc = t(lambda: buf[:, :K // 2].copy_(src))
