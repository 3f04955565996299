"""torchrun --nproc-per-node W tools/check_multi_gpu.py
W-rank run of the hot path (NCCL all-gather / reduce-scatter through ilfdt.functional.AllGather) against a
single-GPU emulation of the same W ranks on rank 0's device (differentiable concatenation, as in
tests/test_parity_gpu.py::test_fdt_step_golden_two_ranks_emulated, which is pinned to the reference's W=2 run)."""
import math
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "iterated-learning-for-vlm_b200"))
from ilfdt import functional as Fn  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ok = True
    for mode, tol in (("fp32", 2e-5), ("bf16", 1e-2)):
        b, P, T, K, D = 64, 49, 77, 2048, 512
        g = torch.Generator().manual_seed(7)
        q_img = torch.randn(world * b, P, D, generator=g)
        q_txt = torch.randn(world * b, T, D, generator=g)
        lens = torch.randint(3, T + 1, (world * b,), generator=g)
        mask = torch.zeros(world * b, T)
        mask[torch.arange(T).view(1, -1) >= lens.view(-1, 1)] = float("-inf")
        sd0 = torch.randn(K, D, generator=g)
        p0 = torch.full((1,), math.log(1 / 0.07))

        def local_feats(r, sd, qi, qt):
            _, f_i = Fn.GroundingFunction.apply(qi, sd, None, 1000.0, mode)
            _, f_t = Fn.GroundingFunction.apply(qt, sd, mask[r * b:(r + 1) * b].to(dev), 1000.0, mode)
            return Fn.L2NormFunction.apply(f_i, 1e-10), Fn.L2NormFunction.apply(f_t, 1e-10)

        # ---- distributed run ----
        sd = sd0.to(dev).requires_grad_(True)
        p = p0.to(dev).requires_grad_(True)
        qi = q_img[rank * b:(rank + 1) * b].to(dev).requires_grad_(True)
        qt = q_txt[rank * b:(rank + 1) * b].to(dev).requires_grad_(True)
        x_i, x_t = local_feats(rank, sd, qi, qt)
        s = p.exp()
        s.data = torch.clamp(s.data, max=100)
        g_i = Fn.AllGather.apply(x_i).view(-1, D)
        g_t = Fn.AllGather.apply(x_t).view(-1, D)
        loss = Fn.InfoNCEFunction.apply(x_i, x_t, g_i, g_t, s, rank * b, mode, True)   # gathered backward: lse all-gather
        (loss / world).backward()

        # ---- emulation of all ranks on this device ----
        sds = [sd0.to(dev).requires_grad_(True) for _ in range(world)]
        ps = [p0.to(dev).requires_grad_(True) for _ in range(world)]
        qis = [q_img[r * b:(r + 1) * b].to(dev).requires_grad_(True) for r in range(world)]
        qts = [q_txt[r * b:(r + 1) * b].to(dev).requires_grad_(True) for r in range(world)]
        feats = [local_feats(r, sds[r], qis[r], qts[r]) for r in range(world)]
        G_i = torch.cat([f[0] for f in feats], 0)
        G_t = torch.cat([f[1] for f in feats], 0)
        total, losses = 0, []
        for r in range(world):
            sr = ps[r].exp()
            sr.data = torch.clamp(sr.data, max=100)
            lr = Fn.InfoNCEFunction.apply(feats[r][0], feats[r][1], G_i, G_t, sr, r * b, mode)
            losses.append(lr)
            total = total + lr / world
        total.backward()

        def rel(a, ref):
            return ((a - ref).abs().max() / ref.abs().max().clamp_min(1e-30)).item()

        errs = dict(loss=rel(loss.detach(), losses[rank].detach()), dq_img=rel(qi.grad, qis[rank].grad),
                    dq_txt=rel(qt.grad, qts[rank].grad), dsd=rel(sd.grad, sds[rank].grad), dp=rel(p.grad, ps[rank].grad))
        good = all(v <= tol for v in errs.values())
        ok = ok and good
        print("rank %d/%d %s loss=%.6f %s %s" % (rank, world, mode, loss.item(), {k: "%.1e" % v for k, v in errs.items()},
                                                 "OK" if good else "MISMATCH"), flush=True)
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    dist.barrier()
    dist.destroy_process_group()
    if flag.item() != 0:
        raise SystemExit(1)


if __name__ == "__main__":
    main()
