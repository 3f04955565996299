"""torchrun --nproc-per-node W tools/check_multi_gpu.py
On-hardware W-rank checks of the hot path (NCCL over NVLink), every one against the ORACLE, not against itself:

  1. Clip_FDT.forward + ClipInfoCELoss + backward on W ranks (async bf16 all-gathers, gathered-form backward with the
     all-gather of the row log-sum-exps) == oracle.fdt_oracle.step emulating the same W ranks in fp64 on this rank's GPU
     (that emulation is pinned to the reference's own 2-process run by tests/golden/clip_fdt_step_w2*.npz).
  2. The same model wrapped in DistributedDataParallel(find_unused_parameters=True) as the reference's solver does
     (prototype/utils/torch_ddp_dist.py:65): two steps, codebook gradient == mean over ranks of the oracle's.
  3. A CUDA graph of the step (forward + backward + collectives) replays to the eager result, and releasing it BEFORE
     destroy_process_group() tears NCCL down without blocking (bench.py used to os._exit(0) here).
"""
import math
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "iterated-learning-for-vlm_b200"))
import ilfdt  # noqa: E402
from ilfdt.graphs import GraphedStep  # noqa: E402
from ilfdt.loss_functions import ClipInfoCELoss  # noqa: E402
from ilfdt.model.clip_fdt import Clip_FDT  # noqa: E402
from oracle import fdt_oracle as O  # noqa: E402


class _Visual(torch.nn.Module):
    def forward(self, image, return_dense=False, return_raw_feature=False):
        return None, image, None


class _Text(torch.nn.Module):
    def forward(self, texts, return_dense=False, return_padmask=False, return_raw_feature=False, raw_text=True):
        return [None, texts[0], None, texts[1]]


class _Identity(torch.nn.Module):
    def forward(self, x):
        return x


def rel(a, ref):
    ref = ref.double().to(a.device)
    return ((a.detach().double() - ref).abs().max() / ref.abs().max().clamp_min(1e-30)).item()


def choices(inp, mode, dev):
    """The discrete choices of the kernels on one rank's inputs (argmax token per (sample, code), sparsemax support):
    gradients are compared with the oracle ON THE SAME PIECE (tests/test_fullsize_gpu.py, module docstring) -- a support
    bit at rounding level otherwise moves whole gradient rows for any two fp32 evaluations, the reference's included."""
    from ilfdt import functional as Fn
    out = {}
    sd = inp["sd"].to(dev)
    for tag, q, mask in (("img", inp["q_img"].to(dev), None), ("txt", inp["q_txt"].to(dev), inp["mask"].to(dev))):
        # the same route selection as GroundingFunction.forward (fp32 mode: the tensor-core split route when it applies)
        out["arg_" + tag], _, out["supp_" + tag] = Fn.grounding_choices(q, sd, mask, 1000.0, mode)
    return out


def build(K, D, dev, sd0, p0):
    m = Clip_FDT(_Visual(), _Text(), True, K, D, D, D, "sparsemax", "max", 1000.0).to(dev)
    m.img_query_model.q_map = _Identity()
    m.txt_query_model.q_map = _Identity()
    m.encode_image = lambda image: m.visual(image, return_dense=True, return_raw_feature=True)
    with torch.no_grad():
        m.space_dict.copy_(sd0)
        m.logit_scale.copy_(p0)
    return m


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ok = True
    crit = ClipInfoCELoss()
    b, P, T, K, D = 64, 49, 77, 2048, 512
    for mode, tol, gtol in (("fp32", 1e-5, 5e-5), ("bf16", 2e-2, 2e-2)):
        ilfdt.set_compute_dtype(mode)
        per_rank = [O.synth_inputs(b, P, T, K, D, seed=3, rank=r) for r in range(world)]
        if mode == "bf16":     # bf16 parity is stated on bf16-exact operands
            for inp in per_rank:
                for k in ("q_img", "q_txt", "sd"):
                    inp[k] = inp[k].to(torch.bfloat16).float()
        mine = {k: v.to(dev) for k, v in per_rank[rank].items()}
        # ---- oracle: all W ranks emulated in fp64 on this GPU
        ranks = [dict(q_img=O.make_leaf(i["q_img"].double().to(dev)), q_txt=O.make_leaf(i["q_txt"].double().to(dev)),
                      mask=i["mask"].double().to(dev), sd=O.make_leaf(i["sd"].double().to(dev)),
                      p=O.make_leaf(i["p"].double().to(dev))) for i in per_rank]
        free = O.step(ranks, temperature=1000.0, upstream=[1.0 / world] * world)      # free-running rule
        force = [choices(i, mode, dev) for i in per_rank]
        flips = sum(int((f["supp_" + t] != (fr["w_" + t] > 0)).sum()) for f, fr in zip(force, free) for t in ("img", "txt"))
        for rk in ranks:
            for v in rk.values():
                if isinstance(v, torch.Tensor) and v.grad is not None:
                    v.grad = None
        ref = O.step(ranks, temperature=1000.0, upstream=[1.0 / world] * world, force=force)
        if rank == 0:
            print("%s: %d support bits (of %d) differ from the free-running fp64 rule over all ranks; loss of rank 0 free %.9f / "
                  "same piece %.9f" % (mode, flips, world * 2 * b * K, free[0]["loss"].item(), ref[0]["loss"].item()), flush=True)
        del ranks, free
        # ---- 1. plain module, W ranks over NCCL
        m = build(K, D, dev, mine["sd"], mine["p"])
        qi = mine["q_img"].clone().requires_grad_(True)
        qt = mine["q_txt"].clone().requires_grad_(True)
        (l_i, l_t), _ = m(qi, (qt, mine["mask"]))
        loss, labels = crit(l_i, l_t)
        (loss / world).backward()                      # train_solver.py:420
        r = ref[rank]
        errs = dict(loss=rel(loss, r["loss"]), dq_img=rel(qi.grad, r["d_q_img"]), dq_txt=rel(qt.grad, r["d_q_txt"]),
                    dsd=rel(m.space_dict.grad, r["d_sd"]), dp=rel(m.logit_scale.grad, r["d_p"]),
                    logits=rel(l_i.materialize(), r["logits_per_image"]))
        good = errs["loss"] <= tol and errs["logits"] <= tol and all(errs[k] <= gtol for k in ("dq_img", "dq_txt", "dsd")) \
            and errs["dp"] <= (1e-4 if mode == "fp32" else tol) and torch.equal(labels.cpu(), r["labels"].cpu())
        ok = ok and good
        print("rank %d/%d %s vs fp64 oracle: loss=%.6f %s %s" % (rank, world, mode, loss.item(),
                                                                {k: "%.1e" % v for k, v in errs.items()},
                                                                "OK" if good else "MISMATCH"), flush=True)
        # ---- 2. DDP(find_unused_parameters=True), two steps
        m2 = build(K, D, dev, mine["sd"], mine["p"])
        ddp = torch.nn.parallel.DistributedDataParallel(m2, device_ids=[local], find_unused_parameters=True)
        for step in range(2):
            ddp.zero_grad(set_to_none=True)
            (l_i, l_t), _ = ddp(mine["q_img"].clone().requires_grad_(True), (mine["q_txt"].clone().requires_grad_(True), mine["mask"]))
            loss2, _ = crit(l_i, l_t)
            (loss2 / world).backward()                 # DDP averages parameter gradients over ranks
        dsd_mean = sum(x["d_sd"] for x in ref) / world
        e_ddp = rel(m2.space_dict.grad, dsd_mean)
        good = e_ddp <= gtol and abs(loss2.item() - loss.item()) <= 1e-6 * abs(loss.item())
        ok = ok and good
        print("rank %d/%d %s DDP(find_unused_parameters=True): d space_dict vs mean of oracle ranks %.1e %s" % (
            rank, world, mode, e_ddp, "OK" if good else "MISMATCH"), flush=True)
        del ddp, m2
        # ---- 3. CUDA graph of the step incl. collectives; release before the process group goes away
        if mode == "bf16":
            # (a fresh model: its parameters' gradient accumulators must first be used on the stream the graph is
            # captured on, as with any torch.cuda.graph capture of a backward pass)
            m = build(K, D, dev, mine["sd"], mine["p"])

            def step_fn(a, c_, mk):
                a = a.detach().requires_grad_(True)
                c_ = c_.detach().requires_grad_(True)
                for prm in m.parameters():
                    prm.grad = None
                (l_i, l_t), _ = m(a, (c_, mk))
                ls, _ = crit(l_i, l_t)
                (ls / world).backward()
                return ls
            gs = GraphedStep(step_fn, (mine["q_img"], mine["q_txt"], mine["mask"]), warmup=2)
            for _ in range(3):
                lg = gs()
            good = abs(lg.item() - loss.item()) <= 1e-6 * abs(loss.item()) and rel(m.space_dict.grad, r["d_sd"]) <= gtol
            ok = ok and good
            print("rank %d/%d graph replay (fwd + bwd + NCCL) loss=%.6f %s" % (rank, world, lg.item(), "OK" if good else "MISMATCH"),
                  flush=True)
            gs.release()
            del gs
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    t0 = time.perf_counter()
    dist.barrier()
    torch.cuda.synchronize()
    dist.destroy_process_group()
    print("rank %d: destroy_process_group after graph release took %.2f s" % (rank, time.perf_counter() - t0), flush=True)
    if flag.item() != 0:
        raise SystemExit(1)


if __name__ == "__main__":
    main()
