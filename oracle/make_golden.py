"""TEST INFRASTRUCTURE ONLY -- writes tests/golden/*.npz by RUNNING THE UNMODIFIED REFERENCE.

Run in the build container (needs /root/reference):   python oracle/make_golden.py
The vectors are small, committed, and are what pins `oracle/fdt_oracle.py` (and, through it, the
CUDA path) to the reference on machines where /root/reference does not exist (the GPU box).
Every array is produced by reference code imported via `oracle/ref_harness.py`; nothing here
restates reference arithmetic.
"""
import math
import os
import sys
import tempfile

import numpy as np
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_harness as H  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def _np(t):
    return t.detach().cpu().numpy()


def _mask(b, T, lens):
    m = torch.zeros(b, T)
    for i, n in enumerate(lens):
        m[i, n:] = float("-inf")
    return m


def gen_sparsemax():
    ref = H.load_reference()
    g = torch.Generator().manual_seed(11)
    out = {}
    cases = {
        "a": (torch.randn(7, 33, generator=g) * 3, -1),
        "b": (torch.randn(4, 5, 16, generator=g), 1),
        "c": (torch.randn(4, 5, 16, generator=g) * 0.2, -1),
        "d": (2.2e-3 + torch.randn(3, 2048, generator=g) * 4e-4, -1),   # near-uniform, large support (F5)
        "e": (torch.tensor([[0.5, 0.5, 0.5, -1.0], [2.0, 2.0, -3.0, 2.0], [0.0, 0.0, 0.0, 0.0], [10.0, -10.0, 0.0, 9.5]]), -1),
        "f": (torch.randn(2, 1, generator=g), -1),
    }
    for name, (x, dim) in cases.items():
        x = x.clone().requires_grad_(True)
        y = ref["Sparsemax"](dim=dim)(x)
        up = torch.randn(y.shape, generator=g)
        (y * up).sum().backward()
        out[f"{name}_x"] = _np(x)
        out[f"{name}_dim"] = np.array(dim)
        out[f"{name}_y"] = _np(y)
        out[f"{name}_up"] = _np(up)
        out[f"{name}_dx"] = _np(x.grad)
    np.savez_compressed(os.path.join(OUT, "sparsemax.npz"), **out)


def gen_query_model():
    ref = H.load_reference()
    out = {}
    cfgs = [
        ("smax", "sparsemax", "max", 1.0),
        ("smaxT", "sparsemax", "max", 50.0),
        ("softsum", "softmax", "sum", 7.0),
        ("sigmean", "sigmoid", "mean", 3.0),
        ("softmax_max", "softmax", "max", 2.0),
        ("smax_mean", "sparsemax", "mean", 0.5),
    ]
    for ci, (name, att, pool, temp) in enumerate(cfgs):
        torch.manual_seed(100 + ci)
        b, N, Fd, D, K = 5, 9, 24, 32, 64
        qm = ref["Query_model"](ft_dim=Fd, sd_dim=D, temperature=temp, att_func_type=att, pool_type=pool)
        g = torch.Generator().manual_seed(200 + ci)
        ft = torch.randn(b, N, Fd, generator=g).requires_grad_(True)
        sd = torch.randn(K, D, generator=g).requires_grad_(True)
        mask = _mask(b, N, [9, 4, 1, 7, 9])
        up_w = torch.randn(b, K, generator=g)
        up_f = torch.randn(b, D, generator=g)
        for tag, m in (("nomask", None), ("mask", mask)):
            for p_ in list(qm.parameters()) + [ft, sd]:
                p_.grad = None
            w, f, sd_out = qm(ft, sd, mask=m)
            ((w * up_w).sum() + (f * up_f).sum()).backward()
            tok, f2, _ = qm(ft, sd, mask=m, return_token_att=True)
            pre = f"{name}_{tag}_"
            out[pre + "w"] = _np(w)
            out[pre + "f"] = _np(f)
            out[pre + "token_att"] = _np(tok)
            out[pre + "d_ft"] = _np(ft.grad)
            out[pre + "d_sd"] = _np(sd.grad)
            for k, v in qm.named_parameters():
                out[pre + "d_" + k] = _np(v.grad)
        out[name + "_ft"] = _np(ft)
        out[name + "_sd"] = _np(sd)
        out[name + "_mask"] = _np(mask)
        out[name + "_up_w"] = _np(up_w)
        out[name + "_up_f"] = _np(up_f)
        out[name + "_temp"] = np.array(temp)
        for k, v in qm.state_dict().items():
            out[name + "_state_" + k] = _np(v)
    np.savez_compressed(os.path.join(OUT, "query_model.npz"), **out)


def _fdt_inputs(seed, b, P, T, Fi, Ft, K, D, rank=0):
    g = torch.Generator().manual_seed(seed + 17 * rank)
    patch = torch.randn(b, P, Fi, generator=g)
    word = torch.randn(b, T, Ft, generator=g)
    lens = torch.randint(1, T + 1, (b,), generator=g).tolist()
    lens[0] = T
    return patch, word, _mask(b, T, lens)


def _run_fdt_rank(rank, world, tmpdir, p_value, temp):
    """One rank of the unmodified reference: Clip_FDT.forward + ClipInfoCELoss + backward."""
    ref = H.load_reference()
    H.ensure_process_group(rank, world, port=29541)
    b, P, T, Fi, Ft, K, D = 6, 5, 7, 24, 20, 96, 32
    patch, word, mask = _fdt_inputs(300, b, P, T, Fi, Ft, K, D, rank)
    patch.requires_grad_(True)
    word.requires_grad_(True)
    torch.manual_seed(7)  # identical parameters on every rank (DDP broadcast, torch_ddp_dist.py:59)
    model = H.build_reference_fdt(patch, word, mask, K, D, temperature=temp)
    with torch.no_grad():
        model.logit_scale.fill_(p_value)
    crit = ref["ClipInfoCELoss"]()
    with H.cpu_cuda_identity():
        (l_i, l_t), (sd_a, sd_b) = model(torch.zeros(b, 1), ["x"] * b)
        loss, labels = crit(l_i, l_t)
    (loss / world).backward()          # example/clip_fdt/train_solver.py:420
    out = dict(patch=_np(patch), word=_np(word), mask=_np(mask), logits_per_image=_np(l_i), logits_per_text=_np(l_t),
               loss=_np(loss), labels=_np(labels), d_patch=_np(patch.grad), d_word=_np(word.grad),
               temp=np.array(temp), p=np.array(p_value, dtype=np.float32))
    for k, v in model.named_parameters():
        if k.startswith(("visual.", "encode_text.")):
            continue
        out["param_" + k] = _np(v)
        out["grad_" + k] = _np(v.grad) if v.grad is not None else np.zeros(0, np.float32)
    np.savez_compressed(os.path.join(tmpdir, f"r{rank}.npz"), **out)
    import torch.distributed as dist
    dist.barrier()
    dist.destroy_process_group()


def gen_fdt_steps():
    for tag, world, p_value, temp in (("w1", 1, math.log(1 / 0.07), 10.0), ("w1_clamp", 1, 5.0, 10.0),
                                      ("w2", 2, math.log(1 / 0.07), 10.0), ("w2_clamp", 2, 5.0, 3.0)):
        with tempfile.TemporaryDirectory() as td:
            mp.spawn(_run_fdt_rank, args=(world, td, p_value, temp), nprocs=world, join=True)
            merged = {}
            for r in range(world):
                z = np.load(os.path.join(td, f"r{r}.npz"))
                for k in z.files:
                    merged[f"r{r}_{k}"] = z[k]
            merged["world"] = np.array(world)
            np.savez_compressed(os.path.join(OUT, f"clip_fdt_step_{tag}.npz"), **merged)


def _run_clip_rank(rank, world, tmpdir, p_value):
    ref = H.load_reference()
    H.ensure_process_group(rank, world, port=29543)
    b, D = 8, 32
    g = torch.Generator().manual_seed(500 + rank)
    f_img = torch.randn(b, D, generator=g).requires_grad_(True)
    f_txt = torch.randn(b, D, generator=g).requires_grad_(True)
    model = ref["CLIP"](H.FakeFeat(f_img), H.FakeFeat(f_txt), use_allgather=True)
    with torch.no_grad():
        model.logit_scale.fill_(p_value)
    crit = ref["ClipInfoCELoss"]()
    with H.cpu_cuda_identity():
        l_i, l_t = model(torch.zeros(b, 1), ["x"] * b)
        loss, labels = crit(l_i, l_t)
    (loss / world).backward()
    np.savez_compressed(os.path.join(tmpdir, f"r{rank}.npz"), f_img=_np(f_img), f_txt=_np(f_txt),
                        logits_per_image=_np(l_i), logits_per_text=_np(l_t), loss=_np(loss), labels=_np(labels),
                        d_f_img=_np(f_img.grad), d_f_txt=_np(f_txt.grad), d_p=_np(model.logit_scale.grad),
                        p=np.array(p_value, dtype=np.float32))
    import torch.distributed as dist
    dist.barrier()
    dist.destroy_process_group()


def gen_clip_steps():
    for tag, world, p_value in (("w1", 1, math.log(1 / 0.07)), ("w2", 2, 4.8)):
        with tempfile.TemporaryDirectory() as td:
            mp.spawn(_run_clip_rank, args=(world, td, p_value), nprocs=world, join=True)
            merged = {}
            for r in range(world):
                z = np.load(os.path.join(td, f"r{r}.npz"))
                for k in z.files:
                    merged[f"r{r}_{k}"] = z[k]
            merged["world"] = np.array(world)
            np.savez_compressed(os.path.join(OUT, f"clip_step_{tag}.npz"), **merged)


def gen_eval_extract():
    """what CLIP_benchmark/clip_benchmark/models/fdt.py:16-91 does with a trained model -- a checkpoint
    `{"model": {"module.<name>": tensor}}` is loaded (prefix stripped, :43-49) and `extract_img_sd_ft` /
    `extract_txt_sd_ft` give the evaluation embeddings (:55-70) -- run on the unmodified reference in eval mode."""
    H.ensure_process_group(0, 1, port=29547)
    b, P, T, Fi, Ft, K, D = 7, 6, 9, 24, 20, 128, 32
    patch, word, mask = _fdt_inputs(900, b, P, T, Fi, Ft, K, D)
    torch.manual_seed(11)
    model = H.build_reference_fdt(patch, word, mask, K, D, temperature=5.0)
    with torch.no_grad():                                   # a "trained" state: nothing at its init value
        for prm in model.parameters():
            prm.add_(0.05 * torch.randn(prm.shape, generator=torch.Generator().manual_seed(prm.numel())))
    model.eval()
    out = dict(patch=_np(patch), word=_np(word), mask=_np(mask), temp=np.array(5.0))
    for k, v in model.state_dict().items():
        if not k.startswith(("visual.", "encode_text.")):
            out["ckpt_module." + k] = _np(v)                # DDP-style key, as the solver saves it
    with torch.no_grad():
        w_i, f_i, _ = model.extract_img_sd_ft(torch.zeros(b, 1))
        w_t, f_t, _ = model.extract_txt_sd_ft(["x"] * b, raw_text=True)
        tok_i, _, _ = model.extract_img_sd_ft(torch.zeros(b, 1), return_token_att=True)
        tok_t, _, _ = model.extract_txt_sd_ft(["x"] * b, return_token_att=True)
        q_i = model.extract_patch_ft(torch.zeros(b, 1))
        q_t, _ = model.extract_word_ft(["x"] * b)
    out.update(img_att_weight=_np(w_i), img_sd_ft=_np(f_i), txt_att_weight=_np(w_t), txt_sd_ft=_np(f_t),
               img_token_att=_np(tok_i), txt_token_att=_np(tok_t), patch_q=_np(q_i), word_q=_np(q_t))
    np.savez_compressed(os.path.join(OUT, "eval_extract.npz"), **out)
    import torch.distributed as dist
    dist.destroy_process_group()


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(4)
    if sys.argv[1:] == ["eval_extract"]:      # added later: leaves the earlier fixtures byte-identical
        gen_eval_extract()
        sys.exit(0)
    gen_sparsemax()
    gen_query_model()
    gen_fdt_steps()
    gen_clip_steps()
    gen_eval_extract()
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))
