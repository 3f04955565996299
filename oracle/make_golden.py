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


def _run_reference_config1(case, dtype):
    """The unmodified reference (Clip_FDT.forward + ClipInfoCELoss + backward) on one regenerated config-1 case."""
    ref = H.load_reference()
    from oracle import fdt_oracle as O
    c = case["cfg"]
    patch = case["patch"].detach().clone().to(dtype).requires_grad_(True)
    word = case["word"].detach().clone().to(dtype).requires_grad_(True)
    mask = case["mask"].to(dtype)
    model = H.build_reference_fdt(patch, word, mask, c["K"], c["D"], temperature=c["temperature"])
    if case["variant"] == "bf16in":       # row a0 (q_map) out of the picture: the kernels' input q IS the feature
        model.img_query_model.q_map = torch.nn.Identity()
        model.txt_query_model.q_map = torch.nn.Identity()
    missing, unexpected = model.load_state_dict({k: v.clone() for k, v in case["state"].items()}, strict=False)
    assert not unexpected and all(k.startswith(("visual.", "encode_text.")) for k in missing), (missing, unexpected)
    model = model.to(dtype)
    crit = ref["ClipInfoCELoss"]()
    b = c["b"]
    with H.cpu_cuda_identity():
        (l_i, l_t), _ = model(torch.zeros(b, 1), ["x"] * b)
        loss, labels = crit(l_i, l_t)
    loss.backward()
    with torch.no_grad():
        w_i, f_i, _ = model.extract_img_sd_ft(torch.zeros(b, 1))
        w_t, f_t, _ = model.extract_txt_sd_ft(["x"] * b)
    out = dict(loss=_np(loss), labels=_np(labels), logits_per_image=_np(l_i), logits_per_text=_np(l_t),
               att_ft_img=_np(f_i), att_ft_txt=_np(f_t),
               support_img=np.packbits(_np(w_i) > 0, axis=1), support_txt=np.packbits(_np(w_t) > 0, axis=1),
               nnz_img=(_np(w_i) > 0).sum(1).astype(np.int32), nnz_txt=(_np(w_t) > 0).sum(1).astype(np.int32),
               wmax_img=_np(w_i).max(1), wmax_txt=_np(w_t).max(1),
               proj_d_patch=_np(O.project(case["proj"]["patch"], patch.grad)),
               proj_d_word=_np(O.project(case["proj"]["word"], word.grad)),
               proj_d_space_dict=_np(O.project(case["proj"]["space_dict"], model.space_dict.grad)),
               d_patch_rows=_np(patch.grad[:1]), d_word_rows=_np(word.grad[:1]),           # one whole sample, exact
               d_space_dict_rows=_np(model.space_dict.grad[:64]),
               d_logit_scale=_np(model.logit_scale.grad))
    for k, v in model.named_parameters():
        if ".q_map." in k:
            out["grad_" + k] = _np(v.grad) if v.dim() == 1 else _np(v.grad[:8])    # 1-D in full, 8 rows of the matrices
    raw = dict(d_patch=patch.grad.detach().double(), d_word=word.grad.detach().double(),
               d_space_dict=model.space_dict.grad.detach().double(), d_logit_scale=model.logit_scale.grad.detach().double(),
               loss=loss.detach().double(), logits_per_image=l_i.detach().double(), att_ft_img=f_i.double())
    return out, raw


def gen_config1():
    """BASELINE.json config 1 (b=64, P=49, T=77, K=16384, D=512): the unmodified reference on CPU in fp32 (the golden)
    and in fp64 (the arbiter), inputs regenerated from seeds (oracle.fdt_oracle.config1_case). Also records how far
    the REFERENCE'S OWN fp32 run sits from its fp64 run: the noise floor any fp32 implementation of this path has."""
    import json
    from oracle import fdt_oracle as O
    H.ensure_process_group(0, 1, port=29549)
    floor = {}
    for variant in ("qmap", "bf16in"):
        case = O.config1_case(variant)
        out32, raw32 = _run_reference_config1(case, torch.float32)
        out64, raw64 = _run_reference_config1(case, torch.float64)
        merged = {"f32_" + k: v for k, v in out32.items()}
        for k, v in out64.items():      # the arbiter run: big arrays stored as float32 (6e-8 relative, far below any gate)
            if k.startswith("support_"):
                merged["f64_flips_" + k] = np.flatnonzero(np.unpackbits(v ^ out32[k], axis=1))   # vs the fp32 run
            elif v.dtype == np.float64 and v.size > 4096 and not k.startswith("proj_"):
                merged["f64_" + k] = v.astype(np.float32)
            else:
                merged["f64_" + k] = v
        merged["fp_patch"] = O.fingerprint(case["patch"])
        merged["fp_word"] = O.fingerprint(case["word"])
        merged["fp_mask"] = O.fingerprint(torch.isinf(case["mask"]).float())
        merged["fp_space_dict"] = O.fingerprint(case["state"]["space_dict"])
        for k, v in case["proj"].items():
            merged["fp_proj_" + k] = O.fingerprint(v)
        for k, v in case["state"].items():
            if ".q_map." in k:
                merged["fp_state_" + k] = O.fingerprint(v)
        np.savez_compressed(os.path.join(OUT, f"config1_{variant}.npz"), **merged)
        rel = lambda a, r: float((a - r).abs().max() / r.abs().max().clamp_min(1e-300))
        floor[variant] = {k: rel(raw32[k], raw64[k]) for k in raw32}
        floor[variant]["support_bits_differing_img"] = int(np.unpackbits(out32["support_img"] ^ out64["support_img"]).sum())
        floor[variant]["support_bits_differing_txt"] = int(np.unpackbits(out32["support_txt"] ^ out64["support_txt"]).sum())
    floor["note"] = ("max-norm relative distance ||fp32 - fp64||_inf / ||fp64||_inf of the UNMODIFIED reference run in "
                     "fp32 against the same code run in fp64 (CPU, torch %s), config 1; written by oracle/make_golden.py "
                     "config1" % torch.__version__)
    with open(os.path.join(OUT, "config1_reference_fp32_noise_floor.json"), "w") as fh:
        json.dump(floor, fh, indent=1, sort_keys=True)
    print(json.dumps(floor, indent=1, sort_keys=True))
    import torch.distributed as dist
    dist.destroy_process_group()



def gen_declip():
    """DeCLIP-FDT's own Query_model copy (declip_fdt.py:120-210, `return_token_att` = activated per-token
    attention) and the eight-logits head (declip_fdt.py:381-412) run on the unmodified reference code. The head lines are
    executed through DECLIP_FDT.__dict__-free helpers: its constructor needs `textaugment` (absent), so the normalise /
    scale / logits lines are evaluated on tensors produced by the reference's Query_model instances."""
    H.load_reference()
    import prototype.model.declip_fdt as D
    out = {}
    b, N, Fd, Dm, K = 6, 7, 24, 32, 96
    g = torch.Generator().manual_seed(77)
    sd = torch.randn(K, Dm, generator=g)
    mask = _mask(b, N, [7, 3, 1, 5, 7, 2])
    for name, att in (("smax", "sparsemax"), ("soft", "softmax")):
        torch.manual_seed(5)
        qm = D.Query_model(ft_dim=Fd, sd_dim=Dm, temperature=4.0, att_func_type=att, pool_type="max")
        ft = torch.randn(b, N, Fd, generator=g)
        with torch.no_grad():
            D.Sparsemax  # noqa: B018 -- the module-level import the activation uses
            import prototype.model.sparsemax as ref_sp
            ref_sp.device = torch.device("cpu")
            for tag, m in (("nomask", None), ("mask", mask)):
                tok, f, _ = qm(ft, sd, mask=m, return_token_att=True)
                w, f2, _ = qm(ft, sd, mask=m)
                out[f"{name}_{tag}_token_att"] = _np(tok)
                out[f"{name}_{tag}_f"] = _np(f)
                out[f"{name}_{tag}_w"] = _np(w)
        out[name + "_ft"] = _np(ft)
        for k, v in qm.state_dict().items():
            out[name + "_state_" + k] = _np(v)
    out["sd"], out["mask"], out["temp"] = _np(sd), _np(mask), np.array(4.0)
    # eight logits (declip_fdt.py:382-408) on given pooled features, single process (gathered == local)
    f1, f2, tm, ta = (torch.randn(b, Dm, generator=g) for _ in range(4))
    p = torch.tensor([2.0])
    i1 = f1 / f1.norm(dim=-1, keepdim=True)                 # :382
    i2 = f2 / f2.norm(dim=-1, keepdim=True)                 # :383
    t1 = tm / (tm.norm(dim=-1, keepdim=True) + 1e-10)       # :384
    t2 = ta / (ta.norm(dim=-1, keepdim=True) + 1e-10)       # :385
    s = p.exp()
    s.data = torch.clamp(s.data, max=100)                   # :388-389
    logits = dict(i1=s * i1 @ t1.t(), i2=s * i2 @ t1.t(), i1a=s * i1 @ t2.t(), i2a=s * i2 @ t2.t(),     # :399-402
                  t1=s * t1 @ i1.t(), t2=s * t1 @ i2.t(), t1a=s * t2 @ i1.t(), t2a=s * t2 @ i2.t())     # :405-408
    for k, v in logits.items():
        out["logits_" + k] = _np(v)
    out.update(head_f1=_np(f1), head_f2=_np(f2), head_tm=_np(tm), head_ta=_np(ta), head_p=_np(p))
    np.savez_compressed(os.path.join(OUT, "declip_fdt.npz"), **out)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(4)
    if sys.argv[1:] == ["eval_extract"]:      # added later: leaves the earlier fixtures byte-identical
        gen_eval_extract()
        sys.exit(0)
    if sys.argv[1:] == ["declip"]:
        gen_declip()
        sys.exit(0)
    if sys.argv[1:] == ["config1"]:           # round 2: full-size golden (BASELINE.json config 1)
        torch.set_num_threads(8)
        gen_config1()
        sys.exit(0)
    gen_sparsemax()
    gen_query_model()
    gen_fdt_steps()
    gen_clip_steps()
    gen_eval_extract()
    gen_config1()
    gen_declip()
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))
