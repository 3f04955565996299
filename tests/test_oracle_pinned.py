"""Pins oracle/fdt_oracle.py to OUTPUTS OF THE REFERENCE (tests/golden/*.npz, written by
oracle/make_golden.py from the unmodified reference). CPU only."""
import math
import os

import numpy as np
import pytest
import torch

from oracle import fdt_oracle as O

TOL = dict(rtol=2e-5, atol=2e-6)


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def _t(a):
    return torch.as_tensor(np.array(a))


def _close(a, b, rtol=2e-5, atol=2e-6):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    scale = max(np.abs(b).max(), 1e-30)
    err = np.abs(a - b).max()
    assert err <= atol + rtol * scale, f"max err {err:.3e} vs scale {scale:.3e}"


@pytest.mark.parametrize("case", list("abcdef"))
def test_sparsemax_golden(golden_dir, case):
    z = _load(golden_dir, "sparsemax.npz")
    x = _t(z[f"{case}_x"]).requires_grad_(True)
    dim = int(z[f"{case}_dim"])
    y = O.sparsemax(x, dim=dim)
    ref = z[f"{case}_y"]
    _close(y.detach().numpy(), ref, rtol=1e-6, atol=1e-7)
    assert np.array_equal(y.detach().numpy() > 0, ref > 0), "support sets differ from the reference"
    (y * _t(z[f"{case}_up"])).sum().backward()
    _close(x.grad.numpy(), z[f"{case}_dx"], rtol=1e-5, atol=1e-6)
    # rows on the simplex
    s = y.detach().movedim(dim, -1).sum(-1)
    assert torch.allclose(s, torch.ones_like(s), atol=1e-5)


QM = [("smax", "sparsemax", "max"), ("smaxT", "sparsemax", "max"), ("softsum", "softmax", "sum"),
      ("sigmean", "sigmoid", "mean"), ("softmax_max", "softmax", "max"), ("smax_mean", "sparsemax", "mean")]


@pytest.mark.parametrize("name,att,pool", QM)
@pytest.mark.parametrize("tag", ["nomask", "mask"])
def test_query_model_golden(golden_dir, name, att, pool, tag):
    z = _load(golden_dir, "query_model.npz")
    state = {k[len(name) + 7:]: z[k] for k in z.files if k.startswith(name + "_state_")}
    ft = _t(z[name + "_ft"]).requires_grad_(True)
    sd = _t(z[name + "_sd"]).requires_grad_(True)
    mask = _t(z[name + "_mask"]) if tag == "mask" else None
    temp = float(z[name + "_temp"])
    q_map = O.make_q_map(ft.shape[-1], sd.shape[-1], state)
    q = q_map(ft)
    w, f, pooled = O.grounding(q, sd, mask, temp, att, pool)
    pre = f"{name}_{tag}_"
    _close(w.detach().numpy(), z[pre + "w"])
    _close(f.detach().numpy(), z[pre + "f"])
    _, tok = O.pooled_scores(q, sd, mask, temp, pool, return_token_scores=True)
    _close(tok.detach().numpy(), z[pre + "token_att"])
    ((w * _t(z[name + "_up_w"])).sum() + (f * _t(z[name + "_up_f"])).sum()).backward()
    _close(ft.grad.numpy(), z[pre + "d_ft"], rtol=1e-4, atol=1e-6)
    _close(sd.grad.numpy(), z[pre + "d_sd"], rtol=1e-4, atol=1e-6)
    for k, v in q_map.named_parameters():
        _close(v.grad.numpy(), z[pre + "d_q_map." + k], rtol=1e-4, atol=1e-6)


def test_eval_extract_golden(golden_dir):
    """The evaluation flow's outputs (extract_*_sd_ft in eval mode on a loaded checkpoint)."""
    z = _load(golden_dir, "eval_extract.npz")
    state = {k[len("ckpt_module."):]: z[k] for k in z.files if k.startswith("ckpt_module.")}
    sd = _t(state["space_dict"])
    temp = float(z["temp"])
    for mod, src, mask, tag in (("img_query_model", "patch", None, "img"), ("txt_query_model", "word", _t(z["mask"]), "txt")):
        x = _t(z[src])
        q_map = O.make_q_map(x.shape[-1], sd.shape[1], state, prefix=mod + ".q_map.")
        with torch.no_grad():
            q = q_map(x)
            w, f, _ = O.grounding(q, sd, mask, temp, "sparsemax", "max")
            _, tok = O.pooled_scores(q, sd, mask, temp, "max", return_token_scores=True)
        _close(q.numpy(), z["patch_q" if tag == "img" else "word_q"])
        _close(w.numpy(), z[tag + "_att_weight"])
        _close(f.numpy(), z[tag + "_sd_ft"])
        _close(tok.numpy(), z[tag + "_token_att"])
        assert np.array_equal(w.numpy() > 0, z[tag + "_att_weight"] > 0)


def _fdt_ranks(z, world):
    ranks, qmaps = [], []
    for r in range(world):
        pre = f"r{r}_"
        params = {k[len(pre) + 6:]: z[k] for k in z.files if k.startswith(pre + "param_")}
        patch = _t(z[pre + "patch"]).requires_grad_(True)
        word = _t(z[pre + "word"]).requires_grad_(True)
        D = params["space_dict"].shape[1]
        qi = O.make_q_map(patch.shape[-1], D, params, prefix="img_query_model.q_map.")
        qt = O.make_q_map(word.shape[-1], D, params, prefix="txt_query_model.q_map.")
        sd = O.make_leaf(_t(params["space_dict"]))
        p = O.make_leaf(_t(params["logit_scale"]))
        ranks.append(dict(q_img=qi(patch), q_txt=qt(word), mask=_t(z[pre + "mask"]), sd=sd, p=p))
        qmaps.append((patch, word, qi, qt, sd, p))
    return ranks, qmaps


@pytest.mark.parametrize("tag", ["w1", "w1_clamp", "w2", "w2_clamp"])
def test_fdt_step_golden(golden_dir, tag):
    z = _load(golden_dir, f"clip_fdt_step_{tag}.npz")
    world = int(z["world"])
    ranks, leaves = _fdt_ranks(z, world)
    for r in ranks:  # q is a non-leaf here; keep its grad for nothing, grads flow to patch/word
        pass
    out = O.step(ranks, temperature=float(z["r0_temp"]), head="fdt", upstream=[1.0 / world] * world)
    for r in range(world):
        pre = f"r{r}_"
        o = out[r]
        patch, word, qi, qt, sd, p = leaves[r]
        _close(o["logits_per_image"].numpy(), z[pre + "logits_per_image"])
        _close(o["logits_per_text"].numpy(), z[pre + "logits_per_text"])
        _close(o["loss"].numpy(), z[pre + "loss"], rtol=1e-6)
        assert np.array_equal(o["labels"].numpy(), z[pre + "labels"])
        _close(patch.grad.numpy(), z[pre + "d_patch"], rtol=2e-4)
        _close(word.grad.numpy(), z[pre + "d_word"], rtol=2e-4)
        _close(sd.grad.numpy(), z[pre + "grad_space_dict"], rtol=2e-4)
        _close(p.grad.numpy(), z[pre + "grad_logit_scale"], rtol=2e-4)
        for k, v in qi.named_parameters():
            _close(v.grad.numpy(), z[pre + "grad_img_query_model.q_map." + k], rtol=2e-4, atol=1e-6)
        for k, v in qt.named_parameters():
            _close(v.grad.numpy(), z[pre + "grad_txt_query_model.q_map." + k], rtol=2e-4, atol=1e-6)
        # logit_scale_sd is declared but unused (clip_fdt.py:227): no gradient in the reference
        assert z[pre + "grad_logit_scale_sd"].size == 0
    if "clamp" in tag:  # F7: forward clamps to 100, backward uses exp(p) = 148.4
        assert float(out[0]["scale"]) == 100.0


@pytest.mark.parametrize("tag", ["w1", "w2"])
def test_clip_step_golden(golden_dir, tag):
    z = _load(golden_dir, f"clip_step_{tag}.npz")
    world = int(z["world"])
    ranks = []
    for r in range(world):
        pre = f"r{r}_"
        ranks.append(dict(f_img=O.make_leaf(_t(z[pre + "f_img"])), f_txt=O.make_leaf(_t(z[pre + "f_txt"])),
                          p=O.make_leaf(_t(z[pre + "p"]).reshape(1))))
    out = O.step(ranks, head="clip", upstream=[1.0 / world] * world, eps_img=0.0, eps_txt=1e-10)
    for r in range(world):
        pre = f"r{r}_"
        _close(out[r]["logits_per_image"].numpy(), z[pre + "logits_per_image"])
        _close(out[r]["logits_per_text"].numpy(), z[pre + "logits_per_text"])
        _close(out[r]["loss"].numpy(), z[pre + "loss"], rtol=1e-6)
        assert np.array_equal(out[r]["labels"].numpy(), z[pre + "labels"])
        _close(out[r]["d_f_img"].numpy(), z[pre + "d_f_img"], rtol=2e-4)
        _close(out[r]["d_f_txt"].numpy(), z[pre + "d_f_txt"], rtol=2e-4)
        _close(out[r]["d_p"].numpy(), z[pre + "d_p"], rtol=2e-4)


def test_sparsemax_closed_form_gradient():
    """autograd through sort/cumsum equals nz * (g - mean_nz(g))."""
    g = torch.Generator().manual_seed(3)
    x = (torch.randn(9, 300, generator=g) * 0.01).requires_grad_(True)
    up = torch.randn(9, 300, generator=g)
    y = O.sparsemax(x)
    (y * up).sum().backward()
    nz = (y > 0).float()
    closed = nz * (up - (up * nz).sum(1, keepdim=True) / nz.sum(1, keepdim=True))
    _close(x.grad.numpy(), closed.numpy(), rtol=1e-4, atol=1e-6)


def test_fp64_arbiter_agrees_on_clear_rows():
    g = torch.Generator().manual_seed(5)
    x = torch.randn(6, 500, generator=g)
    y = O.sparsemax(x).numpy()
    for i in range(6):
        tau, supp, margin = O.sparsemax_tau_exact(x[i].numpy())
        if margin > 1e-5:
            assert np.array_equal(y[i] > 0, supp)


def test_philox_known_answers():
    """Random123 known-answer vectors for Philox4x32-10."""
    kat = [
        ((0, 0, 0, 0), (0, 0), (0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8)),
        ((0xFFFFFFFF,) * 4, (0xFFFFFFFF,) * 2, (0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD)),
        ((0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344), (0xA4093822, 0x299F31D0),
         (0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1)),
    ]
    for ctr, key, exp in kat:
        got = O.philox4x32_10(np.array([ctr], dtype=np.uint32), np.array(key, dtype=np.uint32))[0]
        assert tuple(int(v) for v in got) == exp


def test_philox_randn_moments():
    x = O.philox_randn(1 << 18, seed=1234, generation=3).astype(np.float64)
    assert abs(x.mean()) < 0.01 and abs(x.std() - 1) < 0.01
    assert abs((x ** 3).mean()) < 0.03 and abs((x ** 4).mean() - 3) < 0.08
    y = O.philox_randn(1 << 10, seed=1234, generation=4)
    assert not np.array_equal(x[:1024].astype(np.float32), y)


def test_live_reference_matches_oracle_when_present():
    """Where /root/reference exists (build container), run it live on a fresh seed."""
    from oracle import ref_harness as H
    if not H.reference_available():
        pytest.skip("reference tree not present (GPU box)")
    ref = H.load_reference()
    g = torch.Generator().manual_seed(99)
    x = torch.randn(5, 777, generator=g) * 0.05
    assert torch.equal(ref["Sparsemax"](dim=-1)(x) > 0, O.sparsemax(x) > 0)
    _close(O.sparsemax(x).numpy(), ref["Sparsemax"](dim=-1)(x).numpy(), rtol=1e-6, atol=1e-8)
    torch.manual_seed(1)
    qm = ref["Query_model"](ft_dim=16, sd_dim=32, temperature=1000.0, att_func_type="sparsemax", pool_type="max")
    ft = torch.randn(4, 11, 16, generator=g)
    sd = torch.randn(128, 32, generator=g)
    mask = torch.zeros(4, 11)
    mask[1, 3:] = float("-inf")
    w_ref, f_ref, _ = qm(ft, sd, mask=mask)
    w, f, _ = O.grounding(qm.q_map(ft), sd, mask, 1000.0)
    _close(w.detach().numpy(), w_ref.detach().numpy(), rtol=1e-6, atol=1e-9)
    _close(f.detach().numpy(), f_ref.detach().numpy(), rtol=1e-6, atol=1e-9)


@pytest.mark.parametrize("variant", ["qmap", "bf16in"])
def test_config1_golden_pins_the_oracle_at_full_size(golden_dir, variant):
    """BASELINE.json config 1 (b=64, P=49, T=77, K=16384, D=512): the oracle restatement on the regenerated inputs
    against what the unmodified reference produced (tests/golden/config1_*.npz, oracle/make_golden.py config1)."""
    z = _load(golden_dir, "config1_%s.npz" % variant)
    case = O.config1_case(variant)
    for name, t in (("patch", case["patch"]), ("word", case["word"]), ("mask", torch.isinf(case["mask"]).float()),
                    ("space_dict", case["state"]["space_dict"])):
        assert np.allclose(O.fingerprint(t), z["fp_" + name], rtol=1e-12, atol=0), name
    c = case["cfg"]
    patch = case["patch"].clone().requires_grad_(True)
    word = case["word"].clone().requires_grad_(True)
    if variant == "qmap":
        qi = O.make_q_map(c["Fi"], c["D"], case["state"], prefix="img_query_model.q_map.")
        qt = O.make_q_map(c["Ft"], c["D"], case["state"], prefix="txt_query_model.q_map.")
        q_img, q_txt = qi(patch), qt(word)
    else:
        q_img, q_txt = patch * 1.0, word * 1.0
    sd = O.make_leaf(case["state"]["space_dict"])
    p = O.make_leaf(case["state"]["logit_scale"])
    out = O.step([dict(q_img=q_img, q_txt=q_txt, mask=case["mask"], sd=sd, p=p)], temperature=c["temperature"])[0]
    _close(out["loss"].numpy(), z["f32_loss"], rtol=1e-6)
    assert np.array_equal(out["labels"].numpy(), z["f32_labels"])
    _close(out["logits_per_image"].numpy(), z["f32_logits_per_image"])
    _close(out["logits_per_text"].numpy(), z["f32_logits_per_text"])
    _close(out["f_img"].numpy(), z["f32_att_ft_img"])
    _close(out["f_txt"].numpy(), z["f32_att_ft_txt"])
    for tag in ("img", "txt"):
        bits = np.unpackbits(z["f32_support_" + tag], axis=1)[:, :c["K"]].astype(bool)
        assert (bits != (out["w_" + tag].numpy() > 0)).sum() <= 2, "support sets differ from the reference"
    # projections see every gradient row; the rows stored verbatim are compared with the same rule the GPU tests use
    _close(O.project(case["proj"]["patch"], patch.grad).numpy(), z["f32_proj_d_patch"], rtol=2e-4)
    _close(O.project(case["proj"]["word"], word.grad).numpy(), z["f32_proj_d_word"], rtol=2e-4)
    _close(O.project(case["proj"]["space_dict"], sd.grad).numpy(), z["f32_proj_d_space_dict"], rtol=2e-4)
    _close(p.grad.numpy(), z["f32_d_logit_scale"], rtol=2e-4)
    _close(patch.grad[:1].numpy(), z["f32_d_patch_rows"], rtol=2e-4)
    _close(sd.grad[:64].numpy(), z["f32_d_space_dict_rows"], rtol=2e-4)


def test_reference_fp32_noise_floor_is_recorded(golden_dir):
    """The tolerance argument of tests/test_fullsize_gpu.py rests on these committed numbers: the reference's own fp32
    run vs its fp64 run on config 1."""
    import json
    f = json.load(open(os.path.join(golden_dir, "config1_reference_fp32_noise_floor.json")))
    exact = f["bf16in"]                  # no argmax ties on these inputs: pure fp32 rounding
    assert exact["loss"] < 1e-7 and exact["logits_per_image"] < 2e-6
    assert 1e-6 < max(exact["d_patch"], exact["d_word"], exact["d_space_dict"]) < 1e-5
    assert f["qmap"]["d_space_dict"] > 1e-3   # an argmax / support tie: O(1e-2) even for the reference itself


def test_forced_discrete_choices_reproduce_the_free_running_oracle():
    """The test hooks (force_arg / force_support: evaluate max-pool and sparsemax on GIVEN argmax tokens and support)
    change nothing when handed the free-running rule's own choices -- values and gradients are identical -- and make
    the discontinuity of the sparsemax Jacobian explicit: flipping ONE membership bit of weight ~0 moves the gradient
    by O(1) of a coefficient although the output moves by ~0."""
    inp = O.synth_inputs(6, 9, 11, 96, 32, seed=4)
    def run(force=None):
        r = [dict(q_img=O.make_leaf(inp["q_img"].double()), q_txt=O.make_leaf(inp["q_txt"].double()),
                  mask=inp["mask"].double(), sd=O.make_leaf(inp["sd"].double()), p=O.make_leaf(inp["p"].double()))]
        return O.step(r, temperature=10.0, force=force)[0]
    free = run()
    _, tok_i = O.pooled_scores(inp["q_img"].double(), inp["sd"].double(), None, 10.0, "max", True)
    _, tok_t = O.pooled_scores(inp["q_txt"].double(), inp["sd"].double(), inp["mask"].double(), 10.0, "max", True)
    force = [dict(arg_img=tok_i.argmax(1), arg_txt=tok_t.argmax(1), supp_img=free["w_img"] > 0, supp_txt=free["w_txt"] > 0)]
    forced = run(force)
    for k in ("loss", "w_img", "w_txt", "f_img", "d_q_img", "d_q_txt", "d_sd", "d_p"):
        _close(forced[k].numpy(), free[k].numpy(), rtol=1e-12, atol=1e-15)
    # flip the support entry with the smallest weight of sample 0: output barely moves, gradient moves a lot
    w0 = free["w_img"][0]
    k = int(torch.where(w0 > 0, w0, torch.full_like(w0, 9.0)).argmin())
    supp = (free["w_img"] > 0).clone()
    supp[0, k] = False
    flipped = run([dict(force[0], supp_img=supp)])
    dw = (flipped["w_img"] - free["w_img"]).abs().max().item()
    dg = (flipped["d_q_img"][0] - free["d_q_img"][0]).abs().max().item() / free["d_q_img"][0].abs().max().item()
    assert dg > 0.1 and dg > 5 * dw / w0.max().item(), (dw, dg)   # 44 % of the gradient for 3.5 % of a weight
