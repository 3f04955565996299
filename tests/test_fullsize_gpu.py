"""Full-size parity (-m gpu): the code paths bench.py times, at the sizes that select them, against the ORACLE --
not against themselves. The oracle (oracle/fdt_oracle.py, plain torch) is run in fp64 ON THE GPU here, so BASELINE.json
configs 1 / 3 / 3b / 5 and the 32768-row InfoNCE take seconds. Reference-generated goldens at config 1
(tests/golden/config1_*.npz, written by oracle/make_golden.py from the unmodified reference) are checked in BOTH
compute modes.

Gates (BASELINE.json north_star): fp32 mode <= 1e-5 relative on loss / logits / embeddings; bf16 mode <= 2e-2; integer
outputs exact. Gradients in fp32 mode: <= 1e-5 as well (measured ~2e-6, i.e. at the distance between the REFERENCE'S
OWN fp32 and fp64 runs on config 1: 2.7e-6 .. 4.6e-6, tests/golden/config1_reference_fp32_noise_floor.json); the one
wider gate is 2e-5 on the InfoNCE gradient at B = 32768, a sum of 32768 fp32 products per element (measured 1.04e-5).

The path has two DISCRETE choices per (sample, code): the token the max-pool picks and whether the code is in the
sparsemax support. Outputs are continuous across both, GRADIENTS ARE NOT: the max-pool routes the whole gradient of a
code to one token, and a code that sits at the sparsemax threshold to within rounding carries weight ~0 but a full-size
coefficient g_k - mean_S(g) (and moves mean_S for the whole sample). Two correct evaluations that differ in one such bit
differ in the gradient by O(1) of a row -- the reference's own fp32 run is 2.9e-2 .. 5.7e-2 away from its fp64 run on the
unrounded config-1 inputs for exactly this reason (same JSON; one support bit). Parity of gradients is therefore
stated ON THE SAME PIECE: (1) the implementation's discrete choices are compared with the free-running fp64 oracle and
may differ only at rounding-ambiguous entries (top-2 token gap <= 4e-6 max|score|; |score - tau| <= 1e-6 max|pooled|),
whose count is asserted tiny; (2) the fp64 oracle is then evaluated with those choices given
(oracle.fdt_oracle.step(force=...)) and EVERY gradient row must meet the gate against it, nothing skipped.
"""
import json
import math
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import fdt_oracle as O  # noqa: E402

FP32_TOL = 1e-5
FP32_GRAD_TOL = 1e-5       # on the same piece (see the module docstring); measured 2e-6 at b = 256, K = 16384
BF16_TOL = 2e-2
TIE_EPS = 4e-6            # relative (to max |score|) top-2 token gap below which the argmax is rounding-ambiguous


def _dev():
    return torch.device("cuda:0")


def rel_err(a, ref):
    a = a.detach().double()
    ref = ref.detach().double().to(a.device)
    return (a - ref).abs().max().item() / max(ref.abs().max().item(), 1e-30)


def assert_close(a, ref, tol, what=""):
    e = rel_err(a, ref)
    assert e <= tol, "%s: relative error %.3e > %.1e" % (what, e, tol)
    return e


def bf16_round(t):
    return t.to(torch.bfloat16).to(torch.float32)


@pytest.fixture(scope="module")
def ops():
    from ilfdt import _lib
    from ilfdt import ops as _ops
    _lib.load()
    return _ops


@pytest.fixture(autouse=True)
def _free():
    yield
    torch.cuda.empty_cache()


MARGIN_EPS = 1e-6         # relative (to max |pooled|) distance from the sparsemax threshold below which membership is ambiguous


def _our_choices(ops, q, sd, mask, temp, mode):
    """(arg [b,K] int64, pooled) exactly as GroundingFunction computes them: through the same route selection
    (ilfdt.functional._score_pool), so the fp32 mode's tensor-core split route is what is looked at when it is taken."""
    from ilfdt import functional as Fn
    arg, pooled, _ = Fn.grounding_choices(q, sd, mask, temp, mode)
    return arg, pooled


def _check_choices(q64, sd64, mask, temp, arg, w, w_ref, what, max_ambiguous=64):
    """The implementation's argmax tokens (inside its support) and support set vs the free-running fp64 rule: every
    disagreement must be rounding-ambiguous. Returns (#argmax disagreements, #support disagreements)."""
    b, N, D = q64.shape
    n_arg = 0
    supp = w > 0
    for b0 in range(0, b, 64):
        s = torch.matmul(q64[b0:b0 + 64], sd64.t())
        if mask is not None:
            s = s * (mask[b0:b0 + 64] == 0).to(s.dtype).unsqueeze(-1)
        top, idx = s.topk(2, dim=1)
        mis = (idx[:, 0] != arg[b0:b0 + 64]) & supp[b0:b0 + 64]
        if mask is not None:     # every pad token scores exactly 0 (clip_fdt.py:125-130): which pad holds the max is immaterial
            pad = mask[b0:b0 + 64] != 0
            mis &= ~(torch.gather(pad, 1, arg[b0:b0 + 64]) & torch.gather(pad, 1, idx[:, 0]))
        if mis.any():
            gap = (top[:, 0] - top[:, 1])[mis]
            picked_second = (idx[:, 1] == arg[b0:b0 + 64])[mis]
            assert bool(picked_second.all()) and gap.max().item() <= TIE_EPS * s.abs().max().item(), \
                "%s: argmax differs from the fp64 rule at a gap of %.3e (max|score| %.1f)" % (what, gap.max().item(), s.abs().max().item())
            n_arg += int(mis.sum())
    # support: tau of the free-running rule per row from any support entry: tau = pooled - w
    pooled_ref = O.pooled_scores(q64, sd64, mask, temp, "max")
    pooled_ref = pooled_ref - pooled_ref.max(1, keepdim=True)[0]
    k_star = w_ref.argmax(1, keepdim=True)
    tau = torch.gather(pooled_ref, 1, k_star) - torch.gather(w_ref, 1, k_star)
    mis = supp != (w_ref > 0)
    n_sup = int(mis.sum())
    if n_sup:
        margin = (pooled_ref - tau).abs()[mis].max().item()
        assert margin <= MARGIN_EPS * pooled_ref.abs().max().item(), \
            "%s: support differs at an entry %.3e from the threshold (max|pooled| %.3e)" % (what, margin, pooled_ref.abs().max().item())
    assert n_arg + n_sup <= max_ambiguous, "%s: %d argmax + %d support disagreements" % (what, n_arg, n_sup)
    return n_arg, n_sup


# ---------------------------------------------------------------------------------------------------
# (v) reference-generated golden at BASELINE.json config 1, both compute modes
# ---------------------------------------------------------------------------------------------------
class _FakeVisual(torch.nn.Module):
    def __init__(self, patch_ft):
        super().__init__()
        self.conv1 = torch.nn.Conv2d(1, 1, 1)
        self.patch_ft = patch_ft

    def forward(self, image, return_dense=False, return_raw_feature=False):
        cls = self.patch_ft.mean(1)
        return cls, self.patch_ft, cls


class _FakeText(torch.nn.Module):
    def __init__(self, word_ft, pad_mask):
        super().__init__()
        self.text_projection = torch.nn.Linear(1, 1)
        self.word_ft, self.pad_mask = word_ft, pad_mask

    def forward(self, texts, return_dense=False, return_padmask=False, return_raw_feature=False, raw_text=True):
        cls = self.word_ft.mean(1)
        return [cls, self.word_ft, cls, self.pad_mask]


def _case_model(case, dev):
    from ilfdt.model.clip_fdt import Clip_FDT
    c = case["cfg"]
    patch = case["patch"].to(dev).requires_grad_(True)
    word = case["word"].to(dev).requires_grad_(True)
    m = Clip_FDT(_FakeVisual(patch), _FakeText(word, case["mask"].to(dev)), True, c["K"], c["D"], patch.shape[-1],
                 word.shape[-1], "sparsemax", "max", c["temperature"]).to(dev)
    if case["variant"] == "bf16in":
        m.img_query_model.q_map = torch.nn.Identity()
        m.txt_query_model.q_map = torch.nn.Identity()
    missing, unexpected = m.load_state_dict({k: v.clone() for k, v in case["state"].items()}, strict=False)
    assert not unexpected and all(k.startswith(("visual.", "encode_text.")) for k in missing), (missing, unexpected)
    return m, patch, word


@pytest.mark.parametrize("variant", ["qmap", "bf16in"])
@pytest.mark.parametrize("mode", ["fp32", "bf16"])
def test_config1_reference_golden(ops, golden_dir, variant, mode):
    """Clip_FDT.forward + ClipInfoCELoss + backward through the drop-in modules at b=64, P=49, T=77, K=16384, D=512
    == what the UNMODIFIED reference produced on the same regenerated inputs (f32_* arrays: its fp32 run; f64_*: the
    same code in fp64, the arbiter). "qmap": raw features through the real q_map MLPs; "bf16in": q_map = identity
    and bf16-exact q / codebook, so the tensor-core mode sees the very operands the reference saw."""
    import ilfdt
    from ilfdt.loss_functions import ClipInfoCELoss
    z = np.load(os.path.join(golden_dir, "config1_%s.npz" % variant))
    case = O.config1_case(variant)
    for name, t in (("patch", case["patch"]), ("word", case["word"]), ("mask", torch.isinf(case["mask"]).float()),
                    ("space_dict", case["state"]["space_dict"])):
        assert np.allclose(O.fingerprint(t), z["fp_" + name], rtol=1e-12, atol=0), \
            "regenerated %s differs from the tensor the golden was made from" % name
    for k, v in case["proj"].items():
        assert np.array_equal(O.fingerprint(v), z["fp_proj_" + k])
    dev = _dev()
    c = case["cfg"]
    b = c["b"]
    with ilfdt.compute_dtype(mode):
        m, patch, word = _case_model(case, dev)
        (l_i, l_t), _ = m(torch.zeros(b, 1, device=dev), ["x"] * b)
        loss, labels = ClipInfoCELoss()(l_i, l_t)
        assert not l_i.is_materialized
        loss.backward()
        with torch.no_grad():
            w_i, f_i, _ = m.extract_img_sd_ft(torch.zeros(b, 1, device=dev))
            w_t, f_t, _ = m.extract_txt_sd_ft(["x"] * b)
        li, lt = l_i.materialize(), l_t.materialize()
    T = lambda k: torch.tensor(z[k])
    tol = FP32_TOL if mode == "fp32" else BF16_TOL
    assert np.array_equal(labels.cpu().numpy(), z["f32_labels"])
    report = {}
    for pre in ("f32_", "f64_"):           # the reference's fp32 run and its fp64 run: both within the gate
        report[pre + "loss"] = assert_close(loss, T(pre + "loss"), tol, pre + "loss")
        report[pre + "logits_i"] = assert_close(li, T(pre + "logits_per_image"), tol, pre + "logits_per_image")
        report[pre + "logits_t"] = assert_close(lt, T(pre + "logits_per_text"), tol, pre + "logits_per_text")
        report[pre + "att_ft_i"] = assert_close(f_i, T(pre + "att_ft_img"), tol, pre + "att_ft_img")
        report[pre + "att_ft_t"] = assert_close(f_t, T(pre + "att_ft_txt"), tol, pre + "att_ft_txt")
    # support sets: identical to the reference's, except entries that the reference's own fp32 and fp64 runs disagree on
    # or whose weight is at rounding level (sparsemax is continuous: a flipped entry carries weight ~ 0)
    exact_inputs = mode == "fp32" or variant == "bf16in"
    for tag, w in (("img", w_i), ("txt", w_t)):
        ref_bits = torch.tensor(np.unpackbits(z["f32_support_" + tag], axis=1)[:, :c["K"]].astype(bool))
        mism = ((w > 0).cpu() != ref_bits)
        wmax = float(z["f32_wmax_" + tag].max())
        flipped_w = w.cpu()[mism].max().item() if mism.any() else 0.0
        report["support_bits_differing_" + tag] = int(mism.sum())
        if exact_inputs:
            assert int(mism.sum()) <= 8 and flipped_w <= 1e-6 * wmax, \
                "support (%s) differs from the reference at %d entries, largest weight %.3e" % (tag, int(mism.sum()), flipped_w)
        nnz_ref = z["f32_nnz_" + tag].astype(np.int64)
        assert np.abs((w > 0).sum(1).cpu().numpy() - nnz_ref).max() <= (8 if exact_inputs else 0.05 * nnz_ref.max())
    if exact_inputs:
        # gradients. Projections (16 random +-1 combinations of all rows) see every row and are robust to the handful of
        # argmax ties; one whole sample and 64 codebook rows are compared entry by entry with a gate tied to the
        # reference's own fp32-vs-fp64 distance on those very rows.
        gtol = 5e-5 if mode == "fp32" else BF16_TOL     # free-running golden: 10x the reference's own fp32 noise floor
        ptol = 1e-4 if mode == "fp32" else BF16_TOL
        d_sd = m.space_dict.grad
        for name, g_ in (("patch", patch.grad), ("word", word.grad), ("space_dict", d_sd)):
            ours = O.project(case["proj"][name], g_)
            p32, p64 = T("f32_proj_d_" + name).double(), T("f64_proj_d_" + name).double()
            # how far the REFERENCE'S OWN fp32 run sits from its fp64 run on this very projection: ~1e-6 on the bf16-exact
            # inputs, ~1e-3 on the unrounded ones, where a single argmax tie moves one gradient row (see module docstring)
            ref_floor = ((p32 - p64).abs().max() / p64.abs().max()).item()
            gate = max(ptol, 3 * ref_floor)
            if variant == "qmap":
                # unrounded inputs: the golden cannot be evaluated on our piece, and each of the (<= 8, asserted above)
                # rounding-level membership / argmax bits that differ moves a projection by ~2e-3 (one row of ~3000 with
                # random signs) -- the reference's own fp32 run shows 7.8e-4 .. 1.4e-3 here from ONE bit. Strict gradient
                # parity is asserted on the bf16-exact variant and in the same-piece tests below.
                gate = max(gate, BF16_TOL)
            report["f64_proj_" + name] = assert_close(ours, p64, gate, "proj d_%s vs the fp64 run (reference fp32-vs-fp64: %.1e)" % (name, ref_floor))
            report["f32_proj_" + name] = assert_close(ours, p32, gate + ref_floor, "proj d_%s vs the fp32 run" % name)
            report["ref_floor_proj_" + name] = ref_floor
        for name, g_ in (("patch", patch.grad[:1]), ("word", word.grad[:1]), ("space_dict", d_sd[:64])):
            r32, r64 = T("f32_d_%s_rows" % name).double(), T("f64_d_%s_rows" % name).double()
            ref_floor = ((r32 - r64).abs().max() / r64.abs().max()).item()
            gate = max(gtol, 10 * ref_floor)
            if variant == "qmap":
                gate = max(gate, BF16_TOL)
            report["rows_" + name] = assert_close(g_, r64, gate, "d_%s rows (reference fp32-vs-fp64 on them: %.1e)" % (name, ref_floor))
        report["d_logit_scale"] = assert_close(m.logit_scale.grad, T("f64_d_logit_scale"), 1e-4 if mode == "fp32" else tol,
                                               "d logit_scale")
        if variant == "qmap":
            for k, v in m.named_parameters():
                if ".q_map." in k:
                    ref = T("f64_grad_" + k)
                    got = v.grad if v.dim() == 1 else v.grad[:8]
                    assert_close(got, ref, BF16_TOL if mode == "fp32" else 5e-2, "grad " + k)   # free-running golden, see above
    else:
        # bf16 mode on UNROUNDED queries: q is rounded to bf16 inside, argmax tokens / supports legitimately move
        # (SURVEY H2), so per-row gradients are not comparable; loss, logits and embeddings above are
        report["d_logit_scale"] = assert_close(m.logit_scale.grad, T("f64_d_logit_scale"), 5e-2, "d logit_scale")
    print("config1[%s,%s] %s" % (variant, mode, json.dumps({k: (round(v, 9) if isinstance(v, float) else v)
                                                          for k, v in report.items()})))


# ---------------------------------------------------------------------------------------------------
# (i) whole step at BASELINE.json config 1 / 3 / 3b shapes vs the fp64 oracle run on the GPU
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("b,K", [(64, 16384), (256, 16384), (256, 4096)])
@pytest.mark.parametrize("mode", ["fp32", "bf16"])
def test_whole_step_vs_fp64_oracle(ops, b, K, mode):
    """GroundingFunction x2 -> L2-normalise -> InfoNCE (fused, gathered form) -> backward, P=49, T=77, D=512, through
    the same calls Clip_FDT.forward + ClipInfoCELoss make, against oracle.fdt_oracle.step in fp64 on the same device.
    bf16 mode: q / codebook pre-rounded to bf16 on both sides."""
    from ilfdt import functional as Fn
    dev = _dev()
    P, T, D, temp = 49, 77, 512, 1000.0
    inp = O.synth_inputs(b, P, T, K, D, seed=5)
    if mode == "bf16":
        for k in ("q_img", "q_txt", "sd"):
            inp[k] = bf16_round(inp[k])
    inp = {k: v.to(dev) for k, v in inp.items()}
    def oracle(force=None):
        ranks = [dict(q_img=O.make_leaf(inp["q_img"].double()), q_txt=O.make_leaf(inp["q_txt"].double()),
                      mask=inp["mask"].double(), sd=O.make_leaf(inp["sd"].double()), p=O.make_leaf(inp["p"].double()))]
        out = O.step(ranks, temperature=temp, force=force)[0]
        del ranks
        torch.cuda.empty_cache()
        return out

    ref = oracle()                                   # free-running fp64 rule: values and discrete choices
    qi = inp["q_img"].clone().requires_grad_(True)
    qt = inp["q_txt"].clone().requires_grad_(True)
    sd = inp["sd"].clone().requires_grad_(True)
    p = inp["p"].clone().requires_grad_(True)
    with Fn.codebook_scope():
        w_i, f_i = Fn.GroundingFunction.apply(qi, sd, None, temp, mode, True)
        w_t, f_t = Fn.GroundingFunction.apply(qt, sd, inp["mask"], temp, mode, True)
    x_i, x_t = Fn.L2NormFunction.apply(f_i, 1e-10), Fn.L2NormFunction.apply(f_t, 1e-10)
    s = p.exp()
    s.data = torch.clamp(s.data, max=100)
    loss = Fn.InfoNCEFunction.apply(x_i, x_t, x_i, x_t, s, 0, mode, True)
    loss.backward()
    tol = FP32_TOL if mode == "fp32" else BF16_TOL
    gtol = FP32_GRAD_TOL if mode == "fp32" else BF16_TOL
    e = {}
    e["loss"] = assert_close(loss, ref["loss"], tol, "loss")
    e["w_img"] = assert_close(w_i, ref["w_img"], tol, "att_weight img")
    e["w_txt"] = assert_close(w_t, ref["w_txt"], tol, "att_weight txt")
    e["f_img"] = assert_close(f_i, ref["f_img"], tol, "att_ft img")
    e["f_txt"] = assert_close(f_t, ref["f_txt"], tol, "att_ft txt")
    e["x_img"] = assert_close(x_i, ref["x_img"], tol, "normalised img")
    sc = s.detach().reshape(1).float().contiguous()
    logits = ops.logits(x_i.detach().contiguous(), x_t.detach().contiguous(), sc) if mode == "fp32" else \
        ops.logits(ops.cast_bf16(x_i.detach().contiguous()), ops.cast_bf16(x_t.detach().contiguous()), sc)
    e["logits"] = assert_close(logits, ref["logits_per_image"], tol, "logits_per_image")
    # (1) discrete choices agree with the free-running rule except at rounding-ambiguous entries
    arg_i, _ = _our_choices(ops, inp["q_img"], inp["sd"], None, temp, mode)
    arg_t, _ = _our_choices(ops, inp["q_txt"], inp["sd"], inp["mask"], temp, mode)
    q64i, q64t, sd64, m64 = inp["q_img"].double(), inp["q_txt"].double(), inp["sd"].double(), inp["mask"].double()
    e["ambiguous_img"] = _check_choices(q64i, sd64, None, temp, arg_i, w_i.detach(), ref["w_img"], "img")
    e["ambiguous_txt"] = _check_choices(q64t, sd64, m64, temp, arg_t, w_t.detach(), ref["w_txt"], "txt")
    # (2) gradients against the fp64 oracle evaluated on the same piece: every row, nothing skipped
    ref = oracle([dict(arg_img=arg_i, arg_txt=arg_t, supp_img=w_i.detach() > 0, supp_txt=w_t.detach() > 0)])
    e["dq_img"] = assert_close(qi.grad, ref["d_q_img"], gtol, "d q_img")
    e["dq_txt"] = assert_close(qt.grad, ref["d_q_txt"], gtol, "d q_txt")
    e["dsd"] = assert_close(sd.grad, ref["d_sd"], gtol, "d space_dict")
    e["dp"] = assert_close(p.grad, ref["d_p"], FP32_TOL if mode == "fp32" else tol, "d logit_scale")
    # pad rows beyond the first pad never win the max: exactly zero gradient
    lens = (inp["mask"] == 0).sum(1)
    far = torch.arange(T, device=dev).view(1, -1) > lens.view(-1, 1)
    assert qt.grad[far].abs().max().item() == 0.0
    print("whole_step[b=%d,K=%d,%s] %s" % (b, K, mode, json.dumps({k: (float("%.3g" % v) if isinstance(v, float) else v)
                                                                  for k, v in e.items()})))


# ---------------------------------------------------------------------------------------------------
# (ii) the ranged codebook-gradient sweep (b > 2047) and the per-sample scatter at b = 4096, K = 16384
# ---------------------------------------------------------------------------------------------------
def _scatter_fp64(coef, arg, q_rows64, row_of, sd64, n_rows_out, chunk=128):
    """SURVEY H3 restated with index_add in fp64: dq[row(b,arg)] += coef*sd[k]; dsd[k] += coef*q[row(b,arg)]."""
    b, K = coef.shape
    D = sd64.shape[1]
    dq = torch.zeros(n_rows_out, D, dtype=torch.float64, device=coef.device)
    dsd = torch.zeros(K, D, dtype=torch.float64, device=coef.device)
    for b0 in range(0, b, chunk):
        c = coef[b0:b0 + chunk]
        bi, ki = c.nonzero(as_tuple=True)
        cv = c[bi, ki].unsqueeze(1)
        ti = arg[b0 + bi, ki].long()
        out_row, in_row = row_of(b0 + bi, ti)
        dq.index_add_(0, out_row, cv * sd64[ki])
        dsd.index_add_(0, ki, cv * q_rows64[in_row])
    return dq, dsd


@pytest.mark.parametrize("n_tok,masked", [(49, False), (77, True)])
@pytest.mark.parametrize("lean", [False, True])
def test_scatter_backward_ranged_sweep_full_size(ops, n_tok, masked, lean):
    """ilfdt_ground_bwd / _w16 at b=4096, K=16384, D=512 (the per-rank headline shape; b > 2047 selects the ranged dsd
    sweep + red.global.add.v4.f32 combine) against the fp64 index_add restatement of the formula, on the operands the
    forward kernels produced; and the ranged sweep against the single-sweep, order-deterministic route."""
    dev = _dev()
    b, K, D, temp = 4096, 16384, 512, 1000.0
    g_ = torch.Generator(device="cuda").manual_seed(n_tok)
    q = torch.randn(b, n_tok, D, device=dev, generator=g_)
    sd = torch.randn(K, D, device=dev, generator=g_)
    sd_k = ops.cast_bf16(sd)
    mask = None
    if masked:
        lens = torch.randint(5, n_tok + 1, (b,), device=dev, generator=g_)
        mask = torch.zeros(b, n_tok, device=dev)
        mask[torch.arange(n_tok, device=dev).view(1, -1) >= lens.view(-1, 1)] = float("-inf")
        q_k, plan = ops.text_compact(q, mask)
        row_base = plan.row_base
        pooled, arg = ops.ground_fwd_text(q_k, sd_k, plan, n_tok, temp)
    else:
        q_k = ops.cast_bf16_rows(q, None)
        row_base = None
        pooled, arg = ops.ground_fwd(q_k, sd_k, None, temp, "max")
    w, w_b, _, nnz = ops.sparsemax_fwd(pooled, want_bf16=True)
    g = torch.randn(b, K, device=dev, generator=g_)
    dq_dtype = torch.bfloat16 if lean else torch.float32
    w_arg = w_b if lean else w

    def run(**tune):
        dsd = torch.zeros(K, D, device=dev)
        with ops.tune(**tune):
            dq = ops.ground_bwd(g.clone(), w_arg, arg, q_k, sd_k, mask, temp, dsd, dq_dtype, row_base, n_tok)
        return dq, dsd

    dq, dsd = run()
    dq1, dsd1 = run(dsd_split_rows=0)
    # fp64 restatement
    nz = (w > 0)
    mean = (g.double() * nz).sum(1, keepdim=True) / nz.sum(1, keepdim=True).clamp_min(1)
    keep = torch.ones(b, K, dtype=torch.float64, device=dev)
    if masked:
        keep = (torch.gather(mask, 1, arg.long()) == 0).double()
    coef = nz * (g.double() - mean) * keep / (math.sqrt(D) * temp)
    q_rows64 = q_k.double().reshape(-1, D)
    if masked:
        rb = row_base.long()
        row_of = lambda bi, ti: (bi * n_tok + ti, rb[bi] + ti)
    else:
        row_of = lambda bi, ti: (bi * n_tok + ti, bi * n_tok + ti)
    dq64, dsd64 = _scatter_fp64(coef, arg, q_rows64, row_of, sd_k.double(), b * n_tok)
    tol_dq = 8e-3 if lean else 2e-5        # bf16 storage of dq: one rounding (2^-8 relative per element)
    e = dict(dq=assert_close(dq.reshape(-1, D), dq64, tol_dq, "dq"),
             dsd=assert_close(dsd, dsd64, 2e-5, "dsd (ranged sweep)"),
             dsd_single=assert_close(dsd1, dsd64, 2e-5, "dsd (single sweep)"),
             ranged_vs_single=assert_close(dsd, dsd1, 5e-6, "dsd ranged vs single sweep"))
    assert torch.equal(dq, dq1), "dq must not depend on the dsd sweep plan"
    dq2, dsd2 = run(dsd_split_rows=0)
    assert torch.equal(dsd1, dsd2), "single-sweep dsd is order-deterministic"
    for pipe in (0, 2, 3):     # the v3 kernel, groups of 4 rows, scalar instead of packed (FFMA2) FMAs: same bits
        _, dsd_p = run(dsd_split_rows=0, dsd_pipe=pipe)
        assert torch.equal(dsd_p, dsd1), "dsd differs between gather pipelines (ILFDT_TUNE_DSD_PIPE=%d)" % pipe
    for pipe in (0, 1, 2):
        dq_p, _ = run(dq_pipe=pipe)
        assert torch.equal(dq_p, dq), "dq differs between gather pipelines (ILFDT_TUNE_DQ_PIPE=%d)" % pipe
    if masked:
        far = torch.arange(n_tok, device=dev).view(1, -1) > lens.view(-1, 1)
        assert dq[far].abs().max().item() == 0.0
    print("scatter[n_tok=%d,lean=%s] nnz/row %.0f %s" % (n_tok, lean, nnz.float().mean().item(),
                                                        json.dumps({k: float("%.3g" % v) for k, v in e.items()})))


@pytest.mark.parametrize("n_tok,masked,K", [(49, False, 16384), (77, True, 16384), (49, False, 4096)])
@pytest.mark.parametrize("lean", [False, True])
def test_sparse_regime_backward(ops, n_tok, masked, K, lean):
    """Trained-like, peaked attention (support of tens of codes): ilfdt_ground_bwd with the forward's nnz[b] serves the
    launch with the per-sample sparse kernel (decided on the device). Checked against the fp64 index_add restatement,
    against the dense sweeps (ILFDT_TUNE_SPARSE_BWD_NNZ = 0), and -- the same call on dense random-init statistics --
    that the device-side decision falls back to the dense sweeps."""
    dev = _dev()
    b, D, temp = 1536, 512, 1000.0
    g_ = torch.Generator(device="cuda").manual_seed(n_tok + K)
    sd = torch.randn(K, D, device=dev, generator=g_)
    sd_k = ops.cast_bf16(sd)
    for scale, expect_sparse in ((200.0, True), (1.0, False)):
        q = torch.randn(b, n_tok, D, device=dev, generator=g_) * scale
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
        w, w_b, _, nnz = ops.sparsemax_fwd(pooled, want_bf16=True)
        mean_nnz = nnz.float().mean().item()
        assert (mean_nnz <= 128) == expect_sparse, mean_nnz
        g = torch.randn(b, K, device=dev, generator=g_)
        dq_dtype = torch.bfloat16 if lean else torch.float32
        w_arg = w_b if lean else w

        def run(use_nnz, **tune):
            dsd = torch.zeros(K, D, device=dev)
            gg = g.clone()
            with ops.tune(**tune):
                dq = ops.ground_bwd(gg, w_arg, arg, q_k, sd_k, mask, temp, dsd, dq_dtype, row_base, n_tok,
                                    nnz if use_nnz else None)
            return dq, dsd, gg

        dq_a, dsd_a, g_after = run(True)                       # device-side decision
        dq_d, dsd_d, _ = run(True, sparse_bwd_nnz=0)           # dense sweeps forced
        # the dense sweeps leave coef in g; the sparse kernel never writes it: that is how the taken route is observed
        took_sparse = torch.equal(g_after, g)
        assert took_sparse == expect_sparse
        nz = (w > 0)
        mean = (g.double() * nz).sum(1, keepdim=True) / nz.sum(1, keepdim=True).clamp_min(1)
        keep = torch.ones(b, K, dtype=torch.float64, device=dev)
        if masked:
            keep = (torch.gather(mask, 1, arg.long()) == 0).double()
        coef = nz * (g.double() - mean) * keep / (math.sqrt(D) * temp)
        if masked:
            rb = row_base.long()
            row_of = lambda bi, ti: (bi * n_tok + ti, rb[bi] + ti)
        else:
            row_of = lambda bi, ti: (bi * n_tok + ti, bi * n_tok + ti)
        if expect_sparse:
            dq64, dsd64 = _scatter_fp64(coef, arg, q_k.double().reshape(-1, D), row_of, sd_k.double(), b * n_tok, chunk=b)
            tol_dq = 8e-3 if lean else 2e-5
            e = dict(dq=assert_close(dq_a.reshape(-1, D), dq64, tol_dq, "dq (sparse kernel)"),
                     dsd=assert_close(dsd_a, dsd64, 2e-5, "dsd (sparse kernel)"),
                     dq_vs_dense=assert_close(dq_a.float(), dq_d.float(), tol_dq, "dq sparse vs dense sweeps"),
                     dsd_vs_dense=assert_close(dsd_a, dsd_d, 2e-5, "dsd sparse vs dense sweeps"))
            print("sparse_bwd[n_tok=%d,K=%d,lean=%s] mean support %.1f %s" % (n_tok, K, lean, mean_nnz,
                                                                            json.dumps({k: float("%.3g" % v) for k, v in e.items()})))
        else:
            assert torch.equal(dq_a, dq_d), "dense statistics: the decision must fall back to the dense sweeps"
            assert_close(dsd_a, dsd_d, 5e-6, "dsd")


# ---------------------------------------------------------------------------------------------------
# (iii) InfoNCE at b = 4096, B = 32768, rank 3 of 8: chunked LSE + combine, nce_bwd2 split plan, top-k counts
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("mode", ["fp32", "bf16"])
def test_infonce_rank_of_eight_at_32k(ops, mode):
    """The per-rank InfoNCE of the 8-GPU headline (b=4096 local rows against B=32768 gathered rows, label offset
    3*4096): loss, row log-sum-exps, the gathered-form gradient of the local rows, d scale and the top-1 / top-5 counts
    against fp64 torch on the full [B,B] logits (8.6 GB each; the 8 ranks are emulated on one GPU)."""
    dev = _dev()
    b, W, D = 4096, 8, 512
    B, r = b * W, 3
    g_ = torch.Generator(device="cuda").manual_seed(32)
    # correlated pairs so that top-1 / top-5 are neither 0 nor b
    base = torch.randn(B, D, device=dev, generator=g_)
    # pair cosine a^2/(1+a^2) = 0.17 against the best of 32767 random negatives (~4 sigma of 1/sqrt(512) = 0.18)
    xi = torch.nn.functional.normalize(0.45 * base + torch.randn(B, D, device=dev, generator=g_), dim=1)
    xt = torch.nn.functional.normalize(0.45 * base + torch.randn(B, D, device=dev, generator=g_), dim=1)
    if mode == "bf16":
        xi, xt = bf16_round(xi), bf16_round(xt)
    pval = math.log(1 / 0.07)
    scale = torch.tensor([math.exp(pval)], device=dev)
    # ---- fp64: every rank's loss; total = sum_r loss_r (upstream 1 on every rank)
    Xi, Xt = xi.double().requires_grad_(True), xt.double().requires_grad_(True)
    s64 = scale.double().requires_grad_(True)
    L = (Xi @ Xt.t()) * s64                                           # logits_per_image of all ranks stacked [B,B]
    lse_row = torch.logsumexp(L, 1)
    lse_col = torch.logsumexp(L, 0)                                   # rows of logits_per_text
    diag = L.diagonal()
    per_row = ((lse_row - diag) + (lse_col - diag)) / (2 * b)         # each rank: mean over its b rows of (CE_i + CE_t)/2
    loss_r = per_row[r * b:(r + 1) * b].sum()
    ds_local, = torch.autograd.grad(loss_r, s64, retain_graph=True)
    dXi, dXt = torch.autograd.grad(per_row.sum(), (Xi, Xt))
    rank_of_label = (L[r * b:(r + 1) * b].detach() > diag[r * b:(r + 1) * b].detach().unsqueeze(1)).sum(1)
    top1_ref, top5_ref = int((rank_of_label < 1).sum()), int((rank_of_label < 5).sum())
    lse_row, lse_col = lse_row.detach(), lse_col.detach()
    loss_r = loss_r.detach()
    del L, per_row, diag
    torch.cuda.empty_cache()
    # ---- ours: nce_fwd on every emulated rank (that is what produces the all-gathered lse vectors), backward on rank 3
    cast = (lambda t: ops.cast_bf16(t.contiguous())) if mode == "bf16" else (lambda t: t.contiguous())
    gi, gt = cast(xi), cast(xt)
    lse_i_all = torch.empty(B, device=dev)
    lse_t_all = torch.empty(B, device=dev)
    keep = None
    for rr in range(W):
        out = ops.nce_fwd(gi[rr * b:(rr + 1) * b], gt[rr * b:(rr + 1) * b], gi, gt, scale, rr * b)
        lse_i_all[rr * b:(rr + 1) * b], lse_t_all[rr * b:(rr + 1) * b] = out[1], out[2]
        if rr == r:
            keep = out
    loss, _, _, diag_i, _ = keep
    go_all = torch.ones(W, device=dev)
    dx_i, dx_t, dscale = ops.nce_bwd_gathered(gi[r * b:(r + 1) * b], gt[r * b:(r + 1) * b], gi, gt, scale, r * b,
                                             lse_i_all, lse_t_all, go_all)
    counts = ops.nce_topk_hits(gi[r * b:(r + 1) * b], gt, scale, r * b, diag_i)
    stored = ops.logits(gi[r * b:(r + 1) * b], gt, scale)
    counts_stored = ops.topk_hits(stored, r * b)
    tol = FP32_TOL if mode == "fp32" else BF16_TOL
    e = dict(loss=assert_close(loss, loss_r, tol, "loss of rank 3"),
             lse_i=assert_close(lse_i_all, lse_row, tol, "row lse, logits_per_image, all ranks"),
             lse_t=assert_close(lse_t_all, lse_col, tol, "row lse, logits_per_text, all ranks"),
             logits=assert_close(stored, (xi[r * b:(r + 1) * b].double() @ xt.double().t()) * scale.double(), tol, "logits"),
             dx_img=assert_close(dx_i, dXi[r * b:(r + 1) * b], 2e-5 if mode == "fp32" else tol, "dx_img (gathered form)"),
             dx_txt=assert_close(dx_t, dXt[r * b:(r + 1) * b], 2e-5 if mode == "fp32" else tol, "dx_txt (gathered form)"),
             dscale=assert_close(dscale, ds_local, FP32_TOL if mode == "fp32" else tol, "d scale (rank-local)"))
    c = counts.cpu().tolist()
    assert c == counts_stored.cpu().tolist(), "fused top-k counts != counts on the stored logits"
    assert 0 < c[0] < b and c[0] <= c[1]
    # against fp64 ranks: entries within rounding of the label logit may fall on either side
    assert abs(c[0] - top1_ref) <= 2 and abs(c[1] - top5_ref) <= 2, (c, top1_ref, top5_ref)
    print("infonce32k[%s] top1/top5 %s (fp64 %d/%d) %s" % (mode, c, top1_ref, top5_ref,
                                                          json.dumps({k: float("%.3g" % v) for k, v in e.items()})))


# ---------------------------------------------------------------------------------------------------
# (iv) BASELINE.json config 5: ViT-B/16 grounding (P = 196) at b = 1024, K = 16384 after a codebook reset
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("need_weights", [True, False])
def test_config5_grounding_vs_oracle_in_row_chunks(ops, need_weights):
    """Generation step = codebook reset (clip_fdt.py:281-283) + re-grounding: P=196 patches, b=1024, K=16384, bf16
    mode, forward + backward of the image Query_model against the fp64 oracle evaluated in chunks of 64 samples
    (SURVEY H8); need_weights=False is the route Clip_FDT.forward takes."""
    from ilfdt import functional as Fn
    dev = _dev()
    b, P, K, D, temp = 1024, 196, 16384, 512, 1000.0
    g_ = torch.Generator(device="cuda").manual_seed(196)
    q = bf16_round(torch.randn(b, P, D, device=dev, generator=g_))
    sd = torch.empty(K, D, device=dev)
    ops.randn_fill_(sd, 11, 2)                                       # the reset codebook (Philox, generation 2)
    sd = bf16_round(sd)
    up = torch.randn(b, D, device=dev, generator=g_)
    qd = q.clone().requires_grad_(True)
    sdd = sd.clone().requires_grad_(True)
    w, f = Fn.GroundingFunction.apply(qd, sdd, None, temp, "bf16", need_weights)
    (f * up).sum().backward()
    arg, _ = _our_choices(ops, q, sd, None, temp, "bf16")
    dsd64 = torch.zeros(K, D, dtype=torch.float64, device=dev)
    worst = dict(w=0.0, f=0.0, dq=0.0)
    n_arg = n_sup = 0
    sd64 = sd.double()
    for b0 in range(0, b, 64):
        sl = slice(b0, b0 + 64)
        with torch.no_grad():
            w_ref, f_ref, _ = O.grounding(q[sl].double(), sd64, None, temp)          # free-running rule
        worst["f"] = max(worst["f"], rel_err(f[sl], f_ref))
        worst["w"] = max(worst["w"], rel_err(w[sl].float(), w_ref))
        a_, s_ = _check_choices(q[sl].double(), sd64, None, temp, arg[sl], w[sl].float(), w_ref, "chunk %d" % b0,
                                max_ambiguous=16)
        n_arg += a_
        n_sup += s_
        qo = q[sl].double().requires_grad_(True)
        so = sd64.clone().requires_grad_(True)
        _, f_same, _ = O.grounding(qo, so, None, temp, force_arg=arg[sl], force_support=w[sl] > 0)   # same piece
        (f_same * up[sl].double()).sum().backward()
        dsd64 += so.grad
        worst["dq"] = max(worst["dq"], assert_close(qd.grad[sl], qo.grad, BF16_TOL, "dq chunk %d" % b0))
        del qo, so, w_ref, f_ref, f_same
    assert worst["w"] <= BF16_TOL and worst["f"] <= BF16_TOL
    e_dsd = assert_close(sdd.grad, dsd64, BF16_TOL, "d space_dict")
    if not need_weights:
        assert w.dtype == torch.bfloat16
    print("config5[need_weights=%s] w %.2e f %.2e dq %.2e dsd %.2e; ambiguous: %d argmax, %d support bits of %d" % (
        need_weights, worst["w"], worst["f"], worst["dq"], e_dsd, n_arg, n_sup, b * K))


# ---------------------------------------------------------------------------------------------------
# the bf16 codebook copy must follow in-place writes that do not bump `_version`
# ---------------------------------------------------------------------------------------------------
def test_forward_after_inplace_codebook_reset_uses_new_values(ops):
    from ilfdt.loss_functions import ClipInfoCELoss
    from ilfdt.model.clip_fdt import Clip_FDT
    dev = _dev()
    b, P, T, K, D = 24, 49, 77, 1024, 512
    inp = O.synth_inputs(b, P, T, K, D, seed=9)
    patch, word, mask = inp["q_img"].to(dev), inp["q_txt"].to(dev), inp["mask"].to(dev)

    def make():
        m = Clip_FDT(_FakeVisual(patch), _FakeText(word, mask), True, K, D, D, D, "sparsemax", "max", 1000.0).to(dev)
        m.img_query_model.q_map = torch.nn.Identity()
        m.txt_query_model.q_map = torch.nn.Identity()
        return m

    def run(m):
        m.zero_grad(set_to_none=True)
        (l_i, l_t), _ = m(torch.zeros(b, 1, device=dev), ["x"] * b)
        loss, _ = ClipInfoCELoss()(l_i, l_t)
        loss.backward()
        return loss.detach().clone(), m.space_dict.grad.clone()

    m = make()
    with torch.no_grad():
        m.space_dict.copy_(inp["sd"].to(dev))
    m.eval()
    l0, _ = run(m)                                   # a forward that would have populated a version-keyed cache
    m.store_codebook_value()
    m.reset_codebook(seed=5)                         # raw-pointer write: _version unchanged
    l1, g1 = run(m)
    fresh = make()
    fresh.space_dict = torch.nn.Parameter(m.space_dict.detach().clone())    # the same values in a fresh tensor
    l2, g2 = run(fresh)
    assert l0.item() != l1.item()
    assert torch.equal(l1, l2), "forward after an in-place reset must see the new codebook"
    assert_close(g1, g2, 1e-5, "d space_dict after an in-place reset vs a fresh tensor with the same values")
    m.keep_codebook_value()                          # ops.copy_: also invisible to _version
    l3, _ = run(m)
    assert torch.equal(l3, l0)


def test_graph_replay_follows_inplace_codebook_updates(ops):
    """The cast of the codebook is captured INTO the graph: a replay after an optimizer-style in-place update of the
    parameter equals an eager step on the updated values."""
    from ilfdt import functional as Fn
    from ilfdt.graphs import GraphedStep
    dev = _dev()
    b, P, T, K, D = 32, 49, 77, 1024, 512
    inp = {k: v.to(dev) for k, v in O.synth_inputs(b, P, T, K, D, seed=4).items()}
    sd = torch.nn.Parameter(inp["sd"].clone())
    p = torch.nn.Parameter(inp["p"].clone())

    def step(qi, qt, mk):
        qi = qi.detach().requires_grad_(True)
        qt = qt.detach().requires_grad_(True)
        sd.grad = None
        p.grad = None
        with Fn.codebook_scope():
            _, f_i = Fn.GroundingFunction.apply(qi, sd, None, 1000.0, "bf16", False)
            _, f_t = Fn.GroundingFunction.apply(qt, sd, mk, 1000.0, "bf16", False)
        x_i, x_t = Fn.L2NormFunction.apply(f_i, 1e-10), Fn.L2NormFunction.apply(f_t, 1e-10)
        s = p.exp()
        s.data = torch.clamp(s.data, max=100)
        loss = Fn.InfoNCEFunction.apply(x_i, x_t, x_i, x_t, s, 0, "bf16", True)
        loss.backward()
        return loss

    gstep = GraphedStep(step, (inp["q_img"], inp["q_txt"], inp["mask"]), warmup=2)
    l_a = gstep().clone()
    with torch.no_grad():
        sd.add_(0.5 * torch.randn_like(sd))          # what an optimizer does
    l_b = gstep().clone()
    g_b = sd.grad.clone()
    l_e = step(inp["q_img"], inp["q_txt"], inp["mask"]).detach()
    assert l_a.item() != l_b.item(), "replay ignored the updated codebook"
    assert_close(l_b, l_e, 1e-6, "graph replay vs eager after an in-place codebook update")
    assert_close(g_b, sd.grad, 2e-3, "d space_dict, replay vs eager")
    ops.randn_fill_(sd.data, 3, 1)                   # reset_codebook()'s raw write
    l_c = gstep().clone()
    l_e2 = step(inp["q_img"], inp["q_txt"], inp["mask"]).detach()
    assert_close(l_c, l_e2, 1e-6, "graph replay vs eager after reset_codebook")
    del gstep
