"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's FDT-grounding + InfoNCE hot path.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs may
import this file, and only as the checker / the timed CPU baseline. The product path
(`iterated-learning-for-vlm_b200/ilfdt`) never imports it and has no CPU fallback.

The reference is pure PyTorch, so the restatement uses the same PyTorch CPU
operators (`matmul`, `sort`, `cumsum`, `max`, `cross_entropy`); every function cites the reference
lines it follows (paths relative to /root/reference). It runs in whatever dtype its inputs have,
so the fp64 arbiter run is the same code fed `.double()` tensors.

Parity pin: the reference ships no tests for this path, so the oracle is pinned
against OUTPUTS OF THE REFERENCE ITSELF: `oracle/make_golden.py` imports the unmodified reference
(`oracle/ref_harness.py`) in the build container and writes `tests/golden/*.npz`;
`tests/test_oracle_pinned.py` checks this file against those vectors (and, when /root/reference is
present, against the live reference).
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

# --------------------------------------------------------------------------------------------
# a2  Sparsemax            prototype/model/sparsemax.py:22-71
# --------------------------------------------------------------------------------------------


def sparsemax(x: torch.Tensor, dim: int = -1, force_support=None) -> torch.Tensor:
    """Simplex projection along `dim` by the sorted-prefix rule of sparsemax.py:40-63.

    Layout juggling of sparsemax.py:31-35,66-69 (swap dim with 0, flatten the rest, transpose so the
    logits are the trailing axis) is restated with movedim; the arithmetic per row is identical.

    force_support (TEST HOOK, bool, same shape as x with `dim` last): evaluate the projection ON A GIVEN
    SUPPORT S -- tau = (sum_S z - 1)/|S|, out = (z - tau) on S, 0 elsewhere -- which is what the sorted
    rule computes once S is known. The projection is continuous in z but its JACOBIAN is not: an entry that
    sits at the threshold to within rounding carries weight ~0 and a full-size gradient, so two correct
    evaluations that disagree on that one membership bit disagree on the gradient by O(1). Gradient parity is
    therefore stated on the same piece: the tests pass the support (and argmax tokens) the implementation
    under test chose, after checking that those choices differ from the free-running rule only at
    rounding-ambiguous entries.
    """
    z = x.movedim(dim, -1)
    lead_shape = z.shape
    z = z.reshape(-1, z.shape[-1])
    n = z.shape[-1]
    z = z - z.max(dim=1, keepdim=True)[0]                       # sparsemax.py:40
    if force_support is not None:
        nz = force_support.reshape(z.shape).to(z.dtype)
        tau = ((z * nz).sum(dim=1, keepdim=True) - 1) / nz.sum(dim=1, keepdim=True)
        return ((z - tau) * nz).reshape(lead_shape).movedim(-1, dim)
    zs = torch.sort(z, dim=1, descending=True)[0]               # :45
    ks = torch.arange(1, n + 1, dtype=z.dtype, device=z.device).view(1, -1)      # :46-47
    bound = 1 + ks * zs                                         # :50
    csum = torch.cumsum(zs, dim=1)                              # :51
    is_gt = (bound > csum).to(z.dtype)                          # :52
    k = (is_gt * ks).max(dim=1, keepdim=True)[0]                # :53
    tau = ((is_gt * zs).sum(dim=1, keepdim=True) - 1) / k       # :56-59
    out = torch.clamp_min(z - tau, 0)                           # :63  max(0, z - tau)
    return out.reshape(lead_shape).movedim(-1, dim)


def sparsemax_tau_exact(z_row: np.ndarray):
    """fp64 arbiter for one row: returns (tau, support_mask, margin=|z-tau| of the closest element).

    Same rule as sparsemax.py:50-59 evaluated in float64 on the (already fp32) scores; used by the
    tests to decide which support-set disagreements are rounding-ambiguous.
    """
    z = z_row.astype(np.float64)
    z = z - z.max()
    zs = np.sort(z)[::-1]
    ks = np.arange(1, z.size + 1, dtype=np.float64)
    cs = np.cumsum(zs)
    k = int(ks[(1 + ks * zs) > cs].max())
    tau = (cs[k - 1] - 1.0) / k
    return tau, (z > tau), float(np.abs(z - tau).min())


# --------------------------------------------------------------------------------------------
# a1 + a3  Query_model.forward from q onward      prototype/model/clip_fdt.py:113-157
# --------------------------------------------------------------------------------------------


def pooled_scores(q, sd, mask=None, temperature=1.0, pool_type="max", return_token_scores=False, force_arg=None):
    """q [b,N,D], sd [K,D], mask [b,N] (0 = real token, -inf = pad) -> pooled [b,K].

    Operation ORDER is the reference's (clip_fdt.py:118 matmul, :123 /sqrt(D), :125-130 multiply by
    the 0/1 mask, :137 /temperature, :140-145 pool) -- not a single fused scale.
    """
    s = torch.matmul(q, sd.t().unsqueeze(0))                    # :113-118
    token_att = s
    s = s / math.sqrt(sd.shape[1])                              # :123 (att_dim == sd_dim)
    if mask is not None:
        keep = (mask == 0) * 1                                  # :128  0 -> 1, -inf -> 0
        s = s * keep.unsqueeze(-1)                              # :130  pads contribute score 0 (F6)
        token_att = s
    s = s / temperature                                         # :137
    if pool_type == "sum":
        s = s.sum(1)
    elif pool_type == "mean":
        s = s.mean(1)
    elif pool_type == "max" and force_arg is not None:          # TEST HOOK: the max taken at GIVEN tokens (see sparsemax)
        s = torch.gather(s, 1, force_arg.long().unsqueeze(1)).squeeze(1)
    elif pool_type == "max":
        s = s.max(1)[0]                                         # :145
    else:
        raise ValueError(pool_type)
    if return_token_scores:
        return s, token_att
    return s


def attention_weights(pooled, att_func_type="sparsemax", force_support=None):
    if att_func_type == "sparsemax":
        return sparsemax(pooled, dim=-1, force_support=force_support)   # clip_fdt.py:74,149
    if att_func_type == "softmax":
        return torch.softmax(pooled, dim=-1)                    # :72
    if att_func_type == "sigmoid":
        return torch.sigmoid(pooled)                            # :76
    raise ValueError(att_func_type)


def grounding(q, sd, mask=None, temperature=1.0, att_func_type="sparsemax", pool_type="max", force_arg=None,
              force_support=None):
    """Returns (att_weight [b,K], att_ft [b,D], pooled [b,K])   clip_fdt.py:113-161.
    force_arg / force_support: test hooks, see sparsemax()."""
    pooled = pooled_scores(q, sd, mask, temperature, pool_type, force_arg=force_arg)
    w = attention_weights(pooled, att_func_type, force_support=force_support)
    f = w @ sd                                                  # :154
    if att_func_type == "sigmoid":
        f = f / w.sum(dim=-1, keepdim=True)                     # :156-157
    return w, f, pooled


# --------------------------------------------------------------------------------------------
# a5 a6 a8 a10  contrastive head + ClipInfoCELoss
# --------------------------------------------------------------------------------------------


def l2_normalise(f, eps=1e-10):
    """clip_fdt.py:411-412 (eps 1e-10 on both); clip.py:133 uses eps = 0 for the image side (F10)."""
    return f / (f.norm(dim=-1, keepdim=True) + eps)


def logit_scale_forward(p, max_scale=100.0):
    """clip_fdt.py:415-416: forward value min(exp(p), 100); backward multiplies by UNCLAMPED exp(p) (F7)."""
    s = p.exp()
    s.data = torch.clamp(s.data, max=max_scale)
    return s


def fdt_logits(x_img, x_txt, g_img, g_txt, s):
    """clip_fdt.py:421-422: the PRODUCT is scaled."""
    return x_img @ g_txt.t() * s, x_txt @ g_img.t() * s


def clip_logits(x_img, x_txt, g_img, g_txt, s):
    """clip.py:146-147: the FEATURES are scaled before the matmul (F10)."""
    return s * x_img @ g_txt.t(), s * x_txt @ g_img.t()


def info_nce(logits_per_image, logits_per_text, rank=0):
    """loss.py:40-50. labels = arange(b) when the logits are square, else rank*b + arange(b)."""
    bs, l_bs = logits_per_image.shape
    if l_bs == bs:
        labels = torch.arange(bs, device=logits_per_image.device)
    else:
        labels = rank * bs + torch.arange(0, bs, dtype=torch.long, device=logits_per_image.device)
    loss_i = F.cross_entropy(logits_per_image, labels)
    loss_t = F.cross_entropy(logits_per_text, labels)
    return (loss_i + loss_t) / 2, labels


# --------------------------------------------------------------------------------------------
# whole step, single process emulation of W ranks (a7 AllGather semantics: clip_fdt.py:164-188)
# --------------------------------------------------------------------------------------------


def step(ranks, temperature=1000.0, head="fdt", upstream=None, eps_img=1e-10, eps_txt=1e-10,
         att_func_type="sparsemax", pool_type="max", force=None):
    """Forward + backward of the hot path for W emulated ranks in one process.

    `ranks` is a list (length W) of dicts with leaf tensors:
        FDT  head: q_img [b,P,D], q_txt [b,T,D], mask [b,T] or None, sd [K,D], p [1]
        CLIP head: f_img [b,D], f_txt [b,D], p [1]      (head="clip": no grounding, F10 numerics)
    Each rank gets its OWN leaf copies of sd and p, so that parameter gradients come out as the
    rank-local partials the reference produces (SURVEY.md section 8e "verified semantics").

    AllGather (clip_fdt.py:167-188): forward concatenates all ranks' rows rank-major; backward
    all-reduces (sums) the gradient w.r.t. every rank's gathered copy and returns the own slice.
    Summing the per-rank losses (each weighted by its upstream grad) in one autograd graph where the
    gathered tensor is a differentiable concatenation gives exactly that gradient.
    """
    W = len(ranks)
    upstream = upstream if upstream is not None else [1.0] * W
    inter = []
    xs_img, xs_txt = [], []
    for ri, r in enumerate(ranks):
        d = {}
        if head == "fdt":
            fc = force[ri] if force is not None else {}   # test hook: discrete choices given (see sparsemax())
            w_i, f_i, pooled_i = grounding(r["q_img"], r["sd"], None, temperature, att_func_type, pool_type,
                                           fc.get("arg_img"), fc.get("supp_img"))
            w_t, f_t, pooled_t = grounding(r["q_txt"], r["sd"], r.get("mask"), temperature, att_func_type, pool_type,
                                           fc.get("arg_txt"), fc.get("supp_txt"))
            d.update(w_img=w_i, w_txt=w_t, f_img=f_i, f_txt=f_t, pooled_img=pooled_i, pooled_txt=pooled_t)
        else:
            f_i, f_t = r["f_img"], r["f_txt"]
        x_i = l2_normalise(f_i, eps_img)
        x_t = l2_normalise(f_t, eps_txt)
        d.update(x_img=x_i, x_txt=x_t)
        xs_img.append(x_i)
        xs_txt.append(x_t)
        inter.append(d)
    g_img = torch.cat(xs_img, 0)
    g_txt = torch.cat(xs_txt, 0)
    total = 0
    for rk, (r, d) in enumerate(zip(ranks, inter)):
        s = logit_scale_forward(r["p"])
        if head == "fdt":
            l_i, l_t = fdt_logits(d["x_img"], d["x_txt"], g_img, g_txt, s)
        else:
            l_i, l_t = clip_logits(d["x_img"], d["x_txt"], g_img, g_txt, s)
        loss, labels = info_nce(l_i, l_t, rank=rk)
        d.update(logits_per_image=l_i, logits_per_text=l_t, loss=loss, labels=labels, scale=s)
        total = total + loss * upstream[rk]
    total.backward()
    out = []
    for r, d in zip(ranks, inter):
        o = {k: v.detach() for k, v in d.items()}
        for k, v in r.items():
            if isinstance(v, torch.Tensor) and v.requires_grad and v.is_leaf and v.grad is not None:
                o["d_" + k] = v.grad.detach()
        out.append(o)
    return out


def make_leaf(t, requires_grad=True):
    t = t.detach().clone()
    t.requires_grad_(requires_grad)
    return t


# --------------------------------------------------------------------------------------------
# synthetic inputs -- shared by tests, smoke() and bench.py so that CPU and
# GPU sides see identical bits. Pure torch CPU generator; no reference code involved.
# --------------------------------------------------------------------------------------------


def synth_inputs(b, P, T, K, D, seed=0, rank=0, min_len=5, dtype=torch.float32):
    g = torch.Generator().manual_seed(1000 * seed + 1000 + rank)
    q_img = torch.randn(b, P, D, generator=g, dtype=dtype)
    q_txt = torch.randn(b, T, D, generator=g, dtype=dtype)
    lens = torch.randint(min(min_len, T), T + 1, (b,), generator=g)
    mask = torch.zeros(b, T, dtype=dtype)
    mask[torch.arange(T).view(1, -1) >= lens.view(-1, 1)] = float("-inf")   # text_transformer.py:182-194
    gs = torch.Generator().manual_seed(seed)                                  # codebook identical on all ranks
    sd = torch.randn(K, D, generator=gs, dtype=dtype)                         # clip_fdt.py:215
    p = torch.full((1,), math.log(1 / 0.07), dtype=dtype)                     # clip_fdt.py:228
    return dict(q_img=q_img, q_txt=q_txt, mask=mask, sd=sd, p=p)


# --------------------------------------------------------------------------------------------
# a12  codebook reset RNG: the device fill is Philox4x32-10 + Box-Muller (design decision, see
# DESIGN.md "reset"); this is its integer-exact restatement. The reference draws from the host
# global generator (clip_fdt.py:281-283), which no device RNG can reproduce bit-for-bit; parity is
# distributional (N(0,1)) plus determinism / rank-identity, and bit-exact on the Philox words.
# --------------------------------------------------------------------------------------------

_PHILOX_M0 = np.uint64(0xD2511F53)
_PHILOX_M1 = np.uint64(0xCD9E8D57)
_PHILOX_W0 = np.uint32(0x9E3779B9)
_PHILOX_W1 = np.uint32(0xBB67AE85)


def philox4x32_10(counter: np.ndarray, key: np.ndarray) -> np.ndarray:
    """counter [n,4] uint32, key [2] uint32 -> [n,4] uint32 (Salmon et al. 2011, 10 rounds)."""
    c = counter.astype(np.uint32).copy()
    k0, k1 = np.uint32(key[0]), np.uint32(key[1])
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = _PHILOX_M0 * c[:, 0].astype(np.uint64)
            p1 = _PHILOX_M1 * c[:, 2].astype(np.uint64)
            hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), p0.astype(np.uint32)
            hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), p1.astype(np.uint32)
            c = np.stack([hi1 ^ c[:, 1] ^ k0, lo1, hi0 ^ c[:, 3] ^ k1, lo0], axis=1)
            k0 = np.uint32(k0 + _PHILOX_W0)
            k1 = np.uint32(k1 + _PHILOX_W1)
    return c


def philox_words(n_words: int, seed: int, generation: int) -> np.ndarray:
    """The uint32 stream the device kernel consumes: block i = philox(ctr=(i_lo,i_hi,generation,0), key=seed)."""
    nblk = (n_words + 3) // 4
    idx = np.arange(nblk, dtype=np.uint64)
    ctr = np.stack([(idx & np.uint64(0xFFFFFFFF)).astype(np.uint32), (idx >> np.uint64(32)).astype(np.uint32),
                    np.full(nblk, generation & 0xFFFFFFFF, np.uint32), np.zeros(nblk, np.uint32)], axis=1)
    key = np.array([seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF], dtype=np.uint32)
    return philox4x32_10(ctr, key).reshape(-1)[:n_words]


def philox_randn(n: int, seed: int, generation: int) -> np.ndarray:
    """N(0,1) floats: words (w0,w1,w2,w3) -> u = (w + 0.5) * 2^-32; Box-Muller pairs (u0,u1),(u2,u3)."""
    w = philox_words(4 * ((n + 3) // 4), seed, generation).reshape(-1, 4).astype(np.float64)
    u = (w + 0.5) * (1.0 / 4294967296.0)
    r0 = np.sqrt(-2.0 * np.log(u[:, 0]))
    r1 = np.sqrt(-2.0 * np.log(u[:, 2]))
    t0 = 2.0 * np.pi * u[:, 1]
    t1 = 2.0 * np.pi * u[:, 3]
    out = np.stack([r0 * np.cos(t0), r0 * np.sin(t0), r1 * np.cos(t1), r1 * np.sin(t1)], axis=1)
    return out.reshape(-1)[:n].astype(np.float32)


# --------------------------------------------------------------------------------------------
# a0  q_map (clip_fdt.py:86-92): LN -> Linear -> GELU(erf) -> LN -> Linear. Produces the kernels'
# input q; restated here so the pinned tests can rebuild it from the golden state dicts.
# --------------------------------------------------------------------------------------------


def make_q_map(ft_dim, sd_dim, state=None, prefix="q_map."):
    import torch.nn as nn
    m = nn.Sequential(nn.LayerNorm(ft_dim), nn.Linear(ft_dim, sd_dim), nn.GELU(), nn.LayerNorm(sd_dim),
                      nn.Linear(sd_dim, sd_dim))
    if state is not None:
        m.load_state_dict({k[len(prefix):]: torch.as_tensor(v) for k, v in state.items() if k.startswith(prefix)})
    return m


# --------------------------------------------------------------------------------------------
# BASELINE.json config 1 (b=64, P=49, T=77, K=16384, D=512): inputs, parameters and gradient
# projections REGENERATED FROM SEEDS on both sides -- oracle/make_golden.py feeds them to the
# unmodified reference and stores only small outputs (tests/golden/config1_*.npz); the tests feed
# the same tensors to the CUDA path. `fingerprint` of every regenerated tensor is stored in the
# golden file so a drifting generator fails loudly instead of comparing different inputs.
# --------------------------------------------------------------------------------------------

CFG1 = dict(b=64, P=49, T=77, K=16384, D=512, Fi=768, Ft=512, temperature=1000.0, n_proj=16)


def config1_case(variant, b=None, K=None):
    """variant "qmap": raw features + real q_map MLPs (clip_fdt.py:86-92), fp32 values.
    variant "bf16in": q_map = identity, queries and codebook pre-rounded to bf16 so that the bf16
    tensor-core mode and the reference see bit-identical operands."""
    c = dict(CFG1)
    if b is not None:
        c["b"] = b
    if K is not None:
        c["K"] = K
    b, P, T, K, D = c["b"], c["P"], c["T"], c["K"], c["D"]
    Fi, Ft = (c["Fi"], c["Ft"]) if variant == "qmap" else (D, D)
    g = torch.Generator().manual_seed(20261105 + (0 if variant == "qmap" else 1))
    patch = torch.randn(b, P, Fi, generator=g)
    word = torch.randn(b, T, Ft, generator=g)
    lens = torch.randint(5, T + 1, (b,), generator=g)
    lens[0] = T                                                   # one caption without padding
    mask = torch.zeros(b, T)
    mask[torch.arange(T).view(1, -1) >= lens.view(-1, 1)] = float("-inf")
    state = {"space_dict": torch.randn(K, D, generator=g),
             "logit_scale": torch.full((1,), math.log(1 / 0.07)),
             "logit_scale_sd": torch.full((1,), math.log(1 / 0.07))}
    if variant == "qmap":
        for name, F_in in (("img_query_model", Fi), ("txt_query_model", Ft)):
            pre = name + ".q_map."
            state[pre + "0.weight"] = 1 + 0.1 * torch.randn(F_in, generator=g)
            state[pre + "0.bias"] = 0.1 * torch.randn(F_in, generator=g)
            state[pre + "1.weight"] = torch.randn(D, F_in, generator=g) / math.sqrt(F_in)
            state[pre + "1.bias"] = 0.1 * torch.randn(D, generator=g)
            state[pre + "3.weight"] = 1 + 0.1 * torch.randn(D, generator=g)
            state[pre + "3.bias"] = 0.1 * torch.randn(D, generator=g)
            state[pre + "4.weight"] = torch.randn(D, D, generator=g) / math.sqrt(D)
            state[pre + "4.bias"] = 0.1 * torch.randn(D, generator=g)
    else:
        patch = patch.to(torch.bfloat16).float()
        word = word.to(torch.bfloat16).float()
        state["space_dict"] = state["space_dict"].to(torch.bfloat16).float()
    n = c["n_proj"]
    sign = lambda rows: (torch.randint(0, 2, (n, rows), generator=g).float() * 2 - 1)
    proj = dict(patch=sign(b * P), word=sign(b * T), space_dict=sign(K))
    return dict(cfg=c, variant=variant, patch=patch, word=word, mask=mask, state=state, proj=proj)


def fingerprint(t):
    t = t.detach().double().reshape(-1)
    return np.array([t.sum().item(), t.abs().sum().item(), t[0].item(), t[-1].item(), float(t.numel())])


def project(R, grad):
    """R [n, rows] (+-1) times a gradient viewed as [rows, cols] -> [n, cols], accumulated in fp64."""
    g2 = grad.detach().double().reshape(R.shape[1], -1)
    return (R.double().to(g2.device) @ g2)
