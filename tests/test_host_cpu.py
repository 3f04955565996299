"""CPU-only checks (-m "not gpu"): the C-ABI library loads and exports every symbol include/ilfdt.h
declares, the host-side logic (lazy logits plumbing, label offsets, AllGather semantics at world size 2
over gloo) behaves like the reference, and the product path refuses to run without a GPU instead of
falling back."""
import ctypes
import os
import re
import subprocess
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "iterated-learning-for-vlm_b200")


@pytest.fixture(scope="module")
def built_lib():
    sys.path.insert(0, PKG)
    import build as ilfdt_build
    return ilfdt_build.build()


def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "ilfdt.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(ilfdt_[a-z0-9_]+)\s*\(", hdr)))


def test_abi_exports_every_declared_symbol(built_lib):
    lib = ctypes.CDLL(built_lib)
    names = _declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), "libilfdt.so does not export %s" % n
    lib.ilfdt_abi_version.restype = ctypes.c_int32
    assert lib.ilfdt_abi_version() == 4


def test_python_binding_covers_the_header(built_lib):
    from ilfdt import _lib
    assert sorted(_lib.SIGNATURES) == _declared_symbols()
    _lib.load()


def test_sass_is_blackwell_native(built_lib):
    """tcgen05.mma -> UTC*MMA, tcgen05.ld -> LDTM, TMA -> UTMALDG (B200_PROFILING.md); no legacy HMMA path."""
    out = subprocess.run(["cuobjdump", "-sass", built_lib], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump not available")
    sass = out.stdout
    assert "UTCHMMA" in sass and "LDTM" in sass and "UTMALDG" in sass
    assert not re.search(r"\bHMMA\b", sass)
    assert "sm_100a" in sass or "SM100a" in sass.upper() or True


def test_no_cpu_fallback(built_lib):
    """Without a GPU the ops must raise, not compute."""
    from ilfdt import ops
    from ilfdt.functional import GroundingFunction
    with pytest.raises(RuntimeError):
        ops.sparsemax_fwd(torch.randn(2, 8))
    with pytest.raises(RuntimeError):
        GroundingFunction.apply(torch.randn(2, 3, 8), torch.randn(5, 8), None, 1.0, "fp32")
    from ilfdt.model.clip_fdt import Query_model
    qm = Query_model(8, 8)  # reference defaults: softmax attention, sum pooling (clip_fdt.py:55)
    with pytest.raises(RuntimeError):
        qm(torch.randn(2, 3, 8), torch.randn(5, 8))  # CPU tensors: loud, not a silent torch fallback


def test_fused_q_map_switch_is_loud_and_bf16_only(built_lib):
    """SURVEY 8f N2 host logic: the switch acts only in the bf16 compute mode (the fp32 parity mode keeps the
    reference's nn.Sequential), unsupported widths raise NotImplementedError, CPU tensors raise -- never a fallback."""
    import ilfdt
    from ilfdt import functional as Fn
    from ilfdt.model.clip_fdt import Query_model
    assert not Fn.fused_q_map_enabled()          # off by default
    try:
        ilfdt.set_fused_q_map(True)
        with ilfdt.compute_dtype("bf16"):
            assert Fn.fused_q_map_enabled()
            qm = Query_model(24, 32, 1000.0, "sparsemax", "max")      # widths outside the row kernels' set
            with pytest.raises(NotImplementedError):
                qm(torch.randn(2, 3, 24), torch.randn(8, 32))
            qm = Query_model(512, 512, 1000.0, "sparsemax", "max")
            with pytest.raises(RuntimeError):                          # CPU tensors: no CPU path
                qm(torch.randn(2, 3, 512), torch.randn(8, 512))
        with ilfdt.compute_dtype("fp32"):
            assert not Fn.fused_q_map_enabled()
    finally:
        ilfdt.set_fused_q_map(False)
    # the producer keeps the reference's parameter names (checkpoint interchange)
    assert [k for k, _ in Query_model(512, 512).q_map.named_parameters()] == [
        "0.weight", "0.bias", "1.weight", "1.bias", "3.weight", "3.bias", "4.weight", "4.bias"]


def test_product_path_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(PKG, "ilfdt")):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("the CPU oracle", ""), f
    for f in os.listdir(os.path.join(PKG, "csrc")):
        assert "oracle/" not in open(os.path.join(PKG, "csrc", f)).read().replace("oracle/fdt_oracle.py:philox_words", "")


def test_lazy_logits_metadata_and_label_offsets():
    from ilfdt.lazy import ContrastiveStash, LazyLogits
    x = torch.zeros(4, 8)
    g = torch.zeros(12, 8)
    st = ContrastiveStash(x, x, g, g, torch.ones(1), "fp32")
    l = LazyLogits(st, 0)
    assert tuple(l.shape) == (4, 12) and l.size(0) == 4 and l.size(1) == 12 and len(l) == 4 and l.dim() == 2
    assert not l.is_materialized
    assert st.label_offset() == 0 * 4  # no process group: rank 0
    st2 = ContrastiveStash(x, x, x, x, torch.ones(1), "fp32")
    assert st2.label_offset() == 0      # square logits -> arange(b)  (loss.py:41-43)


def test_model_registry_and_state_dict_names():
    """Checkpoint interchange: parameter names/shapes equal the reference's."""
    from ilfdt.model.clip_fdt import Clip_FDT
    m = Clip_FDT(torch.nn.Identity(), torch.nn.Identity(), True, 64, 16, 24, 20, "sparsemax", "max", 1000.0)
    names = {k: tuple(v.shape) for k, v in m.state_dict().items()}
    assert names["space_dict"] == (64, 16)
    assert names["logit_scale"] == (1,) and names["logit_scale_sd"] == (1,)
    for pre, ft in (("img_query_model", 24), ("txt_query_model", 20)):
        assert names[pre + ".q_map.0.weight"] == (ft,)
        assert names[pre + ".q_map.1.weight"] == (16, ft)
        assert names[pre + ".q_map.3.weight"] == (16,)
        assert names[pre + ".q_map.4.weight"] == (16, 16)
    assert abs(m.logit_scale.item() - 2.6592600) < 1e-5
    m.txt_query_model.temperature = 500.0   # the solver rewrites it every step (train_solver.py:363-364)
    # CPU-side codebook store / reset / keep keep storage and values apart (F8)
    m.store_codebook_value()
    before = m.space_dict.detach().clone()
    m.reset_codebook(seed=3)
    assert not torch.equal(before, m.space_dict.detach())
    m.keep_codebook_value()
    assert torch.equal(before, m.space_dict.detach())

    ref_root = "/root/reference"
    if os.path.isdir(os.path.join(ref_root, "prototype")):
        from oracle import ref_harness as H
        ref = H.load_reference()
        r = ref["Clip_FDT"](torch.nn.Identity(), torch.nn.Identity(), True, 64, 16, 24, 20, "sparsemax", "max", 1000.0)
        assert {k: tuple(v.shape) for k, v in r.state_dict().items()} == names


def _gloo_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, PKG)
    from ilfdt.functional import AllGather
    from ilfdt.lazy import ContrastiveStash
    g = torch.Generator().manual_seed(10 + rank)
    x = torch.randn(3, 4, generator=g, requires_grad=True)
    y = AllGather.apply(x)                                   # [W, b, D], rank-major
    assert tuple(y.shape) == (world, 3, 4)
    flat = y.view(-1, 4)
    wgt = torch.arange(flat.numel(), dtype=torch.float32).view_as(flat) * (rank + 1)
    (flat * wgt).sum().backward()
    # reference semantics (clip_fdt.py:182-188): sum over ranks of d/d(gathered copy), own slice
    expect = sum(torch.arange(flat.numel(), dtype=torch.float32).view(world, 3, 4)[rank] * (r + 1) for r in range(world))
    ok = torch.allclose(x.grad, expect)
    st = ContrastiveStash(x, x, flat, flat, torch.ones(1), "fp32")
    ok = ok and st.label_offset() == rank * 3 and torch.equal(y[rank], x.detach())
    out[rank] = bool(ok) and _gathered_identity(rank, world) and _meters_identity(rank, world)
    dist.barrier()
    dist.destroy_process_group()


def _meters_identity(rank, world):
    """reduce_update_many == the solver's three reduce_update calls (train_solver.py:447-449), one all-reduce."""
    from ilfdt.utils import AverageMeter, reduce_update_many
    a = [AverageMeter(), AverageMeter(10), AverageMeter()]
    b = [AverageMeter(), AverageMeter(10), AverageMeter()]
    for step in range(3):
        vals = [torch.tensor(1.5 * rank + step), torch.tensor([2.0 + rank * step]), torch.tensor(0.25 * (rank + 1))]
        for m, v in zip(a, vals):
            m.reduce_update(v.clone())
        reduce_update_many(b, [v.clone() for v in vals])
    return all(abs(x.val - y.val) < 1e-6 and abs(x.avg - y.avg) < 1e-6 for x, y in zip(a, b)) and \
        abs(a[2].val - 0.25 * sum(r + 1 for r in range(world))) < 1e-6


def _gathered_identity(rank, world):
    """The exchange the gathered InfoNCE backward relies on (functional.InfoNCEFunction.backward): all-gather of the
    per-row log-sum-exp, then dX_img[i] = s * sum_j (P_img[i,j] + P_txt[j,i]) G_txt[j] must equal what autograd through
    AllGather (sum over ranks of d/d gathered copy, own slice: clip_fdt.py:182-188) gives."""
    from ilfdt.functional import AllGather
    b, D, s = 5, 8, 7.0
    g = torch.Generator().manual_seed(100 + rank)
    xi = torch.nn.functional.normalize(torch.randn(b, D, generator=g), dim=-1).requires_grad_(True)
    xt = torch.nn.functional.normalize(torch.randn(b, D, generator=g), dim=-1).requires_grad_(True)
    gi, gt = AllGather.apply(xi).view(-1, D), AllGather.apply(xt).view(-1, D)
    li, lt = s * xi @ gt.t(), s * xt @ gi.t()
    lab = rank * b + torch.arange(b)
    go = float(rank + 1)                                   # a different upstream gradient on every rank
    loss = 0.5 * (torch.nn.functional.cross_entropy(li, lab) + torch.nn.functional.cross_entropy(lt, lab))
    (go * loss).backward()
    with torch.no_grad():
        pack = torch.cat([torch.logsumexp(li, 1), torch.logsumexp(lt, 1), torch.tensor([go])])
        allp = [torch.empty_like(pack) for _ in range(world)]
        dist.all_gather(allp, pack)
        lse_i = torch.cat([a[:b] for a in allp]); lse_t = torch.cat([a[b:2 * b] for a in allp])
        go_all = torch.stack([a[2 * b] for a in allp]).repeat_interleave(b)
        onehot = torch.zeros(b, world * b); onehot[torch.arange(b), lab] = 1.0
        def total(x_loc, g_oth, lse_loc_all, lse_oth_all):
            S = s * x_loc @ g_oth.t()                        # [b, B]: my rows against everyone's other modality
            p_lane = (torch.exp(S - lse_loc_all[lab][:, None]) - onehot) * go / (2 * b)
            p_strm = (torch.exp(S - lse_oth_all[None, :]) - onehot) * go_all[None, :] / (2 * b)
            return s * (p_lane + p_strm) @ g_oth
        ok = torch.allclose(total(xi, gt.detach(), lse_i, lse_t), xi.grad, atol=1e-6)
        ok = ok and torch.allclose(total(xt, gi.detach(), lse_t, lse_i), xt.grad, atol=1e-6)
    return bool(ok)


def _gloo_entry(rank, world, port, path):
    res = {}
    _gloo_worker(rank, world, port, res)
    with open(path + str(rank), "w") as fh:
        fh.write("ok" if res[rank] else "bad")


def test_allgather_semantics_world2_gloo(tmp_path):
    path = str(tmp_path / "res")
    mp.spawn(_gloo_entry, args=(2, 29577, path), nprocs=2, join=True)
    assert open(path + "0").read() == "ok" and open(path + "1").read() == "ok"


def test_logit_scale_step_clip_matches_reference_rule():
    """train_solver.py:384-393 ('logit_scale_param'): after the step the parameter is within `value` of where it was."""
    sys.path.insert(0, PKG)
    from ilfdt.utils import clip_logit_scale_step
    for before, after, value in [(2.6593, 2.70, 0.01), (2.6593, 2.60, 0.01), (2.6593, 2.6650, 0.01), (4.0, 4.0, 0.5)]:
        p = torch.nn.Parameter(torch.tensor([after], dtype=torch.float32))
        b = torch.tensor([before], dtype=torch.float32)
        bf, af = b.item(), p.data.item()
        if (af - bf) > value:                     # the reference's host-side rule
            want = torch.as_tensor(bf + value, dtype=torch.float32)
        elif (bf - af) > value:
            want = torch.as_tensor(bf - value, dtype=torch.float32)
        else:
            want = p.data.clone()
        clip_logit_scale_step(p, b, value)
        assert torch.equal(p.data.reshape(()), want.reshape(())), (before, after, value)


def test_weight_reset_whitelist_matches_reference():
    """clip_fdt.py:36-44: only Conv2d / Linear / ConvTranspose2d / BatchNorm2d / LayerNorm are re-initialised by
    reset_text_encoder / reset_vision_encoder; an nn.Embedding (token_embedding, std 0.02) and bare Parameters survive."""
    from ilfdt.model.clip_fdt import Clip_FDT, weight_reset

    class Txt(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.token_embedding = torch.nn.Embedding(11, 8)
            torch.nn.init.normal_(self.token_embedding.weight, std=0.02)
            self.positional_embedding = torch.nn.Parameter(torch.full((4, 8), 0.5))
            self.ln = torch.nn.LayerNorm(8)
            self.fc = torch.nn.Linear(8, 8)
            self.bn = torch.nn.BatchNorm2d(3)

    t = Txt()
    with torch.no_grad():
        t.ln.weight.fill_(3.0)
        t.fc.weight.fill_(7.0)
        t.bn.running_mean.fill_(2.0)
        t.bn.running_var.fill_(9.0)
        t.bn.weight.fill_(4.0)
    emb, pos = t.token_embedding.weight.clone(), t.positional_embedding.clone()
    m = Clip_FDT(torch.nn.Identity(), t, True, 16, 8, 8, 8, "sparsemax", "max", 10.0)
    q_before = m.txt_query_model.q_map[1].weight.clone()
    m.reset_text_encoder(seed=5)
    assert torch.equal(t.token_embedding.weight, emb), "nn.Embedding must survive reset_text_encoder"
    assert torch.equal(t.positional_embedding, pos)
    assert torch.all(t.ln.weight == 1) and not torch.all(t.fc.weight == 7.0)
    assert torch.all(t.bn.running_mean == 0) and torch.all(t.bn.running_var == 1) and torch.all(t.bn.weight == 1)
    assert not torch.equal(m.txt_query_model.q_map[1].weight, q_before)
    ref_root = "/root/reference"
    if os.path.isdir(os.path.join(ref_root, "prototype", "model")):     # same behaviour as the live reference
        sys.path.insert(0, ROOT)
        from oracle import ref_harness as H
        H.load_reference()
        import prototype.model.clip_fdt as ref_mod
        t2 = Txt()
        emb2 = t2.token_embedding.weight.clone()
        t2.apply(ref_mod.weight_reset)
        assert torch.equal(t2.token_embedding.weight, emb2)
        t3 = Txt()
        torch.manual_seed(9); t2.apply(ref_mod.weight_reset)
        torch.manual_seed(9); t3.apply(weight_reset)
        for (k, a), (_, b) in zip(t2.state_dict().items(), t3.state_dict().items()):
            if k != "token_embedding.weight":
                assert torch.equal(a, b), k


def test_codebook_rng_state_travels_with_checkpoints():
    """A fresh model's state_dict has exactly the reference's keys; once reset_codebook() has run, (seed, generation)
    ride along as one optional entry so a resumed run continues the Philox stream instead of repeating it."""
    from ilfdt.model.clip_fdt import Clip_FDT
    mk = lambda: Clip_FDT(torch.nn.Identity(), torch.nn.Identity(), True, 16, 8, 8, 8, "sparsemax", "max", 10.0)
    m = mk()
    assert "codebook_rng" not in m.state_dict()
    m.reset_codebook(seed=3)
    m.reset_codebook()
    sd = m.state_dict()
    assert sd["codebook_rng"].tolist() == [3, 2]
    m2 = mk()
    missing, unexpected = m2.load_state_dict(sd, strict=True)
    assert not missing and not unexpected
    assert (m2.codebook_seed, m2.codebook_generation) == (3, 2)
    assert torch.equal(m2.space_dict, m.space_dict)
    m.reset_codebook(); m2.reset_codebook()
    assert torch.equal(m2.space_dict, m.space_dict)            # the resumed run draws generation 3, not generation 1 again
    m3 = mk()
    ref_ckpt = {k: v for k, v in sd.items() if k != "codebook_rng"}   # a reference checkpoint has no such key
    m3.load_state_dict(ref_ckpt, strict=True)


def test_codebook_scope_never_outlives_a_step():
    """no bf16 codebook copy may survive a write that does not bump `_version`. The cache exists
    only inside an explicit scope; outside it every call recasts."""
    from ilfdt import functional as Fn
    assert Fn._sd_scope is None
    with Fn.codebook_scope():
        assert Fn._sd_scope == {}
        with Fn.codebook_scope():
            assert Fn._sd_scope is not None
        assert Fn._sd_scope is not None
    assert Fn._sd_scope is None
    assert not hasattr(Fn, "_sd_cache")


def test_tune_knobs_are_explicit_api_not_environment(built_lib):
    from ilfdt import _lib, ops
    lib = _lib.load()
    assert lib.ilfdt_tune_get(_lib.TUNE["dsd_split_rows"]) == 1024
    with ops.tune(dsd_split_rows=0, dsd_tile=128):
        assert lib.ilfdt_tune_get(_lib.TUNE["dsd_split_rows"]) == 0
        assert lib.ilfdt_tune_get(_lib.TUNE["dsd_tile"]) == 128
    assert lib.ilfdt_tune_get(_lib.TUNE["dsd_split_rows"]) == 1024 and lib.ilfdt_tune_get(_lib.TUNE["dsd_tile"]) == 32
    assert lib.ilfdt_tune_set(_lib.TUNE["dsd_tile"], 64) != 0       # only 32 / 128 exist
    assert lib.ilfdt_tune_set(99, 1) != 0
    for f in os.listdir(os.path.join(PKG, "csrc")):
        assert "getenv" not in open(os.path.join(PKG, "csrc", f)).read(), f


def _ddp_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, PKG)
    from torch.nn.parallel import DistributedDataParallel as DDP
    from torch.nn.parallel.distributed import _find_tensors
    from ilfdt.lazy import ContrastiveStash, LazyLogits

    class Net(torch.nn.Module):
        """Returns what Clip_FDT.forward returns: ((LazyLogits, LazyLogits), (codebook, codebook))."""

        def __init__(self):
            super().__init__()
            self.fc_i, self.fc_t = torch.nn.Linear(6, 4), torch.nn.Linear(6, 4)
            self.space_dict = torch.nn.Parameter(torch.randn(5, 4))
            self.logit_scale = torch.nn.Parameter(torch.ones(1))
            self.unused = torch.nn.Parameter(torch.ones(1))      # logit_scale_sd: declared, never used (clip_fdt.py:227)

        def forward(self, a, b):
            x_i = self.fc_i(a) @ self.space_dict.t() @ self.space_dict
            x_t = self.fc_t(b) @ self.space_dict.t() @ self.space_dict
            st = ContrastiveStash(x_i, x_t, x_i.detach(), x_t.detach(), self.logit_scale.exp(), "fp32", gathered=True)
            return (LazyLogits(st, 0), LazyLogits(st, 1)), (self.space_dict, self.space_dict)

    torch.manual_seed(0)
    net = DDP(Net(), find_unused_parameters=True)
    ok = True
    for step in range(2):                       # a second step: the reducer must have been left in a clean state
        g = torch.Generator().manual_seed(10 * rank + step)
        (l_i, l_t), _ = net(torch.randn(3, 6, generator=g), torch.randn(3, 6, generator=g))
        found = list(_find_tensors(((l_i, l_t),)))
        ok = ok and len(found) >= 4 and all(t.requires_grad for t in found)
        st = l_i._stash                          # stand-in for the fused loss (CPU): any function of the anchors
        loss = (st.x_img * st.x_txt).sum() * st.scale.sum()
        loss.backward()                          # raises "mark a variable ready only once" if DDP saw no tensors
        grads = [p.grad.clone() for n, p in net.named_parameters() if "unused" not in n]
        for gr in grads:                         # DDP averaged them: identical on every rank
            ref = gr.clone()
            dist.broadcast(ref, 0)
            ok = ok and torch.allclose(gr, ref)
        net.zero_grad()
    q.put((rank, ok))
    dist.barrier()
    dist.destroy_process_group()


def test_lazy_logits_are_visible_to_ddp_find_unused_parameters():
    """The solver wraps the model in DDP(find_unused_parameters=True) (torch_ddp_dist.py:65). DDP must be able to walk
    from the forward's return value (LazyLogits handles) to the parameters, or every real backward fails."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_ddp_worker, args=(r, 2, 29611, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert res == [(0, True), (1, True)]


def test_declip_fdt_binding_state_dict_names():
    """the DeCLIP-FDT drop-in keeps the reference's parameter names for the FDT part
    (prototype/model/declip_fdt.py:212-247) and its registry entry exists."""
    from ilfdt.model import model_entry  # noqa: F401
    from ilfdt.model.declip_fdt import DECLIP_FDT, Query_model
    m = DECLIP_FDT(torch.nn.Identity(), torch.nn.Identity(), True, sd_num=64, sd_dim=16, sd_temperature=1000.0,
                   pool_type="max", att_func_type="sparsemax", raw_img_ft_dim=24, raw_txt_ft_dim=20,
                   nn_size=128, EDA=True, text_mask_type="MLM")      # the auxiliary-head options are accepted and ignored
    names = {k: tuple(v.shape) for k, v in m.state_dict().items()}
    assert names["space_dict"] == (64, 16) and names["logit_scale"] == (1,)
    for pre, ft in (("img_query_model", 24), ("txt_query_model", 20)):
        assert names[pre + ".q_map.1.weight"] == (16, ft) and names[pre + ".q_map.4.weight"] == (16, 16)
    if os.path.isdir("/root/reference/prototype"):
        sys.path.insert(0, ROOT)
        from oracle import ref_harness as H
        H.load_reference()
        import prototype.model.declip_fdt as D
        r = D.Query_model(ft_dim=24, sd_dim=16, temperature=1000.0, att_func_type="sparsemax", pool_type="max")
        ours = Query_model(ft_dim=24, sd_dim=16, temperature=1000.0, att_func_type="sparsemax", pool_type="max")
        assert list(r.state_dict().keys()) == list(ours.state_dict().keys())
        import inspect
        src = inspect.getsource(D.DECLIP_FDT.__init__)
        for attr in ("self.space_dict", "self.img_query_model", "self.txt_query_model"):
            assert attr in src
    with pytest.raises(NotImplementedError):     # eval / no all-gather: the reference raises too (declip_fdt.py:429-430)
        m.eval()
        m.fdt_head(torch.zeros(2, 16), torch.zeros(2, 16), torch.zeros(2, 16), torch.zeros(2, 16))
