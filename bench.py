#!/usr/bin/env python
"""bench.py -- benchmark of the FDT grounding + ClipInfoCELoss hot path (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--mode bf16|fp32]
                    [--config headline|1|2|3|3b|4|5]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Default (what the driver runs) = `--config headline`: ViT-B/32 shapes (P=49, T=77, D=512), codebook 16384 x 512,
GLOBAL batch 32768 (BASELINE.json metric) split evenly over the N ranks ("strong" scaling). A "step" is one
forward+backward of the whole path over one synthetic batch, through the reference-facing modules
(ilfdt.model.clip_fdt.Clip_FDT.forward -> ilfdt.loss_functions.ClipInfoCELoss -> loss.backward()): for both modalities
q.codebook^T -> token max-pool -> sparsemax -> weighted pooling, then L2-normalise -> all-gather -> global logits ->
bidirectional InfoNCE, and the backward of all of it (gradients for q_img, q_txt, the codebook and logit_scale).

The other BASELINE.json configs (same JSON contract, one line each):
    1   FDT ViT-B/32 grounding + loss, batch 64, one process (the reference's CPU-runnable case)
    2   standard CLIP ClipInfoCELoss only (prototype/model/clip.py:125-149), batch 256, D=512, one GPU
    3   FDT ViT-B/32 grounding + loss, batch 256 PER GPU (run.sh default: 4 GPUs), codebook 16384;  3b: codebook 4096
    4   contrastive loss only, GLOBAL batch 32768 row-sharded over the ranks
    5   iterated-learning generation step: reset_codebook() + re-grounding + loss, ViT-B/16 (P=196), batch 1024 per GPU
Prints ONE JSON line on rank 0 (contract in the task statement). `--impl reference` times the CPU restatement of the
reference's path (oracle/fdt_oracle.py; the reference is pure PyTorch and cannot travel to the GPU box) on the host
cores, on a bounded sample of the same config.
"""
import argparse
import gc
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "iterated-learning-for-vlm_b200"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

D_DIM = 512
TEMPERATURE = 1000.0
DATASHEET_BF16_TFLOPS = 2250.0     # nominal dense bf16 peak of a B200 (task statement); MEASURED_PEAKS.json is the judge's denominator

CONFIGS = {
    # kind: "fdt" grounding + loss | "clip" loss only through ilfdt.model.clip.CLIP; batch: ("global", n) | ("per_gpu", n)
    "headline": dict(kind="fdt", P=49, T=77, K=16384, batch=("global", 32768), reset=False,
                     metric="loss+grounding fwd/bwd pairs/sec",
                     name="FDT ViT-B/32 codebook grounding + ClipInfoCELoss fwd/bwd, global batch 32768"),
    "1": dict(kind="fdt", P=49, T=77, K=16384, batch=("per_gpu", 64), reset=False,
              metric="loss+grounding fwd/bwd pairs/sec",
              name="BASELINE config 1: FDT ViT-B/32 grounding + ClipInfoCELoss fwd/bwd, batch 64, embed 512, one process"),
    "2": dict(kind="clip", batch=("per_gpu", 256), metric="contrastive loss fwd/bwd pairs/sec",
              name="BASELINE config 2: standard CLIP ClipInfoCELoss only, embed 512, batch 256 on 1 B200"),
    "3": dict(kind="fdt", P=49, T=77, K=16384, batch=("per_gpu", 256), reset=False,
              metric="loss+grounding fwd/bwd pairs/sec",
              name="BASELINE config 3: FDT ViT-B/32 grounding (codebook 16384x512, sparsemax) + loss, batch 256/GPU (run.sh: 4 GPUs)"),
    "3b": dict(kind="fdt", P=49, T=77, K=4096, batch=("per_gpu", 256), reset=False,
               metric="loss+grounding fwd/bwd pairs/sec",
               name="BASELINE config 3b: as config 3 with the shipped yaml's codebook 4096x512 (config_cc3m.yaml:18)"),
    "4": dict(kind="clip", batch=("global", 32768), metric="contrastive loss fwd/bwd pairs/sec",
              name="BASELINE config 4: global-batch 32768 contrastive loss, row-sharded logits, all-gather over NVLink"),
    "5": dict(kind="fdt", P=196, T=77, K=16384, batch=("per_gpu", 1024), reset=True,
              metric="generation step (codebook reset + re-grounding + loss) fwd/bwd pairs/sec",
              name="BASELINE config 5: iterated-learning generation step, codebook reset + re-grounding, ViT-B/16 (P=196), batch 1024/GPU"),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--config", default="headline", choices=sorted(CONFIGS))
    ap.add_argument("--global-batch", type=int, default=None, help="override the config's batch (global)")
    ap.add_argument("--per-gpu-batch", type=int, default=None, help="override the config's batch (per GPU: weak scaling)")
    ap.add_argument("--codes", type=int, default=None, help="override the config's codebook size")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--graph", action="store_true", help="replay a CUDA graph of the step (default when N > 1 or the per-GPU batch is small)")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the trained-like (peaked attention) secondary case")
    ap.add_argument("--no-qmap", action="store_true", help="skip the q_map producer timing (SURVEY 8f N2)")
    ap.add_argument("--trained-scale", type=float, default=200.0, help="q multiplier of the trained-like secondary case")
    ap.add_argument("--cpu-sample", type=int, default=None, help="rows of the workload timed on the host cores")
    return ap.parse_args()


def resolve(args):
    """The workload of this run: the named config with the command-line overrides applied."""
    c = dict(CONFIGS[args.config])
    world = max(int(os.environ.get("WORLD_SIZE", 1)), 1) if args.impl == "ours" else max(args.gpus, 1)
    if args.global_batch is not None:
        c["batch"] = ("global", args.global_batch)
    if args.per_gpu_batch is not None:
        c["batch"] = ("per_gpu", args.per_gpu_batch)
    if args.codes is not None and c["kind"] == "fdt":
        c["K"] = args.codes
    how, n = c["batch"]
    c["b"] = n // world if how == "global" else n
    c["B"] = c["b"] * world
    c["scaling"] = "strong" if how == "global" else "weak"
    c["world"] = world
    return c


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return dict(tflops=float(d.get("bf16_tflops_sustained", d.get("bf16_tflops"))), hbm=float(d["hbm_gbs"]),
                    burst=float(d.get("bf16_tflops", 0.0)) or None, src="measured (MEASURED_PEAKS.json, sustained bf16)")
    return dict(tflops=1590.0, hbm=6650.0, burst=1590.0, src="fallback (B200_PROFILING.md)")


def flops_per_pair(c, nnz_img=None, nnz_txt=None):
    """ALGORITHMIC FLOPs per image-text pair: F = 2(P+T)DK + 3(4KD) + 2 nnz 4D + 12 B D for the FDT step,
    12 B D for the loss-only configs. nnz = measured mean sparsemax support per row (0.29 K at random init)."""
    B = c["B"]
    if c["kind"] == "clip":
        return 12.0 * B * D_DIM
    K, P, T = c["K"], c["P"], c["T"]
    nnz = (nnz_img if nnz_img is not None else 0.29 * K) + (nnz_txt if nnz_txt is not None else 0.29 * K)
    return 2.0 * (P + T) * D_DIM * K + 3 * (4.0 * K * D_DIM) + nnz * 4.0 * D_DIM + 12.0 * B * D_DIM


def workload_config(args, c, note=None):
    w = {"workload": c["name"], "bench_config": args.config, "global_batch": c["B"], "per_gpu_batch": c["b"],
         "parallelism": "dp%d (row-sharded logits; bf16 all-gather of the embeddings fwd, all-gather of the row "
                        "log-sum-exps bwd)" % c["world"]}
    if c["kind"] == "fdt":
        w.update(codes=c["K"], P=c["P"], T=c["T"], D=D_DIM, temperature=TEMPERATURE, att="sparsemax", pool="max",
                 codebook_reset_in_step=bool(c["reset"]))
        in_bytes = c["b"] * (c["P"] + c["T"]) * D_DIM * 4
    else:
        w.update(D=D_DIM, head="prototype/model/clip.py:125-149 (features in, no grounding)")
        in_bytes = c["b"] * 2 * D_DIM * 4
    if in_bytes > 256e6:
        w["l2_policy"] = "inputs larger than L2 (%.2f GB per rank per step)" % (in_bytes / 1e9)
    else:
        w["l2_policy"] = "L2 flushed between timed iterations (a 256 MB buffer is overwritten before every step; each " \
                         "step is timed between its own pair of CUDA events, the flushes lie between the pairs)"
    if note:
        w["note"] = note
    return w


# ---------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference path, on the host cores
# ---------------------------------------------------------------------------------------------------
def _oracle_step_fn(c, sample_rows, dev, seed=0):
    """(callable running one fwd+bwd of the oracle on `sample_rows` rows of the config, description). The logits of the
    sample are sample x B gathered rows for the loss-only configs and sample x sample for the FDT configs (the
    grounding dominates there)."""
    from oracle import fdt_oracle as O
    if c["kind"] == "fdt":
        inp = O.synth_inputs(sample_rows, c["P"], c["T"], c["K"], D_DIM, seed=seed)
        inp = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in inp.items()}

        def run():
            ranks = [dict(q_img=O.make_leaf(inp["q_img"]), q_txt=O.make_leaf(inp["q_txt"]), mask=inp["mask"],
                          sd=O.make_leaf(inp["sd"]), p=O.make_leaf(inp["p"]))]
            O.step(ranks, temperature=TEMPERATURE)
        what = "%d rows of the workload (P=%d,T=%d,K=%d,D=%d; logits %dx%d)" % (
            sample_rows, c["P"], c["T"], c["K"], D_DIM, sample_rows, sample_rows)
        return run, what
    g = torch.Generator().manual_seed(seed)
    Bg = c["B"]
    f_img = torch.randn(Bg, D_DIM, generator=g).to(dev)
    f_txt = torch.randn(Bg, D_DIM, generator=g).to(dev)
    p0 = torch.full((1,), math.log(1 / 0.07)).to(dev)

    def run():
        fi, ft, p = O.make_leaf(f_img), O.make_leaf(f_txt), O.make_leaf(p0)
        xi, xt = O.l2_normalise(fi, 0.0), O.l2_normalise(ft, 1e-10)         # clip.py:133-134
        s = O.logit_scale_forward(p)
        l_i, l_t = O.clip_logits(xi[:sample_rows], xt[:sample_rows], xi, xt, s)
        loss, _ = O.info_nce(l_i, l_t, rank=0)
        loss.backward()
    return run, "%d local rows against %d gathered rows (D=%d; logits %dx%d x2)" % (sample_rows, Bg, D_DIM, sample_rows, Bg)


def default_cpu_sample(c):
    if c["kind"] == "clip":
        return min(c["b"], 1024)
    return min(c["b"], 32 if c["P"] > 100 else 256)


def cpu_reference_run(c, sample_rows, steps, warmup):
    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    run, what = _oracle_step_fn(c, sample_rows, torch.device("cpu"))
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        run()
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    t = sum(times) / len(times)
    return dict(value=sample_rows / t, unit="pairs/s", cores=cores, kind="port",
                sample="%s, fp32 torch CPU (oracle/fdt_oracle.py, the restatement of the reference's PyTorch path), "
                       "%d steps, %.3f s/step" % (what, len(times), t)), t


def gpu_eager_reference_run(c, sample_rows, steps, warmup, dev):
    """Context number of SURVEY 8d: the same restatement of the reference's path (fp32 PyTorch eager: dense [b,N,K]
    matmuls, sort-based sparsemax, F.cross_entropy, autograd; TF32 off as in the reference) on the SAME B200. Not the
    target and not the product: the practical bar a user of the reference sees today."""
    torch.backends.cuda.matmul.allow_tf32 = False
    run, what = _oracle_step_fn(c, sample_rows, dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ms = 0.0
    for i in range(warmup + steps):
        e0.record()
        run()
        e1.record()
        torch.cuda.synchronize()
        if i >= warmup:
            ms += e0.elapsed_time(e1)
    ms /= steps
    torch.cuda.empty_cache()
    return dict(value=sample_rows / (ms * 1e-3), unit="pairs/s", ms_per_step=ms,
                sample="%s, fp32 PyTorch eager on the same B200, TF32 off, %d steps" % (what, steps))


def run_reference(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    c = resolve(args)
    steps = max(1, min(args.steps, 5))
    warmup = max(1, min(args.warmup, 2))
    sample = args.cpu_sample or default_cpu_sample(c)
    cb, t = cpu_reference_run(c, sample, steps, warmup)
    line = {
        "impl": "reference", "metric": c["metric"], "value": cb["value"], "unit": "pairs/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": c["scaling"],
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, c, note="CPU arm: each step is a %d-row sample of the workload" % sample),
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------
# q_map producer: reported beside the headline, never inside it (8a row a0 keeps q_map out of the metric)
# ---------------------------------------------------------------------------------------------------
def measure_q_map(c, dev, steps):
    """Forward + backward of both Query_model.q_map MLPs (clip_fdt.py:86-92) at the bench's batch: the library route
    (ilfdt.functional.QMapFunction: tcgen05 Linear layers on bf16 operands, fp32 row kernels, q handed to the score
    kernel in bf16) and the reference's route (fp32 nn.Sequential in PyTorch eager, TF32 off as in the reference) on a
    2048-sample slice scaled to the batch."""
    from ilfdt import functional as Fn
    from ilfdt import ops
    from ilfdt.model.clip_fdt import Query_model
    b = c["b"]
    out = {"ilfdt_bf16_ms": 0.0, "torch_fp32_eager_ms": 0.0, "parts": {}}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for name, n_tok, width in (("image", c["P"], 768), ("text", c["T"], 512)):
        torch.manual_seed(0)
        qm = Query_model(width, D_DIM, TEMPERATURE, "sparsemax", "max").to(dev)
        x = torch.randn(b, n_tok, width, device=dev, requires_grad=True)
        up = torch.randn(b, n_tok, D_DIM, device=dev, dtype=torch.bfloat16)   # dq as the scatter backward returns it

        def fused():
            Fn.q_map_apply(qm.q_map, x).backward(up)

        sample = min(b, 2048)

        def eager():
            xi = x[:sample].detach().requires_grad_(True)
            qm.q_map(xi).backward(up[:sample].float())

        part = {}
        for tag, fn, scale in (("ilfdt_bf16_ms", fused, 1.0), ("torch_fp32_eager_ms", eager, b / float(sample))):
            for _ in range(2):
                fn()
            torch.cuda.synchronize()
            ops.PROFILE.clear()
            ops.PROFILE_ON = tag == "ilfdt_bf16_ms"
            e0.record()
            for _ in range(steps):
                fn()
            e1.record()
            torch.cuda.synchronize()
            ops.PROFILE_ON = False
            part[tag] = e0.elapsed_time(e1) / steps * scale
            out[tag] += part[tag]
            if tag == "ilfdt_bf16_ms":
                part["kernel_ms"] = {k: round(v[1] / steps, 3) for k, v in ops.profile_summary().items()}
            ops.PROFILE.clear()
        rows = b * n_tok
        part["tflops_ilfdt"] = 3 * 2.0 * rows * (width * D_DIM + D_DIM * D_DIM) / (part["ilfdt_bf16_ms"] * 1e-3) / 1e12
        out["parts"][name] = part
        del qm, x, up
        torch.cuda.empty_cache()
    out["note"] = ("both q_map MLPs fwd+bwd at this batch; eager = fp32 nn.Sequential on a %d-sample slice scaled to "
                   "the batch; not part of `value` (SURVEY 8a row a0)" % min(b, 2048))
    return out


# ---------------------------------------------------------------------------------------------------
# clocks sampler
# ---------------------------------------------------------------------------------------------------
class Clocks:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for ts, row in self.rows:
            if ts < t0 or ts > t1 + 0.2:
                continue
            f = [x.strip() for x in row.split(",")]
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------
# host side of the e2e leg: CPU affinity and ONE packed pinned buffer per step
# ---------------------------------------------------------------------------------------------------
def bind_near_gpu(local):
    """Pin this rank to the CPUs of the NUMA node its GPU hangs off BEFORE any pinned allocation, so that the staging
    buffers are node-local. Returns what was done (the bench line carries it). On a virtualised box that exposes one
    NUMA node there is nothing to bind to."""
    info = {"numa_node": None, "cpus": None, "bound": False}
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = int(vis.split(",")[local]) if vis and all(x.strip().isdigit() for x in vis.split(",")) else local
        h = pynvml.nvmlDeviceGetHandleByIndex(idx)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        path = "/sys/bus/pci/devices/%s/numa_node" % bus[-12:].lower()
        node = int(open(path).read().strip()) if os.path.exists(path) else -1
        info["numa_node"] = node
        nodes = [d for d in os.listdir("/sys/devices/system/node") if d.startswith("node") and d[4:].isdigit()]
        info["numa_nodes_visible"] = len(nodes)
        if node >= 0 and len(nodes) > 1:
            cpus = set()
            for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
            cpus &= os.sched_getaffinity(0)
            if cpus:
                os.sched_setaffinity(0, cpus)
                info["bound"] = True
        info["cpus"] = len(os.sched_getaffinity(0))
    except Exception as e:  # noqa: BLE001 -- affinity is an optimisation, never a reason to fail the bench
        info["error"] = str(e).splitlines()[0][:120]
    return info


class PackedHost:
    """One pinned host buffer and two device buffers holding all inputs of a step back to back (256-byte aligned):
    ONE host->device copy per step; the step's tensors are views into the device buffer."""

    def __init__(self, tensors, dev):
        self.specs, off = [], 0
        for t in tensors:
            off = (off + 255) // 256 * 256
            self.specs.append((off, t.shape, t.dtype, t.numel() * t.element_size()))
            off += t.numel() * t.element_size()
        self.nbytes = off
        self.host = torch.empty(off, dtype=torch.uint8, pin_memory=True)
        for (o, shape, dtype, n), t in zip(self.specs, tensors):
            self.host[o:o + n].view(dtype).view(shape).copy_(t)
        self.dev = [torch.empty(off, dtype=torch.uint8, device=dev) for _ in range(2)]

    def views(self, i):
        return tuple(self.dev[i][o:o + n].view(dtype).view(shape) for o, shape, dtype, n in self.specs)

    def copy(self, i):
        self.dev[i].copy_(self.host, non_blocking=True)


# ---------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------
def run_ours(args):
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the ilfdt path has no CPU fallback")
    affinity = bind_near_gpu(local)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    assert world == max(args.gpus, 1) or world == 1, "launch with torchrun --nproc-per-node == --gpus"

    import ilfdt
    from ilfdt import functional as Fn
    from ilfdt import ops
    from ilfdt.loss_functions import ClipInfoCELoss
    from ilfdt.model.clip import CLIP
    from ilfdt.model.clip_fdt import Clip_FDT
    ilfdt.set_compute_dtype(args.mode)

    c = resolve(args)
    b, B, kind = c["b"], c["B"], c["kind"]
    g = torch.Generator(device=dev).manual_seed(1000 + rank)
    p = torch.nn.Parameter(torch.full((1,), math.log(1 / 0.07), device=dev))
    criterion = ClipInfoCELoss()
    lens = None

    class _Identity(torch.nn.Module):
        def forward(self, x):
            return x

    # ---- the step through the reference-facing classes (what example/clip_fdt/train_solver.py calls):
    # model = Clip_FDT(...); (logits, _) = model(images, texts); loss, _ = ClipInfoCELoss()(*logits); loss.backward().
    # The encoders are outside the path: stand-ins hand the step's tensors through, and q_map (not in
    # the metric) is the identity.
    if kind == "fdt":
        P, T, K = c["P"], c["T"], c["K"]
        q_img = torch.randn(b, P, D_DIM, device=dev, generator=g)
        q_txt = torch.randn(b, T, D_DIM, device=dev, generator=g)
        lens = torch.randint(5, T + 1, (b,), device=dev, generator=g)
        mask = torch.zeros(b, T, device=dev)
        mask[torch.arange(T, device=dev).view(1, -1) >= lens.view(-1, 1)] = float("-inf")
        gs = torch.Generator(device=dev).manual_seed(0)   # same codebook on every rank
        sd = torch.nn.Parameter(torch.randn(K, D_DIM, device=dev, generator=gs))
        inputs = (q_img, q_txt, mask)

        class _Visual(torch.nn.Module):
            def forward(self, image, return_dense=False, return_raw_feature=False):
                return None, image, None

        class _Text(torch.nn.Module):
            def forward(self, texts, return_dense=False, return_padmask=False, return_raw_feature=False, raw_text=True):
                return [None, texts[0], None, texts[1]]

        model = Clip_FDT(_Visual(), _Text(), True, K, D_DIM, D_DIM, D_DIM, "sparsemax", "max", TEMPERATURE)
        model.img_query_model.q_map = _Identity()
        model.txt_query_model.q_map = _Identity()
        model.space_dict = sd
        model.logit_scale = p
        model.encode_image = lambda image: model.visual(image, return_dense=True, return_raw_feature=True)
        params = [sd, p]

        def forward(qi, qt, mk):
            if c["reset"]:
                model.reset_codebook()               # clip_fdt.py:281-283, inside the timed generation step (config 5)
            (l_i, l_t), _ = model(qi, (qt, mk))
            return l_i, l_t
        api = "ilfdt.model.clip_fdt.Clip_FDT.forward + ilfdt.loss_functions.ClipInfoCELoss + loss.backward()"
        if c["reset"]:
            api = "Clip_FDT.reset_codebook() + " + api
    else:
        f_img = torch.randn(b, D_DIM, device=dev, generator=g)
        f_txt = torch.randn(b, D_DIM, device=dev, generator=g)
        inputs = (f_img, f_txt)
        model = CLIP(_Identity(), _Identity(), True)
        model.logit_scale = p
        model.encode_image = lambda image: image
        params = [p]

        def forward(fi, ft):
            return model(fi, ft)
        api = "ilfdt.model.clip.CLIP.forward + ilfdt.loss_functions.ClipInfoCELoss + loss.backward()"
    model.to(dev)
    n_grad_inputs = 2
    keep = [None, None]                                  # the gradients handed back to the producer stay reachable

    def model_step(*ins):
        leaves = [t.detach().requires_grad_(True) for t in ins[:n_grad_inputs]]
        keep[0], keep[1] = leaves
        for prm in params:
            prm.grad = None
        l_i, l_t = forward(*leaves, *ins[n_grad_inputs:])
        loss, _labels = criterion(l_i, l_t)
        loss.backward()
        return loss

    in_bytes = sum(t.numel() * t.element_size() for t in inputs)
    flush = None
    if in_bytes <= 256e6:                                # inputs fit L2 (126 MB): flush it between timed iterations
        flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device=dev)

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        model_step(*inputs)
    sync()
    ops.LAUNCHES = 0
    eager_loss = float(model_step(*inputs).item())
    launches_per_step = ops.LAUNCHES

    # One step is ~100-200 launches issued from Python. When the per-GPU work is small (8 GPUs x 4096 rows, or the
    # 64 / 256-row configs) that host work, not the GPU, paces the step, so the step (forward + backward + the NCCL
    # all-gathers) is captured in a CUDA graph once and replayed. The timed region never contains per-call events.
    # At large single-GPU batches the GPU is the bottleneck and eager issue gives the same steady-state time -- but one host
    # stall inside a 5-step timed region (measured: a 216 ms step among 84 ms ones, host issue time 36 ms per step instead of
    # 4) moves the number by 15 %. The value leg therefore replays the graph at every size; the e2e leg keeps eager issue at
    # large batches (it copies straight into the tensors the step reads; a replay would add a device-to-device copy of them).
    use_graph = not args.no_graph
    graph_for_e2e = (world > 1 or args.graph or b <= 8192) and not args.no_graph
    run_step = lambda: model_step(*inputs)
    graph_note = "off (eager issue)"
    gstep = None
    if use_graph:
        try:
            from ilfdt.graphs import GraphedStep
            gstep = GraphedStep(model_step, inputs, warmup=1)
            gl = float(gstep().item())
            if not c.get("reset") and not (abs(gl - eager_loss) <= 1e-6 * max(1.0, abs(eager_loss))):
                raise RuntimeError("graph replay loss %r != eager loss %r" % (gl, eager_loss))
            run_step = gstep
            graph_note = "on (forward + backward + collectives of one step captured once, replayed per step)"
        except Exception as e:  # noqa: BLE001 -- report and measure eagerly
            gstep = None
            graph_note = "fallback to eager: %s" % (str(e).splitlines()[0][:160])
    sync()

    # ---- device-resident timing (value): K steps between two events, no per-call events inside ----
    for _ in range(2):
        run_step()
    sync()
    clocks = Clocks(local)
    if rank == 0:
        clocks.start()
        time.sleep(0.3)
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    t0 = time.perf_counter()
    e0.record()
    for i in range(args.steps):
        if flush is not None:
            flush.zero_()            # evict the previous iteration's working set from the 126 MB L2
        starts[i].record()
        loss = run_step()
        marks[i].record()
    e1.record()
    t_host = time.perf_counter()
    sync()
    t1 = time.perf_counter()
    host_ms_per_step = (t_host - t0) * 1e3 / args.steps
    per_step = [starts[i].elapsed_time(marks[i]) for i in range(args.steps)]
    if flush is None:
        ms = e0.elapsed_time(e1)     # K back-to-back steps between two events
        flush_ms = 0.0
    else:
        ms = sum(per_step)           # K steps, each between its own pair of events; the L2 flushes sit between them
        flush_ms = (e0.elapsed_time(e1) - ms) / args.steps
    if world > 1:
        tms = torch.tensor([ms], device=dev)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        ms = tms.item()
    clk = clocks.stop(t0, t1) if rank == 0 else None
    ms_per_step = ms / args.steps
    value = B / (ms_per_step * 1e-3)
    loss_val = float(loss.item())
    if gstep is not None and not graph_for_e2e:
        gstep.release()              # the e2e leg of a large single-GPU batch issues eagerly: give the graph's pool back first
        gstep = None
        torch.cuda.empty_cache()

    # ---- per-kernel device times: a separate eager pass of the same step with CUDA events around every C-ABI call ----
    ops.PROFILE.clear()
    ops.PROFILE_ON = True
    for _ in range(args.steps):
        model_step(*inputs)
    sync()
    ops.PROFILE_ON = False
    prof = ops.profile_summary()
    kernel_ms = {k: v[1] / args.steps for k, v in prof.items()}
    ops.PROFILE.clear()

    # ---- measured support size (enters the algorithmic FLOP count) ----
    nnz_i = nnz_t = None
    if kind == "fdt":
        with torch.no_grad():
            n_s = min(b, 512)
            w_i, _ = Fn.GroundingFunction.apply(q_img[:n_s], sd.detach(), None, TEMPERATURE, Fn.get_compute_dtype())
            w_t, _ = Fn.GroundingFunction.apply(q_txt[:n_s], sd.detach(), mask[:n_s], TEMPERATURE, Fn.get_compute_dtype())
            nnz_i = float((w_i > 0).sum(1).float().mean().item())
            nnz_t = float((w_t > 0).sum(1).float().mean().item())
            del w_i, w_t
    F_pair = flops_per_pair(c, nnz_i, nnz_t)

    # ---- roofline of the dominant kernel ----
    pk = peaks()
    roof = None
    timed_with = ("CUDA events around this launch in an eager pass of the same step run right after the timed region "
                  "(the timed region itself carries no per-call events)")
    if kind == "fdt" and "ground_fwd" in prof:
        # dominant kernel: ground2_kernel<P> = the image launch of ilfdt_ground_fwd (one per step).
        # ALGORITHMIC FLOPs per launch = 2 * b * P * D * K.
        n_launch, tot_ms = prof["ground_fwd"]
        flops_per_launch = 2.0 * b * c["P"] * D_DIM * c["K"]
        avg_ms = tot_ms / n_launch
        achieved = flops_per_launch / (avg_ms * 1e-3) / 1e12
        roof = {"bound": "tensor", "kernel": "ground2_kernel<%d> (tcgen05 cta_group::2; ilfdt_ground_fwd, image launch)" % c["P"],
                "achieved": achieved, "peak": pk["tflops"], "unit": "TFLOP/s", "frac": achieved / pk["tflops"],
                # dram__bytes_read.sum + dram__bytes_write.sum of this launch from the committed ncu --set full capture
                # (profiles/r01_ncu_step_b32768_ground2_ncebwd.txt: 3.05 GB + 3.20 GB); only valid for that shape
                "traffic": 6.245e9 if (b == 32768 and c["K"] == 16384 and c["P"] == 49) else None,
                "traffic_unit": "bytes per launch (ncu)", "avg_launch_ms": avg_ms,
                "algorithmic_flops_per_launch": flops_per_launch}
        if "ground_fwd_text" in prof:
            # text launches (one per 16-row bucket): algorithmic = full-width 2*b*T*D*K; executed = rows kept by the
            # pad compaction (valid tokens + one pad, rounded to 16)
            n_t, tot_t = prof["ground_fwd_text"]
            T = c["T"]
            rows_kept = torch.clamp(((torch.clamp(lens + 1, max=T) + 15) // 16) * 16, max=T).sum().item()
            ms_t = tot_t / n_t
            roof["text"] = {"kernel": "ground2_kernel<R> x 5 buckets (ilfdt_ground_fwd_text)", "avg_ms": ms_t,
                            "algorithmic_tflops": 2.0 * b * T * D_DIM * c["K"] / (ms_t * 1e-3) / 1e12,
                            "executed_tflops": 2.0 * rows_kept * D_DIM * c["K"] / (ms_t * 1e-3) / 1e12,
                            "rows_kept_frac": rows_kept / float(b * T)}
    elif kind == "clip" and "nce_bwd_gathered" in prof:
        # dominant kernel of the loss-only configs: nce_bwd2_kernel<2>, two launches per ilfdt_nce_bwd_gathered call
        # (one per modality); ALGORITHMIC FLOPs per call = 8 * b * B * D (4 GEMM-equivalents of 2 b B D)
        n_call, tot_ms = prof["nce_bwd_gathered"]
        flops_per_call = 8.0 * b * B * D_DIM
        avg_ms = tot_ms / n_call
        achieved = flops_per_call / (avg_ms * 1e-3) / 1e12
        roof = {"bound": "tensor", "kernel": "nce_bwd2_kernel<2> x 2 launches (tcgen05 cta_group::2; ilfdt_nce_bwd_gathered)",
                "achieved": achieved, "peak": pk["tflops"], "unit": "TFLOP/s", "frac": achieved / pk["tflops"],
                "traffic": None, "traffic_unit": "bytes per launch (ncu)", "avg_launch_ms": avg_ms / 2,
                "algorithmic_flops_per_launch": flops_per_call / 2}
    if roof is not None:
        roof.update(peak_source=pk["src"], timed_with=timed_with,
                    share_of_step=(roof["avg_launch_ms"] * (2 if kind == "clip" else 1)) / ms_per_step,
                    frac_of_burst_peak=(roof["achieved"] / pk["burst"]) if pk.get("burst") else None,
                    frac_of_datasheet_peak=roof["achieved"] / DATASHEET_BF16_TFLOPS)
        if roof["frac"] > 1.0:
            roof["peak_note"] = ("frac > 1: the measured 'sustained' figure (torch.matmul 8192^3 back to back under the "
                                 "power cap) is not a hardware ceiling for a kernel timed alone; see "
                                 "frac_of_burst_peak / frac_of_datasheet_peak")
    step_tflops = value / world * F_pair / 1e12
    step_roofline = {"algorithmic_mflop_per_pair": F_pair / 1e6, "achieved_tflops_per_gpu": step_tflops,
                     "frac_of_sustained_peak": step_tflops / pk["tflops"],
                     "frac_of_datasheet_peak": step_tflops / DATASHEET_BF16_TFLOPS,
                     "mean_support_img": nnz_i, "mean_support_txt": nnz_t}

    # ---- end to end through the public call with HOST buffers ----
    def e2e_measure(host_dtype):
        """The same step through the public call with HOST buffers: every step copies its inputs from ONE packed
        pinned host buffer (on a copy stream, double-buffered against the previous step's compute) and reads the loss
        back to pinned host memory, all inside the timed region."""
        with torch.no_grad():
            host_tensors = [t.to(host_dtype) if i < n_grad_inputs else t for i, t in enumerate(inputs)]
            pack = PackedHost(host_tensors, dev)
            del host_tensors
            h_loss = torch.empty((), dtype=torch.float32, pin_memory=True)
        cs = torch.cuda.Stream()
        ready = [torch.cuda.Event(), torch.cuda.Event()]
        free = [torch.cuda.Event(), torch.cuda.Event()]
        # copy-only probe: the same buffer, the same stream, no compute -- what the box gives this rank when all ranks
        # copy at once (the ceiling of any e2e number at this input size)
        sync()
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(cs):
            pack.copy(0)
            p0.record(cs)
            for i in range(4):
                pack.copy(i % 2)
            p1.record(cs)
        sync()
        probe_ms = p0.elapsed_time(p1) / 4

        def prefetch(i):
            with torch.no_grad(), torch.cuda.stream(cs):
                cs.wait_event(free[i % 2])
                pack.copy(i % 2)
                ready[i % 2].record(cs)

        # When the step is replayed from a CUDA graph (small per-GPU batches, N > 1) the e2e leg replays it too: the
        # packed buffer that arrived from the host is copied device-to-device into the graph's captured input tensors
        # (a few microseconds) and the graph is launched -- otherwise ~150 kernel launches issued from Python, not the
        # GPU or PCIe, would set the end-to-end number.
        e2e_graph = graph_inputs = None
        if gstep is not None:
            try:
                from ilfdt.graphs import GraphedStep
                with torch.no_grad():
                    graph_inputs = tuple(t.to(host_dtype) if i < n_grad_inputs else t.clone() for i, t in enumerate(inputs))
                e2e_graph = GraphedStep(model_step, graph_inputs, warmup=1)
            except Exception:  # noqa: BLE001 -- measure eagerly
                e2e_graph = graph_inputs = None

        def run_on(views):
            if e2e_graph is None:
                return model_step(*views)
            with torch.no_grad():
                for dst, src in zip(graph_inputs, views):
                    dst.copy_(src, non_blocking=True)
            return e2e_graph()

        def e2e_run(n, t0=None, t1=None):
            """n pipelined steps; e0 / e1 are recorded on the compute stream before step t0 / after step t1 - 1."""
            cur = torch.cuda.current_stream()
            for ev in free:
                ev.record(cur)
            prefetch(0)
            for i in range(n):
                if i == t0:
                    e0.record(cur)
                if i + 1 < n:
                    prefetch(i + 1)
                cur.wait_event(ready[i % 2])
                l = run_on(pack.views(i % 2))
                free[i % 2].record(cur)
                with torch.no_grad():
                    h_loss.copy_(l.detach(), non_blocking=True)
                if i + 1 == t1:
                    e1.record(cur)

        def rank_max(ms):
            if world > 1:
                tms = torch.tensor([ms], device=dev)
                dist.all_reduce(tms, op=dist.ReduceOp.MAX)
                ms = tms.item()
            return ms

        # (1) steady state: ONE pipeline of warm-up + K timed + 1 trailing step. The timed region (device events
        # around steps W .. W+K-1) holds exactly K host->device copies (those of steps W+1 .. W+K, overlapping the
        # compute), K steps and K loss read-backs; the trailing step consumes the last prefetch.
        wu = max(1, min(args.warmup, 2))
        sync()
        e2e_run(wu + args.steps + 1, wu, wu + args.steps)
        sync()
        ems = rank_max(e0.elapsed_time(e1))
        # (2) the same K steps started from an idle device: the first copy has nothing to hide behind (pipeline fill)
        e2e_run(args.steps, 0, args.steps)
        sync()
        ems_fill = rank_max(e0.elapsed_time(e1))
        probe_ms = rank_max(probe_ms)
        h2d = pack.nbytes
        del pack
        used_graph = e2e_graph is not None
        if e2e_graph is not None:
            e2e_graph.release()
        return {"value": B / (ems / args.steps * 1e-3), "unit": "pairs/s", "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": 4, "ms_per_step": ems / args.steps,
                "issue": "CUDA graph replay" if used_graph else "eager (Python issues every launch)",
                "h2d_copy_only": {"ms": probe_ms, "gb_per_s_per_gpu": h2d / probe_ms / 1e6,
                                  "pairs_per_s_ceiling": B / (probe_ms * 1e-3),
                                  "note": "the same packed buffer copied with no compute, all ranks at once (slowest rank)"},
                "from_idle": {"value": B / (ems_fill / args.steps * 1e-3), "ms_per_step": ems_fill / args.steps,
                              "note": "same K steps with the pipeline fill (first copy exposed) inside the timed region"}}

    e2e = None
    if not args.no_e2e:
        # bf16 compute mode rounds its inputs to bf16 as its first act, so the host side may hold them in bf16 (a
        # producer under bf16 autocast): half the PCIe bytes, same results. The reference's own producer is fp32: that
        # number is reported in the same object as `fp32_host`.
        if args.mode == "bf16":
            e2e = e2e_measure(torch.bfloat16)
            e2e["host_input_dtype"] = "bf16 (a dtype change at the API boundary the reference does not make; fp32_host = the reference's dtype)"
            e2e["fp32_host"] = e2e_measure(torch.float32)
        else:
            e2e = e2e_measure(torch.float32)
            e2e["host_input_dtype"] = "fp32"
        e2e["api"] = api
        e2e["launches_per_step"] = launches_per_step
        e2e["loss_first_step"] = eager_loss
        e2e["cpu_affinity"] = affinity
        e2e["note"] = ("all inputs of a step are copied from ONE packed pinned host buffer (per rank) on a copy stream, "
                       "double-buffered against the previous step's compute; loss read back to pinned host memory; "
                       "value = steady state of one continuous pipeline (K copies, K steps, K read-backs between the "
                       "device events), from_idle = the same with the first, unhidden copy included")

    # ---- secondary case: "trained-like" peaked attention. Random-init statistics (temperature 1000)
    # leave ~0.29 K codes in every sparsemax support, which is what the sparse scatter backward is paid for; scaling
    # the queries makes the scores peaked (support of tens of codes) as in a trained model. Reported beside the
    # headline, never instead of it.
    secondary = None
    if kind == "fdt" and world == 1 and not args.no_secondary and not c["reset"]:
        with torch.no_grad():
            qi2, qt2 = q_img * args.trained_scale, q_txt * args.trained_scale
        sec_step = lambda: model_step(qi2, qt2, mask)
        sec_graph = None
        if use_graph:                # replayed like the main case (at small batches eager issue would time Python)
            try:
                from ilfdt.graphs import GraphedStep
                sec_graph = GraphedStep(model_step, (qi2, qt2, mask), warmup=1)
                sec_step = sec_graph
            except Exception:  # noqa: BLE001 -- measure eagerly
                sec_graph = None
        for _ in range(2):
            sec_step()
        sync()
        if flush is None:
            e0.record()
            for _ in range(args.steps):
                sec_step()
            e1.record()
            sync()
            sms = e0.elapsed_time(e1) / args.steps
        else:
            sms = 0.0
            for _ in range(args.steps):
                flush.zero_()
                e0.record()
                sec_step()
                e1.record()
                sync()
                sms += e0.elapsed_time(e1) / args.steps
        if sec_graph is not None:
            sec_graph.release()
            del sec_graph
        ops.PROFILE.clear()
        ops.PROFILE_ON = True
        for _ in range(args.steps):
            model_step(qi2, qt2, mask)
        sync()
        ops.PROFILE_ON = False
        sec_kernel_ms = {k: v[1] / args.steps for k, v in ops.profile_summary().items()}
        ops.PROFILE.clear()
        with torch.no_grad():
            w_i, _ = Fn.GroundingFunction.apply(qi2[:min(b, 2048)], sd.detach(), None, TEMPERATURE, Fn.get_compute_dtype())
            support = float((w_i > 0).sum(1).float().mean().item())
        secondary = {"case": "trained-like scores (q x %g): mean sparsemax support %.1f codes of %d" % (
                         args.trained_scale, support, c["K"]),
                     "ms_per_step": sms, "value": B / (sms * 1e-3), "unit": "pairs/s",
                     "mean_support": support, "kernel_ms_per_step": sec_kernel_ms}
        del qi2, qt2, w_i

    qmap = None
    if kind == "fdt" and world == 1 and args.mode == "bf16" and not args.no_qmap and args.config == "headline":
        qmap = measure_q_map(c, dev, args.steps)
        qmap["step_with_q_map_ms"] = ms_per_step + qmap["ilfdt_bf16_ms"]
        qmap["step_with_eager_q_map_ms"] = ms_per_step + qmap["torch_fp32_eager_ms"]

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sample = args.cpu_sample or default_cpu_sample(c)
        cpu, _ = cpu_reference_run(c, sample, 4 if kind == "fdt" else 20, 1)   # ~10-20 s of host work
        try:
            cpu["gpu_eager"] = gpu_eager_reference_run(c, sample, 3, 1, dev)
        except Exception as e:  # noqa: BLE001 -- a context number must not take the bench line down
            cpu["gpu_eager"] = {"unavailable": str(e).splitlines()[0][:160]}

    if rank == 0:
        line = {"metric": c["metric"], "value": value, "unit": "pairs/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": c["scaling"],
                "vs_baseline": None, "dtype": args.mode, "data": "synthetic", "config": workload_config(args, c),
                "roofline": roof, "step_roofline": step_roofline, "cpu_baseline": cpu, "e2e": e2e,
                "gpu_launches": launches_per_step * args.steps, "clocks": clk, "loss": loss_val,
                "kernel_ms_per_step": kernel_ms, "cuda_graph": graph_note, "secondary": secondary, "q_map": qmap,
                "step_ms_spread": {"min": min(per_step), "median": statistics.median(per_step), "max": max(per_step)},
                "l2_flush_ms_per_step": flush_ms, "us_per_step": ms_per_step * 1e3,
                "host_issue_ms_per_step": host_ms_per_step}
        print(json.dumps(line), flush=True)
    # ---- teardown: graphs that captured NCCL collectives must die before their communicator does ----
    if gstep is not None:
        run_step = None
        gstep.release()
        del gstep
    gc.collect()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
