"""python tools/step_once.py [--batch 32768] [--codes 16384] [--ptok 49] [--scale 1.0] [--steps 2] [--mode bf16]
A few steps of the hot path through the reference-facing modules (Clip_FDT.forward + ClipInfoCELoss + backward), sized
for profiling: `ncu --metrics gpu__time_duration.sum --clock-control none ... python tools/step_once.py` gives the launch
list of exactly the kernels bench.py times. --scale 200 = the trained-like (peaked attention) secondary case."""
import argparse, math, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "iterated-learning-for-vlm_b200"))
import ilfdt
from ilfdt.loss_functions import ClipInfoCELoss
from ilfdt.model.clip_fdt import Clip_FDT

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=32768)
ap.add_argument("--codes", type=int, default=16384)
ap.add_argument("--ptok", type=int, default=49)
ap.add_argument("--scale", type=float, default=1.0)
ap.add_argument("--steps", type=int, default=2)
ap.add_argument("--mode", default="bf16")
a = ap.parse_args()
ilfdt.set_compute_dtype(a.mode)
dev = torch.device("cuda:0")
b, P, T, K, D = a.batch, a.ptok, 77, a.codes, 512
g = torch.Generator(device=dev).manual_seed(1000)
q_img = torch.randn(b, P, D, device=dev, generator=g) * a.scale
q_txt = torch.randn(b, T, D, device=dev, generator=g) * a.scale
lens = torch.randint(5, T + 1, (b,), device=dev, generator=g)
mask = torch.zeros(b, T, device=dev)
mask[torch.arange(T, device=dev).view(1, -1) >= lens.view(-1, 1)] = float("-inf")


class _Visual(torch.nn.Module):
    def forward(self, image, return_dense=False, return_raw_feature=False):
        return None, image, None


class _Text(torch.nn.Module):
    def forward(self, texts, return_dense=False, return_padmask=False, return_raw_feature=False, raw_text=True):
        return [None, texts[0], None, texts[1]]


class _Identity(torch.nn.Module):
    def forward(self, x):
        return x


model = Clip_FDT(_Visual(), _Text(), True, K, D, D, D, "sparsemax", "max", 1000.0).to(dev)
model.img_query_model.q_map = _Identity()
model.txt_query_model.q_map = _Identity()
model.encode_image = lambda image: model.visual(image, return_dense=True, return_raw_feature=True)
crit = ClipInfoCELoss()
for i in range(a.steps):
    qi, qt = q_img.detach().requires_grad_(True), q_txt.detach().requires_grad_(True)
    model.zero_grad(set_to_none=True)
    (l_i, l_t), _ = model(qi, (qt, mask))
    loss, _ = crit(l_i, l_t)
    loss.backward()
torch.cuda.synchronize()
print("loss %.6f" % loss.item())
