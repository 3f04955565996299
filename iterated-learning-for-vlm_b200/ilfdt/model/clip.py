"""Drop-in for the contrastive head of prototype/model/clip.py:46-149 (baseline CLIP)."""
import numpy as np
import torch
import torch.nn as nn

from .. import functional as Fn
from ..functional import AllGather
from ..lazy import ContrastiveStash, LazyLogits


class CLIP(nn.Module):
    def __init__(self, image_encode, text_encode, use_allgather):
        super().__init__()
        self.use_allgather = use_allgather
        self.visual = image_encode
        self.encode_text = text_encode
        self.logit_scale = nn.Parameter(torch.ones([1]))
        nn.init.constant_(self.logit_scale, np.log(1 / 0.07))

    @property
    def dtype(self):
        for probe in (lambda: self.visual.conv1.weight, lambda: self.visual.head.weight,
                      lambda: self.visual.stem[0].weight):
            try:
                return probe().dtype
            except Exception:
                continue
        return self.encode_text.text_projection.weight.dtype

    def encode_image(self, image, return_dense=False, return_att=False):
        return self.visual(image.type(self.dtype), return_dense)

    def all_gather(self, input):
        output = AllGather.apply(input)
        return output.view(-1, *(output.shape[2:]))

    def forward(self, images, texts):
        image_features = self.encode_image(images)
        text_features = self.encode_text(texts)
        # clip.py:133-134: the image side is normalised WITHOUT an epsilon, the text side with 1e-10 (F10)
        mode = Fn.get_compute_dtype()
        lp = Fn.lp_embeddings(mode, image_features.shape[1])
        image_features, gathered_image_features = self._normalise_and_gather(image_features, 0.0, lp)
        text_features, gathered_text_features = self._normalise_and_gather(text_features, 1e-10, lp)
        logit_scale = self.logit_scale.exp()                          # clip.py:137-138 (F7)
        logit_scale.data = torch.clamp(logit_scale.data, max=100)
        # clip.py:146-147 scales the features before the matmul; the kernels scale the product, which
        # differs by rounding only (tolerance stated in tests/test_parity_gpu.py)
        stash = ContrastiveStash(image_features, text_features, gathered_image_features, gathered_text_features,
                                 logit_scale, mode, gathered=True)
        return LazyLogits(stash, 0), LazyLogits(stash, 1)

    @staticmethod
    def _normalise_and_gather(ft, eps, lp):
        """clip.py:133-134 + :140-141; the all-gather starts at once and is awaited inside the fused loss."""
        if lp:
            x, xb = Fn.L2NormFunction.apply(ft, eps, True)
            return x, Fn.GatherHandle(xb)
        x = Fn.L2NormFunction.apply(ft, eps)
        return x, Fn.GatherHandle(x.detach())
