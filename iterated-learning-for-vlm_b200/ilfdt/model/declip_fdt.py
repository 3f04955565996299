"""Drop-in for the FDT part of the reference's prototype/model/declip_fdt.py.

DeCLIP-FDT carries its own copy of `Query_model` (declip_fdt.py:120-210) and a contrastive head that forms EIGHT logit
matrices from two image views x {masked, EDA-augmented} captions (declip_fdt.py:322-430). Both are pointed at the same
kernels as Clip_FDT: grounding through `GroundingFunction`, and the eight logits as four lazy pairs
(`ilfdt.lazy.contrastive_logits`) whose InfoNCE runs fused when handed to `ClipInfoCELoss`.

What is NOT rebuilt (outside the hot path): the EDA text augmentation (`textaugment`), the MLM head,
the SimSiam projector / predictor MLPs and the nearest-neighbour banks. `DECLIP_FDT.forward` therefore takes the two
image views and the two caption views already encoded -- `fdt_head(...)` is the part of declip_fdt.py:356-399 from the
encoder outputs to the `logits` / `logits_aug` / `features` entries of its `ret_dict`; a DeCLIP solver keeps its own
augmentation and auxiliary heads and calls this for the FDT contrastive terms.
"""
import math

import torch
import torch.nn as nn

from .. import functional as Fn
from .. import linklink as link
from .. import ops
from ..lazy import contrastive_logits
from .clip import CLIP
from .clip_fdt import Query_model as _ClipFdtQueryModel


class Query_model(_ClipFdtQueryModel):
    """declip_fdt.py:120-210: the same module as clip_fdt.py's (same parameters, same pooled path) EXCEPT for
    `return_token_att`, which here returns the per-token ACTIVATED attention -- `att_activation` applied over the codes
    of every token row of the scaled, pad-zeroed, temperature-divided scores (declip_fdt.py:183-184) -- where clip_fdt.py
    returns the raw scores."""

    def _token_scores(self, q, sd, mask):
        b, n, d = q.shape
        with torch.no_grad():
            s = ops.gemm_f32(q.detach().reshape(b * n, d).contiguous().float(), sd.detach().contiguous().float(),
                             b_t=True).view(b, n, -1)
            s = s / math.sqrt(self.att_dim)                                   # :169
            if mask is not None:
                s = s * ((mask == 0) * 1).unsqueeze(-1)                       # :176-179
            s = (s / self.temperature).contiguous()                           # :181
            flat = s.view(b * n, -1)
            if self.att_func_type == 'sparsemax':
                out = ops.sparsemax_fwd(flat)[0]
            else:
                out = ops.att_fwd(flat, self.att_func_type)[0]
        return out.view(b, n, -1)


class DECLIP_FDT(CLIP):
    """declip_fdt.py:212-470, FDT part. Parameter names (`space_dict`, `img_query_model.*`, `txt_query_model.*`,
    `logit_scale`) equal the reference's, so its checkpoints load (auxiliary heads: strict=False)."""

    def __init__(self, image_encode, text_encode, use_allgather, sd_num, sd_dim, sd_temperature, pool_type, att_func_type,
                 raw_img_ft_dim=768, raw_txt_ft_dim=512, **unused_auxiliary_head_options):
        super().__init__(image_encode, text_encode, use_allgather)
        self.space_dict = nn.Parameter(torch.randn(sd_num, sd_dim))
        self.img_query_model = Query_model(ft_dim=raw_img_ft_dim, sd_dim=sd_dim, temperature=sd_temperature,
                                           att_func_type=att_func_type, pool_type=pool_type)
        self.txt_query_model = Query_model(ft_dim=raw_txt_ft_dim, sd_dim=sd_dim, temperature=sd_temperature,
                                           att_func_type=att_func_type, pool_type=pool_type)

    def encode_image(self, image, return_dense=False):
        return self.visual(image.type(self.dtype), return_dense=True, return_raw_feature=True)   # :257-263

    def extract_img_sd_ft(self, images, return_token_att=False):
        _, patch_ft, _ = self.encode_image(images)
        img_att, sd_img_ft, _ = self.img_query_model(patch_ft, self.space_dict, return_token_att=return_token_att)
        return sd_img_ft, img_att                                                                 # :265-272 (ft first)

    def extract_txt_sd_ft(self, texts, return_token_att=False):
        _, word_ft, _, pad_mask = self.encode_text(texts, return_dense=True, return_padmask=True, return_raw_feature=True)
        txt_att, sd_txt_ft, _ = self.txt_query_model(word_ft, self.space_dict, mask=pad_mask,
                                                     return_token_att=return_token_att)
        return sd_txt_ft, txt_att                                                                 # :284-292

    def ground(self, patch_ft_1, patch_ft_2, mask_word_ft, mask_pad_masks, aug_word_ft, aug_pad_mask):
        """declip_fdt.py:352-368: the four groundings of a step share one bf16 copy of the codebook."""
        with Fn.codebook_scope():
            _, sd_img_ft_1, _ = self.img_query_model(patch_ft_1, self.space_dict, need_weights=False)
            _, sd_img_ft_2, _ = self.img_query_model(patch_ft_2, self.space_dict, need_weights=False)
            _, mask_sd_txt_ft, _ = self.txt_query_model(mask_word_ft, self.space_dict, mask=mask_pad_masks, need_weights=False)
            _, aug_sd_txt_ft, _ = self.txt_query_model(aug_word_ft, self.space_dict, mask=aug_pad_mask, need_weights=False)
        return sd_img_ft_1, sd_img_ft_2, mask_sd_txt_ft, aug_sd_txt_ft

    def fdt_head(self, sd_img_ft_1, sd_img_ft_2, mask_sd_txt_ft, aug_sd_txt_ft):
        """declip_fdt.py:381-412,431-436: normalise (image sides WITHOUT an epsilon, text sides with 1e-10), clamp the
        scale, all-gather, and the eight logits -- returned as the reference's ret_dict entries. Every logits_per_image_X /
        logits_per_text_X couple is one lazy pair: ClipInfoCELoss(pair) runs the fused kernels, no [b,B] matrix is stored."""
        if not (self.training and self.use_allgather):
            raise NotImplementedError('2-View: Not Implemented')                                  # :429-430
        mode = Fn.get_compute_dtype()
        lp = Fn.lp_embeddings(mode, sd_img_ft_1.shape[1])
        link.barrier()                                                                            # :391
        img = [self._normalise_and_gather(f, 0.0, lp) for f in (sd_img_ft_1, sd_img_ft_2)]        # :382-383
        txt = [self._normalise_and_gather(f, 1e-10, lp) for f in (mask_sd_txt_ft, aug_sd_txt_ft)]  # :384-385
        logit_scale = self.logit_scale.exp()                                                      # :388-389
        logit_scale.data = torch.clamp(logit_scale.data, max=100)
        link.barrier()                                                                            # :396
        pair = lambda i, t: contrastive_logits(img[i][0], txt[t][0], img[i][1], txt[t][1], logit_scale, mode, gathered=True)
        (i1, t1), (i2, t2) = pair(0, 0), pair(1, 0)            # :399-400, :405-406  against the masked captions
        (i1a, t1a), (i2a, t2a) = pair(0, 1), pair(1, 1)        # :401-402, :407-408  against the augmented captions
        return {"logits": (i1, i2, t1, t2), "logits_aug": (i1a, i2a, t1a, t2a),
                "features": (txt[0][0], img[0][0], img[1][0])}                                    # :434-437

    def forward(self, patch_ft_1, patch_ft_2, mask_word_ft, mask_pad_masks, aug_word_ft, aug_pad_mask):
        """The FDT contrastive terms of declip_fdt.py:322-437 from the encoder outputs on (see the module docstring)."""
        return self.fdt_head(*self.ground(patch_ft_1, patch_ft_2, mask_word_ft, mask_pad_masks, aug_word_ft, aug_pad_mask))


def _reference_encoders():
    try:
        from prototype.model.image_encoder.visual_transformer import visual_transformer_B16, visual_transformer_B32
        from prototype.model.text_encoder.text_transformer import text_transformers
    except Exception as e:  # pragma: no cover
        raise ImportError("declip_fdt_vitb32 needs the reference's encoders (prototype.model.*) on sys.path") from e
    return visual_transformer_B32, visual_transformer_B16, text_transformers


def declip_fdt_vitb32(**kwargs):
    """declip_fdt.py:470-477"""
    b32, _, txt = _reference_encoders()
    return DECLIP_FDT(b32(**kwargs['image_encode']), txt(**kwargs['text_encode']), **kwargs['clip'])
