"""Drop-in for the reference's prototype/model/clip_fdt.py: same class names, constructor and call
signatures, parameter names and shapes (checkpoints interchange), with the hot path --
Query_model from q onward, L2-normalise, all-gather, logits -- running in libilfdt's sm_100a kernels.

What stays PyTorch: the q_map MLP (LN-Linear-GELU-LN-Linear), which produces
the kernels' input q, and the encoders, which are out of scope.
"""
import math

import numpy as np
import torch
import torch.nn as nn

from .. import functional as Fn
from ..functional import AllGather  # noqa: F401  (re-exported: clip_fdt.py:164-188)
from ..lazy import ContrastiveStash, LazyLogits
from .. import ops
from .sparsemax import Sparsemax


def weight_reset(m):
    """clip_fdt.py:36-44: ONLY Conv2d / Linear / ConvTranspose2d / BatchNorm2d (+ running stats) / LayerNorm are
    re-initialised. Everything else -- nn.Embedding (the text encoder's token_embedding, initialised with std 0.02),
    nn.MultiheadAttention's in_proj, bare nn.Parameters -- survives a reset, exactly as in the reference."""
    if isinstance(m, (nn.Conv2d, nn.Linear, nn.ConvTranspose2d, nn.LayerNorm)):
        m.reset_parameters()
    elif isinstance(m, nn.BatchNorm2d):
        m.reset_parameters()
        m.running_mean.zero_()
        m.running_var.fill_(1)


class Query_model(nn.Module):
    """clip_fdt.py:54-161. forward(ft [b,N,ft_dim], sd [K,sd_dim], mask [b,N]=None,
    return_token_att=False) -> (att_weight [b,K] | token scores [b,N,K], att_ft [b,sd_dim], sd)."""

    def __init__(self, ft_dim, sd_dim, temperature=1, att_func_type='softmax', pool_type='sum'):
        super().__init__()
        assert att_func_type in ['softmax', 'sigmoid', 'sparsemax']
        assert pool_type in ['mean', 'max', 'sum']
        self.att_func_type = att_func_type
        self.pool_type = pool_type
        # kept for state_dict / attribute compatibility (clip_fdt.py:71-78); the arithmetic runs in libilfdt
        self.att_activation = Sparsemax(dim=-1) if att_func_type == 'sparsemax' else (
            nn.Softmax(dim=-1) if att_func_type == 'softmax' else nn.Sigmoid())
        self.att_dim = sd_dim
        self.temperature = temperature   # plain Python float, rewritten by the solver (train_solver.py:363-364)
        self.q_map = nn.Sequential(
            nn.LayerNorm(ft_dim),
            nn.Linear(ft_dim, sd_dim),
            nn.GELU(),
            nn.LayerNorm(sd_dim),
            nn.Linear(sd_dim, sd_dim),
        )

    def forward(self, ft, sd, mask=None, return_token_att=False, need_weights=True):
        # need_weights=False (an addition to the reference signature, used by Clip_FDT.forward which drops
        # att_weight): the tuned route returns att_weight as its bf16 tensor-core copy, not differentiable
        if Fn.fused_q_map_enabled():
            # the MLP on the library's kernels, q handed over in bf16 (same parameters, same names)
            q = Fn.q_map_apply(self.q_map, ft)
        else:
            q = self.q_map(ft)
        if mask is not None:
            assert mask.shape == q.shape[:2]
        if self.att_func_type == 'sparsemax' and self.pool_type == 'max':
            # the shipped configuration (config_cc3m.yaml:14-21): tuned tensor-core path
            att_weight, att_ft = Fn.GroundingFunction.apply(q, sd, mask, float(self.temperature),
                                                            Fn.get_compute_dtype(), need_weights)
        else:
            att_weight, att_ft = Fn.GroundingVariantFunction.apply(q, sd, mask, float(self.temperature),
                                                                   self.att_func_type, self.pool_type)
        if return_token_att:
            return self._token_scores(q, sd, mask), att_ft, sd
        return att_weight, att_ft, sd

    def _token_scores(self, q, sd, mask):
        """Visualisation-only route (clip_fdt.py:120-121,132-133): the un-pooled [b,N,K] scores have
        to exist here by definition. Raw dot products without a mask, /sqrt(d) and pad-zeroed with one."""
        b, n, d = q.shape
        with torch.no_grad():
            s = ops.gemm_f32(q.detach().reshape(b * n, d).contiguous().float(), sd.detach().contiguous().float(),
                             b_t=True).view(b, n, -1)
            if mask is not None:
                s = s / math.sqrt(self.att_dim) * ((mask == 0) * 1).unsqueeze(-1)
        return s


class Clip_FDT(nn.Module):
    """clip_fdt.py:193-428 with the same constructor and attribute names."""

    def __init__(self, image_encode, text_encode, use_allgather, sd_num, sd_dim, raw_img_ft_dim, raw_txt_ft_dim,
                 att_func_type, pool_type, sd_temperature):
        super().__init__()
        self.use_allgather = use_allgather
        self.visual = image_encode
        self.encode_text = text_encode
        self.space_dict = nn.Parameter(torch.randn(sd_num, sd_dim))
        self.img_query_model = Query_model(ft_dim=raw_img_ft_dim, sd_dim=sd_dim, temperature=sd_temperature,
                                           att_func_type=att_func_type, pool_type=pool_type)
        self.txt_query_model = Query_model(ft_dim=raw_txt_ft_dim, sd_dim=sd_dim, temperature=sd_temperature,
                                           att_func_type=att_func_type, pool_type=pool_type)
        self.logit_scale = nn.Parameter(torch.ones([1]))
        self.logit_scale_sd = nn.Parameter(torch.ones([1]))   # declared-but-unused in the reference too (:227)
        nn.init.constant_(self.logit_scale, np.log(1 / 0.07))
        nn.init.constant_(self.logit_scale_sd, np.log(1 / 0.07))
        self.stored_vision_encoder_weight = None
        self.stored_codebook = None
        # (seed, generation) key the Philox stream of reset_codebook(); they travel with checkpoints as the optional
        # state_dict entry `codebook_rng` (see _save_to_state_dict) so that a resumed run does not redraw the
        # codebooks of earlier generations
        self.codebook_generation = 0
        self.codebook_seed = 0

    _RNG_KEY = "codebook_rng"

    def _save_to_state_dict(self, destination, prefix, keep_vars):
        """The key set equals the reference's (checkpoints interchange) until reset_codebook() has run; from then on
        one extra int64[2] entry carries (seed, generation). The reference's loader is non-strict
        (prototype/utils/misc.py:495) and ignores it."""
        super()._save_to_state_dict(destination, prefix, keep_vars)
        if self.codebook_generation or self.codebook_seed:
            destination[prefix + self._RNG_KEY] = torch.tensor([self.codebook_seed, self.codebook_generation],
                                                              dtype=torch.int64)

    def _load_from_state_dict(self, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys,
                              error_msgs):
        rng = state_dict.pop(prefix + self._RNG_KEY, None)   # optional: absent from reference checkpoints
        if rng is not None:
            self.codebook_seed, self.codebook_generation = int(rng[0]), int(rng[1])
        super()._load_from_state_dict(state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys,
                                      error_msgs)

    # ---- helpers the solver / eval adapters call (clip_fdt.py:236-386) --------------------------
    @property
    def dtype(self):
        for probe in (lambda: self.visual.conv1.weight, lambda: self.visual.head.weight,
                      lambda: self.visual.stem[0].weight):
            try:
                return probe().dtype
            except Exception:
                continue
        return self.encode_text.text_projection.weight.dtype

    def find_always_freeze_weight(self):
        self.weight_always_freeze = [n for n, p in self.named_parameters() if not p.requires_grad]
        print("always freeze weight", self.weight_always_freeze)

    def reset_text_encoder(self, seed):
        torch.manual_seed(seed)
        self.encode_text.apply(weight_reset)
        self.txt_query_model.apply(weight_reset)

    def reset_vision_encoder(self):
        self.visual.apply(weight_reset)
        self.img_query_model.apply(weight_reset)

    def swap_vision_encoder(self):
        current = (self.visual.state_dict().copy(), self.img_query_model.state_dict().copy())
        if self.stored_vision_encoder_weight is None:
            self.visual.apply(weight_reset)
            self.img_query_model.apply(weight_reset)
        else:
            self.visual.load_state_dict(self.stored_vision_encoder_weight[0])
            self.img_query_model.load_state_dict(self.stored_vision_encoder_weight[1])
        self.stored_vision_encoder_weight = current

    def reset_codebook(self, seed=None):
        """clip_fdt.py:281-283 re-draws space_dict ~ N(0,1) -- on the HOST, leaving a CPU tensor in a
        CUDA model. Here the refill happens in place on the device with a counter-based
        Philox stream keyed by (seed, generation): every rank draws identical values without a
        broadcast, and the parameter keeps its storage (optimizer state / DDP buckets stay valid)."""
        if seed is not None:
            self.codebook_seed = int(seed)
        self.codebook_generation += 1
        if self.space_dict.is_cuda:
            ops.randn_fill_(self.space_dict.data, self.codebook_seed, self.codebook_generation)
        else:  # model not moved to the GPU yet: nothing of the hot path can run there anyway
            g = torch.Generator().manual_seed(self.codebook_seed * 1000003 + self.codebook_generation)
            self.space_dict.data.copy_(torch.randn(self.space_dict.shape, generator=g))

    def store_codebook_value(self):
        """train_solver.py:211-212 (`.data.clone()`), as a device-to-device copy into a kept buffer."""
        if self.stored_codebook is None or self.stored_codebook.shape != self.space_dict.shape \
                or self.stored_codebook.device != self.space_dict.device:
            self.stored_codebook = torch.empty_like(self.space_dict.data)
        if self.space_dict.is_cuda:
            ops.copy_(self.stored_codebook, self.space_dict.data.contiguous())
        else:
            self.stored_codebook.copy_(self.space_dict.data)

    def keep_codebook_value(self):
        """train_solver.py:214-215 re-binds `.data` to the stored tensor (aliasing it, F8); the intended
        semantics -- pin the codebook to the stored values -- is a copy."""
        if self.stored_codebook is None:
            raise RuntimeError("keep_codebook_value() before store_codebook_value()")
        if self.space_dict.is_cuda:
            ops.copy_(self.space_dict.data, self.stored_codebook)
        else:
            self.space_dict.data.copy_(self.stored_codebook)

    def unfreeze_weights(self, module_names, freeze_codebook=False):
        for encoder_name in module_names:
            for name, param in getattr(self, encoder_name).named_parameters():
                if name not in self.weight_always_freeze:
                    param.requires_grad = True
                else:
                    print("not unfreezing parameter named", name)
        self.logit_scale.requires_grad = True
        self.logit_scale_sd.requires_grad = True
        self.space_dict.requires_grad = not freeze_codebook

    def unfreeze_all_parameters(self):
        for _, param in self.named_parameters():
            param.requires_grad = True

    def freeze_weights(self, module_names, freeze_codebook=False):
        for encoder_name in module_names:
            for _, param in getattr(self, encoder_name).named_parameters():
                param.requires_grad = False
        self.logit_scale.requires_grad = False
        self.logit_scale_sd.requires_grad = False
        self.space_dict.requires_grad = not freeze_codebook

    def freeze_unfreeze_vision_weights(self, unfreeze, freeze_codebook):
        names = ["visual", "img_query_model"]
        (self.unfreeze_weights if unfreeze else self.freeze_weights)(names, freeze_codebook)

    def freeze_unfreeze_text_weights(self, unfreeze, freeze_codebook):
        names = ["encode_text", "txt_query_model"]
        (self.unfreeze_weights if unfreeze else self.freeze_weights)(names, freeze_codebook)

    def _image_features(self, images):
        info = self.encode_image(images)
        return info[1]  # (cls, patch_ft[, raw]) -- visual_transformer.py:72-91; Swin returns 2

    def extract_img_sd_ft(self, images, return_token_att=False):
        patch_ft = self._image_features(images)
        return self.img_query_model(patch_ft, self.space_dict, return_token_att=return_token_att)

    def extract_patch_ft(self, images):
        return self.img_query_model.q_map(self._image_features(images))

    def extract_word_ft(self, texts):
        _, word_ft, _, pad_mask = self.encode_text(texts, return_dense=True, return_padmask=True,
                                                   return_raw_feature=True)
        return self.txt_query_model.q_map(word_ft), pad_mask

    def extract_txt_sd_ft(self, texts, return_token_att=False, raw_text=True):
        _, word_ft, _, pad_mask = self.encode_text(texts, return_dense=True, return_padmask=True,
                                                   return_raw_feature=True, raw_text=raw_text)
        return self.txt_query_model(word_ft, self.space_dict, mask=pad_mask, return_token_att=return_token_att)

    def encode_image(self, image):
        return self.visual(image.type(self.dtype), return_dense=True, return_raw_feature=True)

    def all_gather(self, input):
        output = AllGather.apply(input)
        return output.view(-1, *(output.shape[2:]))

    # ---- the hot path (clip_fdt.py:390-428) --------------------------------------------------------
    def forward(self, images, texts):
        patch_ft = self._image_features(images)
        _, word_ft, _, pad_mask = self.encode_text(texts, return_dense=True, return_padmask=True,
                                                   return_raw_feature=True)
        mode = Fn.get_compute_dtype()
        lp = Fn.lp_embeddings(mode, self.space_dict.shape[1])
        with Fn.codebook_scope():   # both Query_models read the same codebook values: one bf16 copy for the two
            _, sd_img_ft, img_k = self.img_query_model(patch_ft, self.space_dict, need_weights=False)
            # :411 + :418 for the image side right away: its all-gather (bf16 rows in the tensor-core mode) travels
            # over NVLink while the text side is still being grounded
            sd_img_ft, gathered_img = self._normalise_and_gather(sd_img_ft, lp)
            _, sd_txt_ft, txt_k = self.txt_query_model(word_ft, self.space_dict, mask=pad_mask, need_weights=False)
        sd_txt_ft, gathered_txt = self._normalise_and_gather(sd_txt_ft, lp)       # :412, :419

        # :415-416 -- forward value clamped to 100 through .data, backward through the unclamped exp (F7)
        logit_scale = self.logit_scale.exp()
        logit_scale.data = torch.clamp(logit_scale.data, max=100)

        # :421-422 -- the two [b,B] logit matrices are returned lazily: ClipInfoCELoss consumes the stash
        # through the fused kernels; any other use materialises them with the logits kernel.
        stash = ContrastiveStash(sd_img_ft, sd_txt_ft, gathered_img, gathered_txt, logit_scale, mode, gathered=True)
        return (LazyLogits(stash, 0), LazyLogits(stash, 1)), (img_k, txt_k)

    @staticmethod
    def _normalise_and_gather(ft, lp):
        if lp:
            x, xb = Fn.L2NormFunction.apply(ft, 1e-10, True)
            return x, Fn.GatherHandle(xb)
        x = Fn.L2NormFunction.apply(ft, 1e-10)
        return x, Fn.GatherHandle(x.detach())


def _reference_encoders():
    """The encoders are outside the hot path and are taken from the reference tree when this module
    is used under example/clip_fdt/train_solver.py."""
    try:
        from prototype.model.image_encoder.visual_transformer import visual_transformer_B16, visual_transformer_B32
        from prototype.model.text_encoder.text_transformer import text_transformers
    except Exception as e:  # pragma: no cover
        raise ImportError("clip_fdt_vitb32/16 need the reference's encoders (prototype.model.*) on sys.path") from e
    return visual_transformer_B32, visual_transformer_B16, text_transformers


def clip_fdt_vitb32(**kwargs):
    b32, _, txt = _reference_encoders()
    return Clip_FDT(b32(**kwargs['image_encode']), txt(**kwargs['text_encode']), **kwargs['fdt'])


def clip_fdt_vitb16(**kwargs):
    _, b16, txt = _reference_encoders()
    return Clip_FDT(b16(**kwargs['image_encode']), txt(**kwargs['text_encode']), **kwargs['fdt'])
