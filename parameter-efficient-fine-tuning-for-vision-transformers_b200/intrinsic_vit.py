"""Per-component intrinsic-dimension model — drop-in for full_shot/main/lib/models/cls_intrinsic_dimension.py.

Every block owns, next to its frozen `attn` / `mlp` / `adapter`, three Fastfood-wrapped stand-alone sub-modules
(`intrinsic_attn`, `intrinsic_mlp`, `intrinsic_adapter`, reference :466-481); block number `layernum` runs the wrapped module
named by `layerType` ("attention" | "mlp" | "adapter") in place of the frozen one (:483-510, forward_features :779-786), every
other block runs plain.  The trainer (full_shot/main/tools/intrinsic_dimension.py:380-385) trains the parameters whose name
contains "intrinsic", i.e. the d-dimensional `V` vectors; only the one that is used receives a gradient.

Execution: sub-module granularity (eager.py).  Blocks before `layernum` see an input that needs no gradient, so autograd skips
their backward entirely; the wrapped component needs dW for its own weights (theta depends on V: SURVEY §8a-R12), blocks after
it only propagate dX.  As in the reference the wrapped module starts from its OWN initialisation (theta0 is hidden from
state_dict(), so the checkpoint transplant does not reach it).  Same names: `blocks.{i}.intrinsic_attn.V` etc.
"""
from __future__ import annotations

from functools import partial

import torch
import torch.nn as nn

from . import eager
from .fastfood import FastfoodWrap
from .timm_vit import Adapter, Attention, Mlp, PatchEmbed
from .timm_vit import Block as _PlainBlock
from .timm_vit import VisionTransformer as _PlainViT


class Block(_PlainBlock):
    """cls_intrinsic_dimension.py:423-521 (pre-norm)."""

    def __init__(self, dim, num_heads, mlp_ratio=4., qkv_bias=False, qk_scale=None, drop=0., attn_drop=0., drop_path=0.,
                 act_layer=nn.GELU, norm_layer=nn.LayerNorm, pre_norm=True, res_score=False, dintrinsic=100, device=0):
        super().__init__(dim, num_heads, mlp_ratio=mlp_ratio, qkv_bias=qkv_bias, qk_scale=qk_scale, drop=drop, attn_drop=attn_drop,
                         drop_path=drop_path, act_layer=act_layer, norm_layer=norm_layer, pre_norm=pre_norm, res_score=res_score,
                         adapter="houlsby")       # plain attention + mlp + a (frozen, unused: reference :511-515) Adapter
        if drop_path != 0.:
            raise NotImplementedError("stochastic depth is not supported by the intrinsic-dimension blocks")
        dev = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        mk = lambda m: FastfoodWrap(m.to(dev), intrinsic_dimension=dintrinsic, device=device)
        self.intrinsic_attn = mk(Attention(dim, num_heads=num_heads, qkv_bias=qkv_bias, adapter="none"))
        self.intrinsic_mlp = mk(Mlp(in_features=dim, hidden_features=int(dim * mlp_ratio), act_layer=act_layer))
        self.intrinsic_adapter = mk(Adapter(dim, down_sample=64))

    def forward(self, x, prev=None, measure_idim=None):
        for m in (self.attn, self.mlp, self.adapter, self.intrinsic_attn.m[0], self.intrinsic_mlp.m[0], self.intrinsic_adapter.m[0]):
            m.fp32 = self.fp32
        if measure_idim not in (None, "attention", "mlp", "adapter"):
            raise ValueError(f"measure_idim must be attention | mlp | adapter | None, got {measure_idim!r}")
        attn = self.intrinsic_attn if measure_idim == "attention" else self.attn
        x = x + attn(self._norm(self.norm1, x))
        h = (self.intrinsic_mlp if measure_idim == "mlp" else self.mlp)(self._norm(self.norm2, x))
        if measure_idim == "adapter":
            h = self.intrinsic_adapter(h)[0]
        return x + h


class VisionTransformer(nn.Module):
    """cls_intrinsic_dimension.py:591-807; forward(x) = head(final-LayerNorm cls feature) as in the reference (:804-807)."""

    def __init__(self, img_size=224, patch_size=16, in_chans=3, num_classes=1000, embed_dim=768, depth=12, num_heads=12,
                 mlp_ratio=4., qkv_bias=False, qk_scale=None, drop_rate=0., attn_drop_rate=0., drop_path_rate=0.,
                 hybrid_backbone=None, norm_layer=None, use_cls_tocken=True, norm_embed=False, pre_norm=True, res_score=False,
                 init='trunc_norm', dintrinsic=100, layerType="mlp", layernum=100, device=0):
        super().__init__()
        if hybrid_backbone is not None or norm_embed or not use_cls_tocken or in_chans != 3 or not pre_norm:
            raise NotImplementedError("hybrid backbone / norm_embed / no-cls-token / in_chans != 3 / post-norm are not supported")
        if drop_rate != 0. or attn_drop_rate != 0. or drop_path_rate != 0.:
            raise NotImplementedError("dropout / stochastic depth are not supported by the intrinsic-dimension model")
        norm_layer = norm_layer or partial(nn.LayerNorm, eps=1e-6)
        self.num_classes = num_classes
        self.num_features = self.embed_dim = embed_dim
        self.patch_embed = PatchEmbed(img_size=img_size, patch_size=patch_size, in_chans=in_chans, embed_dim=embed_dim)
        self.cls_token = nn.Parameter(torch.zeros(1, 1, embed_dim))
        self.pos_embed = nn.Parameter(torch.zeros(1, self.patch_embed.num_patches + 1, embed_dim))
        self.blocks = nn.ModuleList([
            Block(dim=embed_dim, num_heads=num_heads, mlp_ratio=mlp_ratio, qkv_bias=qkv_bias, qk_scale=qk_scale,
                  norm_layer=norm_layer, pre_norm=pre_norm, res_score=res_score, dintrinsic=dintrinsic, device=device)
            for _ in range(depth)])
        self.norm = norm_layer(embed_dim)
        self.head = nn.Linear(embed_dim, num_classes) if num_classes > 0 else nn.Identity()
        nn.init.trunc_normal_(self.cls_token, std=.02)
        nn.init.trunc_normal_(self.pos_embed, std=.02)
        self.apply(_PlainViT._init_weights_xavier if init == 'xavier' else _PlainViT._init_weights_trunc_normal)
        self.layerType, self.layernum = layerType, layernum
        self.fp32 = False

    @torch.jit.ignore
    def no_weight_decay(self):
        return {'pos_embed', 'cls_token'}

    def forward_features(self, x):
        B = x.shape[0]
        self.patch_embed.fp32 = self.fp32
        x = self.patch_embed(x)
        x = torch.cat((self.cls_token.expand(B, -1, -1), x), dim=1) + self.pos_embed
        for i, blk in enumerate(self.blocks):
            blk.fp32 = self.fp32
            x = blk(x, None, self.layerType if i == self.layernum else None)
        return eager.layer_norm(x[:, 0], self.norm.weight, self.norm.bias, self.norm.eps)      # LayerNorm is per token

    def forward(self, x):
        return self.head(self.forward_features(x))


def get_cls_model(config, dintrinsic, layerType, layernum, **kwargs):
    """cls_intrinsic_dimension.py:810-841 (yacs config or any object with the same attributes)."""
    spec = config.MODEL.SPEC
    return VisionTransformer(img_size=config.TRAIN.IMAGE_SIZE[0], patch_size=spec.PATCH_SIZE, num_classes=config.MODEL.NUM_CLASSES,
                             embed_dim=spec.EMBED_DIM, qkv_bias=spec.QKV_BIAS, depth=spec.DEPTH, num_heads=spec.NUM_HEADS,
                             mlp_ratio=spec.MLP_RATIO, drop_rate=spec.DROP_RATE, attn_drop_rate=spec.ATTN_DROP_RATE,
                             drop_path_rate=spec.DROP_PATH_RATE, norm_layer=partial(nn.LayerNorm, eps=1e-6),
                             use_cls_tocken=spec.USE_CLS_TOKEN, norm_embed=getattr(spec, 'NORM_EMBED', False),
                             pre_norm=getattr(spec, 'PRE_NORM', True), res_score=getattr(spec, 'RES_SCORE', False),
                             init=getattr(spec, 'INIT', 'trunc_norm'), dintrinsic=dintrinsic, layerType=layerType, layernum=layernum,
                             **kwargs)
