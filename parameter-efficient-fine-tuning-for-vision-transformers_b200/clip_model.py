"""CLIP-dialect visual tower with PEFT adapters — drop-in for the reference classes of
few_shot/vision_benchmark/evaluation/{model,lora_model,adapter_model,compacter_model}.py.

Same class names (`CLIP`, `VisionTransformer`, `Transformer`, `ResidualAttentionBlock`, `MultiheadAttention`), same
parameter names and shapes (`visual.conv1.weight`, `visual.class_embedding`, `visual.positional_embedding`,
`visual.ln_pre.*`, `visual.transformer.resblocks.{i}.{ln_1,attn.in_proj_weight,attn.in_proj_bias,attn.out_proj,ln_2,
mlp.c_fc,mlp.c_proj}.*`, `visual.ln_post.*`, `visual.proj`, adapter names per SURVEY §8b), the same builders
(`build_model / build_lora_model / build_adapter_model / build_compacter_model(state_dict)` inferring the geometry from
key shapes, FS/model.py:1378-1420), and `encode_image`.  `VisionTransformer.forward` runs the whole tower through the
CUDA engine; sub-modules only hold parameters and have no eager forward (no fallback).

Reference quirks reproduced in the engine: q is scaled by 1/sqrt(hd) BEFORE the q/v update is added (Q2)
and the update is placed through the raw [L,B,C] -> [B*H,L,hd] reshape (Q1).  adapter="kron" is the C1 definition of
SURVEY Q3 (Kronecker update substituted at FS/model.py:981-982); adapter="lora" is FS/model.py as committed.
The text tower is out of scope: `encode_text` raises.
"""
from __future__ import annotations

from collections import OrderedDict

import torch
import torch.nn as nn

from .timm_vit import (Adapter, HyperComplexAdapter, _no_eager, add_qv_adapter_params, bottleneck_cfg,
                       bottleneck_operands, kron_is_dense, lowrank_operands, lowrank_rank, run_tower)


class LayerNorm(nn.LayerNorm):
    """FS/model.py:156-162 (fp16-safe subclass; eps 1e-5)."""
    forward = _no_eager


class QuickGELU(nn.Module):
    """FS/model.py:165-167: x * sigmoid(1.702 x) — fused into the c_fc GEMM epilogue by the engine."""
    forward = _no_eager


class MultiheadAttention(nn.Module):
    """FS/model.py:629-787 parameters: packed in-projection + out_proj (+ q/v adapter factors)."""

    def __init__(self, embed_dim, num_heads, adapter="none", **adapter_kw):
        super().__init__()
        if embed_dim // num_heads != 64:
            raise NotImplementedError("the attention kernels are specialised for head_dim 64")
        self.embed_dim, self.num_heads = embed_dim, num_heads
        self.in_proj_weight = nn.Parameter(torch.empty(3 * embed_dim, embed_dim))
        self.in_proj_bias = nn.Parameter(torch.zeros(3 * embed_dim))
        self.out_proj = nn.Linear(embed_dim, embed_dim)
        nn.init.xavier_uniform_(self.in_proj_weight)
        nn.init.zeros_(self.out_proj.bias)
        add_qv_adapter_params(self, embed_dim, adapter, **adapter_kw)

    forward = _no_eager


class ResidualAttentionBlock(nn.Module):
    """FS/model.py:1135-1158; Houlsby FS/adapter_model.py:298-334; Compacter FS/compacter_model.py:428-479."""

    def __init__(self, d_model, n_head, attn_mask=None, adapter="none", bottleneck=64, compacter_n_down=32,
                 compacter_n_up=4, **adapter_kw):
        super().__init__()
        if attn_mask is not None:
            raise NotImplementedError("the visual tower has no attention mask (FS/model.py:1197)")
        self.attn = MultiheadAttention(d_model, n_head, adapter=adapter, **adapter_kw)
        self.ln_1 = LayerNorm(d_model)
        self.mlp = nn.Sequential(OrderedDict([("c_fc", nn.Linear(d_model, d_model * 4)), ("gelu", QuickGELU()),
                                              ("c_proj", nn.Linear(d_model * 4, d_model))]))
        self.ln_2 = LayerNorm(d_model)
        if adapter == "houlsby":
            self.adapter = Adapter(d_model, down_sample=bottleneck)
        elif adapter == "compacter":
            self.compacter = HyperComplexAdapter(d_model, down_sample=bottleneck, n_down=compacter_n_down, n_up=compacter_n_up)

    forward = _no_eager


class Transformer(nn.Module):
    """FS/model.py:1161-1178."""

    def __init__(self, width, layers, heads, attn_mask=None, **adapter_kw):
        super().__init__()
        self.width, self.layers, self.heads = width, layers, heads
        self.resblocks = nn.Sequential(*[ResidualAttentionBlock(width, heads, attn_mask, **adapter_kw) for _ in range(layers)])

    forward = _no_eager


class VisionTransformer(nn.Module):
    """FS/model.py:1181-1219.  forward(images [B,3,R,R]) -> [B, output_dim] (or [B, width] if `proj` is None)."""

    _ADAPTER_KEYS = ("phm_rule", "W_left", "W_right", "phmb", "_proj_adapter", ".adapter.", ".compacter.")

    def __init__(self, input_resolution, patch_size, width, layers, heads, output_dim, adapter="none", **adapter_kw):
        super().__init__()
        self.input_resolution = input_resolution
        self.output_dim = output_dim
        self.adapter_kind = adapter
        self.conv1 = nn.Conv2d(3, width, kernel_size=patch_size, stride=patch_size, bias=False)
        scale = width ** -0.5
        self.class_embedding = nn.Parameter(scale * torch.randn(width))
        self.positional_embedding = nn.Parameter(scale * torch.randn((input_resolution // patch_size) ** 2 + 1, width))
        self.ln_pre = LayerNorm(width)
        self.transformer = Transformer(width, layers, heads, adapter=adapter, **adapter_kw)
        self.ln_post = LayerNorm(width)
        self.proj = nn.Parameter(scale * torch.randn(width, output_dim))
        self._geom = dict(img=input_resolution, patch=patch_size, width=width, depth=layers, heads=heads, mlp=4 * width)
        self._engine = None
        self._engine_full = None
        self._frozen_sig = None
        self._tower_keys = [k for k, _ in self._frozen_items()]

    # ---------------------------------------------------------------- engine plumbing
    def _blocks(self):
        return list(self.transformer.resblocks)

    def _frozen_items(self):
        return [(k, v) for k, v in self.state_dict(keep_vars=True).items() if not any(t in k for t in self._ADAPTER_KEYS)]

    def forward(self, x):
        blocks = self._blocks()
        extra = dict(rank=lowrank_rank(blocks[0].attn, self.adapter_kind),
                     out_dim=(self.output_dim if getattr(self, "proj", None) is not None else 0),
                     **bottleneck_cfg(blocks[0], self.adapter_kind))
        if getattr(self, "adapter_layers", None) is not None and extra.get("bottleneck"):
            extra["bottleneck_layers"] = tuple(sorted(int(i) for i in self.adapter_layers))      # AdapterDrop: see timm_vit.py
        if self.adapter_kind == "kron" and kron_is_dense(blocks[0].attn):
            extra["dense_delta"] = True          # large-n KAdaptation (the committed phm_dim = 768 through FS/model.py:170-193)
        return run_tower(self, x, "clip", self._geom, extra,
                         lambda: lowrank_operands([b.attn for b in blocks], self.adapter_kind),
                         lambda: bottleneck_operands(blocks, self.adapter_kind))


class CLIP(nn.Module):
    """FS/model.py:1222-1375.  The visual tower is the hot path (CUDA engine, forward + backward); the text tower
    (`encode_text`, :1311-1327) runs once per task, without gradients, to initialise the classifier head from prompt
    embeddings — it goes through the fp32 C-ABI kernels (LayerNorm, SGEMM with fused bias / QuickGELU / residual epilogues,
    attention with the causal mask of build_attention_mask :1303-1309)."""

    def __init__(self, embed_dim, image_resolution, vision_layers, vision_width, vision_patch_size, context_length=77,
                 vocab_size=49408, transformer_width=512, transformer_heads=8, transformer_layers=12, adapter="none",
                 **adapter_kw):
        super().__init__()
        if isinstance(vision_layers, (tuple, list)):
            raise NotImplementedError("ModifiedResNet visual towers are out of scope (SURVEY §2)")
        self.context_length = context_length
        self.visual = VisionTransformer(input_resolution=image_resolution, patch_size=vision_patch_size, width=vision_width,
                                        layers=vision_layers, heads=vision_width // 64, output_dim=embed_dim,
                                        adapter=adapter, **adapter_kw)
        # text tower parameters, reference names (FS/model.py:1262-1274); no adapter (the LoRA factors model.py also gives
        # the text blocks start at zero and are never trained: the tower is only used under no_grad)
        self.transformer = Transformer(transformer_width, transformer_layers, transformer_heads, adapter="none")
        self.vocab_size = vocab_size
        self.token_embedding = nn.Embedding(vocab_size, transformer_width)
        self.positional_embedding = nn.Parameter(torch.empty(context_length, transformer_width))
        self.ln_final = LayerNorm(transformer_width)
        self.text_projection = nn.Parameter(torch.empty(transformer_width, embed_dim))
        self.logit_scale = nn.Parameter(torch.ones([]) * 2.6592)
        self.initialize_parameters()

    def initialize_parameters(self):
        """FS/model.py:1278-1305."""
        nn.init.normal_(self.token_embedding.weight, std=0.02)
        nn.init.normal_(self.positional_embedding, std=0.01)
        for t in (self.visual.transformer, self.transformer):
            proj_std = (t.width ** -0.5) * ((2 * t.layers) ** -0.5)
            attn_std = t.width ** -0.5
            fc_std = (2 * t.width) ** -0.5
            for block in t.resblocks:
                nn.init.normal_(block.attn.in_proj_weight, std=attn_std)
                nn.init.normal_(block.attn.out_proj.weight, std=proj_std)
                nn.init.normal_(block.mlp.c_fc.weight, std=fc_std)
                nn.init.normal_(block.mlp.c_proj.weight, std=proj_std)
        nn.init.normal_(self.text_projection, std=self.transformer.width ** -0.5)

    @property
    def dtype(self):
        return self.visual.conv1.weight.dtype

    def encode_image(self, image):
        return self.visual(image.type(self.dtype))

    @torch.no_grad()
    def encode_text(self, text):
        """FS/model.py:1311-1327: token + positional embedding -> causal transformer -> ln_final -> the EOT row (argmax of the
        token ids) @ text_projection.  text: CUDA int64 [N, context_length].  Forward only (the reference freezes the text
        tower, FS/kadaptation_clip.py:98-104, and calls this once per task under no_grad)."""
        from . import _lib, ops
        if not text.is_cuda:
            raise RuntimeError("peft_b200 CLIP.encode_text runs on CUDA only (no CPU fallback)")
        N, L = text.shape
        if L != self.context_length:
            raise ValueError(f"text must be [N, {self.context_length}]")
        t = self.transformer
        W, H = t.width, t.heads
        if W // H != 64 or W % 128:
            raise NotImplementedError("text tower kernels need head_dim 64 and width % 128 == 0")
        dev, f32 = text.device, torch.float32
        e = lambda *sh: torch.empty(*sh, device=dev, dtype=f32)
        f = lambda p: p.detach().to(dev, f32).contiguous()
        x = (self.token_embedding.weight.detach().to(dev, f32)[text] + f(self.positional_embedding)).reshape(N * L, W).contiguous()
        xn, qkv, o, x1, lse = e(N * L, W), e(N * L, 3 * W), e(N * L, W), e(N * L, W), e(N * H * L)
        h, act = e(N * L, 4 * W), e(N * L, 4 * W)
        stream = torch.cuda.current_stream().cuda_stream
        for blk in t.resblocks:
            ops.layernorm_fwd(x, f(blk.ln_1.weight), f(blk.ln_1.bias), xn, None, None, 1e-5)
            ops.gemm(xn, f(blk.attn.in_proj_weight), qkv, bias=f(blk.attn.in_proj_bias))
            _lib.check(_lib.load().peft_attention_fwd_f32_ex(qkv.data_ptr(), qkv.stride(0), o.data_ptr(), o.stride(0), lse.data_ptr(),
                                                             N, H, L, 64 ** -0.5, 1, stream), "peft_attention_fwd_f32_ex")
            ops.gemm(o, f(blk.attn.out_proj.weight), x1, bias=f(blk.attn.out_proj.bias), epilogue=ops.EPI_RESID_F32, aux=x)
            ops.layernorm_fwd(x1, f(blk.ln_2.weight), f(blk.ln_2.bias), xn, None, None, 1e-5)
            ops.gemm(xn, f(blk.mlp.c_fc.weight), h, bias=f(blk.mlp.c_fc.bias), epilogue=ops.EPI_QUICKGELU, out2=act)
            ops.gemm(act, f(blk.mlp.c_proj.weight), x, bias=f(blk.mlp.c_proj.bias), epilogue=ops.EPI_RESID_F32, aux=x1)
        eot = x[torch.arange(N, device=dev) * L + text.argmax(dim=-1)].contiguous()     # LayerNorm is per row: gather first
        y = e(N, W)
        ops.layernorm_fwd(eot, f(self.ln_final.weight), f(self.ln_final.bias), y, None, None, 1e-5)
        out = e(N, self.text_projection.shape[1])
        ops.gemm(y, f(self.text_projection).t().contiguous(), out, epilogue=ops.EPI_STORE_F32)
        return out

    @torch.no_grad()
    def zeroshot_head_weights(self, class_token_batches):
        """The classifier initialisation of FS/feature.py:351-512 (extract_text_features) after tokenisation: per class, the
        prompt embeddings are L2-normalised, averaged and normalised again; returns [classes, embed_dim] — what
        FS/kadaptation_clip.py:136-138 stores (transposed there) into `layers[0].weight`.  class_token_batches: iterable of CUDA
        int64 [n_prompts, context_length] tensors, one per class (tokenising is CPU text processing outside this library)."""
        rows = []
        for toks in class_token_batches:
            emb = self.encode_text(toks)
            emb = emb / emb.norm(dim=-1, keepdim=True)
            m = emb.mean(dim=0)
            rows.append(m / m.norm())
        return torch.stack(rows, dim=0)

    def forward(self, image, text=None):
        if text is not None:
            return self.encode_text(text)
        return self.encode_image(image)


def _build(state_dict: dict, adapter: str, **adapter_kw):
    """Geometry inference of FS/model.py:1381-1401 (ViT branch), non-strict load (:1415-1418), returns .eval()."""
    if "visual.proj" not in state_dict:
        raise NotImplementedError("only ViT visual towers are supported (no `visual.proj` in the state dict)")
    vision_width = state_dict["visual.conv1.weight"].shape[0]
    vision_layers = len([k for k in state_dict if k.startswith("visual.") and k.endswith(".attn.in_proj_weight")])
    vision_patch_size = state_dict["visual.conv1.weight"].shape[-1]
    grid_size = round((state_dict["visual.positional_embedding"].shape[0] - 1) ** 0.5)
    image_resolution = vision_patch_size * grid_size
    embed_dim = state_dict["visual.proj"].shape[1]
    text_kw = {}
    if "text_projection" in state_dict:           # text geometry, FS/model.py:1395-1401
        tw = state_dict["ln_final.weight"].shape[0]
        text_kw = dict(context_length=state_dict["positional_embedding"].shape[0], vocab_size=state_dict["token_embedding.weight"].shape[0],
                       transformer_width=tw, transformer_heads=tw // 64,
                       transformer_layers=len({k.split(".")[2] for k in state_dict if k.startswith("transformer.resblocks.")}))
    model = CLIP(embed_dim, image_resolution, vision_layers, vision_width, vision_patch_size, adapter=adapter, **text_kw, **adapter_kw)
    sd = {k: v for k, v in state_dict.items() if k not in ("input_resolution", "context_length", "vocab_size")}
    model.load_state_dict(sd, strict=False)
    return model.eval()


def build_model(state_dict: dict, adapter="lora", **kw):
    """FS/model.py:1378 — the committed file runs the LoRA path; pass adapter="kron" for the C1 definition."""
    return _build(state_dict, adapter, **kw)


def build_lora_model(state_dict: dict, **kw):
    """FS/lora_model.py:1121."""
    return _build(state_dict, "lora", **kw)


def build_adapter_model(state_dict: dict, **kw):
    """FS/adapter_model.py:553."""
    return _build(state_dict, "houlsby", **kw)


def build_compacter_model(state_dict: dict, **kw):
    """FS/compacter_model.py:699.  For width 1024 pass bottleneck=64 (the reference's D//12 = 85 is not divisible by
    its own phm_dim 32)."""
    return _build(state_dict, "compacter", **kw)
