"""Deterministic synthetic weights / inputs (BASELINE.json: "synthetic data, random-init weights") shared by bench.py, the
tests, the golden-vector generator and the oracle.  Pure tensor factories from explicit CPU generators: no model
arithmetic lives here (the CPU restatement of the reference is oracle/peft_oracle.py, which only tests / smoke() /
bench.py's cpu_baseline leg may use).  `oracle/synth.py` re-exports this module for the test code.

Everything is produced from explicit torch CPU generators so that the same tensors can be re-created on any box from
a seed instead of shipping tens of MB of weights.  Standard deviations follow the reference initialisers:
  timm dialect  trunc_normal_(std=.02), zero bias, LayerNorm (1,0)    few_shot/vision_benchmark/evaluation/kadaptation.py:428-446
  CLIP dialect  attn N(0, D^-1), proj N(0, D^-1 (2 layers)^-1), fc N(0,(2D)^-1), class/pos/proj ~ D^-0.5 randn
                                                                       few_shot/vision_benchmark/evaluation/model.py:1192-1200,1278-1305
  adapters      LoRA down N(0,.02^2) (model.py:729); PHM factors N(0,.01^2) (kadaptation.py:179-183);
                Houlsby N(0,.02^2) (adapter_model.py:285-295)
with the SURVEY §8c overrides: uninitialised tensors (phm_rule, compacter rules) ~ N(0,.01^2); the zero-initialised
"up" factors are drawn N(0,.02^2) when `grad_test=True` so that every adapter gradient is exercised.
LayerNorm gains/biases and linear biases are perturbed (not the 1/0 defaults) so that parity tests see them.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import torch


@dataclass
class VitCfg:
    dialect: str = "timm"          # "timm" | "clip"
    img: int = 224
    patch: int = 16
    width: int = 768
    depth: int = 12
    heads: int = 12
    mlp_ratio: int = 4
    out_dim: int = 512             # CLIP visual projection width (clip dialect only)
    adapter: str = "kron"          # "none" | "lora" | "kron" | "houlsby" | "compacter"
    lora_r: int = 4
    lora_alpha: float = 128.0
    phm_dim: int = 4
    phm_rank: int = 1
    bottleneck: int = 64           # Houlsby / Compacter down-projection width
    compacter_n_down: int = 32
    compacter_n_up: int = 4
    num_classes: int = 10
    adapter_layers: tuple = ()     # AdapterDrop: blocks whose bottleneck adapter is applied (() = all)
    block_schedule: tuple = ()     # ((block, adapter applied), ...) per block application; () = every block once

    @property
    def seq(self) -> int:
        return (self.img // self.patch) ** 2 + 1

    @property
    def ln_eps(self) -> float:
        return 1e-6 if self.dialect == "timm" else 1e-5

    @property
    def feat_dim(self) -> int:
        return self.width if self.dialect == "timm" else self.out_dim


def _n(g, shape, std):
    return torch.randn(shape, generator=g, dtype=torch.float32) * std


def make_state_dict(cfg: VitCfg, seed: int = 0, grad_test: bool = True) -> dict:
    """Reference-named state dict for the visual tower + adapters (fp32, CPU)."""
    g = torch.Generator().manual_seed(seed)
    D, L, p = cfg.width, cfg.seq, cfg.patch
    sd = {}
    if cfg.dialect == "timm":
        sd["patch_embed.proj.weight"] = _n(g, (D, 3, p, p), 0.02)
        sd["patch_embed.proj.bias"] = _n(g, (D,), 0.02)
        sd["cls_token"] = _n(g, (1, 1, D), 0.02)
        sd["pos_embed"] = _n(g, (1, L, D), 0.02)
        for i in range(cfg.depth):
            b = f"blocks.{i}."
            sd[b + "norm1.weight"] = 1.0 + _n(g, (D,), 0.1)
            sd[b + "norm1.bias"] = _n(g, (D,), 0.05)
            sd[b + "attn.qkv.weight"] = _n(g, (3 * D, D), 0.02)
            sd[b + "attn.qkv.bias"] = _n(g, (3 * D,), 0.02)
            sd[b + "attn.proj.weight"] = _n(g, (D, D), 0.02)
            sd[b + "attn.proj.bias"] = _n(g, (D,), 0.02)
            sd[b + "norm2.weight"] = 1.0 + _n(g, (D,), 0.1)
            sd[b + "norm2.bias"] = _n(g, (D,), 0.05)
            sd[b + "mlp.fc1.weight"] = _n(g, (cfg.mlp_ratio * D, D), 0.02)
            sd[b + "mlp.fc1.bias"] = _n(g, (cfg.mlp_ratio * D,), 0.02)
            sd[b + "mlp.fc2.weight"] = _n(g, (D, cfg.mlp_ratio * D), 0.02)
            sd[b + "mlp.fc2.bias"] = _n(g, (D,), 0.02)
            _adapter_params(sd, b + "attn.", b, cfg, g, grad_test)
        sd["norm.weight"] = 1.0 + _n(g, (D,), 0.1)
        sd["norm.bias"] = _n(g, (D,), 0.05)
    else:
        scale = D ** -0.5
        proj_std = (D ** -0.5) * ((2 * cfg.depth) ** -0.5)
        attn_std = D ** -0.5
        fc_std = (2 * D) ** -0.5
        sd["conv1.weight"] = _n(g, (D, 3, p, p), 0.02)
        sd["class_embedding"] = _n(g, (D,), scale)
        sd["positional_embedding"] = _n(g, (L, D), scale)
        sd["ln_pre.weight"] = 1.0 + _n(g, (D,), 0.1)
        sd["ln_pre.bias"] = _n(g, (D,), 0.05)
        for i in range(cfg.depth):
            b = f"transformer.resblocks.{i}."
            sd[b + "ln_1.weight"] = 1.0 + _n(g, (D,), 0.1)
            sd[b + "ln_1.bias"] = _n(g, (D,), 0.05)
            sd[b + "attn.in_proj_weight"] = _n(g, (3 * D, D), attn_std)
            sd[b + "attn.in_proj_bias"] = _n(g, (3 * D,), 0.02)
            sd[b + "attn.out_proj.weight"] = _n(g, (D, D), proj_std)
            sd[b + "attn.out_proj.bias"] = _n(g, (D,), 0.02)
            sd[b + "ln_2.weight"] = 1.0 + _n(g, (D,), 0.1)
            sd[b + "ln_2.bias"] = _n(g, (D,), 0.05)
            sd[b + "mlp.c_fc.weight"] = _n(g, (cfg.mlp_ratio * D, D), fc_std)
            sd[b + "mlp.c_fc.bias"] = _n(g, (cfg.mlp_ratio * D,), 0.02)
            sd[b + "mlp.c_proj.weight"] = _n(g, (D, cfg.mlp_ratio * D), proj_std)
            sd[b + "mlp.c_proj.bias"] = _n(g, (D,), 0.02)
            _adapter_params(sd, b + "attn.", b, cfg, g, grad_test)
        sd["ln_post.weight"] = 1.0 + _n(g, (D,), 0.1)
        sd["ln_post.bias"] = _n(g, (D,), 0.05)
        sd["proj"] = _n(g, (D, cfg.out_dim), scale)
    return sd


def _adapter_params(sd, attn_prefix, block_prefix, cfg, g, grad_test):
    D = cfg.width
    up_std = 0.02 if grad_test else 0.0
    if cfg.adapter == "lora":
        r = cfg.lora_r
        for t in ("q", "v"):
            sd[attn_prefix + f"{t}_proj_adapter1.weight"] = _n(g, (r, D), 0.02)
            sd[attn_prefix + f"{t}_proj_adapter2.weight"] = _n(g, (D, r), up_std) if up_std else torch.zeros(D, r)
    elif cfg.adapter == "kron":
        n, rk = cfg.phm_dim, cfg.phm_rank
        assert D % n == 0
        sd[attn_prefix + "phm_rule"] = _n(g, (n, n, n), 0.01)
        for j in (1, 2):
            sd[attn_prefix + f"W_left{j}"] = _n(g, (n, D // n, rk), 0.01)
            sd[attn_prefix + f"W_right{j}"] = _n(g, (n, rk, D // n), 0.01)
        sd[attn_prefix + "phmb"] = torch.zeros(D)
    elif cfg.adapter == "houlsby":
        a = block_prefix + "adapter."
        m = cfg.bottleneck
        sd[a + "adapter_norm_before.weight"] = 1.0 + _n(g, (D,), 0.1)
        sd[a + "adapter_norm_before.bias"] = _n(g, (D,), 0.05)
        sd[a + "adapter_down.1.weight"] = _n(g, (m, D), 0.02)
        sd[a + "adapter_down.1.bias"] = _n(g, (m,), 0.02)
        sd[a + "adapter_up.weight"] = _n(g, (D, m), 0.02)
        sd[a + "adapter_up.bias"] = _n(g, (D,), 0.02)
    elif cfg.adapter == "compacter":
        a = block_prefix + "compacter."
        m, nd, nu = cfg.bottleneck, cfg.compacter_n_down, cfg.compacter_n_up
        sd[a + "adapter_norm_before.weight"] = 1.0 + _n(g, (D,), 0.1)
        sd[a + "adapter_norm_before.bias"] = _n(g, (D,), 0.05)
        # PHMLinear(in=D, out=m, phm_dim=nd): W [nd, D/nd, m/nd], phm_rule [nd,nd,nd], b [m]
        sd[a + "down_PHM.phm_rule"] = _n(g, (nd, nd, nd), 0.01)
        sd[a + "down_PHM.W"] = _n(g, (nd, D // nd, m // nd), math.sqrt(2.0 / (D // nd + m // nd)))
        sd[a + "down_PHM.b"] = _n(g, (m,), 0.02)
        sd[a + "up_sampler.phm_rule"] = _n(g, (nu, nu, nu), 0.01)
        sd[a + "up_sampler.W"] = _n(g, (nu, m // nu, D // nu), math.sqrt(2.0 / (m // nu + D // nu)))
        sd[a + "up_sampler.b"] = _n(g, (D,), 0.02)
    elif cfg.adapter == "none":
        pass
    else:
        raise ValueError(cfg.adapter)


def trainable_names(cfg: VitCfg, sd: dict, train_rule: bool = True) -> list:
    """Names a reference trainer would leave requires_grad=True."""
    out = []
    for k in sd:
        if cfg.adapter == "lora" and "_proj_adapter" in k:
            out.append(k)
        elif cfg.adapter == "kron" and (".W_left" in k or ".W_right" in k or (train_rule and k.endswith("phm_rule"))):
            out.append(k)
        elif cfg.adapter in ("houlsby", "compacter") and (".adapter." in k or ".compacter." in k):
            # AdapterDrop trains only the adapters that are applied (FS/adapterdrop.py:689: 'blocks.11.adapter' in n)
            layer = int(k.split(".adapter.")[0].split(".compacter.")[0].split(".")[-1])
            if cfg.block_schedule:
                if layer == cfg.depth - 1:          # FS/adapterdrop.py:689: only 'blocks.11.adapter' is trainable
                    out.append(k)
            elif not cfg.adapter_layers or layer in cfg.adapter_layers:
                out.append(k)
    return out


def make_head(cfg: VitCfg, seed: int = 7) -> dict:
    """Linear(E, classes) of Classifier.layers (kadaptation_clip.py:126); the BatchNorm1d in front has no parameters."""
    g = torch.Generator().manual_seed(seed)
    E = cfg.feat_dim
    bound = 1.0 / math.sqrt(E)
    w = (torch.rand((cfg.num_classes, E), generator=g) * 2 - 1) * bound
    b = (torch.rand((cfg.num_classes,), generator=g) * 2 - 1) * bound
    return {"layers.0.weight": w, "layers.0.bias": b}


def make_batch(cfg: VitCfg, batch: int, seed_img: int = 1234, seed_lbl: int = 1235):
    """SURVEY §8d synthetic inputs: randn images (seed 1234), randint labels (seed 1235)."""
    gi = torch.Generator().manual_seed(seed_img)
    gl = torch.Generator().manual_seed(seed_lbl)
    images = torch.randn((batch, 3, cfg.img, cfg.img), generator=gi, dtype=torch.float32)
    labels = torch.randint(0, cfg.num_classes, (batch,), generator=gl)
    return images, labels


def make_text_state_dict(width=512, layers=2, vocab=1000, context=77, embed_dim=512, seed=21) -> dict:
    """Reference-named text-tower tensors of CLIP (FS/model.py:1262-1274) with the standard deviations of its
    initialize_parameters (:1278-1305); LayerNorm gains / biases and linear biases are perturbed so that parity tests see them."""
    g = torch.Generator().manual_seed(seed)
    sd = {"token_embedding.weight": _n(g, (vocab, width), 0.02), "positional_embedding": _n(g, (context, width), 0.01)}
    proj_std = (width ** -0.5) * ((2 * layers) ** -0.5)
    for i in range(layers):
        b = f"transformer.resblocks.{i}."
        sd[b + "ln_1.weight"] = 1.0 + _n(g, (width,), 0.1)
        sd[b + "ln_1.bias"] = _n(g, (width,), 0.05)
        sd[b + "attn.in_proj_weight"] = _n(g, (3 * width, width), width ** -0.5)
        sd[b + "attn.in_proj_bias"] = _n(g, (3 * width,), 0.02)
        sd[b + "attn.out_proj.weight"] = _n(g, (width, width), proj_std)
        sd[b + "attn.out_proj.bias"] = _n(g, (width,), 0.02)
        sd[b + "ln_2.weight"] = 1.0 + _n(g, (width,), 0.1)
        sd[b + "ln_2.bias"] = _n(g, (width,), 0.05)
        sd[b + "mlp.c_fc.weight"] = _n(g, (4 * width, width), (2 * width) ** -0.5)
        sd[b + "mlp.c_fc.bias"] = _n(g, (4 * width,), 0.02)
        sd[b + "mlp.c_proj.weight"] = _n(g, (width, 4 * width), proj_std)
        sd[b + "mlp.c_proj.bias"] = _n(g, (width,), 0.02)
    sd["ln_final.weight"] = 1.0 + _n(g, (width,), 0.1)
    sd["ln_final.bias"] = _n(g, (width,), 0.05)
    sd["text_projection"] = _n(g, (width, embed_dim), width ** -0.5)
    return sd


def make_tokens(n, vocab=1000, context=77, seed=31):
    """[n, context] int64 prompts: random ids, one EOT (= vocab - 1, the largest id, which encode_text locates by argmax) at a
    random position >= 3, zero padding after it — the shape of SimpleTokenizer output."""
    g = torch.Generator().manual_seed(seed)
    toks = torch.randint(1, vocab - 1, (n, context), generator=g)
    ends = torch.randint(3, context, (n,), generator=g)
    for i in range(n):
        toks[i, ends[i]] = vocab - 1
        toks[i, ends[i] + 1:] = 0
    return toks


def make_image_u8(h, w, seed=41):
    """Synthetic decoded RGB image uint8 [h, w, 3]: smooth random field + noise (so that resampling sees real gradients)."""
    g = torch.Generator().manual_seed(seed + 1000 * h + w)
    low = torch.rand((1, 3, max(2, h // 16), max(2, w // 16)), generator=g)
    img = torch.nn.functional.interpolate(low, size=(h, w), mode="bilinear", align_corners=False)[0]
    img = (img + 0.25 * torch.rand((3, h, w), generator=g)).clamp(0, 1.2) / 1.2
    return (img * 255).round().to(torch.uint8).permute(1, 2, 0).contiguous()
