"""CPU oracle: a plain-PyTorch fp32 restatement of the reference's PEFT-ViT hot path.

TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product path (peft_b200) never does and fails loudly without its CUDA library.

Every function cites the reference lines it restates (paths relative to the reference root;
FS = few_shot/vision_benchmark/evaluation, FL = full_shot/main).  The reference has no tests or golden vectors for
this path, so parity is pinned by tests/golden/*.pt: outputs of the UNMODIFIED reference modules,
imported in-process by tests/golden/make_golden.py, on the deterministic synthetic weights of oracle/synth.py.
tests/test_oracle_golden.py checks this file against those vectors.

Functions are functional (state-dict in, tensors out) and differentiable through torch autograd, which is the
reference's own backward.
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F

from .synth import VitCfg


# --------------------------------------------------------------------------------------------------------------------
# KAdaptation / PHM algebra
# --------------------------------------------------------------------------------------------------------------------
def kronecker_product_einsum_batched(A: torch.Tensor, B: torch.Tensor) -> torch.Tensor:
    """FS/kadaptation.py:104-115 — (b,a,c) x (b,k,p) -> (b, a*k, c*p)."""
    assert A.dim() == 3 and B.dim() == 3
    return torch.einsum("bac,bkp->bakcp", A, B).reshape(A.size(0), A.size(1) * B.size(1), A.size(2) * B.size(2))


def matvec_product(W: torch.Tensor, x: torch.Tensor, phm_rule: torch.Tensor) -> torch.Tensor:
    """FS/kadaptation.py:81-102 — y = x @ sum_i phm_rule[i] (x) W[i]; the bias add is commented out in the reference."""
    H = kronecker_product_einsum_batched(phm_rule, W).sum(0)
    return torch.matmul(x, H)


def kron_delta(x, phm_rule, W_left, W_right):
    """FS/kadaptation.py:211-215 — W = bmm(W_left, W_right); delta = matvec_product(W, x, phm_rule)."""
    return matvec_product(torch.bmm(W_left, W_right), x, phm_rule)


def kron_lowrank_factors(phm_rule, W_left, W_right):
    """Reshape-trick factorisation used by the CUDA path (not in the reference; checked against kron_delta in tests).

    H[(a,k),(c,p)] = sum_{i,r} A_i[a,c] u_i[k,r] v_i[r,p]  ==  Dn^T @ Up^T with
    Dn[(i,r,c),(a,k)] = A_i[a,c] u_i[k,r]   and   Up[(c,p),(i,r,c')] = delta(c,c') v_i[r,p].
    """
    n, K, r = W_left.shape
    P = W_right.shape[2]
    Dn = torch.einsum("iac,ikr->ircak", phm_rule, W_left).reshape(n * r * n, n * K)
    eye = torch.eye(n, dtype=phm_rule.dtype)
    Up = torch.einsum("cd,irp->cpird", eye, W_right).reshape(n * P, n * r * n)
    return Dn, Up


def lora_delta(x, w1, w2, alpha=128.0, r=4):
    """FS/model.py:754-762 / FS/lora.py adapter_forward — ((x W1^T) W2^T) * alpha / r."""
    return torch.matmul(torch.matmul(x, w1.T), w2.T) * (alpha / r)


# --------------------------------------------------------------------------------------------------------------------
# activations
# --------------------------------------------------------------------------------------------------------------------
def quick_gelu(x):
    """FS/model.py:165-167."""
    return x * torch.sigmoid(1.702 * x)


def gelu_new(x):
    """transformers.activations 'gelu_new' used by FS/compacter_model.py:168-176,376."""
    return 0.5 * x * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (x + 0.044715 * torch.pow(x, 3.0))))


# --------------------------------------------------------------------------------------------------------------------
# bottleneck adapters
# --------------------------------------------------------------------------------------------------------------------
# Test hooks for the discontinuous ReLU gate of the Houlsby bottleneck (tests/houlsby_gate.py).  GATE_RECORD, when a dict,
# receives {adapter prefix: pre-activation}; GATE_OVERRIDE maps an adapter prefix to a 0/1 mask that REPLACES the sign
# test of the ReLU (z = pre * mask): the oracle then differentiates the linear region the CUDA engine selected, which
# separates "the kernels compute the wrong gradient" from "a pre-activation within bf16 noise of 0 flipped its gate".
GATE_RECORD = None
GATE_OVERRIDE = {}
GATE_CALLS = {}


def houlsby_adapter(h, sd, pfx, eps=1e-5):
    """FS/adapter_model.py:204-295 — up(relu(down(LN(h)))) + h  (residual_input = h at the call site :330-334)."""
    D = h.shape[-1]
    z = F.layer_norm(h, (D,), sd[pfx + "adapter_norm_before.weight"], sd[pfx + "adapter_norm_before.bias"], eps)
    pre = F.linear(z, sd[pfx + "adapter_down.1.weight"], sd[pfx + "adapter_down.1.bias"])
    if GATE_RECORD is not None:
        GATE_RECORD.setdefault(pfx, []).append(pre.detach())      # one entry per application (AdapterDrop schedules apply twice)
    gate = GATE_OVERRIDE.get(pfx)
    if isinstance(gate, list):                                    # masks per application, consumed in call order
        n = GATE_CALLS.get(pfx, 0)
        GATE_CALLS[pfx] = n + 1
        gate = gate[n]
    z = F.relu(pre) if gate is None else pre * gate.to(pre.dtype)
    z = F.linear(z, sd[pfx + "adapter_up.weight"], sd[pfx + "adapter_up.bias"])
    return z + h


def phm_linear(x, rule, W, b):
    """FS/compacter_model.py:190-208,213-355 — PHMLinear forward: x @ sum_i rule[i] (x) W[i] + b."""
    H = kronecker_product_einsum_batched(rule, W).sum(0)
    return torch.matmul(x, H) + b


def compacter_adapter(h, sd, pfx, eps=1e-5):
    """FS/compacter_model.py:358-425 — up_PHM(gelu_new(down_PHM(LN(h)))) + h."""
    D = h.shape[-1]
    z = F.layer_norm(h, (D,), sd[pfx + "adapter_norm_before.weight"], sd[pfx + "adapter_norm_before.bias"], eps)
    z = gelu_new(phm_linear(z, sd[pfx + "down_PHM.phm_rule"], sd[pfx + "down_PHM.W"], sd[pfx + "down_PHM.b"]))
    z = phm_linear(z, sd[pfx + "up_sampler.phm_rule"], sd[pfx + "up_sampler.W"], sd[pfx + "up_sampler.b"])
    return z + h


# --------------------------------------------------------------------------------------------------------------------
# timm dialect  (FS/kadaptation.py, FS/lora.py, FS/adapter.py; FL/lib/models/cls_vit_*.py)
# --------------------------------------------------------------------------------------------------------------------
def timm_attention(x, sd, pfx, cfg: VitCfg):
    """FS/kadaptation.py:198-238 (KAdaptation) and the LoRA twin in FS/lora.py: un-scrambled delta, scale after add."""
    B, N, C = x.shape
    H = cfg.heads
    hd = C // H
    qkv = F.linear(x, sd[pfx + "qkv.weight"], sd.get(pfx + "qkv.bias"))
    qkv = qkv.reshape(B, N, 3, H, hd).permute(2, 0, 3, 1, 4)
    q, k, v = qkv[0], qkv[1], qkv[2]
    if cfg.adapter == "kron":
        qd = kron_delta(x, sd[pfx + "phm_rule"], sd[pfx + "W_left1"], sd[pfx + "W_right1"])
        vd = kron_delta(x, sd[pfx + "phm_rule"], sd[pfx + "W_left2"], sd[pfx + "W_right2"])
    elif cfg.adapter == "lora":
        qd = lora_delta(x, sd[pfx + "q_proj_adapter1.weight"], sd[pfx + "q_proj_adapter2.weight"], cfg.lora_alpha, cfg.lora_r)
        vd = lora_delta(x, sd[pfx + "v_proj_adapter1.weight"], sd[pfx + "v_proj_adapter2.weight"], cfg.lora_alpha, cfg.lora_r)
    else:
        qd = vd = None
    if qd is not None:
        q = q + qd.reshape(B, N, H, hd).permute(0, 2, 1, 3)
        v = v + vd.reshape(B, N, H, hd).permute(0, 2, 1, 3)
    attn = (q @ k.transpose(-2, -1)) * (hd ** -0.5)
    attn = F.softmax(attn, dim=-1)
    out = (attn @ v).transpose(1, 2).reshape(B, N, C)
    return F.linear(out, sd[pfx + "proj.weight"], sd[pfx + "proj.bias"])


def timm_mlp(x, sd, pfx):
    """FS/kadaptation.py:57-78 — fc1 -> GELU(erf) -> fc2 (dropout p=0)."""
    return F.linear(F.gelu(F.linear(x, sd[pfx + "fc1.weight"], sd[pfx + "fc1.bias"])), sd[pfx + "fc2.weight"], sd[pfx + "fc2.bias"])


def drop_path(x, keep, keep_prob):
    """timm.models.layers.drop_path (timm 0.4.12, imported at FS/kadaptation.py:30, used :263,276-277) with the Bernoulli
    draw made explicit: `keep` [B] of 0/1 replaces floor(keep_prob + rand(B)); output = x / keep_prob * keep."""
    return x.div(keep_prob) * keep.view(-1, 1, 1).to(x.dtype)


def timm_block(x, sd, pfx, cfg: VitCfg, drop=None, use_adapter=None):
    """FS/kadaptation.py:273-277 (pre-norm branch); Houlsby variant FS/adapter.py (adapter after the MLP).
    drop = (keep_attn [B], keep_mlp [B], keep_prob) applies stochastic depth to the two branches (training mode)."""
    D = x.shape[-1]
    a = timm_attention(F.layer_norm(x, (D,), sd[pfx + "norm1.weight"], sd[pfx + "norm1.bias"], cfg.ln_eps), sd, pfx + "attn.", cfg)
    x = x + (a if drop is None else drop_path(a, drop[0], drop[2]))
    h = timm_mlp(F.layer_norm(x, (D,), sd[pfx + "norm2.weight"], sd[pfx + "norm2.bias"], cfg.ln_eps), sd, pfx + "mlp.")
    layer = int(pfx.split(".")[-2])
    if use_adapter is None:
        use_adapter = not cfg.adapter_layers or layer in cfg.adapter_layers
    if not use_adapter:
        pass        # AdapterDrop (FS/adapterdrop.py:585-588): `blk(x, False, prev)` — the block's adapter is not applied
    elif cfg.adapter == "houlsby":
        h = houlsby_adapter(h, sd, pfx + "adapter.", 1e-5)
    elif cfg.adapter == "compacter":
        h = compacter_adapter(h, sd, pfx + "compacter.", 1e-5)
    return x + (h if drop is None else drop_path(h, drop[1], drop[2]))


def timm_forward_features(images, sd, cfg: VitCfg, drop_keep=None, drop_path_rate=0.0):
    """FS/kadaptation.py:516-544 + PatchEmbed :286-308 — returns the final-LayerNorm cls feature [B, D].
    drop_keep [depth, 2, B] (0/1) + drop_path_rate: stochastic depth with fixed masks, per-block rate
    linspace(0, rate, depth)[i] (:408)."""
    B = images.shape[0]
    x = F.conv2d(images, sd["patch_embed.proj.weight"], sd["patch_embed.proj.bias"], stride=cfg.patch)
    x = x.flatten(2).transpose(1, 2)
    x = torch.cat((sd["cls_token"].expand(B, -1, -1), x), dim=1)
    x = x + sd["pos_embed"]
    dpr = [v.item() for v in torch.linspace(0, drop_path_rate, cfg.depth)]
    if cfg.block_schedule:
        # FS/adapterdrop.py:582-588 as committed: every block is applied twice (see adapterdrop_reference_schedule)
        for b, use in cfg.block_schedule:
            x = timm_block(x, sd, f"blocks.{b}.", cfg, None, use_adapter=bool(use))
    for i in range(0 if cfg.block_schedule else cfg.depth):
        drop = None if drop_keep is None else (drop_keep[i, 0], drop_keep[i, 1], 1.0 - dpr[i])
        x = timm_block(x, sd, f"blocks.{i}.", cfg, drop)
    x = F.layer_norm(x, (cfg.width,), sd["norm.weight"], sd["norm.bias"], cfg.ln_eps)
    return x[:, 0]


# --------------------------------------------------------------------------------------------------------------------
# CLIP dialect  (FS/model.py, FS/lora_model.py, FS/adapter_model.py, FS/compacter_model.py)
# --------------------------------------------------------------------------------------------------------------------
def clip_attention(x, sd, pfx, cfg: VitCfg):
    """FS/model.py:790-1010 multi_head_attention_forward, self-attention, no mask.

    x is LND [L, B, C].  Quirks reproduced: q is scaled by 1/sqrt(hd) BEFORE the delta is added (Q2,
    :967 then :983) and the delta is the RAW reshape of an [L,B,C] tensor to [B*H, L, hd] (Q1, :981-982).
    With adapter == "kron" the Kronecker delta replaces the LoRA call at :981-982 (Q3, SURVEY C1 definition).
    Houlsby / Compacter files use stock nn.MultiheadAttention, i.e. the same math without a delta.
    """
    L, B, C = x.shape
    H = cfg.heads
    hd = C // H
    qkv = F.linear(x, sd[pfx + "in_proj_weight"], sd[pfx + "in_proj_bias"])
    q, k, v = qkv.chunk(3, dim=-1)
    q = q.contiguous().view(L, B * H, hd).transpose(0, 1)
    k = k.contiguous().view(L, B * H, hd).transpose(0, 1)
    v = v.contiguous().view(L, B * H, hd).transpose(0, 1)
    q = q / math.sqrt(hd)
    if cfg.adapter == "lora":
        qd = lora_delta(x, sd[pfx + "q_proj_adapter1.weight"], sd[pfx + "q_proj_adapter2.weight"], cfg.lora_alpha, cfg.lora_r)
        vd = lora_delta(x, sd[pfx + "v_proj_adapter1.weight"], sd[pfx + "v_proj_adapter2.weight"], cfg.lora_alpha, cfg.lora_r)
    elif cfg.adapter == "kron":
        qd = kron_delta(x, sd[pfx + "phm_rule"], sd[pfx + "W_left1"], sd[pfx + "W_right1"])
        vd = kron_delta(x, sd[pfx + "phm_rule"], sd[pfx + "W_left2"], sd[pfx + "W_right2"])
    else:
        qd = vd = None
    if qd is not None:
        q = q + qd.reshape(B * H, L, hd)
        v = v + vd.reshape(B * H, L, hd)
    attn = F.softmax(torch.bmm(q, k.transpose(-2, -1)), dim=-1)
    out = torch.bmm(attn, v)
    out = out.transpose(0, 1).contiguous().view(L * B, C)
    out = F.linear(out, sd[pfx + "out_proj.weight"], sd[pfx + "out_proj.bias"])
    return out.view(L, B, C)


def clip_block(x, sd, pfx, cfg: VitCfg):
    """FS/model.py:1143-1158 ResidualAttentionBlock; Houlsby FS/adapter_model.py:330-334; Compacter :477-479."""
    D = x.shape[-1]
    x = x + clip_attention(F.layer_norm(x, (D,), sd[pfx + "ln_1.weight"], sd[pfx + "ln_1.bias"], 1e-5), sd, pfx + "attn.", cfg)
    h = F.layer_norm(x, (D,), sd[pfx + "ln_2.weight"], sd[pfx + "ln_2.bias"], 1e-5)
    h = F.linear(quick_gelu(F.linear(h, sd[pfx + "mlp.c_fc.weight"], sd[pfx + "mlp.c_fc.bias"])), sd[pfx + "mlp.c_proj.weight"], sd[pfx + "mlp.c_proj.bias"])
    layer = int(pfx.split(".")[-2])
    if cfg.adapter_layers and layer not in cfg.adapter_layers:
        pass
    elif cfg.adapter == "houlsby":
        h = houlsby_adapter(h, sd, pfx + "adapter.", 1e-5)
    elif cfg.adapter == "compacter":
        h = compacter_adapter(h, sd, pfx + "compacter.", 1e-5)
    return x + h


def clip_visual_forward(images, sd, cfg: VitCfg):
    """FS/model.py:1202-1219 VisionTransformer.forward -> [B, out_dim]."""
    B = images.shape[0]
    D = cfg.width
    x = F.conv2d(images, sd["conv1.weight"], None, stride=cfg.patch)
    x = x.reshape(B, D, -1).permute(0, 2, 1)
    x = torch.cat([sd["class_embedding"].to(x.dtype) + torch.zeros(B, 1, D, dtype=x.dtype), x], dim=1)
    x = x + sd["positional_embedding"]
    x = F.layer_norm(x, (D,), sd["ln_pre.weight"], sd["ln_pre.bias"], 1e-5)
    x = x.permute(1, 0, 2)
    for i in range(cfg.depth):
        x = clip_block(x, sd, f"transformer.resblocks.{i}.", cfg)
    x = x.permute(1, 0, 2)
    x = F.layer_norm(x[:, 0, :], (D,), sd["ln_post.weight"], sd["ln_post.bias"], 1e-5)
    return x @ sd["proj"]


def visual_forward(images, sd, cfg: VitCfg):
    return timm_forward_features(images, sd, cfg) if cfg.dialect == "timm" else clip_visual_forward(images, sd, cfg)


# --------------------------------------------------------------------------------------------------------------------
# CLIP text tower  (FS/model.py:1303-1327) and the zero-shot head initialisation (FS/feature.py:351-512)
# --------------------------------------------------------------------------------------------------------------------
def clip_encode_text(text, sd, heads=8):
    """FS/model.py:1311-1327 — token + positional embedding, causal transformer (additive -inf upper-triangular mask,
    build_attention_mask :1303-1309; blocks as in clip_block without adapter), ln_final, EOT row @ text_projection."""
    N, L = text.shape
    W = sd["ln_final.weight"].shape[0]
    hd = W // heads
    x = sd["token_embedding.weight"][text] + sd["positional_embedding"]
    mask = torch.full((L, L), float("-inf")).triu_(1)
    layers = len({k.split(".")[2] for k in sd if k.startswith("transformer.resblocks.")})
    for i in range(layers):
        p = f"transformer.resblocks.{i}."
        y = F.layer_norm(x, (W,), sd[p + "ln_1.weight"], sd[p + "ln_1.bias"], 1e-5)
        q, k, v = F.linear(y, sd[p + "attn.in_proj_weight"], sd[p + "attn.in_proj_bias"]).chunk(3, dim=-1)
        q, k, v = [t.reshape(N, L, heads, hd).transpose(1, 2) for t in (q, k, v)]
        a = torch.softmax(q @ k.transpose(-1, -2) / math.sqrt(hd) + mask, dim=-1) @ v
        x = x + F.linear(a.transpose(1, 2).reshape(N, L, W), sd[p + "attn.out_proj.weight"], sd[p + "attn.out_proj.bias"])
        y = F.layer_norm(x, (W,), sd[p + "ln_2.weight"], sd[p + "ln_2.bias"], 1e-5)
        x = x + F.linear(quick_gelu(F.linear(y, sd[p + "mlp.c_fc.weight"], sd[p + "mlp.c_fc.bias"])), sd[p + "mlp.c_proj.weight"], sd[p + "mlp.c_proj.bias"])
    x = F.layer_norm(x, (W,), sd["ln_final.weight"], sd["ln_final.bias"], 1e-5)
    return x[torch.arange(N), text.argmax(dim=-1)] @ sd["text_projection"]


def zeroshot_weights(class_token_batches, sd, heads=8):
    """FS/feature.py:497-503 — per class: normalise the prompt embeddings, mean, normalise; stacked [classes, E] (the
    reference stacks along dim 1 and transposes at FS/kadaptation_clip.py:137)."""
    rows = []
    for toks in class_token_batches:
        e = clip_encode_text(toks, sd, heads)
        e = e / e.norm(dim=-1, keepdim=True)
        m = e.mean(dim=0)
        rows.append(m / m.norm())
    return torch.stack(rows, dim=0)


# --------------------------------------------------------------------------------------------------------------------
# classifier head, loss, one training step
# --------------------------------------------------------------------------------------------------------------------
def classifier_logits(feature, head, training=True, bn_eps=1e-5):
    """FS/kadaptation_clip.py:159-168 — BatchNorm1d(affine=False) (batch statistics in train mode) -> Linear."""
    if training:
        mean = feature.mean(0, keepdim=True)
        var = feature.var(0, unbiased=False, keepdim=True)
        feature = (feature - mean) / torch.sqrt(var + bn_eps)
    return F.linear(feature, head["layers.0.weight"], head["layers.0.bias"])


def train_step(images, labels, sd, head, cfg: VitCfg, trainable):
    """FS/kadaptation_clip.py:308-314 — forward, CrossEntropy, backward.  Returns (logits, loss, {name: grad})."""
    for k in trainable:
        sd[k].requires_grad_(True)
        sd[k].grad = None
    for k in head:
        head[k].requires_grad_(True)
        head[k].grad = None
    feat = visual_forward(images, sd, cfg)
    logits = classifier_logits(feat, head, True)
    loss = F.cross_entropy(logits, labels)
    loss.backward()
    grads = {k: sd[k].grad.detach().clone() for k in trainable}
    grads.update({k: head[k].grad.detach().clone() for k in head})
    return logits.detach(), loss.detach(), grads, feat.detach()


# --------------------------------------------------------------------------------------------------------------------
# Fastfood  (FL/intrinsic/fastfood.py)
# --------------------------------------------------------------------------------------------------------------------
def fwht(x: torch.Tensor) -> torch.Tensor:
    """FL/intrinsic/fastfood.py:81-121 — unnormalised Walsh-Hadamard transform, natural (Sylvester) order, 1-D."""
    n = x.shape[0]
    ll = int(round(np.log(n) / np.log(2)))
    assert n == 2 ** ll
    ret = x.reshape([2] * ll) if ll > 0 else x
    for dim in range(ll):
        a, b = torch.chunk(ret, 2, dim=dim)
        ret = torch.cat((a + b, a - b), dim=dim)
    return ret.reshape(n)


def fastfood_vars(DD: int, gen_np: np.random.RandomState, gen_t: torch.Generator):
    """FL/intrinsic/fastfood.py:124-148 with explicit generators (the reference uses the global RNGs)."""
    ll = int(np.ceil(np.log(DD) / np.log(2)))
    LL = 2 ** ll
    BB = (torch.rand(LL, generator=gen_t) * 2).long().float() * 2 - 1
    Pi = torch.from_numpy(gen_np.permutation(LL)).long()
    GG = torch.randn(LL, generator=gen_t)
    divisor = torch.sqrt(LL * torch.sum(GG ** 2))
    return [BB, Pi, GG, divisor, LL]


def fastfood_torched(x, DD, params):
    """FL/intrinsic/fastfood.py:151-187 — H G Pi H B pad(x), first DD entries / (divisor * sqrt(DD/LL))."""
    BB, Pi, GG, divisor, LL = params
    dd = x.shape[0]
    pad = F.pad(x, (0, LL - dd))
    m = fwht(BB * pad)
    m = fwht(m[Pi] * GG)
    return m[:DD] / (divisor * np.sqrt(float(DD) / LL))
