"""Frozen-backbone ViT step engine: runs the whole visual tower forward and backward through the C-ABI kernels.

One `VitEngine` owns
  * the prepared frozen weights (bf16 W [N,K] for the forward products, bf16 W^T [K,N] for the dX products — the
    backbone is frozen, so there is never a dW product; fp32 biases and LayerNorm parameters),
  * a per-batch-size workspace holding the activations saved for backward and the scratch buffers,
and exposes the tower as ONE autograd node (`backbone(images, Dn, UpT)`), so Python overhead per step is a few dozen
ctypes calls and torch.autograd only sees the tiny adapter tensors.

Adapter algebra (timm dialect): the q/v update is applied in low-rank form
      qkv = [LN1(x) | p] @ [Wqkv | Up]^T + b,      p = LN1(x) @ Dn^T            (K-extension of the frozen GEMM)
with   LoRA:        Dn = [W1_q ; W1_v],               Up = alpha/r * blockdiag(W2_q, 0, W2_v)
       KAdaptation: Dn[(i,r,c),(a,k)] = A_i[a,c] u_i[k,r],   Up[(c,p),(i,r,c')] = delta(c,c') v_i[r,p]   (reshape trick)
so the [D,D] update matrix is never formed.  The engine consumes Dn [depth, r, D] and UpT [depth, r, 3D] (fp32, built
by tiny differentiable torch ops from the reference-named parameters, see modules.py) and returns their gradients,
which are token reductions p^T dqkv and dp^T LN1(x) done by peft_tn_reduce.

Reference lines replaced: FS/kadaptation.py:198-238,273-277,516-544; FS/lora.py (same lines); FS/model.py:1155-1158,
1202-1219 (CLIP dialect: QuickGELU, ln_pre, ln_post, proj; eps 1e-5).
"""
from __future__ import annotations

from dataclasses import dataclass

import torch

from . import ops


@dataclass
class EngineCfg:
    dialect: str          # "timm" | "clip"
    img: int
    patch: int
    width: int
    depth: int
    heads: int
    mlp: int
    out_dim: int = 0      # clip: visual projection width (0 = no projection)
    rank: int = 0         # low-rank q/v adapter columns r' (0 = no q/v adapter)
    bottleneck: int = 0   # Houlsby / Compacter bottleneck width after the MLP (0 = none; 64 in the reference)
    bottleneck_layers: tuple = ()   # AdapterDrop (FS/adapterdrop.py:585-588): blocks that USE the bottleneck adapter; () = every block
    bn_act: str = "relu"  # "relu" (Houlsby, FS/adapter_model.py:304-311) | "gelu_tanh" (Compacter gelu_new, FS/compacter_model.py:376)
    dense_delta: bool = False  # q/v update given as dense H_q, H_v [D, D] (large-n KAdaptation "fold" variant):
                               # rank == width, Dn is the identity (not trainable) and UpT = [H_q | 0 | H_v]
    train_all: bool = False  # full fine-tuning / intrinsic-dimension mode: backward also emits dL/dW for EVERY tower tensor
    precision: str = "bf16"  # "bf16": tensor-core path | "fp32": verification mode — every buffer fp32, plain CUDA-core kernels
                             # (csrc/f32_sm100.cuh), same engine code; gate 1e-5 relative vs the fp32 oracle

    @property
    def cdt(self):
        return torch.float32 if self.precision == "fp32" else torch.bfloat16

    @property
    def seq(self):
        return (self.img // self.patch) ** 2 + 1

    @property
    def eps(self):
        return 1e-6 if self.dialect == "timm" else 1e-5

    @property
    def patch_k(self):
        return (3 * self.patch * self.patch + 7) // 8 * 8

    def has_bn(self, i):
        return bool(self.bottleneck) and (not self.bottleneck_layers or i in self.bottleneck_layers)

    @property
    def rank_pad(self):
        return ((self.rank + 15) // 16) * 16


_TIMM_KEYS = dict(ln1_w="blocks.{i}.norm1.weight", ln1_b="blocks.{i}.norm1.bias", qkv_w="blocks.{i}.attn.qkv.weight",
                  qkv_b="blocks.{i}.attn.qkv.bias", proj_w="blocks.{i}.attn.proj.weight", proj_b="blocks.{i}.attn.proj.bias",
                  ln2_w="blocks.{i}.norm2.weight", ln2_b="blocks.{i}.norm2.bias", fc1_w="blocks.{i}.mlp.fc1.weight",
                  fc1_b="blocks.{i}.mlp.fc1.bias", fc2_w="blocks.{i}.mlp.fc2.weight", fc2_b="blocks.{i}.mlp.fc2.bias")
_CLIP_KEYS = dict(ln1_w="transformer.resblocks.{i}.ln_1.weight", ln1_b="transformer.resblocks.{i}.ln_1.bias",
                  qkv_w="transformer.resblocks.{i}.attn.in_proj_weight", qkv_b="transformer.resblocks.{i}.attn.in_proj_bias",
                  proj_w="transformer.resblocks.{i}.attn.out_proj.weight", proj_b="transformer.resblocks.{i}.attn.out_proj.bias",
                  ln2_w="transformer.resblocks.{i}.ln_2.weight", ln2_b="transformer.resblocks.{i}.ln_2.bias",
                  fc1_w="transformer.resblocks.{i}.mlp.c_fc.weight", fc1_b="transformer.resblocks.{i}.mlp.c_fc.bias",
                  fc2_w="transformer.resblocks.{i}.mlp.c_proj.weight", fc2_b="transformer.resblocks.{i}.mlp.c_proj.bias")


def _tn_chunks(rp):
    """(offset, width) pieces of the rank_pad adapter columns that peft_tn_reduce accepts (I in {64, 32, 16}); rank_pad is a
    multiple of 16, so a 48-wide tail becomes 32 + 16."""
    out, c = [], 0
    while c < rp:
        cw = 64 if rp - c >= 64 else (32 if rp - c >= 32 else 16)
        out.append((c, cw))
        c += cw
    return out


class _Workspace:
    """Activations saved for backward + scratch, for one batch size."""

    def __init__(self, cfg: EngineCfg, B: int, dev):
        D, L, H, F, dep, rp = cfg.width, cfg.seq, cfg.heads, cfg.mlp, cfg.depth, cfg.rank_pad
        M = B * L
        bf, f32 = cfg.cdt, torch.float32          # "bf" = the compute dtype of the GEMM / attention operands
        fp32_mode = bf == f32
        e = lambda *s, dt=bf: torch.empty(*s, device=dev, dtype=dt)
        self.B, self.M = B, M
        P = L - 1
        self.patches = torch.zeros(B * P, cfg.patch_k, device=dev, dtype=bf)   # K padded to a multiple of 8 (TMA pitch)
        self.pe = e(B * P, D)
        self.tokens = e(M, D, dt=f32) if cfg.dialect == "clip" else None
        # per layer, saved for backward
        self.x0 = [e(M, D, dt=f32) for _ in range(dep + 1)]      # block inputs; x0[dep] = tower output
        self.x1 = [e(M, D, dt=f32) for _ in range(dep)]          # after the attention residual
        self.mean1 = e(dep, M, dt=f32); self.rstd1 = e(dep, M, dt=f32)
        self.mean2 = e(dep, M, dt=f32); self.rstd2 = e(dep, M, dt=f32)
        self.xn1 = [torch.zeros(M, D + rp, device=dev, dtype=bf) for _ in range(dep)]   # [LN1(x) | p]
        self.qkv = [e(M, 3 * D) for _ in range(dep)]
        self.o = [e(M, D) for _ in range(dep)]
        self.lse2 = e(dep, B * H * L, dt=f32)
        self.h = [e(M, F) for _ in range(dep)]                   # MLP pre-activation
        # scratch shared by all layers
        self.xn2 = e(M, D)
        self.act = e(M, F)
        self.dh = e(M, F)
        self.dxn = e(M, D)
        self.d_o = e(M, D)
        self.dqkv = torch.zeros(M, 3 * D + rp, device=dev, dtype=bf)   # [dq dk dv | dp]
        self.delta = e(B * H * L, dt=f32)
        # CLIP dialect with a q/v adapter: the update cannot ride in the K extension (raw-reshape placement)
        self.dlt = e(M, 2 * D) if (cfg.dialect == "clip" and rp) else None
        self.g = e(M, D, dt=f32)                                 # residual-gradient stream (fp32, like the forward stream)
        self.gb = self.g if fp32_mode else e(M, D)               # its bf16 copy: A operand of the dX products
        # final LayerNorm on the cls rows
        self.meanf = e(B, dt=f32); self.rstdf = e(B, dt=f32)
        self.yf = e(B, D)                                        # clip: ln_post output (bf16 operand of the projection)
        self.dyf = e(B, D)
        m = cfg.bottleneck
        if m:
            # bottleneck adapter: saved per layer
            self.hf = [e(M, D, dt=f32) for _ in range(dep)]          # MLP output h (adapter input), fp32
            self.za = [e(M, D) for _ in range(dep)]                  # LN_a(h), bf16 operand of the down projection
            self.mean_a = e(dep, M, dt=f32); self.rstd_a = e(dep, M, dt=f32)
            self.zext = []                                           # [z | 1 0...0]: the ones column turns bias grads into token reductions
            for _ in range(dep):
                z = torch.zeros(M, m + 16, device=dev, dtype=bf)
                z[:, m] = 1.0
                self.zext.append(z)
            self.upre = [e(M, m) for _ in range(dep)] if cfg.bn_act != "relu" else None   # pre-activation (gelu_new)
            self.dz = e(M, m)
            self.gh = e(M, D, dt=f32)                                # gradient w.r.t. h (fp32) and its bf16 copy
            self.ghb = self.gh if fp32_mode else e(M, D)
        self.ones = torch.zeros(M, 16, device=dev, dtype=bf)
        self.ones[:, 0] = 1.0                      # bias sums of the fp32 mode and the (tiny) patch-embed bias
        self.mean_pre = e(M, dt=f32); self.rstd_pre = e(M, dt=f32)      # clip ln_pre statistics (its parameters may be trainable)
        if cfg.train_all:
            # everything a weight-gradient product needs that the PEFT path recomputes or never keeps
            self.xn2_l = [e(M, D) for _ in range(dep)]
            self.act_l = [e(M, F) for _ in range(dep)]
        self.generation = 0


class VitEngine:
    def __init__(self, cfg: EngineCfg, device="cuda"):
        if not torch.cuda.is_available():
            raise RuntimeError("peft_b200.VitEngine needs a CUDA device (sm_100a); there is no CPU fallback")
        if cfg.precision not in ("bf16", "fp32"):
            raise ValueError(f"EngineCfg.precision must be 'bf16' or 'fp32', got {cfg.precision!r}")
        if cfg.bottleneck and cfg.precision == "bf16" and cfg.bottleneck not in (16, 32, 64):
            raise NotImplementedError(f"bottleneck width {cfg.bottleneck}: the adapter-gradient reductions (peft_tn_reduce) take "
                                      "16, 32 or 64 columns (the reference uses 64)")
        self.cfg = cfg
        self.dev = torch.device(device)
        self.ws = {}
        self.frozen_loaded = False
        self._fwd_generation = 0
        self._opgen = 0        # bumped whenever the per-ENGINE operands backward reads are rewritten (adapter K-extension columns,
                               # bottleneck weights, re-packed frozen weights): a backward whose forward is older must not run

    # ------------------------------------------------------------------ frozen weights
    def load_frozen(self, sd: dict):
        """sd: reference-named fp32 state dict of the visual tower (any device)."""
        cfg, dev = self.cfg, self.dev
        D, F, dep, rp = cfg.width, cfg.mlp, cfg.depth, cfg.rank_pad
        bf, f32 = cfg.cdt, torch.float32
        keys = _TIMM_KEYS if cfg.dialect == "timm" else _CLIP_KEYS
        g = lambda k: sd[k].detach().to(dev, f32)
        if keys["qkv_b"].format(i=0) not in sd:           # timm qkv_bias=False
            sd = dict(sd)
            for i in range(dep):
                sd[keys["qkv_b"].format(i=i)] = torch.zeros(3 * D)
        L = {k: [g(v.format(i=i)) for i in range(dep)] for k, v in keys.items()}
        # forward operands: W [N, K]; the qkv weight carries rp extra K columns for scale*Up (refreshed every step).
        # The buffers persist across calls (full mode re-packs every step: theta changes with V) and each layer is
        # converted straight into its slot: one fp32 -> bf16 pass, then W^T from the bf16 copy with the transpose kernel.
        def slot(name, *shape):
            t = getattr(self, name, None)
            if t is None or tuple(t.shape) != shape or t.dtype != bf or t.device != dev:
                t = torch.zeros(*shape, device=dev, dtype=bf)
                setattr(self, name, t)
            return t

        def pack(wname, tname, key, n_pad=0, k_pad=0):
            N, K = L[key][0].shape
            W, WT = slot(wname, dep, N, K + k_pad), slot(tname, dep, K, N + n_pad)
            for i in range(dep):
                W[i, :, :K].copy_(L[key][i])
                if bf == torch.bfloat16 and N % 8 == 0 and K % 8 == 0:
                    ops.transpose(W[i, :, :K], WT[i, :, :N])
                else:
                    WT[i, :, :N].copy_(L[key][i].T)

        pack("wqkv", "wqkvT", "qkv_w", n_pad=rp, k_pad=rp)      # [W | Up^T-columns], [W^T | Dn^T] for dX
        pack("wproj", "wprojT", "proj_w")
        pack("wfc1", "wfc1T", "fc1_w")
        pack("wfc2", "wfc2T", "fc2_w")
        vec = lambda k: torch.stack(L[k]).contiguous()
        self.bqkv, self.bproj, self.bfc1, self.bfc2 = vec("qkv_b"), vec("proj_b"), vec("fc1_b"), vec("fc2_b")
        self.ln1_w, self.ln1_b, self.ln2_w, self.ln2_b = vec("ln1_w"), vec("ln1_b"), vec("ln2_w"), vec("ln2_b")
        if cfg.dialect == "timm":
            self.patch_w = self._pad_k(g("patch_embed.proj.weight").reshape(D, -1)).to(bf).contiguous()
            self.patch_b = g("patch_embed.proj.bias").contiguous()
            self.cls = g("cls_token").reshape(D).contiguous()
            self.pos = g("pos_embed").reshape(cfg.seq, D).contiguous()
            self.lnf_w, self.lnf_b = g("norm.weight").contiguous(), g("norm.bias").contiguous()
            self.vproj = None
        else:
            self.patch_w = self._pad_k(g("conv1.weight").reshape(D, -1)).to(bf).contiguous()
            self.patch_b = None
            self.cls = g("class_embedding").reshape(D).contiguous()
            self.pos = g("positional_embedding").reshape(cfg.seq, D).contiguous()
            self.lnpre_w, self.lnpre_b = g("ln_pre.weight").contiguous(), g("ln_pre.bias").contiguous()
            self.lnf_w, self.lnf_b = g("ln_post.weight").contiguous(), g("ln_post.bias").contiguous()
            pr = sd.get("proj")
            if pr is not None and cfg.out_dim > 0:
                self.vproj = g("proj").to(bf).contiguous()            # [D, E]  == B operand [N=D, K=E] of the backward
                self.vprojT = g("proj").T.to(bf).contiguous()         # [E, D]  == B operand [N=E, K=D] of the forward
            else:
                self.vproj = None
        if cfg.dense_delta:
            if cfg.rank != D:
                raise NotImplementedError("the dense q/v update needs rank == width")
            self.wqkvT[:, :, 3 * D:] = torch.eye(D, device=dev, dtype=bf)       # [W^T | I]: dxn = dqkv W + dp
        if rp > 0:
            self.dn_bf = torch.zeros(dep, rp, D, device=dev, dtype=bf)
            self.upT_bf = torch.zeros(dep, rp, 3 * D, device=dev, dtype=bf)
            if cfg.dialect == "clip":
                self.up2_bf = torch.zeros(dep, 2 * D, rp, device=dev, dtype=bf)     # [Up_q ; Up_v]: B operand [N=2D, K=rp]
                self.up2T_bf = torch.zeros(dep, rp, 2 * D, device=dev, dtype=bf)    # its transpose: B operand of dp
        self.act_fwd = ops.EPI_GELU if cfg.dialect == "timm" else ops.EPI_QUICKGELU
        self.act_bwd = ops.EPI_DGELU if cfg.dialect == "timm" else ops.EPI_DQUICKGELU
        self.frozen_loaded = True
        self._opgen += 1

    # vectors of the tower (biases, LayerNorm gains / biases): the tensors that "bias tuning" / "LayerNorm tuning" train TOGETHER
    # with an adapter (FL/tools/finetune_bias.py, finetune_layernorm.py).  They are re-read every forward (cheap) and their
    # gradients are column sums / LayerNorm parameter gradients — no weight-gradient product is involved.
    SMALL_SLOTS = ("ln1_w", "ln1_b", "qkv_b", "proj_b", "ln2_w", "ln2_b", "fc1_b", "fc2_b")

    def small_keys(self):
        keys = _TIMM_KEYS if self.cfg.dialect == "timm" else _CLIP_KEYS
        out = {keys[s].format(i=i): (s, i) for s in self.SMALL_SLOTS for i in range(self.cfg.depth)}
        fin = ("norm.weight", "norm.bias") if self.cfg.dialect == "timm" else ("ln_post.weight", "ln_post.bias")
        out[fin[0]] = ("lnf_w", -1)
        out[fin[1]] = ("lnf_b", -1)
        if self.cfg.dialect == "timm":
            out["patch_embed.proj.bias"] = ("patch_b", -1)
        else:
            out["ln_pre.weight"] = ("lnpre_w", -1)
            out["ln_pre.bias"] = ("lnpre_b", -1)
        return out

    def load_small(self, named):
        """named = {reference key: tensor} of trainable tower vectors: refresh the engine's copies (every forward)."""
        table = self.small_keys()
        slots = dict(ln1_w=self.ln1_w, ln1_b=self.ln1_b, qkv_b=self.bqkv, proj_b=self.bproj, ln2_w=self.ln2_w, ln2_b=self.ln2_b,
                     fc1_b=self.bfc1, fc2_b=self.bfc2)
        with torch.no_grad():
            for k, t in named.items():
                s_, i = table[k]
                if i < 0:
                    getattr(self, {"patch_b": "patch_b"}.get(s_, s_)).copy_(t.detach().to(self.dev, torch.float32).reshape(-1))
                else:
                    slots[s_][i].copy_(t.detach().to(self.dev, torch.float32).reshape(-1))
        self._opgen += 1

    def _pad_k(self, w2d):
        kp = self.cfg.patch_k
        return w2d if w2d.shape[1] == kp else torch.nn.functional.pad(w2d, (0, kp - w2d.shape[1]))

    def _workspace(self, B):
        w = self.ws.get(B)
        if w is None:
            w = self.ws[B] = _Workspace(self.cfg, B, self.dev)
        return w

    # ------------------------------------------------------------------ adapter operand refresh (per step)
    def _refresh_adapter(self, Dn, UpT):
        cfg = self.cfg
        D, r = cfg.width, cfg.rank
        if r == 0:
            return
        with torch.no_grad():
            self.upT_bf[:, :r].copy_(UpT)
            if cfg.dense_delta and cfg.dialect == "timm":
                self.wqkv[:, :, D:D + r].copy_(UpT.transpose(1, 2))
                return
            if not cfg.dense_delta:
                self.dn_bf[:, :r].copy_(Dn)
                self.wqkvT[:, :, 3 * D:3 * D + r].copy_(Dn.transpose(1, 2))
            if cfg.dialect == "clip":
                self.up2T_bf[:, :r, :D].copy_(UpT[:, :, :D])
                self.up2T_bf[:, :r, D:].copy_(UpT[:, :, 2 * D:])
                self.up2_bf.copy_(self.up2T_bf.transpose(1, 2))
            else:
                self.wqkv[:, :, D:D + r].copy_(UpT.transpose(1, 2))

    # ------------------------------------------------------------------ forward
    def _refresh_bottleneck(self, bn):
        """bn = (Wd [dep,m,D], bd [dep,m], Wu [dep,D,m], bu [dep,D], ln_g [dep,D], ln_b [dep,D]) fp32."""
        Wd, bd, Wu, bu, lg, lb = bn
        bf = self.cfg.cdt
        with torch.no_grad():
            self.wd_bf = Wd.detach().to(bf).contiguous()                        # down: B [N=m, K=D]
            self.wdT_bf = Wd.detach().transpose(1, 2).to(bf).contiguous()       # dza : B [N=D, K=m]
            self.wu_bf = Wu.detach().to(bf).contiguous()                        # up  : B [N=D, K=m]
            self.wuT_bf = Wu.detach().transpose(1, 2).to(bf).contiguous()       # dz  : B [N=m, K=D]
            self.bd_f = bd.detach().float().contiguous()
            self.bu_f = bu.detach().float().contiguous()
            self.lga = lg.detach().float().contiguous()
            self.lba = lb.detach().float().contiguous()

    def forward(self, images, Dn=None, UpT=None, bn=None, drop=None):
        """`drop`: optional fp32 [depth, 2, B*L] per-row stochastic-depth factors keep/keep_prob of the attention ([:, 0]) and
        MLP ([:, 1]) branches (timm DropPath, FS/kadaptation.py:263,276-277); folded into the two residual epilogues."""
        cfg = self.cfg
        if not self.frozen_loaded:
            raise RuntimeError("VitEngine.forward before load_frozen()")
        if images.device != self.dev and images.device.type != "cuda":
            raise RuntimeError("VitEngine: images must be on the CUDA device")
        D, L, H, dep, rp = cfg.width, cfg.seq, cfg.heads, cfg.depth, cfg.rank_pad
        B = images.shape[0]
        w = self._workspace(B)
        self._refresh_adapter(Dn, UpT)
        m = cfg.bottleneck
        if drop is not None:
            if m or cfg.train_all:
                raise NotImplementedError("stochastic depth is supported for the q/v-adapter (PEFT) engine, not with a bottleneck "
                                          "adapter or in full mode")
            drop = drop.detach().to(self.dev, torch.float32).contiguous()
            if tuple(drop.shape) != (dep, 2, B * L):
                raise ValueError(f"drop must be [depth, 2, B*L] = {(dep, 2, B * L)}, got {tuple(drop.shape)}")
        w.drop = drop
        rs = (lambda i, j: None) if drop is None else (lambda i, j: drop[i, j])
        if m:
            if bn is None:
                raise RuntimeError("VitEngine.forward: bottleneck adapter operands missing")
            self._refresh_bottleneck(bn)
        images = images.contiguous().float()
        # patch embedding: unfold -> GEMM (+bias) -> prepend cls, add positions
        ops.im2col_patches(images, w.patches, cfg.patch)
        ops.gemm(w.patches, self.patch_w, w.pe, bias=self.patch_b)
        if cfg.dialect == "clip":
            ops.assemble_tokens(w.pe, self.cls, self.pos, w.tokens, B, L, D)
            ops.layernorm_fwd(w.tokens, self.lnpre_w, self.lnpre_b, w.x0[0], w.mean_pre, w.rstd_pre, cfg.eps)
        else:
            ops.assemble_tokens(w.pe, self.cls, self.pos, w.x0[0], B, L, D)
        scale = 64 ** -0.5
        for i in range(dep):
            xn1 = w.xn1[i]
            ops.layernorm_fwd(w.x0[i], self.ln1_w[i], self.ln1_b[i], xn1[:, :D], w.mean1[i], w.rstd1[i], cfg.eps)
            if rp and cfg.dialect == "clip":
                # reference order (FS/model.py:967,981-986): q/sqrt(hd) first, then += raw-reshaped update (Q1, Q2)
                if cfg.dense_delta:
                    xn1[:, D:].copy_(xn1[:, :D])                                   # Dn = I: p is LN1(x) itself
                else:
                    ops.gemm(xn1[:, :D], self.dn_bf[i], xn1[:, D:])                # p = LN1(x) Dn^T
                ops.gemm(xn1[:, :D], self.wqkv[i][:, :D], w.qkv[i], bias=self.bqkv[i])
                ops.gemm(xn1[:, D:], self.up2_bf[i], w.dlt)                        # [dq | dv] = p Up^T
                ops.clip_delta(w.qkv[i], w.dlt, B, H, L, 1.0 / scale, 1.0)
            elif rp:
                if cfg.dense_delta:
                    xn1[:, D:].copy_(xn1[:, :D])                                   # Dn = I: p is LN1(x) itself
                else:
                    ops.gemm(xn1[:, :D], self.dn_bf[i], xn1[:, D:])                # p = LN1(x) Dn^T
                ops.gemm(xn1, self.wqkv[i], w.qkv[i], bias=self.bqkv[i])           # [x|p] [W|Up]^T + b
            else:
                ops.gemm(xn1[:, :D], self.wqkv[i][:, :D], w.qkv[i], bias=self.bqkv[i])
            ops.attention_fwd(w.qkv[i], w.o[i], w.lse2[i], B, H, L, scale)
            ops.gemm(w.o[i], self.wproj[i], w.x1[i], bias=self.bproj[i], epilogue=ops.EPI_RESID_F32, aux=w.x0[i], rowscale=rs(i, 0))
            xn2 = w.xn2_l[i] if cfg.train_all else w.xn2
            act = w.act_l[i] if cfg.train_all else w.act
            ops.layernorm_fwd(w.x1[i], self.ln2_w[i], self.ln2_b[i], xn2, w.mean2[i], w.rstd2[i], cfg.eps)
            ops.gemm(xn2, self.wfc1[i], w.h[i], bias=self.bfc1[i], epilogue=self.act_fwd, out2=act)
            if not cfg.has_bn(i):
                ops.gemm(act, self.wfc2[i], w.x0[i + 1], bias=self.bfc2[i], epilogue=ops.EPI_RESID_F32, aux=w.x1[i], rowscale=rs(i, 1))
            else:
                # x_{i+1} = x1 + h + up(act(down(LN_a(h)))),  h = mlp(ln2(x1))     (adapter_model.py:330-334 computed once, Q6)
                ops.gemm(act, self.wfc2[i], w.hf[i], bias=self.bfc2[i], epilogue=ops.EPI_STORE_F32)
                ops.layernorm_fwd_add(w.hf[i], self.lga[i], self.lba[i], w.za[i], w.mean_a[i], w.rstd_a[i], w.x1[i], w.x0[i + 1], 1e-5)
                z = w.zext[i][:, :m]
                if cfg.bn_act == "relu":
                    ops.gemm(w.za[i], self.wd_bf[i], z, bias=self.bd_f[i], epilogue=ops.EPI_RELU)
                else:
                    ops.gemm(w.za[i], self.wd_bf[i], w.upre[i], bias=self.bd_f[i], epilogue=ops.EPI_GELU_TANH, out2=z)
                ops.gemm(z, self.wu_bf[i], w.x0[i + 1], bias=self.bu_f[i], epilogue=ops.EPI_RESID_F32, aux=w.x0[i + 1])
        # final LayerNorm on the cls rows (+ CLIP projection)
        xcls = w.x0[dep].view(B, L * D)[:, :D]
        if self.vproj is None:
            feat = torch.empty(B, D, device=self.dev, dtype=torch.float32)
            ops.layernorm_fwd(xcls, self.lnf_w, self.lnf_b, feat, w.meanf, w.rstdf, cfg.eps)
        else:
            ops.layernorm_fwd(xcls, self.lnf_w, self.lnf_b, w.yf, w.meanf, w.rstdf, cfg.eps)
            feat = torch.empty(B, cfg.out_dim, device=self.dev, dtype=torch.float32)
            ops.gemm(w.yf, self.vprojT, feat, epilogue=ops.EPI_STORE_F32)
        w.generation += 1
        if cfg.rank or cfg.bottleneck:
            self._opgen += 1           # _refresh_adapter / _refresh_bottleneck rewrote operands the backward reads
        self._fwd_generation = (B, w.generation, self._opgen)
        return feat

    # ------------------------------------------------------------------ backward
    def backward(self, dfeat, generation=None, small=None):
        """`generation` = the engine's `_fwd_generation` right after the forward this backward belongs to (the autograd
        node records it); a later forward at the same batch size has overwritten the saved activations.
        `small`: reference keys of tower vectors (biases, LayerNorm parameters) whose gradients are wanted as well; they come
        back as the 4th result {key: fp32 grad}."""
        cfg = self.cfg
        D, L, H, dep, r, rp = cfg.width, cfg.seq, cfg.heads, cfg.depth, cfg.rank, cfg.rank_pad
        B = dfeat.shape[0]
        w = self._workspace(B)
        if (generation or self._fwd_generation) != (B, w.generation, self._opgen):
            raise RuntimeError("VitEngine.backward: the saved activations or the adapter / weight operands of this forward were "
                               "overwritten by a later forward (run backward before the next forward of the same engine)")
        dfeat = dfeat.contiguous().float()
        w.g.zero_()
        w.gb.zero_()
        drop = getattr(w, "drop", None)
        # ---- gradients of trainable tower vectors (bias / LayerNorm tuning next to an adapter)
        SG = {}
        table = self.small_keys() if small else {}
        want_slot = {}
        for k in (small or ()):
            want_slot[table[k]] = k
        if small and drop is not None:
            raise NotImplementedError("stochastic depth together with trainable tower biases / LayerNorms is not supported")
        zvec = lambda n_: torch.zeros(n_, device=self.dev, dtype=torch.float32)

        def ln_pg(slot_w, slot_b, i):
            """(dgamma, dbeta) buffers if either LayerNorm parameter of this slot is trainable, else (None, None)."""
            kw_, kb_ = want_slot.get((slot_w, i)), want_slot.get((slot_b, i))
            if kw_ is None and kb_ is None:
                return None, None
            dg_, db_ = zvec(D), zvec(D)
            if kw_ is not None:
                SG[kw_] = dg_
            if kb_ is not None:
                SG[kb_] = db_
            return dg_, db_

        def bias_grad(slot, i, dY):
            k_ = want_slot.get((slot, i))
            if k_ is None:
                return
            n_ = dY.shape[1]
            if cfg.precision == "fp32":
                c16 = torch.zeros(16, n_, device=self.dev, dtype=torch.float32)
                ops.tn_reduce(w.ones[:dY.shape[0]], dY, c16)
                SG[k_] = c16[0]
            else:
                SG[k_] = ops.colsum(dY, zvec(n_))

        need_x0 = any(sl in ("patch_b", "lnpre_w", "lnpre_b") for sl, _ in want_slot)     # gradient w.r.t. the tower input
        xcls = w.x0[dep].view(B, L * D)[:, :D]
        gcls = w.g.view(B, L * D)[:, :D]
        gbcls = w.gb.view(B, L * D)[:, :D]
        dgf, dbf = ln_pg("lnf_w", "lnf_b", -1)
        if self.vproj is None:
            ops.layernorm_bwd_f32g(dfeat, xcls, w.meanf, w.rstdf, self.lnf_w, None, gcls, gbcls, dgf, dbf)
        else:
            ops.gemm(dfeat.to(cfg.cdt), self.vproj, w.dyf)                       # dy = dfeat proj^T
            ops.layernorm_bwd_f32g(w.dyf, xcls, w.meanf, w.rstdf, self.lnf_w, None, gcls, gbcls, dgf, dbf)
        rs = (lambda i, j: None) if drop is None else (lambda i, j: drop[i, j])
        dDn = dUpT = None
        if r:
            dDn = torch.zeros(dep, rp, D, device=self.dev, dtype=torch.float32)
            dUpT = torch.zeros(dep, rp, 3 * D, device=self.dev, dtype=torch.float32)
        scale = 64 ** -0.5
        m = cfg.bottleneck
        bng = None
        if m:
            f32 = torch.float32
            z0 = lambda *sh: torch.zeros(*sh, device=self.dev, dtype=f32)
            dWuT, dWd = z0(dep, m, D), z0(dep, m, D)
            dbu16, dbd16 = z0(dep, 16, D), z0(dep, 16, m)
            dlg, dlb = z0(dep, D), z0(dep, D)
            bn_dact = ops.EPI_DRELU if cfg.bn_act == "relu" else ops.EPI_DGELU_TANH
        for i in reversed(range(dep)):
            gmlp = w.gb                  # gradient w.r.t. the MLP output h (bf16 operand)
            if cfg.has_bn(i):
                # bottleneck adapter backward; g = dL/dx_{i+1} is also dL/dx1 (residual) and part of dL/dh
                z, ones = w.zext[i][:, :m], w.zext[i][:, m:m + 16]
                bsum = (lambda x, o16: ops.tn_reduce(ones, x, o16)) if cfg.precision == "fp32" else (lambda x, o16: ops.colsum(x, o16[0]))
                bsum(w.gb, dbu16[i])                                                  # dbu  = colsum(g)       (row 0)
                ops.tn_reduce(z, w.gb, dWuT[i])                                       # dWu^T = z^T g
                ops.gemm(w.gb, self.wuT_bf[i], w.dz, epilogue=bn_dact, aux=(z if cfg.bn_act == "relu" else w.upre[i]))
                bsum(w.dz, dbd16[i])                                                  # dbd  = colsum(dz)
                ops.tn_reduce(w.dz, w.za[i], dWd[i])                                  # dWd  = dz^T LN_a(h)
                ops.gemm(w.dz, self.wdT_bf[i], w.dxn)                                 # dza  = dz Wd
                ops.layernorm_bwd_f32g(w.dxn, w.hf[i], w.mean_a[i], w.rstd_a[i], self.lga[i], w.g, w.gh, w.ghb, dlg[i], dlb[i])
                gmlp = w.ghb
            # MLP: dh = (g W2) * act'(h) ; dxn2 = dh W1 ; g += LN2'(dxn2)
            bias_grad("fc2_b", i, gmlp)
            ops.gemm(gmlp, self.wfc2T[i], w.dh, epilogue=self.act_bwd, aux=w.h[i], rowscale=rs(i, 1))
            bias_grad("fc1_b", i, w.dh)
            ops.gemm(w.dh, self.wfc1T[i], w.dxn)
            dg2, db2 = ln_pg("ln2_w", "ln2_b", i)
            ops.layernorm_bwd_f32g(w.dxn, w.x1[i], w.mean2[i], w.rstd2[i], self.ln2_w[i], w.g, w.g, w.gb, dg2, db2)
            if i == 0 and not rp and not small:
                break       # nothing trainable is upstream of block 0's attention when there is no q/v adapter
            # attention: d_o = g Wproj ; dqkv = attn'(d_o)
            bias_grad("proj_b", i, w.gb)
            ops.gemm(w.gb, self.wprojT[i], w.d_o, rowscale=rs(i, 0))
            ops.attention_bwd(w.qkv[i], w.o[i], w.d_o, w.lse2[i], w.delta, w.dqkv, B, H, L, scale)
            if rp and cfg.dialect == "clip":
                dq = w.dqkv[:, :3 * D]
                ops.clip_delta(dq, w.dlt, B, H, L, 1.0 / scale, 1.0, backward=True)  # d[dq|dv] gathered back (Q1 inverse)
                ops.gemm(w.dlt, self.up2T_bf[i], w.dqkv[:, 3 * D:])                 # dp = d[dq|dv] [Up_q;Up_v]
                if rp > 128 and cfg.precision != "fp32":
                    d2 = torch.empty(rp, 2 * D, device=self.dev, dtype=torch.float32)
                    self._dw(w, w.xn1[i][:, D:], w.dlt, d2)                         # d[Up_q|Up_v]^T = p^T d[dq|dv]: one tensor-core product
                    dUpT[i, :, :D].copy_(d2[:, :D])
                    dUpT[i, :, 2 * D:].copy_(d2[:, D:])
                else:
                    for c, cw in _tn_chunks(rp):
                        pc = w.xn1[i][:, D + c:D + c + cw]
                        ops.tn_reduce(pc, w.dlt[:, :D], dUpT[i, c:c + cw, :D])
                        ops.tn_reduce(pc, w.dlt[:, D:], dUpT[i, c:c + cw, 2 * D:])
                if not cfg.dense_delta:
                    for c, cw in _tn_chunks(rp):
                        ops.tn_reduce(w.dqkv[:, 3 * D + c:3 * D + c + cw], w.xn1[i][:, :D], dDn[i, c:c + cw])
            elif rp > 128 and cfg.precision != "fp32":
                dq = w.dqkv[:, :3 * D]
                ops.gemm(dq, self.upT_bf[i], w.dqkv[:, 3 * D:])                    # dp = dqkv Up
                self._dw(w, w.xn1[i][:, D:], dq, dUpT[i])                           # dUp^T = p^T dqkv as ONE tensor-core product
                if not cfg.dense_delta:
                    for c, cw in _tn_chunks(rp):
                        ops.tn_reduce(w.dqkv[:, 3 * D + c:3 * D + c + cw], w.xn1[i][:, :D], dDn[i, c:c + cw])
            elif rp:
                dq = w.dqkv[:, :3 * D]
                ops.gemm(dq, self.upT_bf[i], w.dqkv[:, 3 * D:])                    # dp = dqkv Up
                for c, cw in _tn_chunks(rp):                                        # adapter-gradient token reductions
                    ops.tn_reduce(w.xn1[i][:, D + c:D + c + cw], dq, dUpT[i, c:c + cw])            # dUp^T = p^T dqkv
                    if not cfg.dense_delta:
                        ops.tn_reduce(w.dqkv[:, 3 * D + c:3 * D + c + cw], w.xn1[i][:, :D], dDn[i, c:c + cw])  # dDn = dp^T x
            bias_grad("qkv_b", i, w.dqkv[:, :3 * D])
            dg1, db1 = ln_pg("ln1_w", "ln1_b", i)
            if i == 0 and dg1 is None and not need_x0:
                break       # block 0's input gradient feeds only the frozen patch embedding: no dX product, no LN1'
            if rp:
                ops.gemm(w.dqkv, self.wqkvT[i], w.dxn)                              # dxn1 = [dqkv|dp] [W^T|Dn^T]^T
            else:
                ops.gemm(w.dqkv[:, :3 * D], self.wqkvT[i][:, :3 * D], w.dxn)
            ops.layernorm_bwd_f32g(w.dxn, w.x0[i], w.mean1[i], w.rstd1[i], self.ln1_w[i], w.g, w.g, w.gb, dg1, db1)
        # the patch embedding / ln_pre weights are frozen and the images need no gradient: only their trainable VECTORS go on
        if need_x0:
            g3 = w.g.view(B, L, D)
            if cfg.dialect == "clip":
                dgp, dbp = ln_pg("lnpre_w", "lnpre_b", -1)
                ops.layernorm_bwd_f32g(w.g, w.tokens, w.mean_pre, w.rstd_pre, self.lnpre_w, None, w.x1[0], None, dgp, dbp)
            else:
                SG[want_slot[("patch_b", -1)]] = g3[:, 1:, :].sum((0, 1))        # conv bias: sum over the patch tokens (glue)
        if m:
            bng = (dWd, dbd16[:, 0], dWuT.transpose(1, 2), dbu16[:, 0], dlg, dlb)
        if r:
            return (None if cfg.dense_delta else dDn[:, :r]), dUpT[:, :r], bng, SG
        return None, None, bng, SG



    # ------------------------------------------------------------------ backward, full fine-tuning mode
    def _dw(self, w, dY, X, out):
        """out[N, K] (fp32) = dY[M, N]^T X[M, K], contraction over the M tokens: one tcgen05 product straight from the
        token-major activations (MN-major operands, split-K when the output has few tiles) — no transposes."""
        if self.cfg.precision == "fp32":           # verification mode: one fp32 token reduction (any shape)
            out.zero_()
            ops.tn_reduce(dY, X, out)
            return
        ops.gemm_tn(dY, X, out)

    def _dw_db(self, w, G, want, kw, kb, dY, X):
        """Weight and bias gradient of one linear layer (either may be unwanted): G[kw] = dY^T X, G[kb] = colsum(dY)."""
        N, K = dY.shape[1], X.shape[1]
        if want(kb):
            if self.cfg.precision == "fp32":
                c16 = torch.zeros(16, N, device=self.dev, dtype=torch.float32)
                ops.tn_reduce(w.ones[:dY.shape[0]], dY, c16)
                G[kb] = c16[0]
            else:
                G[kb] = ops.colsum(dY, torch.zeros(N, device=self.dev, dtype=torch.float32))
        if want(kw):
            dw = torch.empty(N, K, device=self.dev, dtype=torch.float32)
            self._dw(w, dY, X, dw)
            G[kw] = dw

    def backward_full(self, dfeat, need=None, generation=None):
        """dL/d(every tower tensor) -> {reference key: fp32 grad}.  Only for cfg.train_all (no adapters): the Fastfood /
        full-fine-tuning configuration C5, where theta is a function of V and autograd needs every dW
        (FL/intrinsic/fastfood.py:61-77).  Same kernels as the PEFT backward plus, per linear layer, one weight-gradient
        product (two transposes + GEMM) and one bias column-sum; LayerNorm backward also emits dgamma / dbeta."""
        cfg = self.cfg
        if not cfg.train_all or cfg.rank or cfg.bottleneck:
            raise RuntimeError("backward_full needs an engine built with train_all=True and no adapter")
        D, L, H, dep, F = cfg.width, cfg.seq, cfg.heads, cfg.depth, cfg.mlp
        B = dfeat.shape[0]
        w = self._workspace(B)
        if (generation or self._fwd_generation) != (B, w.generation, self._opgen):
            raise RuntimeError("VitEngine.backward: the saved activations or the adapter / weight operands of this forward were "
                               "overwritten by a later forward (run backward before the next forward of the same engine)")
        dev, f32 = self.dev, torch.float32
        z = lambda *sh: torch.zeros(*sh, device=dev, dtype=f32)
        e = lambda *sh: torch.empty(*sh, device=dev, dtype=f32)
        dfeat = dfeat.contiguous().float()
        keys = _TIMM_KEYS if cfg.dialect == "timm" else _CLIP_KEYS
        G = {}
        want = (lambda name: True) if need is None else (lambda name: name in need)   # bias-only / LN-only tuning skip the dW products
        w.g.zero_()
        w.gb.zero_()
        xcls = w.x0[dep].view(B, L * D)[:, :D]
        gcls = w.g.view(B, L * D)[:, :D]
        gbcls = w.gb.view(B, L * D)[:, :D]
        dlnf_w, dlnf_b = z(D), z(D)
        if self.vproj is None:
            ops.layernorm_bwd_f32g(dfeat, xcls, w.meanf, w.rstdf, self.lnf_w, None, gcls, gbcls, dlnf_w, dlnf_b)
        else:
            dfb = dfeat.to(cfg.cdt)
            if want("proj"):
                dproj = e(D, cfg.out_dim)
                self._dw(w, w.yf, dfb, dproj)                                      # proj [D, E] = yf^T dfeat
                G["proj"] = dproj
            ops.gemm(dfb, self.vproj, w.dyf)
            ops.layernorm_bwd_f32g(w.dyf, xcls, w.meanf, w.rstdf, self.lnf_w, None, gcls, gbcls, dlnf_w, dlnf_b)
        if cfg.dialect == "timm":
            G["norm.weight"], G["norm.bias"] = dlnf_w, dlnf_b
        else:
            G["ln_post.weight"], G["ln_post.bias"] = dlnf_w, dlnf_b
        scale = 64 ** -0.5
        dq = w.dqkv[:, :3 * D]
        for i in reversed(range(dep)):
            k = lambda n: keys[n].format(i=i)
            # MLP
            self._dw_db(w, G, want, k("fc2_w"), k("fc2_b"), w.gb, w.act_l[i])
            ops.gemm(w.gb, self.wfc2T[i], w.dh, epilogue=self.act_bwd, aux=w.h[i])
            self._dw_db(w, G, want, k("fc1_w"), k("fc1_b"), w.dh, w.xn2_l[i])
            ops.gemm(w.dh, self.wfc1T[i], w.dxn)
            dg, db = z(D), z(D)
            ops.layernorm_bwd_f32g(w.dxn, w.x1[i], w.mean2[i], w.rstd2[i], self.ln2_w[i], w.g, w.g, w.gb, dg, db)
            G[k("ln2_w")], G[k("ln2_b")] = dg, db
            # attention
            self._dw_db(w, G, want, k("proj_w"), k("proj_b"), w.gb, w.o[i])
            ops.gemm(w.gb, self.wprojT[i], w.d_o)
            ops.attention_bwd(w.qkv[i], w.o[i], w.d_o, w.lse2[i], w.delta, w.dqkv, B, H, L, scale)
            self._dw_db(w, G, want, k("qkv_w"), k("qkv_b"), dq, w.xn1[i][:, :D])
            ops.gemm(dq, self.wqkvT[i][:, :3 * D], w.dxn)
            dg, db = z(D), z(D)
            ops.layernorm_bwd_f32g(w.dxn, w.x0[i], w.mean1[i], w.rstd1[i], self.ln1_w[i], w.g, w.g, w.gb, dg, db)
            G[k("ln1_w")], G[k("ln1_b")] = dg, db
        # embedding: w.g = dL/dx0 (fp32 [B, L, D])
        if cfg.dialect == "clip":
            dg, db = z(D), z(D)
            fp32_mode = cfg.precision == "fp32"
            ops.layernorm_bwd_f32g(w.g, w.tokens, w.mean_pre, w.rstd_pre, self.lnpre_w, None, w.x1[0],
                                   None if fp32_mode else w.gb, dg, db)
            G["ln_pre.weight"], G["ln_pre.bias"] = dg, db
            g3 = w.x1[0].view(B, L, D)          # x1[0] is dead by now: reuse it for dL/d(tokens before ln_pre)
            gop = g3 if fp32_mode else w.gb.view(B, L, D)    # operand copy of it in the compute dtype
        else:
            g3 = w.g.view(B, L, D)
            gop = w.gb.view(B, L, D)            # (fp32 mode: gb aliases g)
        dpos = g3.sum(0)                                     # [L, D] batch reduction of the token gradient (glue)
        P = L - 1
        dpe = gop[:, 1:, :].reshape(B * P, D)
        kk = 3 * cfg.patch * cfg.patch
        dconv = e(D, kk)
        if want("patch_embed.proj.weight" if cfg.dialect == "timm" else "conv1.weight"):
            self._dw(w, dpe, w.patches[:, :kk], dconv)
        else:
            dconv.zero_()
        if cfg.dialect == "timm":
            c16 = z(16, D)
            ops.tn_reduce(w.ones[:B * P], dpe, c16)
            G["patch_embed.proj.weight"] = dconv.view(D, 3, cfg.patch, cfg.patch)
            G["patch_embed.proj.bias"] = c16[0]
            G["cls_token"] = dpos[0].view(1, 1, D)
            G["pos_embed"] = dpos.view(1, L, D)
        else:
            G["conv1.weight"] = dconv.view(D, 3, cfg.patch, cfg.patch)
            G["class_embedding"] = dpos[0]
            G["positional_embedding"] = dpos
        return G


class _BackboneFullFn(torch.autograd.Function):
    """Tower with EVERY tensor differentiable: forward re-reads the weights (they are functions of V under FastfoodWrap),
    backward returns dL/dW for each of them in the order they were passed."""

    @staticmethod
    def forward(ctx, engine, images, names, *tensors):
        ctx.engine, ctx.names = engine, names
        ctx.need = {n for n, t in zip(names, tensors) if t.requires_grad}
        engine.load_frozen({n: t.detach() for n, t in zip(names, tensors)})
        out = engine.forward(images)
        ctx.generation = engine._fwd_generation
        return out

    @staticmethod
    def backward(ctx, dfeat):
        G = ctx.engine.backward_full(dfeat, ctx.need, ctx.generation)
        return (None, None, None) + tuple(G.get(n) if n in ctx.need else None for n in ctx.names)


def backbone_full(engine: VitEngine, images, named_tensors):
    """Differentiable w.r.t. every tower tensor; `named_tensors` = [(reference key, tensor)]."""
    names = tuple(n for n, _ in named_tensors)
    return _BackboneFullFn.apply(engine, images, names, *[t for _, t in named_tensors])


class _BackboneFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, engine, images, Dn, UpT, drop, nbn, small_names, *rest):
        ctx.engine = engine
        ctx.nbn = nbn
        bn, small = rest[:nbn], rest[nbn:]
        ctx.small_names = small_names
        if small_names:
            engine.load_small(dict(zip(small_names, small)))
        out = engine.forward(images, Dn, UpT, bn if bn else None, drop)
        ctx.generation = engine._fwd_generation
        return out

    @staticmethod
    def backward(ctx, dfeat):
        dDn, dUpT, bng, SG = ctx.engine.backward(dfeat, ctx.generation, ctx.small_names or None)
        extra = tuple(bng) if bng is not None else (None,) * ctx.nbn
        small = tuple(SG.get(k) for k in ctx.small_names)
        return (None, None, dDn, dUpT, None, None, None) + extra + small


def backbone(engine: VitEngine, images, Dn=None, UpT=None, bn=None, drop=None, small=None):
    """Differentiable (w.r.t. Dn / UpT / the bottleneck-adapter operands and, optionally, trainable tower vectors `small` =
    [(reference key, tensor)]: biases and LayerNorm parameters) visual-tower feature extraction."""
    bn = tuple(bn) if bn is not None else ()
    small = list(small or [])
    names = tuple(k for k, _ in small)
    return _BackboneFn.apply(engine, images, Dn, UpT, drop, len(bn), names, *bn, *[t for _, t in small])
