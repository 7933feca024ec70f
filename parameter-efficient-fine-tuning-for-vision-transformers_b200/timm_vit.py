"""timm-dialect ViT with PEFT adapters — drop-in for the reference classes of
few_shot/vision_benchmark/evaluation/{kadaptation,lora,adapter}.py (= full_shot/main/lib/models/cls_vit_*.py).

Same class names, constructor arguments, parameter names and shapes (so `state_dict()` / `load_state_dict()` and the
trainers' name filters — `'W' in name`, `'adapter' in name` — work unchanged), but `VisionTransformer.forward` runs the
whole tower through the CUDA engine (engine.py) instead of ATen.  `Mlp`, the adapter-free `Attention`, `Adapter`, `Block` and
`PatchEmbed` also have stand-alone forwards (eager.py: one autograd node each over the same C-ABI kernels, used by the per-
component intrinsic-dimension model, intrinsic_vit.py); the q/v-adapted attention and the PHM modules run inside the engine
only and raise when called alone.  There is no PyTorch-math fallback path anywhere.

Adapter hyper-parameters are constructor kwargs with the reference's values as defaults where the reference is
usable: LoRA r=4, alpha=128 (FS/lora.py); KAdaptation phm_rank=1, init 0.01, and phm_dim=4 instead of the committed
768 (north_star formula / headline config; the reference leaves phm_rule uninitialised — here it is
N(0, 0.01^2)); Houlsby bottleneck 64 with ReLU (FS/adapter.py).
"""
from __future__ import annotations

from functools import partial

import torch
import torch.nn as nn

from . import eager, ops
from .engine import EngineCfg, VitEngine, backbone, backbone_full


def _no_eager(self, *a, **k):
    raise RuntimeError(f"{type(self).__name__} has no eager forward in peft_b200: run the enclosing VisionTransformer "
                       "(CUDA engine); there is no PyTorch fallback path")


class Mlp(nn.Module):
    """FS/kadaptation.py:57-78."""

    def __init__(self, in_features, hidden_features=None, out_features=None, act_layer=nn.GELU, drop=0.):
        super().__init__()
        out_features = out_features or in_features
        hidden_features = hidden_features or in_features
        if drop != 0.:
            raise NotImplementedError("dropout > 0 is not supported (reference configs use 0)")
        self.fc1 = nn.Linear(in_features, hidden_features)
        self.act = act_layer()
        self.fc2 = nn.Linear(hidden_features, out_features)
        self.fp32 = False

    def forward(self, x):
        """Stand-alone forward (eager.py): one autograd node, weights read through attribute access on every call so that a
        FastfoodWrap around this module works (FL/lib/models/cls_intrinsic_dimension.py:470-473)."""
        act = ops.EPI_GELU if isinstance(self.act, nn.GELU) else ops.EPI_QUICKGELU
        return eager.mlp(x, self.fc1.weight, self.fc1.bias, self.fc2.weight, self.fc2.bias, act, self.fp32)


class Attention(nn.Module):
    """FS/kadaptation.py:117-238 (adapter='kron'), FS/lora.py Attention (adapter='lora'), plain (adapter='none')."""

    def __init__(self, dim, num_heads=8, qkv_bias=False, qk_scale=None, attn_drop=0., proj_drop=0., res_score=False,
                 adapter="kron", phm_dim=4, phm_rank=1, phm_init_range=0.01, lora_attn_dim=4, lora_attn_alpha=128):
        super().__init__()
        if attn_drop != 0. or proj_drop != 0. or res_score or qk_scale is not None:
            raise NotImplementedError("attn/proj dropout, res_score and qk_scale overrides are not supported")
        if dim // num_heads != 64:
            raise NotImplementedError("the attention kernels are specialised for head_dim 64")
        self.num_heads = num_heads
        self.scale = (dim // num_heads) ** -0.5
        self.adapter_kind = adapter
        self.qkv = nn.Linear(dim, dim * 3, bias=qkv_bias)
        self.proj = nn.Linear(dim, dim)
        if adapter == "kron":
            assert dim % phm_dim == 0
            self.phm_rank = phm_rank
            self.phm_init_range = phm_init_range
            n, k = phm_dim, dim // phm_dim
            self.phm_rule = nn.Parameter(torch.empty(n, n, n).normal_(0, phm_init_range))
            self.W_left1 = nn.Parameter(torch.empty(n, k, phm_rank).normal_(0, phm_init_range))
            self.W_right1 = nn.Parameter(torch.empty(n, phm_rank, k).normal_(0, phm_init_range))
            self.W_left2 = nn.Parameter(torch.empty(n, k, phm_rank).normal_(0, phm_init_range))
            self.W_right2 = nn.Parameter(torch.empty(n, phm_rank, k).normal_(0, phm_init_range))
            self.phmb = nn.Parameter(torch.zeros(dim))      # allocated but unused by the reference forward (:100-101)
        elif adapter == "lora":
            self.lora_attn_dim = lora_attn_dim
            self.lora_attn_alpha = lora_attn_alpha
            self.q_proj_adapter1 = nn.Linear(dim, lora_attn_dim, bias=False)
            nn.init.normal_(self.q_proj_adapter1.weight, std=0.02)
            self.q_proj_adapter2 = nn.Linear(lora_attn_dim, dim, bias=False)
            self.q_proj_adapter2.weight.data.zero_()
            self.v_proj_adapter1 = nn.Linear(dim, lora_attn_dim, bias=False)
            nn.init.normal_(self.v_proj_adapter1.weight, std=0.02)
            self.v_proj_adapter2 = nn.Linear(lora_attn_dim, dim, bias=False)
            self.v_proj_adapter2.weight.data.zero_()
        elif adapter not in ("none", "houlsby", "compacter"):
            raise ValueError(adapter)
        self.fp32 = False

    def forward(self, x, prev=None):
        """Stand-alone forward of the adapter-free attention (FL/lib/models/cls_intrinsic_dimension.py:466-469 wraps exactly
        this); the q/v-adapted variants run inside the whole-tower engine only."""
        if self.adapter_kind in ("kron", "lora"):
            raise RuntimeError("Attention with a q/v adapter has no stand-alone forward in peft_b200: run the enclosing "
                               "VisionTransformer (CUDA engine)")
        return eager.attention(x, self.qkv.weight, self.qkv.bias, self.proj.weight, self.proj.bias, self.num_heads, self.scale,
                               self.fp32)


class Adapter(nn.Module):
    """Houlsby bottleneck adapter — FS/adapter.py:308-400 (= FS/adapter_model.py:204-295): LayerNorm -> Linear(D, m) -> ReLU
    -> Linear(m, D) + residual.  As in the reference, the LayerNorm is registered both as `adapter_norm_before` and as
    `adapter_down.0` (state_dict carries both keys; named_parameters de-duplicates)."""

    def __init__(self, input_size, down_sample=64, non_linearity="relu", init_bert_weights=True):
        super().__init__()
        if non_linearity != "relu":
            raise NotImplementedError("the Houlsby adapter kernels implement the reference's ReLU bottleneck")
        self.input_size, self.down_sample = input_size, down_sample
        self.adapter_norm_before = nn.LayerNorm(input_size)
        self.adapter_down = nn.Sequential(self.adapter_norm_before, nn.Linear(input_size, down_sample), nn.ReLU())
        self.adapter_up = nn.Linear(down_sample, input_size)
        if init_bert_weights:
            for lin in (self.adapter_down[1], self.adapter_up):
                nn.init.normal_(lin.weight, std=0.02)
                nn.init.zeros_(lin.bias)
        self.fp32 = False

    def forward(self, x, residual_input=None):
        """(output, down, None): output = adapter_up(adapter_down(x)) + residual_input (default x, as in the copy of
        FL/lib/models/cls_intrinsic_dimension.py:340-359); `down` = the post-ReLU bottleneck activations (compute dtype, not
        differentiable).  The reference's third value `up` is read by no caller (`[0]` everywhere) and is not materialised."""
        ln, down = self.adapter_norm_before, self.adapter_down[1]
        out, z = eager.adapter(x, x if residual_input is None else residual_input, ln.weight, ln.bias, down.weight, down.bias,
                               self.adapter_up.weight, self.adapter_up.bias, ln.eps, 1.0, self.fp32)
        return out, z, None


class PHMLinear(nn.Module):
    """FS/compacter_model.py:213-355 — y = x (sum_i phm_rule[i] kron W[i]) + b; parameters only (phm_rule is initialised
    N(0, 0.01^2): the reference leaves it uninitialised)."""

    def __init__(self, in_features, out_features, phm_dim):
        super().__init__()
        assert in_features % phm_dim == 0 and out_features % phm_dim == 0
        self.in_features, self.out_features, self.phm_dim = in_features, out_features, phm_dim
        self.phm_rule = nn.Parameter(torch.empty(phm_dim, phm_dim, phm_dim).normal_(0, 0.01))
        self.W = nn.Parameter(torch.empty(phm_dim, in_features // phm_dim, out_features // phm_dim))
        nn.init.xavier_uniform_(self.W, gain=2 ** 0.5)
        self.b = nn.Parameter(torch.zeros(out_features))

    def dense(self):
        """H [in, out] = sum_i phm_rule[i] (x) W[i]  (kronecker_product_einsum_batched, FS/compacter_model.py:190-208)."""
        n = self.phm_dim
        return torch.einsum("iac,ikp->akcp", self.phm_rule, self.W).reshape(self.in_features, self.out_features)

    forward = _no_eager


class HyperComplexAdapter(nn.Module):
    """Compacter — FS/compacter_model.py:358-425: LayerNorm -> PHMLinear(D, m, n=32) -> gelu_new -> PHMLinear(m, D, n=4)."""

    def __init__(self, input_size, down_sample=64, n_down=32, n_up=4):
        super().__init__()
        self.input_size, self.down_sample = input_size, down_sample
        self.adapter_norm_before = nn.LayerNorm(input_size)
        self.down_PHM = PHMLinear(input_size, down_sample, n_down)
        self.up_sampler = PHMLinear(down_sample, input_size, n_up)

    forward = _no_eager


class Block(nn.Module):
    """FS/kadaptation.py:241-283 (pre-norm only)."""

    def __init__(self, dim, num_heads, mlp_ratio=4., qkv_bias=False, qk_scale=None, drop=0., attn_drop=0., drop_path=0.,
                 act_layer=nn.GELU, norm_layer=nn.LayerNorm, pre_norm=True, res_score=False, bottleneck=64,
                 compacter_n_down=32, compacter_n_up=4, **adapter_kw):
        super().__init__()
        if not pre_norm:
            raise NotImplementedError("post-norm blocks are not supported")
        # stochastic depth (timm DropPath, FS/kadaptation.py:263,276-277): the per-sample keep mask / keep_prob is folded into
        # the two residual epilogues by the engine; the rate lives here like in the reference block
        self.drop_path_rate = float(drop_path)
        self.norm1 = norm_layer(dim)
        self.attn = Attention(dim, num_heads=num_heads, qkv_bias=qkv_bias, qk_scale=qk_scale, attn_drop=attn_drop,
                              proj_drop=drop, res_score=res_score, **adapter_kw)
        self.norm2 = norm_layer(dim)
        self.mlp = Mlp(in_features=dim, hidden_features=int(dim * mlp_ratio), act_layer=act_layer, drop=drop)
        kind = adapter_kw.get("adapter", "kron")
        if kind == "houlsby":
            self.adapter = Adapter(dim, down_sample=bottleneck)
        elif kind == "compacter":
            self.compacter = HyperComplexAdapter(dim, down_sample=bottleneck, n_down=compacter_n_down, n_up=compacter_n_up)
        self.fp32 = False

    def _norm(self, ln, x):
        return eager.layer_norm(x, ln.weight, ln.bias, ln.eps)

    def forward(self, x, prev=None):
        """Stand-alone block forward composed of the sub-module nodes (FS/kadaptation.py:273-277; with the Houlsby adapter
        FS/adapter.py:452-456).  For throughput run the enclosing VisionTransformer: the engine fuses the same arithmetic."""
        if self.drop_path_rate > 0. and self.training:
            raise RuntimeError("stochastic depth runs inside the whole-tower engine only")
        if hasattr(self, "compacter"):
            raise RuntimeError("the Compacter block has no stand-alone forward in peft_b200 (whole-tower engine only)")
        for m in (self.attn, self.mlp):
            m.fp32 = self.fp32
        x = x + self.attn(self._norm(self.norm1, x))
        h = self.mlp(self._norm(self.norm2, x))
        if hasattr(self, "adapter"):
            self.adapter.fp32 = self.fp32
            h = self.adapter(h, h)[0]
        return x + h


class PatchEmbed(nn.Module):
    """FS/kadaptation.py:286-308."""

    def __init__(self, img_size=224, patch_size=16, in_chans=3, embed_dim=768):
        super().__init__()
        self.img_size = (img_size, img_size)
        self.patch_size = (patch_size, patch_size)
        self.num_patches = (img_size // patch_size) ** 2
        self.proj = nn.Conv2d(in_chans, embed_dim, kernel_size=patch_size, stride=patch_size)
        self.fp32 = False

    def forward(self, x):
        """[B, 3, S, S] -> [B, num_patches, D]: im2col + tensor-core product (no cuDNN convolution)."""
        return eager.patch_embed(x, self.proj.weight, self.proj.bias, self.fp32)


def lowrank_factors_kron(rule, w_left, w_right):
    """[dep,n,n,n], [dep,n,K,r], [dep,n,r,P] -> Dn [dep, n*r*n, n*K], UpT [dep, n*r*n, n*P]  (reshape trick).

    H[(a,k),(c,p)] = sum_{i,r} A_i[a,c] u_i[k,r] v_i[r,p] = (Dn^T UpT)[(a,k),(c,p)]  (FS/kadaptation.py:81-115).
    """
    dep, n, K, r = w_left.shape
    P = w_right.shape[3]
    Dn = torch.einsum("liac,likr->lircak", rule, w_left).reshape(dep, n * r * n, n * K)
    eye = torch.eye(n, device=rule.device, dtype=rule.dtype)
    UpT = torch.einsum("dc,lirp->lirdcp", eye, w_right).reshape(dep, n * r * n, n * P)
    return Dn, UpT


class _StackParams(torch.autograd.Function):
    """Zero-copy torch.stack of parameters that already lie stacked in memory (see stack_params)."""

    @staticmethod
    def forward(ctx, pview, gview, *params):
        ctx.gview = gview
        ctx.n = len(params)
        return pview.detach()

    @staticmethod
    def backward(ctx, g):
        ctx.gview.add_(g)                      # one kernel for all layers; the parameters' own AccumulateGrad nodes get nothing
        return (None, None) + (None,) * ctx.n


STACK_FAST_HITS = 0      # how often stack_params took the zero-copy path (tests)


def stack_params(params):
    """torch.stack(params) — as a VIEW with a fused gradient accumulation when the parameters (and their .grad tensors) are
    consecutive slices of one buffer each, which is how dist.FusedSGDBucket / GradBucket lay them out under dist.kind_major.
    Forward then launches nothing and backward adds the stacked gradient to the bucket with one kernel; the result is identical
    to torch.stack + per-parameter accumulation.  Falls back to torch.stack whenever any condition does not hold."""
    global STACK_FAST_HITS
    p0 = params[0]
    n, k = len(params), p0.numel()
    ok = n > 1 and p0.is_cuda and p0.dtype == torch.float32 and torch.is_grad_enabled()
    if ok:
        d0, st = p0.data_ptr(), p0.untyped_storage().data_ptr()
        for i, q in enumerate(params):
            g = q.grad
            if (not q.requires_grad or g is None or q.shape != p0.shape or not q.is_contiguous() or not g.is_contiguous()
                    or g.dtype != torch.float32 or q.data_ptr() != d0 + 4 * k * i or q.untyped_storage().data_ptr() != st
                    or g.data_ptr() != params[0].grad.data_ptr() + 4 * k * i
                    or g.untyped_storage().data_ptr() != params[0].grad.untyped_storage().data_ptr()):
                ok = False
                break
    if not ok:
        return torch.stack(list(params))
    shape, strides = (n,) + tuple(p0.shape), (k,) + tuple(p0.stride())
    pview = torch.as_strided(p0.detach(), shape, strides)
    gview = torch.as_strided(p0.grad, shape, strides)
    STACK_FAST_HITS += 1
    return _StackParams.apply(pview, gview, *params)


class _KronFactors(torch.autograd.Function):
    """(phm_rule [dep,n,n,n], W_left1, W_right1, W_left2, W_right2 stacked over the layers) -> (Dn, UpT) through
    peft_kron_factors: one launch per direction instead of ~40 tiny ATen launches (csrc/factors_sm100.cuh)."""

    @staticmethod
    def forward(ctx, rule, wl1, wr1, wl2, wr2):
        from . import _lib
        dep, n, K, r = wl1.shape
        P = wr1.shape[3]
        rule, wl1, wr1, wl2, wr2 = (t.detach().contiguous() for t in (rule, wl1, wr1, wl2, wr2))
        Dn = torch.empty(dep, 2 * n * r * n, n * K, device=wl1.device, dtype=torch.float32)
        UpT = torch.empty(dep, 2 * n * r * n, 3 * n * P, device=wl1.device, dtype=torch.float32)
        _lib.check(_lib.load().peft_kron_factors(rule.data_ptr(), wl1.data_ptr(), wr1.data_ptr(), wl2.data_ptr(), wr2.data_ptr(),
                                                 Dn.data_ptr(), UpT.data_ptr(), dep, n, K, r, P, 0,
                                                 torch.cuda.current_stream().cuda_stream), "peft_kron_factors")
        ctx.save_for_backward(rule, wl1, wl2)
        ctx.dims = (dep, n, K, r, P)
        return Dn, UpT

    @staticmethod
    def backward(ctx, dDn, dUpT):
        from . import _lib
        rule, wl1, wl2 = ctx.saved_tensors
        dep, n, K, r, P = ctx.dims
        dDn, dUpT = dDn.contiguous(), dUpT.contiguous()
        gl1, gl2 = (torch.empty(dep, n, K, r, device=rule.device, dtype=torch.float32) for _ in range(2))
        gr1, gr2 = (torch.empty(dep, n, r, P, device=rule.device, dtype=torch.float32) for _ in range(2))
        _lib.check(_lib.load().peft_kron_factors(rule.data_ptr(), gl1.data_ptr(), gr1.data_ptr(), gl2.data_ptr(), gr2.data_ptr(),
                                                 dDn.data_ptr(), dUpT.data_ptr(), dep, n, K, r, P, 1,
                                                 torch.cuda.current_stream().cuda_stream), "peft_kron_factors")
        grule = None
        if ctx.needs_input_grad[0]:
            grule = torch.empty_like(rule)
            _lib.check(_lib.load().peft_kron_factors_drule(dDn.data_ptr(), wl1.data_ptr(), wl2.data_ptr(), grule.data_ptr(), dep, n, K, r,
                                                           torch.cuda.current_stream().cuda_stream), "peft_kron_factors_drule")
        return grule, gl1, gr1, gl2, gr2


_RULE_CACHE = {}     # id(first rule) -> (versions, stacked frozen phm_rule): the rule is frozen, stack it once


def _stack_frozen(ts):
    if torch.cuda.is_current_stream_capturing():      # a tensor allocated during graph capture must not outlive / precede its graph
        return torch.stack([t.detach() for t in ts])
    key = tuple((t.data_ptr(), t._version) for t in ts)
    hit = _RULE_CACHE.get(id(ts[0]))
    if hit is not None and hit[0] == key:
        return hit[1]
    st = torch.stack([t.detach() for t in ts])
    _RULE_CACHE[id(ts[0])] = (key, st)
    return st


def lowrank_operands(attns, kind):
    """Dn [depth, r', D], UpT [depth, r', 3D] from the attention modules' reference-named adapter parameters
    (`phm_rule / W_left{1,2} / W_right{1,2}` or `{q,v}_proj_adapter{1,2}.weight`); shared by both dialects."""
    if kind == "kron" and kron_is_dense(attns[0]):
        return dense_kron_operands(attns)
    if kind == "kron" and attns[0].phm_rule.is_cuda and attns[0].phm_rule.dtype == torch.float32:
        # both operands of every layer from one kernel per direction (csrc/factors_sm100.cuh)
        rules = [a.phm_rule for a in attns]
        return _KronFactors.apply(stack_params(rules) if any(t.requires_grad for t in rules) else _stack_frozen(rules),
                                  stack_params([a.W_left1 for a in attns]), stack_params([a.W_right1 for a in attns]),
                                  stack_params([a.W_left2 for a in attns]), stack_params([a.W_right2 for a in attns]))
    if kind == "kron":
        rule = torch.stack([a.phm_rule for a in attns])
        dq, uq = lowrank_factors_kron(rule, stack_params([a.W_left1 for a in attns]), stack_params([a.W_right1 for a in attns]))
        dv, uv = lowrank_factors_kron(rule, stack_params([a.W_left2 for a in attns]), stack_params([a.W_right2 for a in attns]))
    elif kind == "lora":
        s = attns[0].lora_attn_alpha / attns[0].lora_attn_dim
        dq = stack_params([a.q_proj_adapter1.weight for a in attns])
        dv = stack_params([a.v_proj_adapter1.weight for a in attns])
        uq = stack_params([a.q_proj_adapter2.weight for a in attns]).transpose(1, 2) * s
        uv = stack_params([a.v_proj_adapter2.weight for a in attns]).transpose(1, 2) * s
    else:
        return None, None
    z = torch.zeros_like(uq)
    Dn = torch.cat([dq, dv], dim=1)
    UpT = torch.cat([torch.cat([uq, z, z], dim=2), torch.cat([z, z, uv], dim=2)], dim=1)
    return Dn, UpT


DENSE_DELTA_MIN_RANK = 512     # reshape-trick rank 2 n r n above which the dense ("fold") form of the update is used


def kron_is_dense(attn):
    n = attn.phm_rule.shape[0]
    return 2 * n * attn.phm_rank * n > DENSE_DELTA_MIN_RANK


class _PhmFoldFn(torch.autograd.Function):
    """(phm_rule [n,n,n] frozen, s [2,n]) -> H [2, n, n] through peft_phm_fold (K = P = 1 case of the dense update)."""

    @staticmethod
    def forward(ctx, rule, s):
        from . import ops
        if rule.requires_grad:
            raise NotImplementedError("phm_rule is frozen in the large-n fold path (the reference's 'W' filter freezes it too)")
        n = rule.shape[0]
        H = torch.empty(2, n, n, device=rule.device, dtype=torch.float32)
        ops.phm_fold(rule.detach().contiguous(), s.detach().float().contiguous(), H)
        ctx.rule = rule
        return H

    @staticmethod
    def backward(ctx, dH):
        from . import ops
        n = ctx.rule.shape[0]
        ds = torch.empty(2, n, device=dH.device, dtype=torch.float32)
        ops.phm_fold(ctx.rule.detach().contiguous(), dH.float().contiguous(), ds, backward=True)
        return None, ds


def dense_kron_operands(attns):
    """Large-n KAdaptation (the committed phm_dim = 768): H = sum_i A_i (x) (u_i v_i^T) is formed as a dense
    [D, D] matrix per projection by a differentiable contraction over i (it reads phm_rule once: n^3 floats, HBM-bound),
    and handed to the engine as UpT = [H_q | 0 | H_v] with Dn = I.  dH comes back as ONE tensor-core product per layer and
    autograd maps it to W_left / W_right (and to phm_rule if it requires grad; the reference freezes it)."""
    outs = []
    for a in attns:
        n, K = a.W_left1.shape[0], a.W_left1.shape[1]
        D = n * K
        if K == 1 and a.phm_rule.is_cuda and not a.phm_rule.requires_grad:
            # the committed hyper-parameters (n = D): one streaming pass over phm_rule for q and v together
            s = torch.stack([torch.bmm(a.W_left1, a.W_right1).reshape(n), torch.bmm(a.W_left2, a.W_right2).reshape(n)])
            H = _PhmFoldFn.apply(a.phm_rule, s)
            hs = [H[0], H[1]]
        else:
            hs = []
            for wl, wr in ((a.W_left1, a.W_right1), (a.W_left2, a.W_right2)):
                B = torch.bmm(wl, wr)                                                   # [n, K, P]
                hs.append(torch.einsum("iac,ikp->akcp", a.phm_rule, B).reshape(D, D))   # FS/kadaptation.py:104-115
        outs.append(torch.cat([hs[0], torch.zeros_like(hs[0]), hs[1]], dim=1))      # [D, 3D]
    return None, torch.stack(outs)


def lowrank_rank(attn, kind):
    if kind == "kron":
        n = attn.phm_rule.shape[0]
        if kron_is_dense(attn):
            return attn.W_left1.shape[0] * attn.W_left1.shape[1]                    # dense update: rank == width
        return 2 * n * attn.phm_rank * n
    if kind == "lora":
        return 2 * attn.lora_attn_dim
    return 0


def bottleneck_operands(blocks, kind):
    def st(xs):      # parameters: zero-copy view when they lie stacked in the bucket (stack_params); computed tensors: torch.stack
        return stack_params(xs) if all(isinstance(x, torch.nn.Parameter) for x in xs) else torch.stack(xs)
    if kind == "houlsby":
        ad = [b.adapter for b in blocks]
        return (st([a.adapter_down[1].weight for a in ad]), st([a.adapter_down[1].bias for a in ad]),
                st([a.adapter_up.weight for a in ad]), st([a.adapter_up.bias for a in ad]),
                st([a.adapter_norm_before.weight for a in ad]), st([a.adapter_norm_before.bias for a in ad]))
    if kind == "compacter":
        ad = [b.compacter for b in blocks]
        return (st([a.down_PHM.dense().t() for a in ad]), st([a.down_PHM.b for a in ad]),
                st([a.up_sampler.dense().t() for a in ad]), st([a.up_sampler.b for a in ad]),
                st([a.adapter_norm_before.weight for a in ad]), st([a.adapter_norm_before.bias for a in ad]))
    return None


def bottleneck_cfg(block0, kind):
    if kind == "houlsby":
        return dict(bottleneck=block0.adapter.down_sample, bn_act="relu")
    if kind == "compacter":
        return dict(bottleneck=block0.compacter.down_sample, bn_act="gelu_tanh")
    return {}


def add_qv_adapter_params(mod, dim, adapter, phm_dim=4, phm_rank=1, phm_init_range=0.01, lora_attn_dim=4, lora_attn_alpha=128):
    """Registers the reference-named q/v adapter parameters on an attention module (both dialects use the same names:
    FS/kadaptation.py:143-187, FS/model.py:719-736)."""
    mod.adapter_kind = adapter
    if adapter == "kron":
        assert dim % phm_dim == 0
        mod.phm_rank = phm_rank
        mod.phm_init_range = phm_init_range
        n, k = phm_dim, dim // phm_dim
        mod.phm_rule = nn.Parameter(torch.empty(n, n, n).normal_(0, phm_init_range))
        mod.W_left1 = nn.Parameter(torch.empty(n, k, phm_rank).normal_(0, phm_init_range))
        mod.W_right1 = nn.Parameter(torch.empty(n, phm_rank, k).normal_(0, phm_init_range))
        mod.W_left2 = nn.Parameter(torch.empty(n, k, phm_rank).normal_(0, phm_init_range))
        mod.W_right2 = nn.Parameter(torch.empty(n, phm_rank, k).normal_(0, phm_init_range))
        mod.phmb = nn.Parameter(torch.zeros(dim))      # allocated but unused by the reference forward (:100-101)
    elif adapter == "lora":
        mod.lora_attn_dim = lora_attn_dim
        mod.lora_attn_alpha = lora_attn_alpha
        mod.q_proj_adapter1 = nn.Linear(dim, lora_attn_dim, bias=False)
        nn.init.normal_(mod.q_proj_adapter1.weight, std=0.02)
        mod.q_proj_adapter2 = nn.Linear(lora_attn_dim, dim, bias=False)
        mod.q_proj_adapter2.weight.data.zero_()
        mod.v_proj_adapter1 = nn.Linear(dim, lora_attn_dim, bias=False)
        nn.init.normal_(mod.v_proj_adapter1.weight, std=0.02)
        mod.v_proj_adapter2 = nn.Linear(lora_attn_dim, dim, bias=False)
        mod.v_proj_adapter2.weight.data.zero_()
    elif adapter not in ("none", "houlsby", "compacter"):
        raise ValueError(adapter)


def tower_tensors(root, keys):
    """[(key, tensor)] read through ATTRIBUTE ACCESS on every call: under FastfoodWrap the parameters have been deleted
    and are re-set as plain (non-leaf) tensors each forward (FL/intrinsic/fastfood.py:52-57,61-77)."""
    out = []
    for key in keys:
        obj = root
        *path, leaf = key.split(".")
        for p in path:
            obj = getattr(obj, p)
        t = getattr(obj, leaf)
        if t is not None:
            out.append((key, t))
    return out


def sample_drop_path(model, B, L, device):
    """fp32 [depth, 2, B*L] row factors keep/keep_prob of the attention and MLP residual branches, or None (eval mode /
    all rates 0).  One Bernoulli(keep_prob) draw per sample and branch, exactly timm's `drop_path`:
    floor(keep_prob + U[0,1)) / keep_prob.  `model.drop_path_keep` ([depth, 2, B] of 0/1), when set, replaces the random
    draw (parity tests against the oracle with a fixed mask)."""
    rates = [getattr(b, "drop_path_rate", 0.0) for b in model.blocks]
    if not model.training or not any(r > 0 for r in rates):
        return None
    keep_prob = 1.0 - torch.tensor(rates, device=device, dtype=torch.float32).view(-1, 1, 1)
    fixed = getattr(model, "drop_path_keep", None)
    if fixed is not None:
        keep = fixed.to(device=device, dtype=torch.float32)
        if tuple(keep.shape) != (len(rates), 2, B):
            raise ValueError(f"drop_path_keep must be [depth, 2, B] = {(len(rates), 2, B)}")
    else:
        keep = torch.floor(keep_prob + torch.rand(len(rates), 2, B, device=device))
    return (keep / keep_prob).repeat_interleave(L, dim=2).contiguous()


def adapterdrop_reference_schedule(depth=12, last=11):
    """The block schedule of the COMMITTED FS/adapterdrop.py:582-588 (like Q1-Q8): its loop runs every block
    TWICE — `x = blk(x, add_adapter=True)` and then `blk(x, True)` for block 11 / `blk(x, False)` for the others — so the
    tower has 2 x depth block applications with shared weights, the first of each pair with the adapter.  Returns
    [(block index, adapter applied)] for `VisionTransformer(block_schedule=...)`; `adapter_layers=(11,)` gives the intended
    AdapterDrop (each block once, adapter in block 11 only)."""
    out = []
    for i in range(depth):
        out += [(i, True), (i, i == last)]
    return out


def expand_schedule(tensors, schedule):
    """[(reference key, tensor)] -> the same list with every `blocks.{b}.` entry repeated for each application j of block b,
    renamed `blocks.{j}.` (the engine sees a tower of len(schedule) layers; shared tensors appear several times and autograd
    sums their gradients)."""
    out = []
    for k, t in tensors:
        if k.startswith("blocks."):
            b, rest = k[len("blocks."):].split(".", 1)
            out += [(f"blocks.{j}.{rest}", t) for j, (bb, _) in enumerate(schedule) if bb == int(b)]
        else:
            out.append((k, t))
    return out


def run_tower(model, x, dialect, geom, extra_cfg, lowrank, bottleneck, drop=None):
    """Shared forward of both dialects' VisionTransformer: picks the PEFT engine (frozen tower, adapters trainable) or the
    full engine (any tower tensor requires grad: full fine-tuning, bias/LayerNorm tuning, Fastfood)."""
    if not x.is_cuda:
        raise RuntimeError("peft_b200 VisionTransformer runs on CUDA only (no CPU fallback)")
    tensors = tower_tensors(model, model._tower_keys)
    if getattr(model, "block_schedule", None):
        tensors = expand_schedule(tensors, model.block_schedule)
    grad = torch.is_grad_enabled()
    trainable = [(k, t) for k, t in tensors if t.requires_grad]
    with_adapter = model.adapter_kind != "none"
    full = grad and bool(trainable) and not with_adapter
    slot = "_engine_full" if full else "_engine"
    eng = getattr(model, slot, None)
    precision = getattr(model, "precision", "bf16")      # "fp32": verification mode (plain fp32 kernels, 1e-5 gate)
    if eng is None or eng.dev != x.device or eng.cfg.precision != precision:
        eng = VitEngine(EngineCfg(dialect=dialect, train_all=full, precision=precision, **extra_cfg, **geom), x.device)
        setattr(model, slot, eng)
        model._frozen_sig = None
    if full:
        if drop is not None:
            raise NotImplementedError("stochastic depth is not supported in full (train-all) mode")
        return backbone_full(eng, x, tensors)
    # adapter + bias tuning / LayerNorm tuning (FL/tools/finetune_bias.py, finetune_layernorm.py on an adapter model): trainable
    # tower VECTORS (biases, LayerNorm gains / biases) are re-read every forward and get their gradients inside the PEFT
    # backward (column sums, LayerNorm parameter gradients).  Weight matrices next to an adapter would need dW products.
    small = [(k, t) for k, t in trainable if k in eng.small_keys()] if with_adapter else []
    if grad and with_adapter and len(small) != len(trainable):
        other = [k for k, _ in trainable if k not in eng.small_keys()]
        raise NotImplementedError(f"training the tower tensors {other[:4]}... together with an adapter is not supported: weight "
                                  "matrices need dW products — freeze them (bias / LayerNorm tuning works) or build the model "
                                  "with adapter='none' (full fine-tuning)")
    small_names = {k for k, _ in small}
    cacheable = all(isinstance(t, nn.Parameter) for _, t in tensors)
    sig = (slot,) + tuple((k, t.data_ptr(), t._version) for k, t in tensors if k not in small_names) if cacheable else None
    if sig is None or sig != model._frozen_sig:
        eng.load_frozen({k: t.detach() for k, t in tensors})
        model._frozen_sig = sig
    if small and not grad:                        # eval / no_grad forward: still re-read the trained vectors, no autograd inputs
        eng.load_small({k: t for k, t in small})
        small = None
    Dn, UpT = lowrank()
    return backbone(eng, x, Dn, UpT, bottleneck(), drop, small)


class VisionTransformer(nn.Module):
    """FS/kadaptation.py:352-551.  forward(x) returns the final-LayerNorm cls feature (the head is applied by the
    trainer's Classifier, as in the reference: `# x = self.head(x)`)."""

    def __init__(self, img_size=224, patch_size=16, in_chans=3, num_classes=1000, embed_dim=768, depth=12, num_heads=12,
                 mlp_ratio=4., qkv_bias=False, qk_scale=None, drop_rate=0., attn_drop_rate=0., drop_path_rate=0.,
                 hybrid_backbone=None, norm_layer=None, use_cls_tocken=True, norm_embed=False, pre_norm=True,
                 res_score=False, init='trunc_norm', adapter="kron", adapter_layers=None, block_schedule=None, **adapter_kw):
        super().__init__()
        # AdapterDrop (FS/adapterdrop.py:585-588,689: every block owns an adapter, only block 11 calls and trains it):
        # indices of the blocks whose bottleneck adapter is applied; None = all blocks
        self.adapter_layers = None if adapter_layers is None else tuple(sorted(int(i) for i in adapter_layers))
        # [(block index, adapter applied)] per block APPLICATION; None = every block once (see adapterdrop_reference_schedule)
        self.block_schedule = None if block_schedule is None else [(int(b), bool(u)) for b, u in block_schedule]
        if self.block_schedule is not None and self.adapter_layers is not None:
            raise ValueError("give either adapter_layers or block_schedule")
        if hybrid_backbone is not None or norm_embed or not use_cls_tocken or in_chans != 3:
            raise NotImplementedError("hybrid backbone / norm_embed / no-cls-token / in_chans != 3 are not supported")
        if drop_rate != 0.:
            raise NotImplementedError("dropout > 0 is not supported (the reference configs use 0)")
        norm_layer = norm_layer or partial(nn.LayerNorm, eps=1e-6)
        self.num_classes = num_classes
        self.num_features = self.embed_dim = embed_dim
        self.adapter_kind = adapter
        self.patch_embed = PatchEmbed(img_size=img_size, patch_size=patch_size, in_chans=in_chans, embed_dim=embed_dim)
        num_patches = self.patch_embed.num_patches
        self.cls_token = nn.Parameter(torch.zeros(1, 1, embed_dim))
        self.pos_embed = nn.Parameter(torch.zeros(1, num_patches + 1, embed_dim))
        dpr = [v.item() for v in torch.linspace(0, drop_path_rate, depth)]      # stochastic depth decay rule (FS/kadaptation.py:408)
        self.blocks = nn.ModuleList([
            Block(dim=embed_dim, num_heads=num_heads, mlp_ratio=mlp_ratio, qkv_bias=qkv_bias, qk_scale=qk_scale,
                  drop=drop_rate, attn_drop=attn_drop_rate, drop_path=dpr[i], norm_layer=norm_layer, pre_norm=pre_norm,
                  res_score=res_score, adapter=adapter, **adapter_kw)
            for i in range(depth)])
        self.norm = norm_layer(embed_dim)
        self.head = nn.Linear(embed_dim, num_classes) if num_classes > 0 else nn.Identity()
        nn.init.trunc_normal_(self.cls_token, std=.02)
        nn.init.trunc_normal_(self.pos_embed, std=.02)
        self.apply(self._init_weights_xavier if init == 'xavier' else self._init_weights_trunc_normal)
        eps = self.norm.eps
        if abs(eps - 1e-6) > 1e-12:
            raise NotImplementedError("timm dialect uses LayerNorm eps 1e-6")
        self._cfg = dict(img=img_size, patch=patch_size, width=embed_dim, heads=num_heads, mlp=int(embed_dim * mlp_ratio),
                         depth=depth if self.block_schedule is None else len(self.block_schedule))
        self._engine = None
        self._engine_full = None
        self._frozen_sig = None
        self._tower_keys = [k for k, _ in self._frozen_items()]
        self._qkv_bias = qkv_bias

    @staticmethod
    def _init_weights_trunc_normal(m):
        if isinstance(m, nn.Linear):
            nn.init.trunc_normal_(m.weight, std=0.02)
            if m.bias is not None:
                nn.init.constant_(m.bias, 0)
        elif isinstance(m, nn.LayerNorm):
            nn.init.constant_(m.bias, 0)
            nn.init.constant_(m.weight, 1.0)

    @staticmethod
    def _init_weights_xavier(m):
        if isinstance(m, nn.Linear):
            nn.init.xavier_uniform_(m.weight)
            if m.bias is not None:
                nn.init.constant_(m.bias, 0)
        elif isinstance(m, nn.LayerNorm):
            nn.init.constant_(m.bias, 0)
            nn.init.constant_(m.weight, 1.0)

    @torch.jit.ignore
    def no_weight_decay(self):
        return {'pos_embed', 'cls_token'}

    # ---------------------------------------------------------------- adapter -> low-rank operands
    _ADAPTER_KEYS = ("phm_rule", "W_left", "W_right", "phmb", "_proj_adapter", ".adapter.", ".compacter.")

    def _rank(self):
        return lowrank_rank(self.blocks[0].attn, self.adapter_kind)

    def _applied_blocks(self):
        return list(self.blocks) if self.block_schedule is None else [self.blocks[b] for b, _ in self.block_schedule]

    def lowrank_operands(self):
        """Dn [depth, r', D], UpT [depth, r', 3D] built differentiably from the reference-named parameters."""
        return lowrank_operands([b.attn for b in self._applied_blocks()], self.adapter_kind)

    def bottleneck_operands(self):
        """(Wd [dep,m,D], bd [dep,m], Wu [dep,D,m], bu [dep,D], ln_g [dep,D], ln_b [dep,D]) built differentiably from the
        reference-named Houlsby / Compacter parameters; None when the model has no bottleneck adapter."""
        return bottleneck_operands(self._applied_blocks(), self.adapter_kind)

    def _bottleneck_cfg(self):
        return bottleneck_cfg(self.blocks[0], self.adapter_kind)

    # ---------------------------------------------------------------- engine plumbing
    def _frozen_items(self):
        return [(k, v) for k, v in self.state_dict(keep_vars=True).items()
                if not any(t in k for t in self._ADAPTER_KEYS) and not k.startswith("head.")]

    def forward_features(self, x):
        extra = dict(rank=self._rank(), **self._bottleneck_cfg())
        if self.adapter_layers is not None and extra.get("bottleneck"):
            extra["bottleneck_layers"] = self.adapter_layers
        if self.block_schedule is not None and extra.get("bottleneck"):
            extra["bottleneck_layers"] = tuple(j for j, (_, use) in enumerate(self.block_schedule) if use)
        if self.adapter_kind == "kron" and kron_is_dense(self.blocks[0].attn):
            extra["dense_delta"] = True
        drop = sample_drop_path(self, x.shape[0], self.patch_embed.num_patches + 1, x.device)
        return run_tower(self, x, "timm", self._cfg, extra, self.lowrank_operands, self.bottleneck_operands, drop)

    def forward(self, x):
        return self.forward_features(x)
