"""Stand-alone sub-module forwards (eager.py) and the per-component intrinsic-dimension model (intrinsic_vit.py, SURVEY §8a-R12 /
FL/lib/models/cls_intrinsic_dimension.py) against the oracle restatements run in float64 on the CPU.

Tolerances (relative Frobenius error per tensor): fp32 kernels 2e-5 (north_star fp32 gate 1e-5 on the features; weight gradients
sum over tokens in a different order), bf16 tensor-core path 2e-2 (operands rounded to bf16, fp32 accumulation).
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import peft_oracle as O

pytestmark = pytest.mark.gpu
TOL = {True: 2e-5, False: 2e-2}


def rel(a, b):
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


def ref_sd(module, pfx=""):
    return {pfx + k: v.detach().double().cpu().clone().requires_grad_(True) for k, v in module.state_dict().items()}


def check_grads(module, sd, pfx, tol, skip=()):
    for k, p in module.named_parameters():
        if any(s in k for s in skip):
            continue
        assert p.grad is not None, k
        assert rel(p.grad, sd[pfx + k].grad) < tol, (k, rel(p.grad, sd[pfx + k].grad))


def gated(fn, downs, fp32):
    """Oracle evaluated on the linear region the kernels selected (tests/houlsby_gate.py): the bf16 path may put a ReLU
    pre-activation that lies within rounding of 0 on the other side; the gate mask is taken from the module's own `down`
    output and at most 1 % of the units may differ from the oracle's own decision.  fp32: the oracle's own gates."""
    if fp32:
        return fn()
    masks = {k: (d > 0).cpu() for k, d in downs.items()}
    O.GATE_RECORD = {}
    try:
        with torch.no_grad():
            fn()
        for k, m in masks.items():
            own = O.GATE_RECORD[k][0] > 0
            assert float((own != m).double().mean()) < 0.01
    finally:
        O.GATE_RECORD = None
    O.GATE_OVERRIDE = masks
    try:
        return fn()
    finally:
        O.GATE_OVERRIDE = {}


def rand(*shape, seed=0, scale=1.0):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed)) * scale


@pytest.mark.parametrize("fp32", [True, False])
def test_mlp(fp32):
    from peft_b200.timm_vit import Mlp
    torch.manual_seed(1)
    m = Mlp(128, 512).cuda()
    m.fp32 = fp32
    x = rand(2, 10, 128).cuda().requires_grad_(True)
    c = rand(2, 10, 128, seed=3)
    y = m(x)
    (y * c.cuda()).sum().backward()
    sd = ref_sd(m)
    xr = x.detach().double().cpu().requires_grad_(True)
    yr = O.timm_mlp(xr, sd, "")
    (yr * c.double()).sum().backward()
    assert rel(y, yr) < TOL[fp32]
    assert rel(x.grad, xr.grad) < TOL[fp32]
    check_grads(m, sd, "", TOL[fp32])


@pytest.mark.parametrize("fp32,qkv_bias", [(True, True), (False, True), (False, False)])
def test_attention(fp32, qkv_bias):
    from peft_b200.timm_vit import Attention
    torch.manual_seed(2)
    m = Attention(128, num_heads=2, qkv_bias=qkv_bias, adapter="none").cuda()
    m.fp32 = fp32
    x = rand(3, 17, 128).cuda().requires_grad_(True)
    c = rand(3, 17, 128, seed=4)
    y = m(x)
    (y * c.cuda()).sum().backward()
    sd = ref_sd(m)
    xr = x.detach().double().cpu().requires_grad_(True)
    yr = O.timm_attention(xr, sd, "", O.VitCfg(width=128, heads=2, adapter="none"))
    (yr * c.double()).sum().backward()
    assert rel(y, yr) < TOL[fp32]
    assert rel(x.grad, xr.grad) < TOL[fp32]
    check_grads(m, sd, "", TOL[fp32])


def test_attention_with_qv_adapter_has_no_standalone_forward():
    from peft_b200.timm_vit import Attention
    m = Attention(128, num_heads=2, adapter="kron").cuda()
    with pytest.raises(RuntimeError):
        m(torch.zeros(1, 5, 128, device="cuda"))


@pytest.mark.parametrize("fp32", [True, False])
def test_adapter(fp32):
    from peft_b200.timm_vit import Adapter
    torch.manual_seed(3)
    m = Adapter(128, down_sample=64).cuda()
    with torch.no_grad():
        m.adapter_norm_before.weight.add_(rand(128, seed=8).cuda() * 0.1)
        m.adapter_down[1].bias.add_(rand(64, seed=9).cuda() * 0.02)
        m.adapter_up.bias.add_(rand(128, seed=10).cuda() * 0.02)
    m.fp32 = fp32
    x = rand(2, 9, 128).cuda().requires_grad_(True)
    c = rand(2, 9, 128, seed=5)
    out, down, _ = m(x)                                     # residual_input defaults to x
    (out * c.cuda()).sum().backward()
    sd = ref_sd(m)
    xr = x.detach().double().cpu().requires_grad_(True)
    outr = gated(lambda: O.houlsby_adapter(xr, sd, ""), {"": down}, fp32)
    (outr * c.double()).sum().backward()
    assert rel(out, outr) < TOL[fp32]
    assert rel(x.grad, xr.grad) < TOL[fp32]
    check_grads(m, sd, "", TOL[fp32], skip=("adapter_down.0",))   # same LayerNorm registered twice; named_parameters de-duplicates


@pytest.mark.parametrize("fp32,kind", [(True, "houlsby"), (False, "houlsby"), (False, "none")])
def test_block(fp32, kind):
    from peft_b200.timm_vit import Block
    torch.manual_seed(4)
    m = Block(128, 2, qkv_bias=True, norm_layer=lambda d: torch.nn.LayerNorm(d, eps=1e-6), adapter=kind).cuda()
    m.fp32 = fp32
    x = rand(2, 10, 128).cuda().requires_grad_(True)
    c = rand(2, 10, 128, seed=6)
    downs = {}
    if kind == "houlsby":
        m.adapter.register_forward_hook(lambda mod, a, out: downs.__setitem__("blocks.0.adapter.", out[1]))
    y = m(x)
    (y * c.cuda()).sum().backward()
    sd = ref_sd(m, "blocks.0.")
    xr = x.detach().double().cpu().requires_grad_(True)
    yr = gated(lambda: O.timm_block(xr, sd, "blocks.0.", O.VitCfg(width=128, heads=2, adapter=kind)), downs, fp32)
    (yr * c.double()).sum().backward()
    assert rel(y, yr) < TOL[fp32]
    assert rel(x.grad, xr.grad) < TOL[fp32]
    check_grads(m, sd, "blocks.0.", TOL[fp32], skip=("adapter_down.0",))


@pytest.mark.parametrize("fp32,patch", [(True, 16), (False, 16), (False, 14)])
def test_patch_embed(fp32, patch):
    from peft_b200.timm_vit import PatchEmbed
    torch.manual_seed(5)
    S = patch * 3
    m = PatchEmbed(img_size=S, patch_size=patch, embed_dim=128).cuda()
    m.fp32 = fp32
    img = rand(2, 3, S, S, seed=7).cuda()
    c = rand(2, 9, 128, seed=8)
    y = m(img)
    (y * c.cuda()).sum().backward()
    w = m.proj.weight.detach().double().cpu().requires_grad_(True)
    b = m.proj.bias.detach().double().cpu().requires_grad_(True)
    yr = F.conv2d(img.double().cpu(), w, b, stride=patch).flatten(2).transpose(1, 2)       # FS/kadaptation.py:302-307
    (yr * c.double()).sum().backward()
    assert rel(y, yr) < TOL[fp32]
    assert rel(m.proj.weight.grad, w.grad) < TOL[fp32]
    assert rel(m.proj.bias.grad, b.grad) < TOL[fp32]


def wrapped_theta(wrap, V):
    """oracle: name -> theta0 + Fastfood(V) in float64-free fp32 (the reference arithmetic), differentiable w.r.t. V"""
    out = {}
    for name, _, _ in wrap.name_base_localname:
        t0 = wrap.initial_value[name].detach().cpu()
        DD = t0.numel()
        p = [q.cpu() if torch.is_tensor(q) else q for q in wrap.fastfood_params[name]]
        out[name] = t0 + O.fastfood_torched(V, DD, p).view(t0.shape)
    return out


def test_fastfood_wrapped_mlp():
    """FastfoodWrap(Mlp): the wrapped weights are re-set as plain tensors before every forward and the stand-alone forward
    reads them through attribute access; dV flows through dW of both layers (fp32 kernels, 5e-5)."""
    from peft_b200.fastfood import FastfoodWrap
    from peft_b200.timm_vit import Mlp
    torch.manual_seed(6)
    np.random.seed(6)
    inner = Mlp(128, 512)
    inner.fp32 = True
    wrap = FastfoodWrap(inner.cuda(), intrinsic_dimension=40, device=0)
    with torch.no_grad():
        wrap.V.copy_(rand(40, seed=11).cuda() * 0.05)
    x = rand(2, 10, 128).cuda()
    c = rand(2, 10, 128, seed=12)
    for _ in range(2):                                       # second pass: weights are re-projected, not cached
        wrap.V.grad = None
        y = wrap(x)
        (y * c.cuda()).sum().backward()
    Vr = wrap.V.detach().cpu().clone().requires_grad_(True)
    th = wrapped_theta(wrap, Vr)
    yr = O.timm_mlp(x.cpu(), th, "")
    (yr * c).sum().backward()
    assert rel(y, yr) < 5e-5
    assert rel(wrap.V.grad, Vr.grad) < 5e-5, rel(wrap.V.grad, Vr.grad)


@pytest.mark.parametrize("layer_type", ["attention", "mlp", "adapter"])
def test_intrinsic_vit(layer_type):
    """cls_intrinsic_dimension.VisionTransformer: block `layernum` runs its Fastfood-wrapped component; only that V gets a
    gradient; logits and dV against the oracle composition (fp32 kernels)."""
    from peft_b200.intrinsic_vit import VisionTransformer
    torch.manual_seed(7)
    np.random.seed(7)
    D, H, depth, ln = 128, 2, 3, 1
    model = VisionTransformer(img_size=32, patch_size=16, num_classes=10, embed_dim=D, depth=depth, num_heads=H, qkv_bias=True,
                              dintrinsic=50, layerType=layer_type, layernum=ln).cuda()
    model.fp32 = True
    for n, p in model.named_parameters():                   # FL/tools/intrinsic_dimension.py:380-385
        p.requires_grad = "intrinsic" in n
    key = {"attention": "intrinsic_attn", "mlp": "intrinsic_mlp", "adapter": "intrinsic_adapter"}[layer_type]
    wrap = getattr(model.blocks[ln], key)
    with torch.no_grad():
        wrap.V.copy_(rand(50, seed=13).cuda() * 0.05)
    img = rand(4, 3, 32, 32, seed=14).cuda()
    c = rand(4, 10, seed=15)
    logits = model(img)
    (logits * c.cuda()).sum().backward()
    for n, p in model.named_parameters():
        if "intrinsic" in n:
            assert (p.grad is not None) == (n == f"blocks.{ln}.{key}.V"), n

    sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
    Vr = wrap.V.detach().cpu().clone().requires_grad_(True)
    th = wrapped_theta(wrap, Vr)
    cfg = O.VitCfg(img=32, patch=16, width=D, heads=H, depth=depth, adapter="none")
    x = F.conv2d(img.cpu(), sd["patch_embed.proj.weight"], sd["patch_embed.proj.bias"], stride=16).flatten(2).transpose(1, 2)
    x = torch.cat((sd["cls_token"].expand(4, -1, -1), x), dim=1) + sd["pos_embed"]
    for i in range(depth):
        pfx = f"blocks.{i}."
        lt = layer_type if i == ln else None
        n1 = F.layer_norm(x, (D,), sd[pfx + "norm1.weight"], sd[pfx + "norm1.bias"], 1e-6)
        x = x + (O.timm_attention(n1, th, "", cfg) if lt == "attention" else O.timm_attention(n1, sd, pfx + "attn.", cfg))
        n2 = F.layer_norm(x, (D,), sd[pfx + "norm2.weight"], sd[pfx + "norm2.bias"], 1e-6)
        h = O.timm_mlp(n2, th, "") if lt == "mlp" else O.timm_mlp(n2, sd, pfx + "mlp.")
        if lt == "adapter":
            h = O.houlsby_adapter(h, th, "")
        x = x + h
    feat = F.layer_norm(x, (D,), sd["norm.weight"], sd["norm.bias"], 1e-6)[:, 0]
    ref = F.linear(feat, sd["head.weight"], sd["head.bias"])
    (ref * c).sum().backward()
    assert rel(logits, ref) < 5e-5, rel(logits, ref)
    assert rel(wrap.V.grad, Vr.grad) < 2e-4, rel(wrap.V.grad, Vr.grad)
