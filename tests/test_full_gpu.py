"""Full fine-tuning / intrinsic-dimension mode (SURVEY config C5, row R12): dL/dW for EVERY tower tensor from the CUDA
engine, and FastfoodWrap around the engine-backed VisionTransformer, against the CPU oracle + torch autograd.
Tolerance: bf16 mode, <= 2e-2 relative on the whole gradient vector (north_star), 4e-2 on any single tensor."""
import numpy as np
import pytest
import torch

from oracle import peft_oracle as O
from oracle import synth
from oracle.synth import VitCfg

pytestmark = pytest.mark.gpu
TOL = 2e-2


def rel(a, b):
    a = a.detach().float().cpu()
    b = b.detach().float().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


def build(cfg, sd):
    if cfg.dialect == "timm":
        from peft_b200.timm_vit import VisionTransformer
        m = VisionTransformer(img_size=cfg.img, patch_size=cfg.patch, num_classes=0, embed_dim=cfg.width, depth=cfg.depth,
                              num_heads=cfg.heads, mlp_ratio=cfg.mlp_ratio, qkv_bias=True, adapter="none")
    else:
        from peft_b200.clip_model import VisionTransformer
        m = VisionTransformer(cfg.img, cfg.patch, cfg.width, cfg.depth, cfg.heads, cfg.out_dim, adapter="none")
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected and not missing, (missing, unexpected)
    return m.cuda()


CASES = {
    "timm_L5": (VitCfg(dialect="timm", adapter="none", img=64, patch=32, depth=2), 8),
    "timm_L50": (VitCfg(dialect="timm", adapter="none", img=224, patch=32, depth=2), 6),
    "clip_L50": (VitCfg(dialect="clip", adapter="none", img=224, patch=32, depth=2), 5),
    "clip_L5_d3": (VitCfg(dialect="clip", adapter="none", img=64, patch=32, depth=3), 16),
}


@pytest.mark.parametrize("name", list(CASES))
def test_all_tower_gradients_match_oracle(name):
    cfg, B = CASES[name]
    sd = synth.make_state_dict(cfg, seed=0, grad_test=True)
    model = build(cfg, sd)
    images, _ = synth.make_batch(cfg, B)
    w = torch.randn(B, cfg.feat_dim, generator=torch.Generator().manual_seed(3))
    feat = model(images.cuda())
    (feat * w.cuda()).sum().backward()
    sd_o = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    rfeat = O.visual_forward(images, sd_o, cfg)
    (rfeat * w).sum().backward()
    assert rel(feat, rfeat) < TOL
    params = dict(model.named_parameters())
    names = list(sd)
    for k in names:
        assert params[k].grad is not None, k
    got = torch.cat([params[k].grad.float().cpu().reshape(-1) for k in names])
    ref = torch.cat([sd_o[k].grad.reshape(-1) for k in names])
    assert rel(got, ref) < TOL, rel(got, ref)
    worst = max((rel(params[k].grad, sd_o[k].grad), k) for k in names)
    assert worst[0] < 2 * TOL, worst


def test_bias_only_tuning_uses_full_engine():
    """BitFit-style selection: only biases require grad; the others get no .grad."""
    cfg = VitCfg(dialect="timm", adapter="none", img=64, patch=32, depth=2)
    sd = synth.make_state_dict(cfg, seed=0, grad_test=True)
    model = build(cfg, sd)
    for n, p in model.named_parameters():
        p.requires_grad_(n.endswith(".bias"))
    images, _ = synth.make_batch(cfg, 4)
    model(images.cuda()).sum().backward()
    sd_o = {k: v.clone().requires_grad_(k.endswith(".bias")) for k, v in sd.items()}
    O.visual_forward(images, sd_o, cfg).sum().backward()
    for n, p in model.named_parameters():
        assert (p.grad is not None) == n.endswith(".bias"), n
        if p.grad is not None:
            assert rel(p.grad, sd_o[n].grad) < 2 * TOL, (n, rel(p.grad, sd_o[n].grad))


def test_fastfood_wrap_vit_end_to_end():
    """C5 at reduced size: FastfoodWrap(VisionTransformer, d) -> features and dL/dV against the oracle (reference formula
    for the projection with the wrapper's own B / Pi / G, oracle tower, torch autograd)."""
    from peft_b200.fastfood import FastfoodWrap
    cfg = VitCfg(dialect="timm", adapter="none", img=64, patch=32, depth=2)
    sd = synth.make_state_dict(cfg, seed=0, grad_test=True)
    model = build(cfg, sd)
    torch.manual_seed(0)
    np.random.seed(0)
    d = 200
    wrap = FastfoodWrap(model, d, device=0)
    assert [n for n, _ in wrap.named_parameters()] == ["V"]
    with torch.no_grad():
        wrap.V.copy_(torch.randn(d, generator=torch.Generator().manual_seed(5)) * 0.5)
    B = 8
    images, _ = synth.make_batch(cfg, B)
    w = torch.randn(B, cfg.feat_dim, generator=torch.Generator().manual_seed(3))
    feat = wrap(images.cuda())
    (feat * w.cuda()).sum().backward()
    Vr = wrap.V.detach().cpu().clone().requires_grad_(True)
    sd_o = {}
    for k, v in sd.items():
        BB, Pi, GG, divisor, LL = wrap.fastfood_params[k]
        sd_o[k] = v + O.fastfood_torched(Vr, v.numel(), [BB, Pi, GG, divisor, LL]).view(v.shape)
    rfeat = O.visual_forward(images, sd_o, cfg)
    (rfeat * w).sum().backward()
    assert rel(feat, rfeat) < TOL, rel(feat, rfeat)
    assert rel(wrap.V.grad, Vr.grad) < TOL, rel(wrap.V.grad, Vr.grad)
    # a second step re-reads the re-projected weights (no stale pre-packed copies, SURVEY §8b last row)
    with torch.no_grad():
        wrap.V.mul_(-1.0)
        f2 = wrap(images.cuda())
    sd2 = {}
    for k, v in sd.items():
        sd2[k] = v + O.fastfood_torched(-Vr.detach(), v.numel(), wrap.fastfood_params[k]).view(v.shape)
    assert rel(f2, O.visual_forward(images, sd2, cfg)) < TOL
