"""Stochastic depth (timm DropPath; FS/kadaptation.py:263,276-277, rate rule :408; full_shot yaml DROP_PATH_RATE 0.1) folded
into the residual GEMM epilogues: parity against the oracle under FIXED per-sample keep masks, in the bf16 path (2e-2),
in the fp32 verification mode (1e-5), and at a geometry where the 2-CTA kernel carries the row scale."""
import pytest
import torch

from oracle import peft_oracle as O
from oracle import synth
from oracle.synth import VitCfg

pytestmark = pytest.mark.gpu


def rel(a, b):
    a = a.detach().float().cpu()
    b = b.detach().float().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


def build(cfg, sd, rate):
    from peft_b200.timm_vit import VisionTransformer
    kw = dict(phm_dim=cfg.phm_dim, phm_rank=cfg.phm_rank) if cfg.adapter == "kron" else dict(lora_attn_dim=cfg.lora_r, lora_attn_alpha=cfg.lora_alpha)
    m = VisionTransformer(img_size=cfg.img, patch_size=cfg.patch, num_classes=10, embed_dim=cfg.width, depth=cfg.depth,
                          num_heads=cfg.heads, qkv_bias=True, adapter=cfg.adapter, drop_path_rate=rate, **kw)
    m.load_state_dict(sd, strict=False)
    return m.cuda()


CASES = {
    "kron_L5": (VitCfg(dialect="timm", adapter="kron", phm_dim=4, img=64, patch=32, depth=3), 8, 0.3),
    "lora_L50": (VitCfg(dialect="timm", adapter="lora", img=224, patch=32, depth=3), 6, 0.2),
    "kron_L197_B32": (VitCfg(dialect="timm", adapter="kron", phm_dim=4, img=224, patch=16, depth=2), 32, 0.1),
}


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
@pytest.mark.parametrize("name", list(CASES))
def test_drop_path_fixed_mask_matches_oracle(name, precision):
    cfg, B, rate = CASES[name]
    if precision == "fp32" and B > 8:
        pytest.skip("fp32 verification kernels: small cases only")
    sd = synth.make_state_dict(cfg, seed=0, grad_test=True)
    names = synth.trainable_names(cfg, sd)
    model = build(cfg, sd, rate)
    model.precision = precision
    model.train()
    for n, p in model.named_parameters():
        p.requires_grad_(n in names)
    g = torch.Generator().manual_seed(11)
    keep = (torch.rand(cfg.depth, 2, B, generator=g) > 0.35).float()       # dropped samples in every block
    keep[0] = 1.0                                                           # block 0 has rate 0 (linspace rule)
    assert float(keep.min()) == 0.0
    model.drop_path_keep = keep
    images, _ = synth.make_batch(cfg, B)
    w = torch.randn(B, cfg.feat_dim, generator=torch.Generator().manual_seed(3))
    feat = model(images.cuda())
    (feat * w.cuda()).sum().backward()
    sd_o = {k: v.clone() for k, v in sd.items()}
    for k in names:
        sd_o[k].requires_grad_(True)
    rfeat = O.timm_forward_features(images, sd_o, cfg, drop_keep=keep, drop_path_rate=rate)
    (rfeat * w).sum().backward()
    tol = 2e-2 if precision == "bf16" else 1e-5
    assert rel(feat, rfeat) < tol
    params = dict(model.named_parameters())
    got = torch.cat([params[k].grad.detach().float().cpu().reshape(-1) for k in names])
    ref = torch.cat([sd_o[k].grad.detach().float().reshape(-1) for k in names])
    assert rel(got, ref) < tol, rel(got, ref)
    # and it is NOT the no-drop result (the mask really acted)
    with torch.no_grad():
        model.eval()
        feat_eval = model(images.cuda())
    assert rel(feat_eval, rfeat) > 10 * tol


def test_drop_path_random_draw_statistics():
    """Training mode draws a fresh Bernoulli(keep_prob) mask per sample, block and branch; eval mode is the identity."""
    from peft_b200.timm_vit import sample_drop_path
    cfg = VitCfg(dialect="timm", adapter="kron", phm_dim=4, img=64, patch=32, depth=4)
    model = build(cfg, synth.make_state_dict(cfg, seed=0, grad_test=True), 0.3)
    model.train()
    B, L = 4096, 5
    d = sample_drop_path(model, B, L, torch.device("cuda"))
    assert tuple(d.shape) == (4, 2, B * L)
    rates = torch.linspace(0, 0.3, 4)
    for i in range(4):
        kp = 1.0 - float(rates[i])
        vals = d[i].unique()
        assert all(abs(float(v)) < 1e-6 or abs(float(v) - 1.0 / kp) < 1e-5 for v in vals)
        frac = float((d[i] > 0).float().mean())
        assert abs(frac - kp) < 0.03
        rows = d[i, 0].view(B, L)
        assert bool((rows == rows[:, :1]).all())          # one draw per SAMPLE (all L token rows share it)
    model.eval()
    assert sample_drop_path(model, B, L, torch.device("cuda")) is None
