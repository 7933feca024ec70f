"""fp32 verification mode (`model.precision = "fp32"`): the SAME engine code with fp32 buffers and the plain CUDA-core
kernels of csrc/f32_sm100.cuh, against the fp32 CPU oracle.

Gate (BASELINE.json north_star): <= 1e-5 relative (Frobenius) on the features and on the adapter gradient.  This pins the
algebra the bf16 path shares with it — K-extension, reshape trick / low-rank factorisation, dense fold, CLIP Q1/Q2
placement, token-reduction gradients, bottleneck adapters — at fp32 accuracy, where the bf16 tests can only show 2e-2.
"""
import pytest
import torch

from oracle import peft_oracle as O
from oracle import synth
from oracle.synth import VitCfg

from tests.test_clip_gpu import build_clip
from tests.test_model_gpu import build_timm

pytestmark = pytest.mark.gpu
TOL = 1e-5


def rel(a, b):
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-300))


CASES = {
    "timm_kron_n4": (VitCfg(dialect="timm", adapter="kron", phm_dim=4, img=64, patch=32, depth=2), 8),
    "timm_kron_n8_r2": (VitCfg(dialect="timm", adapter="kron", phm_dim=8, phm_rank=2, img=64, patch=32, depth=2), 5),
    "timm_lora": (VitCfg(dialect="timm", adapter="lora", img=64, patch=32, depth=2), 8),
    "timm_none": (VitCfg(dialect="timm", adapter="none", img=64, patch=32, depth=2), 3),
    "timm_houlsby": (VitCfg(dialect="timm", adapter="houlsby", img=64, patch=32, depth=2), 8),
    "timm_compacter": (VitCfg(dialect="timm", adapter="compacter", img=64, patch=32, depth=2), 6),
    "timm_kron_n64_dense": (VitCfg(dialect="timm", adapter="kron", phm_dim=64, img=64, patch=32, depth=2), 6),
    "timm_kron_n4_L197": (VitCfg(dialect="timm", adapter="kron", phm_dim=4, img=224, patch=16, depth=3), 3),
    "timm_kron_n4_L50_d12": (VitCfg(dialect="timm", adapter="kron", phm_dim=4, img=224, patch=32, depth=12), 4),
    "clip_c1_kron_n4": (VitCfg(dialect="clip", adapter="kron", phm_dim=4, img=224, patch=32, depth=3), 5),
    "clip_c1_lora": (VitCfg(dialect="clip", adapter="lora", img=224, patch=32, depth=3), 7),
    "clip_none": (VitCfg(dialect="clip", adapter="none", img=64, patch=32, depth=2), 3),
    "clip_kron_n64_dense": (VitCfg(dialect="clip", adapter="kron", phm_dim=64, img=64, patch=32, depth=2), 4),
    "clip_houlsby": (VitCfg(dialect="clip", adapter="houlsby", img=64, patch=32, depth=2), 8),
    "clip_compacter": (VitCfg(dialect="clip", adapter="compacter", img=64, patch=32, depth=2), 6),
    "clip_c4_houlsby_L257": (VitCfg(dialect="clip", adapter="houlsby", img=224, patch=14, width=1024, heads=16, out_dim=768, depth=2), 2),
}


@pytest.mark.parametrize("name", list(CASES))
def test_fp32_mode_matches_oracle(name):
    cfg, B = CASES[name]
    sd = synth.make_state_dict(cfg, seed=0, grad_test=True)
    if cfg.dialect == "clip" and cfg.adapter == "lora":
        # same conditioning fix as tests/test_clip_gpu.py: with N(0, .02^2) "up" factors the CLIP-dialect q update
        # (added after q / sqrt(hd), x32 LoRA scale) dominates q and saturates the softmax, which amplifies fp32
        # round-off on BOTH sides (unscaled, this case measures 1.5e-5 between two fp32 implementations)
        for k in sd:
            if "_proj_adapter2" in k:
                sd[k] = sd[k] * 0.125
    names = synth.trainable_names(cfg, sd)
    model = (build_timm if cfg.dialect == "timm" else build_clip)(cfg, sd)
    model.precision = "fp32"
    for n, p in model.named_parameters():
        p.requires_grad_(n in names)
    images, _ = synth.make_batch(cfg, B)
    w = torch.randn(B, cfg.feat_dim, generator=torch.Generator().manual_seed(3))
    if names:
        feat = model(images.cuda())
        (feat * w.cuda()).sum().backward()
    else:
        with torch.no_grad():
            feat = model(images.cuda())
    assert model._engine.cfg.precision == "fp32" and model._engine.wproj.dtype == torch.float32
    params = dict(model.named_parameters())
    sd_o = {k: v.clone() for k, v in sd.items()}
    for k in names:
        sd_o[k].requires_grad_(True)
    rfeat = O.visual_forward(images, sd_o, cfg)
    if names:
        (rfeat * w).sum().backward()
    assert rel(feat, rfeat) < TOL, rel(feat, rfeat)
    if names:
        got = torch.cat([params[k].grad.detach().float().cpu().reshape(-1) for k in names])
        ref = torch.cat([sd_o[k].grad.detach().float().reshape(-1) for k in names])
        assert rel(got, ref) < TOL, rel(got, ref)
    # per tensor: the same gate, with headroom only for tensors whose gradient is itself a small difference of large
    # terms (reported, not hidden: the assertion message carries the value)
    for k in names:
        r = rel(params[k].grad, sd_o[k].grad)
        assert r < 5 * TOL, (k, r)


FULL_CASES = {
    "timm_L5": (VitCfg(dialect="timm", adapter="none", img=64, patch=32, depth=2), 8),
    "timm_L50": (VitCfg(dialect="timm", adapter="none", img=224, patch=32, depth=2), 6),
    "clip_L50": (VitCfg(dialect="clip", adapter="none", img=224, patch=32, depth=2), 5),
}


@pytest.mark.parametrize("name", list(FULL_CASES))
def test_fp32_full_mode_all_tower_gradients(name):
    """train_all (full fine-tuning / Fastfood C5 backward): dL/dW of every tower tensor in fp32 mode vs the oracle."""
    from tests.test_full_gpu import build
    cfg, B = FULL_CASES[name]
    sd = synth.make_state_dict(cfg, seed=0, grad_test=True)
    model = build(cfg, sd)
    model.precision = "fp32"
    images, _ = synth.make_batch(cfg, B)
    w = torch.randn(B, cfg.feat_dim, generator=torch.Generator().manual_seed(3))
    feat = model(images.cuda())
    (feat * w.cuda()).sum().backward()
    assert model._engine_full.cfg.precision == "fp32"
    sd_o = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    rfeat = O.visual_forward(images, sd_o, cfg)
    (rfeat * w).sum().backward()
    assert rel(feat, rfeat) < TOL, rel(feat, rfeat)
    params = dict(model.named_parameters())
    got = torch.cat([params[k].grad.float().cpu().reshape(-1) for k in sd])
    ref = torch.cat([sd_o[k].grad.reshape(-1) for k in sd])
    assert rel(got, ref) < TOL, rel(got, ref)
    worst = max((rel(params[k].grad, sd_o[k].grad), k) for k in sd)
    assert worst[0] < 5 * TOL, worst


def test_fp32_and_bf16_engines_coexist():
    """Switching `precision` rebuilds the engine; the bf16 result stays within its own gate of the fp32 one."""
    cfg = VitCfg(dialect="timm", adapter="kron", phm_dim=4, img=64, patch=32, depth=2)
    sd = synth.make_state_dict(cfg, seed=0, grad_test=True)
    model = build_timm(cfg, sd)
    images, _ = synth.make_batch(cfg, 4)
    with torch.no_grad():
        f16 = model(images.cuda()).clone()
        model.precision = "fp32"
        f32 = model(images.cuda()).clone()
        model.precision = "bf16"
        f16b = model(images.cuda()).clone()
    assert rel(f16, f32) < 2e-2 and rel(f16, f32) > 1e-6
    assert torch.equal(f16, f16b)
