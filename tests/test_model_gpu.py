"""Model-level parity on the GPU: features, logits, loss and every adapter gradient of one training step of the CUDA
engine against (a) the committed golden outputs of the unmodified reference and (b) the CPU oracle at larger shapes.

Tolerance (north_star): bf16 mode <= 2e-2 relative (Frobenius) on logits and adapter gradients.
"""
import pytest
import torch
import torch.nn.functional as F

from oracle import peft_oracle as O
from oracle import synth
from oracle.synth import VitCfg

pytestmark = pytest.mark.gpu
TOL = 2e-2


def rel(a, b):
    a = a.detach().float().cpu()
    b = b.detach().float().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


def build_timm(cfg: VitCfg, sd):
    from peft_b200.timm_vit import VisionTransformer
    kw = {}
    if cfg.adapter == "kron":
        kw = dict(phm_dim=cfg.phm_dim, phm_rank=cfg.phm_rank)
    elif cfg.adapter == "lora":
        kw = dict(lora_attn_dim=cfg.lora_r, lora_attn_alpha=cfg.lora_alpha)
    elif cfg.adapter in ("houlsby", "compacter"):
        kw = dict(bottleneck=cfg.bottleneck, compacter_n_down=cfg.compacter_n_down, compacter_n_up=cfg.compacter_n_up)
    if getattr(cfg, "adapter_layers", ()):
        kw["adapter_layers"] = cfg.adapter_layers
    if getattr(cfg, "block_schedule", ()):
        kw["block_schedule"] = cfg.block_schedule
    m = VisionTransformer(img_size=cfg.img, patch_size=cfg.patch, num_classes=cfg.num_classes, embed_dim=cfg.width,
                          depth=cfg.depth, num_heads=cfg.heads, mlp_ratio=cfg.mlp_ratio, qkv_bias=True,
                          adapter=cfg.adapter, **kw)
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all(k.startswith("head.") or ".adapter_down.0." in k for k in missing), missing
    return m.cuda()


def run_step(model, cfg, batch, names):
    from peft_b200.classifier import Classifier
    head = synth.make_head(cfg)
    clf = Classifier(model, cfg.feat_dim, cfg.num_classes, trainable_filter=lambda n: n in names).cuda()
    clf.layers[0].load_state_dict({"weight": head["layers.0.weight"], "bias": head["layers.0.bias"]})
    clf.train()
    images, labels = synth.make_batch(cfg, batch)
    feat = model(images.cuda())
    logits = clf.layers(clf.channel_bn(feat))
    loss = F.cross_entropy(logits, labels.cuda())
    loss.backward()
    params = dict(model.named_parameters())
    grads = {k: params[k].grad for k in names}
    hg = {"layers.0.weight": clf.layers[0].weight.grad, "layers.0.bias": clf.layers[0].bias.grad}
    return feat, logits, loss, grads, hg


@pytest.mark.parametrize("name", ["timm_kron_n4", "timm_kron_n8_r2", "timm_lora", "timm_houlsby", "timm_adapterdrop"])
def test_features_match_reference_golden(golden, name):
    """Forward features against the committed outputs of the unmodified reference modules."""
    c = golden[name]
    cfg = VitCfg(**c["cfg"])
    sd = synth.make_state_dict(cfg, seed=0, grad_test=True)
    model = build_timm(cfg, sd)
    images, _ = synth.make_batch(cfg, c["batch"])
    with torch.no_grad():
        feat = model(images.cuda())
    assert rel(feat, c["feature"]) < TOL


BACKBONE_CASES = {
    "kron_n4": (VitCfg(dialect="timm", adapter="kron", phm_dim=4, img=64, patch=32, depth=2), 8),
    "kron_n8_r2": (VitCfg(dialect="timm", adapter="kron", phm_dim=8, phm_rank=2, img=64, patch=32, depth=2), 5),
    "lora": (VitCfg(dialect="timm", adapter="lora", img=64, patch=32, depth=2), 8),
    "none": (VitCfg(dialect="timm", adapter="none", img=64, patch=32, depth=2), 3),
    "houlsby": (VitCfg(dialect="timm", adapter="houlsby", img=64, patch=32, depth=2), 8),
    "compacter": (VitCfg(dialect="timm", adapter="compacter", img=64, patch=32, depth=2), 6),
    "houlsby_L197": (VitCfg(dialect="timm", adapter="houlsby", img=224, patch=16, depth=2), 2),
    "kron_n4_L197": (VitCfg(dialect="timm", adapter="kron", phm_dim=4, img=224, patch=16, depth=3), 3),
    "lora_L197": (VitCfg(dialect="timm", adapter="lora", img=224, patch=16, depth=2), 2),
    # large-n KAdaptation (the committed phm_dim is 768): dense "fold" form of the update, dH as one GEMM
    "kron_n64_dense": (VitCfg(dialect="timm", adapter="kron", phm_dim=64, img=64, patch=32, depth=2), 6),
    "kron_n192_dense_L50": (VitCfg(dialect="timm", adapter="kron", phm_dim=192, img=224, patch=32, depth=2), 3),
    # the COMMITTED hyper-parameters (phm_dim = 768 = D, K = P = 1): peft_phm_fold path, phm_rule frozen as in few_shot
    "kron_n768_committed": (VitCfg(dialect="timm", adapter="kron", phm_dim=768, img=64, patch=32, depth=1), 2),
    "kron_n4_L50_d12": (VitCfg(dialect="timm", adapter="kron", phm_dim=4, img=224, patch=32, depth=12), 4),
    # AdapterDrop (FS/adapterdrop.py): the Houlsby adapter is applied and trained in the listed blocks only
    "adapterdrop_last": (VitCfg(dialect="timm", adapter="houlsby", adapter_layers=(3,), img=64, patch=32, depth=4), 6),
    "adapterdrop_two": (VitCfg(dialect="timm", adapter="houlsby", adapter_layers=(1, 2), img=224, patch=32, depth=3), 4),
    "compacter_drop": (VitCfg(dialect="timm", adapter="compacter", adapter_layers=(1,), img=64, patch=32, depth=2), 5),
    # the COMMITTED FS/adapterdrop.py loop: every block applied twice (adapter on the first pass, on the second only in the last
    # block), only the last block's adapter trainable (pinned by the golden case timm_adapterdrop at depth 12)
    "adapterdrop_committed": (VitCfg(dialect="timm", adapter="houlsby", img=64, patch=32, depth=3,
                                     block_schedule=((0, True), (0, False), (1, True), (1, False), (2, True), (2, True))), 6),
}


@pytest.mark.parametrize("name", list(BACKBONE_CASES))
def test_backbone_grads_match_oracle(name):
    """Features and every adapter gradient under a fixed random cotangent (loss = <feat, w>): isolates the tower from
    the conditioning of the BatchNorm head."""
    cfg, B = BACKBONE_CASES[name]
    sd = synth.make_state_dict(cfg, seed=0, grad_test=True)
    names = synth.trainable_names(cfg, sd, train_rule=cfg.phm_dim <= 256)
    if cfg.phm_dim == 768:
        for k in sd:                      # n = 768 sums 768 terms per entry: keep |H| in the range of the small-n cases
            if "W_left" in k or "W_right" in k:
                sd[k] = sd[k] * 3.0
    model = build_timm(cfg, sd)
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
    params = dict(model.named_parameters())
    sd_o = {k: v.clone() for k, v in sd.items()}
    for k in names:
        sd_o[k].requires_grad_(True)
    if cfg.adapter == "houlsby":
        # discontinuous ReLU gate: the oracle differentiates the linear region the engine selected, and the gate decisions
        # themselves may differ only within bf16 noise of 0 (tests/houlsby_gate.py) — then EVERY tensor holds 2e-2
        from tests.houlsby_gate import engine_gate_masks, oracle_with_engine_gates
        rfeat, st = oracle_with_engine_gates(images, sd_o, cfg, engine_gate_masks(model, cfg, B), w)
        assert st["flip_frac"] < 1e-2 and st["worst_flip_sigma"] < 0.1, st
    else:
        rfeat = O.visual_forward(images, sd_o, cfg)
        if names:
            (rfeat * w).sum().backward()
    assert rel(feat, rfeat) < TOL
    if names:
        # north_star gate: <= 2e-2 relative on the adapter gradient (all trainable tensors taken as one vector) ...
        got = torch.cat([params[k].grad.detach().float().cpu().reshape(-1) for k in names])
        ref = torch.cat([sd_o[k].grad.detach().float().reshape(-1) for k in names])
        assert rel(got, ref) < TOL, rel(got, ref)
    # ... and per tensor: 2e-2 for the shallow towers; at depth 12 the bf16 operand rounding of the forward (feature
    # error ~8e-3) shows up as 1-2e-2 on the smallest-norm factors (|grad| ~ 1e-2), so the per-tensor bound is 3e-2 there
    per_tensor = TOL if cfg.depth <= 3 else 1.5 * TOL
    for k in names:
        assert rel(params[k].grad, sd_o[k].grad) < per_tensor, (k, rel(params[k].grad, sd_o[k].grad))


@pytest.mark.parametrize("adapter", ["kron", "lora"])
def test_full_step_with_bn_head(adapter):
    """train_one semantics (BatchNorm1d head, CrossEntropy) at a batch large enough for the batch statistics to be
    well conditioned: logits, loss, adapter and head gradients against the CPU oracle."""
    cfg = VitCfg(dialect="timm", adapter=adapter, img=64, patch=32, depth=2)
    B = 32
    sd = synth.make_state_dict(cfg, seed=0, grad_test=True)
    names = synth.trainable_names(cfg, sd)
    model = build_timm(cfg, sd)
    feat, logits, loss, grads, hg = run_step(model, cfg, B, names)
    images, labels = synth.make_batch(cfg, B)
    sd_o = {k: v.clone() for k, v in sd.items()}
    r_logits, r_loss, r_grads, r_feat = O.train_step(images, labels, sd_o, synth.make_head(cfg), cfg, names)
    assert rel(feat, r_feat) < TOL
    assert rel(logits, r_logits) < TOL
    assert abs(float(loss) - float(r_loss)) < 1e-2
    for k in names:
        assert rel(grads[k], r_grads[k]) < TOL, (k, rel(grads[k], r_grads[k]))
    for k in hg:
        assert rel(hg[k], r_grads[k]) < TOL, k


def test_no_cpu_fallback():
    from peft_b200.timm_vit import VisionTransformer
    m = VisionTransformer(img_size=64, patch_size=32, embed_dim=768, depth=1, num_heads=12, qkv_bias=True)
    with pytest.raises(RuntimeError):
        m(torch.zeros(1, 3, 64, 64))
    with pytest.raises(RuntimeError):
        m.blocks[0](torch.zeros(1, 5, 768))
