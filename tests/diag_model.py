"""Diagnostic (not a test): per-tensor relative errors of the CUDA backbone vs the CPU oracle, no BN head
(loss = <feat, w> with a fixed random w), so conditioning of the tiny-batch BatchNorm does not enter."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import synth, peft_oracle as O
from oracle.synth import VitCfg
from tests.test_model_gpu import build_timm, rel

cases = {
    "kron_n4": VitCfg(dialect="timm", adapter="kron", phm_dim=4, img=64, patch=32, depth=2),
    "kron_n8_r2": VitCfg(dialect="timm", adapter="kron", phm_dim=8, phm_rank=2, img=64, patch=32, depth=2),
    "lora": VitCfg(dialect="timm", adapter="lora", img=64, patch=32, depth=2),
    "kron_n4_L197": VitCfg(dialect="timm", adapter="kron", phm_dim=4, img=224, patch=16, depth=2),
    "houlsby": VitCfg(dialect="timm", adapter="houlsby", img=64, patch=32, depth=2),
    "compacter": VitCfg(dialect="timm", adapter="compacter", img=64, patch=32, depth=2),
    "kron_n4_L50_d12": VitCfg(dialect="timm", adapter="kron", phm_dim=4, img=224, patch=32, depth=12),
}
B = int(os.environ.get("DIAG_B", "8"))
for name in sys.argv[1:] or list(cases):
    cfg = cases[name]
    sd = synth.make_state_dict(cfg, seed=0, grad_test=True)
    names = synth.trainable_names(cfg, sd)
    model = build_timm(cfg, sd)
    for n, p in model.named_parameters():
        p.requires_grad_(n in names)
    images, _ = synth.make_batch(cfg, B)
    w = torch.randn(B, cfg.feat_dim, generator=torch.Generator().manual_seed(3))
    feat = model(images.cuda())
    (feat * w.cuda()).sum().backward()
    params = dict(model.named_parameters())
    sd_o = {k: v.clone() for k, v in sd.items()}
    for k in names:
        sd_o[k].requires_grad_(True)
    rfeat = O.visual_forward(images, sd_o, cfg)
    (rfeat * w).sum().backward()
    print(f"== {name} B={B}: feat rel {rel(feat, rfeat):.2e}")
    for k in names:
        print(f"   {k:45s} rel {rel(params[k].grad, sd_o[k].grad):.3e}  |ref| {float(sd_o[k].grad.norm()):.3e}")

if sys.argv[1:]:
    sys.exit(0)
# full step with the BatchNorm head at a larger batch
from tests.test_model_gpu import run_step
for name in ["kron_n4", "lora"]:
    for Bb in (32, 128):
        cfg = cases[name]
        sd = synth.make_state_dict(cfg, seed=0, grad_test=True)
        names = synth.trainable_names(cfg, sd)
        model = build_timm(cfg, sd)
        feat, logits, loss, grads, hg = run_step(model, cfg, Bb, names)
        images, labels = synth.make_batch(cfg, Bb)
        sd_o = {k: v.clone() for k, v in sd.items()}
        r_logits, r_loss, r_grads, r_feat = O.train_step(images, labels, sd_o, synth.make_head(cfg), cfg, names)
        print(f"== BN-head {name} B={Bb}: feat {rel(feat, r_feat):.2e} logits {rel(logits, r_logits):.2e} loss {float(loss):.5f}/{float(r_loss):.5f}")
        print("   grads:", " ".join(f"{rel(grads[k], r_grads[k]):.1e}" for k in names))
        fc = r_feat - r_feat.mean(0, keepdim=True)
        print(f"   |feat| {float(r_feat.norm()):.2f}  |feat - batchmean| {float(fc.norm()):.2f}")
