"""Edge cases of the drop-in boundary and of the fp32 verification kernels: batch 1, strided inputs, stale activations,
in-place edits of frozen weights, ragged operand shapes."""
import pytest
import torch

from oracle import peft_oracle as O
from oracle import synth
from oracle.synth import VitCfg

from tests.test_model_gpu import build_timm

pytestmark = pytest.mark.gpu


def rel(a, b):
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-300))


def _model(adapter="kron", depth=2):
    cfg = VitCfg(dialect="timm", adapter=adapter, phm_dim=4, img=64, patch=32, depth=depth)
    sd = synth.make_state_dict(cfg, seed=0, grad_test=True)
    names = synth.trainable_names(cfg, sd)
    model = build_timm(cfg, sd)
    for n, p in model.named_parameters():
        p.requires_grad_(n in names)
    return cfg, sd, names, model


@pytest.mark.parametrize("precision,tol", [("bf16", 2e-2), ("fp32", 1e-5)])
def test_batch_of_one(precision, tol):
    cfg, sd, names, model = _model()
    model.precision = precision
    images, _ = synth.make_batch(cfg, 1)
    feat = model(images.cuda())
    feat.sum().backward()
    sd_o = {k: v.clone().requires_grad_(k in names) for k, v in sd.items()}
    rfeat = O.visual_forward(images, sd_o, cfg)
    rfeat.sum().backward()
    assert rel(feat, rfeat) < tol
    params = dict(model.named_parameters())
    got = torch.cat([params[k].grad.float().cpu().reshape(-1) for k in names])
    ref = torch.cat([sd_o[k].grad.reshape(-1) for k in names])
    assert rel(got, ref) < tol, rel(got, ref)


def test_strided_images_and_eval_mode():
    """Non-contiguous image tensors are accepted (made contiguous at the boundary); eval()/no_grad gives the same features
    as the training forward (the tower has no dropout / stochastic depth in the reference configs)."""
    cfg, sd, names, model = _model()
    images, _ = synth.make_batch(cfg, 4)
    x = images.cuda()
    strided = x.permute(0, 2, 3, 1).contiguous().permute(0, 3, 1, 2)        # channels-last memory, NCHW view
    assert not strided.is_contiguous()
    f_train = model(x).detach().clone()
    f_strided = model(strided).detach().clone()
    model.eval()
    with torch.no_grad():
        f_eval = model(x).clone()
    assert torch.equal(f_train, f_strided)
    assert torch.equal(f_train, f_eval)


def test_backward_after_a_later_forward_raises():
    """The engine keeps ONE set of saved activations per batch size: backward through a stale graph must fail loudly,
    never silently use the activations of the later forward."""
    cfg, sd, names, model = _model()
    images, _ = synth.make_batch(cfg, 4)
    f1 = model(images.cuda())
    _ = model(images.cuda() * 0.5)
    with pytest.raises(RuntimeError, match="overwritten"):
        f1.sum().backward()


def test_backward_after_a_forward_at_another_batch_size_raises():
    """the adapter operands the backward reads (K-extension columns of W_qkv, Up^T, Dn) are per ENGINE, not per
    batch size — a forward at a different batch size with changed adapter weights rewrites them, so the older graph's
    backward must be refused as well (it would silently use the new weights)."""
    cfg, sd, names, model = _model()
    images, _ = synth.make_batch(cfg, 8)
    f8 = model(images.cuda())
    with torch.no_grad():
        for n, p in model.named_parameters():
            if n in names:
                p.mul_(1.5)
    _ = model(images[:4].cuda())
    with pytest.raises(RuntimeError, match="overwritten"):
        f8.sum().backward()


def test_lora_rank_24_tail_chunk():
    """rank_pad % 64 == 48 (LoRA r = 24 -> 48 adapter columns) made the engine hand peft_tn_reduce a 48-wide
    chunk, which the kernel rejects; the tail is now split 32 + 16.  Forward AND backward against the oracle."""
    cfg = VitCfg(dialect="timm", adapter="lora", lora_r=24, img=64, patch=32, depth=2)
    sd = synth.make_state_dict(cfg, seed=0, grad_test=True)
    for k in sd:
        if "_proj_adapter2" in k:
            sd[k] = sd[k] * 0.4          # 6x the default rank: keep |delta q| in the range of the r = 4 cases
    names = synth.trainable_names(cfg, sd)
    model = build_timm(cfg, sd)
    for n, p in model.named_parameters():
        p.requires_grad_(n in names)
    images, _ = synth.make_batch(cfg, 6)
    w = torch.randn(6, cfg.feat_dim, generator=torch.Generator().manual_seed(3))
    feat = model(images.cuda())
    (feat * w.cuda()).sum().backward()
    sd_o = {k: v.clone().requires_grad_(k in names) for k, v in sd.items()}
    rfeat = O.visual_forward(images, sd_o, cfg)
    (rfeat * w).sum().backward()
    assert rel(feat, rfeat) < 2e-2
    params = dict(model.named_parameters())
    got = torch.cat([params[k].grad.float().cpu().reshape(-1) for k in names])
    ref = torch.cat([sd_o[k].grad.reshape(-1) for k in names])
    assert rel(got, ref) < 2e-2, rel(got, ref)


def test_frozen_weight_edit_is_picked_up():
    """load_state_dict / in-place edits of a frozen tensor after the first forward re-pack the engine's operands."""
    cfg, sd, names, model = _model()
    images, _ = synth.make_batch(cfg, 3)
    with torch.no_grad():
        f0 = model(images.cuda()).clone()
        model.blocks[1].mlp.fc2.weight.mul_(0.5)
        f1 = model(images.cuda()).clone()
    sd2 = {k: v.clone() for k, v in sd.items()}
    sd2["blocks.1.mlp.fc2.weight"] *= 0.5
    r1 = O.visual_forward(images, sd2, cfg)
    assert rel(f0, f1) > 1e-3
    assert rel(f1, r1) < 2e-2


def test_fp32_kernels_ragged_shapes():
    """peft_gemm_f32 / peft_tn_reduce_f32 / peft_attention_*_f32 on shapes that are multiples of nothing."""
    from peft_b200 import ops
    g = torch.Generator().manual_seed(0)
    a, b, bias = torch.randn(37, 13, generator=g), torch.randn(19, 13, generator=g), torch.randn(19, generator=g)
    out = torch.empty(37, 19, device="cuda")
    ops.gemm(a.cuda(), b.cuda(), out, bias=bias.cuda(), epilogue=ops.EPI_STORE_F32)
    assert rel(out, a @ b.T + bias) < 1e-6
    x, y = torch.randn(2500, 7, generator=g), torch.randn(2500, 21, generator=g)
    acc = torch.ones(7, 21, device="cuda")
    ops.tn_reduce(x.cuda(), y.cuda(), acc)
    assert rel(acc, 1.0 + x.T @ y) < 1e-5
    for B, H, L in [(2, 3, 5), (1, 2, 197), (2, 1, 130)]:
        D = H * 64
        qkv = torch.randn(B * L, 3 * D, generator=g)
        d_o = torch.randn(B * L, D, generator=g)
        qr = qkv.clone().requires_grad_(True)
        q, k, v = [t.view(B, L, H, 64).transpose(1, 2) for t in qr.split(D, dim=1)]
        ref = torch.softmax(q @ k.transpose(-1, -2) * 0.125, -1) @ v
        ref = ref.transpose(1, 2).reshape(B * L, D)
        (ref * d_o).sum().backward()
        o = torch.empty(B * L, D, device="cuda")
        lse = torch.empty(B * H * L, device="cuda")
        delta = torch.empty(B * H * L, device="cuda")
        dqkv = torch.zeros(B * L, 3 * D, device="cuda")
        qc = qkv.cuda()
        ops.attention_fwd(qc, o, lse, B, H, L, 0.125)
        ops.attention_bwd(qc, o, d_o.cuda(), lse, delta, dqkv, B, H, L, 0.125)
        assert rel(o, ref) < 1e-5, (B, H, L, rel(o, ref))
        assert rel(dqkv, qr.grad) < 1e-5, (B, H, L, rel(dqkv, qr.grad))
