"""Mask-conditioned oracle for the Houlsby bottleneck (FS/adapter_model.py:204-295).

The adapter's ReLU is discontinuous: a pre-activation that lies within the bf16 forward error of 0 can come out on the
other side in the bf16 engine than in the fp32 oracle, and the derivative of everything below the gate then differs by
O(1) at that unit.  No amount of precision in one product removes this (the forward error of the whole tower feeds the
pre-activation), so the parity statement is made in two parts:
  (1) the engine's gate decisions differ from the oracle's only where the oracle's pre-activation is within a small
      multiple of the bf16 noise of 0, and on a small fraction of the units;
  (2) on the linear region the engine selected (oracle evaluated with the ENGINE's gate mask) every adapter gradient —
      including adapter_down / adapter_norm_before — meets the north_star 2e-2 bound.
"""
import torch

from oracle import peft_oracle as O


def engine_gate_masks(model, cfg, B):
    """{oracle adapter prefix: [0/1 mask per application, shaped like the oracle's pre-activation]} from the engine's saved
    ReLU outputs (engine layer j = j-th block application; AdapterDrop schedules apply a block more than once)."""
    ws = model._engine.ws[B]
    m, L = cfg.bottleneck, cfg.seq
    masks = {}
    sched = list(cfg.block_schedule) if cfg.block_schedule else [(i, (not cfg.adapter_layers) or i in cfg.adapter_layers) for i in range(cfg.depth)]
    for j, (b, use) in enumerate(sched):
        if not use:
            continue                                            # this application does not call the adapter
        z = (ws.zext[j][:, :m] > 0).cpu()                       # rows b*L + l
        if cfg.dialect == "timm":
            masks.setdefault(f"blocks.{b}.adapter.", []).append(z.view(B, L, m))
        else:
            masks.setdefault(f"transformer.resblocks.{b}.adapter.", []).append(z.view(B, L, m).permute(1, 0, 2).contiguous())   # LND
    return masks


def oracle_with_engine_gates(images, sd_o, cfg, masks, cotangent):
    """Oracle forward/backward on the engine's linear region; returns (features, flip statistics)."""
    O.GATE_RECORD = {}
    try:
        with torch.no_grad():
            O.visual_forward(images, sd_o, cfg)                 # plain ReLU: records the oracle's own pre-activations
        pre = dict(O.GATE_RECORD)
    finally:
        O.GATE_RECORD = None
    flipped = total = 0
    worst = 0.0
    for pfx, mlist in masks.items():
        for mask, p in zip(mlist, pre[pfx]):
            diff = mask != (p > 0)
            flipped += int(diff.sum())
            total += diff.numel()
            if diff.any():
                worst = max(worst, float(p[diff].abs().max() / p.std()))
    O.GATE_OVERRIDE = masks
    O.GATE_CALLS = {}
    try:
        feat = O.visual_forward(images, sd_o, cfg)
        (feat * cotangent).sum().backward()
    finally:
        O.GATE_OVERRIDE = {}
        O.GATE_CALLS = {}
    return feat, dict(flip_frac=flipped / max(1, total), worst_flip_sigma=worst)
