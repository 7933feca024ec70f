"""Trainer-side pieces kept API-compatible with the reference (few_shot/vision_benchmark/evaluation/kadaptation_clip.py).

`Classifier` = backbone -> BatchNorm1d(affine=False) -> Linear (:83-168); `train_one_step` = the body of train_one
(:308-314): zero_grad, forward, CrossEntropy, backward, optimizer step.  The head is [B, E] x [E, classes] — negligible
work — and stays in torch; everything under `backbone` is the CUDA engine.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


class Classifier(nn.Module):
    def __init__(self, backbone: nn.Module, feat_dim: int, num_classes: int, trainable_filter=None, sync_bn: bool = False):
        super().__init__()
        self.backbone = backbone
        if trainable_filter is not None:
            # the reference freezes everything whose name does not match ('adapter' / 'compacter' / 'W'), :99-116
            for n, p in self.backbone.named_parameters():
                p.requires_grad = bool(trainable_filter(n))
        if sync_bn:     # multi-GPU option: batch statistics over the global batch (peft_b200.dist.GlobalBatchNorm1d)
            from .dist import GlobalBatchNorm1d
            self.channel_bn = GlobalBatchNorm1d(feat_dim)
        else:
            self.channel_bn = nn.BatchNorm1d(feat_dim, affine=False)
        self.layers = nn.Sequential(nn.Linear(feat_dim, num_classes))

    @torch.no_grad()
    def init_head_with_text_encoder(self, zeroshot_weights, logit_scale=None):
        """TRAIN.INIT_HEAD_WITH_TEXT_ENCODER (every *_clip.sh script sets it; FS/kadaptation_clip.py:128-138,154-155): the
        Linear head starts from the zero-shot classifier — `zeroshot_weights` [classes, E] from
        `CLIP.zeroshot_head_weights(...)`, bias 0, optionally scaled by exp(logit_scale) (INIT_HEAD_WITH_LOGIT_SCALE)."""
        lin = self.layers[0]
        if tuple(zeroshot_weights.shape) != tuple(lin.weight.shape):
            raise ValueError(f"zeroshot_weights must be {tuple(lin.weight.shape)}, got {tuple(zeroshot_weights.shape)}")
        lin.weight.copy_(zeroshot_weights.to(lin.weight.dtype))
        lin.bias.fill_(0.0)
        if logit_scale is not None:
            lin.weight.mul_(torch.as_tensor(logit_scale, dtype=lin.weight.dtype, device=lin.weight.device).exp())

    def forward(self, img):
        feature = self.backbone(img).to(img.dtype)
        return self.layers(self.channel_bn(feature))


def train_one_step(model: Classifier, images, target, optimizer, criterion=None):
    """One iteration of train_one (kadaptation_clip.py:308-314). Returns (loss tensor, logits)."""
    criterion = criterion or F.cross_entropy
    optimizer.zero_grad(set_to_none=True)
    output = model(images)
    loss = criterion(output, target)
    loss.backward()
    optimizer.step()
    return loss.detach(), output.detach()
