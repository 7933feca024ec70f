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
