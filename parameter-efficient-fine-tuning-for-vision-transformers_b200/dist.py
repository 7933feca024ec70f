"""Batch data parallelism for the PEFT step: one process per GPU, ONE flat all-reduce bucket per step.

Only the trainable tensors (adapter factors + classifier head, 0.15-13 MB fp32: SURVEY.md §8e) ever cross NVLink; the
frozen backbone is replicated.  Every trainable parameter's `.grad` is a view into one flat fp32 buffer, so backward
accumulates straight into the bucket and the exchange is a single `all_reduce` (NCCL on GPUs, gloo in the CPU tests)
with no packing copies.  The reference has no multi-GPU PEFT path (its trainers are single-GPU,
few_shot/vision_benchmark/evaluation/kadaptation_clip.py:239-243); semantics here: gradients are averaged over ranks
(global-batch mean), BatchNorm head statistics stay per-replica, LR scaling by world size is left to the caller
(few_shot/vision_benchmark/config/default.py:248).
"""
from __future__ import annotations

import torch
import torch.distributed as dist


class GradBucket:
    def __init__(self, params, process_group=None):
        self.params = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("GradBucket: no trainable parameters")
        dev = self.params[0].device
        self.numel = sum(p.numel() for p in self.params)
        self.flat = torch.zeros(self.numel, device=dev, dtype=torch.float32)
        self.group = process_group
        off = 0
        for p in self.params:
            if p.dtype != torch.float32:
                raise TypeError("GradBucket expects fp32 trainable parameters")
            p.grad = self.flat[off:off + p.numel()].view_as(p)
            off += p.numel()

    @property
    def nbytes(self):
        return self.numel * 4

    def zero(self):
        """Replaces optimizer.zero_grad(): grads stay attached to the bucket."""
        self.flat.zero_()

    def check_attached(self):
        base = self.flat.untyped_storage().data_ptr()
        for p in self.params:
            if p.grad is None or p.grad.untyped_storage().data_ptr() != base:
                raise RuntimeError("a .grad was re-allocated outside the bucket (use bucket.zero(), not zero_grad(set_to_none=True))")

    def all_reduce_mean(self):
        """Average the bucket over the process group (no-op for a single process)."""
        if not dist.is_available() or not dist.is_initialized():
            return
        ws = dist.get_world_size(self.group)
        if ws == 1:
            return
        dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group)
        self.flat.mul_(1.0 / ws)
