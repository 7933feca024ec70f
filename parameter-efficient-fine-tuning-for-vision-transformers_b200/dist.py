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


class GlobalBatchNorm1d(torch.nn.Module):
    """BatchNorm1d(affine=False) whose TRAIN-mode statistics are taken over the global batch (all ranks), so that an N-rank
    step equals the 1-rank step on the concatenated batch (SURVEY §8e "BN head").  The reference head is
    `nn.BatchNorm1d(E, affine=False)` (few_shot/vision_benchmark/evaluation/kadaptation_clip.py:122-125) on a single GPU;
    this module is the optional multi-GPU counterpart (`Classifier(..., sync_bn=True)`), differentiable through the
    exchange (`torch.distributed.nn.functional.all_reduce` of [sum x, sum x^2, count]: 2E + 1 floats).  Works on NCCL and on
    gloo (CPU tests).  Running statistics follow nn.BatchNorm1d (momentum 0.1, unbiased running variance)."""

    def __init__(self, num_features, eps=1e-5, momentum=0.1, process_group=None):
        super().__init__()
        self.num_features, self.eps, self.momentum, self.group = num_features, eps, momentum, process_group
        self.register_buffer("running_mean", torch.zeros(num_features))
        self.register_buffer("running_var", torch.ones(num_features))
        self.register_buffer("num_batches_tracked", torch.tensor(0, dtype=torch.long))

    def forward(self, x):
        if not self.training:
            return (x - self.running_mean) / torch.sqrt(self.running_var + self.eps)
        n_local = x.shape[0]
        stats = torch.cat([x.sum(0), (x * x).sum(0), x.new_full((1,), float(n_local))])
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1:
            import torch.distributed.nn.functional as dfn
            stats = dfn.all_reduce(stats, op=dist.ReduceOp.SUM, group=self.group)
        E = self.num_features
        n = stats[2 * E]
        mean = stats[:E] / n
        var = stats[E:2 * E] / n - mean * mean
        with torch.no_grad():
            self.running_mean.mul_(1 - self.momentum).add_(self.momentum * mean)
            self.running_var.mul_(1 - self.momentum).add_(self.momentum * var * n / (n - 1).clamp(min=1))
            self.num_batches_tracked += 1
        return (x - mean) / torch.sqrt(var + self.eps)
