"""Whole-step CUDA-graph capture.

One fine-tuning step of the reference (`train_one` body, few_shot/vision_benchmark/evaluation/kadaptation_clip.py:308-314:
zero grads -> forward -> criterion -> backward -> [all-reduce] -> optimizer step) is ~230 kernel launches of this library
plus a few dozen small torch kernels.  At ViT-B/32 shapes or small per-GPU batches the step is launch-bound, so the step is
captured ONCE into a CUDA graph and replayed: the Python / ctypes / autograd work of a step disappears from the steady state.

Everything the step touches is static by construction: the engine's workspaces and frozen-weight operands are persistent,
TMA descriptors are encoded on the host at capture time and baked into the kernel parameters, gradients live in the
`GradBucket`'s flat buffer, and tensors created inside the step come from the graph's private pool.  Inputs are copied
into static device buffers before each replay; the loss is read from a static scalar.

Caller contract (torch's own rule for capturing a backward pass): if the model has already been run eagerly, that must
have happened on a NON-default stream (`torch.cuda.set_stream(torch.cuda.Stream())` at program start, as bench.py does) and
no tensor of an earlier step may still hold its autograd graph — autograd binds each parameter's AccumulateGrad node to the
stream of its first use and a node bound to the legacy default stream invalidates the capture.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


class GraphedTrainStep:
    def __init__(self, model, bucket, optimizer, images, labels, criterion=None, warmup=3):
        """`model(images) -> logits`; `bucket` = peft_b200.dist.GradBucket over the trainables (`optimizer` = a torch
        optimizer) or a peft_b200.dist.FusedSGDBucket (`optimizer` = None: its step() is the optimiser); `images` / `labels` are
        example CUDA tensors fixing shape and dtype (their contents are used for the warm-up steps)."""
        if not images.is_cuda:
            raise RuntimeError("GraphedTrainStep needs CUDA tensors (no CPU fallback)")
        self.model, self.bucket, self.optimizer = model, bucket, optimizer
        self.criterion = criterion or F.cross_entropy
        self.images = images.clone()
        self.labels = labels.clone()
        self.loss = None
        self.logits = None
        # warm-up on a side stream (torch's capture recipe): initialises optimizer state, engine workspaces, the library's
        # one-time cudaFuncSetAttribute calls and the TMA-descriptor cache, so that capture sees no such side effects
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(max(1, warmup)):
                self._step()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        self.bucket.check_attached()
        from . import _lib
        n0 = _lib.launch_count()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.loss, self.logits = self._step()
        self.launches_per_step = _lib.launch_count() - n0      # kernels of libpeftb200 inside one replay
        self.bucket.check_attached()

    def _step(self):
        self.bucket.zero()
        out = self.model(self.images)
        loss = self.criterion(out, self.labels)
        loss.backward()
        if self.optimizer is None:
            self.bucket.step()                  # FusedSGDBucket: all-reduce + SGD + zeroing in one kernel
        else:
            self.bucket.all_reduce_mean()
            self.optimizer.step()
        return loss.detach(), out.detach()

    def release(self):
        """Destroys the captured graph (and the NCCL kernels recorded in it).  Call before
        torch.distributed.destroy_process_group(): a live graph holding collective nodes can block the teardown."""
        if self.graph is not None:
            self.graph.reset()
            self.graph = None

    def __call__(self, images=None, labels=None):
        """Copies the batch into the static buffers (device-to-device or host-to-device, on the current stream) and replays
        the step.  Returns the static loss tensor (read it with .item() only when needed)."""
        if images is not None:
            self.images.copy_(images, non_blocking=True)
        if labels is not None:
            self.labels.copy_(labels, non_blocking=True)
        self.graph.replay()
        return self.loss
