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


def kind_major(named_params):
    """[(name, parameter)] -> parameters ordered so that the same tensor of consecutive layers is adjacent in the bucket
    (`blocks.0.attn.W_left1, blocks.1.attn.W_left1, ..., blocks.0.attn.W_right1, ...`), kinds in order of first appearance.

    With this order a bucket (and, for FusedSGDBucket, the flat parameter buffer) holds `[depth, ...]` of every per-layer adapter
    tensor contiguously, which lets `timm_vit.stack_params` hand the tower a zero-copy stacked VIEW in forward and add the stacked
    gradient to the bucket with ONE kernel in backward, instead of a cat kernel + one AccumulateGrad add per layer and tensor
    (48 of the ~150 small launches of a KAdaptation step).  Any other order stays correct: stack_params falls back to torch.stack."""
    import re
    keyed, first = [], {}
    for idx, (n, q) in enumerate(named_params):
        m = re.search(r"\.(\d+)\.", n)
        kind = n[:m.start()] + ".#." + n[m.end():] if m else n
        first.setdefault(kind, idx)
        keyed.append((first[kind], int(m.group(1)) if m else 0, idx, q))
    keyed.sort(key=lambda k: k[:3])
    return [k[3] for k in keyed]


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


class FusedSGDBucket(GradBucket):
    """GradBucket + flat parameter / momentum replicas + ONE kernel per step for `all-reduce(mean) -> SGD-momentum update ->
    zero the bucket` ("fuse SGD-momentum into the all-reduce").

    replaces, in the data-parallel step: `bucket.all_reduce_mean()` (an NCCL kernel + a scaling kernel), `torch.optim.SGD.step()`
    (multi_tensor_apply kernels) and `bucket.zero()` of the next step.  Optimiser semantics = torch.optim.SGD(lr, momentum,
    dampening=0, weight_decay, nesterov=False) as built by few_shot/vision_benchmark/optim/build.py:120-159.

    * every trainable parameter becomes a view of one flat fp32 buffer (`p.data` is re-pointed; names, shapes, state_dict are
      unchanged), its `.grad` a view of the flat bucket, and the momentum lives in a third flat buffer;
    * with world size > 1 the bucket is allocated in torch symmetric memory (NVLink peer mapping): the kernel
      (csrc/allreduce_sgd_sm100.cuh) reads every peer's bucket directly and synchronises through the signal pads, so no NCCL
      call is on the step path; all ranks compute bit-identical parameters;
    * `lr` is a device scalar: `set_lr()` changes it without re-capturing a CUDA graph of the step.
    CPU tensors (the gloo tests of the host logic) take the unfused torch path with the same arithmetic.
    """

    MAX_BLOCKS = 32                     # csrc/allreduce_sgd_sm100.cuh kArMaxBlocks

    def __init__(self, params, lr, momentum=0.9, weight_decay=0.0, process_group=None):
        params = [p for p in params if p.requires_grad]
        if not params:
            raise ValueError("FusedSGDBucket: no trainable parameters")
        dev = params[0].device
        self.world = dist.get_world_size(process_group) if (dist.is_available() and dist.is_initialized()) else 1
        self.rank = dist.get_rank(process_group) if self.world > 1 else 0
        n = sum(p.numel() for p in params)
        self.padded = (n + 3) // 4 * 4
        self._symm = None
        flat = None
        if dev.type == "cuda" and self.world > 1:
            import torch.distributed._symmetric_memory as symm_mem
            group = process_group if process_group is not None else dist.group.WORLD
            need = 2 * self.MAX_BLOCKS * self.world          # 2 barriers x blocks x ranks, uint32 slots
            if symm_mem.get_signal_pad_size() // 4 < 2 * need:
                symm_mem.set_signal_pad_size(8 * need)        # bytes; must precede the first symmetric allocation
            flat = symm_mem.empty(self.padded, dtype=torch.float32, device=dev)
            flat.zero_()
            self._symm = symm_mem.rendezvous(flat, group)
            self.pad_slots = symm_mem.get_signal_pad_size() // 4
            if self.pad_slots < 2 * need:
                raise RuntimeError(f"symmetric-memory signal pad too small ({self.pad_slots} slots, need {2 * need})")
            # the first slots of the pad belong to torch's own barrier / put_signal helpers: this kernel uses the tail
            self.pad_base = self.pad_slots - need
        # GradBucket.__init__ allocates its own flat buffer: build the views by hand on the (possibly symmetric) buffer
        self.params = params
        self.numel = n
        self.flat = flat if flat is not None else torch.zeros(self.padded, device=dev, dtype=torch.float32)
        self.group = process_group
        self.flat_param = torch.zeros(self.padded, device=dev, dtype=torch.float32)
        self.flat_mom = torch.zeros(self.padded, device=dev, dtype=torch.float32)
        self.lr_dev = torch.full((1,), float(lr), device=dev, dtype=torch.float32)
        self.momentum, self.weight_decay = float(momentum), float(weight_decay)
        off = 0
        with torch.no_grad():
            for p in params:
                if p.dtype != torch.float32:
                    raise TypeError("FusedSGDBucket expects fp32 trainable parameters")
                k = p.numel()
                self.flat_param[off:off + k].copy_(p.detach().reshape(-1))
                p.data = self.flat_param[off:off + k].view_as(p)
                p.grad = self.flat[off:off + k].view_as(p)
                off += k
        if self._symm is not None:
            self._symm.barrier()        # every rank's bucket is zeroed and mapped before the first step

    @property
    def nbytes(self):
        return self.numel * 4

    def set_lr(self, lr):
        self.lr_dev.fill_(float(lr))

    def step(self):
        """all-reduce(mean) + SGD-momentum update + zero the bucket, in one kernel (CUDA) — call right after backward()."""
        if self.flat.is_cuda:
            from . import _lib
            sm = self._symm
            _lib.check(_lib.load().peft_allreduce_sgd(
                sm.buffer_ptrs_dev if sm is not None else 0, sm.signal_pad_ptrs_dev if sm is not None else 0, self.rank, self.world,
                self.flat.data_ptr(), self.flat_param.data_ptr(), self.flat_mom.data_ptr(), self.padded, self.lr_dev.data_ptr(),
                self.momentum, self.weight_decay, 1, getattr(self, "pad_base", 0), getattr(self, "pad_slots", 0),
                torch.cuda.current_stream().cuda_stream), "peft_allreduce_sgd")
            self._stepped = True
            return
        # CPU tensors (gloo tests): the same arithmetic, unfused
        self.all_reduce_mean()
        with torch.no_grad():
            g = self.flat + self.weight_decay * self.flat_param
            self.flat_mom.mul_(self.momentum).add_(g)
            self.flat_param.sub_(self.lr_dev * self.flat_mom)
            self.flat.zero_()
        self._stepped = True

    def zero(self, force=False):
        """step() leaves the bucket zeroed, so in the `backward -> step` rhythm of train_one this is a no-op after the first
        step (one memset launch less per step).  `force=True` clears gradients accumulated without a step()."""
        if force or not getattr(self, "_stepped", False):
            self.flat.zero_()


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
