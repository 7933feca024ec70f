"""Intrinsic-dimension fine-tuning: drop-in for full_shot/main/intrinsic/fastfood.py (`FastfoodWrap`).

Same constructor, same single trainable parameter `V` [d], same random draws in the same order (so a run seeded like
the reference gets the same B, Pi, G), same forward contract: every wrapped parameter is deleted from its module and
re-set on each forward as  theta0 + Fastfood(V)  (fastfood.py:36-57,61-77).  The projection of ALL wrapped tensors
and its backward are two C-ABI calls (`peft_fastfood_fwd/_bwd`, csrc/fastfood_sm100.cuh) instead of
2*log2(LL) chunk/cat kernels per tensor; there is no torch fallback.

Layout in HBM (ViT-B/32 visual tower: 152 tensors, sum DD = 87.8 M, sum LL = 143.4 M):
  pi_lo uint16 [sum LL] (Pi mod wn, 287 MB), gg fp32 [sum LL] (574 MB), scratch fp32 [sum LL], bsign fp32 [T, wn],
  theta0 / theta / dtheta fp32 flat with every tensor starting on a 256-byte boundary.
"""
from __future__ import annotations

import math
import os

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib

_TILE = 8192
_ALIGN = 64      # elements: every tensor starts on a 128-byte line in the uint16 table, 256 bytes in the fp32 ones
_LAG = 640       # items between the two phases of a tensor (~2 grids of 296 CTAs)
_FF_DTYPE = np.dtype([("off_ll", "<i8"), ("off_dd", "<i8"), ("DD", "<i4"), ("ll", "<i4"), ("coef", "<f4"), ("reserved", "<i4")])


def fastfood_vars(DD, device="cpu"):
    """fastfood.py:124-148 — same draws, same order (torch global RNG for B and G, numpy global RNG for Pi)."""
    ll = int(np.ceil(np.log(DD) / np.log(2)))
    LL = 2 ** ll
    BB = torch.empty(LL, dtype=torch.float32).uniform_(0, 2).type(torch.LongTensor)
    BB = (BB * 2 - 1).type(torch.FloatTensor)
    Pi = torch.LongTensor(np.random.permutation(LL))
    GG = torch.empty(LL, dtype=torch.float32).normal_()
    divisor = torch.sqrt(LL * torch.sum(torch.pow(GG, 2)))
    return [BB, Pi, GG, divisor, LL]


_BWD_GROUP = max(1, min(256, int(os.environ.get("PEFT_FF_BWD_GROUP", "4"))))   # tiles per backward high item


def _schedule(phases, lag=_LAG):
    """Work list of one direction.  The kernels are persistent (occupancy x SMs CTAs take items round-robin) and an item
    only waits on lower-numbered ones, so any order with producers first is valid; the second phase of a tensor is issued
    `lag` items (about two full grids) after its first phase, so that a CTA reaching it normally finds the counter already
    complete instead of spinning through the tail of the first phase.  The intermediate of the tensors in flight
    (lag x 32 KB + the tensor itself) stays far below the 126 MB L2."""
    out, pending = [], []
    for first, second in phases:
        while pending and pending[0][0] <= len(out):
            out += pending.pop(0)[1]
        out += first
        if second:
            pending.append((len(out) + lag, second))
    for _, second in pending:
        out += second
    return out


class FastfoodPlan:
    """Device-resident tables for a set of wrapped tensors; built once from the per-tensor (B, Pi, G, divisor, LL)."""

    def __init__(self, shapes, params, d, device):
        if not torch.cuda.is_available():
            raise RuntimeError("peft_b200 Fastfood needs a CUDA device (no CPU fallback)")
        _lib.load()
        self.d = int(d)
        self.wn = 1 << max(0, math.ceil(math.log2(max(1, self.d))))
        if self.wn > 4096:
            raise NotImplementedError("Fastfood kernels support intrinsic dimension <= 4096")
        self.device = torch.device(device)
        T = len(shapes)
        self.T = T
        desc = np.zeros(T, dtype=_FF_DTYPE)
        off_ll = off_dd = 0
        self.offsets, self.numels, self.shapes = [], [], [tuple(s) for s in shapes]
        fwd_ph, bwd_ph = [], []          # per tensor: (first-phase items, second-phase items that wait for them)
        sum_ll = 0
        bsign = torch.zeros(T, self.wn, dtype=torch.float32)
        pi_parts, gg_parts = [], []
        for t, (shape, (BB, Pi, GG, divisor, LL)) in enumerate(zip(shapes, params)):
            DD = int(np.prod(shape)) if len(shape) else 1
            ll = int(LL).bit_length() - 1
            assert (1 << ll) == LL and LL >= DD
            if ll > 26:
                raise NotImplementedError("Fastfood kernels support tensors up to 2^26 elements")
            if LL < self.wn and DD < self.d:
                # the reference pads V to LL with a NEGATIVE pad here (truncation); keep that behaviour
                pass
            n1 = min(LL, self.wn)
            coef = 1.0 / (float(divisor) * math.sqrt(float(DD) / LL))
            desc[t] = (off_ll, off_dd, DD, ll, coef, 0)
            bsign[t, :n1] = BB[:n1]
            pad = (-LL) % _ALIGN                                  # tensor starts stay 256-byte aligned (cp.async staging)
            pi_parts.append(F.pad((Pi & (n1 - 1)).to(torch.int16), (0, pad)))   # n1 <= 4096 fits; uint16 on device
            gg_parts.append(F.pad(GG.float(), (0, pad)))
            hi = [(t, k | (1 << 30)) for k in range(LL // _TILE)] if ll > 13 else []
            lo = [(t, k) for k in range((DD + _TILE - 1) // _TILE)]
            fwd_ph.append((hi, lo) if hi else (lo, []))   # forward: high tiles, then the low blocks that wait for them
            # backward: the other way round; its high items are RUNS of `grp` consecutive tiles (count - 1 in bits 20..27) that
            # share one set of bins and one round of global atomics (csrc/fastfood_sm100.cuh: ff_bwd_kernel)
            grp = _BWD_GROUP
            nt = LL // _TILE
            hi_b = [(t, k | ((min(grp, nt - k) - 1) << 20) | (1 << 30)) for k in range(0, nt, grp)] if ll > 13 else []
            bwd_ph.append((lo, hi_b) if hi_b else (lo, []))
            self.offsets.append(off_dd)
            self.numels.append(DD)
            off_ll += LL + pad
            sum_ll += LL
            off_dd += (DD + _ALIGN - 1) // _ALIGN * _ALIGN
        self.total_ll, self.total_dd, self.sum_ll = off_ll, off_dd, sum_ll
        fwd_items, bwd_items = _schedule(fwd_ph), _schedule(bwd_ph)
        dev = self.device
        self.desc = torch.from_numpy(desc.view(np.uint8).copy()).to(dev)
        self.pi_lo = torch.cat(pi_parts).to(dev)
        self.gg = torch.cat(gg_parts).to(dev)
        self.bsign = bsign.to(dev)
        mk = lambda items: torch.tensor(items, dtype=torch.int32).reshape(-1, 2).contiguous().to(dev)
        self.items_fwd, self.items_bwd, self.n_items, self.n_items_bwd = mk(fwd_items), mk(bwd_items), len(fwd_items), len(bwd_items)
        self.counters = torch.zeros(T + 1, dtype=torch.int32, device=dev)      # per-tensor completion counters + the backward's item ticket
        self.scratch = torch.empty(off_ll, dtype=torch.float32, device=dev)
        self.wu = torch.zeros(T, self.wn, dtype=torch.float32, device=dev)
        self.theta0 = torch.zeros(off_dd, dtype=torch.float32, device=dev)
        self.theta = torch.zeros(off_dd, dtype=torch.float32, device=dev)
        self.dtheta = torch.zeros(off_dd, dtype=torch.float32, device=dev)

    # algorithmic HBM bytes of one forward / backward: 6 B per LL element + 8 / 4 B per DD element
    def algorithmic_bytes(self):
        dd = sum(self.numels)
        return 6 * self.sum_ll + 8 * dd, 6 * self.sum_ll + 4 * dd

    def set_theta0(self, tensors):
        for off, n, t in zip(self.offsets, self.numels, tensors):
            self.theta0[off:off + n].copy_(t.detach().reshape(-1).to(self.device, torch.float32))

    def views(self, flat):
        return [flat[off:off + n].view(shape) for off, n, shape in zip(self.offsets, self.numels, self.shapes)]

    def project(self, V):
        """theta (flat) = theta0 + Fastfood(V); returns the persistent flat buffer."""
        if not V.is_cuda or V.dtype != torch.float32 or V.numel() != self.d or not V.is_contiguous():
            raise RuntimeError("Fastfood: V must be a contiguous CUDA fp32 vector of the intrinsic dimension")
        s = torch.cuda.current_stream().cuda_stream
        _lib.check(_lib.load().peft_fastfood_fwd(V.data_ptr(), self.d, self.wn, self.desc.data_ptr(), self.T,
                                                 self.items_fwd.data_ptr(), self.n_items,
                                                 self.pi_lo.data_ptr(), self.gg.data_ptr(), self.bsign.data_ptr(),
                                                 self.theta0.data_ptr(), self.theta.data_ptr(), self.wu.data_ptr(),
                                                 self.scratch.data_ptr(), self.counters.data_ptr(), s), "peft_fastfood_fwd")
        return self.theta

    def project_backward(self, dtheta_flat):
        """dV [d] from the flat gradient w.r.t. theta."""
        dV = torch.zeros(self.d, dtype=torch.float32, device=self.device)
        s = torch.cuda.current_stream().cuda_stream
        _lib.check(_lib.load().peft_fastfood_bwd(dtheta_flat.data_ptr(), self.d, self.wn, self.desc.data_ptr(), self.T,
                                                 self.items_bwd.data_ptr(), self.n_items_bwd,
                                                 self.pi_lo.data_ptr(), self.gg.data_ptr(), self.bsign.data_ptr(),
                                                 self.wu.data_ptr(), self.scratch.data_ptr(), self.counters.data_ptr(),
                                                 dV.data_ptr(), s), "peft_fastfood_bwd")
        return dV


class _FastfoodFn(torch.autograd.Function):
    """V -> tuple of per-tensor views of the flat theta buffer."""

    @staticmethod
    def forward(ctx, V, plan: FastfoodPlan):
        ctx.plan = plan
        flat = plan.project(V.contiguous())
        return tuple(plan.views(flat))

    @staticmethod
    def backward(ctx, *grads):
        plan = ctx.plan
        dst = plan.views(plan.dtheta)
        for g, d in zip(grads, dst):
            if g is None:
                d.zero_()
            else:
                d.copy_(g)
        return plan.project_backward(plan.dtheta), None


class FastfoodWrap(nn.Module):
    def __init__(self, module, intrinsic_dimension, device=0):
        super().__init__()
        self.m = [module]                    # hidden from parameters(), as in the reference (:19)
        self.name_base_localname = []
        self.initial_value = dict()
        self.fastfood_params = {}
        dev = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        if dev.type != "cuda":
            raise RuntimeError("peft_b200 FastfoodWrap runs on CUDA only (no CPU fallback)")
        self.V = nn.Parameter(torch.zeros(intrinsic_dimension, device=dev))
        shapes, params = [], []
        for name, param in module.named_parameters():
            if param.requires_grad:
                self.initial_value[name] = v0 = param.clone().detach().requires_grad_(False).to(dev)
                DD = int(np.prod(v0.size()))
                self.fastfood_params[name] = fp = fastfood_vars(DD)
                shapes.append(tuple(v0.size()))
                params.append(fp)
                base, localname = module, name
                while "." in localname:
                    prefix, localname = localname.split(".", 1)
                    base = base.__getattr__(prefix)
                self.name_base_localname.append((name, base, localname))
        for name, base, localname in self.name_base_localname:
            delattr(base, localname)
        self.plan = FastfoodPlan(shapes, params, intrinsic_dimension, dev)
        self.plan.set_theta0([self.initial_value[n] for n, _, _ in self.name_base_localname])

    def project(self):
        """The wrapped tensors theta0 + Fastfood(V), in wrap order (differentiable w.r.t. V)."""
        return _FastfoodFn.apply(self.V, self.plan)

    def forward(self, x):
        thetas = self.project()
        for (name, base, localname), th in zip(self.name_base_localname, thetas):
            setattr(base, localname, th)
        return self.m[0](x)
