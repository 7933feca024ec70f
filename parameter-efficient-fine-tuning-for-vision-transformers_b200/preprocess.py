"""GPU input pipeline of the CLIP trainers: `_transform(n_px)` of
few_shot/vision_benchmark/evaluation/clip_load.py:80-87 —

    Resize(n_px, interpolation=BICUBIC) -> CenterCrop(n_px) -> convert("RGB") -> ToTensor() -> Normalize(mean, std)

on a decoded uint8 RGB image [H, W, 3] that already lives in device memory (JPEG decoding stays outside).  The reference runs
this on PIL images, i.e. through Pillow's two-pass 8-bit resampler: antialiased bicubic (a = -0.5, support 2 x scale), the
coefficients of every output pixel normalised and rounded to 22-bit fixed point, a horizontal pass whose result is rounded and
clipped to uint8, then a vertical pass.  The kernels (csrc/preprocess_sm100.cuh) do exactly that integer arithmetic, so the
result is bit-identical to the reference pipeline, not merely close; this module only builds the coefficient tables (host
logic, numpy double arithmetic as in Pillow's precompute_coeffs / normalize_coeffs_8bpc) and caches them per image size.
"""
from __future__ import annotations

import math

import numpy as np
import torch

from . import _lib

CLIP_MEAN = (0.48145466, 0.4578275, 0.40821073)
CLIP_STD = (0.26862954, 0.26130258, 0.27577711)
_PRECISION_BITS = 32 - 8 - 2


def _bicubic(x):
    a = -0.5
    x = np.abs(x)
    return np.where(x < 1.0, ((a + 2.0) * x - (a + 3.0)) * x * x + 1.0, np.where(x < 2.0, (((x - 5.0) * x + 8.0) * x - 4.0) * a, 0.0))


def resample_coeffs(in_size: int, out_size: int):
    """(bounds int32 [out, 2] = (first input index, count), coefficients int32 [out, ksize]) of one resampling axis."""
    scale = in_size / out_size
    filterscale = max(scale, 1.0)
    support = 2.0 * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    bounds = np.zeros((out_size, 2), dtype=np.int32)
    kk = np.zeros((out_size, ksize), dtype=np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        xmin = max(int(center - support + 0.5), 0)
        xmax = min(int(center + support + 0.5), in_size) - xmin
        w = _bicubic((np.arange(xmax, dtype=np.float64) + xmin - center + 0.5) * ss)
        ww = w.sum()
        if ww != 0.0:
            w = w / ww
        fixed = np.where(w < 0, -0.5 + w * (1 << _PRECISION_BITS), 0.5 + w * (1 << _PRECISION_BITS))
        kk[xx, :xmax] = np.trunc(fixed).astype(np.int64).astype(np.int32)
        bounds[xx] = (xmin, xmax)
    return bounds, kk


def resized_size(h: int, w: int, n_px: int):
    """torchvision Resize(int): the shorter edge becomes n_px, the longer one int(n_px * long / short)."""
    if h <= w:
        return n_px, int(n_px * w / h)
    return int(n_px * h / w), n_px


class ClipPreprocess:
    """preprocess(image uint8 [H, W, 3] CUDA) -> fp32 [3, n_px, n_px], bit-identical to clip_load._transform(n_px) on the same pixels."""

    def __init__(self, n_px: int = 224, mean=CLIP_MEAN, std=CLIP_STD, device="cuda"):
        if not torch.cuda.is_available():
            raise RuntimeError("peft_b200 ClipPreprocess needs a CUDA device (no CPU fallback)")
        self.n_px, self.device = int(n_px), torch.device(device)
        self.mean = torch.tensor(mean, dtype=torch.float32, device=self.device)
        self.std = torch.tensor(std, dtype=torch.float32, device=self.device)
        self._tables = {}

    def _plan(self, H, W):
        key = (H, W)
        if key not in self._tables:
            n = self.n_px
            Ho, Wo = resized_size(H, W, n)
            hb, hk = resample_coeffs(W, Wo)
            vb, vk = resample_coeffs(H, Ho)
            top, left = int(round((Ho - n) / 2.0)), int(round((Wo - n) / 2.0))
            dev = self.device
            t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
            # rows of the input the vertical pass of the CROPPED output needs (Pillow also restricts its horizontal pass)
            r0 = int(vb[top, 0])
            r1 = int(vb[top + n - 1, 0] + vb[top + n - 1, 1])
            self._tables[key] = dict(Ho=Ho, Wo=Wo, top=top, left=left, hb=t(hb), hk=t(hk), vb=t(vb), vk=t(vk), r0=r0, r1=r1,
                                     tmp=torch.empty((r1 - r0, n, 3), dtype=torch.uint8, device=dev))
        return self._tables[key]

    def __call__(self, image_u8: torch.Tensor, out: torch.Tensor = None) -> torch.Tensor:
        if not image_u8.is_cuda or image_u8.dtype != torch.uint8 or image_u8.dim() != 3 or image_u8.shape[2] != 3 or not image_u8.is_contiguous():
            raise RuntimeError("ClipPreprocess: image must be a contiguous CUDA uint8 tensor [H, W, 3] (no CPU fallback)")
        H, W, _ = image_u8.shape
        p = self._plan(H, W)
        n = self.n_px
        if min(p["Ho"], p["Wo"]) < n:
            raise ValueError("image smaller than the crop after resizing")
        if out is None:
            out = torch.empty((3, n, n), dtype=torch.float32, device=self.device)
        _lib.check(_lib.load().peft_preprocess_clip(
            image_u8.data_ptr(), H, W, p["hb"].data_ptr(), p["hk"].data_ptr(), p["hk"].shape[1], p["vb"].data_ptr(), p["vk"].data_ptr(),
            p["vk"].shape[1], p["top"], p["left"], n, p["r0"], p["r1"], self.mean.data_ptr(), self.std.data_ptr(), p["tmp"].data_ptr(),
            out.data_ptr(), torch.cuda.current_stream().cuda_stream), "peft_preprocess_clip")
        return out

    def batch(self, images):
        """list of uint8 [H_i, W_i, 3] CUDA images (sizes may differ) -> fp32 [B, 3, n, n]"""
        out = torch.empty((len(images), 3, self.n_px, self.n_px), dtype=torch.float32, device=self.device)
        for i, im in enumerate(images):
            self(im, out[i])
        return out
