"""CPU oracle of the CLIP input pipeline — TEST INFRASTRUCTURE (only tests/ and the golden generator may import it).

Restates `_transform(n_px)` of few_shot/vision_benchmark/evaluation/clip_load.py:80-87, i.e. torchvision's
Resize(n_px, BICUBIC) -> CenterCrop(n_px) -> ToTensor -> Normalize applied to a PIL image.  The arithmetic lives in
third-party code absent from /root/reference: Pillow (pinned `Pillow` via torchvision in few_shot/requirements.txt; here 12.2.0)
`src/libImaging/Resample.c` — bicubic_filter (a = -0.5), precompute_coeffs, normalize_coeffs_8bpc (PRECISION_BITS = 22),
ImagingResampleHorizontal_8bpc / Vertical_8bpc (horizontal pass first, intermediate rounded to uint8).  Pinned by
tests/golden/reference_outputs.pt["preprocess_*"], which tests/golden/make_golden.py produces with the REAL torchvision-on-PIL
pipeline; tests/test_oracle_golden.py requires bit-equality.
"""
import math

import numpy as np
import torch

MEAN = (0.48145466, 0.4578275, 0.40821073)
STD = (0.26862954, 0.26130258, 0.27577711)


def _filter(x):
    x = abs(x)
    a = -0.5
    if x < 1.0:
        return ((a + 2.0) * x - (a + 3.0)) * x * x + 1
    if x < 2.0:
        return (((x - 5) * x + 8) * x - 4) * a
    return 0.0


def _coeffs(in_size, out_size):
    scale = filterscale = in_size / out_size
    if filterscale < 1.0:
        filterscale = 1.0
    support = 2.0 * filterscale
    out = []
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        ss = 1.0 / filterscale
        xmin = int(center - support + 0.5)
        if xmin < 0:
            xmin = 0
        xmax = int(center + support + 0.5)
        if xmax > in_size:
            xmax = in_size
        xmax -= xmin
        k = [_filter((x + xmin - center + 0.5) * ss) for x in range(xmax)]
        ww = sum(k)
        if ww != 0.0:
            k = [v / ww for v in k]
        ki = [int(-0.5 + v * (1 << 22)) if v < 0 else int(0.5 + v * (1 << 22)) for v in k]
        out.append((xmin, np.array(ki, dtype=np.int64)))
    return out


def _resample_axis0(img, out_size):
    """img uint8 [N, ...]: resample along axis 0 with Pillow's 8-bit arithmetic."""
    co = _coeffs(img.shape[0], out_size)
    res = np.empty((out_size,) + img.shape[1:], dtype=np.uint8)
    src = img.astype(np.int64)
    for i, (xmin, k) in enumerate(co):
        ss = (1 << 21) + np.tensordot(k, src[xmin:xmin + len(k)], axes=(0, 0))
        res[i] = np.clip(ss >> 22, 0, 255).astype(np.uint8)
    return res


def clip_transform(image_u8: np.ndarray, n_px: int) -> torch.Tensor:
    """image uint8 [H, W, 3] -> fp32 [3, n_px, n_px]"""
    H, W, _ = image_u8.shape
    if H <= W:
        Ho, Wo = n_px, int(n_px * W / H)
    else:
        Ho, Wo = int(n_px * H / W), n_px
    x = _resample_axis0(np.ascontiguousarray(image_u8.transpose(1, 0, 2)), Wo).transpose(1, 0, 2)      # horizontal pass first
    x = _resample_axis0(np.ascontiguousarray(x), Ho)
    top, left = int(round((Ho - n_px) / 2.0)), int(round((Wo - n_px) / 2.0))
    x = x[top:top + n_px, left:left + n_px]
    t = torch.from_numpy(np.ascontiguousarray(x.transpose(2, 0, 1))).to(torch.float32).div(255)        # ToTensor
    mean = torch.tensor(MEAN, dtype=torch.float32).view(3, 1, 1)
    std = torch.tensor(STD, dtype=torch.float32).view(3, 1, 1)
    return (t - mean) / std                                                                             # Normalize
