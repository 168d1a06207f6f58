"""Whole-tensor PSNR — drop-in for the reference `metrics/psnr.py` (calculate_psnr, psnr.py:39-88).

Both inputs are promoted to float64 inside the kernel, the squared error is averaged over EVERY element
(batch included) and `inf` is returned when the MSE is within 1e-12 of zero.  Returns a Python float
(one device->host read, as in the reference's `.item()`).
"""
from __future__ import annotations

import torch

from .. import _functional as Fn
from .. import _native as nat


def calculate_psnr(img_true: torch.Tensor, img_pred: torch.Tensor, data_range: float) -> float:
    if img_true.shape != img_pred.shape:
        raise ValueError(f"Input shapes must match exactly, got {img_true.shape=} and {img_pred.shape=}.")
    if data_range <= 0:
        raise ValueError(f"`data_range` must be positive, received {data_range}.")
    dev = nat.cuda_device_for(img_true, img_pred)
    work = torch.float64 if torch.float64 in (img_true.dtype, img_pred.dtype) else torch.float32
    a = nat.stage(img_true.detach(), dev, work)
    b = nat.stage(img_pred.detach(), dev, work)
    numel = a.numel()
    if numel == 0:
        return float("nan")
    table, _, _ = Fn.psnr_stats(a, b, 1, 1, numel, cast_first=True, clamp_hi=None, data_range=float(data_range),
                                eps=1e-12)
    return float(table[0, 4].item())
