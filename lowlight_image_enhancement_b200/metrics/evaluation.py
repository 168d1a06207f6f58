"""Evaluation helpers — drop-in for the tensor helpers of the reference `metrics/evaluation.py`.

`_linear_to_srgb` (evaluation.py:27-38), `_norm_to_lpips_range` (:50-52), `_to_3ch` (:41-47),
`_compute_channelwise_psnr_sse` (:55-62) and `_sobel_magnitude` (:65-86, the manual-Sobel branch: kornia is not a
dependency here) keep their names and argument meaning and run on the sm_100a kernels; `srgb_and_lpips_inputs`
is the fused form of the two calls `compute_metrics` makes back to back (:175-178): one read of the linear image,
both outputs.  `compute_metrics` itself (a model / dataloader / torchmetrics / lpips driver, and upstream it imports
two classes its sibling modules no longer define) is out of scope: the sharded eval driver of this
package is `lowlight_image_enhancement_b200.eval`.
"""
from __future__ import annotations

from typing import Tuple

import torch
from torch import Tensor

from .. import _functional as Fn
from .. import _native as nat

__all__ = ["srgb_and_lpips_inputs"]


def _linear_to_srgb(x: Tensor) -> Tensor:
    """Linear RGB -> sRGB (IEC 61966-2-1) after clamp to [0, 1]; shape and device preserved."""
    srgb, _ = Fn.linear_to_srgb(x, want_srgb=True, want_lpips=False)
    return srgb


def srgb_and_lpips_inputs(x_linear: Tensor) -> Tuple[Tensor, Tensor]:
    """``(_linear_to_srgb(x), _norm_to_lpips_range(_linear_to_srgb(x)))`` from one pass over ``x``."""
    return Fn.linear_to_srgb(x_linear, want_srgb=True, want_lpips=True)


def _to_3ch(x: Tensor) -> Tensor:
    if x.size(1) == 3:
        return x
    if x.size(1) == 1:
        return x.repeat(1, 3, 1, 1)
    return x[:, :3, ...]


def _norm_to_lpips_range(x: Tensor) -> Tensor:
    """[0, 1] -> [-1, 1].  Pure scale-and-shift: stays a torch op (use `srgb_and_lpips_inputs` for the fused form)."""
    return x.mul(2.0).sub(1.0)


def _compute_channelwise_psnr_sse(pred: Tensor, target: Tensor) -> Tensor:
    """Per-channel sum of squared errors over N, H, W -> [C] (float32 like the reference's fp32 `.sum`; the kernel
    accumulates in float64, so it is the more accurate of the two)."""
    if pred.shape != target.shape or pred.dim() != 4:
        raise ValueError(f"expected two [N, C, H, W] tensors of one shape, got {tuple(pred.shape)} / {tuple(target.shape)}")
    dev = nat.cuda_device_for(pred, target)
    a = nat.stage(pred.detach(), dev, torch.float32)
    b = nat.stage(target.detach(), dev, torch.float32)
    n, c, h, w = a.shape
    table, _, _ = Fn.psnr_stats(a, b, n, c, h * w, cast_first=False, clamp_hi=None, data_range=1.0, eps=1e-12)
    return (table[:n, :c].sum(dim=0) * float(h * w)).to(pred.dtype)  # columns [0, c) hold the per-channel MSE


def _sobel_magnitude(gray: Tensor) -> Tensor:
    """[N, 1, H, W] -> sqrt(gx^2 + gy^2 + 1e-12), zero padding."""
    dev = nat.cuda_device_for(gray)
    return Fn.sobel_magnitude(nat.stage(gray.detach(), dev, torch.float32))
