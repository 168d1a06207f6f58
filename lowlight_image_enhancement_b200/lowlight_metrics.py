"""Metric adapters — drop-in for `NAFNet_base/basicsr/metrics/lowlight_metrics.py:211-272`
(`getattr(basicsr.metrics, metric_type)(pred, gt, **opt)` in image_restoration_model.py:424-428).

Same names, `(pred, target, **kw) -> float`, same argument-order swap towards the underlying functions
and the same clamp to [0,1] before the colour metrics (folded into the kernel's load instead of two
extra elementwise passes).  `lpips_distance` stays with the reference (pretrained network; out of scope).
"""
from __future__ import annotations

from typing import Literal

import torch

from .metrics.color_error import deltaE2000_summary, edge_deltaE2000_stats
from .metrics.psnr import calculate_psnr
from .metrics.ssim import calculate_ssim

__all__ = ["linear_psnr", "linear_ssim", "deltae2000_mean", "deltae2000_p95", "edge_deltae2000_mean"]


def linear_psnr(pred: torch.Tensor, target: torch.Tensor, *, data_range: float = 1.0) -> float:
    return calculate_psnr(target.to(dtype=torch.float32), pred.to(dtype=torch.float32), data_range=data_range)


def linear_ssim(pred: torch.Tensor, target: torch.Tensor, *, data_range: float = 1.0) -> float:
    return calculate_ssim(target.to(dtype=torch.float32), pred.to(dtype=torch.float32), data_range=data_range)


def deltae2000_mean(pred: torch.Tensor, target: torch.Tensor, *,
                    whitepoint: Literal["D65-2", "D50-2"] = "D65-2") -> float:
    # the reference also computes (and discards) the p95 through a full sort; only the sum is needed
    return deltaE2000_summary(pred, target, whitepoint=whitepoint, percentiles=(), clamp01=True)["mean"]


def deltae2000_p95(pred: torch.Tensor, target: torch.Tensor, *,
                   whitepoint: Literal["D65-2", "D50-2"] = "D65-2") -> float:
    return deltaE2000_summary(pred, target, whitepoint=whitepoint, percentiles=(95.0,), clamp01=True)["p95"]


def edge_deltae2000_mean(pred: torch.Tensor, target: torch.Tensor, *,
                         whitepoint: Literal["D65-2", "D50-2"] = "D65-2", q: float = 0.85) -> float:
    # the reference also computes (and discards) the masked p95 through a second full sort
    return edge_deltaE2000_stats(pred, target, whitepoint=whitepoint, q=q, clamp01=True, want_p95=False)["mean"]
