"""TEST INFRASTRUCTURE ONLY — CPU restatement of the BasicSR pixel losses.

Follows `NAFNet_base/basicsr/models/losses/losses.py:18-31` (element-wise terms), `:110-138` (PSNRLoss) and
`loss_util.py:9-60` (weighting / reduction).  Pinned against the unmodified reference by
`tests/golden/make_golden_basicsr.py` -> `tests/golden/basicsr_scalars.json`.  Only tests/, smoke() and the
benchmark's CPU legs may import this module.
"""
from __future__ import annotations

import math

import torch


def _reduce(loss, weight, reduction):
    """loss_util.py:31-60."""
    if weight is not None:
        assert weight.dim() == loss.dim()
        assert weight.size(1) == 1 or weight.size(1) == loss.size(1)
        loss = loss * weight
    if weight is None or reduction == "sum":
        if reduction == "none":
            return loss
        return loss.mean() if reduction == "mean" else loss.sum()
    if reduction == "mean":
        denom = weight.sum() if weight.size(1) > 1 else weight.sum() * loss.size(1)
        return loss.sum() / denom
    return loss


def l1_loss(pred, target, weight=None, reduction="mean"):
    """losses.py:18-20."""
    return _reduce((pred - target).abs(), weight, reduction)


def mse_loss(pred, target, weight=None, reduction="mean"):
    """losses.py:23-25."""
    return _reduce((pred - target) ** 2, weight, reduction)


def charbonnier_loss(pred, target, weight=None, reduction="mean", eps=1e-12):
    """losses.py:28-31 (eps is added un-squared)."""
    d = pred - target
    return _reduce(torch.sqrt(d * d + eps), weight, reduction)


def psnr_loss(pred, target, loss_weight=1.0, toY=False):
    """losses.py:112-138."""
    if toY:
        coef = torch.tensor([65.481, 128.553, 24.966], dtype=pred.dtype, device=pred.device).reshape(1, 3, 1, 1)
        pred = ((pred * coef).sum(dim=1).unsqueeze(1) + 16.0) / 255.0
        target = ((target * coef).sum(dim=1).unsqueeze(1) + 16.0) / 255.0
    scale = 10 / math.log(10)
    return loss_weight * scale * torch.log(((pred - target) ** 2).mean(dim=(1, 2, 3)) + 1e-8).mean()
