"""BasicSR pixel losses — drop-in for `NAFNet_base/basicsr/models/losses/losses.py:18-138` and the weighting /
reduction rules of `loss_util.py:9-95` (the losses of the non-NewBP baseline configs).

`l1_loss`, `mse_loss`, `charbonnier_loss` keep the `@weighted_loss` call form `(pred, target, weight=None,
reduction='mean', **kwargs)`; `L1Loss`, `MSELoss`, `CharbonnierLoss`, `PSNRLoss` keep their constructors.
For reduction 'mean' / 'sum' the value and the gradient come from ONE streaming pass (`llie_pixel_loss_grad`:
8 B/element read + 4 written, against ~6 elementwise passes of autograd); backward is a scaling by the upstream
scalar.  Reduction 'none' (an element-wise map with an element-wise upstream gradient) and planes whose H*W is not a
multiple of 4 are composed from stock device tensor ops.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
from torch.autograd.function import once_differentiable
import torch.nn as nn
from torch import Tensor

from . import _functional as Fn
from . import _native as nat

_reduction_modes = ["none", "mean", "sum"]
_MODES = {"l1": 0, "mse": 1, "charbonnier": 2}


def _elementwise(mode: str, pred: Tensor, target: Tensor, eps: float) -> Tensor:
    d = pred - target
    if mode == "l1":
        return d.abs()
    if mode == "mse":
        return d * d
    return torch.sqrt(d * d + eps)


class _PixelLossFn(torch.autograd.Function):
    """sum_e w_e psi(x_e - y_e) (float32 scalar) with the unit gradient stored by the same pass."""

    @staticmethod
    def forward(ctx, pred: Tensor, target: Tensor, weight: Optional[Tensor], mode: int, eps: float, scale: float):
        x = pred.detach().contiguous()
        y = target.detach().contiguous()
        dev = x.device
        n, c = x.shape[0], x.shape[1]
        plane = x.numel() // (n * c)
        sums = torch.empty(1, dtype=torch.float64, device=dev)
        flags = torch.empty(1, dtype=torch.int32, device=dev)
        grad = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        w = weight.detach().contiguous() if weight is not None else None
        per_c = 1 if (w is not None and w.shape[1] == c and c > 1) or (w is not None and c == 1) else 0
        with torch.cuda.device(dev):
            rc = nat.lib().llie_pixel_loss_grad(mode, nat.ptr(x), nat.ptr(y), nat.ptr(w), per_c, None, n, c, plane,
                                                float(eps), float(scale), 0, nat.ptr(sums), nat.ptr(grad),
                                                nat.ptr(flags), nat.ptr(nat.workspace(dev)), nat.stream_ptr(dev))
        nat.check(rc, "llie_pixel_loss_grad")
        nat.count_launch()
        if grad is not None:
            ctx.save_for_backward(grad)
        return (sums[0] * scale).to(torch.float32)

    @staticmethod
    @once_differentiable
    def backward(ctx, gout):
        (grad,) = ctx.saved_tensors
        return grad * gout, None, None, None, None, None


def _weighted(mode: str, pred: Tensor, target: Tensor, weight: Optional[Tensor], reduction: str, eps: float) -> Tensor:
    """loss_util.py:31-95 (`weight_reduce_loss` behind `@weighted_loss`)."""
    if reduction not in _reduction_modes:
        raise ValueError(f"{reduction} is not a valid value for reduction")
    if pred.shape != target.shape:
        raise ValueError(f"pred and target must share a shape, got {tuple(pred.shape)} vs {tuple(target.shape)}")
    if weight is not None:
        assert weight.dim() == pred.dim()
        assert weight.size(1) == 1 or weight.size(1) == pred.size(1)
    dev = nat.cuda_device_for(pred, target)
    home = pred.device
    x = nat.stage(pred, dev, torch.float32)
    y = nat.stage(target.detach(), dev, torch.float32)
    w = nat.stage(weight.detach(), dev, torch.float32) if weight is not None else None
    fusable = (reduction != "none" and x.ndim == 4 and (x.shape[2] * x.shape[3]) % 4 == 0
               and (w is None or tuple(w.shape[2:]) == tuple(x.shape[2:]) and w.shape[0] == x.shape[0]))
    if not fusable:
        loss = _elementwise(mode, x, y, eps)
        if w is not None:
            loss = loss * w
        if w is None or reduction == "sum":
            out = loss if reduction == "none" else (loss.mean() if reduction == "mean" else loss.sum())
        elif reduction == "mean":
            out = loss.sum() / (w.sum() if w.size(1) > 1 else w.sum() * loss.size(1))
        else:
            out = loss
        return out if home == dev else out.to(home)
    if w is None:
        scale = 1.0 / x.numel() if reduction == "mean" else 1.0
        out = _PixelLossFn.apply(x, y, None, _MODES[mode], eps, scale)
    else:
        out = _PixelLossFn.apply(x, y, w, _MODES[mode], eps, 1.0)
        if reduction == "mean":  # mean over the weighted region (loss_util.py:52-57)
            out = out / (w.sum() if w.size(1) > 1 else w.sum() * x.size(1))
    return out if home == dev else out.to(home)


def l1_loss(pred, target, weight=None, reduction="mean", **kwargs):
    return _weighted("l1", pred, target, weight, reduction, 0.0)


def mse_loss(pred, target, weight=None, reduction="mean", **kwargs):
    return _weighted("mse", pred, target, weight, reduction, 0.0)


def charbonnier_loss(pred, target, weight=None, reduction="mean", eps=1e-12, **kwargs):
    return _weighted("charbonnier", pred, target, weight, reduction, float(eps))


class _Base(nn.Module):
    def __init__(self, loss_weight=1.0, reduction="mean"):
        super().__init__()
        if reduction not in _reduction_modes:
            raise ValueError(f"Unsupported reduction mode: {reduction}. Supported ones are: {_reduction_modes}")
        self.loss_weight = loss_weight
        self.reduction = reduction


class L1Loss(_Base):
    """L1 (MAE) loss (losses.py:33-59)."""

    def forward(self, pred, target, weight=None, **kwargs):
        return self.loss_weight * l1_loss(pred, target, weight, reduction=self.reduction)


class MSELoss(_Base):
    """MSE (L2) loss (losses.py:61-87)."""

    def forward(self, pred, target, weight=None, **kwargs):
        return self.loss_weight * mse_loss(pred, target, weight, reduction=self.reduction)


class CharbonnierLoss(_Base):
    """Charbonnier loss sqrt(d^2 + eps) (losses.py:90-108; eps enters un-squared)."""

    def __init__(self, loss_weight=1.0, reduction="mean", eps=1e-12):
        super().__init__(loss_weight, reduction)
        self.eps = eps

    def forward(self, pred, target, weight=None, **kwargs):
        return self.loss_weight * charbonnier_loss(pred, target, weight, reduction=self.reduction, eps=self.eps)


class _PSNRLossFn(torch.autograd.Function):
    """scale * mean_n log(mse_n + 1e-8): per-sample MSE in float64 from the PSNR statistics kernel, gradient
    2 (x - y) * coef_n from the pixel-loss kernel with a per-sample device coefficient."""

    @staticmethod
    def forward(ctx, pred: Tensor, target: Tensor, scale: float):
        x = pred.detach().contiguous()
        y = target.detach().contiguous()
        n = x.shape[0]
        chw = x.numel() // n
        table, _, _ = Fn.psnr_stats(x, y, n, 1, chw, cast_first=False, clamp_hi=None, data_range=1.0, eps=1e-12)
        mse = table[:n, 0]                                        # float64 per-sample MSE
        ctx.save_for_backward(x, y, mse)
        ctx.scale = scale
        return (scale * torch.log(mse + 1e-8).mean()).to(torch.float32)

    @staticmethod
    @once_differentiable
    def backward(ctx, gout):
        x, y, mse = ctx.saved_tensors
        dev = x.device
        n, c = x.shape[0], x.shape[1]
        plane = x.numel() // (n * c)
        coef = (gout.double() * ctx.scale / (n * (mse + 1e-8) * (c * plane))).to(torch.float32).contiguous()
        if plane % 4 != 0:
            return (2.0 * (x - y)) * coef.view(n, 1, 1, 1), None, None
        grad = torch.empty_like(x)
        sums = torch.empty(1, dtype=torch.float64, device=dev)
        flags = torch.empty(1, dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            rc = nat.lib().llie_pixel_loss_grad(1, nat.ptr(x), nat.ptr(y), None, 0, nat.ptr(coef), n, c, plane, 0.0, 1.0,
                                                0, nat.ptr(sums), nat.ptr(grad), nat.ptr(flags),
                                                nat.ptr(nat.workspace(dev)), nat.stream_ptr(dev))
        nat.check(rc, "llie_pixel_loss_grad")
        nat.count_launch()
        return grad, None, None


class PSNRLoss(nn.Module):
    """loss_weight * 10 / ln 10 * mean_n ln(mse_n + 1e-8), optionally on the BT.601 luma (losses.py:110-138)."""

    def __init__(self, loss_weight=1.0, reduction="mean", toY=False):
        super().__init__()
        assert reduction == "mean"
        self.loss_weight = loss_weight
        self.scale = 10 / math.log(10)
        self.toY = toY
        self.coef = torch.tensor([65.481, 128.553, 24.966]).reshape(1, 3, 1, 1)

    def forward(self, pred, target):
        assert len(pred.size()) == 4
        dev = nat.cuda_device_for(pred, target)
        home = pred.device
        x = nat.stage(pred, dev, torch.float32)
        y = nat.stage(target.detach(), dev, torch.float32)
        if self.toY:
            coef = self.coef.to(dev)
            x = ((x * coef).sum(dim=1).unsqueeze(dim=1) + 16.0) / 255.0
            y = ((y * coef).sum(dim=1).unsqueeze(dim=1) + 16.0) / 255.0
        out = self.loss_weight * _PSNRLossFn.apply(x, y, float(self.scale))
        return out if home == dev else out.to(home)
