"""Linear-domain PSNR / SSIM — drop-in for the reference `metrics/linear.py`.

`psnr_linear` keeps the reference signature (linear.py:140-148) and numerics: the difference is taken
in the input dtype (fp32 or fp64), squared and accumulated in float64, one fused streaming kernel
(csrc/pointwise.cu) instead of sub + cast (an 8 B/elem temporary) + pow + mean.
`ssim_linear` is the separable-Gaussian SSIM kernel of csrc/ssim.cu.
"""
from __future__ import annotations

from typing import Literal

import torch
from torch import Tensor

from .. import _functional as Fn
from .. import _native as nat

__all__ = ["psnr_linear", "ssim_linear"]

_Reduction = Literal["mean", "sum", "none"]
_ChannelAggregate = Literal["mean", "none"]
_PaddingMode = Literal["reflect", "replicate", "circular", "constant"]


def _check_pair(pred: Tensor, target: Tensor) -> tuple[Tensor, Tensor]:
    """Type / shape contract of linear.py:27-65 (finiteness is checked inside the kernel)."""
    if not isinstance(pred, Tensor) or not isinstance(target, Tensor):
        raise TypeError("psnr_linear/ssim_linear expect torch.Tensor inputs.")
    if pred.dtype not in {torch.float32, torch.float64}:
        raise TypeError(f"Expected pred dtype float32/float64, received {pred.dtype}.")
    if target.dtype != pred.dtype:
        raise TypeError("pred and target must share the same dtype.")
    if pred.device != target.device:
        raise ValueError("pred and target must live on the same device.")
    if pred.shape != target.shape:
        raise ValueError(f"pred and target must share identical shape, got {pred.shape} vs {target.shape}.")
    if pred.ndim == 3:
        pred, target = pred.unsqueeze(0), target.unsqueeze(0)
    elif pred.ndim != 4:
        raise ValueError(
            "Inputs must have 3 (C,H,W) or 4 (N,C,H,W) dimensions; "
            f"received tensor with shape {tuple(pred.shape)}.")
    if pred.shape[0] == 0:
        raise ValueError("Batch dimension cannot be zero.")
    if pred.shape[1] == 0:
        raise ValueError("Channel dimension cannot be zero.")
    if pred.shape[2] == 0 or pred.shape[3] == 0:
        raise ValueError("Spatial dimensions must be strictly positive.")
    return pred.detach(), target.detach()


def _reduce(values: Tensor, reduction: str) -> Tensor:
    if reduction == "none":
        return values
    if reduction == "mean":
        return values.mean(dim=0)
    if reduction == "sum":
        return values.sum(dim=0)
    raise ValueError(f"Unsupported reduction='{reduction}'. Expected 'mean', 'sum', or 'none'.")


def _raise_nonfinite(flags: Tensor, first: str, second: str) -> None:
    f = int(flags.item())
    if f & 1:
        raise ValueError(f"{first} contains NaN or Inf values.")
    if f & 2:
        raise ValueError(f"{second} contains NaN or Inf values.")


@torch.no_grad()
def psnr_linear(
    pred: Tensor,
    target: Tensor,
    *,
    data_range: float = 1.0,
    reduction: _Reduction = "mean",
    clamp: bool | float = False,
    eps: float = 1e-12,
) -> Tensor:
    """PSNR = 10 log10(L^2 / MSE) per sample (MSE over C*H*W in float64), `inf` where MSE <= eps,
    then `reduction` over the batch.  Returns a float64 tensor like the reference."""
    if data_range <= 0:
        raise ValueError(f"`data_range` must be positive, received {data_range}.")
    if eps <= 0:
        raise ValueError(f"`eps` must be positive, received {eps}.")
    if reduction not in {"mean", "sum", "none"}:
        raise ValueError(f"Unsupported reduction='{reduction}'. Expected 'mean', 'sum', or 'none'.")
    pred, target = _check_pair(pred, target)
    hi = None
    if isinstance(clamp, bool):
        if clamp:
            hi = float(data_range)
    else:
        hi = float(clamp)
    home = pred.device
    dev = nat.cuda_device_for(pred, target)
    p, t = nat.stage(pred, dev), nat.stage(target, dev)
    n, c, h, w = p.shape
    table, flags, _ = Fn.psnr_stats(p, t, n, c, h * w, cast_first=False, clamp_hi=hi, data_range=data_range, eps=eps)
    _raise_nonfinite(flags, "pred", "target")
    col = 2 * c + 2
    out = table[:n, col] if reduction == "none" else (table[n, col] if reduction == "mean" else table[n + 1, col])
    return out if home == dev else out.to(home)


def ssim_linear(*args, **kwargs) -> Tensor:  # replaced below once the SSIM kernel module is imported
    from ._ssim_impl import ssim_linear as impl
    return impl(*args, **kwargs)
