"""Channel-wise RGB metrics — drop-in for the reference `metrics/channelwise.py`.

rgb_psnr / cpsnr_rgb: inputs are promoted to float64 (and optionally clamped) BEFORE the difference
(channelwise.py:146-150, :205-210); one streaming kernel yields the per-channel MSEs and the finished
PSNR table.  rgb_ssim: per-channel `ssim_linear` scores (channelwise.py:262-295) from one SSIM launch.
"""
from __future__ import annotations

from typing import Dict, Literal, Tuple, TypedDict, Union

import torch
from torch import Tensor

from .. import _functional as Fn
from .. import _native as nat

__all__ = ["rgb_psnr", "cpsnr_rgb", "rgb_ssim", "MetricMeta", "PSNRResult", "CPSNRResult", "SSIMResult"]


class MetricMeta(TypedDict):
    """channelwise.py:32-34."""
    domain: Literal["linear", "srgb"]
    data_range: float


class PSNRResult(TypedDict):
    R: Tensor
    G: Tensor
    B: Tensor
    mean: Tensor
    meta: MetricMeta


class CPSNRResult(TypedDict):
    cpsnr: Tensor
    meta: MetricMeta


class SSIMResult(TypedDict):
    R: Tensor
    G: Tensor
    B: Tensor
    mean: Tensor
    meta: MetricMeta

_Reduction = Literal["mean", "sum", "none"]


def _check_rgb(pred: Tensor, target: Tensor) -> Tuple[Tensor, Tensor]:
    """channelwise.py:62-92 (finiteness is checked inside the kernel)."""
    if not isinstance(pred, Tensor) or not isinstance(target, Tensor):
        raise TypeError("Expected torch.Tensor inputs for rgb metrics.")
    if pred.shape != target.shape:
        raise ValueError(
            f"`pred` and `target` must share identical shape, got {tuple(pred.shape)} vs {tuple(target.shape)}.")
    if pred.device != target.device:
        raise ValueError("`pred` and `target` must reside on the same device.")
    if pred.ndim == 3:
        pred, target = pred.unsqueeze(0), target.unsqueeze(0)
    elif pred.ndim != 4:
        raise ValueError(
            "Inputs must have 3 (C,H,W) or 4 (N,C,H,W) dimensions; "
            f"received tensor with shape {tuple(pred.shape)}.")
    if pred.shape[1] != 3:
        raise ValueError(f"Channel-wise metrics require exactly 3 channels (RGB). Received {pred.shape[1]}.")
    return pred.detach(), target.detach()


def _check_reduction(reduction: str) -> None:
    if reduction not in {"mean", "sum", "none"}:
        raise ValueError(f"Unsupported reduction '{reduction}'. Choose from 'mean', 'sum', 'none'.")


def _pick(table: Tensor, col: int, n: int, reduction: str) -> Tensor:
    return table[:n, col] if reduction == "none" else (table[n, col] if reduction == "mean" else table[n + 1, col])


def _clamp_hi(clamp, data_range):
    # channelwise.py:104-108: bool is an int, so clamp=True clamps to [0, 1.0] (float(True)), NOT to data_range
    if not clamp:
        return None
    return float(clamp) if isinstance(clamp, (int, float)) else float(data_range)


def _table(pred, target, data_range, clamp, eps):
    home = pred.device
    dev = nat.cuda_device_for(pred, target)
    work = torch.float64 if pred.dtype == torch.float64 else torch.float32
    p, t = nat.stage(pred, dev, work), nat.stage(target, dev, work)
    n, c, h, w = p.shape
    table, flags, _ = Fn.psnr_stats(p, t, n, c, h * w, cast_first=True, clamp_hi=_clamp_hi(clamp, data_range),
                                    data_range=data_range, eps=eps)
    f = int(flags.item())
    if f & 1:
        raise ValueError("`pred` contains NaN or Inf values.")
    if f & 2:
        raise ValueError("`target` contains NaN or Inf values.")
    return table, n, home, dev


@torch.no_grad()
def rgb_psnr(pred: Tensor, target: Tensor, *, data_range: float = 1.0, reduction: _Reduction = "mean",
             domain: Literal["linear", "srgb"] = "linear", clamp: bool | float = False, meta: bool = False,
             eps: float = 1e-12) -> Dict[str, Union[Tensor, dict]]:
    """Per-channel PSNR {'R','G','B','mean'} (+ 'meta'), float64, eps-floored MSE (no inf)."""
    if data_range <= 0:
        raise ValueError(f"`data_range` must be positive; received {data_range}.")
    if eps <= 0:
        raise ValueError(f"`eps` must be positive; received {eps}.")
    _check_reduction(reduction)
    pred_n, target_n = _check_rgb(pred, target)
    table, n, home, dev = _table(pred_n, target_n, data_range, clamp, eps)
    res = {k: _pick(table, col, n, reduction) for k, col in (("R", 3), ("G", 4), ("B", 5), ("mean", 6))}
    if home != dev:
        res = {k: v.to(home) for k, v in res.items()}
    if meta:
        res["meta"] = {"domain": domain, "data_range": float(data_range)}
    return res


@torch.no_grad()
def cpsnr_rgb(pred: Tensor, target: Tensor, *, data_range: float = 1.0, reduction: _Reduction = "mean",
              domain: Literal["linear", "srgb"] = "linear", clamp: bool | float = False, meta: bool = False,
              eps: float = 1e-12):
    """Colour PSNR: 10 log10(L^2 / mean_c MSE_c) per sample, then `reduction` (float64)."""
    if data_range <= 0:
        raise ValueError(f"`data_range` must be positive; received {data_range}.")
    if eps <= 0:
        raise ValueError(f"`eps` must be positive; received {eps}.")
    _check_reduction(reduction)
    pred_n, target_n = _check_rgb(pred, target)
    table, n, home, dev = _table(pred_n, target_n, data_range, clamp, eps)
    metric = _pick(table, 7, n, reduction)
    if home != dev:
        metric = metric.to(home)
    if not meta:
        return metric
    return {"cpsnr": metric, "meta": {"domain": domain, "data_range": float(data_range)}}


def rgb_ssim(*args, **kwargs):
    from ._ssim_impl import rgb_ssim as impl
    return impl(*args, **kwargs)
