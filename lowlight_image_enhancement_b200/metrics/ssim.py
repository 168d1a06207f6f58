"""SSIM evaluator — drop-in for the reference `metrics/ssim.py` (SSIMEvaluator, calculate_ssim,
calculate_ssim_per_image, evaluate_pairs_ssim, SSIMMetric; ssim.py:188-538).

The reference delegates the arithmetic to torchmetrics 1.2.0 `StructuralSimilarityIndexMeasure`
(gaussian_kernel=True): Gaussian support derived from sigma (2*int(3.5*sigma+0.5)+1, NOT from
kernel_size), reflect padding, SSIM map without variance clamp / eps, mean over the interior
(pad pixels cropped per side) of all channels jointly.  Here that is one launch of the fused SSIM
kernel (csrc/ssim.cu, `interior=1`), BT.601 luma computed inside the kernel's tile load, and the
auto data range taken from the streaming min/max kernel.  No torchmetrics dependency.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Any, Dict, List, Literal, Optional, Tuple

import torch
import torch.nn.functional as F
from torch import Tensor

from .. import _functional as Fn
from .. import _native as nat

__all__ = ["SSIMEvaluator", "SSIMMetricUnified", "calculate_ssim", "calculate_ssim_per_image",
           "evaluate_pairs_ssim", "SSIMMetric"]


def _ensure_batch_dim(tensor: Tensor) -> Tensor:
    if tensor.ndim == 4:
        return tensor
    if tensor.ndim == 3:
        return tensor.unsqueeze(0)
    raise ValueError(
        "SSIM expects tensors with 3 (C,H,W) or 4 (N,C,H,W) dimensions. "
        f"Received shape {tuple(tensor.shape)}.")


def _valid_kernel_size(height: int, width: int, kernel_size: int) -> int:
    """ssim.py:74-89: clip to an odd integer not exceeding the smallest dimension."""
    k = int(kernel_size)
    if k <= 0:
        raise ValueError(f"kernel_size must be positive, received {kernel_size}.")
    if k % 2 == 0:
        k -= 1
    k = max(1, min(k, height, width))
    if k % 2 == 0:
        k -= 1
    if k < 1:
        raise ValueError(f"kernel_size cannot be adjusted to a valid value for shape {(height, width)}.")
    return k


def _percentile_95(values: Tensor) -> float:
    if values.numel() == 0:
        return float("nan")
    if values.numel() == 1:
        return float(values.item())
    return float(torch.quantile(values, 0.95).item())


def _align_pair(target: Tensor, prediction: Tensor, policy, mode="bilinear") -> Tuple[Tensor, Tensor]:
    """ssim.py:130-167 (plumbing: stock torch ops, not on the hot path)."""
    if policy is None:
        if target.shape[-2:] != prediction.shape[-2:]:
            raise ValueError(
                "SSIM requires equal spatial dimensions when no resize_policy is set. "
                f"Got target={target.shape[-2:]}, prediction={prediction.shape[-2:]}")
        return target, prediction
    if policy == "resize":
        prediction = F.interpolate(prediction, size=target.shape[-2:], mode=mode, align_corners=False)
        return target, prediction
    if policy == "center_crop":
        h = min(target.shape[-2], prediction.shape[-2])
        w = min(target.shape[-1], prediction.shape[-1])

        def _crop(x: Tensor, hh: int, ww: int) -> Tensor:
            _, _, H, W = x.shape
            top, left = max((H - hh) // 2, 0), max((W - ww) // 2, 0)
            return x[:, :, top:top + hh, left:left + ww]
        return _crop(target, h, w), _crop(prediction, h, w)
    raise ValueError(f"Unknown resize_policy '{policy}'. Use None, 'resize', or 'center_crop'.")


@dataclass
class _SSIMConfig:
    kernel_size: int = 11
    sigma: float = 1.5
    k1: float = 0.01
    k2: float = 0.03
    color_space: Literal["rgb", "y"] = "rgb"
    resize_policy: Optional[Literal["resize", "center_crop"]] = None
    resize_mode: Literal["bilinear", "bicubic"] = "bilinear"
    reduction: Literal["none", "elementwise_mean"] = "none"
    domain: Optional[Literal["linear", "srgb"]] = None


def torchmetrics_ssim_scores(pred: Tensor, target: Tensor, *, data_range: float, sigma: float, k1: float,
                             k2: float, luma: bool = False) -> Tensor:
    """Per-image SSIM exactly as torchmetrics 1.2.0 `_ssim_update` (gaussian kernel) computes it;
    contiguous fp32 CUDA inputs [N,C,H,W] -> float32 [N] on the same device."""
    n, c, h, w = pred.shape
    gk = int(3.5 * sigma + 0.5) * 2 + 1
    pad = (gk - 1) // 2
    if gk > nat.SSIM_MAX_K:
        raise NotImplementedError(f"the CUDA SSIM kernels take Gaussian supports up to {nat.SSIM_MAX_K} taps (sigma <= 8.9)")
    if h <= 2 * pad or w <= 2 * pad:
        raise ValueError(f"Images of size {(h, w)} are too small for a {gk}x{gk} SSIM window.")
    sums, count, _ = ssim_sums_torchmetrics(pred, target, data_range=data_range, sigma=sigma, k1=k1, k2=k2, luma=luma)
    return (sums.sum(dim=1) / float(count)).to(torch.float32)


def ssim_sums_torchmetrics(pred: Tensor, target: Tensor, *, data_range: float, sigma: float, k1: float, k2: float,
                           luma: bool = False):
    """The raw form of `torchmetrics_ssim_scores`: (float64 sums [N, Co] of the interior SSIM map, elements per image
    those sums cover, kernel flags) -- score_n = float32(sums[n].sum() / count)."""
    n, c, h, w = pred.shape
    gk = int(3.5 * sigma + 0.5) * 2 + 1
    pad = (gk - 1) // 2
    if gk > nat.SSIM_MAX_K:
        raise NotImplementedError(f"the CUDA SSIM kernels take Gaussian supports up to {nat.SSIM_MAX_K} taps (sigma <= 8.9)")
    if h <= 2 * pad or w <= 2 * pad:
        raise ValueError(f"Images of size {(h, w)} are too small for a {gk}x{gk} SSIM window.")
    win = Fn.window_1d("torchmetrics", gk, float(sigma), pred.device)
    if luma:
        # the luma planes first (one streaming pass each, rounded exactly like the reference's eager expression), then
        # the one-channel statistics on the fast kernels: 3.5x faster than the in-kernel luma of the tile kernel
        pred, target, luma = Fn.luma_bt601(pred), Fn.luma_bt601(target), False
    sums, flags, _ = Fn.ssim_stats(pred, target, win, pad_mode=nat.PAD["reflect"], c1=(k1 * data_range) ** 2,
                                   c2=(k2 * data_range) ** 2, eps=0.0, luma=luma, interior=True)
    return sums, sums.shape[1] * (h - 2 * pad) * (w - 2 * pad), flags


class SSIMEvaluator:
    """Model-agnostic SSIM evaluator returning per-image scores and summaries (ssim.py:188-329)."""

    def __init__(self, *, kernel_size: int = 11, sigma: float = 1.5, k1: float = 0.01, k2: float = 0.03,
                 data_range: Optional[float] = None, color_space: Literal["rgb", "y"] = "rgb",
                 resize_policy: Optional[Literal["resize", "center_crop"]] = None,
                 resize_mode: Literal["bilinear", "bicubic"] = "bilinear",
                 reduction: Literal["none", "elementwise_mean"] = "none",
                 domain: Optional[Literal["linear", "srgb"]] = None,
                 device: Optional[torch.device | str] = None) -> None:
        self.cfg = _SSIMConfig(kernel_size=kernel_size, sigma=sigma, k1=k1, k2=k2, color_space=color_space,
                               resize_policy=resize_policy, resize_mode=resize_mode, reduction=reduction,
                               domain=domain)
        self._explicit_data_range = data_range
        self._device = torch.device(device) if device is not None else None

    def _prepare_inputs(self, target: Tensor, prediction: Tensor):
        target = _ensure_batch_dim(target)
        prediction = _ensure_batch_dim(prediction)
        if target.shape != prediction.shape:
            if target.shape[0] != prediction.shape[0] or target.shape[1] != prediction.shape[1]:
                raise ValueError(
                    "SSIM requires the same batch size and channel count for target and prediction. "
                    f"Got target={target.shape}, prediction={prediction.shape}.")
        if self._device is not None and self._device.type == "cuda":
            dev = self._device if self._device.index is not None else nat.cuda_device_for()
        else:
            dev = nat.cuda_device_for(target, prediction)  # a 'cpu' request still computes on the GPU
        target = nat.stage(target.detach(), dev, torch.float32)
        prediction = nat.stage(prediction.detach(), dev, torch.float32)
        target, prediction = _align_pair(target, prediction, self.cfg.resize_policy, self.cfg.resize_mode)
        luma = False
        if self.cfg.color_space == "y":
            if target.shape[1] == 3:
                luma = True
            elif target.shape[1] != 1:
                raise ValueError(f"color_space='y' expects inputs with 1 or 3 channels, got C={target.shape[1]}.")
        elif self.cfg.color_space != "rgb":
            raise ValueError(f"Unsupported color_space '{self.cfg.color_space}'. Use 'rgb' or 'y'.")
        return target.contiguous(), prediction.contiguous(), luma

    @torch.no_grad()
    def __call__(self, ground_truth: Tensor, prediction: Tensor) -> Dict[str, Any]:
        target, pred, luma = self._prepare_inputs(ground_truth, prediction)
        n, c, h, w = target.shape
        if self._explicit_data_range is not None:
            data_range = float(self._explicit_data_range)
        else:
            # ssim.py:62-71 on the tensors the metric sees (luma when color_space='y'); one streaming pass
            a, b = (target, pred)
            if luma:
                a, b = Fn.luma_bt601(target), Fn.luma_bt601(pred)
            _, _, mm = Fn.psnr_stats(a, b, a.shape[0], a.shape[1], h * w, cast_first=False, clamp_hi=None,
                                     data_range=1.0, eps=1e-12, want_minmax=True)
            lo, hi = (float(v) for v in mm.tolist())
            data_range = 1.0 if (hi <= 1.0 and lo >= 0.0) else (2.0 if (hi <= 1.0 and lo >= -1.0) else 255.0)
        kernel = _valid_kernel_size(h, w, self.cfg.kernel_size)
        self.cfg.kernel_size = kernel
        scores = torchmetrics_ssim_scores(pred, target, data_range=data_range, sigma=float(self.cfg.sigma),
                                          k1=float(self.cfg.k1), k2=float(self.cfg.k2), luma=luma)
        if self.cfg.reduction == "elementwise_mean":
            scores = scores.mean().view(1)
        scores = scores.view(-1).cpu()
        mean = float(scores.mean().item())
        std = float(scores.std(unbiased=False).item()) if scores.numel() > 1 else 0.0
        median = float(scores.median().item())
        return {
            "per_image": [float(v) for v in scores.tolist()], "mean": mean, "std": std, "p50": median,
            "p95": _percentile_95(scores), "count": int(scores.numel()), "kernel_size": int(kernel),
            "sigma": float(self.cfg.sigma), "k1": float(self.cfg.k1), "k2": float(self.cfg.k2),
            "data_range": float(data_range), "color_space": self.cfg.color_space,
            "resize_policy": self.cfg.resize_policy, "domain": self.cfg.domain,
        }


SSIMMetricUnified = SSIMEvaluator


def _evaluator(data_range, kernel_size, sigma, k1, k2, win_size, color_space, resize_policy, resize_mode, domain,
               reduction, device=None) -> SSIMEvaluator:
    if data_range <= 0:
        raise ValueError(f"data_range must be positive, received {data_range}.")
    if win_size is not None:
        kernel_size = int(win_size)
    return SSIMEvaluator(kernel_size=kernel_size, sigma=sigma, k1=k1, k2=k2, data_range=data_range,
                         color_space=color_space, resize_policy=resize_policy, resize_mode=resize_mode,
                         domain=domain, reduction=reduction, device=device)


def calculate_ssim(img_true: Tensor, img_pred: Tensor, data_range: float, *, kernel_size: int = 11,
                   sigma: float = 1.5, k1: float = 0.01, k2: float = 0.03, win_size: Optional[int] = None,
                   color_space: Literal["rgb", "y"] = "rgb", resize_policy=None, resize_mode="bilinear",
                   domain=None) -> float:
    """Average SSIM between two tensors (ssim.py:341-377)."""
    ev = _evaluator(data_range, kernel_size, sigma, k1, k2, win_size, color_space, resize_policy, resize_mode,
                    domain, "elementwise_mean")
    return ev(img_true, img_pred)["mean"]


@torch.no_grad()
def calculate_ssim_per_image(img_true: Tensor, img_pred: Tensor, data_range: float, *, kernel_size: int = 11,
                             sigma: float = 1.5, k1: float = 0.01, k2: float = 0.03,
                             win_size: Optional[int] = None, color_space: Literal["rgb", "y"] = "rgb",
                             resize_policy=None, resize_mode="bilinear", domain=None) -> Tensor:
    """Per-image SSIM scores of a batch (ssim.py:381-417)."""
    ev = _evaluator(data_range, kernel_size, sigma, k1, k2, win_size, color_space, resize_policy, resize_mode,
                    domain, "none")
    return torch.tensor(ev(img_true, img_pred)["per_image"], dtype=torch.float32)


def evaluate_pairs_ssim(pairs: List[Tuple[Tensor, Tensor]], **kwargs: Any) -> Dict[str, Any]:
    """Evaluate multiple (ground_truth, prediction) pairs under one configuration (ssim.py:425-476)."""
    evaluator = SSIMEvaluator(**kwargs)
    scores: List[float] = []
    last: Optional[Dict[str, Any]] = None
    for gt, pred in pairs:
        last = evaluator(gt, pred)
        scores.extend(last["per_image"])
    if not scores:
        return {"per_image": [], "mean": float("nan"), "std": float("nan"), "p50": float("nan"),
                "p95": float("nan"), "count": 0}
    ts = torch.tensor(scores, dtype=torch.float32)
    summary: Dict[str, Any] = {
        "per_image": [float(v) for v in ts.tolist()], "mean": float(ts.mean().item()),
        "std": float(ts.std(unbiased=False).item()) if ts.numel() > 1 else 0.0,
        "p50": float(ts.median().item()), "p95": _percentile_95(ts), "count": int(ts.numel()),
    }
    if last is not None:
        summary.update({k: last[k] for k in ("kernel_size", "sigma", "k1", "k2", "color_space", "resize_policy",
                                             "data_range", "domain")})
    return summary


class SSIMMetric:
    """Streaming SSIM aggregator (ssim.py:479-538)."""

    def __init__(self, *, data_range: float, kernel_size: int = 11, sigma: float = 1.5, k1: float = 0.01,
                 k2: float = 0.03, color_space: Literal["rgb", "y"] = "rgb",
                 device: Optional[torch.device | str] = None, win_size: Optional[int] = None,
                 resize_policy=None, resize_mode="bilinear", domain=None) -> None:
        if data_range <= 0:
            raise ValueError("data_range must be positive.")
        self._evaluator = _evaluator(data_range, kernel_size, sigma, k1, k2, win_size, color_space, resize_policy,
                                     resize_mode, domain, "none", device)
        self.reset()

    def reset(self) -> None:
        self._sum, self._sumsq, self._count = 0.0, 0.0, 0

    @torch.no_grad()
    def update(self, img_true: Tensor, img_pred: Tensor) -> None:
        scores = torch.tensor(self._evaluator(img_true, img_pred)["per_image"], dtype=torch.float64)
        self._sum += float(scores.sum().item())
        self._sumsq += float((scores ** 2).sum().item())
        self._count += int(scores.numel())

    def compute(self) -> Dict[str, float]:
        if self._count == 0:
            return {"mean": float("nan"), "std": float("nan"), "count": 0}
        mean = self._sum / self._count
        var = max(self._sumsq / self._count - mean * mean, 0.0)
        return {"mean": mean, "std": math.sqrt(var), "count": int(self._count)}
