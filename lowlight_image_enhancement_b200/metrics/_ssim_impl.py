"""ssim_linear / rgb_ssim on the fused separable SSIM kernel (csrc/ssim.cu).

Reference: metrics/linear.py:213-324 (ssim_linear) and metrics/channelwise.py:226-302 (rgb_ssim).
The K x K window of the reference is an outer product, so it is applied separably; the five moment
maps never leave shared memory.  Inputs in float64 are computed in float32 (the kernel's arithmetic
type) and the score is returned in the input dtype.
"""
from __future__ import annotations

from typing import Dict, Literal, Union

import torch
from torch import Tensor

from .. import _functional as Fn
from .. import _native as nat

_PADS = {"reflect", "replicate", "circular", "constant"}


def _stats(pred: Tensor, target: Tensor, *, data_range, kernel_size, sigma, k1, k2, gaussian, padding, eps):
    """-> per-(n, c) SSIM means [N, C] float32 on the compute device, plus (home, dev)."""
    if kernel_size <= 0 or kernel_size % 2 == 0:
        raise ValueError(f"kernel_size must be a positive odd integer; received {kernel_size}.")
    if gaussian and sigma <= 0:
        raise ValueError("sigma must be positive when gaussian=True.")
    if padding not in _PADS:
        raise ValueError(f"Unsupported padding mode '{padding}'.")
    n, c, h, w = pred.shape
    if h < kernel_size or w < kernel_size:
        raise ValueError(
            f"Spatial dimensions must be >= kernel_size. Got H={h}, W={w}, kernel_size={kernel_size}.")
    if kernel_size > nat.SSIM_MAX_K:
        raise NotImplementedError(f"the CUDA SSIM kernels take odd windows up to {nat.SSIM_MAX_K}x{nat.SSIM_MAX_K}")
    if padding == "reflect" and (kernel_size // 2 >= h or kernel_size // 2 >= w):
        raise RuntimeError("Padding size should be less than the corresponding input dimension")
    home = pred.device
    dev = nat.cuda_device_for(pred, target)
    p = nat.stage(pred, dev, torch.float32)
    t = nat.stage(target, dev, torch.float32)
    win = Fn.window_1d("linear" if gaussian else "uniform", int(kernel_size), float(sigma), dev)
    c1 = (k1 * float(data_range)) ** 2
    c2 = (k2 * float(data_range)) ** 2
    sums, flags, _ = Fn.ssim_stats(p, t, win, pad_mode=nat.PAD[padding], c1=c1, c2=c2, eps=eps, clamp_var=True)
    return sums / float(h * w), flags, home, dev


def _reduce(values: Tensor, reduction: str) -> Tensor:
    if reduction == "none":
        return values
    if reduction == "mean":
        return values.mean(dim=0)
    if reduction == "sum":
        return values.sum(dim=0)
    raise ValueError(f"Unsupported reduction='{reduction}'. Expected 'mean', 'sum', or 'none'.")


@torch.no_grad()
def ssim_linear(pred: Tensor, target: Tensor, *, data_range: float = 1.0, kernel_size: int = 11,
                sigma: float = 1.5, k1: float = 0.01, k2: float = 0.03, gaussian: bool = True,
                reduction: str = "mean", channel_aggregate: str = "mean", padding: str = "reflect",
                eps: float = 1e-12) -> Tensor:
    """SSIM in the linear domain, same signature / validation order as linear.py:213-280."""
    from .linear import _check_pair, _raise_nonfinite
    if data_range <= 0:
        raise ValueError(f"`data_range` must be positive, received {data_range}.")
    if eps <= 0:
        raise ValueError(f"`eps` must be positive, received {eps}.")
    if k1 < 0 or k2 < 0:
        raise ValueError("k1 and k2 must be non-negative.")
    if channel_aggregate not in {"mean", "none"}:
        raise ValueError(f"channel_aggregate must be 'mean' or 'none', received {channel_aggregate}.")
    if reduction not in {"mean", "sum", "none"}:
        raise ValueError(f"Unsupported reduction='{reduction}'. Expected 'mean', 'sum', or 'none'.")
    pred, target = _check_pair(pred, target)
    per_c, flags, home, dev = _stats(pred, target, data_range=data_range, kernel_size=kernel_size, sigma=sigma,
                                     k1=k1, k2=k2, gaussian=gaussian, padding=padding, eps=eps)
    _raise_nonfinite(flags, "pred", "target")
    per_c = per_c.to(pred.dtype)
    per = per_c.mean(dim=1) if channel_aggregate == "mean" else per_c
    out = _reduce(per, reduction)
    return out if home == dev else out.to(home)


def rgb_ssim(pred: Tensor, target: Tensor, *, data_range: float = 1.0, kernel_size: int = 11, sigma: float = 1.5,
             k1: float = 0.01, k2: float = 0.03, gaussian: bool = True, reduction: str = "mean",
             channel_aggregate: Literal["none", "mean"] = "none", padding: str = "reflect",
             domain: Literal["linear", "srgb"] = "linear", meta: bool = False,
             eps: float = 1e-12) -> Union[Dict[str, Tensor], dict]:
    """Per-channel SSIM {R, G, B, mean} (channelwise.py:226-302): one kernel launch over the three planes
    instead of three ssim_linear calls."""
    from .channelwise import _check_rgb
    if data_range <= 0:
        raise ValueError(f"`data_range` must be positive; received {data_range}.")
    if eps <= 0:
        raise ValueError(f"`eps` must be positive, received {eps}.")
    if k1 < 0 or k2 < 0:
        raise ValueError("k1 and k2 must be non-negative.")
    if reduction not in {"mean", "sum", "none"}:
        raise ValueError(f"Unsupported reduction='{reduction}'. Expected 'mean', 'sum', or 'none'.")
    pred_n, target_n = _check_rgb(pred, target)
    with torch.no_grad():
        per_c, flags, home, dev = _stats(pred_n, target_n, data_range=data_range, kernel_size=kernel_size,
                                         sigma=sigma, k1=k1, k2=k2, gaussian=gaussian, padding=padding, eps=eps)
        f = int(flags.item())
        if f & 1:
            raise ValueError("pred contains NaN or Inf values.")
        if f & 2:
            raise ValueError("target contains NaN or Inf values.")
        mean_scores = per_c.mean(dim=1)
        agg = _reduce(mean_scores, reduction)
        if channel_aggregate == "mean":
            res = {"R": agg, "G": agg, "B": agg, "mean": agg}
        else:
            res = {"R": _reduce(per_c[:, 0], reduction), "G": _reduce(per_c[:, 1], reduction),
                   "B": _reduce(per_c[:, 2], reduction), "mean": agg}
        if home != dev:
            res = {k: v.to(home) for k, v in res.items()}
    if meta:
        res["meta"] = {"domain": domain, "data_range": float(data_range)}
    return res
