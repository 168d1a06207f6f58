"""Physics-consistency metrics — drop-in for the reference `metrics/phys_consistency.py`.

    A ~= clip(rho * (K * B_hat)) + n

`phys_cons_raw` / `phys_cons_srgb` keep the reference signatures (phys_consistency.py:260-274,
:323-338), validation order, exception types and messages, but the whole projector
(pad + PSF conv + rho + clip + crop + |.|/Charbonnier + per-sample mean + finiteness check) is ONE
fused sm_100a kernel (csrc/phys.cu) instead of ~10 eager passes.

Supersets of the reference:
  * `psf` may also be [1,1,kh,kw] / 2-D (mono, broadcast over channels) or [C,1,kh,kw]
    (per-channel depthwise) — the forms README.md:142 of the reference advertises but its
    `_prepare_psf` rejects; dense [C_out,C_in,kh,kw] keeps working.
  * `differentiable=True` keeps the autograd graph through `pred` (the reference metrics are
    no-grad); the gradient is the fused conv_transpose2d adjoint kernel.
  * host tensors are accepted: they are uploaded, reduced on the GPU, and the result is returned on
    the input's device.  Without CUDA the call raises (no CPU implementation exists).
"""
from __future__ import annotations

from typing import Literal, Tuple

import torch
from torch import Tensor

from .. import _functional as Fn
from .. import _native as nat

__all__ = ["phys_cons_raw", "phys_cons_srgb"]

_Reduction = Literal["mean", "sum", "none"]
_Padding = Literal["reflect", "replicate", "zeros"]
_Crop = Literal["valid", "same"]
_Robust = Literal["none", "charbonnier"]


def _check_pair(pred: Tensor, obs: Tensor) -> Tuple[Tensor, Tensor]:
    """Shape/device contract of phys_consistency.py:37-62 (finiteness is checked inside the kernel)."""
    if not isinstance(pred, Tensor) or not isinstance(obs, Tensor):
        raise TypeError("phys_consistency metrics expect torch.Tensor inputs.")
    if pred.shape != obs.shape:
        raise ValueError(
            f"`pred` and `obs` must share identical shape, got {tuple(pred.shape)} vs {tuple(obs.shape)}.")
    if pred.device != obs.device:
        raise ValueError("`pred` and `obs` must reside on the same device.")
    if pred.ndim == 3:
        pred, obs = pred.unsqueeze(0), obs.unsqueeze(0)
    elif pred.ndim != 4:
        raise ValueError(
            "Inputs must have 3 (C,H,W) or 4 (N,C,H,W) dimensions; "
            f"received tensor with shape {tuple(pred.shape)}.")
    if pred.shape[1] == 0:
        raise ValueError("Channel dimension must be positive.")
    return pred, obs


def _core(pred, obs, *, psf, expo_ratio, reduction, padding, normalize_psf, enforce_nonnegative, crop, robust,
          return_map, clamp01, eps, differentiable):
    if eps <= 0:
        raise ValueError(f"`eps` must be positive, received {eps}.")
    if robust not in {"none", "charbonnier"}:
        raise ValueError(f"Unsupported robust loss '{robust}'.")
    if crop not in {"valid", "same"}:
        raise ValueError(f"Unsupported crop mode '{crop}'.")
    if padding not in {"reflect", "replicate", "zeros"}:
        raise ValueError(f"Unsupported padding mode '{padding}'.")
    if reduction not in {"mean", "sum", "none"}:
        raise ValueError(f"Unsupported reduction '{reduction}'. Choose from 'mean', 'sum', 'none'.")

    home = pred.device
    dev = nat.cuda_device_for(pred, obs)
    p = nat.stage(pred if differentiable else pred.detach(), dev, torch.float32)
    o = nat.stage(obs.detach(), dev, torch.float32)
    n, c, h, w = p.shape
    prepared = Fn.prepare_psf(psf, c, o.shape[1], dev, normalize=normalize_psf, nonneg=enforce_nonnegative, eps=eps)
    if prepared.zero_sum:
        Fn.warn_zero_sum()
    if padding == "reflect" and (prepared.kh // 2 >= h or prepared.kw // 2 >= w):
        raise RuntimeError("Padding size should be less than the corresponding input dimension")
    rho_t, rho_s, rho_st = Fn.normalise_rho(expo_ratio, n, c, h, w, dev)
    cfg = Fn.PhysCfg(nat.PAD[padding], crop == "valid", robust == "charbonnier", float(eps), 0, bool(clamp01), False)

    amap = None
    if differentiable and p.requires_grad:
        # same non-finite input contract and the same optional |residual| map as the no-grad metric
        vec, flags, amap = Fn.PhysProjector.apply(p, o, prepared, rho_t, rho_s, rho_st, cfg, return_map)
    else:
        vec, flags, amap = Fn.phys_forward(p.detach(), o, prepared, rho_t, rho_s, rho_st, cfg, return_map)
    if flags is not None:
        f = int(flags.item())  # the one host sync (the reference pays two isfinite().all() syncs)
        if f & 4 and not f & 3:
            # un-clipped paths detect non-finite inputs through the (non-finite) sum, which cannot say
            # which tensor it came from: resolve on this error path only
            f |= 1 if not bool(torch.isfinite(p).all()) else 0
            f |= 2 if not bool(torch.isfinite(o).all()) else 0
        if f & 1:
            raise ValueError("`pred` contains NaN or Inf values.")
        if f & 2:
            raise ValueError("`obs` contains NaN or Inf values.")
    metric = vec[:n] if reduction == "none" else (vec[n] if reduction == "mean" else vec[n + 1])
    if home != dev:
        metric = metric.to(home)
        amap = amap.to(home) if amap is not None else None
    return (metric, amap) if return_map else metric


def phys_cons_raw(
    pred_linear: Tensor,
    obs_short_linear: Tensor,
    psf: Tensor,
    expo_ratio: float | Tensor,
    *,
    reduction: _Reduction = "mean",
    padding: _Padding = "reflect",
    normalize_psf: bool = True,
    enforce_nonnegative: bool = False,
    crop: _Crop = "valid",
    robust: _Robust = "none",
    return_map: bool = False,
    eps: float = 1e-12,
    differentiable: bool = False,
) -> Tensor | Tuple[Tensor, Tensor]:
    """Physics-consistency error in the RAW / linear domain (lower is better):
    mean |rho * (K * B_hat) - A| (or Charbonnier) per sample, then `reduction` over the batch.
    Same arguments and defaults as the reference (phys_consistency.py:260-319)."""
    pred, obs = _check_pair(pred_linear, obs_short_linear)
    with torch.set_grad_enabled(differentiable and torch.is_grad_enabled()):
        return _core(pred, obs, psf=psf, expo_ratio=expo_ratio, reduction=reduction, padding=padding,
                     normalize_psf=normalize_psf, enforce_nonnegative=enforce_nonnegative, crop=crop,
                     robust=robust, return_map=return_map, clamp01=False, eps=eps, differentiable=differentiable)


def phys_cons_srgb(
    pred_srgb: Tensor,
    obs_short_srgb: Tensor,
    psf: Tensor,
    expo_ratio: float | Tensor = 1.0,
    *,
    reduction: _Reduction = "mean",
    padding: _Padding = "reflect",
    normalize_psf: bool = True,
    enforce_nonnegative: bool = False,
    crop: _Crop = "valid",
    robust: _Robust = "none",
    clamp01: bool = True,
    return_map: bool = False,
    eps: float = 1e-12,
    differentiable: bool = False,
) -> Tensor | Tuple[Tensor, Tensor]:
    """sRGB-domain variant (phys_consistency.py:323-368): the synthesised observation
    rho * (K * pred) is clipped to [0,1] when `clamp01` (default)."""
    pred, obs = _check_pair(pred_srgb, obs_short_srgb)
    with torch.set_grad_enabled(differentiable and torch.is_grad_enabled()):
        return _core(pred, obs, psf=psf, expo_ratio=expo_ratio, reduction=reduction, padding=padding,
                     normalize_psf=normalize_psf, enforce_nonnegative=enforce_nonnegative, crop=crop,
                     robust=robust, return_map=return_map, clamp01=clamp01, eps=eps,
                     differentiable=differentiable)
