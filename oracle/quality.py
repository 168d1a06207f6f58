"""Oracle (TEST INFRASTRUCTURE ONLY): PSNR family and the three SSIM policies on CPU torch.

Follows (paths relative to /root/reference):
  * metrics/linear.py:139-210       psnr_linear  (fp32 subtract -> fp64 square/mean)
  * metrics/linear.py:82-121,213-324 ssim_linear (2-D window built in fp64, cast, normalised;
                                     pad; 5 depthwise convs; var clamp; eps in denominator)
  * metrics/channelwise.py:113-302  rgb_psnr / cpsnr_rgb (fp64 inputs) / rgb_ssim (per-channel ssim_linear)
  * metrics/psnr.py:39-88           calculate_psnr (fp64 global mean, isclose(0, atol=1e-12) -> inf)
  * metrics/ssim.py:246-311         SSIMEvaluator -> torchmetrics 1.2.0 StructuralSimilarityIndexMeasure
  * NewBP_model/losses.py:146-155   SSIMLoss -> kornia 0.6.12 kornia.losses.SSIMLoss

Third-party arithmetic (NOT under /root/reference, not installed here -> parity unpinned):
  * torchmetrics==1.2.0 functional/image/ssim.py::_ssim_update, helper._gaussian
  * kornia==0.6.12 metrics/ssim.py::ssim, losses/ssim.py::ssim_loss, filters.get_gaussian_kernel2d, filter2d
restated below from the published algorithms of those releases.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F


# ------------------------------------------------------------------ PSNR family
def psnr_linear(pred, target, *, data_range=1.0, reduction="mean", clamp=False, eps=1e-12):
    if pred.ndim == 3:
        pred, target = pred[None], target[None]
    hi = None
    if isinstance(clamp, bool):
        hi = float(data_range) if clamp else None
    else:
        hi = float(clamp)
    if hi is not None:
        pred, target = pred.clamp(0.0, hi), target.clamp(0.0, hi)
    mse = (pred - target).to(torch.float64).pow(2).flatten(1).mean(dim=1)
    val = 10.0 * torch.log10(float(data_range) ** 2 / torch.maximum(mse, torch.full_like(mse, eps)))
    val = torch.where(mse <= eps, torch.full_like(val, float("inf")), val)
    return _batch(val, reduction)


def _batch(v, reduction):
    if reduction == "none":
        return v
    return v.mean(dim=0) if reduction == "mean" else v.sum(dim=0)


def _clamp_opt(t, clamp, data_range):
    if not clamp:
        return t
    # channelwise.py:104-108: bool is an int in Python, so clamp=True -> hi = float(True) = 1.0
    # (NOT data_range) — quirk kept.
    hi = float(clamp) if isinstance(clamp, (int, float)) else float(data_range)
    return t.clamp(0.0, hi)


def channel_mse64(pred, target, clamp=False, data_range=1.0):
    if pred.ndim == 3:
        pred, target = pred[None], target[None]
    p = _clamp_opt(pred.to(torch.float64), clamp, data_range)
    t = _clamp_opt(target.to(torch.float64), clamp, data_range)
    return (p - t).pow(2).flatten(2).mean(dim=2)


def rgb_psnr(pred, target, *, data_range=1.0, reduction="mean", clamp=False, eps=1e-12):
    mse = channel_mse64(pred, target, clamp, data_range)
    ps = 10.0 * torch.log10(float(data_range) ** 2 / torch.maximum(mse, torch.full_like(mse, eps)))
    return {"R": _batch(ps[:, 0], reduction), "G": _batch(ps[:, 1], reduction),
            "B": _batch(ps[:, 2], reduction), "mean": _batch(ps.mean(dim=1), reduction)}


def cpsnr_rgb(pred, target, *, data_range=1.0, reduction="mean", clamp=False, eps=1e-12):
    cm = channel_mse64(pred, target, clamp, data_range).mean(dim=1)
    return _batch(10.0 * torch.log10(float(data_range) ** 2 / torch.maximum(cm, torch.full_like(cm, eps))), reduction)


def calculate_psnr(img_true, img_pred, data_range):
    mse = torch.mean((img_true.to(torch.float64) - img_pred.to(torch.float64)) ** 2)
    if torch.isclose(mse, mse.new_tensor(0.0), atol=1e-12):
        return float("inf")
    return float((10.0 * torch.log10((data_range ** 2) / mse)).item())


# ------------------------------------------------------------------ ssim_linear
def window2d_linear(k, sigma, gaussian, dtype):
    """linear.py:82-121."""
    if gaussian:
        c = torch.arange(k, dtype=torch.float64) - (k - 1) / 2.0
        g = torch.exp(-(c ** 2) / (2 * sigma ** 2))
        w = g[:, None] * g[None, :]
    else:
        w = torch.ones((k, k), dtype=torch.float64)
    w = w.to(dtype)
    return (w / w.sum()).view(1, 1, k, k)


def ssim_linear(pred, target, *, data_range=1.0, kernel_size=11, sigma=1.5, k1=0.01, k2=0.03,
                gaussian=True, reduction="mean", channel_aggregate="mean", padding="reflect", eps=1e-12,
                per_channel_raw=False):
    if pred.ndim == 3:
        pred, target = pred[None], target[None]
    n, c, h, w = pred.shape
    win = window2d_linear(kernel_size, sigma, gaussian, pred.dtype).repeat(c, 1, 1, 1).to(pred.device)
    p = kernel_size // 2

    def padf(t):
        if p == 0:
            return t
        if padding == "constant":
            return F.pad(t, (p, p, p, p), mode="constant", value=0.0)
        return F.pad(t, (p, p, p, p), mode=padding)

    xp, yp = padf(pred), padf(target)
    mx = F.conv2d(xp, win, groups=c)
    my = F.conv2d(yp, win, groups=c)
    mx2, my2, mxy = mx.pow(2), my.pow(2), mx * my
    sx = torch.clamp(F.conv2d(xp * xp, win, groups=c) - mx2, min=0.0)
    sy = torch.clamp(F.conv2d(yp * yp, win, groups=c) - my2, min=0.0)
    sxy = F.conv2d(xp * yp, win, groups=c) - mxy
    c1 = (k1 * float(data_range)) ** 2
    c2 = (k2 * float(data_range)) ** 2
    smap = ((2 * mx * my + c1) * (2 * sxy + c2)) / ((mx2 + my2 + c1) * (sx + sy + c2) + eps)
    per_c = smap.flatten(2).mean(dim=2)
    if per_channel_raw:
        return per_c
    per = per_c.mean(dim=1) if channel_aggregate == "mean" else per_c
    return _batch(per, reduction)


def rgb_ssim(pred, target, *, reduction="mean", **kw):
    """channelwise.py:262-295 with channel_aggregate='none'."""
    if pred.ndim == 3:
        pred, target = pred[None], target[None]
    p32, t32 = pred.to(torch.float32), target.to(torch.float32)
    cols = [ssim_linear(p32[:, c:c + 1], t32[:, c:c + 1], reduction="none", channel_aggregate="mean", **kw).view(-1)
            for c in range(3)]
    st = torch.stack(cols, dim=1)
    return {"R": _batch(st[:, 0], reduction), "G": _batch(st[:, 1], reduction),
            "B": _batch(st[:, 2], reduction), "mean": _batch(st.mean(dim=1), reduction)}


# ------------------------------------------------------------------ torchmetrics 1.2.0 SSIM (restated; unpinned)
def tm_gaussian_1d(k, sigma, dtype):
    dist = torch.arange(start=(1 - k) / 2, end=(1 + k) / 2, step=1, dtype=dtype)
    g = torch.exp(-torch.pow(dist / sigma, 2) / 2)
    return (g / g.sum()).unsqueeze(0)


def ssim_torchmetrics(preds, target, *, data_range=1.0, sigma=1.5, kernel_size=11, k1=0.01, k2=0.03,
                      compute_dtype=torch.float32):
    """Per-image scores, as StructuralSimilarityIndexMeasure(gaussian_kernel=True, reduction='none').
    The Gaussian support is derived from sigma (2*int(3.5*sigma+0.5)+1), NOT from kernel_size.
    `compute_dtype=torch.float64` evaluates the same formula in double (separates "we differ"
    from "fp32 eager is itself noisy" — E[x^2] - mu^2 cancels badly on smooth 8-bit images)."""
    if preds.ndim == 3:
        preds, target = preds[None], target[None]
    preds, target = preds.to(compute_dtype), target.to(compute_dtype)
    c = preds.shape[1]
    gk = int(3.5 * sigma + 0.5) * 2 + 1
    pad = (gk - 1) // 2
    c1 = (k1 * data_range) ** 2
    c2 = (k2 * data_range) ** 2
    g = tm_gaussian_1d(gk, sigma, preds.dtype)
    win = torch.matmul(g.t(), g).expand(c, 1, gk, gk).to(preds.device)
    xp = F.pad(preds, (pad, pad, pad, pad), mode="reflect")
    yp = F.pad(target, (pad, pad, pad, pad), mode="reflect")
    stack = torch.cat((xp, yp, xp * xp, yp * yp, xp * yp))
    out = F.conv2d(stack, win, groups=c).split(preds.shape[0])
    mx2, my2, mxy = out[0].pow(2), out[1].pow(2), out[0] * out[1]
    sx, sy, sxy = out[2] - mx2, out[3] - my2, out[4] - mxy
    full = ((2 * mxy + c1) * (2 * sxy + c2)) / ((mx2 + my2 + c1) * (sx + sy + c2))
    inner = full[..., pad:-pad, pad:-pad]
    return inner.reshape(inner.shape[0], -1).mean(-1)


def luma_bt601(x):
    """ssim.py:116-127."""
    return 0.2989 * x[:, 0:1] + 0.5870 * x[:, 1:2] + 0.1140 * x[:, 2:3]


def auto_data_range(x, y):
    """ssim.py:62-71."""
    lo = min(float(x.min()), float(y.min()))
    hi = max(float(x.max()), float(y.max()))
    if hi <= 1.0 and lo >= 0.0:
        return 1.0
    if hi <= 1.0 and lo >= -1.0:
        return 2.0
    return 255.0


# ------------------------------------------------------------------ kornia 0.6.12 SSIM loss (restated; unpinned)
def kornia_gaussian_1d(k, sigma, dtype=torch.float32):
    x = torch.arange(k, dtype=dtype) - k // 2
    if k % 2 == 0:
        x = x + 0.5
    g = torch.exp(-x.pow(2.0) / (2 * sigma ** 2))
    return g / g.sum()


def kornia_ssim_map(x, y, window_size=11, max_val=1.0, eps=1e-12):
    c = x.shape[1]
    g = kornia_gaussian_1d(window_size, 1.5, x.dtype)
    win = torch.matmul(g.unsqueeze(-1), g.unsqueeze(-1).t()).expand(c, 1, window_size, window_size).to(x.device)
    p = window_size // 2

    def filt(t):
        return F.conv2d(F.pad(t, (p, p, p, p), mode="reflect"), win, groups=c)

    c1 = (0.01 * max_val) ** 2
    c2 = (0.03 * max_val) ** 2
    mx, my = filt(x), filt(y)
    mx2, my2, mxy = mx ** 2, my ** 2, mx * my
    sx = filt(x ** 2) - mx2
    sy = filt(y ** 2) - my2
    sxy = filt(x * y) - mxy
    return ((2.0 * mxy + c1) * (2.0 * sxy + c2)) / ((mx2 + my2 + c1) * (sx + sy + c2) + eps)


def ssim_loss_kornia(gen, tgt, window_size=11, max_val=1.0):
    """losses.py:154-155 — inputs clamped to [0,1]; mean(clamp((1-S)/2, 0, 1))."""
    s = kornia_ssim_map(gen.clamp(0, 1), tgt.clamp(0, 1), window_size, max_val)
    return torch.clamp((1.0 - s) / 2, min=0, max=1).mean()
